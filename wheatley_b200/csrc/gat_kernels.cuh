// Internal kernel-family entry points behind b200gat_conv_fwd / b200gat_conv_bwd.
#pragma once
#include "common.cuh"

namespace b200gat {

int generic_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                const float* Pi, const float* Pj, const float* V, float* out, float* s_save,
                float* m_save, float* l_save, float* agg_save, cudaStream_t st);
int generic_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                const float* Pi, const float* Pj, const float* V, const float* s_save,
                const float* m_save, const float* l_save, const float* g_out, float* gPi, float* gPj,
                float* gV, float* g_attn, float* g_attn_bias, float* g_table, float* g_dense,
                float* g_affine, float* gs, cudaStream_t st);

// Specialised kernels (gat_fast.cu); return B200GAT_ERR_UNSUPPORTED without touching
// anything when the shape / alignment is not covered, so the caller falls through to the
// generic path.
bool fast_supported(const b200gat_attn* a, const b200gat_edge_feats* ef);
// node_blk != NULL: nodes with node_blk[v] >= 0 keep their un-normalised softmax state (clique members)
int fast_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
             const float* Pi, const float* Pj, const float* V, float* out, float* s_save,
             float* m_save, float* l_save, float* agg_save, const int32_t* node_blk, cudaStream_t st);
int fast_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
             const float* Pi, const float* Pj, const float* V, const float* s_save,
             const float* m_save, const float* l_save, const float* agg_save, const float* g_out,
             float* gPi, float* gPj, float* gV, float* g_attn, float* g_attn_bias, float* g_table,
             float* g_dense, float* g_affine, float* gs, int64_t n_pos, cudaStream_t st);
// n_pos: number of edge positions the gs / alpha scratch is indexed by (the FULL graph's edge count when
// g is a residual graph); 0 = g->n_edges.

// Dense clique-block kernels (gat_clique.cu): run after the sparse kernels on g->residual.
bool clique_supported(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef);
int clique_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef, const float* Pi,
               const float* Pj, const float* V, float* out, float* s_dense, float* m_save, float* l_save,
               float* agg_save, cudaStream_t st);
int clique_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef, const float* Pi,
               const float* Pj, const float* V, const float* s_dense, const float* m_save, const float* l_save,
               const float* agg_save, const float* g_out, float* gPi, float* gPj, float* gV, float* g_attn,
               float* g_table, float* scratch, cudaStream_t st);

}  // namespace b200gat
