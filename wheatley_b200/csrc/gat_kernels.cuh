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
int fast_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
             const float* Pi, const float* Pj, const float* V, float* out, float* s_save,
             float* m_save, float* l_save, float* agg_save, cudaStream_t st);
int fast_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
             const float* Pi, const float* Pj, const float* V, const float* s_save,
             const float* m_save, const float* l_save, const float* agg_save, const float* g_out,
             float* gPi, float* gPj, float* gV, float* g_attn, float* g_attn_bias, float* g_table,
             float* g_dense, float* g_affine, float* gs, cudaStream_t st);

}  // namespace b200gat
