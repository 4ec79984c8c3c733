// C-ABI dispatch for the fused attention layer:
//   dense clique path  (blocks detected, additive flavour, D == 64, edge mode NONE/TABLE):
//       sparse kernels on g->residual (clique members keep a partial softmax state) + clique kernels
//   specialised sparse kernels when the shape is covered (and B200GAT_FORCE_GENERIC is not set),
//   generic kernels otherwise.
#include <stdlib.h>
#include "gat_kernels.cuh"

using namespace b200gat;

static bool force_generic() {
  const char* e = getenv("B200GAT_FORCE_GENERIC");
  return e && e[0] == '1';
}
static bool no_cliques() {
  const char* e = getenv("B200GAT_NO_CLIQUES");
  return e && e[0] == '1';
}
static bool use_cliques(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef) {
  return !force_generic() && !no_cliques() && fast_supported(a, ef) && clique_supported(g, a, ef);
}

extern "C" int b200gat_conv_uses_cliques(const b200gat_graph* g, const b200gat_attn* a,
                                         const b200gat_edge_feats* ef) {
  return (g && a && use_cliques(g, a, ef)) ? 1 : 0;
}

extern "C" int b200gat_conv_fwd(const b200gat_graph* g, const b200gat_attn* a,
                                const b200gat_edge_feats* ef, const float* Pi, const float* Pj,
                                const float* V, float* out, float* s_save, float* m_save,
                                float* l_save, float* agg_save, float* s_dense, void* stream) {
  B200_CHECK_ARG(g && a, "null descriptor");
  if (g->n_nodes > 0) B200_CHECK_ARG(Pi && Pj && V && out && m_save && l_save, "null tensor pointer");
  if (g->n_edges > 0) B200_CHECK_ARG(s_save, "null s_save");
  cudaStream_t st = (cudaStream_t)stream;
  if (use_cliques(g, a, ef)) {
    B200_CHECK_ARG(s_dense, "null s_dense (the graph has clique blocks)");
    int rc = fast_fwd(g->residual, a, ef, Pi, Pj, V, out, s_save, m_save, l_save, agg_save,
                      g->cliques->node_blk, st);
    if (rc == B200GAT_ERR_UNSUPPORTED) set_error("clique path: tensors must be 16-byte aligned");
    if (rc) return rc;
    return clique_fwd(g, a, ef, Pi, Pj, V, out, s_dense, m_save, l_save, agg_save, st);
  }
  if (!force_generic() && fast_supported(a, ef)) {
    int rc = fast_fwd(g, a, ef, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, nullptr, st);
    if (rc != B200GAT_ERR_UNSUPPORTED) return rc;
  }
  return generic_fwd(g, a, ef, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, st);
}

extern "C" int b200gat_conv_bwd(const b200gat_graph* g, const b200gat_attn* a,
                                const b200gat_edge_feats* ef, const float* Pi, const float* Pj,
                                const float* V, const float* s_save, const float* m_save,
                                const float* l_save, const float* agg_save, const float* s_dense,
                                const float* g_out, float* gPi, float* gPj, float* gV, float* g_attn,
                                float* g_attn_bias, float* g_table, float* g_dense, float* g_affine,
                                float* gs_scratch, void* stream) {
  B200_CHECK_ARG(g && a, "null descriptor");
  if (g->n_nodes > 0)
    B200_CHECK_ARG(Pi && Pj && V && m_save && l_save && g_out && gPi && gPj && gV && g_attn,
                   "null tensor pointer");
  if (g->n_edges > 0) B200_CHECK_ARG(s_save && gs_scratch, "null edge buffer");
  if (a->flavour == B200GAT_FLAVOUR_DOT) B200_CHECK_ARG(g_attn_bias, "null g_attn_bias");
  cudaStream_t st = (cudaStream_t)stream;
  if (use_cliques(g, a, ef)) {
    B200_CHECK_ARG(s_dense && agg_save, "null s_dense / agg_save (the graph has clique blocks)");
    int rc = fast_bwd(g->residual, a, ef, Pi, Pj, V, s_save, m_save, l_save, agg_save, g_out, gPi, gPj, gV,
                      g_attn, g_attn_bias, g_table, g_dense, g_affine, gs_scratch, g->n_edges, st);
    if (rc == B200GAT_ERR_UNSUPPORTED) set_error("clique path: tensors must be 16-byte aligned");
    if (rc) return rc;
    float* extra = gs_scratch + 2 * (int64_t)g->n_edges * a->heads;
    return clique_bwd(g, a, ef, Pi, Pj, V, s_dense, m_save, l_save, agg_save, g_out, gPi, gPj, gV, g_attn,
                      g_table, extra, st);
  }
  if (!force_generic() && fast_supported(a, ef)) {
    int rc = fast_bwd(g, a, ef, Pi, Pj, V, s_save, m_save, l_save, agg_save, g_out, gPi, gPj, gV, g_attn,
                      g_attn_bias, g_table, g_dense, g_affine, gs_scratch, 0, st);
    if (rc != B200GAT_ERR_UNSUPPORTED) return rc;
  }
  return generic_bwd(g, a, ef, Pi, Pj, V, s_save, m_save, l_save, g_out, gPi, gPj, gV, g_attn,
                     g_attn_bias, g_table, g_dense, g_affine, gs_scratch, st);
}
