// C-ABI dispatch for the fused attention layer: specialised kernels when the shape is
// covered (and B200GAT_FORCE_GENERIC is not set), the generic kernels otherwise.
#include <stdlib.h>
#include "gat_kernels.cuh"

using namespace b200gat;

static bool force_generic() {
  const char* e = getenv("B200GAT_FORCE_GENERIC");
  return e && e[0] == '1';
}

extern "C" int b200gat_conv_fwd(const b200gat_graph* g, const b200gat_attn* a,
                                const b200gat_edge_feats* ef, const float* Pi, const float* Pj,
                                const float* V, float* out, float* s_save, float* m_save,
                                float* l_save, float* agg_save, void* stream) {
  B200_CHECK_ARG(g && a, "null descriptor");
  if (g->n_nodes > 0) B200_CHECK_ARG(Pi && Pj && V && out && m_save && l_save, "null tensor pointer");
  if (g->n_edges > 0) B200_CHECK_ARG(s_save, "null s_save");
  if (!force_generic() && fast_supported(a, ef)) {
    int rc = fast_fwd(g, a, ef, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, (cudaStream_t)stream);
    if (rc != B200GAT_ERR_UNSUPPORTED) return rc;
  }
  return generic_fwd(g, a, ef, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, (cudaStream_t)stream);
}

extern "C" int b200gat_conv_bwd(const b200gat_graph* g, const b200gat_attn* a,
                                const b200gat_edge_feats* ef, const float* Pi, const float* Pj,
                                const float* V, const float* s_save, const float* m_save,
                                const float* l_save, const float* agg_save, const float* g_out,
                                float* gPi, float* gPj, float* gV, float* g_attn, float* g_attn_bias,
                                float* g_table, float* g_dense, float* g_affine, float* gs_scratch,
                                void* stream) {
  B200_CHECK_ARG(g && a, "null descriptor");
  if (g->n_nodes > 0)
    B200_CHECK_ARG(Pi && Pj && V && m_save && l_save && g_out && gPi && gPj && gV && g_attn,
                   "null tensor pointer");
  if (g->n_edges > 0) B200_CHECK_ARG(s_save && gs_scratch, "null edge buffer");
  if (a->flavour == B200GAT_FLAVOUR_DOT) B200_CHECK_ARG(g_attn_bias, "null g_attn_bias");
  if (!force_generic() && fast_supported(a, ef)) {
    int rc = fast_bwd(g, a, ef, Pi, Pj, V, s_save, m_save, l_save, agg_save, g_out, gPi, gPj, gV, g_attn,
                      g_attn_bias, g_table, g_dense, g_affine, gs_scratch, (cudaStream_t)stream);
    if (rc != B200GAT_ERR_UNSUPPORTED) return rc;
  }
  return generic_bwd(g, a, ef, Pi, Pj, V, s_save, m_save, l_save, g_out, gPi, gPj, gV, g_attn,
                     g_attn_bias, g_table, g_dense, g_affine, gs_scratch, (cudaStream_t)stream);
}
