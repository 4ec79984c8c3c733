// Specialised attention kernels (placeholder until the tuned variants land).
#include "gat_kernels.cuh"

namespace b200gat {
bool fast_supported(const b200gat_attn*, const b200gat_edge_feats*) { return false; }
int fast_fwd(const b200gat_graph*, const b200gat_attn*, const b200gat_edge_feats*, const float*,
             const float*, const float*, float*, float*, float*, float*, cudaStream_t) {
  set_error("no specialised kernel for this shape");
  return B200GAT_ERR_UNSUPPORTED;
}
int fast_bwd(const b200gat_graph*, const b200gat_attn*, const b200gat_edge_feats*, const float*,
             const float*, const float*, const float*, const float*, const float*, const float*,
             float*, float*, float*, float*, float*, float*, float*, float*, float*, cudaStream_t) {
  set_error("no specialised kernel for this shape");
  return B200GAT_ERR_UNSUPPORTED;
}
}  // namespace b200gat
