// Fused GELU MLP passes: the GEMM sequence of generic/mlp.py:70-77 (forward) and of its autograd (backward)
// issued from one C-ABI call each on b200gat_gemm_tf32 (gemm_tf32.cu).  Host code only: the eager PPO step is
// bound by the host launching ~350 kernels, so one binding call per pass instead of one per GEMM matters.
#include <string.h>
#include "common.cuh"

using namespace b200gat;

namespace {

bool check_mlp(const b200gat_mlp* m, const char* who) {
  if (!m || m->n_layers < 1 || m->n_layers > B200GAT_MLP_MAX_LAYERS || m->M <= 0 || !m->x || !m->y) {
    set_error("%s: bad descriptor (layers / rows / null x or y)", who);
    return false;
  }
  for (int i = 0; i <= m->n_layers; ++i)
    if (m->dims[i] <= 0 || m->dims[i] % 4 != 0) {
      set_error("%s: width %d of layer boundary %d is not a positive multiple of 4", who, m->dims[i], i);
      return false;
    }
  for (int i = 0; i < m->n_layers; ++i) {
    if (!m->W[i]) {
      set_error("%s: null weight %d", who, i);
      return false;
    }
    if (i < m->n_layers - 1 && !m->h[i]) {
      set_error("%s: null activation buffer %d", who, i);
      return false;
    }
  }
  return true;
}

void base_gemm(b200gat_gemm* g) { memset(g, 0, sizeof(*g)); }

// dW_i = dz^T h_in (+ row sums): the descriptor of layer i's weight gradient
void dw_desc(b200gat_gemm* g, const b200gat_mlp* m, const b200gat_mlp_grads* gr, int i, const float* dz) {
  base_gemm(g);
  g->a_major = 1; g->b_major = 1;
  g->M = m->dims[i + 1]; g->N = m->dims[i]; g->K = m->M;
  g->A = dz; g->lda = m->dims[i + 1];
  g->B = i == 0 ? m->x : m->h[i - 1]; g->ldb = m->dims[i];
  g->D = gr->dW[i]; g->ldd = m->dims[i];
  g->a_rowsum = gr->db[i];
}

}  // namespace

extern "C" int b200gat_mlp_fwd(const b200gat_mlp* m, void* stream) {
  if (!check_mlp(m, "b200gat_mlp_fwd")) return B200GAT_ERR_INVALID;
  const int n = m->n_layers;
  const float* in = m->x;
  for (int i = 0; i < n; ++i) {
    b200gat_gemm g;
    base_gemm(&g);
    g.M = m->M; g.N = m->dims[i + 1]; g.K = m->dims[i];
    g.A = in; g.lda = m->dims[i];
    g.B = m->W[i]; g.ldb = m->dims[i];
    g.bias = m->b[i];
    if (i < n - 1) {
      g.D = m->z[i]; g.ldd = m->dims[i + 1];
      g.epilogue = B200GAT_EPI_GELU; g.D2 = m->h[i]; g.ld2 = m->dims[i + 1];
      in = m->h[i];
    } else {
      g.D = m->y; g.ldd = m->dims[i + 1];
      g.residual = m->residual; g.ldr = m->dims[i + 1];
    }
    const int rc = b200gat_gemm_tf32(&g, stream);
    if (rc) return rc;
  }
  return B200GAT_OK;
}

extern "C" int64_t b200gat_mlp_bwd_workspace_elems(const b200gat_mlp* m, const b200gat_mlp_grads* gr) {
  if (!m || !gr || m->n_layers < 1 || m->n_layers > B200GAT_MLP_MAX_LAYERS) return 0;
  int64_t need = 0;
  for (int i = 0; i < m->n_layers; ++i) {
    if (!gr->dW[i]) continue;
    b200gat_gemm g;
    dw_desc(&g, m, gr, i, gr->gy);            // the operand pointer does not change the sizing
    const int64_t e = b200gat_gemm_workspace_elems(&g);
    if (e > need) need = e;
  }
  return need;
}

int b200gat::mlp_bwd_dw(const b200gat_mlp* m, const b200gat_mlp_grads* gr, void* stream) {
  if (!check_mlp(m, "mlp_bwd_dw")) return B200GAT_ERR_INVALID;
  B200_CHECK_ARG(gr && gr->gy, "null gradient descriptor / gy");
  const int n = m->n_layers;
  for (int i = n - 1; i >= 0; --i) {
    if (!gr->dW[i]) {
      B200_CHECK_ARG(!gr->db[i], "a bias gradient needs its weight gradient (row sums ride on that GEMM)");
      continue;
    }
    const float* dz = i == n - 1 ? gr->gy : gr->dz[i];
    B200_CHECK_ARG(dz, "weight gradients need dz of every hidden layer");
    b200gat_gemm g;
    dw_desc(&g, m, gr, i, dz);
    g.workspace = gr->workspace; g.workspace_elems = gr->workspace_elems;
    if (b200gat_gemm_workspace_elems(&g) > gr->workspace_elems) g.workspace = nullptr;   // runs unsplit
    const int rc = b200gat_gemm_tf32(&g, stream);
    if (rc) return rc;
  }
  return B200GAT_OK;
}

extern "C" int b200gat_mlp_bwd(const b200gat_mlp* m, const b200gat_mlp_grads* gr, void* stream) {
  if (!check_mlp(m, "b200gat_mlp_bwd")) return B200GAT_ERR_INVALID;
  B200_CHECK_ARG(gr && gr->gy, "null gradient descriptor / gy");
  const int n = m->n_layers;
  for (int i = 0; i < n - 1; ++i) B200_CHECK_ARG(m->z[i] && gr->dz[i], "backward needs z and dz of every hidden layer");
  const float* dz = gr->gy;
  for (int i = n - 1; i >= 0; --i) {
    if (gr->dW[i]) {
      b200gat_gemm g;
      dw_desc(&g, m, gr, i, dz);
      g.workspace = gr->workspace; g.workspace_elems = gr->workspace_elems;
      if (b200gat_gemm_workspace_elems(&g) > gr->workspace_elems) g.workspace = nullptr;   // runs unsplit
      const int rc = b200gat_gemm_tf32(&g, stream);
      if (rc) return rc;
    } else {
      B200_CHECK_ARG(!gr->db[i], "a bias gradient needs its weight gradient (row sums ride on that GEMM)");
    }
    float* out = i > 0 ? gr->dz[i - 1] : gr->dx;
    if (!out) break;                            // no input gradient wanted
    b200gat_gemm g;
    base_gemm(&g);
    g.a_major = 0; g.b_major = 1;
    g.M = m->M; g.N = m->dims[i]; g.K = m->dims[i + 1];
    g.A = dz; g.lda = m->dims[i + 1];
    g.B = m->W[i]; g.ldb = m->dims[i];
    g.D = out; g.ldd = m->dims[i];
    if (i > 0) {
      g.epilogue = B200GAT_EPI_DGELU; g.Zin = m->z[i - 1]; g.ld2 = m->dims[i];
    }
    const int rc = b200gat_gemm_tf32(&g, stream);
    if (rc) return rc;
    dz = out;
  }
  return B200GAT_OK;
}
