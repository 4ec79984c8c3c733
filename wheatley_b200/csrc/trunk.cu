// Whole-trunk passes: the L (projection GEMM -> fused attention -> GELU MLP (+ residual)) layers of
// jssp/models/gnn_dgl.py:366-399 / psp/models/gnn_flat.py:170-253 issued from ONE C-ABI call per pass.
// Host code plus two small glue kernels; every heavy launch is one of the library's own entry points
// (b200gat_gemm_tf32, b200gat_conv_fwd/bwd, b200gat_mlp_fwd/bwd, b200gat_colsum).
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include "common.cuh"

namespace b200gat {

struct CopySegs {                       // up to 6 (weights + biases of one layer) x 16 layers
  const float* src[96];
  int32_t dst_off[96];                  // float offset into wcat (multiples of 4)
  int32_t n4[96];                       // float4 count
  int32_t n;
};

// gridDim.x = segments; every segment is a whole number of float4 (H*D*in_dim and H*D are multiples of 4)
__global__ void __launch_bounds__(256) k_concat_segs(CopySegs S, float* __restrict__ dst) {
  const int s = blockIdx.y;
  const float4* __restrict__ src = reinterpret_cast<const float4*>(S.src[s]);
  float4* __restrict__ d = reinterpret_cast<float4*>(dst + S.dst_off[s]);
  for (int i = blockIdx.x * 256 + threadIdx.x; i < S.n4[s]; i += gridDim.x * 256) d[i] = __ldg(src + i);
}

__global__ void __launch_bounds__(256) k_add2(const float4* __restrict__ a, const float4* __restrict__ b,
                                              float4* __restrict__ o, int64_t n4) {
  pdl_sync();
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < n4) {
    const float4 x = a[i], y = b[i];
    o[i] = make_float4(x.x + y.x, x.y + y.y, x.z + y.z, x.w + y.w);
  }
}

static inline int64_t up4(int64_t x) { return (x + 3) & ~3ll; }

// Side stream of the backward pass (declared in common.cuh, shared with agent.cu).  The weight-gradient GEMMs (split-K
// over the node dimension + their reduce, 13 % of the step's kernel time) feed nothing but the optimizer, so they are
// issued on a second stream, forked from and joined back into the caller's stream with events: they run under the
// attention backward of the same layer, whose clique kernels leave a third of the SMs idle in their last wave.  Works
// the same under CUDA graph capture (the side stream joins the capture through the event).  One stream + four events
// per device, created on the first call that is not being captured (creating them is not a capturable operation);
// B200GAT_DW_OVERLAP=0 disables it.
AuxStream* aux_stream_for(cudaStream_t main) {
  static const bool enabled = [] {
    const char* e = getenv("B200GAT_DW_OVERLAP");
    return !(e && e[0] == '0');
  }();
  if (!enabled || timing_enabled()) return nullptr;      // per-class timing wants the serial order
  constexpr int kMaxDev = 64;
  static AuxStream aux[kMaxDev];
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) return nullptr;
  std::lock_guard<std::mutex> lk(mu);
  AuxStream& a = aux[dev];
  if (!a.ok) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(main, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return nullptr;
    if (cudaStreamCreateWithFlags(&a.s, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
    cudaEvent_t* ev[4] = {&a.m1, &a.m2, &a.a1, &a.a2};
    for (auto e : ev)
      if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    a.ok = true;
  }
  return &a;
}
bool aux_fork(cudaStream_t main, cudaStream_t aux, cudaEvent_t ev) {
  return cudaEventRecord(ev, main) == cudaSuccess && cudaStreamWaitEvent(aux, ev, 0) == cudaSuccess;
}
bool aux_join(cudaStream_t main, cudaStream_t aux, cudaEvent_t ev) {
  return cudaEventRecord(ev, aux) == cudaSuccess && cudaStreamWaitEvent(main, ev, 0) == cudaSuccess;
}

struct TrunkShape {
  int HD;
  int64_t wcat_off[B200GAT_TRUNK_MAX_LAYERS];   // per layer: [n_proj*HD*in_dim weights | n_proj*HD biases]
  int64_t wcat_elems;
  // backward workspace offsets (floats)
  int64_t o_G[2], o_R, o_gP, o_gconv, o_dz[B200GAT_MLP_MAX_LAYERS], o_gs, o_gemm, o_colsum, total;
  int64_t gemm_elems, colsum_elems;
};

static bool aligned16t(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Validation shared by every entry point; fills the shape when `sh` is given.
static int check_trunk(const b200gat_trunk* t, TrunkShape* sh, bool report) {
#define TR_REQ(cond, msg)                                   \
  do {                                                      \
    if (!(cond)) {                                          \
      if (report) set_error("b200gat_trunk: %s", msg);      \
      return B200GAT_ERR_UNSUPPORTED;                       \
    }                                                       \
  } while (0)
  TR_REQ(t && t->graph, "null descriptor / graph");
  TR_REQ(t->n_layers >= 1 && t->n_layers <= B200GAT_TRUNK_MAX_LAYERS, "1 <= n_layers <= 16");
  TR_REQ(t->n_nodes > 0 && t->n_nodes == t->graph->n_nodes, "n_nodes must match the graph");
  TR_REQ(t->in_dim > 0 && t->in_dim % 32 == 0, "in_dim must be a multiple of 32");
  const b200gat_trunk_layer& l0 = t->layer[0];
  const int HD = l0.attn.heads * l0.attn.dim;
  TR_REQ(HD > 0 && HD % 128 == 0, "H*D must be a multiple of 128");
  int64_t off = 0, max_hidden = 0;
  for (int l = 0; l < t->n_layers; ++l) {
    const b200gat_trunk_layer& L = t->layer[l];
    TR_REQ(L.n_proj == 2 || L.n_proj == 3, "n_proj is 2 or 3");
    TR_REQ(L.attn.flavour == B200GAT_FLAVOUR_ADDITIVE && L.attn.heads * L.attn.dim == HD, "additive flavour, same H*D");
    TR_REQ(L.ef.mode != B200GAT_EDGE_DENSE, "dense edge features are not covered");
    TR_REQ(L.n_mlp >= 1 && L.n_mlp <= B200GAT_MLP_MAX_LAYERS, "1 <= n_mlp <= 8");
    TR_REQ(L.dims[0] == HD && L.dims[L.n_mlp] == t->in_dim, "MLP maps H*D to in_dim");
    for (int i = 0; i <= L.n_mlp; ++i) TR_REQ(L.dims[i] > 0 && L.dims[i] % 4 == 0, "MLP widths must be multiples of 4");
    const bool has_b = L.bp[0] != nullptr;
    for (int p = 0; p < L.n_proj; ++p) {
      TR_REQ(L.Wp[p] && aligned16t(L.Wp[p]), "projection weights (16-byte aligned)");
      TR_REQ((L.bp[p] != nullptr) == has_b && aligned16t(L.bp[p]), "projection biases: all or none, 16-byte aligned");
    }
    if (sh) {
      sh->wcat_off[l] = off;
      for (int i = 1; i < L.n_mlp; ++i) max_hidden = L.dims[i] > max_hidden ? L.dims[i] : max_hidden;
    }
    off += up4((int64_t)L.n_proj * HD * t->in_dim) + up4((int64_t)L.n_proj * HD);
  }
  if (sh) {
    sh->HD = HD;
    sh->wcat_elems = off;
  }
  return B200GAT_OK;
#undef TR_REQ
}

static void proj_fwd_desc(b200gat_gemm* g, const b200gat_trunk* t, const TrunkShape& sh, int l, const float* x) {
  const b200gat_trunk_layer& L = t->layer[l];
  memset(g, 0, sizeof(*g));
  g->M = t->n_nodes; g->N = (int64_t)L.n_proj * sh.HD; g->K = t->in_dim;
  g->A = x; g->lda = t->in_dim;
  g->B = t->wcat + sh.wcat_off[l]; g->ldb = t->in_dim;
  g->D = L.P; g->ldd = sh.HD; g->d_parts = L.n_proj; g->d_part_stride = t->n_nodes * sh.HD;
  if (L.bp[0]) g->bias = t->wcat + sh.wcat_off[l] + up4((int64_t)L.n_proj * sh.HD * t->in_dim);
}

static void proj_dw_desc(b200gat_gemm* g, const b200gat_trunk* t, const TrunkShape& sh, int l, const float* x,
                         const float* gP, float* dWp) {
  const b200gat_trunk_layer& L = t->layer[l];
  memset(g, 0, sizeof(*g));
  g->a_major = 1; g->b_major = 1;
  g->M = (int64_t)L.n_proj * sh.HD; g->N = t->in_dim; g->K = t->n_nodes;
  g->A = gP; g->lda = sh.HD; g->a_parts = L.n_proj; g->a_part_stride = t->n_nodes * sh.HD;
  g->B = x; g->ldb = t->in_dim;
  g->D = dWp; g->ldd = t->in_dim;
}

static void mlp_desc(b200gat_mlp* m, const b200gat_trunk* t, int l, const float* residual) {
  const b200gat_trunk_layer& L = t->layer[l];
  memset(m, 0, sizeof(*m));
  m->n_layers = L.n_mlp; m->M = t->n_nodes;
  for (int i = 0; i <= L.n_mlp; ++i) m->dims[i] = L.dims[i];
  for (int i = 0; i < L.n_mlp; ++i) { m->W[i] = L.W[i]; m->b[i] = L.b[i]; m->z[i] = L.z[i]; m->h[i] = L.h[i]; }
  m->x = L.conv_out; m->residual = residual; m->y = L.y;
}

static int bwd_shape(const b200gat_trunk* t, TrunkShape* sh) {
  int rc = check_trunk(t, sh, true);
  if (rc) return rc;
  const int64_t N = t->n_nodes, HD = sh->HD, E = t->graph->n_edges, H = t->layer[0].attn.heads;
  int64_t off = 0;
  auto take = [&](int64_t n) { const int64_t o = off; off += up4(n); return o; };
  sh->o_G[0] = take(N * t->in_dim);
  sh->o_G[1] = take(N * t->in_dim);
  sh->o_R = take(N * t->in_dim);
  sh->o_gP = take(3 * N * HD);
  sh->o_gconv = take(N * HD);
  int64_t dz_w[B200GAT_MLP_MAX_LAYERS] = {0};
  int64_t gemm_ws = 0;
  for (int l = 0; l < t->n_layers; ++l) {
    const b200gat_trunk_layer& L = t->layer[l];
    for (int i = 0; i < L.n_mlp - 1; ++i) dz_w[i] = L.dims[i + 1] > dz_w[i] ? L.dims[i + 1] : dz_w[i];
    b200gat_mlp m;
    mlp_desc(&m, t, l, nullptr);
    b200gat_mlp_grads gr;
    memset(&gr, 0, sizeof(gr));
    alignas(16) static float dummy[4];
    for (int i = 0; i < L.n_mlp; ++i) { gr.dW[i] = dummy; gr.db[i] = dummy; }   // sizing only (plan_gemm checks alignment)
    gr.gy = dummy;
    int64_t e = b200gat_mlp_bwd_workspace_elems(&m, &gr);
    gemm_ws = e > gemm_ws ? e : gemm_ws;
    b200gat_gemm g;
    proj_dw_desc(&g, t, *sh, l, dummy, dummy, dummy);
    e = b200gat_gemm_workspace_elems(&g);
    gemm_ws = e > gemm_ws ? e : gemm_ws;
  }
  for (int i = 0; i < B200GAT_MLP_MAX_LAYERS; ++i) sh->o_dz[i] = dz_w[i] ? take(N * dz_w[i]) : -1;
  const int64_t s_el = t->graph->cliques ? t->graph->cliques->s_elems : 0;
  sh->o_gs = take(2 * E * H + up4(H * s_el) + 2 * N * H);
  sh->gemm_elems = gemm_ws;
  sh->o_gemm = take(gemm_ws > 4 ? gemm_ws : 4);
  sh->colsum_elems = b200gat_colsum_scratch_elems(N, HD);
  sh->o_colsum = take(sh->colsum_elems > 4 ? sh->colsum_elems : 4);
  sh->total = off;
  return B200GAT_OK;
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int b200gat_trunk_supported(const b200gat_trunk* t) {
  TrunkShape sh;
  if (check_trunk(t, &sh, false) != B200GAT_OK) return 0;
  // the GEMM shapes themselves (the MLP ones are checked by b200gat_mlp_fwd's own requirements: widths % 4)
  alignas(16) static float dummy[4];
  b200gat_trunk probe = *t;
  probe.wcat = dummy;
  for (int l = 0; l < t->n_layers; ++l) {
    b200gat_gemm g;
    probe.layer[l].P = probe.wcat;
    proj_fwd_desc(&g, &probe, sh, l, probe.wcat);
    g.B = probe.wcat; g.bias = nullptr;
    if (!b200gat_gemm_supported(&g)) return 0;
  }
  return 1;
}

extern "C" int64_t b200gat_trunk_wcat_elems(const b200gat_trunk* t) {
  TrunkShape sh;
  return check_trunk(t, &sh, false) == B200GAT_OK ? sh.wcat_elems : 0;
}

extern "C" int64_t b200gat_trunk_bwd_workspace_elems(const b200gat_trunk* t) {
  TrunkShape sh;
  return bwd_shape(t, &sh) == B200GAT_OK ? sh.total : 0;
}

extern "C" int b200gat_trunk_fwd(const b200gat_trunk* t, void* stream) {
  TrunkShape sh;
  int rc = check_trunk(t, &sh, true);
  if (rc) return rc;
  B200_CHECK_ARG(t->x0 && t->wcat && aligned16t(t->x0) && aligned16t(t->wcat), "x0 / wcat (16-byte aligned)");
  cudaStream_t st = (cudaStream_t)stream;
  // all layers' projection weights (and biases) side by side: one launch
  {
    CopySegs S;
    S.n = 0;
    int max4 = 0;
    for (int l = 0; l < t->n_layers; ++l) {
      const b200gat_trunk_layer& L = t->layer[l];
      const int64_t wsz = (int64_t)sh.HD * t->in_dim;
      for (int p = 0; p < L.n_proj; ++p) {
        S.src[S.n] = L.Wp[p]; S.dst_off[S.n] = (int32_t)(sh.wcat_off[l] + p * wsz); S.n4[S.n] = (int32_t)(wsz / 4);
        max4 = S.n4[S.n] > max4 ? S.n4[S.n] : max4;
        ++S.n;
      }
      if (L.bp[0])
        for (int p = 0; p < L.n_proj; ++p) {
          S.src[S.n] = L.bp[p];
          S.dst_off[S.n] = (int32_t)(sh.wcat_off[l] + up4((int64_t)L.n_proj * wsz) + (int64_t)p * sh.HD);
          S.n4[S.n] = sh.HD / 4;
          ++S.n;
        }
    }
    B200_CHECK_ARG(sh.wcat_elems < (1ll << 31), "projection weights too large");
    dim3 grid((unsigned)ceil_div(max4, 256 * 4), (unsigned)S.n);
    k_concat_segs<<<grid, 256, 0, st>>>(S, t->wcat);
    B200_CHECK_LAUNCH("k_concat_segs");
  }
  const float* x = t->x0;
  for (int l = 0; l < t->n_layers; ++l) {
    const b200gat_trunk_layer& L = t->layer[l];
    B200_CHECK_ARG(L.P && L.conv_out && L.y && L.m_save && L.l_save, "null layer buffer");
    b200gat_gemm g;
    proj_fwd_desc(&g, t, sh, l, x);
    if ((rc = b200gat_gemm_tf32(&g, stream))) return rc;
    const int64_t part = t->n_nodes * sh.HD;
    const float* Pi = L.P;
    const float* Pj = L.P + part;
    const float* V = L.n_proj == 3 ? L.P + 2 * part : Pi;
    if ((rc = b200gat_conv_fwd(t->graph, &L.attn, &L.ef, Pi, Pj, V, L.conv_out, L.s_save, L.m_save, L.l_save,
                               L.agg_save, L.s_dense, stream)))
      return rc;
    b200gat_mlp m;
    mlp_desc(&m, t, l, t->residual ? x : nullptr);
    if ((rc = b200gat_mlp_fwd(&m, stream))) return rc;
    x = L.y;
  }
  return B200GAT_OK;
}

extern "C" int b200gat_trunk_bwd(const b200gat_trunk* t, const b200gat_trunk_grads* gr, void* stream) {
  TrunkShape sh;
  int rc = bwd_shape(t, &sh);
  if (rc) return rc;
  B200_CHECK_ARG(gr && gr->workspace && aligned16t(gr->workspace), "null / unaligned workspace");
  B200_CHECK_ARG(gr->workspace_elems >= sh.total, "workspace too small (b200gat_trunk_bwd_workspace_elems)");
  B200_CHECK_ARG(gr->layer[t->n_layers - 1].g_y, "the last layer's output needs a gradient");
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = gr->workspace;
  const int64_t N = t->n_nodes, HD = sh.HD, part = N * HD, n_in = N * t->in_dim;
  const float* Gy = gr->layer[t->n_layers - 1].g_y;          // total gradient of the current layer's output
  AuxStream* aux = aux_stream_for(st);
  bool mlp_dw_pending = false, proj_dw_pending = false;      // side-stream work main has not waited for yet
#define AUX_REQ(call)                                                        \
  do {                                                                       \
    if (!(call)) {                                                           \
      set_error("b200gat_trunk_bwd: event fork / join failed: %s", cudaGetErrorString(cudaGetLastError()));  \
      return B200GAT_ERR_CUDA;                                               \
    }                                                                        \
  } while (0)
  for (int l = t->n_layers - 1; l >= 0; --l) {
    const b200gat_trunk_layer& L = t->layer[l];
    const b200gat_trunk_layer_grads& G = gr->layer[l];
    const float* x = l == 0 ? t->x0 : t->layer[l - 1].y;
    B200_CHECK_ARG(G.dWp && G.g_attn, "null dWp / g_attn");
    // ---- MLP backward: Gy -> dW_i, db_i, gradient of the attention output
    b200gat_mlp m;
    mlp_desc(&m, t, l, nullptr);
    m.y = const_cast<float*>(Gy);                             // not read backward; must be non-null
    b200gat_mlp_grads mg;
    memset(&mg, 0, sizeof(mg));
    mg.gy = Gy;
    mg.dx = ws + sh.o_gconv;
    for (int i = 0; i < L.n_mlp - 1; ++i) mg.dz[i] = ws + sh.o_dz[i];
    mg.workspace = ws + sh.o_gemm; mg.workspace_elems = sh.gemm_elems;
    if (aux) {
      // the chain below overwrites dz[] (read by the previous layer's MLP weight gradients on the side stream)
      if (mlp_dw_pending) AUX_REQ(cudaStreamWaitEvent(st, aux->a1, 0) == cudaSuccess);
      mlp_dw_pending = false;
      if ((rc = b200gat_mlp_bwd(&m, &mg, stream))) return rc;             // input-gradient chain only
      for (int i = 0; i < L.n_mlp; ++i) { mg.dW[i] = G.dW[i]; mg.db[i] = G.db[i]; }
      AUX_REQ(aux_fork(st, aux->s, aux->m1));
      if ((rc = mlp_bwd_dw(&m, &mg, aux->s))) return rc;
      AUX_REQ(cudaEventRecord(aux->a1, aux->s) == cudaSuccess);
      mlp_dw_pending = true;
    } else {
      for (int i = 0; i < L.n_mlp; ++i) { mg.dW[i] = G.dW[i]; mg.db[i] = G.db[i]; }
      if ((rc = b200gat_mlp_bwd(&m, &mg, stream))) return rc;
    }
    const float* g_conv = ws + sh.o_gconv;
    if (G.g_out_bias && (rc = b200gat_colsum(g_conv, N, HD, HD, G.g_out_bias, ws + sh.o_colsum, stream))) return rc;
    // ---- fused attention backward: gPi | gPj | gV side by side
    float* gP = ws + sh.o_gP;
    const float* Pi = L.P;
    const float* Pj = L.P + part;
    const float* V = L.n_proj == 3 ? L.P + 2 * part : Pi;
    b200gat_attn a = L.attn;
    a.out_bias = nullptr;
    // the attention backward overwrites gP (read by the previous layer's projection weight gradient)
    if (proj_dw_pending) AUX_REQ(cudaStreamWaitEvent(st, aux->a2, 0) == cudaSuccess);
    proj_dw_pending = false;
    if ((rc = b200gat_conv_bwd(t->graph, &a, &L.ef, Pi, Pj, V, L.s_save, L.m_save, L.l_save, L.agg_save, L.s_dense,
                               g_conv, gP, gP + part, L.n_proj == 3 ? gP + 2 * part : gP, G.g_attn, nullptr,
                               G.g_table, nullptr, G.g_affine, ws + sh.o_gs, stream)))
      return rc;
    if (G.dbp)
      for (int p = 0; p < L.n_proj; ++p)
        if ((rc = b200gat_colsum(gP + p * part, N, HD, HD, G.dbp + (int64_t)p * HD, ws + sh.o_colsum, stream)))
          return rc;
    // ---- projection weight gradient
    b200gat_gemm g;
    proj_dw_desc(&g, t, sh, l, x, gP, G.dWp);
    g.workspace = ws + sh.o_gemm; g.workspace_elems = sh.gemm_elems;
    if (b200gat_gemm_workspace_elems(&g) > sh.gemm_elems) g.workspace = nullptr;
    if (aux) {
      AUX_REQ(aux_fork(st, aux->s, aux->m2));
      if ((rc = b200gat_gemm_tf32(&g, aux->s))) return rc;
      AUX_REQ(cudaEventRecord(aux->a2, aux->s) == cudaSuccess);
      proj_dw_pending = true;
    } else if ((rc = b200gat_gemm_tf32(&g, stream))) {
      return rc;
    }
    // ---- gradient of the layer's input = gP Wcat (+ residual path) (+ what reaches that tensor from outside)
    float* out = l > 0 ? ws + sh.o_G[(l - 1) & 1] : gr->g_x0;
    if (!out) break;
    const float* r1 = t->residual ? Gy : nullptr;
    const float* r2 = l > 0 ? gr->layer[l - 1].g_y : gr->g_x0_in;
    const float* R = r1 ? r1 : r2;
    if (r1 && r2) {
      B200_CHECK_ARG(aligned16t(r1) && aligned16t(r2), "gradients must be 16-byte aligned");
      launch_k(k_add2, dim3((unsigned)ceil_div(n_in / 4, 256)), dim3(256), 0, st, reinterpret_cast<const float4*>(r1),
               reinterpret_cast<const float4*>(r2), reinterpret_cast<float4*>(ws + sh.o_R), n_in / 4);
      B200_CHECK_LAUNCH("k_add2");
      R = ws + sh.o_R;
    }
    memset(&g, 0, sizeof(g));
    g.a_major = 0; g.b_major = 1;
    g.M = N; g.N = t->in_dim; g.K = (int64_t)L.n_proj * HD;
    g.A = gP; g.lda = HD; g.a_parts = L.n_proj; g.a_part_stride = part;
    g.B = t->wcat + sh.wcat_off[l]; g.ldb = t->in_dim;
    g.D = out; g.ldd = t->in_dim;
    g.residual = R; g.ldr = t->in_dim;
    if ((rc = b200gat_gemm_tf32(&g, stream))) return rc;
    Gy = out;
  }
  // join: everything queued on the side stream precedes whatever the caller queues next
  if (mlp_dw_pending) AUX_REQ(cudaStreamWaitEvent(st, aux->a1, 0) == cudaSuccess);
  if (proj_dw_pending) AUX_REQ(cudaStreamWaitEvent(st, aux->a2, 0) == cudaSuccess);
#undef AUX_REQ
  return B200GAT_OK;
}
