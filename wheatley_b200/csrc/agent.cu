// Whole PPO-minibatch evaluation from one C-ABI call per pass (include/b200gat.h "whole PPO-minibatch evaluation"):
// alg/ppo.py:575-711 = agent.get_action_and_value (generic/agent.py:190-231) + the clipped-PPO loss, forward and
// backward.  Host sequencing of the library's own entry points plus four small glue kernels.
#include <string.h>
#include "common.cuh"

namespace b200gat {

struct PadJobs {
  b200gat_pad_job j[B200GAT_AGENT_MAX_PAD];
  int n;
};

// grid (x: row chunks, y: job)
__global__ void __launch_bounds__(256) k_pad_jobs(PadJobs P) {
  const b200gat_pad_job& J = P.j[blockIdx.y];
  const int rows_dst = J.rows_dst > J.rows ? J.rows_dst : J.rows;
  const int64_t total = (int64_t)rows_dst * J.ld_dst;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int64_t r = i / J.ld_dst;
    const int c = (int)(i - r * J.ld_dst);
    J.dst[i] = (c < J.cols && r < J.rows) ? __ldg(J.src + r * J.ld_src + c) : 0.f;
  }
}

// logits[m] = ya[m * lda] (m < Ma) ; value[b] = yv[b * ldv] (b < B)
__global__ void __launch_bounds__(256) k_take_col0(const float* __restrict__ ya, int lda, int64_t Ma,
                                                   float* __restrict__ logits, const float* __restrict__ yv, int ldv,
                                                   int B, float* __restrict__ value) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < Ma) logits[i] = ya[i * lda];
  if (i < B) value[i] = yv[i * ldv];
}

// logprob[b] = logp[b, action[b]]
__global__ void __launch_bounds__(256) k_pick_action(const float* __restrict__ logp, const int64_t* __restrict__ action,
                                                     int B, int n, float* __restrict__ logprob) {
  const int b = blockIdx.x * 256 + threadIdx.x;
  if (b < B) {
    const int64_t a = action[b];
    logprob[b] = (a >= 0 && a < n) ? logp[(int64_t)b * n + a] : 0.f;
  }
}

// gy_a[m, :] = (g_logits[m] * s, 0, 0, 0) ; gy_v[b, :] = (g_value[b] * s, 0, ...) ; g_lp / g_ent scaled in place
__global__ void __launch_bounds__(256) k_spread_col0(const float* __restrict__ g_logits, int64_t Ma, int lda,
                                                     float* __restrict__ gy_a, const float* __restrict__ g_value, int B,
                                                     int ldv, float* __restrict__ gy_v) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i < Ma) {
    gy_a[i * lda] = g_logits[i];
    for (int c = 1; c < lda; ++c) gy_a[i * lda + c] = 0.f;
  }
  if (i < B) {
    gy_v[i * ldv] = g_value[i];
    for (int c = 1; c < ldv; ++c) gy_v[i * ldv + c] = 0.f;
  }
}

__global__ void __launch_bounds__(256) k_scale_head(const float* __restrict__ g_head, const float* __restrict__ g_loss,
                                                    int n, float* __restrict__ out) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) out[i] = g_head[i] * (g_loss ? g_loss[0] : 1.f);
}

static inline int64_t up4a(int64_t x) { return (x + 3) & ~3ll; }

struct AgentShape {
  int B, max_n, F, ld, n_src;
  int64_t Ma;                 // rows of the action head = B * max_n
  int64_t N;                  // node rows of the sources
  int wa, wv;                 // padded output widths of the two heads
  // backward workspace
  int64_t o_ghead, o_glogits, o_gya, o_gyv, o_dxa, o_dxv, o_dz_a[B200GAT_MLP_MAX_LAYERS], o_dz_v[B200GAT_MLP_MAX_LAYERS],
      o_dz_e[B200GAT_MLP_MAX_LAYERS], o_gsrc[B200GAT_READOUT_MAX_SRC], o_gx0, o_gemm, o_trunk, total;
  int64_t gemm_elems, trunk_elems;
};

static int mlp_ws(const b200gat_mlp* m, int64_t* need) {
  b200gat_mlp_grads gr;
  memset(&gr, 0, sizeof(gr));
  alignas(16) static float dummy[4];
  for (int i = 0; i < m->n_layers; ++i) { gr.dW[i] = dummy; gr.db[i] = dummy; }
  gr.gy = dummy;
  const int64_t e = b200gat_mlp_bwd_workspace_elems(m, &gr);
  if (e > *need) *need = e;
  return 0;
}

static int agent_shape(const b200gat_agent_step* a, AgentShape* sh, bool bwd) {
  B200_CHECK_ARG(a, "null descriptor");
  const b200gat_readout& r = a->readout;
  B200_CHECK_ARG(r.n_src > 0 && r.n_src <= B200GAT_READOUT_MAX_SRC, "readout sources");
  sh->B = r.n_graphs; sh->max_n = r.max_n_nodes; sh->n_src = r.n_src;
  int F = 0;
  for (int i = 0; i < r.n_src; ++i) F += r.width[i];
  sh->F = F;
  sh->ld = r.out_ld ? (int)r.out_ld : F;
  B200_CHECK_ARG(sh->ld % 4 == 0, "readout.out_ld must be a multiple of 4 (it is the heads' input width)");
  sh->Ma = (int64_t)sh->B * sh->max_n;
  sh->N = a->trunk.n_nodes;
  B200_CHECK_ARG(a->action_net.n_layers >= 1 && a->value_net.n_layers >= 1, "heads");
  B200_CHECK_ARG(a->action_net.M == sh->Ma && a->action_net.dims[0] == sh->ld, "action_net input is node_features");
  B200_CHECK_ARG(a->value_net.M == sh->B && a->value_net.dims[0] == sh->ld, "value_net input is graph_embedding");
  sh->wa = a->action_net.dims[a->action_net.n_layers];
  sh->wv = a->value_net.dims[a->value_net.n_layers];
  B200_CHECK_ARG(a->loss.n == sh->B, "loss.n must be the number of graphs");
  if (!bwd) return B200GAT_OK;
  int64_t off = 0;
  auto take = [&](int64_t n) { const int64_t o = off; off += up4a(n > 0 ? n : 4); return o; };
  sh->o_ghead = take(3 * sh->B);
  sh->o_glogits = take(sh->Ma);
  sh->o_gya = take(sh->Ma * sh->wa);
  sh->o_gyv = take((int64_t)sh->B * sh->wv);
  sh->o_dxa = take(sh->Ma * sh->ld);
  sh->o_dxv = take((int64_t)sh->B * sh->ld);
  for (int i = 0; i < a->action_net.n_layers - 1; ++i) sh->o_dz_a[i] = take(sh->Ma * a->action_net.dims[i + 1]);
  for (int i = 0; i < a->value_net.n_layers - 1; ++i) sh->o_dz_v[i] = take((int64_t)sh->B * a->value_net.dims[i + 1]);
  if (a->has_embedder)
    for (int i = 0; i < a->embedder.n_layers - 1; ++i) sh->o_dz_e[i] = take(sh->N * a->embedder.dims[i + 1]);
  for (int i = 0; i < r.n_src; ++i)
    sh->o_gsrc[i] = a->src_kind[i] == B200GAT_SRC_EXTERNAL ? -1 : take(sh->N * r.width[i]);
  sh->o_gx0 = take(sh->N * a->trunk.in_dim);
  int64_t gemm = 0;
  mlp_ws(&a->action_net, &gemm);
  mlp_ws(&a->value_net, &gemm);
  if (a->has_embedder) mlp_ws(&a->embedder, &gemm);
  sh->gemm_elems = gemm;
  sh->o_gemm = take(gemm);
  sh->trunk_elems = b200gat_trunk_bwd_workspace_elems(&a->trunk);
  B200_CHECK_ARG(sh->trunk_elems > 0, "trunk configuration not covered");
  sh->o_trunk = take(sh->trunk_elems);
  sh->total = off;
  return B200GAT_OK;
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int64_t b200gat_agent_bwd_workspace_elems(const b200gat_agent_step* a) {
  AgentShape sh;
  return agent_shape(a, &sh, true) == B200GAT_OK ? sh.total : 0;
}

extern "C" int b200gat_agent_fwd(const b200gat_agent_step* a, void* stream) {
  AgentShape sh;
  int rc = agent_shape(a, &sh, false);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CHECK_ARG(a->n_pad >= 0 && a->n_pad <= B200GAT_AGENT_MAX_PAD, "n_pad");
  B200_CHECK_ARG(a->node_features && a->logits && a->logp && a->head && a->loss_out && a->g_head && a->action,
                 "null buffer");
  if (a->n_pad) {
    PadJobs P;
    P.n = a->n_pad;
    int64_t mx = 0;
    for (int i = 0; i < a->n_pad; ++i) {
      P.j[i] = a->pad[i];
      B200_CHECK_ARG(P.j[i].src && P.j[i].dst && P.j[i].ld_dst >= P.j[i].cols && P.j[i].rows > 0, "pad job");
      const int64_t t = (int64_t)(P.j[i].rows_dst > P.j[i].rows ? P.j[i].rows_dst : P.j[i].rows) * P.j[i].ld_dst;
      mx = t > mx ? t : mx;
    }
    int gx = (int)ceil_div(mx, 256 * 8);
    gx = gx < 1 ? 1 : (gx > 1184 ? 1184 : gx);
    k_pad_jobs<<<dim3((unsigned)gx, (unsigned)a->n_pad), 256, 0, st>>>(P);
    B200_CHECK_LAUNCH("k_pad_jobs");
  }
  if (a->has_embedder) {
    B200_CHECK_ARG(a->embedder.y == a->trunk.x0, "the embedder writes the trunk's input");
    if ((rc = b200gat_mlp_fwd(&a->embedder, stream))) return rc;
  }
  if ((rc = b200gat_trunk_fwd(&a->trunk, stream))) return rc;
  if ((rc = b200gat_readout_fwd(&a->readout, a->node_features, a->graph_embedding, a->argmax, stream))) return rc;
  if ((rc = b200gat_mlp_fwd(&a->action_net, stream))) return rc;
  if ((rc = b200gat_mlp_fwd(&a->value_net, stream))) return rc;
  float* logprob = a->head;
  float* entropy = a->head + sh.B;
  float* value = a->head + 2 * sh.B;
  k_take_col0<<<(unsigned)ceil_div(sh.Ma, 256), 256, 0, st>>>(a->action_net.y, sh.wa, sh.Ma, a->logits,
                                                              a->value_net.y, sh.wv, sh.B, value);
  B200_CHECK_LAUNCH("k_take_col0");
  if ((rc = b200gat_mask_head_fwd(a->logits, sh.max_n, a->mask, a->ld_mask, sh.B, sh.max_n, a->fill, a->logp, entropy,
                                  nullptr, stream)))
    return rc;
  k_pick_action<<<(unsigned)ceil_div(sh.B, 256), 256, 0, st>>>(a->logp, a->action, sh.B, sh.max_n, logprob);
  B200_CHECK_LAUNCH("k_pick_action");
  b200gat_ppo_loss_args L = a->loss;
  L.logprob = logprob; L.entropy = entropy; L.value = value;
  return b200gat_ppo_loss(&L, a->loss_out, a->loss_out + 1, a->g_head, a->g_head + sh.B, a->g_head + 2 * sh.B, stream);
}

extern "C" int b200gat_agent_bwd(const b200gat_agent_step* a, const b200gat_agent_step_grads* g, void* stream) {
  AgentShape sh;
  int rc = agent_shape(a, &sh, true);
  if (rc) return rc;
  B200_CHECK_ARG(g && g->workspace && (reinterpret_cast<uintptr_t>(g->workspace) & 15) == 0, "workspace");
  B200_CHECK_ARG(g->workspace_elems >= sh.total, "workspace too small (b200gat_agent_bwd_workspace_elems)");
  cudaStream_t st = (cudaStream_t)stream;
  float* ws = g->workspace;
  // ---- d loss / d (logprob, entropy, value) scaled by the upstream gradient
  float* gh = ws + sh.o_ghead;
  k_scale_head<<<(unsigned)ceil_div(3 * sh.B, 256), 256, 0, st>>>(a->g_head, g->g_loss, 3 * sh.B, gh);
  B200_CHECK_LAUNCH("k_scale_head");
  // ---- masked head backward -> gradient of the logits; spread into the heads' padded output columns
  float* g_logits = ws + sh.o_glogits;
  if ((rc = b200gat_mask_head_bwd(a->logp, a->mask, a->ld_mask, a->action, a->head + sh.B, gh, gh + sh.B, sh.B,
                                  sh.max_n, g_logits, stream)))
    return rc;
  float* gya = ws + sh.o_gya;
  float* gyv = ws + sh.o_gyv;
  k_spread_col0<<<(unsigned)ceil_div(sh.Ma, 256), 256, 0, st>>>(g_logits, sh.Ma, sh.wa, gya, gh + 2 * sh.B, sh.B, sh.wv,
                                                                gyv);
  B200_CHECK_LAUNCH("k_spread_col0");
  // ---- heads
  b200gat_mlp_grads ga = g->action_net;
  ga.gy = gya; ga.dx = ws + sh.o_dxa;
  for (int i = 0; i < a->action_net.n_layers - 1; ++i) ga.dz[i] = ws + sh.o_dz_a[i];
  ga.workspace = ws + sh.o_gemm; ga.workspace_elems = sh.gemm_elems;
  // weight gradients on the side stream (trunk.cu "Side stream of the backward pass"): input-gradient chain here,
  // then the dW GEMMs over there, in the order of the calls (they share the split-K workspace)
  AuxStream* aux = aux_stream_for(st);
  auto mlp_bwd = [&](const b200gat_mlp* m, b200gat_mlp_grads* gr) -> int {
    if (!aux) return b200gat_mlp_bwd(m, gr, stream);
    b200gat_mlp_grads chain = *gr;
    for (int i = 0; i < m->n_layers; ++i) { chain.dW[i] = nullptr; chain.db[i] = nullptr; }
    int r = b200gat_mlp_bwd(m, &chain, stream);
    if (r) return r;
    if (!aux_fork(st, aux->s, aux->m1)) {
      set_error("b200gat_agent_bwd: event fork failed: %s", cudaGetErrorString(cudaGetLastError()));
      return B200GAT_ERR_CUDA;
    }
    return mlp_bwd_dw(m, gr, aux->s);
  };
  if ((rc = mlp_bwd(&a->action_net, &ga))) return rc;
  b200gat_mlp_grads gv = g->value_net;
  gv.gy = gyv; gv.dx = ws + sh.o_dxv;
  for (int i = 0; i < a->value_net.n_layers - 1; ++i) gv.dz[i] = ws + sh.o_dz_v[i];
  gv.workspace = ws + sh.o_gemm; gv.workspace_elems = sh.gemm_elems;
  if ((rc = mlp_bwd(&a->value_net, &gv))) return rc;
  // ---- readout backward: every source's gradient
  float* g_src[B200GAT_READOUT_MAX_SRC];
  for (int i = 0; i < B200GAT_READOUT_MAX_SRC; ++i) g_src[i] = nullptr;
  for (int i = 0; i < sh.n_src; ++i)
    g_src[i] = a->src_kind[i] == B200GAT_SRC_EXTERNAL ? g->g_src_ext[i] : ws + sh.o_gsrc[i];
  if ((rc = b200gat_readout_bwd(&a->readout, ws + sh.o_dxa, sh.ld, ws + sh.o_dxv, sh.ld, a->argmax, g_src, sh.N,
                                stream)))
    return rc;
  // ---- trunk backward
  b200gat_trunk_grads tg = g->trunk;
  for (int l = 0; l < a->trunk.n_layers; ++l) tg.layer[l].g_y = nullptr;
  tg.g_x0_in = nullptr;
  alignas(16) static float zero_dummy[4] = {0.f, 0.f, 0.f, 0.f};
  (void)zero_dummy;
  for (int i = 0; i < sh.n_src; ++i) {
    const int k = a->src_kind[i];
    if (k == B200GAT_SRC_X0) tg.g_x0_in = g_src[i];
    else if (k >= B200GAT_SRC_LAYER0) {
      B200_CHECK_ARG(k - B200GAT_SRC_LAYER0 < a->trunk.n_layers, "src_kind names a layer the trunk does not have");
      tg.layer[k - B200GAT_SRC_LAYER0].g_y = g_src[i];
    }
  }
  B200_CHECK_ARG(tg.layer[a->trunk.n_layers - 1].g_y, "the readout must read the last layer's output");
  const bool need_x0 = a->has_embedder || g->g_x0;
  tg.g_x0 = need_x0 ? (a->has_embedder ? ws + sh.o_gx0 : g->g_x0) : nullptr;
  tg.workspace = ws + sh.o_trunk; tg.workspace_elems = sh.trunk_elems;
  if ((rc = b200gat_trunk_bwd(&a->trunk, &tg, stream))) return rc;
  // ---- embedder
  if (a->has_embedder) {
    b200gat_mlp_grads ge = g->embedder;
    ge.gy = ws + sh.o_gx0;
    for (int i = 0; i < a->embedder.n_layers - 1; ++i) ge.dz[i] = ws + sh.o_dz_e[i];
    ge.workspace = ws + sh.o_gemm; ge.workspace_elems = sh.gemm_elems;
    if ((rc = mlp_bwd(&a->embedder, &ge))) return rc;
  }
  // everything queued on the side stream precedes whatever the caller queues next
  if (aux && !aux_join(st, aux->s, aux->a1)) {
    set_error("b200gat_agent_bwd: event join failed: %s", cudaGetErrorString(cudaGetLastError()));
    return B200GAT_ERR_CUDA;
  }
  return B200GAT_OK;
}
