// Generic (any H, any D <= 128) fused graph-attention kernels: one warp per
// (destination, head) for the forward and the destination half of the backward, one warp
// per (source, head) for the source half of the backward.  All segmented reductions run
// in CSR order -> bit-reproducible; only the parameter gradients (attn, edge table) use
// float atomics.  These kernels are the always-available path and the cross-check for
// the specialised kernels in gat_fast.cu.
#include <math.h>
#include "common.cuh"
#include "gat_kernels.cuh"

namespace b200gat {

constexpr int kMaxK = 4;  // D <= 128

struct GenericArgs {
  // graph
  int N, E;
  const int32_t *in_ptr, *in_src, *in_eid, *out_ptr, *out_dst, *out_pos;
  // attention
  int H, D;
  float slope, eps, inv_sqrt_d;
  const float *attn, *attn_bias, *out_bias;
  // edge features
  int emode;
  const float* table;
  const int32_t* etype;
  const float* dense;
  int K;
  const float* attr;
  const float* affine;
};

__device__ __forceinline__ float edge_feat(const GenericArgs& A, int pos, int f) {
  if (A.emode == B200GAT_EDGE_TABLE) return __ldg(A.table + (int64_t)A.etype[pos] * (A.H * A.D) + f);
  if (A.emode == B200GAT_EDGE_AFFINE) {
    float r = __ldg(A.table + (int64_t)A.etype[pos] * (A.H * A.D) + f);
    for (int k = 0; k < A.K; ++k)
      r = fmaf(__ldg(A.attr + (int64_t)pos * A.K + k), __ldg(A.affine + (int64_t)k * (A.H * A.D) + f), r);
    return r;
  }
  if (A.emode == B200GAT_EDGE_DENSE) return __ldg(A.dense + (int64_t)A.in_eid[pos] * (A.H * A.D) + f);
  return 0.f;
}

template <int FLAV>
__global__ void __launch_bounds__(128) k_fwd_generic(GenericArgs A, const float* __restrict__ Pi,
                                                     const float* __restrict__ Pj,
                                                     const float* __restrict__ V,
                                                     float* __restrict__ out, float* __restrict__ s_save,
                                                     float* __restrict__ m_save, float* __restrict__ l_save,
                                                     float* __restrict__ agg_save) {
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (wid >= (int64_t)A.N * A.H) return;
  const int v = (int)(wid / A.H), h = (int)(wid % A.H);
  const int HD = A.H * A.D, hoff = h * A.D;
  const int b = A.in_ptr[v], en = A.in_ptr[v + 1];
  float pj[kMaxK], at[kMaxK];
#pragma unroll
  for (int k = 0; k < kMaxK; ++k) {
    int d = lane + 32 * k;
    bool ok = d < A.D;
    pj[k] = ok ? Pj[(int64_t)v * HD + hoff + d] : 0.f;
    if (FLAV == B200GAT_FLAVOUR_ADDITIVE) at[k] = ok ? A.attn[hoff + d] : 0.f;
    else at[k] = ok ? A.attn[d] : 0.f;
  }
  const float sbias = (FLAV == B200GAT_FLAVOUR_DOT) ? A.attn_bias[0] : 0.f;
  float mx = -INFINITY;
  for (int e = b; e < en; ++e) {
    const int u = A.in_src[e];
    float part = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      int d = lane + 32 * k;
      if (d < A.D) {
        float pi = __ldg(Pi + (int64_t)u * HD + hoff + d);
        if (FLAV == B200GAT_FLAVOUR_ADDITIVE) {
          float z = pi + pj[k] + edge_feat(A, e, hoff + d);
          part += at[k] * leaky(z, A.slope);
        } else {
          part += at[k] * (pj[k] * pi);
        }
      }
    }
    float sc = warp_sum(part);
    if (FLAV == B200GAT_FLAVOUR_DOT) sc = (sc + sbias) * A.inv_sqrt_d;
    if (lane == 0) s_save[(int64_t)e * A.H + h] = sc;
    mx = fmaxf(mx, sc);
  }
  __syncwarp();
  float l = 0.f, acc[kMaxK] = {0.f, 0.f, 0.f, 0.f};
  for (int e = b; e < en; ++e) {
    const int u = A.in_src[e];
    const float p = expf(s_save[(int64_t)e * A.H + h] - mx);
    l += p;
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      int d = lane + 32 * k;
      if (d < A.D) acc[k] = fmaf(p, __ldg(V + (int64_t)u * HD + hoff + d), acc[k]);
    }
  }
  const float inv = en > b ? 1.f / (l + A.eps) : 0.f;
#pragma unroll
  for (int k = 0; k < kMaxK; ++k) {
    int d = lane + 32 * k;
    if (d < A.D) {
      float o = acc[k] * inv;
      if (agg_save) agg_save[(int64_t)v * HD + hoff + d] = o;
      if (A.out_bias) o += A.out_bias[hoff + d];
      out[(int64_t)v * HD + hoff + d] = o;
    }
  }
  if (lane == 0) {
    m_save[(int64_t)v * A.H + h] = en > b ? mx : 0.f;
    l_save[(int64_t)v * A.H + h] = l;
  }
}

template <int FLAV>
__global__ void __launch_bounds__(128) k_bwd_dst_generic(
    GenericArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj,
    const float* __restrict__ V, const float* __restrict__ s_save, const float* __restrict__ m_save,
    const float* __restrict__ l_save, const float* __restrict__ G, float* __restrict__ gPj,
    float* __restrict__ g_attn, float* __restrict__ g_attn_bias, float* __restrict__ g_table,
    float* __restrict__ g_dense, float* __restrict__ g_affine, float* __restrict__ gs) {
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (wid >= (int64_t)A.N * A.H) return;
  const int v = (int)(wid / A.H), h = (int)(wid % A.H);
  const int HD = A.H * A.D, hoff = h * A.D;
  const int b = A.in_ptr[v], en = A.in_ptr[v + 1];
  float pj[kMaxK], at[kMaxK], g[kMaxK];
#pragma unroll
  for (int k = 0; k < kMaxK; ++k) {
    int d = lane + 32 * k;
    bool ok = d < A.D;
    pj[k] = ok ? Pj[(int64_t)v * HD + hoff + d] : 0.f;
    g[k] = ok ? G[(int64_t)v * HD + hoff + d] : 0.f;
    if (FLAV == B200GAT_FLAVOUR_ADDITIVE) at[k] = ok ? A.attn[hoff + d] : 0.f;
    else at[k] = ok ? A.attn[d] : 0.f;
  }
  const float mx = m_save[(int64_t)v * A.H + h];
  const float linv = en > b ? 1.f / (l_save[(int64_t)v * A.H + h] + A.eps) : 0.f;
  // pass A: ga_e = <G[v], V[u]>,  dot = sum_e a_e ga_e
  float dot = 0.f;
  for (int e = b; e < en; ++e) {
    const int u = A.in_src[e];
    float part = 0.f;
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      int d = lane + 32 * k;
      if (d < A.D) part = fmaf(g[k], __ldg(V + (int64_t)u * HD + hoff + d), part);
    }
    const float ga = warp_sum(part);
    const float a = expf(s_save[(int64_t)e * A.H + h] - mx) * linv;
    dot = fmaf(a, ga, dot);
    if (lane == 0) gs[(int64_t)e * A.H + h] = ga;
  }
  __syncwarp();
  // pass B
  float gat[kMaxK] = {0.f, 0.f, 0.f, 0.f}, gpj[kMaxK] = {0.f, 0.f, 0.f, 0.f};
  float gb = 0.f;
  float gtab[kMaxK] = {0.f, 0.f, 0.f, 0.f};
  const bool typed = A.emode == B200GAT_EDGE_TABLE || A.emode == B200GAT_EDGE_AFFINE;
  int t_cur = -1;
  auto flush_tab = [&]() {
    if (t_cur < 0) return;
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      int d = lane + 32 * k;
      if (d < A.D && gtab[k] != 0.f) atomicAdd(g_table + (int64_t)t_cur * HD + hoff + d, gtab[k]);
      gtab[k] = 0.f;
    }
  };
  for (int e = b; e < en; ++e) {
    const int u = A.in_src[e];
    if (typed && A.etype[e] != t_cur) { flush_tab(); t_cur = A.etype[e]; }
    const float ga = gs[(int64_t)e * A.H + h];
    const float a = expf(s_save[(int64_t)e * A.H + h] - mx) * linv;
    const float gsv = a * (ga - dot);
    __syncwarp();
    if (lane == 0) gs[(int64_t)e * A.H + h] = gsv;
    if (FLAV == B200GAT_FLAVOUR_DOT) gb += gsv * A.inv_sqrt_d;
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      int d = lane + 32 * k;
      if (d < A.D) {
        float pi = __ldg(Pi + (int64_t)u * HD + hoff + d);
        if (FLAV == B200GAT_FLAVOUR_ADDITIVE) {
          float z = pi + pj[k] + edge_feat(A, e, hoff + d);
          gat[k] = fmaf(gsv, leaky(z, A.slope), gat[k]);
          float gz = gsv * at[k] * (z > 0.f ? 1.f : A.slope);
          gpj[k] += gz;
          if (A.emode == B200GAT_EDGE_DENSE) g_dense[(int64_t)A.in_eid[e] * HD + hoff + d] = gz;
          else if (A.emode == B200GAT_EDGE_TABLE || A.emode == B200GAT_EDGE_AFFINE) {
            gtab[k] += gz;                     // flushed once per run of equal edge types
            if (A.emode == B200GAT_EDGE_AFFINE)
              for (int q = 0; q < A.K; ++q) {
                float av = __ldg(A.attr + (int64_t)e * A.K + q);
                if (av != 0.f) atomicAdd(g_affine + (int64_t)q * HD + hoff + d, gz * av);
              }
          }
        } else {
          float c = gsv * A.inv_sqrt_d;
          gat[k] = fmaf(c, pj[k] * pi, gat[k]);
          gpj[k] = fmaf(c * at[k], pi, gpj[k]);
        }
      }
    }
  }
  if (typed) flush_tab();
#pragma unroll
  for (int k = 0; k < kMaxK; ++k) {
    int d = lane + 32 * k;
    if (d < A.D) {
      gPj[(int64_t)v * HD + hoff + d] = gpj[k];
      if (en > b) atomicAdd(g_attn + (FLAV == B200GAT_FLAVOUR_ADDITIVE ? hoff + d : d), gat[k]);
    }
  }
  if (FLAV == B200GAT_FLAVOUR_DOT && lane == 0 && en > b) atomicAdd(g_attn_bias, gb);
}

template <int FLAV>
__global__ void __launch_bounds__(128) k_bwd_src_generic(
    GenericArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj,
    const float* __restrict__ s_save, const float* __restrict__ m_save,
    const float* __restrict__ l_save, const float* __restrict__ G, const float* __restrict__ gs,
    float* __restrict__ gPi, float* __restrict__ gV) {
  const int64_t wid = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (wid >= (int64_t)A.N * A.H) return;
  const int u = (int)(wid / A.H), h = (int)(wid % A.H);
  const int HD = A.H * A.D, hoff = h * A.D;
  const int b = A.out_ptr[u], en = A.out_ptr[u + 1];
  float pi[kMaxK], at[kMaxK];
#pragma unroll
  for (int k = 0; k < kMaxK; ++k) {
    int d = lane + 32 * k;
    bool ok = d < A.D;
    pi[k] = ok ? Pi[(int64_t)u * HD + hoff + d] : 0.f;
    if (FLAV == B200GAT_FLAVOUR_ADDITIVE) at[k] = ok ? A.attn[hoff + d] : 0.f;
    else at[k] = ok ? A.attn[d] : 0.f;
  }
  float gpi[kMaxK] = {0.f, 0.f, 0.f, 0.f}, gv[kMaxK] = {0.f, 0.f, 0.f, 0.f};
  for (int q = b; q < en; ++q) {
    const int pos = A.out_pos[q], v = A.out_dst[q];
    const float gsv = gs[(int64_t)pos * A.H + h];
    const float a = expf(s_save[(int64_t)pos * A.H + h] - m_save[(int64_t)v * A.H + h]) /
                    (l_save[(int64_t)v * A.H + h] + A.eps);
#pragma unroll
    for (int k = 0; k < kMaxK; ++k) {
      int d = lane + 32 * k;
      if (d < A.D) {
        float pjv = __ldg(Pj + (int64_t)v * HD + hoff + d);
        if (FLAV == B200GAT_FLAVOUR_ADDITIVE) {
          float z = pi[k] + pjv + edge_feat(A, pos, hoff + d);
          gpi[k] += gsv * at[k] * (z > 0.f ? 1.f : A.slope);
        } else {
          gpi[k] = fmaf(gsv * A.inv_sqrt_d * at[k], pjv, gpi[k]);
        }
        gv[k] = fmaf(a, __ldg(G + (int64_t)v * HD + hoff + d), gv[k]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kMaxK; ++k) {
    int d = lane + 32 * k;
    if (d < A.D) {
      int64_t o = (int64_t)u * HD + hoff + d;
      if (gV == gPi) gPi[o] = gpi[k] + gv[k];
      else { gPi[o] = gpi[k]; gV[o] = gv[k]; }
    }
  }
}

// Clique members (cl != NULL, node_blk[v] >= 0) keep the logits of their block in-edges
// [run, run + n_b - 1) in s_dense[blk_soff[b]*H + h*n_b*npad + rank(v)*npad + rank(u)].
__global__ void k_attention(int N, int H, float eps, const int32_t* __restrict__ in_ptr,
                            const int32_t* __restrict__ in_eid, const float* __restrict__ s_save,
                            const float* __restrict__ m_save, const float* __restrict__ l_save,
                            const float* __restrict__ s_dense, const int32_t* __restrict__ node_blk,
                            const int32_t* __restrict__ node_run, const int32_t* __restrict__ node_rank,
                            const int32_t* __restrict__ blk_ptr, const int32_t* __restrict__ blk_soff,
                            float* __restrict__ alpha) {
  const int64_t v = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (v >= N) return;
  const int b = in_ptr[v], en = in_ptr[v + 1];
  int run = 0, nb = 0, rank = 0, npad = 0;
  int64_t soff = 0;
  if (s_dense && node_blk && node_blk[v] >= 0) {
    const int blk = node_blk[v];
    nb = blk_ptr[blk + 1] - blk_ptr[blk];
    npad = (nb + 3) & ~3;
    run = node_run[v];
    rank = node_rank[v];
    soff = (int64_t)blk_soff[blk] * H + (int64_t)rank * npad;
  }
  const int64_t n = (int64_t)(en - b) * H;
  for (int64_t i = lane; i < n; i += 32) {
    int e = b + (int)(i / H), h = (int)(i % H);
    float s;
    if (nb > 0 && e >= run && e < run + nb - 1) {
      const int k = e - run, ur = k + (k >= rank ? 1 : 0);
      s = s_dense[soff + (int64_t)h * nb * npad + ur];
    } else {
      s = s_save[(int64_t)e * H + h];
    }
    alpha[(int64_t)in_eid[e] * H + h] = expf(s - m_save[v * H + h]) / (l_save[v * H + h] + eps);
  }
}

__global__ void k_edge_out(GenericArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj,
                           float* __restrict__ f_out) {
  const int64_t v = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (v >= A.N) return;
  const int HD = A.H * A.D;
  for (int e = A.in_ptr[v]; e < A.in_ptr[v + 1]; ++e) {
    const int u = A.in_src[e];
    const int64_t eid = A.in_eid[e];
    for (int f = lane; f < HD; f += 32)
      f_out[eid * HD + f] = leaky(Pi[(int64_t)u * HD + f] + Pj[v * HD + f] + edge_feat(A, e, f), A.slope);
  }
}

static int fill_args(GenericArgs& A, const b200gat_graph* g, const b200gat_attn* a,
                     const b200gat_edge_feats* ef) {
  B200_CHECK_ARG(g && a, "null graph/attn descriptor");
  B200_CHECK_ARG(g->n_nodes >= 0 && g->n_edges >= 0, "negative graph size");
  B200_CHECK_ARG(a->heads > 0 && a->dim > 0, "heads/dim must be positive");
  B200_CHECK_ARG(a->flavour == B200GAT_FLAVOUR_ADDITIVE || a->flavour == B200GAT_FLAVOUR_DOT,
                 "unknown flavour");
  B200_CHECK_ARG(g->n_nodes == 0 || (g->in_ptr && g->out_ptr), "null CSR pointers");
  B200_CHECK_ARG(g->n_edges == 0 || (g->in_src && g->in_eid && g->out_dst && g->out_pos),
                 "null CSR edge arrays");
  B200_CHECK_ARG(a->attn, "null attn");
  B200_CHECK_ARG(a->flavour != B200GAT_FLAVOUR_DOT || a->attn_bias, "dot flavour needs attn_bias");
  A.N = g->n_nodes; A.E = g->n_edges;
  A.in_ptr = g->in_ptr; A.in_src = g->in_src; A.in_eid = g->in_eid;
  A.out_ptr = g->out_ptr; A.out_dst = g->out_dst; A.out_pos = g->out_pos;
  A.H = a->heads; A.D = a->dim; A.slope = a->slope; A.eps = a->softmax_eps;
  A.inv_sqrt_d = 1.0f / sqrtf((float)a->dim);
  A.attn = a->attn; A.attn_bias = a->attn_bias; A.out_bias = a->out_bias;
  A.emode = ef ? ef->mode : B200GAT_EDGE_NONE;
  A.table = ef ? ef->table : nullptr; A.etype = ef ? ef->etype : nullptr;
  A.dense = ef ? ef->dense : nullptr;
  A.K = ef ? ef->n_attr : 0; A.attr = ef ? ef->attr : nullptr; A.affine = ef ? ef->affine : nullptr;
  if (a->flavour == B200GAT_FLAVOUR_DOT) A.emode = B200GAT_EDGE_NONE;  // GATv3 ignores edge_attr
  if (A.emode == B200GAT_EDGE_TABLE || A.emode == B200GAT_EDGE_AFFINE)
    B200_CHECK_ARG(A.table && (A.E == 0 || A.etype), "null edge table/etype");
  if (A.emode == B200GAT_EDGE_AFFINE)
    B200_CHECK_ARG(A.K >= 0 && A.K <= B200GAT_MAX_ATTR && (A.K == 0 || (A.affine && (A.E == 0 || A.attr))),
                   "bad affine edge attributes");
  if (A.emode == B200GAT_EDGE_DENSE) B200_CHECK_ARG(A.E == 0 || A.dense, "null dense edge features");
  return B200GAT_OK;
}

int generic_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                const float* Pi, const float* Pj, const float* V, float* out, float* s_save,
                float* m_save, float* l_save, float* agg_save, cudaStream_t st) {
  GenericArgs A;
  int rc = fill_args(A, g, a, ef);
  if (rc) return rc;
  if (A.D > 32 * kMaxK) {
    set_error("generic kernels support dim <= %d (got %d)", 32 * kMaxK, A.D);
    return B200GAT_ERR_UNSUPPORTED;
  }
  if (A.N == 0) return B200GAT_OK;
  unsigned blocks = (unsigned)ceil_div((int64_t)A.N * A.H * 32, 128);
  ScopedTimer tm(B200GAT_TIMING_CONV_FWD, st);
  if (a->flavour == B200GAT_FLAVOUR_ADDITIVE)
    k_fwd_generic<B200GAT_FLAVOUR_ADDITIVE><<<blocks, 128, 0, st>>>(A, Pi, Pj, V, out, s_save, m_save, l_save, agg_save);
  else
    k_fwd_generic<B200GAT_FLAVOUR_DOT><<<blocks, 128, 0, st>>>(A, Pi, Pj, V, out, s_save, m_save, l_save, agg_save);
  B200_CHECK_LAUNCH("k_fwd_generic");
  return B200GAT_OK;
}

int generic_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef,
                const float* Pi, const float* Pj, const float* V, const float* s_save,
                const float* m_save, const float* l_save, const float* g_out, float* gPi, float* gPj,
                float* gV, float* g_attn, float* g_attn_bias, float* g_table, float* g_dense,
                float* g_affine, float* gs, cudaStream_t st) {
  GenericArgs A;
  int rc = fill_args(A, g, a, ef);
  if (rc) return rc;
  if (A.D > 32 * kMaxK) {
    set_error("generic kernels support dim <= %d (got %d)", 32 * kMaxK, A.D);
    return B200GAT_ERR_UNSUPPORTED;
  }
  if (A.emode == B200GAT_EDGE_TABLE || A.emode == B200GAT_EDGE_AFFINE) B200_CHECK_ARG(g_table, "null g_table");
  if (A.emode == B200GAT_EDGE_AFFINE && A.K > 0) B200_CHECK_ARG(g_affine, "null g_affine");
  if (A.emode == B200GAT_EDGE_DENSE) B200_CHECK_ARG(A.E == 0 || g_dense, "null g_dense");
  if (A.N == 0) return B200GAT_OK;
  unsigned blocks = (unsigned)ceil_div((int64_t)A.N * A.H * 32, 128);
  {
    ScopedTimer tm(B200GAT_TIMING_CONV_BWD_DST, st);
    if (a->flavour == B200GAT_FLAVOUR_ADDITIVE)
      k_bwd_dst_generic<B200GAT_FLAVOUR_ADDITIVE><<<blocks, 128, 0, st>>>(
          A, Pi, Pj, V, s_save, m_save, l_save, g_out, gPj, g_attn, g_attn_bias, g_table, g_dense, g_affine, gs);
    else
      k_bwd_dst_generic<B200GAT_FLAVOUR_DOT><<<blocks, 128, 0, st>>>(
          A, Pi, Pj, V, s_save, m_save, l_save, g_out, gPj, g_attn, g_attn_bias, g_table, g_dense, g_affine, gs);
  }
  {
    ScopedTimer tm(B200GAT_TIMING_CONV_BWD_SRC, st);
    if (a->flavour == B200GAT_FLAVOUR_ADDITIVE)
      k_bwd_src_generic<B200GAT_FLAVOUR_ADDITIVE><<<blocks, 128, 0, st>>>(A, Pi, Pj, s_save, m_save, l_save,
                                                                        g_out, gs, gPi, gV);
    else
      k_bwd_src_generic<B200GAT_FLAVOUR_DOT><<<blocks, 128, 0, st>>>(A, Pi, Pj, s_save, m_save, l_save,
                                                                   g_out, gs, gPi, gV);
  }
  B200_CHECK_LAUNCHES("k_bwd_generic", 2);
  return B200GAT_OK;
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int b200gat_attention(const b200gat_graph* g, int32_t heads, float eps, const float* s_save,
                                 const float* m_save, const float* l_save, const float* s_dense,
                                 float* alpha, void* stream) {
  B200_CHECK_ARG(g && heads > 0, "bad arguments");
  if (g->n_nodes == 0 || g->n_edges == 0) return B200GAT_OK;
  B200_CHECK_ARG(s_save && m_save && l_save && alpha && g->in_eid, "null pointer");
  const b200gat_cliques* c = (s_dense && g->cliques && g->cliques->n_blocks > 0) ? g->cliques : nullptr;
  k_attention<<<(unsigned)ceil_div((int64_t)g->n_nodes * 32, 256), 256, 0, (cudaStream_t)stream>>>(
      g->n_nodes, heads, eps, g->in_ptr, g->in_eid, s_save, m_save, l_save, c ? s_dense : nullptr,
      c ? c->node_blk : nullptr, c ? c->node_run : nullptr, c ? c->node_rank : nullptr,
      c ? c->blk_ptr : nullptr, c ? c->blk_soff : nullptr, alpha);
  B200_CHECK_LAUNCH("k_attention");
  return B200GAT_OK;
}

extern "C" int b200gat_edge_out(const b200gat_graph* g, const b200gat_attn* a,
                                const b200gat_edge_feats* ef, const float* Pi, const float* Pj,
                                float* f_out, void* stream) {
  GenericArgs A;
  int rc = fill_args(A, g, a, ef);
  if (rc) return rc;
  B200_CHECK_ARG(a->flavour == B200GAT_FLAVOUR_ADDITIVE, "edge_out is defined for the additive flavour");
  if (A.N == 0 || A.E == 0) return B200GAT_OK;
  B200_CHECK_ARG(Pi && Pj && f_out, "null pointer");
  k_edge_out<<<(unsigned)ceil_div((int64_t)A.N * 32, 256), 256, 0, (cudaStream_t)stream>>>(A, Pi, Pj, f_out);
  B200_CHECK_LAUNCH("k_edge_out");
  return B200GAT_OK;
}
