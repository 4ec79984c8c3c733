// Dense clique-block kernels for the additive flavour (EGAT / GATv2), D = 64 features per head.
//
// A block (b200gat_cliques) is a complete digraph over n <= 128 nodes with one edge type, so for a head h
//   s[u->v] = sum_d a_d * leaky(Pi[u,d] + Pj'[v,d]),   Pj' = Pj + table[type]          (bias folded in table)
// is needed for ALL ordered pairs u != v.  With leaky(z) = slope*z + (1-slope)*relu(z) and
// relu(x + y) = max(x, -y) + y the per-pair, per-feature work is ONE max and ONE fma:
//   s = (1-slope) * sum_d a_d * max(Pi[u,d], -Pj'[v,d])  +  slope * <a,Pi[u]>  +  <a,Pj'[v]>
// (the two dot products are per-node scalars computed while staging).
//
// Forward, one work item = (block, head):
//   stage   : Pi_h[S], -Pj'_h[S], V_h[S] of the block's members into shared memory (rows padded so that
//             the access patterns below are bank-conflict free), per-node scalars, attn_h
//   logits  : warp = 16 destinations (an mma M tile); lane (g = lane/4, t = lane%4) owns destinations
//             {g, g+8} x sources {8k+t, 8k+t+4} - exactly the A-fragment layout of mma.m16n8k8 - and loops
//             over the 64 features with 16-byte shared loads; no shuffles, no index traffic
//   softmax : online (running max / rescale), continuing the partial state the sparse kernel left for
//             the node's non-clique in-edges (self loop, job-chain precedence, pool edges)
//   P.V     : numerators stay in registers as A fragments; out[v] += P[v,u] V[u] on the tensor cores,
//             3xTF32 split operands (a_hi*b_hi + a_hi*b_lo + a_lo*b_hi) for fp32-level accuracy
//   final   : normalise, write out / agg_save / m_save / l_save; raw logits go to s_dense[h][v][u]
#include <atomic>
#include <math.h>
#include <stdlib.h>
#include "common.cuh"
#include "gat_kernels.cuh"

namespace b200gat {

constexpr int kD = 64;        // features per head
constexpr int kSP = 72;       // row stride of the Pi / V tiles   (== 8 mod 32: B-fragment loads conflict free)
constexpr int kSQ = 68;       // row stride of the -Pj' tile      (== 4 mod 32: 8 distinct rows x 16 B per load)

struct CliqueArgs {
  int H, HD;
  int n_items;   // n_blocks * H
  int tiles;     // 16-row destination tiles per item slot = ceil(max_block / 16)
  int ipc;       // item slots per CTA
  int ru, rv;    // padded rows: sources (multiple of 8), destinations (tiles * 16)
  int item_floats;
  const int32_t *blk_ptr, *blk_nodes, *blk_type, *blk_soff;
  float slope, eps;
  const float *attn, *out_bias, *table;
};

__host__ __device__ inline int clique_item_floats(int ru, int rv, bool alias) {
  return ru * kSP * (alias ? 1 : 2) + rv * kSQ + ru + rv + kD;
}

// 3xTF32 operand split.  The tensor core reads only the upper 19 bits of an fp32 operand, so hi = x itself
// (the hardware truncates it to x & 0xffffe000) and lo = x - trunc(x) (exact in fp32, |lo| < 2^-10 |x|, itself
// truncated by the hardware with a relative error < 2^-10): a_hi*b_hi + a_hi*b_lo + a_lo*b_hi reproduces the
// fp32 product to ~2^-20 relative.  Two instructions per operand (LOP3 + FADD) instead of two cvt.rna.
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = __float_as_uint(x);
  lo = __float_as_uint(x - __uint_as_float(hi & 0xffffe000u));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Packed fp32x2 helpers (sm_100 FFMA2).
__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ float lo2(unsigned long long v) { return __uint_as_float((unsigned)(v & 0xffffffffull)); }
__device__ __forceinline__ float hi2(unsigned long long v) { return __uint_as_float((unsigned)(v >> 32)); }
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}

// Shared-memory views of one item slot.
struct ItemSmem {
  float *pi, *v, *npj, *ai, *aj, *at;
};
template <bool ALIAS>
__device__ __forceinline__ ItemSmem item_views(float* base, int ru, int rv) {
  ItemSmem s;
  s.pi = base;
  s.v = ALIAS ? base : base + ru * kSP;
  s.npj = s.v + ru * kSP;
  s.ai = s.npj + rv * kSQ;
  s.aj = s.ai + ru;
  s.at = s.aj + rv;
  return s;
}

// Stage the rows of one (block, head) item.  Called by the nthr = tiles*32 threads of the slot with the
// same trip count for every thread (rv*16 float4 slots, rv = tiles*16).
template <bool ALIAS, bool WITH_V>
__device__ __forceinline__ void stage_item(const CliqueArgs& A, const ItemSmem& S, int tid, int nthr, int n,
                                           const int32_t* __restrict__ nodes, int type, int h,
                                           const float* __restrict__ Pi, const float* __restrict__ Pj,
                                           const float* __restrict__ V) {
  const int hoff = h * kD;
  const int c4 = (tid & 15) * 4;
  const float4 a4 = __ldg(reinterpret_cast<const float4*>(A.attn + hoff + c4));
  float4 tb = make_float4(0.f, 0.f, 0.f, 0.f);
  if (A.table) tb = __ldg(reinterpret_cast<const float4*>(A.table + (int64_t)type * A.HD + hoff + c4));
  // rv * 16 == 8 * nthr float4 slots: two batches of four, all global loads of a batch in flight together
#pragma unroll 1
  for (int half = 0; half < 2; ++half) {
    float4 pi[4], pj[4], vv[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int row = (tid + (half * 4 + k) * nthr) >> 4;
      pi[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      pj[k] = pi[k];
      vv[k] = pi[k];
      if (row < n) {
        const int64_t node = nodes[row];
        pi[k] = __ldg(reinterpret_cast<const float4*>(Pi + node * A.HD + hoff + c4));
        pj[k] = __ldg(reinterpret_cast<const float4*>(Pj + node * A.HD + hoff + c4));
        if (WITH_V && !ALIAS) vv[k] = __ldg(reinterpret_cast<const float4*>(V + node * A.HD + hoff + c4));
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int row = (tid + (half * 4 + k) * nthr) >> 4;
      if (row < n) { pj[k].x += tb.x; pj[k].y += tb.y; pj[k].z += tb.z; pj[k].w += tb.w; }
      if (row < A.ru) {
        *reinterpret_cast<float4*>(S.pi + row * kSP + c4) = pi[k];
        if (WITH_V && !ALIAS) *reinterpret_cast<float4*>(S.v + row * kSP + c4) = vv[k];
      }
      *reinterpret_cast<float4*>(S.npj + row * kSQ + c4) = make_float4(-pj[k].x, -pj[k].y, -pj[k].z, -pj[k].w);
      float di = a4.x * pi[k].x + a4.y * pi[k].y + a4.z * pi[k].z + a4.w * pi[k].w;
      float dj = a4.x * pj[k].x + a4.y * pj[k].y + a4.z * pj[k].z + a4.w * pj[k].w;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        di += __shfl_xor_sync(0xffffffffu, di, o);
        dj += __shfl_xor_sync(0xffffffffu, dj, o);
      }
      if ((tid & 15) == 0) {
        if (row < A.ru) S.ai[row] = A.slope * di;
        S.aj[row] = dj;
      }
    }
  }
  if (tid < 16) *reinterpret_cast<float4*>(S.at + tid * 4) = __ldg(reinterpret_cast<const float4*>(A.attn + hoff + tid * 4));
}

// sum_d a_d * max(Pi[u,d], -Pj'[v,d]) for destinations {r0, r1} x the 2*KB sources ub + 4*iu + t.
template <int KB>
__device__ __forceinline__ void relu_dots(const ItemSmem& S, int r0, int r1, int ub, int t, float (&s0)[8],
                                          float (&s1)[8]) {
  const float4* q0p = reinterpret_cast<const float4*>(S.npj + r0 * kSQ);
  const float4* q1p = reinterpret_cast<const float4*>(S.npj + r1 * kSQ);
  const float4* atp = reinterpret_cast<const float4*>(S.at);
  const float* pp = S.pi + (ub + t) * kSP;
#pragma unroll
  for (int i = 0; i < 2 * KB; ++i) { s0[i] = 0.f; s1[i] = 0.f; }
#pragma unroll 4
  for (int d4 = 0; d4 < kD / 4; ++d4) {
    const float4 a = atp[d4], q0 = q0p[d4], q1 = q1p[d4];
#pragma unroll
    for (int i = 0; i < 2 * KB; ++i) {
      const float4 p = *reinterpret_cast<const float4*>(pp + i * 4 * kSP + d4 * 4);
      s0[i] = fmaf(a.x, fmaxf(p.x, q0.x), s0[i]);
      s0[i] = fmaf(a.y, fmaxf(p.y, q0.y), s0[i]);
      s0[i] = fmaf(a.z, fmaxf(p.z, q0.z), s0[i]);
      s0[i] = fmaf(a.w, fmaxf(p.w, q0.w), s0[i]);
      s1[i] = fmaf(a.x, fmaxf(p.x, q1.x), s1[i]);
      s1[i] = fmaf(a.y, fmaxf(p.y, q1.y), s1[i]);
      s1[i] = fmaf(a.z, fmaxf(p.z, q1.z), s1[i]);
      s1[i] = fmaf(a.w, fmaxf(p.w, q1.w), s1[i]);
    }
  }
}

struct FwdState {
  float m0, m1, l0, l1;
  float acc[8][4];
};

template <int KB>
__device__ __forceinline__ void fwd_chunk(const CliqueArgs& A, const ItemSmem& S, FwdState& st, int n, int r0,
                                          int r1, int ub, int g, int t, float aj0, float aj1,
                                          float* __restrict__ sd, int npad) {
  float s0[8], s1[8];
  relu_dots<KB>(S, r0, r1, ub, t, s0, s1);
  const float c1 = 1.f - A.slope;
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int i = 0; i < 2 * KB; ++i) {
    const int u = ub + 4 * i + t;
    const float ai = S.ai[u];
    float x0 = fmaf(c1, s0[i], ai + aj0), x1 = fmaf(c1, s1[i], ai + aj1);
    if (u >= n || u == r0) x0 = -INFINITY;
    if (u >= n || u == r1) x1 = -INFINITY;
    if (u < n) {
      if (r0 < n) sd[r0 * npad + u] = x0;
      if (r1 < n) sd[r1 * npad + u] = x1;
    }
    s0[i] = x0; s1[i] = x1;
    mx0 = fmaxf(mx0, x0); mx1 = fmaxf(mx1, x1);
  }
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
  const float mn0 = fmaxf(st.m0, mx0), mn1 = fmaxf(st.m1, mx1);
  const float ms0 = mn0 == -INFINITY ? 0.f : mn0, ms1 = mn1 == -INFINITY ? 0.f : mn1;
  const float sc0 = __expf(st.m0 - ms0), sc1 = __expf(st.m1 - ms1);     // exp(-inf) = 0 on the first chunk
  st.m0 = mn0; st.m1 = mn1;
  float ps0 = 0.f, ps1 = 0.f;
#pragma unroll
  for (int i = 0; i < 2 * KB; ++i) {
    s0[i] = __expf(s0[i] - ms0); s1[i] = __expf(s1[i] - ms1);
    ps0 += s0[i]; ps1 += s1[i];
  }
  st.l0 = fmaf(st.l0, sc0, ps0);
  st.l1 = fmaf(st.l1, sc1, ps1);
#pragma unroll
  for (int nb = 0; nb < 8; ++nb) {
    st.acc[nb][0] *= sc0; st.acc[nb][1] *= sc0; st.acc[nb][2] *= sc1; st.acc[nb][3] *= sc1;
  }
#pragma unroll
  for (int k = 0; k < KB; ++k) {
    uint32_t ah[4], al[4];
    split_tf32(s0[2 * k], ah[0], al[0]);
    split_tf32(s1[2 * k], ah[1], al[1]);
    split_tf32(s0[2 * k + 1], ah[2], al[2]);
    split_tf32(s1[2 * k + 1], ah[3], al[3]);
    const float* vb = S.v + (ub + 8 * k + t) * kSP + g;
    // four column blocks at a time; the three split products are issued product-major so that consecutive
    // mma's target different accumulators (no back-to-back dependent tensor instructions)
#pragma unroll
    for (int nq = 0; nq < 8; nq += 4) {
      uint32_t bh0[4], bl0[4], bh1[4], bl1[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        split_tf32(vb[(nq + j) * 8], bh0[j], bl0[j]);
        split_tf32(vb[4 * kSP + (nq + j) * 8], bh1[j], bl1[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_tf32(st.acc[nq + j], al[0], al[1], al[2], al[3], bh0[j], bh1[j]);
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_tf32(st.acc[nq + j], ah[0], ah[1], ah[2], ah[3], bl0[j], bl1[j]);
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_tf32(st.acc[nq + j], ah[0], ah[1], ah[2], ah[3], bh0[j], bh1[j]);
    }
  }
}

template <bool ALIAS>
__global__ void __launch_bounds__(256, 2) k_clq_fwd(CliqueArgs A, const float* __restrict__ Pi,
                                                     const float* __restrict__ Pj, const float* __restrict__ V,
                                                     float* __restrict__ out, float* __restrict__ s_dense,
                                                     float* __restrict__ m_save, float* __restrict__ l_save,
                                                     float* __restrict__ agg_save) {
  pdl_sync();
  extern __shared__ __align__(16) float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = warp / A.tiles, vt = warp - slot * A.tiles;
  const int item = blockIdx.x * A.ipc + slot;
  const bool item_ok = item < A.n_items;
  const ItemSmem S = item_views<ALIAS>(smem + (size_t)slot * A.item_floats, A.ru, A.rv);
  int b = 0, h = 0, n = 0;
  const int32_t* nodes = nullptr;
  if (item_ok) {
    b = item / A.H;
    h = item - b * A.H;
    n = A.blk_ptr[b + 1] - A.blk_ptr[b];
    nodes = A.blk_nodes + A.blk_ptr[b];
    stage_item<ALIAS, true>(A, S, threadIdx.x - slot * A.tiles * 32, A.tiles * 32, n, nodes, A.blk_type[b], h, Pi,
                            Pj, V);
  }
  __syncthreads();
  if (!item_ok || vt * 16 >= n) return;

  const int g = lane >> 2, t = lane & 3;
  const int r0 = vt * 16 + g, r1 = r0 + 8;
  const bool ok0 = r0 < n, ok1 = r1 < n;
  const int64_t v0 = ok0 ? nodes[r0] : 0, v1 = ok1 ? nodes[r1] : 0;
  const int hoff = h * kD;
  FwdState st;
  st.m0 = ok0 ? m_save[v0 * A.H + h] : -INFINITY;
  st.m1 = ok1 ? m_save[v1 * A.H + h] : -INFINITY;
  st.l0 = (ok0 && t == 0) ? l_save[v0 * A.H + h] : 0.f;
  st.l1 = (ok1 && t == 0) ? l_save[v1 * A.H + h] : 0.f;
#pragma unroll
  for (int nb = 0; nb < 8; ++nb) {
    float2 c0 = make_float2(0.f, 0.f), c1 = c0;
    if (ok0) c0 = *reinterpret_cast<const float2*>(out + v0 * A.HD + hoff + nb * 8 + 2 * t);
    if (ok1) c1 = *reinterpret_cast<const float2*>(out + v1 * A.HD + hoff + nb * 8 + 2 * t);
    st.acc[nb][0] = c0.x; st.acc[nb][1] = c0.y; st.acc[nb][2] = c1.x; st.acc[nb][3] = c1.y;
  }
  const float aj0 = S.aj[r0], aj1 = S.aj[r1];
  const int npad = (n + 3) & ~3;
  float* sd = s_dense + (int64_t)A.blk_soff[b] * A.H + (int64_t)h * n * npad;
  const int nkb = (n + 7) >> 3;
  int kb = 0;
  for (; kb + 4 <= nkb; kb += 4) fwd_chunk<4>(A, S, st, n, r0, r1, kb * 8, g, t, aj0, aj1, sd, npad);
  switch (nkb - kb) {
    case 3: fwd_chunk<3>(A, S, st, n, r0, r1, kb * 8, g, t, aj0, aj1, sd, npad); break;
    case 2: fwd_chunk<2>(A, S, st, n, r0, r1, kb * 8, g, t, aj0, aj1, sd, npad); break;
    case 1: fwd_chunk<1>(A, S, st, n, r0, r1, kb * 8, g, t, aj0, aj1, sd, npad); break;
    default: break;
  }
  // row sums live in four partials (one per t lane)
  st.l0 += __shfl_xor_sync(0xffffffffu, st.l0, 1);
  st.l0 += __shfl_xor_sync(0xffffffffu, st.l0, 2);
  st.l1 += __shfl_xor_sync(0xffffffffu, st.l1, 1);
  st.l1 += __shfl_xor_sync(0xffffffffu, st.l1, 2);
  const float inv0 = 1.f / (st.l0 + A.eps), inv1 = 1.f / (st.l1 + A.eps);
#pragma unroll
  for (int nb = 0; nb < 8; ++nb) {
    const int col = hoff + nb * 8 + 2 * t;
    float2 ob = make_float2(0.f, 0.f);
    if (A.out_bias) ob = __ldg(reinterpret_cast<const float2*>(A.out_bias + col));
    if (ok0) {
      const float2 a = make_float2(st.acc[nb][0] * inv0, st.acc[nb][1] * inv0);
      if (agg_save) *reinterpret_cast<float2*>(agg_save + v0 * A.HD + col) = a;
      *reinterpret_cast<float2*>(out + v0 * A.HD + col) = make_float2(a.x + ob.x, a.y + ob.y);
    }
    if (ok1) {
      const float2 a = make_float2(st.acc[nb][2] * inv1, st.acc[nb][3] * inv1);
      if (agg_save) *reinterpret_cast<float2*>(agg_save + v1 * A.HD + col) = a;
      *reinterpret_cast<float2*>(out + v1 * A.HD + col) = make_float2(a.x + ob.x, a.y + ob.y);
    }
  }
  if (t == 0) {
    if (ok0) { m_save[v0 * A.H + h] = st.m0; l_save[v0 * A.H + h] = st.l0; }
    if (ok1) { m_save[v1 * A.H + h] = st.m1; l_save[v1 * A.H + h] = st.l1; }
  }
}


// ------------------------------------------------------------------------------------------ backward, part 1
// Tensor-core part of the block backward, one work item = (block, head), warp = 16 sources:
//   ga^T[u,v] = <V[u], G[v]>                      mma #1 (M = u, N = v, K = 64 features)
//   alpha     = exp(s - m_v) / (l_v + eps)         from the saved logits
//   gs        = alpha * (ga - <G[v], agg[v]>)      -> gs_dense[h][v][u], row sums gsum_u, column sums gsum_v
//   gV[u]    += sum_v alpha[u,v] * G[v]            mma #2 (M = u, K = v, N = 64): the C fragment of mma #1 is
//                                                  re-used as the A fragment with the K index permuted
//                                                  (k-index t <-> v = 2t, t+4 <-> v = 2t+1) and the rows of G
//                                                  loaded in the same permuted order
constexpr int kSB = 68;   // row stride of the V / G tiles here (== 4 mod 32: A-fragment row loads conflict free)

struct BwdSmem {
  float *v, *g, *m, *linv, *dot, *gsv;
};
__host__ __device__ inline int clique_bwd1_floats(int rv) { return 2 * rv * kSB + 4 * rv; }
__device__ __forceinline__ BwdSmem bwd_views(float* base, int rv) {
  BwdSmem s;
  s.v = base;
  s.g = base + rv * kSB;
  s.m = s.g + rv * kSB;
  s.linv = s.m + rv;
  s.dot = s.linv + rv;
  s.gsv = s.dot + rv;
  return s;
}

template <int VB>
__device__ __forceinline__ void bwd1_chunk(const CliqueArgs& A, const BwdSmem& S, float (&acc)[8][4], float& gsu0,
                                           float& gsu1, int n, int u0, int u1, int vbase, int g, int t,
                                           const float* __restrict__ sd, float* __restrict__ gsd, int npad) {
  float c[VB][4], sv[VB][4];
#pragma unroll
  for (int j = 0; j < VB; ++j) {          // saved logits first: their latency hides behind mma #1
    c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int u = (q & 2) ? u1 : u0, v = vbase + 8 * j + 2 * t + (q & 1);
      sv[j][q] = (u < n && v < n) ? __ldg(sd + v * npad + u) : 0.f;
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    uint32_t ah[4], al[4];
    split_tf32(S.v[u0 * kSB + k * 8 + t], ah[0], al[0]);
    split_tf32(S.v[u1 * kSB + k * 8 + t], ah[1], al[1]);
    split_tf32(S.v[u0 * kSB + k * 8 + t + 4], ah[2], al[2]);
    split_tf32(S.v[u1 * kSB + k * 8 + t + 4], ah[3], al[3]);
    uint32_t bh0[VB], bl0[VB], bh1[VB], bl1[VB];
#pragma unroll
    for (int j = 0; j < VB; ++j) {
      const float* gp = S.g + (vbase + 8 * j + g) * kSB + k * 8 + t;
      split_tf32(gp[0], bh0[j], bl0[j]);
      split_tf32(gp[4], bh1[j], bl1[j]);
    }
#pragma unroll
    for (int j = 0; j < VB; ++j) mma_tf32(c[j], al[0], al[1], al[2], al[3], bh0[j], bh1[j]);
#pragma unroll
    for (int j = 0; j < VB; ++j) mma_tf32(c[j], ah[0], ah[1], ah[2], ah[3], bl0[j], bl1[j]);
#pragma unroll
    for (int j = 0; j < VB; ++j) mma_tf32(c[j], ah[0], ah[1], ah[2], ah[3], bh0[j], bh1[j]);
  }
  // alpha / gs on the C fragments: c[j][0]=(u0,va) [1]=(u0,vb) [2]=(u1,va) [3]=(u1,vb)
#pragma unroll
  for (int j = 0; j < VB; ++j) {
    const int va = vbase + 8 * j + 2 * t, vb = va + 1;
    const float ma = S.m[va], mb = S.m[vb], ia = S.linv[va], ib = S.linv[vb], da = S.dot[va], db = S.dot[vb];
    float al4[4], gs4[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int u = (q & 2) ? u1 : u0, v = (q & 1) ? vb : va;
      const bool ok = u < n && v < n && u != v;
      float a = 0.f, gsv = 0.f;
      if (ok) {
        a = __expf(sv[j][q] - ((q & 1) ? mb : ma)) * ((q & 1) ? ib : ia);
        gsv = a * (c[j][q] - ((q & 1) ? db : da));
      }
      if (u < n && v < n) gsd[v * npad + u] = gsv;
      al4[q] = a; gs4[q] = gsv;
    }
    gsu0 += gs4[0] + gs4[1];
    gsu1 += gs4[2] + gs4[3];
    float ca = gs4[0] + gs4[2], cb = gs4[1] + gs4[3];
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      ca += __shfl_xor_sync(0xffffffffu, ca, o);
      cb += __shfl_xor_sync(0xffffffffu, cb, o);
    }
    if (g == 0) {
      atomicAdd(S.gsv + va, ca);
      atomicAdd(S.gsv + vb, cb);
    }
    // mma #2: A = alpha^T (rows u0/u1, k-index t -> va, t+4 -> vb), B = G rows va / vb
    uint32_t ah[4], al[4];
    split_tf32(al4[0], ah[0], al[0]);
    split_tf32(al4[2], ah[1], al[1]);
    split_tf32(al4[1], ah[2], al[2]);
    split_tf32(al4[3], ah[3], al[3]);
    const float* ga = S.g + va * kSB + g;
#pragma unroll
    for (int nq = 0; nq < 8; nq += 4) {
      uint32_t bh0[4], bl0[4], bh1[4], bl1[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        split_tf32(ga[(nq + i) * 8], bh0[i], bl0[i]);
        split_tf32(ga[kSB + (nq + i) * 8], bh1[i], bl1[i]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) mma_tf32(acc[nq + i], al[0], al[1], al[2], al[3], bh0[i], bh1[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i) mma_tf32(acc[nq + i], ah[0], ah[1], ah[2], ah[3], bl0[i], bl1[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i) mma_tf32(acc[nq + i], ah[0], ah[1], ah[2], ah[3], bh0[i], bh1[i]);
    }
  }
}

__global__ void __launch_bounds__(256, 2) k_clq_bwd_mma(CliqueArgs A, const float* __restrict__ V,
                                                         const float* __restrict__ G, const float* __restrict__ agg,
                                                         const float* __restrict__ s_dense,
                                                         const float* __restrict__ m_save,
                                                         const float* __restrict__ l_save,
                                                         float* __restrict__ gs_dense, float* __restrict__ gsum_u,
                                                         float* __restrict__ gsum_v, float* __restrict__ gV) {
  extern __shared__ __align__(16) float smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = warp / A.tiles, ut = warp - slot * A.tiles;
  const int item = blockIdx.x * A.ipc + slot;
  const bool item_ok = item < A.n_items;
  const BwdSmem S = bwd_views(smem + (size_t)slot * A.item_floats, A.rv);
  int b = 0, h = 0, n = 0;
  const int32_t* nodes = nullptr;
  const int tid = threadIdx.x - slot * A.tiles * 32, nthr = A.tiles * 32;
  if (item_ok) {
    b = item / A.H;
    h = item - b * A.H;
    n = A.blk_ptr[b + 1] - A.blk_ptr[b];
    nodes = A.blk_nodes + A.blk_ptr[b];
    const int hoff = h * kD;
    const int c4 = (tid & 15) * 4;
#pragma unroll 1
    for (int half = 0; half < 2; ++half) {      // rv * 16 == 8 * nthr float4 slots, loads of a batch in flight together
      float4 vv[4], gg[4], ag[4];
      float mm[4], ll[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int row = (tid + (half * 4 + k) * nthr) >> 4;
        vv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        gg[k] = vv[k];
        ag[k] = vv[k];
        mm[k] = 0.f;
        ll[k] = 0.f;
        if (row < n) {
          const int64_t node = nodes[row];
          vv[k] = __ldg(reinterpret_cast<const float4*>(V + node * A.HD + hoff + c4));
          gg[k] = __ldg(reinterpret_cast<const float4*>(G + node * A.HD + hoff + c4));
          ag[k] = __ldg(reinterpret_cast<const float4*>(agg + node * A.HD + hoff + c4));
          if ((tid & 15) == 0) {
            mm[k] = m_save[node * A.H + h];
            ll[k] = 1.f / (l_save[node * A.H + h] + A.eps);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int row = (tid + (half * 4 + k) * nthr) >> 4;
        *reinterpret_cast<float4*>(S.v + row * kSB + c4) = vv[k];
        *reinterpret_cast<float4*>(S.g + row * kSB + c4) = gg[k];
        float d = gg[k].x * ag[k].x + gg[k].y * ag[k].y + gg[k].z * ag[k].z + gg[k].w * ag[k].w;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        if ((tid & 15) == 0) {
          S.dot[row] = d;
          S.m[row] = mm[k];
          S.linv[row] = ll[k];
          S.gsv[row] = 0.f;
        }
      }
    }
  }
  __syncthreads();
  if (item_ok && ut * 16 < n) {
    const int g = lane >> 2, t = lane & 3;
    const int u0 = ut * 16 + g, u1 = u0 + 8;
    const int npad = (n + 3) & ~3;
    const int64_t soff = (int64_t)A.blk_soff[b] * A.H + (int64_t)h * n * npad;
    const float* sd = s_dense + soff;
    float* gsd = gs_dense + soff;
    float acc[8][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) { acc[nb][0] = acc[nb][1] = acc[nb][2] = acc[nb][3] = 0.f; }
    float gsu0 = 0.f, gsu1 = 0.f;
    const int nvb = (n + 7) >> 3;
    int vb = 0;
    for (; vb + 4 <= nvb; vb += 4) bwd1_chunk<4>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, gsd, npad);
    switch (nvb - vb) {
      case 3: bwd1_chunk<3>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, gsd, npad); break;
      case 2: bwd1_chunk<2>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, gsd, npad); break;
      case 1: bwd1_chunk<1>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, gsd, npad); break;
      default: break;
    }
    gsu0 += __shfl_xor_sync(0xffffffffu, gsu0, 1);
    gsu0 += __shfl_xor_sync(0xffffffffu, gsu0, 2);
    gsu1 += __shfl_xor_sync(0xffffffffu, gsu1, 1);
    gsu1 += __shfl_xor_sync(0xffffffffu, gsu1, 2);
    const bool ok0 = u0 < n, ok1 = u1 < n;
    const int64_t n0 = ok0 ? nodes[u0] : 0, n1 = ok1 ? nodes[u1] : 0;
    if (t == 0) {
      if (ok0) gsum_u[n0 * A.H + h] = gsu0;
      if (ok1) gsum_u[n1 * A.H + h] = gsu1;
    }
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const int col = h * kD + nb * 8 + 2 * t;
      if (ok0) {
        float2* p = reinterpret_cast<float2*>(gV + n0 * A.HD + col);
        float2 o = *p;
        *p = make_float2(o.x + acc[nb][0], o.y + acc[nb][1]);
      }
      if (ok1) {
        float2* p = reinterpret_cast<float2*>(gV + n1 * A.HD + col);
        float2 o = *p;
        *p = make_float2(o.x + acc[nb][2], o.y + acc[nb][3]);
      }
    }
  }
  __syncthreads();
  if (item_ok)
    for (int row = tid; row < n; row += nthr) gsum_v[(int64_t)nodes[row] * A.H + h] = S.gsv[row];
}


// ------------------------------------------------------------------------------------------ backward, part 2
// Element-wise part, one work item = (block, head) on one 8-warp CTA.  Warp w owns the features
// d in [8w, 8w+8); lane = (dl = lane/4, ug = lane%4) owns feature d = 8w + dl and the GU sources
// u in [ug*GU, (ug+1)*GU) (4*GU >= n), so a thread holds Pi[u,d] and Qu[u,d] of its sources in registers:
//   step[u,v,d] = Pi[u,d] > -Pj'[v,d]                       (== z > 0)
//   Qu[u,d] = sum_v gs[u,v] step ,  Qv[v,d] = sum_u gs[u,v] step        (one compare + two predicated adds)
//   gPi'[u,d] = slope*gsum_u[u] + (1-slope)*Qu[u,d]   ->  gPi[u,d] += a_d gPi'
//   gPj'[v,d] = slope*gsum_v[v] + (1-slope)*Qv[v,d]   ->  gPj[v,d] += a_d gPj' ; g_table[type,d] += a_d gPj'
//   g_attn[d] += sum_u Pi[u,d] gPi'[u,d] + sum_v Pj'[v,d] gPj'[v,d]
// (leaky'(z) = slope + (1-slope) step ; leaky(z) = leaky'(z) z ; z = Pi[u] + Pj'[v].)
// Qv[v,d] is complete inside the CTA's four ug lanes (two shuffles): no atomics on node rows, no second phase.
__host__ __device__ inline int clique_bwd2_floats(int n_max, int gu) {
  return n_max * kD + n_max * (4 * gu > kD ? 4 * gu : kD) + 3 * n_max;
}

template <int GU>
__global__ void __launch_bounds__(256, 2) k_clq_bwd_elem(CliqueArgs A, const float* __restrict__ Pi,
                                                          const float* __restrict__ Pj,
                                                          const float* __restrict__ gs_dense,
                                                          const float* __restrict__ gsum_u,
                                                          const float* __restrict__ gsum_v, float* __restrict__ gPi,
                                                          float* __restrict__ gPj, float* __restrict__ g_attn,
                                                          float* __restrict__ g_table) {
  extern __shared__ __align__(16) float smem[];
  constexpr int SG = 4 * GU;                 // row stride of the gs tile
  const int item = blockIdx.x;
  const int b = item / A.H, h = item - b * A.H;
  const int n = A.blk_ptr[b + 1] - A.blk_ptr[b];
  const int32_t* __restrict__ nodes = A.blk_nodes + A.blk_ptr[b];
  const int type = A.blk_type[b];
  const int hoff = h * kD, npad = (n + 3) & ~3;
  float* sNPj = smem;                        // [n][64]   -(Pj + table)
  float* sGs = sNPj + A.ru * kD;             // [n][SG]   gs[v][u], zero beyond the block
  float* sGv = sGs + A.ru * (SG > kD ? SG : kD);   // [n]  slope * gsum_v
  float* sGu = sGv + A.ru;                   // [n]       slope * gsum_u
  int* sNode = reinterpret_cast<int*>(sGu + A.ru);   // [n]
  const float* gsd = gs_dense + (int64_t)A.blk_soff[b] * A.H + (int64_t)h * n * npad;

#pragma unroll 4
  for (int idx = threadIdx.x; idx < n * 16; idx += 256) {
    const int row = idx >> 4, c4 = (idx & 15) * 4;
    const int64_t node = nodes[row];
    float4 pj = __ldg(reinterpret_cast<const float4*>(Pj + node * A.HD + hoff + c4));
    if (A.table) {
      const float4 tb = __ldg(reinterpret_cast<const float4*>(A.table + (int64_t)type * A.HD + hoff + c4));
      pj.x += tb.x; pj.y += tb.y; pj.z += tb.z; pj.w += tb.w;
    }
    *reinterpret_cast<float4*>(sNPj + row * kD + c4) = make_float4(-pj.x, -pj.y, -pj.z, -pj.w);
  }
#pragma unroll 4
  for (int idx = threadIdx.x; idx < n * GU; idx += 256) {       // GU float4 per row
    const int row = idx / GU, c = (idx - row * GU) * 4;
    float4 gq = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < n) gq = __ldg(reinterpret_cast<const float4*>(gsd + (int64_t)row * npad + c));
    if (c + 1 >= n) gq.y = 0.f;
    if (c + 2 >= n) gq.z = 0.f;
    if (c + 3 >= n) gq.w = 0.f;
    *reinterpret_cast<float4*>(sGs + row * SG + c) = gq;
  }
  for (int row = threadIdx.x; row < n; row += 256) {
    const int node = nodes[row];
    sNode[row] = node;
    sGv[row] = A.slope * __ldg(gsum_v + (int64_t)node * A.H + h);
    sGu[row] = A.slope * __ldg(gsum_u + (int64_t)node * A.H + h);
  }

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dl = lane >> 2, ug = lane & 3;
  const int d = warp * 8 + dl, col = hoff + d;
  const int u0 = ug * GU;
  const float c1 = 1.f - A.slope;
  const float at = __ldg(A.attn + col);
  float pi[GU], qu[GU];
#pragma unroll
  for (int j = 0; j < GU; ++j) {
    qu[j] = 0.f;
    pi[j] = (u0 + j < n) ? __ldg(Pi + (int64_t)nodes[u0 + j] * A.HD + col) : 0.f;
  }
  __syncthreads();

  float gat = 0.f, tab = 0.f;
  // The cross-lane sum of Qv[v,d] (two shuffles over the four ug lanes) is software-pipelined: the shuffles of
  // destination v-1 are issued between the two halves of destination v's compare/add block.
  float qv_pend = 0.f, q_pend = 0.f;
  auto finish = [&](int v, float qv, float q) {       // d Pj'[v,d], parked over the dead -Pj'[v,d]
    if (ug == 0) {
      const float gq = fmaf(c1, qv, sGv[v]);
      gat = fmaf(-q, gq, gat);
      const float t = at * gq;
      tab += t;
      sNPj[v * kD + d] = t;
    }
  };
  for (int v = 0; v < n; ++v) {
    const float q = sNPj[v * kD + d];
    float4 g4[GU / 4];
#pragma unroll
    for (int k = 0; k < GU / 4; ++k) g4[k] = *reinterpret_cast<const float4*>(sGs + v * SG + u0 + 4 * k);
    float qa[4] = {0.f, 0.f, 0.f, 0.f};
    const float r1 = __shfl_xor_sync(0xffffffffu, qv_pend, 1);
#pragma unroll
    for (int j = 0; j < GU / 2; ++j) {
      const float4 g = g4[j >> 2];
      const float gsv = (j & 3) == 0 ? g.x : (j & 3) == 1 ? g.y : (j & 3) == 2 ? g.z : g.w;
      if (pi[j] > q) { qu[j] += gsv; qa[j & 3] += gsv; }
    }
    const float qv1 = qv_pend + r1;
    const float r2 = __shfl_xor_sync(0xffffffffu, qv1, 2);
#pragma unroll
    for (int j = GU / 2; j < GU; ++j) {
      const float4 g = g4[j >> 2];
      const float gsv = (j & 3) == 0 ? g.x : (j & 3) == 1 ? g.y : (j & 3) == 2 ? g.z : g.w;
      if (pi[j] > q) { qu[j] += gsv; qa[j & 3] += gsv; }
    }
    if (v > 0) finish(v - 1, qv1 + r2, q_pend);
    qv_pend = (qa[0] + qa[1]) + (qa[2] + qa[3]);
    q_pend = q;
  }
  {
    float qv = qv_pend + __shfl_xor_sync(0xffffffffu, qv_pend, 1);
    qv += __shfl_xor_sync(0xffffffffu, qv, 2);
    if (n > 0) finish(n - 1, qv, q_pend);
  }
  __syncthreads();                 // every warp is done with the gs tile: reuse it for the d Pi rows [u][64]
#pragma unroll
  for (int j = 0; j < GU; ++j) {
    if (u0 + j < n) {
      const float g0 = fmaf(c1, qu[j], sGu[u0 + j]);
      gat = fmaf(pi[j], g0, gat);
      sGs[(u0 + j) * kD + d] = at * g0;
    }
  }
  gat += __shfl_xor_sync(0xffffffffu, gat, 1);
  gat += __shfl_xor_sync(0xffffffffu, gat, 2);
  if (ug == 0) {
    atomicAdd(g_attn + col, gat);
    if (g_table) atomicAdd(g_table + (int64_t)type * A.HD + col, tab);
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < n * 16; idx += 256) {     // coalesced read-modify-write of the node rows
    const int row = idx >> 4, c4 = (idx & 15) * 4;
    const int64_t node = sNode[row];
    float4* pj = reinterpret_cast<float4*>(gPj + node * A.HD + hoff + c4);
    float4* pi4 = reinterpret_cast<float4*>(gPi + node * A.HD + hoff + c4);
    const float4 aj = *reinterpret_cast<const float4*>(sNPj + row * kD + c4);
    const float4 ai = *reinterpret_cast<const float4*>(sGs + row * kD + c4);
    float4 oj = *pj, oi = *pi4;
    oj.x += aj.x; oj.y += aj.y; oj.z += aj.z; oj.w += aj.w;
    *pj = oj;
    if (pi4 == pj) oi = oj;        // never: gPi and gPj are distinct tensors
    oi.x += ai.x; oi.y += ai.y; oi.z += ai.z; oi.w += ai.w;
    *pi4 = oi;
  }
}

// ------------------------------------------------------------------------------------------ backward, fused
// Parts 1 and 2 in ONE kernel for blocks of more than 64 members (one item per CTA): d loss / d logit never leaves
// the CTA (part 1 writes the [v][u] tile part 2 reads straight into shared memory: no 26 MB gs_dense round trip per
// layer, no gsum_u / gsum_v arrays, one staging of the block instead of two) and, with two CTAs per SM running out
// of phase, the latency-bound tensor-core phase of one overlaps the issue-bound element-wise phase of the other.
// Shared memory: [V | G | m | linv | dot | gsv] (phase 1; -Pj' overlays V in phase 2) | gs tile | gsum_v | gsum_u | ids.
__host__ __device__ inline int clique_bwdf_floats(int ru, int rv, int gu) {
  const int sg = 4 * gu > kD ? 4 * gu : kD;
  return 2 * rv * kSB + 4 * rv + ru * sg + 3 * ru;
}

template <int VB>
__device__ __forceinline__ void bwdf_chunk(const CliqueArgs& A, const BwdSmem& S, float (&acc)[8][4], float& gsu0,
                                           float& gsu1, int n, int u0, int u1, int vbase, int g, int t,
                                           const float* __restrict__ sd, float* __restrict__ sGs, int SG, int npad) {
  float c[VB][4], sv[VB][4];
#pragma unroll
  for (int j = 0; j < VB; ++j) {          // saved logits first: their latency hides behind mma #1
    c[j][0] = c[j][1] = c[j][2] = c[j][3] = 0.f;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int u = (q & 2) ? u1 : u0, v = vbase + 8 * j + 2 * t + (q & 1);
      sv[j][q] = (u < n && v < n) ? __ldg(sd + v * npad + u) : 0.f;
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    uint32_t ah[4], al[4];
    split_tf32(S.v[u0 * kSB + k * 8 + t], ah[0], al[0]);
    split_tf32(S.v[u1 * kSB + k * 8 + t], ah[1], al[1]);
    split_tf32(S.v[u0 * kSB + k * 8 + t + 4], ah[2], al[2]);
    split_tf32(S.v[u1 * kSB + k * 8 + t + 4], ah[3], al[3]);
    uint32_t bh0[VB], bl0[VB], bh1[VB], bl1[VB];
#pragma unroll
    for (int j = 0; j < VB; ++j) {
      const float* gp = S.g + (vbase + 8 * j + g) * kSB + k * 8 + t;
      split_tf32(gp[0], bh0[j], bl0[j]);
      split_tf32(gp[4], bh1[j], bl1[j]);
    }
#pragma unroll
    for (int j = 0; j < VB; ++j) mma_tf32(c[j], al[0], al[1], al[2], al[3], bh0[j], bh1[j]);
#pragma unroll
    for (int j = 0; j < VB; ++j) mma_tf32(c[j], ah[0], ah[1], ah[2], ah[3], bl0[j], bl1[j]);
#pragma unroll
    for (int j = 0; j < VB; ++j) mma_tf32(c[j], ah[0], ah[1], ah[2], ah[3], bh0[j], bh1[j]);
  }
#pragma unroll
  for (int j = 0; j < VB; ++j) {
    const int va = vbase + 8 * j + 2 * t, vb = va + 1;
    const float ma = S.m[va], mb = S.m[vb], ia = S.linv[va], ib = S.linv[vb], da = S.dot[va], db = S.dot[vb];
    float al4[4], gs4[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int u = (q & 2) ? u1 : u0, v = (q & 1) ? vb : va;
      const bool ok = u < n && v < n && u != v;
      float a = 0.f, gsv = 0.f;
      if (ok) {
        a = __expf(sv[j][q] - ((q & 1) ? mb : ma)) * ((q & 1) ? ib : ia);
        gsv = a * (c[j][q] - ((q & 1) ? db : da));
      }
      if (u < n && v < n) sGs[v * SG + u] = gsv;
      al4[q] = a; gs4[q] = gsv;
    }
    gsu0 += gs4[0] + gs4[1];
    gsu1 += gs4[2] + gs4[3];
    float ca = gs4[0] + gs4[2], cb = gs4[1] + gs4[3];
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
      ca += __shfl_xor_sync(0xffffffffu, ca, o);
      cb += __shfl_xor_sync(0xffffffffu, cb, o);
    }
    if (g == 0) {
      atomicAdd(S.gsv + va, ca);
      atomicAdd(S.gsv + vb, cb);
    }
    uint32_t ah[4], al[4];
    split_tf32(al4[0], ah[0], al[0]);
    split_tf32(al4[2], ah[1], al[1]);
    split_tf32(al4[1], ah[2], al[2]);
    split_tf32(al4[3], ah[3], al[3]);
    const float* ga = S.g + va * kSB + g;
#pragma unroll
    for (int nq = 0; nq < 8; nq += 4) {
      uint32_t bh0[4], bl0[4], bh1[4], bl1[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        split_tf32(ga[(nq + i) * 8], bh0[i], bl0[i]);
        split_tf32(ga[kSB + (nq + i) * 8], bh1[i], bl1[i]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) mma_tf32(acc[nq + i], al[0], al[1], al[2], al[3], bh0[i], bh1[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i) mma_tf32(acc[nq + i], ah[0], ah[1], ah[2], ah[3], bl0[i], bl1[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i) mma_tf32(acc[nq + i], ah[0], ah[1], ah[2], ah[3], bh0[i], bh1[i]);
    }
  }
}

template <int GU>
__global__ void __launch_bounds__(256, 2) k_clq_bwd_fused(CliqueArgs A, const float* __restrict__ Pi,
                                                           const float* __restrict__ Pj, const float* __restrict__ V,
                                                           const float* __restrict__ G, const float* __restrict__ agg,
                                                           const float* __restrict__ s_dense,
                                                           const float* __restrict__ m_save,
                                                           const float* __restrict__ l_save, float* __restrict__ gPi,
                                                           float* __restrict__ gPj, float* __restrict__ gV,
                                                           float* __restrict__ g_attn, float* __restrict__ g_table) {
  pdl_sync();
  extern __shared__ __align__(16) float smem[];
  constexpr int SG = 4 * GU;
  constexpr int SGW = SG > kD ? SG : kD;
  const int item = blockIdx.x;
  const int b = item / A.H, h = item - b * A.H;
  const int n = A.blk_ptr[b + 1] - A.blk_ptr[b];
  const int32_t* __restrict__ nodes = A.blk_nodes + A.blk_ptr[b];
  const int type = A.blk_type[b];
  const int hoff = h * kD, npad = (n + 3) & ~3;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const BwdSmem S = bwd_views(smem, A.rv);
  float* sGs = smem + 2 * A.rv * kSB + 4 * A.rv;   // [ru][SGW] gs[v][u], zero beyond the block
  float* sGv = sGs + A.ru * SGW;                  // [ru] slope * gsum_v
  float* sGu = sGv + A.ru;                        // [ru] slope * gsum_u
  int* sNode = reinterpret_cast<int*>(sGu + A.ru);
  const float* sd = s_dense + (int64_t)A.blk_soff[b] * A.H + (int64_t)h * n * npad;

  // ---------------------------------------------------------------- phase 1: stage V, G, softmax state
  for (int idx = tid; idx < A.rv * 16; idx += 256) {
    const int row = idx >> 4, c4 = (idx & 15) * 4;
    float4 vv = make_float4(0.f, 0.f, 0.f, 0.f), gg = vv, ag = vv;
    float mm = 0.f, ll = 0.f;
    if (row < n) {
      const int64_t node = nodes[row];
      vv = __ldg(reinterpret_cast<const float4*>(V + node * A.HD + hoff + c4));
      gg = __ldg(reinterpret_cast<const float4*>(G + node * A.HD + hoff + c4));
      ag = __ldg(reinterpret_cast<const float4*>(agg + node * A.HD + hoff + c4));
      if ((idx & 15) == 0) {
        mm = m_save[node * A.H + h];
        ll = 1.f / (l_save[node * A.H + h] + A.eps);
      }
    }
    *reinterpret_cast<float4*>(S.v + row * kSB + c4) = vv;
    *reinterpret_cast<float4*>(S.g + row * kSB + c4) = gg;
    float d = gg.x * ag.x + gg.y * ag.y + gg.z * ag.z + gg.w * ag.w;
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
    if ((idx & 15) == 0) {
      S.dot[row] = d;
      S.m[row] = mm;
      S.linv[row] = ll;
      S.gsv[row] = 0.f;
    }
  }
  for (int idx = tid; idx < A.ru * SGW / 4; idx += 256)
    reinterpret_cast<float4*>(sGs)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int row = tid; row < A.ru; row += 256) {
    sNode[row] = row < n ? nodes[row] : 0;
    sGu[row] = 0.f;
  }
  __syncthreads();
  if (warp * 16 < n) {
    const int g = lane >> 2, t = lane & 3;
    const int u0 = warp * 16 + g, u1 = u0 + 8;
    float acc[8][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) { acc[nb][0] = acc[nb][1] = acc[nb][2] = acc[nb][3] = 0.f; }
    float gsu0 = 0.f, gsu1 = 0.f;
    const int nvb = (n + 7) >> 3;
    int vb = 0;
    for (; vb + 4 <= nvb; vb += 4) bwdf_chunk<4>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, sGs, SGW, npad);
    switch (nvb - vb) {
      case 3: bwdf_chunk<3>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, sGs, SGW, npad); break;
      case 2: bwdf_chunk<2>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, sGs, SGW, npad); break;
      case 1: bwdf_chunk<1>(A, S, acc, gsu0, gsu1, n, u0, u1, vb * 8, g, t, sd, sGs, SGW, npad); break;
      default: break;
    }
    gsu0 += __shfl_xor_sync(0xffffffffu, gsu0, 1);
    gsu0 += __shfl_xor_sync(0xffffffffu, gsu0, 2);
    gsu1 += __shfl_xor_sync(0xffffffffu, gsu1, 1);
    gsu1 += __shfl_xor_sync(0xffffffffu, gsu1, 2);
    const bool ok0 = u0 < n, ok1 = u1 < n;
    if (t == 0) {
      if (ok0) sGu[u0] = A.slope * gsu0;
      if (ok1) sGu[u1] = A.slope * gsu1;
    }
    const int64_t n0 = ok0 ? sNode[u0] : 0, n1 = ok1 ? sNode[u1] : 0;
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const int col = hoff + nb * 8 + 2 * t;
      if (ok0) {
        float2* p = reinterpret_cast<float2*>(gV + n0 * A.HD + col);
        float2 o = *p;
        *p = make_float2(o.x + acc[nb][0], o.y + acc[nb][1]);
      }
      if (ok1) {
        float2* p = reinterpret_cast<float2*>(gV + n1 * A.HD + col);
        float2 o = *p;
        *p = make_float2(o.x + acc[nb][2], o.y + acc[nb][3]);
      }
    }
  }
  __syncthreads();
  // ---------------------------------------------------------------- phase 2 (the body of k_clq_bwd_elem)
  float* sNPj = smem;                              // [n][64] -(Pj + table), over the dead V tile
  for (int row = tid; row < n; row += 256) sGv[row] = A.slope * S.gsv[row];
  __syncthreads();                                 // S.gsv lives inside the region sNPj is about to overwrite
#pragma unroll 4
  for (int idx = tid; idx < n * 16; idx += 256) {
    const int row = idx >> 4, c4 = (idx & 15) * 4;
    const int64_t node = sNode[row];
    float4 pj = __ldg(reinterpret_cast<const float4*>(Pj + node * A.HD + hoff + c4));
    if (A.table) {
      const float4 tb = __ldg(reinterpret_cast<const float4*>(A.table + (int64_t)type * A.HD + hoff + c4));
      pj.x += tb.x; pj.y += tb.y; pj.z += tb.z; pj.w += tb.w;
    }
    *reinterpret_cast<float4*>(sNPj + row * kD + c4) = make_float4(-pj.x, -pj.y, -pj.z, -pj.w);
  }
  const int dl = lane >> 2, ug = lane & 3;
  const int d = warp * 8 + dl, col = hoff + d;
  const int u0 = ug * GU;
  const float c1 = 1.f - A.slope;
  const float at = __ldg(A.attn + col);
  float pi[GU], qu[GU];
#pragma unroll
  for (int j = 0; j < GU; ++j) {
    qu[j] = 0.f;
    pi[j] = (u0 + j < n) ? __ldg(Pi + (int64_t)sNode[u0 + j] * A.HD + col) : 0.f;
  }
  __syncthreads();

  float gat = 0.f, tab = 0.f;
  float qv_pend = 0.f, q_pend = 0.f;
  auto finish = [&](int v, float qv, float q) {
    if (ug == 0) {
      const float gq = fmaf(c1, qv, sGv[v]);
      gat = fmaf(-q, gq, gat);
      const float t = at * gq;
      tab += t;
      sNPj[v * kD + d] = t;
    }
  };
  for (int v = 0; v < n; ++v) {
    const float q = sNPj[v * kD + d];
    float4 g4[GU / 4];
#pragma unroll
    for (int k = 0; k < GU / 4; ++k) g4[k] = *reinterpret_cast<const float4*>(sGs + v * SGW + u0 + 4 * k);
    float qa[4] = {0.f, 0.f, 0.f, 0.f};
    const float r1 = __shfl_xor_sync(0xffffffffu, qv_pend, 1);
#pragma unroll
    for (int j = 0; j < GU / 2; ++j) {
      const float4 g = g4[j >> 2];
      const float gsv = (j & 3) == 0 ? g.x : (j & 3) == 1 ? g.y : (j & 3) == 2 ? g.z : g.w;
      if (pi[j] > q) { qu[j] += gsv; qa[j & 3] += gsv; }
    }
    const float qv1 = qv_pend + r1;
    const float r2 = __shfl_xor_sync(0xffffffffu, qv1, 2);
#pragma unroll
    for (int j = GU / 2; j < GU; ++j) {
      const float4 g = g4[j >> 2];
      const float gsv = (j & 3) == 0 ? g.x : (j & 3) == 1 ? g.y : (j & 3) == 2 ? g.z : g.w;
      if (pi[j] > q) { qu[j] += gsv; qa[j & 3] += gsv; }
    }
    if (v > 0) finish(v - 1, qv1 + r2, q_pend);
    qv_pend = (qa[0] + qa[1]) + (qa[2] + qa[3]);
    q_pend = q;
  }
  {
    float qv = qv_pend + __shfl_xor_sync(0xffffffffu, qv_pend, 1);
    qv += __shfl_xor_sync(0xffffffffu, qv, 2);
    if (n > 0) finish(n - 1, qv, q_pend);
  }
  __syncthreads();                 // every warp is done with the gs tile: reuse it for the d Pi rows [u][64]
#pragma unroll
  for (int j = 0; j < GU; ++j) {
    if (u0 + j < n) {
      const float g0 = fmaf(c1, qu[j], sGu[u0 + j]);
      gat = fmaf(pi[j], g0, gat);
      sGs[(u0 + j) * kD + d] = at * g0;
    }
  }
  gat += __shfl_xor_sync(0xffffffffu, gat, 1);
  gat += __shfl_xor_sync(0xffffffffu, gat, 2);
  if (ug == 0) {
    atomicAdd(g_attn + col, gat);
    if (g_table) atomicAdd(g_table + (int64_t)type * A.HD + col, tab);
  }
  __syncthreads();
  for (int idx = tid; idx < n * 16; idx += 256) {     // coalesced read-modify-write of the node rows
    const int row = idx >> 4, c4 = (idx & 15) * 4;
    const int64_t node = sNode[row];
    float4* pj = reinterpret_cast<float4*>(gPj + node * A.HD + hoff + c4);
    float4* pi4 = reinterpret_cast<float4*>(gPi + node * A.HD + hoff + c4);
    const float4 aj = *reinterpret_cast<const float4*>(sNPj + row * kD + c4);
    const float4 ai = *reinterpret_cast<const float4*>(sGs + row * kD + c4);
    float4 oj = *pj, oi = *pi4;
    oj.x += aj.x; oj.y += aj.y; oj.z += aj.z; oj.w += aj.w;
    *pj = oj;
    oi.x += ai.x; oi.y += ai.y; oi.z += ai.z; oi.w += ai.w;
    *pi4 = oi;
  }
}

// ------------------------------------------------------------------------------------------ host side
static bool aligned16p(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

bool clique_supported(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef) {
  if (!g || !g->cliques || !g->residual || g->cliques->n_blocks <= 0) return false;
  if (a->flavour != B200GAT_FLAVOUR_ADDITIVE || a->dim != kD || a->heads <= 0) return false;
  const int mode = ef ? ef->mode : B200GAT_EDGE_NONE;
  if (mode != B200GAT_EDGE_NONE && mode != B200GAT_EDGE_TABLE) return false;
  const int mb = g->cliques->max_block;
  return mb >= 2 && mb <= B200GAT_MAX_CLIQUE;
}

static int fill_clique(CliqueArgs& A, const b200gat_cliques* c, const b200gat_attn* a,
                       const b200gat_edge_feats* ef) {
  B200_CHECK_ARG(c->blk_ptr && c->blk_nodes && c->blk_type && c->blk_soff, "null clique tables");
  B200_CHECK_ARG(a->attn && aligned16p(a->attn), "attn must be 16-byte aligned");
  A.H = a->heads; A.HD = a->heads * kD;
  A.n_items = c->n_blocks * a->heads;
  A.tiles = (c->max_block + 15) / 16;
  A.ipc = A.tiles >= 8 ? 1 : 8 / A.tiles;
  A.ru = (c->max_block + 7) & ~7;
  A.rv = A.tiles * 16;
  A.blk_ptr = c->blk_ptr; A.blk_nodes = c->blk_nodes; A.blk_type = c->blk_type; A.blk_soff = c->blk_soff;
  A.slope = a->slope; A.eps = a->softmax_eps; A.attn = a->attn; A.out_bias = a->out_bias;
  A.table = (ef && ef->mode == B200GAT_EDGE_TABLE) ? ef->table : nullptr;
  if (A.table) B200_CHECK_ARG(aligned16p(A.table), "edge table must be 16-byte aligned");
  return B200GAT_OK;
}

template <auto Kernel>
static int set_smem(size_t bytes, const char* name) {
  if (bytes > 227 * 1024) {
    set_error("%s: %zu bytes of shared memory needed", name, bytes);
    return B200GAT_ERR_UNSUPPORTED;
  }
  // one instantiation of this function (and of `granted`) per kernel: the attribute is raised only when a
  // launch needs more than any before it (the call costs a few microseconds on a host-bound step)
  static std::atomic<size_t> granted{48 * 1024};
  if (bytes > granted.load(std::memory_order_relaxed)) {
    if (cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) {
      set_error("%s: cudaFuncSetAttribute(%zu) failed", name, bytes);
      return B200GAT_ERR_CUDA;
    }
    granted.store(bytes, std::memory_order_relaxed);
  }
  return B200GAT_OK;
}

int clique_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef, const float* Pi,
               const float* Pj, const float* V, float* out, float* s_dense, float* m_save, float* l_save,
               float* agg_save, cudaStream_t st) {
  CliqueArgs A;
  int rc = fill_clique(A, g->cliques, a, ef);
  if (rc) return rc;
  B200_CHECK_ARG(s_dense, "null s_dense");
  B200_CHECK_ARG(aligned16p(Pi) && aligned16p(Pj) && aligned16p(V) && aligned16p(out) && aligned16p(agg_save) &&
                     aligned16p(a->out_bias),
                 "clique kernels need 16-byte aligned tensors");
  const bool alias = (V == Pi);
  A.item_floats = clique_item_floats(A.ru, A.rv, alias);
  const size_t smem = (size_t)A.ipc * A.item_floats * sizeof(float);
  const unsigned grid = (unsigned)ceil_div(A.n_items, A.ipc), block = 32u * A.ipc * A.tiles;
  ScopedTimer tm(B200GAT_TIMING_CLQ_FWD, st);
  if (alias) {
    if ((rc = set_smem<k_clq_fwd<true>>(smem, "k_clq_fwd"))) return rc;
    launch_k(k_clq_fwd<true>, dim3(grid), dim3(block), smem, st, A, Pi, Pj, V, out, s_dense, m_save, l_save, agg_save);
  } else {
    if ((rc = set_smem<k_clq_fwd<false>>(smem, "k_clq_fwd"))) return rc;
    launch_k(k_clq_fwd<false>, dim3(grid), dim3(block), smem, st, A, Pi, Pj, V, out, s_dense, m_save, l_save, agg_save);
  }
  B200_CHECK_LAUNCH("k_clq_fwd");
  return B200GAT_OK;
}

// scratch: gs_dense [H * s_elems] | gsum_u [N * H] | gsum_v [N * H]
int clique_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef, const float* Pi,
               const float* Pj, const float* V, const float* s_dense, const float* m_save, const float* l_save,
               const float* agg_save, const float* g_out, float* gPi, float* gPj, float* gV, float* g_attn,
               float* g_table, float* scratch, cudaStream_t st) {
  CliqueArgs A;
  int rc = fill_clique(A, g->cliques, a, ef);
  if (rc) return rc;
  B200_CHECK_ARG(s_dense && agg_save && scratch, "null s_dense / agg_save / scratch");
  B200_CHECK_ARG(aligned16p(Pi) && aligned16p(Pj) && aligned16p(V) && aligned16p(g_out) && aligned16p(agg_save) &&
                     aligned16p(gPi) && aligned16p(gPj) && aligned16p(gV) && aligned16p(scratch),
                 "clique kernels need 16-byte aligned tensors");
  if (A.table) B200_CHECK_ARG(g_table, "null g_table");
  {
    // blocks of more than 64 members: parts 1 and 2 fused (gs stays in shared memory)
    const int gu = (((g->cliques->max_block + 3) / 4) + 3) & ~3;
    const size_t smem_f = (size_t)clique_bwdf_floats(A.ru, A.rv, gu) * sizeof(float);
    static const bool no_fused = getenv("B200GAT_NO_FUSED_CLQ_BWD") != nullptr;
    if (A.ipc == 1 && A.tiles <= 8 && smem_f <= 227 * 1024 && !no_fused) {
      ScopedTimer tm(B200GAT_TIMING_CLQ_BWD_ELEM, st);
      float* gt = A.table ? g_table : nullptr;
#define CLQ_BWDF(GU_)                                                                                          \
  case GU_:                                                                                                    \
    if ((rc = set_smem<k_clq_bwd_fused<GU_>>(smem_f, "k_clq_bwd_fused"))) return rc;                           \
    launch_k(k_clq_bwd_fused<GU_>, dim3(A.n_items), dim3(256), smem_f, st, A, Pi, Pj, V, g_out, agg_save, s_dense, \
             m_save, l_save, gPi, gPj, gV, g_attn, gt);                                                        \
    break;
      switch (gu) {
        CLQ_BWDF(20) CLQ_BWDF(24) CLQ_BWDF(28) CLQ_BWDF(32)
        default: set_error("k_clq_bwd_fused: block size %d not covered", g->cliques->max_block); return B200GAT_ERR_UNSUPPORTED;
      }
#undef CLQ_BWDF
      B200_CHECK_LAUNCH("k_clq_bwd_fused");
      return B200GAT_OK;
    }
  }
  float* gs_dense = scratch;
  float* gsum_u = gs_dense + (((int64_t)g->cliques->s_elems * A.H + 3) & ~3ll);
  float* gsum_v = gsum_u + (int64_t)g->n_nodes * A.H;
  const unsigned grid = (unsigned)ceil_div(A.n_items, A.ipc), block = 32u * A.ipc * A.tiles;
  {
    ScopedTimer tm(B200GAT_TIMING_CLQ_BWD_MMA, st);
    A.item_floats = clique_bwd1_floats(A.rv);
    const size_t smem = (size_t)A.ipc * A.item_floats * sizeof(float);
    if ((rc = set_smem<k_clq_bwd_mma>(smem, "k_clq_bwd_mma"))) return rc;
    k_clq_bwd_mma<<<grid, block, smem, st>>>(A, V, g_out, agg_save, s_dense, m_save, l_save, gs_dense, gsum_u,
                                            gsum_v, gV);
    B200_CHECK_LAUNCH("k_clq_bwd_mma");
  }
  {
    ScopedTimer tm(B200GAT_TIMING_CLQ_BWD_ELEM, st);
    const int gu = (((g->cliques->max_block + 3) / 4) + 3) & ~3;          // sources per lane group, multiple of 4
    const size_t smem = (size_t)clique_bwd2_floats(A.ru, gu) * sizeof(float);
    float* gt = A.table ? g_table : nullptr;
#define CLQ_BWD2(GU_)                                                                                         \
  case GU_:                                                                                                   \
    if ((rc = set_smem<k_clq_bwd_elem<GU_>>(smem, "k_clq_bwd_elem"))) return rc;                              \
    k_clq_bwd_elem<GU_><<<A.n_items, 256, smem, st>>>(A, Pi, Pj, gs_dense, gsum_u, gsum_v, gPi, gPj, g_attn, gt); \
    break;
    switch (gu) {
      CLQ_BWD2(4) CLQ_BWD2(8) CLQ_BWD2(12) CLQ_BWD2(16) CLQ_BWD2(20) CLQ_BWD2(24) CLQ_BWD2(28) CLQ_BWD2(32)
      default: set_error("k_clq_bwd_elem: block size %d not covered", g->cliques->max_block); return B200GAT_ERR_UNSUPPORTED;
    }
#undef CLQ_BWD2
    B200_CHECK_LAUNCH("k_clq_bwd_elem");
  }
  return B200GAT_OK;
}

}  // namespace b200gat
