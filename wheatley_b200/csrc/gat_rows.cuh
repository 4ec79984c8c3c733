// Shared pieces of the warp-per-node attention kernels (gat_fast.cu) and their windowed variants (gat_win.cu):
// argument block, row load / store helpers (lane l owns VPL contiguous features of a row), edge-term helpers,
// work assignment and the forward's per-destination epilogue.
#pragma once
#include <math.h>
#include "common.cuh"
#include "gat_kernels.cuh"

namespace b200gat {

struct FastArgs {
  int N, E, flags;
  const int32_t *in_ptr, *in_src, *out_ptr, *out_dst, *out_pos, *dst_order, *src_order;
  const int32_t *in_pos;      // residual graphs: position of each in-edge in the full dst-sorted order
  const int32_t *node_blk;    // clique members (>= 0): the forward leaves their softmax state un-normalised
  const int32_t *hub_dst, *n_hub_dst, *hub_src, *n_hub_src;
  int hub_degree, hub_blocks;   // hub_blocks = 0 disables segment splitting
  float* hub_scratch;           // [hub_cap][hub_parts][HD + 64] partial states of split hub segments, or NULL
  int hub_parts, hub_cap;
  int* hub_cnt;                 // [hub_cap] arrival counters of the slices (zero between launches), or NULL
  int H;
  float slope, eps;
  const float *attn, *out_bias;
  int emode, T, K;
  const float* table;
  const int32_t* etype;
  const float* attr;
  const float* affine;
};

template <int VPL>
__device__ __forceinline__ void ldrow(float (&r)[VPL], const float* __restrict__ p) {
  if constexpr (VPL == 8) {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  } else if constexpr (VPL == 4) {
    float4 a = __ldg(reinterpret_cast<const float4*>(p));
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
  } else if constexpr (VPL == 2) {
    float2 a = __ldg(reinterpret_cast<const float2*>(p));
    r[0] = a.x; r[1] = a.y;
  } else {
    r[0] = __ldg(p);
  }
}
template <int VPL>
__device__ __forceinline__ void ldrow_s(float (&r)[VPL], const float* p) {     // plain (coherent) loads
  if constexpr (VPL == 8) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  } else if constexpr (VPL == 4) {
    const float4 a = reinterpret_cast<const float4*>(p)[0];
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
  } else if constexpr (VPL == 2) {
    const float2 a = reinterpret_cast<const float2*>(p)[0];
    r[0] = a.x; r[1] = a.y;
  } else {
    r[0] = p[0];
  }
}
template <int VPL>
__device__ __forceinline__ void ldrow_cg(float (&r)[VPL], const float* p) {    // L2-coherent (written by other CTAs)
  if constexpr (VPL == 8) {
    const float4 a = __ldcg(reinterpret_cast<const float4*>(p)), b = __ldcg(reinterpret_cast<const float4*>(p) + 1);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w; r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  } else if constexpr (VPL == 4) {
    const float4 a = __ldcg(reinterpret_cast<const float4*>(p));
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
  } else if constexpr (VPL == 2) {
    const float2 a = __ldcg(reinterpret_cast<const float2*>(p));
    r[0] = a.x; r[1] = a.y;
  } else {
    r[0] = __ldcg(p);
  }
}
template <int VPL>
__device__ __forceinline__ void strow(float* __restrict__ p, const float (&r)[VPL]) {
  if constexpr (VPL == 8) {
    reinterpret_cast<float4*>(p)[0] = make_float4(r[0], r[1], r[2], r[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(r[4], r[5], r[6], r[7]);
  } else if constexpr (VPL == 4) {
    reinterpret_cast<float4*>(p)[0] = make_float4(r[0], r[1], r[2], r[3]);
  } else if constexpr (VPL == 2) {
    reinterpret_cast<float2*>(p)[0] = make_float2(r[0], r[1]);
  } else {
    p[0] = r[0];
  }
}
template <int LPH>
__device__ __forceinline__ float head_sum(float v) {
#pragma unroll
  for (int o = LPH >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Edge term for the current type: qt = base + table[t]  (base = Pj[v] or Pi[u]).
template <int VPL>
__device__ __forceinline__ void type_row(const FastArgs& A, int t, int c0, const float (&base)[VPL],
                                         float (&qt)[VPL]) {
  float pe[VPL];
  ldrow<VPL>(pe, A.table + (int64_t)t * (32 * VPL) + c0);
#pragma unroll
  for (int i = 0; i < VPL; ++i) qt[i] = base[i] + pe[i];
}
// Affine part of the edge term: z += sum_k attr[pos,k] * affine[k].  The K <= 8 attributes of an edge are fetched by
// the first K lanes in ONE load, issued together with the edge's row gathers (affine_fetch), and broadcast by shuffles
// here; the rows of the affine map stay in L1.  (A first version read attr[pos,k] inside this loop: K dependent
// global round trips per edge, which made the PSP workloads latency-bound.)
__device__ __forceinline__ float affine_fetch(const FastArgs& A, int64_t pos, int lane) {
  return lane < A.K ? __ldg(A.attr + pos * A.K + lane) : 0.f;
}
template <int VPL>
__device__ __forceinline__ void affine_add(const FastArgs& A, float my_attr, int c0, float (&z)[VPL]) {
  for (int k = 0; k < A.K; ++k) {
    const float av = __shfl_sync(0xffffffffu, my_attr, k);
    if (av != 0.f) {                                   // warp-uniform
      float w[VPL];
      ldrow<VPL>(w, A.affine + (int64_t)k * (32 * VPL) + c0);
#pragma unroll
      for (int i = 0; i < VPL; ++i) z[i] = fmaf(av, w[i], z[i]);
    }
  }
}

#ifndef B200GAT_FAST_MINB
#define B200GAT_FAST_MINB 2      // resident CTAs per SM the node kernels are compiled for (register cap)
#endif
constexpr int kFastThreads = 256;
constexpr int kWarps = kFastThreads / 32;
constexpr int kHubBlocks = 148;  // CTAs at the front of the grid that walk the hub list
constexpr int kHubRec = 64;      // extra floats per partial record: m[32] + l[32] (one per lane)
constexpr int kUnroll = 4;

// Work assignment shared by the three kernels.  Blocks [0, hub_blocks) walk the hub list: the
// whole CTA takes one hub node at a time and each warp a contiguous slice of its adjacency
// segment.  The other blocks take one node per warp (in the locality order) and skip hubs.
struct WorkItem {
  int node, rb, re, seg_b, seg_e;
  bool active;
};
// A hub segment is cut into `parts` slices (one CTA each) of kWarps sub-slices (one warp each).
__device__ __forceinline__ WorkItem hub_item(const int32_t* __restrict__ ptr, int node, int wi, int part = 0,
                                             int parts = 1) {
  WorkItem it;
  it.node = node;
  it.seg_b = ptr[node];
  it.seg_e = ptr[node + 1];
  const int ns = kWarps * parts;
  const int per = (((it.seg_e - it.seg_b) + ns - 1) / ns + kUnroll - 1) & ~(kUnroll - 1);
  it.rb = min(it.seg_e, it.seg_b + (part * kWarps + wi) * per);
  it.re = min(it.seg_e, it.rb + per);
  it.active = true;
  return it;
}
// Number of slices per hub: the scratch must exist and hold every hub of the list.
__device__ __forceinline__ int hub_parts_of(const FastArgs& A, int n_hub) {
  return (A.hub_scratch && A.hub_parts > 1 && n_hub <= A.hub_cap) ? A.hub_parts : 1;
}
__device__ __forceinline__ WorkItem node_item(const FastArgs& A, const int32_t* __restrict__ ptr,
                                              const int32_t* __restrict__ order, int wi, int group) {
  WorkItem it;
  const int w = group * kWarps + wi;
  it.active = w < A.N;
  it.node = 0; it.seg_b = it.seg_e = 0;
  if (it.active) {
    it.node = order ? order[w] : w;
    it.seg_b = ptr[it.node];
    it.seg_e = ptr[it.node + 1];
    if (A.hub_blocks > 0 && it.seg_e - it.seg_b > A.hub_degree) it.active = false;   // a hub CTA has it
  }
  it.rb = it.seg_b;
  it.re = it.seg_e;
  return it;
}

// Last step of the forward for one destination: clique members keep the un-normalised state (the
// dense kernel continues and finalises it); everybody else is normalised, biased and saved.
template <int VPL>
__device__ __forceinline__ void fwd_finalize(const FastArgs& A, int v, bool any, float (&acc)[VPL], float m,
                                             float l, int c0, int h, bool head_lead, float* __restrict__ out,
                                             float* __restrict__ m_save, float* __restrict__ l_save,
                                             float* __restrict__ agg_save) {
  constexpr int HD = 32 * VPL;
  if (A.node_blk && A.node_blk[v] >= 0) {
    strow<VPL>(out + (int64_t)v * HD + c0, acc);
    if (head_lead) {
      m_save[(int64_t)v * A.H + h] = m;
      l_save[(int64_t)v * A.H + h] = l;
    }
    return;
  }
  const float inv = any ? 1.f / (l + A.eps) : 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) acc[i] *= inv;
  if (agg_save) strow<VPL>(agg_save + (int64_t)v * HD + c0, acc);
  if (A.out_bias) {
    float ob[VPL];
    ldrow<VPL>(ob, A.out_bias + c0);
#pragma unroll
    for (int i = 0; i < VPL; ++i) acc[i] += ob[i];
  }
  strow<VPL>(out + (int64_t)v * HD + c0, acc);
  if (head_lead) {
    m_save[(int64_t)v * A.H + h] = any ? m : 0.f;
    l_save[(int64_t)v * A.H + h] = l;
  }
}

// Merge of the slices of split hub `w` (its index in the hub list), by ONE warp, in slice order (so the result does
// not depend on which CTA runs it): the last slice to arrive (hub_arrive) does it inside the hub CTA; without arrival
// counters the k_hub_merge_* kernels do it in a launch of their own.
template <int VPL, int LPH>
__device__ __forceinline__ void hub_merge_fwd_one(const FastArgs& A, int w, int P, int lane, float* __restrict__ out,
                                                  float* __restrict__ m_save, float* __restrict__ l_save,
                                                  float* __restrict__ agg_save) {
  constexpr int HD = 32 * VPL;
  const int c0 = lane * VPL, h = lane / LPH;
  const int v = A.hub_dst[w];
  const float* rec0 = A.hub_scratch + (int64_t)w * P * (HD + kHubRec);
  float M = -INFINITY;
#pragma unroll 8
  for (int p = 0; p < P; ++p) M = fmaxf(M, __ldcg(rec0 + (int64_t)p * (HD + kHubRec) + HD + lane));
  float acc[VPL], l = 0.f;
#pragma unroll
  for (int i = 0; i < VPL; ++i) acc[i] = 0.f;
#pragma unroll 4                         // four slices' loads in flight: one warp per hub, pure latency
  for (int p = 0; p < P; ++p) {
    const float* rec = rec0 + (int64_t)p * (HD + kHubRec);
    const float mp = __ldcg(rec + HD + lane);
    const float f = mp == -INFINITY ? 0.f : __expf(mp - M);  // empty slice
    l = fmaf(__ldcg(rec + HD + 32 + lane), f, l);
    float pa[VPL];
    ldrow_cg<VPL>(pa, rec + c0);
#pragma unroll
    for (int i = 0; i < VPL; ++i) acc[i] = fmaf(pa[i], f, acc[i]);
  }
  fwd_finalize<VPL>(A, v, true, acc, M, l, c0, h, (lane % LPH) == 0, out, m_save, l_save, agg_save);
}
template <int VPL>
__device__ __forceinline__ void hub_merge_bwd_one(const FastArgs& A, int w, int P, int lane, float* __restrict__ gPj) {
  constexpr int HD = 32 * VPL;
  const int c0 = lane * VPL;
  float acc[VPL];
#pragma unroll
  for (int i = 0; i < VPL; ++i) acc[i] = 0.f;
#pragma unroll 4
  for (int p = 0; p < P; ++p) {
    float pa[VPL];
    ldrow_cg<VPL>(pa, A.hub_scratch + ((int64_t)w * P + p) * (HD + kHubRec) + c0);
#pragma unroll
    for (int i = 0; i < VPL; ++i) acc[i] += pa[i];
  }
  strow<VPL>(gPj + (int64_t)A.hub_dst[w] * HD + c0, acc);
}
// Called by warp 0 of a hub CTA after it has parked the partial of one slice of hub `w`: true (for the whole warp) in
// the warp whose slice arrived last.  The counter is back at zero when the kernel ends.
__device__ __forceinline__ bool hub_arrive(const FastArgs& A, int w, int P, int lane) {
  __threadfence();                       // the partial is visible device-wide before the arrival is
  __syncwarp();
  int last = 0;
  if (lane == 0) {
    last = atomicAdd(A.hub_cnt + w, 1) == P - 1;
    if (last) A.hub_cnt[w] = 0;
  }
  last = __shfl_sync(0xffffffffu, last, 0);
  if (last) __threadfence();             // the other slices' partials are read after their arrivals were seen
  return last != 0;
}

// Template dispatch on (VPL, LPH, ALIAS).
#define FAST_DISPATCH_LPH(VPL_, lph, ALIAS_, CALL)        \
  switch (lph) {                                          \
    case 4: { constexpr int VPL = VPL_, LPH = 4; constexpr bool ALIAS = ALIAS_; CALL; } break;   \
    case 8: { constexpr int VPL = VPL_, LPH = 8; constexpr bool ALIAS = ALIAS_; CALL; } break;   \
    case 16: { constexpr int VPL = VPL_, LPH = 16; constexpr bool ALIAS = ALIAS_; CALL; } break; \
    default: { constexpr int VPL = VPL_, LPH = 32; constexpr bool ALIAS = ALIAS_; CALL; } break; \
  }
#define FAST_DISPATCH_VPL(vpl, lph, ALIAS_, CALL)                   \
  switch (vpl) {                                                    \
    case 2: FAST_DISPATCH_LPH(2, lph, ALIAS_, CALL) break;          \
    case 4: FAST_DISPATCH_LPH(4, lph, ALIAS_, CALL) break;          \
    default: FAST_DISPATCH_LPH(8, lph, ALIAS_, CALL) break;         \
  }
#define FAST_DISPATCH(vpl, lph, alias, CALL)                        \
  if (alias) { FAST_DISPATCH_VPL(vpl, lph, true, CALL) } else { FAST_DISPATCH_VPL(vpl, lph, false, CALL) }


// Windowed variants (gat_win.cu): a CTA stages the Pi / V (or Pj / G) rows of a run of consecutive nodes and the
// run's adjacency lists in shared memory; hub nodes are left to the hub CTAs of the warp-per-node kernels.
bool win_enabled(const FastArgs& A);
int win_fwd(const FastArgs& A, int vpl, int lph, bool alias, const float* Pi, const float* Pj, const float* V,
            float* out, float* s_save, float* m_save, float* l_save, float* agg_save, cudaStream_t st);
int win_bwd_dst(const FastArgs& A, int vpl, int lph, bool alias, const float* Pi, const float* Pj, const float* V,
                const float* s_save, const float* m_save, const float* l_save, const float* agg_save,
                const float* G, float* gPj, float* g_attn, float* g_table, float* g_affine, float* gs_out,
                float* alpha_out, cudaStream_t st);
int win_bwd_src(const FastArgs& A, int vpl, int lph, bool alias, const float* Pi, const float* Pj, const float* G,
                const float* gs_in, const float* alpha_in, float* gPi, float* gV, cudaStream_t st);

}  // namespace b200gat
