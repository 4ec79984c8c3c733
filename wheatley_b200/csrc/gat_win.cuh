// Windowed variants of the warp-per-node attention kernels (gat_fast.cu) for LOW-DEGREE, BANDED graphs: the residual
// of a clique view (self loops, job-chain precedence u = v +- 1, pool star) and the small PSP graphs.
//
// The warp-per-node kernels are latency-bound there (ncu r2s: long-scoreboard 8 stalls per issue, issue-active 27 %):
// a warp walks order[w] -> in_ptr -> in_src / in_pos -> etype -> rows, five dependent global round trips for ~4 edges,
// and only ~3 nodes per warp to overlap them with.  Here a CTA takes a run of W consecutive nodes [w0, w0 + W):
//
//   rows    : the run's Pi / V rows (backward by source: Pj / G) are one contiguous block of W * HD floats each:
//             ONE bulk async copy (cp.async.bulk, mbarrier complete_tx) per tensor, issued by one thread before
//             anything else
//   indices : in_ptr[w0 .. w0 + W] is one coalesced load; the run's adjacency lists are the contiguous range
//             [in_ptr[w0], in_ptr[w0 + W]) of in_src / in_pos: one edge per thread, coalesced, with the edge type
//             gathered by the same thread -> three round trips per RUN instead of five per node
//   math    : warp per node as before, same edge order (the online softmax rescales every 1-2 edges instead of
//             every 4: same result up to rounding); a source row comes from the window when w0 <= u < w0 + W
//             (self loops and job neighbours) and from global memory otherwise (pool node, run boundaries)
//
// Hub nodes (degree > hub_degree) are skipped by the runs: the first hub_blocks CTAs of each grid are the hub CTAs of
// gat_fast.cu (fwd_fast_cta / bwd_dst_fast_cta / bwd_src_fast_cta), so the hubs overlap with the runs instead of
// costing a launch of their own (measured as separate launches: +19 us per kernel at 8 x Taillard 100x20).
// Included at the end of gat_fast.cu (one translation unit: the hub CTAs are the *_fast_cta functions defined there).
#include <stdlib.h>
#include "gat_rows.cuh"

namespace b200gat {

namespace {

#ifndef B200GAT_WIN_W
#define B200GAT_WIN_W 16
#endif
constexpr int kWinW = 16;                 // nodes per run, backward by destination (its accumulators need the room)
constexpr int kWinWide = B200GAT_WIN_W;   // nodes per run, forward and backward by source (32 was measured: 3-8 % slower)
constexpr int kWinEdges = 768;            // staged adjacency entries per run (longer runs read them from global)

__device__ __forceinline__ uint32_t w_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void w_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void w_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol error traps (the launch fails) instead of hanging the GPU.
__device__ __forceinline__ void w_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0, spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (++spins > 20000000u) __trap();
  }
}
// 1-D bulk copy global -> shared (16-byte aligned addresses, size a multiple of 16), completion on the mbarrier.
__device__ __forceinline__ void w_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void w_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// The adjacency of one run, staged in shared memory.
template <int W>
struct WinIdxT {
  int ptr[W + 1];
  int nbr[kWinEdges * (W / 16)], pos[kWinEdges * (W / 16)], typ[kWinEdges * (W / 16)];
};

// Start of a run: thread 0 queues the row copies, the first threads fetch the run's segment bounds.
template <int W>
__device__ __forceinline__ void win_begin(uint32_t bar, int w0, int nw, int HD, float* w_a, const float* ga,
                                          float* w_b, const float* gb, WinIdxT<W>& ix,
                                          const int32_t* __restrict__ ptr) {
  if (threadIdx.x == 0) {
    const uint32_t bytes = (uint32_t)nw * HD * 4u;
    w_fence_async();                 // the previous run's reads of the window precede these asynchronous writes
    w_mbar_expect_tx(bar, w_b ? 2 * bytes : bytes);
    w_bulk_g2s(w_smem_u32(w_a), ga + (int64_t)w0 * HD, bytes, bar);
    if (w_b) w_bulk_g2s(w_smem_u32(w_b), gb + (int64_t)w0 * HD, bytes, bar);
  }
  if ((int)threadIdx.x <= nw) ix.ptr[threadIdx.x] = ptr[w0 + threadIdx.x];
}
// After a __syncthreads: every thread stages one adjacency entry per 256 (neighbour, edge position, edge type).
template <int W>
__device__ __forceinline__ bool win_stage_idx(const FastArgs& A, WinIdxT<W>& ix, int nw, const int32_t* __restrict__ nbr,
                                              const int32_t* __restrict__ pos, bool typed) {
  const int eb = ix.ptr[0], ne = ix.ptr[nw] - eb;
  if (ne > kWinEdges * (W / 16)) return false;
  for (int t = threadIdx.x; t < ne; t += kFastThreads) {
    const int p = pos ? pos[eb + t] : eb + t;
    ix.nbr[t] = nbr[eb + t];
    ix.pos[t] = p;
    ix.typ[t] = typed ? A.etype[p] : 0;
  }
  return true;
}
// One lane's adjacency entry of the chunk [e0, e0 + n) of a segment.
template <int W>
__device__ __forceinline__ void win_entry(const FastArgs& A, const WinIdxT<W>& ix, bool staged, int eb, int e0, int n,
                                          int lane, const int32_t* __restrict__ nbr,
                                          const int32_t* __restrict__ pos, bool typed, int& my_u, int& my_p,
                                          int& my_t) {
  my_u = 0; my_p = 0; my_t = 0;
  if (lane >= n) return;
  if (staged) {
    my_u = ix.nbr[e0 - eb + lane]; my_p = ix.pos[e0 - eb + lane]; my_t = ix.typ[e0 - eb + lane];
  } else {
    my_u = nbr[e0 + lane];
    my_p = pos ? pos[e0 + lane] : e0 + lane;
    my_t = typed ? A.etype[my_p] : 0;
  }
}

// ------------------------------------------------------------------------------ forward
template <int VPL, int LPH, bool ALIAS, int OCC>
__global__ void __launch_bounds__(kFastThreads, OCC) k_fwd_win(
    FastArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ V,
    float* __restrict__ out, float* __restrict__ s_save, float* __restrict__ m_save,
    float* __restrict__ l_save, float* __restrict__ agg_save) {
  pdl_sync();
  constexpr int HD = 32 * VPL;
  constexpr int U = OCC >= 3 ? 1 : 2;     // edges in flight (register budget: 80 / thread at 3 CTAs per SM)
  extern __shared__ __align__(128) float w_rows[];
  __shared__ WinIdxT<kWinWide> ix;
  __shared__ __align__(8) unsigned long long s_bar;
  float* w_pi = w_rows;
  float* w_v = ALIAS ? w_rows : w_rows + kWinWide * HD;
  if ((int)blockIdx.x < A.hub_blocks) {      // hub CTAs (whole CTA: uniform branch)
    fwd_fast_cta<VPL, LPH, ALIAS>(A, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, (int)blockIdx.x, A.hub_blocks);
    return;
  }
  const uint32_t bar = w_smem_u32(&s_bar);
  if (threadIdx.x == 0) w_mbar_init(bar, 1);
  __syncthreads();
  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31, c0 = lane * VPL, h = lane / LPH;
  const bool head_lead = (lane % LPH) == 0;
  const bool typed = A.emode != B200GAT_EDGE_NONE;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  float at[VPL];
  ldrow<VPL>(at, A.attn + c0);
  uint32_t phase = 0;
  const int n_runs = (A.N + kWinWide - 1) / kWinWide;

  for (int run = blockIdx.x - A.hub_blocks; run < n_runs; run += gridDim.x - A.hub_blocks) {
    const int w0 = run * kWinWide, nw = min(kWinWide, A.N - w0);
    win_begin(bar, w0, nw, HD, w_pi, Pi, ALIAS ? nullptr : w_v, V, ix, A.in_ptr);
    constexpr int kPre = kWinWide / kWarps < 2 ? kWinWide / kWarps : 2;
    float qpre[kPre][VPL];                                   // first own rows: in flight under the index round trips
#pragma unroll
    for (int j = 0; j < kPre; ++j) {
      const int vl = wi + j * kWarps;
      if (vl < nw) ldrow<VPL>(qpre[j], Pj + (int64_t)(w0 + vl) * HD + c0);
    }
    __syncthreads();
    const bool staged = win_stage_idx(A, ix, nw, A.in_src, A.in_pos, typed);
    __syncthreads();
    w_mbar_wait(bar, phase);
    phase ^= 1u;
    const int eb = ix.ptr[0];

#pragma unroll
    for (int j = 0; j < (kWinWide / kWarps); ++j) {
      const int vl = wi + j * kWarps;
      if (vl >= nw) continue;
      const int v = w0 + vl;
      const int rb = ix.ptr[vl], re = ix.ptr[vl + 1];
      if (A.hub_blocks > 0 && re - rb > A.hub_degree) continue;             // a hub CTA has it
      float acc[VPL];
#pragma unroll
      for (int i = 0; i < VPL; ++i) acc[i] = 0.f;
      float m = -INFINITY, l = 0.f;
      if (re > rb) {
        float q[VPL], qt[VPL];
        if (j < kPre) {
#pragma unroll
          for (int i = 0; i < VPL; ++i) q[i] = qpre[j < kPre ? j : 0][i];
        } else {
          ldrow<VPL>(q, Pj + (int64_t)v * HD + c0);
        }
#pragma unroll
        for (int i = 0; i < VPL; ++i) qt[i] = q[i];
        int t_cur = -1;
        for (int e0 = rb; e0 < re; e0 += 32) {
          const int n = min(32, re - e0);
          int my_u, my_p, my_t;
          win_entry(A, ix, staged, eb, e0, n, lane, A.in_src, A.in_pos, typed, my_u, my_p, my_t);
          for (int j0 = 0; j0 < n; j0 += U) {
            float pi[U][VPL], vv[ALIAS ? 1 : U][VPL];
            int tt[U], pp[U];
            float aa[U];
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const int idx = min(j0 + k, n - 1);
              const int u = __shfl_sync(0xffffffffu, my_u, idx);
              tt[k] = __shfl_sync(0xffffffffu, my_t, idx);
              pp[k] = __shfl_sync(0xffffffffu, my_p, idx);
              aa[k] = affine ? affine_fetch(A, pp[k], lane) : 0.f;
              const int ul = u - w0;
              if ((unsigned)ul < (unsigned)nw) {
                ldrow_s<VPL>(pi[k], w_pi + ul * HD + c0);
                if constexpr (!ALIAS) ldrow_s<VPL>(vv[k], w_v + ul * HD + c0);
              } else {
                ldrow<VPL>(pi[k], Pi + (int64_t)u * HD + c0);
                if constexpr (!ALIAS) ldrow<VPL>(vv[k], V + (int64_t)u * HD + c0);
              }
            }
            float sc[U];
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const bool valid = j0 + k < n;
              if (typed && tt[k] != t_cur) {
                t_cur = tt[k];
                type_row<VPL>(A, t_cur, c0, q, qt);
              }
              float z[VPL];
#pragma unroll
              for (int i = 0; i < VPL; ++i) z[i] = pi[k][i] + qt[i];
              if (affine && valid) affine_add<VPL>(A, aa[k], c0, z);
              float part = 0.f;
#pragma unroll
              for (int i = 0; i < VPL; ++i) part = fmaf(at[i], leaky(z[i], A.slope), part);
              part = head_sum<LPH>(part);
              if (valid && head_lead) s_save[(int64_t)pp[k] * A.H + h] = part;
              sc[k] = valid ? part : -INFINITY;
            }
            float mnew = m;
#pragma unroll
            for (int k = 0; k < U; ++k) mnew = fmaxf(mnew, sc[k]);
            const float scale = __expf(m - mnew);       // m = -inf on the first chunk -> 0
            l *= scale;
#pragma unroll
            for (int i = 0; i < VPL; ++i) acc[i] *= scale;
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const float p = __expf(sc[k] - mnew);     // invalid slot: exp(-inf) = 0
              l += p;
#pragma unroll
              for (int i = 0; i < VPL; ++i) acc[i] = fmaf(p, ALIAS ? pi[k][i] : vv[ALIAS ? 0 : k][i], acc[i]);
            }
            m = mnew;
          }
        }
      }
      fwd_finalize<VPL>(A, v, re > rb, acc, m, l, c0, h, head_lead, out, m_save, l_save, agg_save);
    }
    __syncthreads();       // every read of the window and of the staged indices precedes the next run's writes
  }
}

// ------------------------------------------------------------------------------ backward, by destination
// Dynamic shared memory: the window rows, then sh_gat[HD] + the CTA's d table / d affine accumulators exactly as in
// k_bwd_dst_fast (tab_in_smem / aff_in_smem have the same meaning).
template <int VPL, int LPH, bool ALIAS>
__global__ void __launch_bounds__(kFastThreads, 2) k_bwd_dst_win(
    FastArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ V,
    const float* __restrict__ s_save, const float* __restrict__ m_save, const float* __restrict__ l_save,
    const float* __restrict__ agg_save, const float* __restrict__ G, float* __restrict__ gPj,
    float* __restrict__ g_attn, float* __restrict__ g_table, float* __restrict__ g_affine,
    float* __restrict__ gs_out, float* __restrict__ alpha_out, int tab_in_smem, int aff_in_smem) {
  pdl_sync();
  constexpr int HD = 32 * VPL;
  constexpr int U = 2;
  extern __shared__ __align__(128) float w_rows[];
  __shared__ WinIdxT<kWinW> ix;
  __shared__ __align__(8) unsigned long long s_bar;
  float* w_pi = w_rows;
  float* w_v = ALIAS ? w_rows : w_rows + kWinW * HD;
  float* sh = w_rows + (ALIAS ? 1 : 2) * kWinW * HD;
  float* sh_gat = sh;
  float* sh_tab = sh + HD;
  const bool typed = A.emode != B200GAT_EDGE_NONE;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  const int n_tab = tab_in_smem == 2 ? kWarps * A.T * HD : tab_in_smem == 1 ? A.T * HD : 0;
  float* sh_aff = sh + HD + n_tab;
  const int n_sh = HD + n_tab + (aff_in_smem ? kWarps * A.K * HD : 0);
  if ((int)blockIdx.x < A.hub_blocks) {      // hub CTAs: same accumulators, at the same place of the dynamic array
    bwd_dst_fast_cta<VPL, LPH, ALIAS>(A, Pi, Pj, V, s_save, m_save, l_save, agg_save, G, gPj, g_attn, g_table,
                                      g_affine, gs_out, alpha_out, tab_in_smem, aff_in_smem, sh, (int)blockIdx.x,
                                      A.hub_blocks);
    return;
  }
  for (int i = threadIdx.x; i < n_sh; i += kFastThreads) sh[i] = 0.f;
  const uint32_t bar = w_smem_u32(&s_bar);
  if (threadIdx.x == 0) w_mbar_init(bar, 1);
  __syncthreads();

  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31, c0 = lane * VPL, h = lane / LPH;
  const bool head_lead = (lane % LPH) == 0;
  float at[VPL], gat[VPL];
  ldrow<VPL>(at, A.attn + c0);
#pragma unroll
  for (int i = 0; i < VPL; ++i) gat[i] = 0.f;
  uint32_t phase = 0;
  const int n_runs = (A.N + kWinW - 1) / kWinW;

  for (int run = blockIdx.x - A.hub_blocks; run < n_runs; run += gridDim.x - A.hub_blocks) {
    const int w0 = run * kWinW, nw = min(kWinW, A.N - w0);
    win_begin(bar, w0, nw, HD, w_pi, Pi, ALIAS ? nullptr : w_v, V, ix, A.in_ptr);
    __syncthreads();
    const bool staged = win_stage_idx(A, ix, nw, A.in_src, A.in_pos, typed);
    __syncthreads();
    w_mbar_wait(bar, phase);
    phase ^= 1u;
    const int eb = ix.ptr[0];

    for (int j = 0; j < (kWinW / kWarps); ++j) {
      const int vl = wi + j * kWarps;
      if (vl >= nw) continue;
      const int v = w0 + vl;
      const int rb = ix.ptr[vl], re = ix.ptr[vl + 1];
      if (A.hub_blocks > 0 && re - rb > A.hub_degree) continue;             // a hub CTA has it
      float gpj[VPL];
#pragma unroll
      for (int i = 0; i < VPL; ++i) gpj[i] = 0.f;
      if (re > rb) {
        float q[VPL], qt[VPL], g[VPL], gtab[VPL];
        ldrow<VPL>(q, Pj + (int64_t)v * HD + c0);
        ldrow<VPL>(g, G + (int64_t)v * HD + c0);
        float dot = 0.f;
        {
          float ag[VPL];
          ldrow<VPL>(ag, agg_save + (int64_t)v * HD + c0);
#pragma unroll
          for (int i = 0; i < VPL; ++i) dot = fmaf(g[i], ag[i], dot);
          dot = head_sum<LPH>(dot);
        }
#pragma unroll
        for (int i = 0; i < VPL; ++i) { qt[i] = q[i]; gtab[i] = 0.f; }
        const float mh = m_save[(int64_t)v * A.H + h];
        const float linv = 1.f / (l_save[(int64_t)v * A.H + h] + A.eps);
        int t_cur = -1;
        auto flush_tab = [&]() {
          if (t_cur < 0) return;
          if (tab_in_smem == 2) {
            float* dst = sh_tab + ((wi * A.T + t_cur) * HD + c0);
            float old[VPL];
            ldrow_s<VPL>(old, dst);
#pragma unroll
            for (int i = 0; i < VPL; ++i) { old[i] += gtab[i]; gtab[i] = 0.f; }
            strow<VPL>(dst, old);
            return;
          }
          float* dst = tab_in_smem ? sh_tab + t_cur * HD + c0 : g_table + (int64_t)t_cur * HD + c0;
#pragma unroll
          for (int i = 0; i < VPL; ++i) {
            if (gtab[i] != 0.f) atomicAdd(dst + i, gtab[i]);
            gtab[i] = 0.f;
          }
        };
        for (int e0 = rb; e0 < re; e0 += 32) {
          const int n = min(32, re - e0);
          int my_u, my_p, my_t;
          win_entry(A, ix, staged, eb, e0, n, lane, A.in_src, A.in_pos, typed, my_u, my_p, my_t);
          for (int j0 = 0; j0 < n; j0 += U) {
            float pi[U][VPL], vv[ALIAS ? 1 : U][VPL], sv[U];
            int tt[U], pp[U];
            float aa[U];
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const int idx = min(j0 + k, n - 1);
              const int u = __shfl_sync(0xffffffffu, my_u, idx);
              tt[k] = __shfl_sync(0xffffffffu, my_t, idx);
              pp[k] = __shfl_sync(0xffffffffu, my_p, idx);
              aa[k] = affine ? affine_fetch(A, pp[k], lane) : 0.f;
              const int ul = u - w0;
              if ((unsigned)ul < (unsigned)nw) {
                ldrow_s<VPL>(pi[k], w_pi + ul * HD + c0);
                if constexpr (!ALIAS) ldrow_s<VPL>(vv[k], w_v + ul * HD + c0);
              } else {
                ldrow<VPL>(pi[k], Pi + (int64_t)u * HD + c0);
                if constexpr (!ALIAS) ldrow<VPL>(vv[k], V + (int64_t)u * HD + c0);
              }
              sv[k] = __ldg(s_save + (int64_t)pp[k] * A.H + h);
            }
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const bool valid = j0 + k < n;
              const int64_t pos = pp[k];
              float ga = 0.f;
#pragma unroll
              for (int i = 0; i < VPL; ++i) ga = fmaf(g[i], ALIAS ? pi[k][i] : vv[ALIAS ? 0 : k][i], ga);
              ga = head_sum<LPH>(ga);
              const float alpha = __expf(sv[k] - mh) * linv;
              float gsv = alpha * (ga - dot);
              if (valid && head_lead) {
                gs_out[pos * A.H + h] = gsv;
                alpha_out[pos * A.H + h] = alpha;
              }
              if (!valid) gsv = 0.f;
              if (typed && tt[k] != t_cur) {
                flush_tab();
                t_cur = tt[k];
                type_row<VPL>(A, t_cur, c0, q, qt);
              }
              float z[VPL];
#pragma unroll
              for (int i = 0; i < VPL; ++i) z[i] = pi[k][i] + qt[i];
              if (affine && valid) affine_add<VPL>(A, aa[k], c0, z);
              float gz[VPL];
#pragma unroll
              for (int i = 0; i < VPL; ++i) {
                const bool posz = z[i] > 0.f;
                gat[i] = fmaf(gsv, posz ? z[i] : z[i] * A.slope, gat[i]);
                gz[i] = gsv * (posz ? at[i] : at[i] * A.slope);
                gpj[i] += gz[i];
                gtab[i] += gz[i];
              }
              if (affine && valid) {
                for (int kk = 0; kk < A.K; ++kk) {
                  const float av = __shfl_sync(0xffffffffu, aa[k], kk);
                  if (av != 0.f) {
                    if (aff_in_smem) {
                      float* dst = sh_aff + ((wi * A.K + kk) * HD + c0);
                      float old[VPL];
                      ldrow_s<VPL>(old, dst);
#pragma unroll
                      for (int i = 0; i < VPL; ++i) old[i] = fmaf(gz[i], av, old[i]);
                      strow<VPL>(dst, old);
                    } else {
#pragma unroll
                      for (int i = 0; i < VPL; ++i) atomicAdd(g_affine + (int64_t)kk * HD + c0 + i, gz[i] * av);
                    }
                  }
                }
              }
            }
          }
        }
        if (typed) flush_tab();
      }
      strow<VPL>(gPj + (int64_t)v * HD + c0, gpj);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < VPL; ++i)
    if (gat[i] != 0.f) atomicAdd(sh_gat + c0 + i, gat[i]);
  __syncthreads();
  for (int i = threadIdx.x; i < HD; i += kFastThreads)
    if (sh_gat[i] != 0.f) atomicAdd(g_attn + i, sh_gat[i]);
  if (tab_in_smem == 2) {
    for (int i = threadIdx.x; i < A.T * HD; i += kFastThreads) {
      float sum = 0.f;
#pragma unroll
      for (int w2 = 0; w2 < kWarps; ++w2) sum += sh_tab[w2 * A.T * HD + i];
      if (sum != 0.f) atomicAdd(g_table + i, sum);
    }
  } else if (tab_in_smem) {
    for (int i = threadIdx.x; i < A.T * HD; i += kFastThreads)
      if (sh_tab[i] != 0.f) atomicAdd(g_table + i, sh_tab[i]);
  }
  if (aff_in_smem) {
    for (int i = threadIdx.x; i < A.K * HD; i += kFastThreads) {
      float sum = 0.f;
#pragma unroll
      for (int w2 = 0; w2 < kWarps; ++w2) sum += sh_aff[w2 * A.K * HD + i];
      if (sum != 0.f) atomicAdd(g_affine + i, sum);
    }
  }
}

// ------------------------------------------------------------------------------ backward, by source
// The window holds the Pj and G rows of the run (the out-neighbours of a banded graph); u's own Pi row is loaded
// directly.
template <int VPL, int LPH, bool ALIAS, int OCC>
__global__ void __launch_bounds__(kFastThreads, OCC) k_bwd_src_win(
    FastArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ G,
    const float* __restrict__ gs_in, const float* __restrict__ alpha_in, float* __restrict__ gPi,
    float* __restrict__ gV) {
  pdl_sync();
  constexpr int HD = 32 * VPL;
  constexpr int U = OCC >= 3 ? 1 : 2;
  extern __shared__ __align__(128) float w_rows[];
  __shared__ WinIdxT<kWinWide> ix;
  __shared__ __align__(8) unsigned long long s_bar;
  float* w_pj = w_rows;
  float* w_g = w_rows + kWinWide * HD;
  if ((int)blockIdx.x < A.hub_blocks) {      // hub CTAs
    bwd_src_fast_cta<VPL, LPH, ALIAS>(A, Pi, Pj, G, gs_in, alpha_in, gPi, gV, (int)blockIdx.x, A.hub_blocks);
    return;
  }
  const uint32_t bar = w_smem_u32(&s_bar);
  if (threadIdx.x == 0) w_mbar_init(bar, 1);
  __syncthreads();
  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31, c0 = lane * VPL, h = lane / LPH;
  const bool typed = A.emode != B200GAT_EDGE_NONE;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  float at[VPL], ats[VPL];
  ldrow<VPL>(at, A.attn + c0);
#pragma unroll
  for (int i = 0; i < VPL; ++i) ats[i] = at[i] * A.slope;
  uint32_t phase = 0;
  const int n_runs = (A.N + kWinWide - 1) / kWinWide;

  for (int run = blockIdx.x - A.hub_blocks; run < n_runs; run += gridDim.x - A.hub_blocks) {
    const int w0 = run * kWinWide, nw = min(kWinWide, A.N - w0);
    win_begin(bar, w0, nw, HD, w_pj, Pj, w_g, G, ix, A.out_ptr);
    constexpr int kPre = kWinWide / kWarps < 2 ? kWinWide / kWarps : 2;
    float ppre[kPre][VPL];
#pragma unroll
    for (int j = 0; j < kPre; ++j) {
      const int ul = wi + j * kWarps;
      if (ul < nw) ldrow<VPL>(ppre[j], Pi + (int64_t)(w0 + ul) * HD + c0);
    }
    __syncthreads();
    const bool staged = win_stage_idx(A, ix, nw, A.out_dst, A.out_pos, typed);
    __syncthreads();
    w_mbar_wait(bar, phase);
    phase ^= 1u;
    const int eb = ix.ptr[0];

#pragma unroll
    for (int j = 0; j < (kWinWide / kWarps); ++j) {
      const int ul = wi + j * kWarps;
      if (ul >= nw) continue;
      const int u = w0 + ul;
      const int rb = ix.ptr[ul], re = ix.ptr[ul + 1];
      if (A.hub_blocks > 0 && re - rb > A.hub_degree) continue;             // a hub CTA has it
      float gpi[VPL], gv[VPL];
#pragma unroll
      for (int i = 0; i < VPL; ++i) { gpi[i] = 0.f; gv[i] = 0.f; }
      if (re > rb) {
        float pr[VPL], pt[VPL];
        if (j < kPre) {
#pragma unroll
          for (int i = 0; i < VPL; ++i) pr[i] = ppre[j < kPre ? j : 0][i];
        } else {
          ldrow<VPL>(pr, Pi + (int64_t)u * HD + c0);
        }
#pragma unroll
        for (int i = 0; i < VPL; ++i) pt[i] = pr[i];
        int t_cur = -1;
        for (int q0 = rb; q0 < re; q0 += 32) {
          const int n = min(32, re - q0);
          int my_v, my_pos, my_t;
          win_entry(A, ix, staged, eb, q0, n, lane, A.out_dst, A.out_pos, typed, my_v, my_pos, my_t);
          for (int j0 = 0; j0 < n; j0 += U) {
            float pj[U][VPL], gg[U][VPL], gsv[U], al[U];
            int tt[U], pp[U];
            float aa[U];
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const int idx = min(j0 + k, n - 1);
              const int v = __shfl_sync(0xffffffffu, my_v, idx);
              pp[k] = __shfl_sync(0xffffffffu, my_pos, idx);
              aa[k] = affine ? affine_fetch(A, pp[k], lane) : 0.f;
              tt[k] = __shfl_sync(0xffffffffu, my_t, idx);
              const int vl = v - w0;
              if ((unsigned)vl < (unsigned)nw) {
                ldrow_s<VPL>(pj[k], w_pj + vl * HD + c0);
                ldrow_s<VPL>(gg[k], w_g + vl * HD + c0);
              } else {
                ldrow<VPL>(pj[k], Pj + (int64_t)v * HD + c0);
                ldrow<VPL>(gg[k], G + (int64_t)v * HD + c0);
              }
              const bool valid = j0 + k < n;
              gsv[k] = valid ? __ldg(gs_in + (int64_t)pp[k] * A.H + h) : 0.f;
              al[k] = valid ? __ldg(alpha_in + (int64_t)pp[k] * A.H + h) : 0.f;
            }
#pragma unroll
            for (int k = 0; k < U; ++k) {
              const bool valid = j0 + k < n;
              if (typed && tt[k] != t_cur) {
                t_cur = tt[k];
                type_row<VPL>(A, t_cur, c0, pr, pt);
              }
              float z[VPL];
#pragma unroll
              for (int i = 0; i < VPL; ++i) z[i] = pj[k][i] + pt[i];
              if (affine && valid) affine_add<VPL>(A, aa[k], c0, z);
#pragma unroll
              for (int i = 0; i < VPL; ++i) {
                gpi[i] = fmaf(gsv[k], z[i] > 0.f ? at[i] : ats[i], gpi[i]);
                gv[i] = fmaf(al[k], gg[k][i], gv[i]);
              }
            }
          }
        }
      }
      if constexpr (ALIAS) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) gpi[i] += gv[i];
        strow<VPL>(gPi + (int64_t)u * HD + c0, gpi);
      } else {
        strow<VPL>(gPi + (int64_t)u * HD + c0, gpi);
        strow<VPL>(gV + (int64_t)u * HD + c0, gv);
      }
    }
    __syncthreads();
  }
}

unsigned win_ctas(int64_t N, int per_sm, int W) {
  const int64_t runs = ceil_div(N, W), cap = (int64_t)per_sm * kNumSMs;
  return (unsigned)(runs < cap ? runs : cap);
}

template <auto Kernel>
int win_set_smem(size_t bytes) {
  if (bytes <= 16 * 1024) return B200GAT_OK;      // static (indices, hub merge buffers) + dynamic within 48 KB
  cudaError_t e = cudaFuncSetAttribute(Kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e != cudaSuccess) {
    set_error("gat_win: cudaFuncSetAttribute(%zu bytes): %s", bytes, cudaGetErrorString(e));
    return B200GAT_ERR_CUDA;
  }
  return B200GAT_OK;
}

// Resident CTAs per SM the forward / by-source kernels are built for: 2 (128 registers, two edges in flight) or
// 3 (80 registers, one edge in flight).  B200GAT_WIN_OCC=2|3 overrides.
int win_occ(int vpl) {
  static const int forced = [] {
    const char* e = getenv("B200GAT_WIN_OCC");
    return e ? atoi(e) : 0;
  }();
  if (forced == 2 || forced == 3) return forced;
  return vpl <= 4 ? 3 : 2;      // H*D <= 128: 80 registers hold a row pair without spills (psp272: 2.92 vs 3.00 ms)
}

}  // namespace

// Graphs flagged B200GAT_GRAPH_LOW_DEGREE by the plan (short adjacency lists everywhere except the hubs): the per-node
// latency chain, not the row traffic, bounds the warp-per-node kernels there.  A long non-hub list would be walked by
// one warp two edges at a time (measured on the PSP node-conflict graphs, ~100 in-edges per resource node: 1.4x
// slower than the warp-per-node kernels), hence a flag from the host and not a mean-degree test here.
// B200GAT_WIN=0 keeps the warp-per-node kernels; B200GAT_WIN=1 forces the windowed ones on every graph.
bool win_enabled(const FastArgs& A) {
  static const int mode = [] {
    const char* e = getenv("B200GAT_WIN");
    return e ? atoi(e) : -1;
  }();
  if (mode == 0) return false;
  if (mode == 1) return true;
  return (A.flags & B200GAT_GRAPH_LOW_DEGREE) != 0;
}

int win_fwd(const FastArgs& A, int vpl, int lph, bool alias, const float* Pi, const float* Pj, const float* V,
            float* out, float* s_save, float* m_save, float* l_save, float* agg_save, cudaStream_t st) {
  const size_t smem = sizeof(float) * (size_t)(alias ? 1 : 2) * kWinWide * 32 * vpl;
  const int occ = win_occ(vpl);
  const unsigned blocks = win_ctas(A.N, occ, kWinWide) + A.hub_blocks;
  int rc = B200GAT_OK;
  if (occ == 3) {
    FAST_DISPATCH(vpl, lph, alias, (rc = win_set_smem<k_fwd_win<VPL, LPH, ALIAS, 3>>(smem)));
    if (rc) return rc;
    FAST_DISPATCH(vpl, lph, alias,
                  (launch_k(k_fwd_win<VPL, LPH, ALIAS, 3>, dim3(blocks), dim3(kFastThreads), smem, st, A, Pi, Pj, V, out,
                            s_save, m_save, l_save, agg_save)));
  } else {
    FAST_DISPATCH(vpl, lph, alias, (rc = win_set_smem<k_fwd_win<VPL, LPH, ALIAS, 2>>(smem)));
    if (rc) return rc;
    FAST_DISPATCH(vpl, lph, alias,
                  (launch_k(k_fwd_win<VPL, LPH, ALIAS, 2>, dim3(blocks), dim3(kFastThreads), smem, st, A, Pi, Pj, V, out,
                            s_save, m_save, l_save, agg_save)));
  }
  B200_CHECK_LAUNCH("k_fwd_win");
  return B200GAT_OK;
}

int win_bwd_dst(const FastArgs& A, int vpl, int lph, bool alias, const float* Pi, const float* Pj, const float* V,
                const float* s_save, const float* m_save, const float* l_save, const float* agg_save,
                const float* G, float* gPj, float* g_attn, float* g_table, float* g_affine, float* gs_out,
                float* alpha_out, cudaStream_t st) {
  const int HD = 32 * vpl;
  const int64_t tab_bytes = (int64_t)A.T * HD * 4;
  const int tab_in_smem = A.emode == B200GAT_EDGE_NONE ? 0 : kWarps * tab_bytes <= 64 * 1024 ? 2
                          : tab_bytes <= 40 * 1024 ? 1 : 0;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  const int aff_in_smem = affine && (int64_t)kWarps * A.K * HD * 4 <= 64 * 1024 ? 1 : 0;
  const size_t smem = sizeof(float) * ((size_t)(alias ? 1 : 2) * kWinW * HD + HD +
                                       (tab_in_smem == 2 ? (size_t)kWarps * A.T * HD
                                        : tab_in_smem == 1 ? (size_t)A.T * HD : 0) +
                                       (aff_in_smem ? (size_t)kWarps * A.K * HD : 0));
  const unsigned blocks = win_ctas(A.N, 2, kWinW) + A.hub_blocks;
  int rc = B200GAT_OK;
  FAST_DISPATCH(vpl, lph, alias, (rc = win_set_smem<k_bwd_dst_win<VPL, LPH, ALIAS>>(smem)));
  if (rc) return rc;
  FAST_DISPATCH(vpl, lph, alias,
                (launch_k(k_bwd_dst_win<VPL, LPH, ALIAS>, dim3(blocks), dim3(kFastThreads), smem, st, A, Pi, Pj, V, s_save,
                          m_save, l_save, agg_save, G, gPj, g_attn, g_table, g_affine, gs_out, alpha_out, tab_in_smem,
                          aff_in_smem)));
  B200_CHECK_LAUNCH("k_bwd_dst_win");
  return B200GAT_OK;
}

int win_bwd_src(const FastArgs& A, int vpl, int lph, bool alias, const float* Pi, const float* Pj, const float* G,
                const float* gs_in, const float* alpha_in, float* gPi, float* gV, cudaStream_t st) {
  const size_t smem = sizeof(float) * (size_t)2 * kWinWide * 32 * vpl;
  const int occ = win_occ(vpl);
  const unsigned blocks = win_ctas(A.N, occ, kWinWide) + A.hub_blocks;
  int rc = B200GAT_OK;
  if (occ == 3) {
    FAST_DISPATCH(vpl, lph, alias, (rc = win_set_smem<k_bwd_src_win<VPL, LPH, ALIAS, 3>>(smem)));
    if (rc) return rc;
    FAST_DISPATCH(vpl, lph, alias,
                  (launch_k(k_bwd_src_win<VPL, LPH, ALIAS, 3>, dim3(blocks), dim3(kFastThreads), smem, st, A, Pi, Pj, G,
                            gs_in, alpha_in, gPi, gV)));
  } else {
    FAST_DISPATCH(vpl, lph, alias, (rc = win_set_smem<k_bwd_src_win<VPL, LPH, ALIAS, 2>>(smem)));
    if (rc) return rc;
    FAST_DISPATCH(vpl, lph, alias,
                  (launch_k(k_bwd_src_win<VPL, LPH, ALIAS, 2>, dim3(blocks), dim3(kFastThreads), smem, st, A, Pi, Pj, G,
                            gs_in, alpha_in, gPi, gV)));
  }
  B200_CHECK_LAUNCH("k_bwd_src_win");
  return B200GAT_OK;
}

}  // namespace b200gat
