// Specialised fused graph-attention kernels for the additive flavour (EGAT / GATv2) with
// H*D in {32,64,128,256} and factored (none / table / affine) edge features - the shapes
// Wheatley runs (H=4, D=64 default; D=32 in 272.sh).
//
// Work decomposition: ONE WARP PER NODE, ALL HEADS.  Lane l owns the VPL = H*D/32 contiguous
// features [l*VPL, (l+1)*VPL) of a row, so a row gather is one or two 16-byte loads per lane
// (a full 128-byte line per 4-8 lanes, fully coalesced) and a head is LPH = D/VPL adjacent
// lanes (a log2(LPH)-step xor-shuffle reduction instead of a 5-step one per head).
// Nodes are visited in the plan's locality order (b200gat_node_order) so the warps of a CTA
// work on destinations of the same machine clique and re-use each other's gathered rows in L1.
//
//   forward      : single pass over the in-edges with an online (running max / rescale) softmax,
//                  UNROLL edges in flight; V aliasing Pi (GATv2) is gathered once.
//   backward dst : per destination: d logit from <G[v], V[u]> and <G[v], agg[v]> (one pass, the
//                  softmax Jacobian's dot product comes from the saved aggregate), d attn,
//                  d Pj[v], d edge table (accumulated per type run in registers -> shared -> global)
//   backward src : per source: d Pi[u], d V[u] over the out-edges (CSR by source), deterministic.
#include <math.h>
#include "common.cuh"
#include "gat_kernels.cuh"
#include "gat_rows.cuh"

namespace b200gat {

// ------------------------------------------------------------------------------ forward
// The work of CTA `block` of a grid of `nblocks` (the kernel below; the windowed kernels of gat_win.cuh run the hub
// CTAs of this function at the front of their own grids: block < hub_blocks == nblocks).
template <int VPL, int LPH, bool ALIAS>
__device__ __forceinline__ void fwd_fast_cta(
    const FastArgs& A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ V,
    float* __restrict__ out, float* __restrict__ s_save, float* __restrict__ m_save,
    float* __restrict__ l_save, float* __restrict__ agg_save, int block, int nblocks) {
  constexpr int HD = 32 * VPL;
  __shared__ float sm_acc[kWarps][HD];
  __shared__ float sm_m[kWarps][32], sm_l[kWarps][32];
  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31, c0 = lane * VPL, h = lane / LPH;
  const bool head_lead = (lane % LPH) == 0;
  const bool hub_block = block < A.hub_blocks;
  const bool typed = A.emode != B200GAT_EDGE_NONE;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  const int n_hub = hub_block ? *A.n_hub_dst : 0;
  const int P = hub_block ? hub_parts_of(A, n_hub) : 1;
  const int n_items = hub_block ? n_hub * P : (A.N + kWarps - 1) / kWarps;      // node CTAs: groups of kWarps nodes
  const int item0 = hub_block ? block : block - A.hub_blocks;
  const int istep = hub_block ? A.hub_blocks : nblocks - A.hub_blocks;
  float at[VPL];
  ldrow<VPL>(at, A.attn + c0);

  for (int item = item0; item < n_items; item += istep) {
    const WorkItem it = hub_block ? hub_item(A.in_ptr, A.hub_dst[item / P], wi, item % P, P)
                                  : node_item(A, A.in_ptr, A.dst_order, wi, item);
    const int v = it.node;
    float acc[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) acc[i] = 0.f;
    float m = -INFINITY, l = 0.f;
    if (it.active && it.re > it.rb) {
      float q[VPL], qt[VPL];
      ldrow<VPL>(q, Pj + (int64_t)v * HD + c0);
#pragma unroll
      for (int i = 0; i < VPL; ++i) qt[i] = q[i];
      int t_cur = -1;
      for (int e0 = it.rb; e0 < it.re; e0 += 32) {
        const int n = min(32, it.re - e0);
        const int my_u = lane < n ? A.in_src[e0 + lane] : 0;
        const int my_p = lane < n ? (A.in_pos ? A.in_pos[e0 + lane] : e0 + lane) : 0;
        const int my_t = (typed && lane < n) ? A.etype[my_p] : 0;
        for (int j0 = 0; j0 < n; j0 += kUnroll) {
          float pi[kUnroll][VPL], vv[ALIAS ? 1 : kUnroll][VPL];
          int tt[kUnroll], pp[kUnroll];
          float aa[kUnroll];
#pragma unroll
          for (int k = 0; k < kUnroll; ++k) {
            const int idx = min(j0 + k, n - 1);
            const int u = __shfl_sync(0xffffffffu, my_u, idx);
            tt[k] = __shfl_sync(0xffffffffu, my_t, idx);
            pp[k] = __shfl_sync(0xffffffffu, my_p, idx);
            aa[k] = affine ? affine_fetch(A, pp[k], lane) : 0.f;
            ldrow<VPL>(pi[k], Pi + (int64_t)u * HD + c0);
            if constexpr (!ALIAS) ldrow<VPL>(vv[k], V + (int64_t)u * HD + c0);
          }
          float sc[kUnroll];
#pragma unroll
          for (int k = 0; k < kUnroll; ++k) {
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
          for (int k = 0; k < kUnroll; ++k) mnew = fmaxf(mnew, sc[k]);
          const float scale = __expf(m - mnew);       // m = -inf on the first chunk -> 0
          l *= scale;
#pragma unroll
          for (int i = 0; i < VPL; ++i) acc[i] *= scale;
#pragma unroll
          for (int k = 0; k < kUnroll; ++k) {
            const float p = __expf(sc[k] - mnew);     // invalid slot: exp(-inf) = 0
            l += p;
#pragma unroll
            for (int i = 0; i < VPL; ++i) acc[i] = fmaf(p, ALIAS ? pi[k][i] : vv[ALIAS ? 0 : k][i], acc[i]);
          }
          m = mnew;
        }
      }
    }
    bool finalize = it.active;
    if (hub_block) {                       // merge the warps' partial softmax states in warp order
      strow<VPL>(&sm_acc[wi][c0], acc);
      sm_m[wi][lane] = m;
      sm_l[wi][lane] = l;
      __syncthreads();
      finalize = wi == 0;
      if (finalize) {
        float M = -INFINITY;
#pragma unroll
        for (int w2 = 0; w2 < kWarps; ++w2) M = fmaxf(M, sm_m[w2][lane]);
        l = 0.f;
#pragma unroll
        for (int i = 0; i < VPL; ++i) acc[i] = 0.f;
        for (int w2 = 0; w2 < kWarps; ++w2) {
          const float f = sm_m[w2][lane] == -INFINITY ? 0.f : __expf(sm_m[w2][lane] - M);   // empty slice
          l = fmaf(sm_l[w2][lane], f, l);
          float pa[VPL];
#pragma unroll
          for (int i = 0; i < VPL; ++i) pa[i] = sm_acc[w2][c0 + i];
#pragma unroll
          for (int i = 0; i < VPL; ++i) acc[i] = fmaf(pa[i], f, acc[i]);
        }
        m = M;
      }
      __syncthreads();
      if (finalize && P > 1) {             // one slice of a split hub: park the partial state
        float* rec = A.hub_scratch + (int64_t)item * (HD + kHubRec);
        strow<VPL>(rec + c0, acc);
        rec[HD + lane] = m;
        rec[HD + 32 + lane] = l;
        finalize = false;
        if (A.hub_cnt && hub_arrive(A, item / P, P, lane))
          hub_merge_fwd_one<VPL, LPH>(A, item / P, P, lane, out, m_save, l_save, agg_save);
      }
    }
    if (finalize)
      fwd_finalize<VPL>(A, v, it.seg_e > it.seg_b, acc, m, l, c0, h, head_lead, out, m_save, l_save, agg_save);
  }
}

template <int VPL, int LPH, bool ALIAS>
__global__ void __launch_bounds__(kFastThreads, B200GAT_FAST_MINB) k_fwd_fast(
    FastArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ V,
    float* __restrict__ out, float* __restrict__ s_save, float* __restrict__ m_save,
    float* __restrict__ l_save, float* __restrict__ agg_save) {
  fwd_fast_cta<VPL, LPH, ALIAS>(A, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, (int)blockIdx.x, (int)gridDim.x);
}

// Merge the partial softmax states of the slices of each split hub (in slice order) and finalise.
template <int VPL, int LPH>
__global__ void __launch_bounds__(kFastThreads) k_hub_merge_fwd(FastArgs A, float* __restrict__ out,
                                                                float* __restrict__ m_save,
                                                                float* __restrict__ l_save,
                                                                float* __restrict__ agg_save) {
  const int n_hub = *A.n_hub_dst, P = hub_parts_of(A, n_hub);
  const int w = (blockIdx.x * kFastThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (P == 1 || w >= n_hub) return;
  hub_merge_fwd_one<VPL, LPH>(A, w, P, lane, out, m_save, l_save, agg_save);
}

// ------------------------------------------------------------------------------ backward, by destination
// Dynamic shared memory: sh_gat[HD] (d attn of the CTA) + sh_tab[T][HD] (d edge table of the CTA,
// when it fits).  Static: per-warp partial d Pj rows for the hub merge.
template <int VPL, int LPH, bool ALIAS>
__device__ __forceinline__ void bwd_dst_fast_cta(
    const FastArgs& A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ V,
    const float* __restrict__ s_save, const float* __restrict__ m_save, const float* __restrict__ l_save,
    const float* __restrict__ agg_save, const float* __restrict__ G, float* __restrict__ gPj,
    float* __restrict__ g_attn, float* __restrict__ g_table, float* __restrict__ g_affine,
    float* __restrict__ gs_out, float* __restrict__ alpha_out, int tab_in_smem, int aff_in_smem, float* sh,
    int block, int nblocks) {
  constexpr int HD = 32 * VPL;
  constexpr int U = 2;                      // edges in flight (register budget: 128/thread)
  __shared__ float sm_part[kWarps][HD];
  float* sh_gat = sh;
  float* sh_tab = sh + HD;
  const bool typed = A.emode != B200GAT_EDGE_NONE;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  // tab_in_smem: 0 = atomics on the global table, 1 = one shared table per CTA (shared fp32 atomics, which are
  // CAS loops on sm_100), 2 = one private table per warp (plain read-modify-write: lane l owns its VPL columns)
  const int n_tab = tab_in_smem == 2 ? kWarps * A.T * HD : tab_in_smem == 1 ? A.T * HD : 0;
  // aff_in_smem: one private [K][HD] block per warp for the gradient of the affine map (lane l owns its VPL columns:
  // plain read-modify-write).  Global atomics here were 2.0 of 2.4 ms per layer on the PSP node-conflict graphs:
  // every resource-conflict edge carries two non-zero attributes, i.e. 2 * HD atomics on the same 2 * HD addresses.
  float* sh_aff = sh + HD + n_tab;
  const int n_sh = HD + n_tab + (aff_in_smem ? kWarps * A.K * HD : 0);
  for (int i = threadIdx.x; i < n_sh; i += kFastThreads) sh[i] = 0.f;
  __syncthreads();

  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31, c0 = lane * VPL, h = lane / LPH;
  const bool head_lead = (lane % LPH) == 0;
  const bool hub_block = block < A.hub_blocks;
  const int n_hub = hub_block ? *A.n_hub_dst : 0;
  const int P = hub_block ? hub_parts_of(A, n_hub) : 1;
  const int n_items = hub_block ? n_hub * P : (A.N + kWarps - 1) / kWarps;
  const int item0 = hub_block ? block : block - A.hub_blocks;
  const int istep = hub_block ? A.hub_blocks : nblocks - A.hub_blocks;
  float at[VPL], gat[VPL];
  ldrow<VPL>(at, A.attn + c0);
#pragma unroll
  for (int i = 0; i < VPL; ++i) gat[i] = 0.f;

  for (int item = item0; item < n_items; item += istep) {
    const WorkItem it = hub_block ? hub_item(A.in_ptr, A.hub_dst[item / P], wi, item % P, P)
                                  : node_item(A, A.in_ptr, A.dst_order, wi, item);
    const int v = it.node;
    float gpj[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) gpj[i] = 0.f;
    if (it.active && it.re > it.rb) {
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
      for (int e0 = it.rb; e0 < it.re; e0 += 32) {
        const int n = min(32, it.re - e0);
        const int my_u = lane < n ? A.in_src[e0 + lane] : 0;
        const int my_p = lane < n ? (A.in_pos ? A.in_pos[e0 + lane] : e0 + lane) : 0;
        const int my_t = (typed && lane < n) ? A.etype[my_p] : 0;
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
            ldrow<VPL>(pi[k], Pi + (int64_t)u * HD + c0);
            if constexpr (!ALIAS) ldrow<VPL>(vv[k], V + (int64_t)u * HD + c0);
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
    if (hub_block) {                       // ordered sum of the warps' partial d Pj rows
      strow<VPL>(&sm_part[wi][c0], gpj);
      __syncthreads();
      if (wi == 0) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) gpj[i] = 0.f;
        for (int w2 = 0; w2 < kWarps; ++w2) {
#pragma unroll
          for (int i = 0; i < VPL; ++i) gpj[i] += sm_part[w2][c0 + i];
        }
        if (P > 1) {
          strow<VPL>(A.hub_scratch + (int64_t)item * (HD + kHubRec) + c0, gpj);               // slice partial
          if (A.hub_cnt && hub_arrive(A, item / P, P, lane)) hub_merge_bwd_one<VPL>(A, item / P, P, lane, gPj);
        } else {
          strow<VPL>(gPj + (int64_t)v * HD + c0, gpj);
        }
      }
      __syncthreads();
    } else if (it.active) {
      strow<VPL>(gPj + (int64_t)v * HD + c0, gpj);
    }
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

template <int VPL, int LPH, bool ALIAS>
__global__ void __launch_bounds__(kFastThreads, B200GAT_FAST_MINB) k_bwd_dst_fast(
    FastArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ V,
    const float* __restrict__ s_save, const float* __restrict__ m_save, const float* __restrict__ l_save,
    const float* __restrict__ agg_save, const float* __restrict__ G, float* __restrict__ gPj,
    float* __restrict__ g_attn, float* __restrict__ g_table, float* __restrict__ g_affine,
    float* __restrict__ gs_out, float* __restrict__ alpha_out, int tab_in_smem, int aff_in_smem) {
  extern __shared__ float sh[];
  bwd_dst_fast_cta<VPL, LPH, ALIAS>(A, Pi, Pj, V, s_save, m_save, l_save, agg_save, G, gPj, g_attn, g_table, g_affine,
                                    gs_out, alpha_out, tab_in_smem, aff_in_smem, sh, (int)blockIdx.x, (int)gridDim.x);
}

// Sum the partial d Pj rows of the slices of each split hub (in slice order).
template <int VPL>
__global__ void __launch_bounds__(kFastThreads) k_hub_merge_bwd(FastArgs A, float* __restrict__ gPj) {
  const int n_hub = *A.n_hub_dst, P = hub_parts_of(A, n_hub);
  const int w = (blockIdx.x * kFastThreads + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (P == 1 || w >= n_hub) return;
  hub_merge_bwd_one<VPL>(A, w, P, lane, gPj);
}

// ------------------------------------------------------------------------------ backward, by source
template <int VPL, int LPH, bool ALIAS>
__device__ __forceinline__ void bwd_src_fast_cta(
    const FastArgs& A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ G,
    const float* __restrict__ gs_in, const float* __restrict__ alpha_in, float* __restrict__ gPi,
    float* __restrict__ gV, int block, int nblocks) {
  constexpr int HD = 32 * VPL;
  __shared__ float sm_part[kWarps][2 * HD];
  const int wi = threadIdx.x >> 5, lane = threadIdx.x & 31, c0 = lane * VPL, h = lane / LPH;
  const bool hub_block = block < A.hub_blocks;
  const bool typed = A.emode != B200GAT_EDGE_NONE;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  const int n_items = hub_block ? *A.n_hub_src : (A.N + kWarps - 1) / kWarps;
  const int item0 = hub_block ? block : block - A.hub_blocks;
  const int istep = hub_block ? A.hub_blocks : nblocks - A.hub_blocks;
  float at[VPL], ats[VPL];
  ldrow<VPL>(at, A.attn + c0);
#pragma unroll
  for (int i = 0; i < VPL; ++i) ats[i] = at[i] * A.slope;

  for (int item = item0; item < n_items; item += istep) {
    const WorkItem it = hub_block ? hub_item(A.out_ptr, A.hub_src[item], wi)
                                  : node_item(A, A.out_ptr, A.src_order, wi, item);
    const int u = it.node;
    float gpi[VPL], gv[VPL];
#pragma unroll
    for (int i = 0; i < VPL; ++i) { gpi[i] = 0.f; gv[i] = 0.f; }
    if (it.active && it.re > it.rb) {
      float p[VPL], pt[VPL];
      ldrow<VPL>(p, Pi + (int64_t)u * HD + c0);
#pragma unroll
      for (int i = 0; i < VPL; ++i) pt[i] = p[i];
      int t_cur = -1;
      for (int q0 = it.rb; q0 < it.re; q0 += 32) {
        const int n = min(32, it.re - q0);
        const int my_pos = lane < n ? A.out_pos[q0 + lane] : 0;
        const int my_v = lane < n ? A.out_dst[q0 + lane] : 0;
        const int my_t = (typed && lane < n) ? A.etype[my_pos] : 0;
        for (int j0 = 0; j0 < n; j0 += kUnroll) {
          float pj[kUnroll][VPL], gg[kUnroll][VPL], gsv[kUnroll], al[kUnroll];
          int tt[kUnroll], pp[kUnroll];
          float aa[kUnroll];
#pragma unroll
          for (int k = 0; k < kUnroll; ++k) {
            const int idx = min(j0 + k, n - 1);
            const int v = __shfl_sync(0xffffffffu, my_v, idx);
            pp[k] = __shfl_sync(0xffffffffu, my_pos, idx);
            aa[k] = affine ? affine_fetch(A, pp[k], lane) : 0.f;
            tt[k] = __shfl_sync(0xffffffffu, my_t, idx);
            ldrow<VPL>(pj[k], Pj + (int64_t)v * HD + c0);
            ldrow<VPL>(gg[k], G + (int64_t)v * HD + c0);
            const bool valid = j0 + k < n;
            gsv[k] = valid ? __ldg(gs_in + (int64_t)pp[k] * A.H + h) : 0.f;
            al[k] = valid ? __ldg(alpha_in + (int64_t)pp[k] * A.H + h) : 0.f;
          }
#pragma unroll
          for (int k = 0; k < kUnroll; ++k) {
            const bool valid = j0 + k < n;
            if (typed && tt[k] != t_cur) {
              t_cur = tt[k];
              type_row<VPL>(A, t_cur, c0, p, pt);
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
    bool finalize = it.active;
    if (hub_block) {
      strow<VPL>(&sm_part[wi][c0], gpi);
      strow<VPL>(&sm_part[wi][HD + c0], gv);
      __syncthreads();
      finalize = wi == 0;
      if (finalize) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) { gpi[i] = 0.f; gv[i] = 0.f; }
        for (int w2 = 0; w2 < kWarps; ++w2) {
#pragma unroll
          for (int i = 0; i < VPL; ++i) { gpi[i] += sm_part[w2][c0 + i]; gv[i] += sm_part[w2][HD + c0 + i]; }
        }
      }
      __syncthreads();
    }
    if (finalize) {
      if constexpr (ALIAS) {
#pragma unroll
        for (int i = 0; i < VPL; ++i) gpi[i] += gv[i];
        strow<VPL>(gPi + (int64_t)u * HD + c0, gpi);
      } else {
        strow<VPL>(gPi + (int64_t)u * HD + c0, gpi);
        strow<VPL>(gV + (int64_t)u * HD + c0, gv);
      }
    }
  }
}

template <int VPL, int LPH, bool ALIAS>
__global__ void __launch_bounds__(kFastThreads, B200GAT_FAST_MINB) k_bwd_src_fast(
    FastArgs A, const float* __restrict__ Pi, const float* __restrict__ Pj, const float* __restrict__ G,
    const float* __restrict__ gs_in, const float* __restrict__ alpha_in, float* __restrict__ gPi,
    float* __restrict__ gV) {
  bwd_src_fast_cta<VPL, LPH, ALIAS>(A, Pi, Pj, G, gs_in, alpha_in, gPi, gV, (int)blockIdx.x, (int)gridDim.x);
}

// ------------------------------------------------------------------------------ dispatch
static bool shape_of(const b200gat_attn* a, int& vpl, int& lph) {
  const int HD = a->heads * a->dim;
  if (HD != 64 && HD != 128 && HD != 256) return false;
  vpl = HD / 32;
  if (a->dim % vpl) return false;
  lph = a->dim / vpl;
  return lph == 4 || lph == 8 || lph == 16 || lph == 32;
}

bool fast_supported(const b200gat_attn* a, const b200gat_edge_feats* ef) {
  if (a->flavour != B200GAT_FLAVOUR_ADDITIVE) return false;
  const int mode = ef ? ef->mode : B200GAT_EDGE_NONE;
  if (mode == B200GAT_EDGE_DENSE) return false;
  if (mode == B200GAT_EDGE_AFFINE && (ef->n_attr < 0 || ef->n_attr > B200GAT_MAX_ATTR)) return false;
  int vpl, lph;
  return shape_of(a, vpl, lph);
}

// Node CTAs are persistent: at most two waves of the resident set (2 CTAs/SM), each looping over groups of
// kWarps nodes, so the per-CTA flushes of d attn / d table stay few.
static unsigned node_ctas(int64_t N) {
  const int64_t need = ceil_div(N * 32, kFastThreads), cap = 4 * kNumSMs;
  return (unsigned)(need < cap ? need : cap);
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static int fill_fast(FastArgs& A, const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef) {
  B200_CHECK_ARG(g->n_nodes >= 0 && g->n_edges >= 0, "negative graph size");
  B200_CHECK_ARG(g->n_nodes == 0 || (g->in_ptr && g->out_ptr), "null CSR pointers");
  B200_CHECK_ARG(g->n_edges == 0 || (g->in_src && g->out_dst && g->out_pos), "null CSR edge arrays");
  B200_CHECK_ARG(a->attn, "null attn");
  A.N = g->n_nodes; A.E = g->n_edges; A.flags = g->flags;
  A.in_ptr = g->in_ptr; A.in_src = g->in_src; A.out_ptr = g->out_ptr; A.out_dst = g->out_dst;
  A.out_pos = g->out_pos; A.dst_order = g->dst_order; A.src_order = g->src_order;
  A.in_pos = g->in_pos; A.node_blk = nullptr;
  A.hub_dst = g->hub_dst; A.n_hub_dst = g->n_hub_dst; A.hub_src = g->hub_src; A.n_hub_src = g->n_hub_src;
  A.hub_degree = g->hub_degree; A.hub_blocks = 0;
  A.hub_scratch = g->hub_scratch; A.hub_parts = g->hub_parts; A.hub_cap = g->hub_cap;
  A.hub_cnt = (g->hub_scratch && (g->flags & B200GAT_GRAPH_HUB_COUNTERS) && g->hub_parts > 1 && g->hub_cap > 0)
                  ? reinterpret_cast<int*>(g->hub_scratch + (int64_t)g->hub_cap * g->hub_parts * B200GAT_HUB_RECORD)
                  : nullptr;
  A.H = a->heads; A.slope = a->slope; A.eps = a->softmax_eps; A.attn = a->attn; A.out_bias = a->out_bias;
  A.emode = ef ? ef->mode : B200GAT_EDGE_NONE;
  A.T = ef ? ef->n_types : 0; A.K = ef ? ef->n_attr : 0;
  A.table = ef ? ef->table : nullptr; A.etype = ef ? ef->etype : nullptr;
  A.attr = ef ? ef->attr : nullptr; A.affine = ef ? ef->affine : nullptr;
  if (A.emode != B200GAT_EDGE_NONE)
    B200_CHECK_ARG(A.table && A.T > 0 && (A.E == 0 || A.etype), "null edge table/etype");
  if (A.emode == B200GAT_EDGE_AFFINE && A.K > 0)
    B200_CHECK_ARG(A.affine && (A.E == 0 || A.attr), "null affine edge attributes");
  return B200GAT_OK;
}

int fast_fwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef, const float* Pi,
             const float* Pj, const float* V, float* out, float* s_save, float* m_save, float* l_save,
             float* agg_save, const int32_t* node_blk, cudaStream_t st) {
  int vpl, lph;
  if (!shape_of(a, vpl, lph) || !aligned16(Pi) || !aligned16(Pj) || !aligned16(V) || !aligned16(out) ||
      !aligned16(a->attn) || !aligned16(a->out_bias) || !aligned16(agg_save) ||
      (ef && (!aligned16(ef->table) || !aligned16(ef->affine))))
    return B200GAT_ERR_UNSUPPORTED;
  FastArgs A;
  int rc = fill_fast(A, g, a, ef);
  if (rc) return rc;
  if (A.N == 0) return B200GAT_OK;
  A.node_blk = node_blk;
  A.hub_blocks = (A.hub_dst && A.n_hub_dst && A.hub_degree > 0) ? kHubBlocks : 0;
  const bool alias = (V == Pi);
  ScopedTimer tm(B200GAT_TIMING_CONV_FWD, st);
  // graphs flagged low-degree: the windowed kernel (its first CTAs are this kernel's hub CTAs)
  const bool win = win_enabled(A);
  if (win) {
    rc = win_fwd(A, vpl, lph, alias, Pi, Pj, V, out, s_save, m_save, l_save, agg_save, st);
    if (rc) return rc;
  }
  const unsigned blocks = win ? 0 : node_ctas(A.N) + A.hub_blocks;
  if (blocks) {
    FAST_DISPATCH(vpl, lph, alias,
                  (k_fwd_fast<VPL, LPH, ALIAS><<<blocks, kFastThreads, 0, st>>>(A, Pi, Pj, V, out, s_save, m_save,
                                                                                l_save, agg_save)));
    B200_CHECK_LAUNCH("k_fwd_fast");
  }
  if (A.hub_blocks && A.hub_scratch && A.hub_parts > 1 && A.hub_cap > 0 && !A.hub_cnt) {
    const unsigned mb = (unsigned)ceil_div((int64_t)A.hub_cap * 32, kFastThreads);
    switch (vpl) {
      case 2: FAST_DISPATCH_LPH(2, lph, false, (k_hub_merge_fwd<VPL, LPH><<<mb, kFastThreads, 0, st>>>(A, out, m_save, l_save, agg_save))) break;
      case 4: FAST_DISPATCH_LPH(4, lph, false, (k_hub_merge_fwd<VPL, LPH><<<mb, kFastThreads, 0, st>>>(A, out, m_save, l_save, agg_save))) break;
      default: FAST_DISPATCH_LPH(8, lph, false, (k_hub_merge_fwd<VPL, LPH><<<mb, kFastThreads, 0, st>>>(A, out, m_save, l_save, agg_save))) break;
    }
    B200_CHECK_LAUNCH("k_hub_merge_fwd");
  }
  return B200GAT_OK;
}

int fast_bwd(const b200gat_graph* g, const b200gat_attn* a, const b200gat_edge_feats* ef, const float* Pi,
             const float* Pj, const float* V, const float* s_save, const float* m_save, const float* l_save,
             const float* agg_save, const float* g_out, float* gPi, float* gPj, float* gV, float* g_attn,
             float* g_attn_bias, float* g_table, float* g_dense, float* g_affine, float* gs,
             int64_t n_pos, cudaStream_t st) {
  (void)g_attn_bias; (void)g_dense;
  int vpl, lph;
  if (!agg_save || !shape_of(a, vpl, lph) || !aligned16(Pi) || !aligned16(Pj) || !aligned16(V) ||
      !aligned16(g_out) || !aligned16(gPi) || !aligned16(gPj) || !aligned16(gV) || !aligned16(a->attn) ||
      !aligned16(agg_save) || (ef && (!aligned16(ef->table) || !aligned16(ef->affine))))
    return B200GAT_ERR_UNSUPPORTED;
  FastArgs A;
  int rc = fill_fast(A, g, a, ef);
  if (rc) return rc;
  if (A.emode != B200GAT_EDGE_NONE) B200_CHECK_ARG(g_table, "null g_table");
  if (A.emode == B200GAT_EDGE_AFFINE && A.K > 0) B200_CHECK_ARG(g_affine, "null g_affine");
  if (A.N == 0) return B200GAT_OK;
  const int HD = 32 * vpl;
  const unsigned nblocks = node_ctas(A.N);
  const bool alias = (V == Pi);
  const bool win = win_enabled(A);
  B200_CHECK_ARG(alias == (gV == gPi), "gV must alias gPi exactly when V aliases Pi");
  const int64_t tab_bytes = (int64_t)A.T * HD * 4;
  const int tab_in_smem = A.emode == B200GAT_EDGE_NONE ? 0 : kWarps * tab_bytes <= 64 * 1024 ? 2
                          : tab_bytes <= 40 * 1024 ? 1 : 0;
  const bool affine = A.emode == B200GAT_EDGE_AFFINE && A.K > 0;
  const int aff_in_smem = affine && (int64_t)kWarps * A.K * HD * 4 <= 64 * 1024 ? 1 : 0;
  const size_t smem = sizeof(float) * (HD + (tab_in_smem == 2 ? (size_t)kWarps * A.T * HD
                                             : tab_in_smem == 1 ? (size_t)A.T * HD : 0) +
                                       (aff_in_smem ? (size_t)kWarps * A.K * HD : 0));
  float* alpha = gs + (n_pos > 0 ? n_pos : (int64_t)A.E) * A.H;   // edge positions may be those of the full graph
  {
    ScopedTimer tm(B200GAT_TIMING_CONV_BWD_DST, st);
    A.hub_blocks = (A.hub_dst && A.n_hub_dst && A.hub_degree > 0) ? kHubBlocks : 0;
    if (win) {
      rc = win_bwd_dst(A, vpl, lph, alias, Pi, Pj, V, s_save, m_save, l_save, agg_save, g_out, gPj, g_attn, g_table,
                       g_affine, gs, alpha, st);
      if (rc) return rc;
    }
    const unsigned blocks = win ? 0 : nblocks + A.hub_blocks;
    if (smem > 48 * 1024) {
      FAST_DISPATCH(vpl, lph, alias,
                    (cudaFuncSetAttribute(k_bwd_dst_fast<VPL, LPH, ALIAS>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)));
    }
    if (blocks) {
      FAST_DISPATCH(vpl, lph, alias,
                    (k_bwd_dst_fast<VPL, LPH, ALIAS><<<blocks, kFastThreads, smem, st>>>(
                        A, Pi, Pj, V, s_save, m_save, l_save, agg_save, g_out, gPj, g_attn, g_table, g_affine,
                        gs, alpha, tab_in_smem, aff_in_smem)));
      note_launch(1);
    }
    if (A.hub_blocks && A.hub_scratch && A.hub_parts > 1 && A.hub_cap > 0 && !A.hub_cnt) {
      const unsigned mb = (unsigned)ceil_div((int64_t)A.hub_cap * 32, kFastThreads);
      if (vpl == 2) k_hub_merge_bwd<2><<<mb, kFastThreads, 0, st>>>(A, gPj);
      else if (vpl == 4) k_hub_merge_bwd<4><<<mb, kFastThreads, 0, st>>>(A, gPj);
      else k_hub_merge_bwd<8><<<mb, kFastThreads, 0, st>>>(A, gPj);
      note_launch(1);
    }
  }
  {
    ScopedTimer tm(B200GAT_TIMING_CONV_BWD_SRC, st);
    A.hub_blocks = (A.hub_src && A.n_hub_src && A.hub_degree > 0) ? kHubBlocks : 0;
    if (win) {
      rc = win_bwd_src(A, vpl, lph, alias, Pi, Pj, g_out, gs, alpha, gPi, gV, st);
      if (rc) return rc;
    }
    const unsigned blocks = win ? 0 : nblocks + A.hub_blocks;
    if (blocks) {
      FAST_DISPATCH(vpl, lph, alias,
                    (k_bwd_src_fast<VPL, LPH, ALIAS><<<blocks, kFastThreads, 0, st>>>(A, Pi, Pj, g_out, gs, alpha,
                                                                                     gPi, gV)));
      note_launch(1);
    }
  }
  B200_CHECK_LAUNCHES("k_bwd_fast", 0);
  return B200GAT_OK;
}

}  // namespace b200gat

// The windowed kernels call the *_fast_cta functions above for their hub CTAs: same translation unit.
#include "gat_win.cuh"
