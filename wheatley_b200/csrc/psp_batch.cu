// Device-side PSP / PyG-flavour batch builder (include/b200gat.h "PSP batch builder").
//
// Input: an arena of fixed-layout observation slots (the wire format of N2, one slot per schedule graph: header,
// feat [max_nodes][16], resources [max_nodes][max_res], res_cal [max_res][3], prec [2][max_prec], rp [2][max_rp],
// rp_r [max_rp][4]) copied to the device as it is.  Output: the homogeneous batched graph psp/models/agent.py
// Agent.preprocess hands to the GNN - same nodes, same edges IN THE SAME ORDER, same per-edge attributes:
//   for every base edge type t in the batch's store order (prec | rprec | rp | rrp | rc):
//       graph 0's edges of type t, graph 1's, ...                           (Batch.from_data_list per store)
//   self loops of all task nodes                                             (pyg_graph.py:331-335)
//   nodeconf | rnodeconf | selfres          (conflicts == "node",            rewiring.py:104-253)
//   pool | [rpool] | poolres | [rpoolres] | selfpool   (learned pooling,    rewiring.py:27-101)
// rc edges of a graph = non-zeros of used[i, r] & used[j, r] & (i != j) in (i, j, r) order
// (psp/utils/utils.py:73-108); nodeconf edges = non-zeros of resources[node, r] in (node, r) order over the
// batch (rewiring.py:146-160).
//
// Integer / byte work: 48 bytes written per edge (int64 src, dst, type; int32 rid; 5 fp32 attributes), the
// observation is read once.  One CTA column per graph; the rows of a graph (rc rows, feature rows) are split over
// gridDim.x CTAs which each recompute the graph's resource masks and row offsets in shared memory.
#include <algorithm>
#include "common.cuh"

namespace b200gat {
namespace {

constexpr int kThreads = 256;
constexpr int kWarpsPerCta = kThreads / 32;
constexpr int kF = B200GAT_PSP_N_FEAT;

enum { T_SELF = 0, T_PREC = 1, T_RPREC = 2, T_RC = 3, T_RP = 4, T_RRP = 5, T_POOL = 6, T_RPOOL = 7,
       T_NODECONF = 10, T_RNODECONF = 11, T_POOLRES = 12, T_RPOOLRES = 13, T_SELFPOOL = 14, T_SELFRES = 15 };

struct Out {
  int64_t* src;
  int64_t* dst;
  int64_t* type;
  int32_t* rid;
  float* attr;      // [E, 5] = att_rc (2) | att_rp (3)
};

__device__ __forceinline__ void put(const Out& o, int64_t e, int64_t s, int64_t d, int type, int rid, float a0,
                                    float a1, float p0, float p1, float p2) {
  o.src[e] = s;
  o.dst[e] = d;
  o.type[e] = type;
  o.rid[e] = rid;
  float* a = o.attr + e * 5;
  a[0] = a0; a[1] = a1; a[2] = p0; a[3] = p1; a[4] = p2;
}

__device__ __forceinline__ int block_excl_scan(int v, int* warp_sums, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) warp_sums[warp] = x;
  __syncthreads();
  int base = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < kWarpsPerCta; ++w) {
    const int s = warp_sums[w];
    if (w < warp) base += s;
    tot += s;
  }
  __syncthreads();
  total = tot;
  return base + x - v;
}

// grid (parts, B); dynamic smem: mask [max_nodes] u32 + rowoff [max_nodes + 1] i32
__global__ void __launch_bounds__(kThreads)
k_psp_graph(b200gat_psp_obs o, Out out, float* __restrict__ feat, float* __restrict__ res_cal_id) {
  extern __shared__ int smem_i[];
  uint32_t* mask = reinterpret_cast<uint32_t*>(smem_i);
  int* rowoff = smem_i + o.layout.max_nodes;
  __shared__ int warp_sums[kWarpsPerCta];
  __shared__ int cnt_r[B200GAT_PSP_MAX_RES];
  const int g = blockIdx.y, part = blockIdx.x, parts = gridDim.x, t = threadIdx.x;
  const int lane = t & 31, warp = t >> 5;
  const int B = o.n_graphs;
  const b200gat_psp_layout& L = o.layout;
  const int32_t* slot = o.arena + (int64_t)g * L.slot_words;
  const int n = slot[1], Rg = slot[2], np_ = slot[3], nrp = slot[4];
  const float* s_feat = reinterpret_cast<const float*>(slot + B200GAT_PSP_HEADER_WORDS);
  const float* s_res = s_feat + (int64_t)L.max_nodes * kF;
  const float* s_cal = s_res + (int64_t)L.max_nodes * L.max_res;
  const int32_t* s_prec = slot + B200GAT_PSP_HEADER_WORDS + (int64_t)L.max_nodes * kF +
                          (int64_t)L.max_nodes * L.max_res + (int64_t)L.max_res * 3;
  const int32_t* s_rp = s_prec + 2 * (int64_t)L.max_prec;
  const float* s_rpr = reinterpret_cast<const float*>(s_rp + 2 * (int64_t)L.max_rp);
  const int64_t* meta = o.meta;
  const int64_t noff = meta[g];
  const int64_t NT = o.n_task_nodes;
  const int R = o.n_res;
  const bool node_mode = o.conflicts == B200GAT_CONFLICTS_NODE;
  const bool clique = o.conflicts == B200GAT_CONFLICTS_CLIQUE;
  const int64_t res_base = NT;
  const int64_t pool_base = NT + (node_mode ? (int64_t)B * R : 0);

  // ---- feature rows of the task nodes (all parts), extra rows zero-filled (part 0)
  for (int idx = part * kThreads + t; idx < n * (kF / 4); idx += parts * kThreads) {
    const int row = idx / (kF / 4), c4 = (idx - row * (kF / 4)) * 4;
    *reinterpret_cast<float4*>(feat + (noff + row) * kF + c4) =
        *reinterpret_cast<const float4*>(s_feat + (int64_t)row * kF + c4);
  }

  // ---- resource masks, per-resource user counts
  if (clique || node_mode) {
    if (t < B200GAT_PSP_MAX_RES) cnt_r[t] = 0;
    __syncthreads();
    for (int i = t; i < n; i += kThreads) {
      uint32_t m = 0;
      for (int r = 0; r < Rg; ++r)
        if (s_res[(int64_t)i * L.max_res + r] != 0.f) m |= 1u << r;
      mask[i] = m;
      if (clique)
        for (uint32_t mm = m; mm; mm &= mm - 1) atomicAdd(&cnt_r[__ffs(mm) - 1], 1);
    }
    __syncthreads();
  }

  if (clique && o.seg_start[T_RC] >= 0) {
    // rows of the (i, j, r) enumeration: row i holds sum_{r in mask_i} (cnt_r - 1) edges
    int carry = 0;
    for (int i0 = 0; i0 < n; i0 += kThreads) {
      const int i = i0 + t;
      int c = 0;
      if (i < n)
        for (uint32_t mm = mask[i]; mm; mm &= mm - 1) c += cnt_r[__ffs(mm) - 1] - 1;
      int total;
      const int ex = block_excl_scan(c, warp_sums, total);
      if (i < n) rowoff[i] = carry + ex;
      carry += total;
    }
    __syncthreads();
    const int64_t base = o.seg_start[T_RC] + meta[3 * (B + 1) + g];
    for (int i = part * kWarpsPerCta + warp; i < n; i += parts * kWarpsPerCta) {
      const uint32_t mi = mask[i];
      if (!mi) continue;
      int64_t run = base + rowoff[i];
      for (int j0 = 0; j0 < n; j0 += 32) {
        const int j = j0 + lane;
        const uint32_t both = (j < n && j != i) ? (mi & mask[j]) : 0u;
        const int c = __popc(both);
        int x = c;
#pragma unroll
        for (int s = 1; s < 32; s <<= 1) {
          const int y = __shfl_up_sync(0xffffffffu, x, s);
          if (lane >= s) x += y;
        }
        const int tot = __shfl_sync(0xffffffffu, x, 31);
        int64_t e = run + x - c;
        for (uint32_t mm = both; mm; mm &= mm - 1) {
          const int r = __ffs(mm) - 1;
          put(out, e++, noff + i, noff + j, T_RC, r + 1, s_res[(int64_t)i * L.max_res + r],
              s_res[(int64_t)j * L.max_res + r], 0.f, 0.f, 0.f);
        }
        run += tot;
      }
    }
  }
  if (part != 0) return;

  // ---- prec / rprec / rp / rrp                                     (agent_graph_observation.py:107-118)
  if (o.seg_start[T_PREC] >= 0 || o.seg_start[T_RPREC] >= 0) {
    const int64_t pb = meta[(B + 1) + g];
    for (int e = t; e < np_; e += kThreads) {
      const int64_t u = noff + s_prec[e], v = noff + s_prec[L.max_prec + e];
      if (o.seg_start[T_PREC] >= 0) put(out, o.seg_start[T_PREC] + pb + e, u, v, T_PREC, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
      if (o.seg_start[T_RPREC] >= 0) put(out, o.seg_start[T_RPREC] + pb + e, v, u, T_RPREC, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
    }
  }
  if (o.seg_start[T_RP] >= 0 || o.seg_start[T_RRP] >= 0) {
    const int64_t rb = meta[2 * (B + 1) + g];
    for (int e = t; e < nrp; e += kThreads) {
      const int64_t u = noff + s_rp[e], v = noff + s_rp[L.max_rp + e];
      const float* rr = s_rpr + (int64_t)e * 4;
      const float4 r = make_float4(rr[0], rr[1], rr[2], rr[3]);
      const int rid = (int)(r.x + 1.f);                                  // (r[:, 0] + 1).int()  rewiring.py:417
      const float p2 = r.w + 1.f;                                        // att_rp[:, -1] + 1     :420
      if (o.seg_start[T_RP] >= 0) put(out, o.seg_start[T_RP] + rb + e, u, v, T_RP, rid, 0.f, 0.f, r.y, r.z, p2);
      if (o.seg_start[T_RRP] >= 0) put(out, o.seg_start[T_RRP] + rb + e, v, u, T_RRP, rid, 0.f, 0.f, r.y, r.z, p2);
    }
  }
  // ---- self loops, pool edges of the task nodes
  for (int k = t; k < n; k += kThreads) {
    const int64_t v = noff + k;
    if (o.seg_start[T_SELF] >= 0) put(out, o.seg_start[T_SELF] + v, v, v, T_SELF, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
    if (o.seg_start[T_POOL] >= 0) put(out, o.seg_start[T_POOL] + v, v, pool_base + g, T_POOL, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
    if (o.seg_start[T_RPOOL] >= 0) put(out, o.seg_start[T_RPOOL] + v, pool_base + g, v, T_RPOOL, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
  }
  if (o.pooling != B200GAT_POOL_NONE) {
    if (t == 0) put(out, o.seg_start[T_SELFPOOL] + g, pool_base + g, pool_base + g, T_SELFPOOL, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
    if (t < kF) feat[(pool_base + g) * kF + t] = 0.f;
  }
  // ---- resource nodes                                               (rewiring.py:104-253)
  if (node_mode) {
    int carry = 0;
    const int64_t cb = meta[4 * (B + 1) + g];
    for (int i0 = 0; i0 < n; i0 += kThreads) {
      const int i = i0 + t;
      const uint32_t m = i < n ? mask[i] : 0u;
      int total;
      int ex = block_excl_scan(__popc(m), warp_sums, total);
      int64_t e = cb + carry + ex;
      for (uint32_t mm = m; mm; mm &= mm - 1) {
        const int r = __ffs(mm) - 1;
        const float v = s_res[(int64_t)i * L.max_res + r];
        const int64_t rn = res_base + (int64_t)g * R + r;
        put(out, o.seg_start[T_NODECONF] + e, noff + i, rn, T_NODECONF, r, v, v, 0.f, 0.f, 0.f);
        put(out, o.seg_start[T_RNODECONF] + e, rn, noff + i, T_RNODECONF, r, v, v, 0.f, 0.f, 0.f);
        ++e;
      }
      carry += total;
    }
    for (int idx = t; idx < R * kF; idx += kThreads) feat[(res_base + (int64_t)g * R) * kF + idx] = 0.f;
    if (t < R) {
      if (o.seg_start[T_POOLRES] >= 0)
        put(out, o.seg_start[T_POOLRES] + (int64_t)g * R + t, res_base + (int64_t)g * R + t, pool_base + g,
            T_POOLRES, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
      if (o.seg_start[T_RPOOLRES] >= 0)
        put(out, o.seg_start[T_RPOOLRES] + (int64_t)g * R + t, pool_base + g, res_base + (int64_t)g * R + t,
            T_RPOOLRES, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
    }
    // selfres: the reference indexes the [B * R] resource-node list with the COLUMN ids of the resources no graph
    // of the batch uses (rewiring.py:218-228), i.e. always nodes of graph 0's block
    if (g == 0 && t < R && ((o.unused_res_mask >> t) & 1u)) {
      const int k = __popc(o.unused_res_mask & ((1u << t) - 1u));
      put(out, o.seg_start[T_SELFRES] + k, res_base + t, res_base + t, T_SELFRES, 0, 0.f, 0.f, 0.f, 0.f, 0.f);
    }
  }
  // ---- res_cal ids, padded to the batch's resource count           (agent_graph_observation.py:264-269)
  for (int idx = t; idx < R * 3; idx += kThreads) {
    const int r = idx / 3;
    res_cal_id[((int64_t)g * R) * 3 + idx] = r < Rg ? s_cal[idx] : 0.f;
  }
}

}  // namespace
}  // namespace b200gat

using namespace b200gat;

extern "C" int64_t b200gat_psp_slot_words(const b200gat_psp_layout* l) {
  if (!l) return 0;
  int64_t w = B200GAT_PSP_HEADER_WORDS + (int64_t)l->max_nodes * kF + (int64_t)l->max_nodes * l->max_res +
              (int64_t)l->max_res * 3 + 2 * (int64_t)l->max_prec + 2 * (int64_t)l->max_rp + 4 * (int64_t)l->max_rp;
  return (w + 3) & ~3ll;
}

// Host: segment starts / totals from the per-graph prefix table (host copy of `meta`).
extern "C" int b200gat_psp_batch_sizes(b200gat_psp_obs* o, const int64_t* meta_host) {
  B200_CHECK_ARG(o && meta_host, "null descriptor");
  B200_CHECK_ARG(o->n_graphs >= 0 && o->n_res >= 0 && o->n_res <= B200GAT_PSP_MAX_RES, "bad sizes");
  const int B = o->n_graphs;
  const int64_t NT = meta_host[B], NP = meta_host[(B + 1) + B], NRP = meta_host[2 * (B + 1) + B];
  const int64_t NRC = meta_host[3 * (B + 1) + B], NC = meta_host[4 * (B + 1) + B];
  const bool node_mode = o->conflicts == B200GAT_CONFLICTS_NODE;
  for (int i = 0; i < 16; ++i) o->seg_start[i] = -1;
  int64_t e = 0;
  bool seen[5] = {false, false, false, false, false};
  for (int k = 0; k < 5; ++k) {
    const int c = o->base_order[k];
    if (c < 0) continue;
    B200_CHECK_ARG(c < 5 && !seen[c], "base_order must list each of prec/rprec/rp/rrp/rc at most once");
    seen[c] = true;
    switch (c) {
      case 0: o->seg_start[T_PREC] = e; e += NP; break;
      case 1: o->seg_start[T_RPREC] = e; e += NP; break;
      case 2: o->seg_start[T_RP] = e; e += NRP; break;
      case 3: o->seg_start[T_RRP] = e; e += NRP; break;
      case 4:
        if (o->conflicts == B200GAT_CONFLICTS_CLIQUE) { o->seg_start[T_RC] = e; e += NRC; }
        break;
    }
  }
  if (o->self_loops) { o->seg_start[T_SELF] = e; e += NT; }
  int64_t n_total = NT;
  if (node_mode) {
    o->seg_start[T_NODECONF] = e; e += NC;
    o->seg_start[T_RNODECONF] = e; e += NC;
    o->seg_start[T_SELFRES] = e;
    e += __builtin_popcount(o->unused_res_mask & (o->n_res >= 32 ? 0xffffffffu : ((1u << o->n_res) - 1u)));
    n_total += (int64_t)B * o->n_res;
  }
  if (o->pooling != B200GAT_POOL_NONE) {
    o->seg_start[T_POOL] = e; e += NT;
    if (o->pooling == B200GAT_POOL_LEARNINV) { o->seg_start[T_RPOOL] = e; e += NT; }
    if (node_mode) {
      o->seg_start[T_POOLRES] = e; e += (int64_t)B * o->n_res;
      if (o->pooling == B200GAT_POOL_LEARNINV) { o->seg_start[T_RPOOLRES] = e; e += (int64_t)B * o->n_res; }
    }
    o->seg_start[T_SELFPOOL] = e; e += B;
    n_total += B;
  }
  o->n_edges = e;
  o->n_task_nodes = NT;
  o->n_total_nodes = n_total;
  return B200GAT_OK;
}

extern "C" int b200gat_psp_batch_build(const b200gat_psp_obs* o, int64_t* edge_index, int64_t* type, int32_t* rid,
                                       float* attr, float* feat, float* res_cal_id, void* stream) {
  B200_CHECK_ARG(o, "null descriptor");
  if (o->n_graphs == 0) return B200GAT_OK;
  B200_CHECK_ARG(o->arena && o->meta && feat && res_cal_id, "null pointer");
  if (o->n_edges > 0) B200_CHECK_ARG(edge_index && type && rid && attr, "null edge output");
  B200_CHECK_ARG(o->layout.slot_words == b200gat_psp_slot_words(&o->layout), "slot_words does not match the layout");
  B200_CHECK_ARG(o->layout.max_res <= B200GAT_PSP_MAX_RES && o->n_res <= o->layout.max_res, "too many resources");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(o->arena) & 15) == 0 && (reinterpret_cast<uintptr_t>(feat) & 15) == 0,
                 "arena / feat must be 16-byte aligned");
  Out out{edge_index, edge_index + o->n_edges, type, rid, attr};
  const int parts = (int)std::min<int64_t>(32, std::max<int64_t>(1, ceil_div(o->layout.max_nodes, 128)));
  const size_t smem = (size_t)(2 * o->layout.max_nodes + 1) * sizeof(int);
  if (smem > 48 * 1024) {
    if (smem > 200 * 1024) {
      set_error("psp batch builder: %d nodes per graph exceed the shared-memory row tables", o->layout.max_nodes);
      return B200GAT_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(k_psp_graph, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  ScopedTimer tm(B200GAT_TIMING_BATCH, (cudaStream_t)stream);
  k_psp_graph<<<dim3(parts, o->n_graphs), kThreads, smem, (cudaStream_t)stream>>>(*o, out, feat, res_cal_id);
  B200_CHECK_LAUNCH("k_psp_graph");
  return B200GAT_OK;
}
