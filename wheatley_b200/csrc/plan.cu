// Graph plan: int64 COO -> int32 CSR by destination + CSR by source, both STABLE
// (edges keep their original relative order inside a segment), so every segmented
// reduction downstream runs in a fixed order and is bit-reproducible.
//
// Pipeline (all on `stream`, no host sync):
//   count   : 16-byte loads of the int64 endpoints, atomic histogram into ptr[i+1]
//   scan    : 3-phase inclusive scan over N+1 counters (2048 per block)
//   fill    : atomic slot claim inside each segment (order arbitrary)
//   sort    : per-segment rank-by-counting on the original edge id (warp per node for
//             degree <= 256, block per node above) -> stable order
//   finish  : in_src / inverse permutation / out_dst / out_pos, degree statistics
#include "common.cuh"

namespace b200gat {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048
constexpr int kWarpSortMax = 256;

// Warp-aggregated atomics: consecutive edges of a batched schedule graph share an endpoint (the conflict-clique
// edges of one task are stored together), so a warp's 32 atomics on one counter become one.
__device__ __forceinline__ void agg_inc(int32_t* base, int key, bool on) {
  const unsigned act = __ballot_sync(0xffffffffu, on);
  if (!on) return;
  const unsigned peers = __match_any_sync(act, key);
  if ((threadIdx.x & 31) == __ffs(peers) - 1) atomicAdd(base + key, __popc(peers));
}
// slot = (old value of base[key]) + rank among the warp's lanes with the same key
__device__ __forceinline__ int agg_slot(int32_t* base, int key, bool on) {
  const unsigned act = __ballot_sync(0xffffffffu, on);
  if (!on) return 0;
  const unsigned peers = __match_any_sync(act, key);
  const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
  int old = 0;
  if (lane == leader) old = atomicAdd(base + key, __popc(peers));
  old = __shfl_sync(peers, old, leader);
  return old + __popc(peers & ((1u << lane) - 1u));
}

__global__ void k_count(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t E,
                        int32_t N, int32_t* __restrict__ in_ptr, int32_t* __restrict__ out_ptr,
                        int32_t* __restrict__ stats, int vec) {
  // two edges per thread: one 16-byte load per endpoint array when both are 16-byte aligned
  int64_t i = (int64_t)(blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * 2;
  int bad = 0;
  longlong2 s = make_longlong2(-1, -1), d = make_longlong2(-1, -1);
  const bool has0 = i < E, has1 = i + 1 < E;
  if (has1) {
    if (vec) {
      s = *reinterpret_cast<const longlong2*>(src + i);
      d = *reinterpret_cast<const longlong2*>(dst + i);
    } else {
      s = make_longlong2(src[i], src[i + 1]);
      d = make_longlong2(dst[i], dst[i + 1]);
    }
  } else if (has0) {
    s.x = src[i]; d.x = dst[i];
  }
  const bool ok0 = has0 && s.x >= 0 && s.x < N && d.x >= 0 && d.x < N;
  const bool ok1 = has1 && s.y >= 0 && s.y < N && d.y >= 0 && d.y < N;
  bad = (has0 && !ok0) + (has1 && !ok1);
  agg_inc(in_ptr + 1, (int)d.x, ok0);
  agg_inc(out_ptr + 1, (int)s.x, ok0);
  agg_inc(in_ptr + 1, (int)d.y, ok1);
  agg_inc(out_ptr + 1, (int)s.y, ok1);
  if (bad) atomicAdd(&stats[2], bad);
}

// ---- scan -----------------------------------------------------------------------------
__device__ __forceinline__ int block_inclusive_scan(int v, int* smem_warp) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) smem_warp[w] = v;
  __syncthreads();
  if (w == 0) {
    int x = lane < (kScanThreads / 32) ? smem_warp[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += t;
    }
    smem_warp[lane] = x;
  }
  __syncthreads();
  if (w > 0) v += smem_warp[w - 1];
  __syncthreads();
  return v;
}

// mode 0: block sums only; mode 1: in-place inclusive scan with block offsets.
__global__ void k_scan_tiles(int32_t* __restrict__ a, int64_t n, int32_t* __restrict__ bsum,
                             int mode) {
  __shared__ int sw[32];
  int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int v[kScanItems];
  int tsum = 0;
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    v[k] = (base + k < n) ? a[base + k] : 0;
    tsum += v[k];
  }
  int incl = block_inclusive_scan(tsum, sw);
  if (mode == 0) {
    if (threadIdx.x == kScanThreads - 1) bsum[blockIdx.x] = incl;
    return;
  }
  int run = incl - tsum + (blockIdx.x > 0 ? bsum[blockIdx.x - 1] : 0);
#pragma unroll
  for (int k = 0; k < kScanItems; ++k) {
    run += v[k];
    if (base + k < n) a[base + k] = run;
  }
}

// single block: inclusive scan of bsum[0..nb) in place
__global__ void k_scan_bsums(int32_t* __restrict__ bsum, int nb) {
  __shared__ int sw[32];
  __shared__ int carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  for (int base = 0; base < nb; base += kScanThreads) {
    int i = base + threadIdx.x;
    int v = i < nb ? bsum[i] : 0;
    int incl = block_inclusive_scan(v, sw);
    int carry = carry_s;
    if (i < nb) bsum[i] = incl + carry;
    __syncthreads();
    if (threadIdx.x == kScanThreads - 1) carry_s = carry + incl;
    __syncthreads();
  }
}

// ---- fill -----------------------------------------------------------------------------
__global__ void k_fill(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t E,
                       int32_t N, const int32_t* __restrict__ in_ptr,
                       const int32_t* __restrict__ out_ptr, int32_t* __restrict__ fill_in,
                       int32_t* __restrict__ fill_out, int32_t* __restrict__ tmp_in,
                       int32_t* __restrict__ tmp_out) {
  int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int64_t s = -1, d = -1;
  if (e < E) { s = src[e]; d = dst[e]; }
  const bool ok = e < E && s >= 0 && s < N && d >= 0 && d < N;
  const int si = agg_slot(fill_in, (int)d, ok);
  const int so = agg_slot(fill_out, (int)s, ok);
  if (!ok) return;
  tmp_in[in_ptr[d] + si] = (int32_t)e;
  tmp_out[out_ptr[s] + so] = (int32_t)e;
}

// ---- stable order inside segments: rank by counting -------------------------------------
// One warp per node.  Segments longer than kWarpSortMax are pushed to `large` instead.
__global__ void k_sort_small(const int32_t* __restrict__ ptr, int32_t N,
                             const int32_t* __restrict__ tmp, int32_t* __restrict__ sorted,
                             int32_t* __restrict__ large, int32_t* __restrict__ n_large) {
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  int lane = threadIdx.x & 31;
  if (warp >= N) return;
  int b = ptr[warp], deg = ptr[warp + 1] - b;
  if (deg <= 0) return;
  if (deg > kWarpSortMax) {
    if (lane == 0) large[atomicAdd(n_large, 1)] = warp;
    return;
  }
  if (deg <= 32) {
    int v = lane < deg ? tmp[b + lane] : 0x7fffffff;
    int rank = 0;
    for (int j = 0; j < deg; ++j) rank += (__shfl_sync(0xffffffffu, v, j) < v);
    if (lane < deg) sorted[b + rank] = v;
    return;
  }
  for (int i = lane; i < deg; i += 32) {
    int v = tmp[b + i];
    int rank = 0;
    for (int j = 0; j < deg; ++j) rank += (__ldg(tmp + b + j) < v);
    sorted[b + rank] = v;
  }
}

// Persistent blocks over the list of long segments; shared-memory tiles of the segment.
__global__ void k_sort_large(const int32_t* __restrict__ ptr, const int32_t* __restrict__ tmp,
                             int32_t* __restrict__ sorted, const int32_t* __restrict__ large,
                             const int32_t* __restrict__ n_large) {
  constexpr int kTile = 2048;
  __shared__ int tile[kTile];
  int nl = *n_large;
  for (int li = blockIdx.x; li < nl; li += gridDim.x) {
    int node = large[li];
    int b = ptr[node], deg = ptr[node + 1] - b;
    for (int i0 = 0; i0 < deg; i0 += blockDim.x) {
      int i = i0 + threadIdx.x;
      int v = i < deg ? tmp[b + i] : 0;
      int rank = 0;
      for (int t0 = 0; t0 < deg; t0 += kTile) {
        int tn = min(kTile, deg - t0);
        __syncthreads();
        for (int k = threadIdx.x; k < tn; k += blockDim.x) tile[k] = tmp[b + t0 + k];
        __syncthreads();
        if (i < deg)
          for (int k = 0; k < tn; ++k) rank += (tile[k] < v);
      }
      if (i < deg) sorted[b + rank] = v;
    }
    __syncthreads();
  }
}

// ---- finish -----------------------------------------------------------------------------
// Edges with an endpoint outside [0, N) are counted (stats[2]) and left out of both CSRs: only the first
// in_ptr[N] (= out_ptr[N]) positions of the sorted arrays are defined; the rest is never written and must not be
// dereferenced (an uninitialised edge id there was an out-of-bounds read whenever the scratch held stale data).
__global__ void k_finish_in(const int64_t* __restrict__ src, int64_t E, const int32_t* __restrict__ in_ptr, int32_t N,
                            const int32_t* __restrict__ in_eid, int32_t* __restrict__ in_src,
                            int32_t* __restrict__ inv) {
  int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (p >= E) return;
  if (p >= in_ptr[N]) { in_src[p] = 0; return; }
  int e = in_eid[p];
  in_src[p] = (int32_t)src[e];
  inv[e] = (int32_t)p;
}
__global__ void k_finish_out(const int64_t* __restrict__ dst, int64_t E, const int32_t* __restrict__ out_ptr, int32_t N,
                             const int32_t* __restrict__ out_eid, const int32_t* __restrict__ inv,
                             int32_t* __restrict__ out_dst, int32_t* __restrict__ out_pos) {
  int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (q >= E) return;
  if (q >= out_ptr[N]) { out_dst[q] = 0; out_pos[q] = 0; return; }
  int e = out_eid[q];
  out_dst[q] = (int32_t)dst[e];
  out_pos[q] = inv[e];
}
__global__ void k_degree_stats(const int32_t* __restrict__ in_ptr,
                               const int32_t* __restrict__ out_ptr, int32_t N,
                               int32_t* __restrict__ stats) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int din = 0, dout = 0, zero = 0;
  if (i < N) {
    din = in_ptr[i + 1] - in_ptr[i];
    dout = out_ptr[i + 1] - out_ptr[i];
    zero = din == 0;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    din = max(din, __shfl_xor_sync(0xffffffffu, din, o));
    dout = max(dout, __shfl_xor_sync(0xffffffffu, dout, o));
    zero += __shfl_xor_sync(0xffffffffu, zero, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (zero) atomicAdd(&stats[0], zero);
    atomicMax(&stats[1], din);
    atomicMax(&stats[3], dout);
  }
}

// ---- locality order of the nodes --------------------------------------------------------
// key[v] = neighbour at the middle of v's adjacency segment (v itself when the segment is
// empty).  Nodes whose neighbourhoods overlap heavily (members of one machine clique) get
// nearly equal keys, so sorting the nodes by key puts them next to each other and the warps of
// one CTA then gather the same rows (L1 hits instead of L2 round trips).
__global__ void k_order_key(const int32_t* __restrict__ ptr, const int32_t* __restrict__ nbr, int32_t N,
                            int32_t* __restrict__ key, int32_t* __restrict__ hist) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  int b = ptr[v], e = ptr[v + 1];
  int k = e > b ? nbr[b + ((e - b) >> 1)] : v;
  key[v] = k;
  atomicAdd(&hist[k + 1], 1);
}
__global__ void k_order_fill(const int32_t* __restrict__ key, int32_t N, const int32_t* __restrict__ hist,
                             int32_t* __restrict__ fill, int32_t* __restrict__ tmp) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  int k = key[v];
  tmp[hist[k] + atomicAdd(&fill[k], 1)] = v;
}

__global__ void k_hub_list(const int32_t* __restrict__ ptr, int32_t N, int32_t thresh,
                           int32_t* __restrict__ list, int32_t* __restrict__ count) {
  int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  if (ptr[v + 1] - ptr[v] > thresh) list[atomicAdd(count, 1)] = v;
}

__global__ void k_gather_i32(const int32_t* __restrict__ table, const int32_t* __restrict__ idx,
                             int64_t n, int32_t* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = table[idx[i]];
}
// gather + range of the gathered values: range[0] = min, range[1] = max (caller initialises INT_MAX / INT_MIN)
__global__ void k_gather_i32_range(const int32_t* __restrict__ table, const int32_t* __restrict__ idx,
                                   int64_t n, int32_t* __restrict__ out, int32_t* __restrict__ range) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  int lo = 0x7fffffff, hi = (int)0x80000000;
  if (i < n) {
    const int v = table[idx[i]];
    out[i] = v;
    lo = v; hi = v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  // 1.6 M edges = 50 k warps hammering two addresses cost 79 us; the range converges after the first few warps, so
  // look (a plain, possibly stale read: the atomic re-checks) before touching it
  if ((threadIdx.x & 31) == 0 && lo <= hi) {
    if (lo < *reinterpret_cast<volatile int32_t*>(range)) atomicMin(range, lo);
    if (hi > *reinterpret_cast<volatile int32_t*>(range + 1)) atomicMax(range + 1, hi);
  }
}
__global__ void k_init_range(int32_t* range) {
  range[0] = 0x7fffffff;
  range[1] = (int)0x80000000;
}
__global__ void k_gather_rows_f32(const float* __restrict__ table, const int32_t* __restrict__ idx,
                                  int64_t n, int width, float* __restrict__ out) {
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n * width) return;
  int64_t r = i / width;
  int c = (int)(i - r * width);
  out[i] = table[(int64_t)idx[r] * width + c];
}
__global__ void k_narrow_i64(const int64_t* __restrict__ in, int64_t n, int32_t* __restrict__ out) {
  int64_t i = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) * 2;
  if (i + 1 < n) {
    longlong2 v = *reinterpret_cast<const longlong2*>(in + i);
    *reinterpret_cast<int2*>(out + i) = make_int2((int)v.x, (int)v.y);
  } else if (i < n) {
    out[i] = (int)in[i];
  }
}

int64_t scan_bsum_elems(int64_t n) { return ceil_div(n, kScanTile) + 1; }

int scan_inplace(int32_t* a, int64_t n, int32_t* bsum, cudaStream_t st) {
  int nb = (int)ceil_div(n, kScanTile);
  if (nb > 1) {
    k_scan_tiles<<<nb, kScanThreads, 0, st>>>(a, n, bsum, 0);
    k_scan_bsums<<<1, kScanThreads, 0, st>>>(bsum, nb);
  }
  k_scan_tiles<<<nb, kScanThreads, 0, st>>>(a, n, bsum, 1);
  note_launch(nb > 1 ? 3 : 1);
  return 0;
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int64_t b200gat_plan_scratch_elems(int64_t N, int64_t E) {
  int64_t nb = ceil_div(N + 1, kScanTile) + 1;
  return 4 * E + 4 * N + 2 * nb + 64;
}

extern "C" int b200gat_plan_build(const int64_t* edge_src, const int64_t* edge_dst, int64_t E,
                                  int64_t N, int32_t* in_ptr, int32_t* in_src, int32_t* in_eid,
                                  int32_t* out_ptr, int32_t* out_dst, int32_t* out_pos,
                                  int32_t* scratch, int32_t* stats, void* stream) {
  B200_CHECK_ARG(E >= 0 && N >= 0, "negative size");
  B200_CHECK_ARG(E < (1ll << 31) - 2048 && N < (1ll << 31) - 2048, "graph too large for int32 ids");
  B200_CHECK_ARG(in_ptr && out_ptr && stats && scratch, "null output pointer");
  B200_CHECK_ARG(E == 0 || (edge_src && edge_dst && in_src && in_eid && out_dst && out_pos),
                 "null edge pointer");
  const int vec = (reinterpret_cast<uintptr_t>(edge_src) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(edge_dst) & 15) == 0;
  cudaStream_t st = (cudaStream_t)stream;
  int64_t nb = ceil_div(N + 1, kScanTile) + 1;
  int32_t* tmp_in = scratch;
  int32_t* tmp_out = tmp_in + E;
  int32_t* out_eid = tmp_out + E;
  int32_t* inv = out_eid + E;
  int32_t* fill_in = inv + E;
  int32_t* fill_out = fill_in + N;
  int32_t* large_in = fill_out + N;
  int32_t* large_out = large_in + N;
  int32_t* bsum_in = large_out + N;
  int32_t* bsum_out = bsum_in + nb;
  int32_t* counters = bsum_out + nb;  // [0] n_large_in, [1] n_large_out

  ScopedTimer tm(B200GAT_TIMING_PLAN, st);
  cudaMemsetAsync(in_ptr, 0, sizeof(int32_t) * (N + 1), st);
  cudaMemsetAsync(out_ptr, 0, sizeof(int32_t) * (N + 1), st);
  cudaMemsetAsync(stats, 0, sizeof(int32_t) * 4, st);
  cudaMemsetAsync(fill_in, 0, sizeof(int32_t) * 2 * N, st);
  cudaMemsetAsync(counters, 0, sizeof(int32_t) * 8, st);
  if (E > 0) {
    int64_t pairs = ceil_div(E, 2);
    k_count<<<(unsigned)ceil_div(pairs, 256), 256, 0, st>>>(edge_src, edge_dst, E, (int32_t)N,
                                                            in_ptr, out_ptr, stats, vec);
    B200_CHECK_LAUNCH("k_count");
  }
  scan_inplace(in_ptr, N + 1, bsum_in, st);
  scan_inplace(out_ptr, N + 1, bsum_out, st);
  if (cudaGetLastError() != cudaSuccess) { set_error("k_scan: launch failed"); return B200GAT_ERR_CUDA; }
  if (E > 0) {
    unsigned eb = (unsigned)ceil_div(E, 256);
    k_fill<<<eb, 256, 0, st>>>(edge_src, edge_dst, E, (int32_t)N, in_ptr, out_ptr, fill_in,
                               fill_out, tmp_in, tmp_out);
    B200_CHECK_LAUNCH("k_fill");
    unsigned wb = (unsigned)ceil_div(N * 32, 256);
    k_sort_small<<<wb, 256, 0, st>>>(in_ptr, (int32_t)N, tmp_in, in_eid, large_in, counters + 0);
    k_sort_small<<<wb, 256, 0, st>>>(out_ptr, (int32_t)N, tmp_out, out_eid, large_out, counters + 1);
    k_sort_large<<<2 * kNumSMs, 256, 0, st>>>(in_ptr, tmp_in, in_eid, large_in, counters + 0);
    k_sort_large<<<2 * kNumSMs, 256, 0, st>>>(out_ptr, tmp_out, out_eid, large_out, counters + 1);
    B200_CHECK_LAUNCHES("k_sort", 4);
    k_finish_in<<<eb, 256, 0, st>>>(edge_src, E, in_ptr, (int32_t)N, in_eid, in_src, inv);
    k_finish_out<<<eb, 256, 0, st>>>(edge_dst, E, out_ptr, (int32_t)N, out_eid, inv, out_dst, out_pos);
    B200_CHECK_LAUNCHES("k_finish", 2);
  }
  if (N > 0) {
    k_degree_stats<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(in_ptr, out_ptr, (int32_t)N, stats);
    B200_CHECK_LAUNCH("k_degree_stats");
  }
  return B200GAT_OK;
}

extern "C" int64_t b200gat_node_order_scratch_elems(int64_t N) {
  int64_t nb = ceil_div(N + 1, kScanTile) + 1;
  return 5 * N + nb + 72;
}

extern "C" int b200gat_node_order(const int32_t* ptr, const int32_t* nbr, int64_t N, int32_t* order,
                                  int32_t* scratch, void* stream) {
  B200_CHECK_ARG(N >= 0, "negative size");
  if (N == 0) return B200GAT_OK;
  B200_CHECK_ARG(ptr && order && scratch, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  ScopedTimer tm(B200GAT_TIMING_PLAN, st);
  int64_t nb = ceil_div(N + 1, kScanTile) + 1;
  int32_t* key = scratch;
  int32_t* hist = key + N;          // [N+1]
  int32_t* fill = hist + N + 1;     // [N]
  int32_t* tmp = fill + N;          // [N]
  int32_t* large = tmp + N;         // [N]
  int32_t* bsum = large + N;        // [nb]
  int32_t* counter = bsum + nb;     // [1]
  cudaMemsetAsync(hist, 0, sizeof(int32_t) * (2 * N + 1), st);   // hist + fill
  cudaMemsetAsync(counter, 0, sizeof(int32_t) * 8, st);
  unsigned nbk = (unsigned)ceil_div(N, 256);
  k_order_key<<<nbk, 256, 0, st>>>(ptr, nbr, (int32_t)N, key, hist);
  scan_inplace(hist, N + 1, bsum, st);
  k_order_fill<<<nbk, 256, 0, st>>>(key, (int32_t)N, hist, fill, tmp);
  k_sort_small<<<(unsigned)ceil_div(N * 32, 256), 256, 0, st>>>(hist, (int32_t)N, tmp, order, large, counter);
  k_sort_large<<<2 * kNumSMs, 256, 0, st>>>(hist, tmp, order, large, counter);
  B200_CHECK_LAUNCHES("b200gat_node_order", 4);
  return B200GAT_OK;
}

extern "C" int b200gat_hub_list(const int32_t* ptr, int64_t N, int32_t hub_degree, int32_t* list,
                                int32_t* count, void* stream) {
  B200_CHECK_ARG(N >= 0 && hub_degree > 0, "bad arguments");
  B200_CHECK_ARG(count, "null count");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(count, 0, sizeof(int32_t), st);
  if (N == 0) return B200GAT_OK;
  B200_CHECK_ARG(ptr && list, "null pointer");
  k_hub_list<<<(unsigned)ceil_div(N, 256), 256, 0, st>>>(ptr, (int32_t)N, hub_degree, list, count);
  B200_CHECK_LAUNCH("k_hub_list");
  return B200GAT_OK;
}

extern "C" int b200gat_gather_i32(const int32_t* table, const int32_t* idx, int64_t n, int32_t* out,
                                  void* stream) {
  B200_CHECK_ARG(n >= 0, "negative size");
  if (n == 0) return B200GAT_OK;
  B200_CHECK_ARG(table && idx && out, "null pointer");
  k_gather_i32<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(table, idx, n, out);
  B200_CHECK_LAUNCH("k_gather_i32");
  return B200GAT_OK;
}

extern "C" int b200gat_gather_i32_range(const int32_t* table, const int32_t* idx, int64_t n, int32_t* out,
                                        int32_t* range, void* stream) {
  B200_CHECK_ARG(n >= 0, "negative size");
  B200_CHECK_ARG(range, "null range");
  k_init_range<<<1, 1, 0, (cudaStream_t)stream>>>(range);
  if (n == 0) {
    B200_CHECK_LAUNCH("k_init_range");
    return B200GAT_OK;
  }
  B200_CHECK_ARG(table && idx && out, "null pointer");
  k_gather_i32_range<<<(unsigned)ceil_div(n, 256), 256, 0, (cudaStream_t)stream>>>(table, idx, n, out, range);
  B200_CHECK_LAUNCHES("k_gather_i32_range", 2);
  return B200GAT_OK;
}

extern "C" int b200gat_gather_rows_f32(const float* table, const int32_t* idx, int64_t n, int32_t width,
                                       float* out, void* stream) {
  B200_CHECK_ARG(n >= 0 && width >= 0, "negative size");
  if (n == 0 || width == 0) return B200GAT_OK;
  B200_CHECK_ARG(table && idx && out, "null pointer");
  k_gather_rows_f32<<<(unsigned)ceil_div(n * width, 256), 256, 0, (cudaStream_t)stream>>>(table, idx, n,
                                                                                         width, out);
  B200_CHECK_LAUNCH("k_gather_rows_f32");
  return B200GAT_OK;
}

extern "C" int b200gat_narrow_i64(const int64_t* in, int64_t n, int32_t* out, void* stream) {
  B200_CHECK_ARG(n >= 0, "negative size");
  if (n == 0) return B200GAT_OK;
  B200_CHECK_ARG(in && out, "null pointer");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 7) == 0,
                 "misaligned pointer");
  k_narrow_i64<<<(unsigned)ceil_div(ceil_div(n, 2), 256), 256, 0, (cudaStream_t)stream>>>(in, n, out);
  B200_CHECK_LAUNCH("k_narrow_i64");
  return B200GAT_OK;
}
