// Clique-block detection and the residual CSR (b200gat_clique_build).
//
// A block is a node set S (min_size <= |S| <= B200GAT_MAX_CLIQUE) such that every v in S has, as ONE
// contiguous run of its destination-sorted in-edges, exactly the sources S \ {v} in ascending node id,
// all with the same edge type - the machine-conflict cliques of a schedule graph
// (jssp/utils/utils.py:460-478 generates them in that order).  Detection is exact:
//   runs    : per node, the longest run of same-typed in-edges with strictly ascending sources != v
//   insert  : commutative 64-bit hash of (run sources + v, type, length) into an open-addressing table;
//             slot -> (count, smallest member = leader)
//   verify  : every node of a slot compares its run with the member list derived from the leader's
//             run, element by element; any mismatch (or a hash collision) invalidates the slot
//   number  : blocks are numbered by ascending leader id (prefix sum) -> deterministic layout
// Everything outside the runs of valid members goes to a compact residual CSR (both directions,
// stable order, `pos` = position of the edge in the full destination-sorted order).
#include "common.cuh"

namespace b200gat {

int scan_inplace(int32_t* a, int64_t n, int32_t* bsum, cudaStream_t st);   // plan.cu
int64_t scan_bsum_elems(int64_t n);

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

constexpr unsigned kFull = 0xffffffffu;

// ---- runs: one warp per node ------------------------------------------------------------------
__global__ void k_clq_runs(const int32_t* __restrict__ in_ptr, const int32_t* __restrict__ in_src,
                           const int32_t* __restrict__ etype, int N, int min_run, int max_run,
                           int32_t* __restrict__ run_b, int32_t* __restrict__ run_len,
                           unsigned long long* __restrict__ key) {
  const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (v >= N) return;
  const int b = in_ptr[v], e = in_ptr[v + 1];
  int cur = b, best_b = b, best_len = 0;
  int carry_s = -1, carry_t = -1;
  for (int e0 = b; e0 < e; e0 += 32) {
    const int p = e0 + lane;
    const bool in = p < e;
    const int s = in ? in_src[p] : -1;
    const int t = (in && etype) ? etype[p] : 0;
    int ps = __shfl_up_sync(kFull, s, 1), pt = __shfl_up_sync(kFull, t, 1);
    if (lane == 0) { ps = carry_s; pt = carry_t; }
    const bool self = in && s == v;
    const bool brk = in && (p == b || t != pt || s <= ps || self || ps == v);
    unsigned mb = __ballot_sync(kFull, brk);
    const unsigned ms = __ballot_sync(kFull, self);
    while (mb) {
      const int i = __ffs(mb) - 1;
      mb &= mb - 1;
      const int pos = e0 + i, len = pos - cur;
      if (len > best_len) { best_len = len; best_b = cur; }
      cur = ((ms >> i) & 1u) ? pos + 1 : pos;
    }
    carry_s = __shfl_sync(kFull, s, 31);
    carry_t = __shfl_sync(kFull, t, 31);
  }
  if (e - cur > best_len) { best_len = e - cur; best_b = cur; }
  const bool ok = best_len >= min_run && best_len <= max_run;
  unsigned long long h = 0;
  if (ok) {
    for (int i = lane; i < best_len; i += 32) h += mix64((uint64_t)in_src[best_b + i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(kFull, h, o);
    h += mix64((uint64_t)v);
    const int t = etype ? etype[best_b] : 0;
    h = mix64(h ^ mix64(((uint64_t)(uint32_t)t << 32) | (uint32_t)best_len));
    if (h == 0) h = 1;
  }
  if (lane == 0) {
    run_b[v] = best_b;
    run_len[v] = ok ? best_len : 0;
    key[v] = h;
  }
}

// ---- insert: one thread per node ----------------------------------------------------------------
__global__ void k_clq_insert(const int32_t* __restrict__ run_len, const unsigned long long* __restrict__ key,
                             int N, unsigned ts_mask, unsigned long long* __restrict__ tkey,
                             int32_t* __restrict__ leader, int32_t* __restrict__ cnt,
                             int32_t* __restrict__ slot_of) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  if (run_len[v] == 0) { slot_of[v] = -1; return; }
  const unsigned long long k = key[v];
  unsigned slot = (unsigned)(k >> 17) & ts_mask;
  for (;;) {
    const unsigned long long old = atomicCAS(&tkey[slot], 0ull, k);
    if (old == 0ull || old == k) break;
    slot = (slot + 1) & ts_mask;
  }
  atomicMin(&leader[slot], v);
  atomicAdd(&cnt[slot], 1);
  slot_of[v] = (int)slot;
}

// ---- verify: one warp per node ------------------------------------------------------------------
__global__ void k_clq_verify(const int32_t* __restrict__ in_src, const int32_t* __restrict__ etype,
                             const int32_t* __restrict__ run_b, const int32_t* __restrict__ run_len, int N,
                             const int32_t* __restrict__ slot_of, const int32_t* __restrict__ leader,
                             const int32_t* __restrict__ cnt, int32_t* __restrict__ bad,
                             int32_t* __restrict__ rank) {
  const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (v >= N) return;
  const int len = run_len[v];
  if (len == 0) return;
  const int slot = slot_of[v], g = cnt[slot], L = leader[slot];
  const int bv = run_b[v], bL = run_b[L];
  bool ok = (g == len + 1) && (run_len[L] == len);
  if (ok && etype) ok = etype[bv] == etype[bL];
  int r = 0;
  if (ok && v != L) {                       // rank of v = 1 + its index in the leader's run
    r = -1;
    for (int i0 = 0; i0 < len; i0 += 32) {
      const int i = i0 + lane;
      const unsigned hit = __ballot_sync(kFull, i < len && in_src[bL + i] == v);
      if (hit) { r = i0 + __ffs(hit); break; }
    }
    ok = r > 0;
  }
  if (ok) {
    if (v == L) ok = in_src[bL] > L;        // members ascending: leader is the smallest
    bool mine = true;
    for (int i = lane; i < g; i += 32) {
      if (i == r) continue;
      const int Mi = (i == 0) ? L : in_src[bL + i - 1];
      const int got = in_src[bv + i - (i > r ? 1 : 0)];
      mine = mine && (Mi == got);
    }
    ok = ok && __all_sync(kFull, mine);
  }
  if (lane == 0) {
    if (!ok) atomicExch(&bad[slot], 1);
    rank[v] = ok ? r : 0;
  }
}

// lead[v+1] = 1 when v is the leader of a valid slot (then scanned -> block ids)
__global__ void k_clq_leaders(const int32_t* __restrict__ run_len, const int32_t* __restrict__ slot_of,
                              const int32_t* __restrict__ leader, const int32_t* __restrict__ bad, int N,
                              int32_t* __restrict__ lead) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  int f = 0;
  if (run_len[v] > 0) {
    const int slot = slot_of[v];
    f = (!bad[slot] && leader[slot] == v) ? 1 : 0;
  }
  lead[v + 1] = f;
  if (v == 0) lead[0] = 0;
}

__global__ void k_clq_assign(const int32_t* __restrict__ run_b, const int32_t* __restrict__ run_len,
                             const int32_t* __restrict__ etype, const int32_t* __restrict__ slot_of,
                             const int32_t* __restrict__ leader, const int32_t* __restrict__ bad,
                             const int32_t* __restrict__ lead, int N, int32_t* __restrict__ node_blk,
                             int32_t* __restrict__ node_run, int2* __restrict__ runinfo,
                             int32_t* __restrict__ sizes, int32_t* __restrict__ soff,
                             int32_t* __restrict__ blk_type, int32_t* __restrict__ counts) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  int blk = -1, len = 0;
  if (run_len[v] > 0) {
    const int slot = slot_of[v];
    if (!bad[slot]) {
      const int L = leader[slot];
      blk = lead[L];                         // exclusive prefix = number of leaders below L
      len = run_len[v];
      if (v == L) {
        const int g = len + 1;
        sizes[blk + 1] = g;
        soff[blk + 1] = g * ((g + 3) & ~3);
        blk_type[blk] = etype ? etype[run_b[v]] : 0;
        atomicMax(&counts[1], g);
      }
    }
  }
  node_blk[v] = blk;
  node_run[v] = blk >= 0 ? run_b[v] : -1;
  runinfo[v] = make_int2(run_b[v], len);
  if (v == 0) counts[0] = lead[N];
}

__global__ void k_clq_members(const int32_t* __restrict__ node_blk, const int32_t* __restrict__ rank,
                              const int32_t* __restrict__ blk_ptr, const int32_t* __restrict__ blk_soff, int N,
                              int32_t* __restrict__ blk_nodes, int32_t* __restrict__ counts) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  const int b = node_blk[v];
  if (b >= 0) blk_nodes[blk_ptr[b] + rank[v]] = v;
  if (v == 0) {
    counts[3] = blk_ptr[counts[0]];
    counts[4] = blk_soff[counts[0]];
  }
}

// ---- residual CSR ---------------------------------------------------------------------------------
__global__ void k_res_count_in(const int32_t* __restrict__ in_ptr, const int2* __restrict__ runinfo, int N,
                               int32_t* __restrict__ r_in_ptr) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= N) return;
  r_in_ptr[v + 1] = in_ptr[v + 1] - in_ptr[v] - runinfo[v].y;
  if (v == 0) r_in_ptr[0] = 0;
}
__device__ __forceinline__ bool in_block_run(const int2* __restrict__ runinfo, int dst, int pos) {
  const int2 ri = __ldg(runinfo + dst);
  return (unsigned)(pos - ri.x) < (unsigned)ri.y;
}
__global__ void k_res_count_out(const int32_t* __restrict__ out_ptr, const int32_t* __restrict__ out_dst,
                                const int32_t* __restrict__ out_pos, const int2* __restrict__ runinfo, int N,
                                int32_t* __restrict__ r_out_ptr) {
  const int u = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (u >= N) return;
  const int b = out_ptr[u], e = out_ptr[u + 1];
  int c = 0;
  for (int q = b + lane; q < e; q += 32) c += in_block_run(runinfo, out_dst[q], out_pos[q]) ? 0 : 1;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(kFull, c, o);
  if (lane == 0) {
    r_out_ptr[u + 1] = c;
    if (u == 0) r_out_ptr[0] = 0;
  }
}
__global__ void k_res_fill_in(const int32_t* __restrict__ in_ptr, const int32_t* __restrict__ in_src,
                              const int2* __restrict__ runinfo, const int32_t* __restrict__ r_in_ptr, int N,
                              int32_t* __restrict__ r_in_src, int32_t* __restrict__ r_in_pos,
                              int32_t* __restrict__ counts) {
  const int v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (v >= N) return;
  const int b = in_ptr[v], e = in_ptr[v + 1];
  const int2 ri = runinfo[v];
  const int rb = ri.x, re = ri.x + ri.y, ob = r_in_ptr[v];
  for (int p = b + lane; p < e; p += 32) {
    if (p >= rb && p < re) continue;
    const int j = ob + (p - b) - (p >= re ? ri.y : 0);
    r_in_src[j] = in_src[p];
    r_in_pos[j] = p;
  }
  if (v == N - 1 && lane == 0) counts[2] = r_in_ptr[N];
}
__global__ void k_res_fill_out(const int32_t* __restrict__ out_ptr, const int32_t* __restrict__ out_dst,
                               const int32_t* __restrict__ out_pos, const int2* __restrict__ runinfo,
                               const int32_t* __restrict__ r_out_ptr, int N, int32_t* __restrict__ r_out_dst,
                               int32_t* __restrict__ r_out_pos) {
  const int u = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (u >= N) return;
  const int b = out_ptr[u], e = out_ptr[u + 1];
  int o = r_out_ptr[u];
  for (int q0 = b; q0 < e; q0 += 32) {
    const int q = q0 + lane;
    int d = 0, p = 0;
    bool keep = false;
    if (q < e) {
      d = out_dst[q];
      p = out_pos[q];
      keep = !in_block_run(runinfo, d, p);
    }
    const unsigned m = __ballot_sync(kFull, keep);
    if (keep) {
      const int j = o + __popc(m & ((1u << lane) - 1u));
      r_out_dst[j] = d;
      r_out_pos[j] = p;
    }
    o += __popc(m);
  }
}

}  // namespace b200gat

using namespace b200gat;

static int64_t table_size(int64_t N) {
  int64_t ts = 64;
  while (ts < 2 * N) ts <<= 1;
  return ts;
}

extern "C" int64_t b200gat_clique_scratch_elems(int64_t N, int64_t E) {
  (void)E;
  const int64_t ts = table_size(N);
  // key[N] u64, tkey[ts] u64, leader/cnt/bad [ts], run_b/run_len/slot_of [N], lead/sizes/soff [N+1],
  // runinfo [N] int2, scan block sums
  return 2 * N + 2 * ts + 3 * ts + 3 * N + 3 * (N + 1) + 2 * N + 3 * scan_bsum_elems(N + 1) + 64;
}

extern "C" int b200gat_clique_build(const b200gat_graph* g, const int32_t* etype_sorted, int32_t min_size,
                                    int32_t* node_blk, int32_t* node_run, int32_t* node_rank,
                                    int32_t* blk_ptr, int32_t* blk_nodes, int32_t* blk_type,
                                    int32_t* blk_soff, int32_t* r_in_ptr, int32_t* r_in_src,
                                    int32_t* r_in_pos, int32_t* r_out_ptr, int32_t* r_out_dst,
                                    int32_t* r_out_pos, int32_t* counts, int32_t* scratch, void* stream) {
  B200_CHECK_ARG(g, "null graph");
  const int64_t N = g->n_nodes, E = g->n_edges;
  B200_CHECK_ARG(N >= 0 && E >= 0, "negative size");
  B200_CHECK_ARG(min_size >= 3 && min_size <= B200GAT_MAX_CLIQUE, "min_size out of range");
  B200_CHECK_ARG(counts && scratch, "null counts/scratch");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(counts, 0, sizeof(int32_t) * 8, st);
  if (N == 0) return B200GAT_OK;
  B200_CHECK_ARG(g->in_ptr && g->out_ptr && node_blk && node_run && node_rank && blk_ptr && blk_nodes &&
                     blk_type && blk_soff && r_in_ptr && r_out_ptr,
                 "null pointer");
  B200_CHECK_ARG(E == 0 || (g->in_src && g->out_dst && g->out_pos && r_in_src && r_in_pos && r_out_dst &&
                            r_out_pos),
                 "null edge pointer");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(scratch) & 7) == 0, "scratch must be 8-byte aligned");
  ScopedTimer tm(B200GAT_TIMING_PLAN, st);
  const int64_t ts = table_size(N), nbs = scan_bsum_elems(N + 1);
  unsigned long long* key = reinterpret_cast<unsigned long long*>(scratch);
  unsigned long long* tkey = key + N;
  int2* runinfo = reinterpret_cast<int2*>(tkey + ts);
  int32_t* leader = reinterpret_cast<int32_t*>(runinfo + N);
  int32_t* cnt = leader + ts;
  int32_t* bad = cnt + ts;
  int32_t* run_b = bad + ts;
  int32_t* run_len = run_b + N;
  int32_t* slot_of = run_len + N;
  int32_t* lead = slot_of + N;          // [N+1]
  int32_t* sizes = lead + N + 1;        // [N+1]
  int32_t* soff = sizes + N + 1;        // [N+1]
  int32_t* bsum = soff + N + 1;         // [3 * nbs]

  cudaMemsetAsync(tkey, 0, sizeof(unsigned long long) * ts, st);
  cudaMemsetAsync(leader, 0x7f, sizeof(int32_t) * ts, st);
  cudaMemsetAsync(cnt, 0, sizeof(int32_t) * 2 * ts, st);                 // cnt + bad
  cudaMemsetAsync(sizes, 0, sizeof(int32_t) * 2 * (N + 1), st);          // sizes + soff
  const unsigned tb = (unsigned)ceil_div(N, 256), wb = (unsigned)ceil_div(N * 32, 256);
  k_clq_runs<<<wb, 256, 0, st>>>(g->in_ptr, g->in_src, etype_sorted, (int)N, min_size - 1,
                                 B200GAT_MAX_CLIQUE - 1, run_b, run_len, key);
  k_clq_insert<<<tb, 256, 0, st>>>(run_len, key, (int)N, (unsigned)(ts - 1), tkey, leader, cnt, slot_of);
  k_clq_verify<<<wb, 256, 0, st>>>(g->in_src, etype_sorted, run_b, run_len, (int)N, slot_of, leader, cnt,
                                   bad, node_rank);
  k_clq_leaders<<<tb, 256, 0, st>>>(run_len, slot_of, leader, bad, (int)N, lead);
  B200_CHECK_LAUNCHES("k_clq_detect", 4);
  scan_inplace(lead, N + 1, bsum, st);
  k_clq_assign<<<tb, 256, 0, st>>>(run_b, run_len, etype_sorted, slot_of, leader, bad, lead, (int)N, node_blk,
                                   node_run, runinfo, sizes, soff, blk_type, counts);
  B200_CHECK_LAUNCH("k_clq_assign");
  scan_inplace(sizes, N + 1, bsum + nbs, st);
  scan_inplace(soff, N + 1, bsum + 2 * nbs, st);
  cudaMemcpyAsync(blk_ptr, sizes, sizeof(int32_t) * (N + 1), cudaMemcpyDeviceToDevice, st);
  cudaMemcpyAsync(blk_soff, soff, sizeof(int32_t) * (N + 1), cudaMemcpyDeviceToDevice, st);
  k_clq_members<<<tb, 256, 0, st>>>(node_blk, node_rank, blk_ptr, blk_soff, (int)N, blk_nodes, counts);
  // residual CSR
  k_res_count_in<<<tb, 256, 0, st>>>(g->in_ptr, runinfo, (int)N, r_in_ptr);
  k_res_count_out<<<wb, 256, 0, st>>>(g->out_ptr, g->out_dst, g->out_pos, runinfo, (int)N, r_out_ptr);
  B200_CHECK_LAUNCHES("k_res_count", 3);
  scan_inplace(r_in_ptr, N + 1, bsum, st);
  scan_inplace(r_out_ptr, N + 1, bsum + nbs, st);
  k_res_fill_in<<<wb, 256, 0, st>>>(g->in_ptr, g->in_src, runinfo, r_in_ptr, (int)N, r_in_src, r_in_pos,
                                    counts);
  k_res_fill_out<<<wb, 256, 0, st>>>(g->out_ptr, g->out_dst, g->out_pos, runinfo, r_out_ptr, (int)N,
                                     r_out_dst, r_out_pos);
  B200_CHECK_LAUNCHES("k_res_fill", 2);
  return B200GAT_OK;
}
