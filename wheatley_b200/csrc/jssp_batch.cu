// Device-side JSSP batch builder (include/b200gat.h "batch builder").
//
// Edge layout, identical to the reference's host builder (jssp/models/agent_observation.py:110-229 followed by
// jssp/models/gnn_dgl.py:263-286), graph after graph:
//   [ precedence fwd (ne) | precedence rev (ne) | clique conflicts (K) | self loops (nn) ]
// then once for the batch:
//   [ pool self loops (B) | pool -> node (N, learninv only) | node -> pool (N) ]
// Clique edges of one graph are the row-major non-zeros of cond[r, c] = (mid[c] == mid[r]) & (r != c), mid != -1,
// emitted as (src = r, dst = c) (jssp/utils/utils.py:460-478): rows ascending, columns ascending inside a row.
//
// HBM-bound byte work: 24 bytes written per edge, nothing re-read from HBM (the machine ids of a graph sit in
// shared memory).  One warp per row r scans the graph's machine ids 32 at a time and writes the matching columns
// in order with a ballot prefix; CTAs take 32 rows of one graph and recompute the graph's machine histogram and
// their rows' offsets locally (a few thousand shared-memory operations) instead of a separate scan kernel.
#include "common.cuh"

namespace b200gat {
namespace {

constexpr int kThreads = 256;
constexpr int kRowsPerCta = 32;

struct ObsArgs {
  int B, max_nodes, F, max_edges;
  const float* features;
  const int64_t* edge_index;
  const int64_t* meta;
  int clique, mid_in_edges, bidir, self_loops, no_tct, pooling, n_mid;
  int64_t N, Eg;
  int node_mode, M;
  int64_t A;
};

__global__ void __launch_bounds__(kThreads)
k_jssp_prec(ObsArgs a, int64_t* __restrict__ src, int64_t* __restrict__ dst, int64_t* __restrict__ et) {
  const int g = blockIdx.y;
  const int64_t ne = a.meta[2 * (a.B + 1) + g];
  const int64_t j = (int64_t)blockIdx.x * kThreads + threadIdx.x;
  if (j >= ne) return;
  const int64_t noff = a.meta[g], base = a.meta[(a.B + 1) + g];
  const int64_t* ei = a.edge_index + (int64_t)g * 2 * a.max_edges;
  const int64_t u = ei[j] + noff, v = ei[a.max_edges + j] + noff;
  src[base + j] = u;
  dst[base + j] = v;
  et[base + j] = 1;
  if (a.bidir) {
    src[base + ne + j] = v;
    dst[base + ne + j] = u;
    et[base + ne + j] = 2;
  }
}

// grid (ceil(max_nodes / kRowsPerCta), B); dynamic smem: mid [max_nodes] + hist [n_mid] + rowoff [kRowsPerCta + 1]
__global__ void __launch_bounds__(kThreads)
k_jssp_nodes(ObsArgs a, int64_t* __restrict__ src, int64_t* __restrict__ dst, int64_t* __restrict__ et,
             float* __restrict__ feat) {
  extern __shared__ int smem_i[];
  int* mid = smem_i;
  int* hist = mid + a.max_nodes;
  int* rowoff = hist + a.n_mid;
  __shared__ int red[kThreads / 32];
  const int g = blockIdx.y, t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t noff = a.meta[g];
  const int nn = (int)(a.meta[g + 1] - noff);
  const int r0 = blockIdx.x * kRowsPerCta;
  if (r0 >= nn && !(r0 == 0)) return;                       // (an empty graph still owns its pool node)
  const int64_t ne = a.meta[2 * (a.B + 1) + g];
  const int64_t K = a.clique ? a.meta[3 * (a.B + 1) + g] : 0;
  const int64_t ebase = a.meta[(a.B + 1) + g];
  const int64_t clq_base = ebase + (a.bidir ? 2 : 1) * ne;
  const int64_t self_base = clq_base + K;
  const float* fg = a.features + (int64_t)g * a.max_nodes * a.F;

  const bool need_mid = a.clique || a.node_mode;
  if (need_mid) {
    for (int i = t; i < a.n_mid; i += kThreads) hist[i] = 0;
    __syncthreads();
    for (int i = t; i < nn; i += kThreads) {
      const int m = (int)fg[(int64_t)i * a.F + 6];          // .long() of the float id: truncation
      mid[i] = m;
      if (m >= 0 && m < a.n_mid) atomicAdd(&hist[m], 1);
    }
    __syncthreads();
    // clique edges of the rows before this CTA's first row
    int part = 0;
    for (int i = t; i < r0 && i < nn; i += kThreads) {
      const int m = mid[i];
      part += (m >= 0 && m < a.n_mid) ? (a.clique ? hist[m] - 1 : 1) : 0;
    }
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) red[warp] = part;
    __syncthreads();
    if (t == 0) {
      int acc = 0;
      for (int w = 0; w < kThreads / 32; ++w) acc += red[w];
      for (int k = 0; k < kRowsPerCta; ++k) {
        rowoff[k] = acc;
        const int r = r0 + k;
        if (r < nn) {
          const int m = mid[r];
          acc += (m >= 0 && m < a.n_mid) ? (a.clique ? hist[m] - 1 : 1) : 0;
        }
      }
      rowoff[kRowsPerCta] = acc;
    }
    __syncthreads();
  }

  const int64_t pool_self = a.Eg;                                     // [B]
  const int64_t pool_to_node = a.Eg + a.B;                            // [N] (learninv)
  const int64_t node_to_pool = pool_to_node + (a.pooling == B200GAT_POOL_LEARNINV ? a.N : 0);
  const int64_t pool_id = a.N + g;
  // machine-node block (conflicts == "node"), after the pool wiring
  const int64_t mach_blk = a.pooling != B200GAT_POOL_NONE ? node_to_pool + a.N : a.Eg;
  const int64_t MB = (int64_t)a.M * a.B;
  const int64_t mach_node0 = a.N + (a.pooling != B200GAT_POOL_NONE ? a.B : 0) + (int64_t)g * a.M;

  for (int k = warp; k < kRowsPerCta; k += kThreads / 32) {
    const int r = r0 + k;
    if (r >= nn) break;
    const int64_t gid = noff + r;
    if (a.clique) {
      const int m = mid[r];
      if (m >= 0 && m < a.n_mid) {
        int64_t pos = clq_base + rowoff[k];
        const int64_t ty = a.mid_in_edges ? (int64_t)m + 5 : 5;
        for (int c0 = 0; c0 < nn; c0 += 32) {
          const int c = c0 + lane;
          const bool hit = c < nn && c != r && mid[c] == m;
          const unsigned bal = __ballot_sync(0xffffffffu, hit);
          if (hit) {
            const int64_t p = pos + __popc(bal & ((1u << lane) - 1u));
            src[p] = gid;
            dst[p] = noff + c;
            et[p] = ty;
          }
          pos += __popc(bal);
        }
      }
    }
    if (lane == 0 && a.node_mode) {
      const int m = mid[r];
      if (m >= 0 && m < a.n_mid) {
        const int64_t p = mach_blk + MB + a.meta[3 * (a.B + 1) + g] + rowoff[k];
        src[p] = gid;
        dst[p] = mach_node0 + m;
        et[p] = (int64_t)m + 5;
        src[p + a.A] = mach_node0 + m;
        dst[p + a.A] = gid;
        et[p + a.A] = (int64_t)m + 5 + a.M;
      }
    }
    if (lane == 0) {
      if (a.self_loops) {
        src[self_base + r] = gid;
        dst[self_base + r] = gid;
        et[self_base + r] = 0;
      }
      if (a.pooling != B200GAT_POOL_NONE) {
        src[node_to_pool + gid] = gid;
        dst[node_to_pool + gid] = pool_id;
        et[node_to_pool + gid] = 4;
        if (a.pooling == B200GAT_POOL_LEARNINV) {
          src[pool_to_node + gid] = pool_id;
          dst[pool_to_node + gid] = gid;
          et[pool_to_node + gid] = 3;
        }
      }
    }
    // feature row with the reference's fix-ups (agent_observation.py:133-139)
    const float* fr = fg + (int64_t)r * a.F;
    float* out = feat + gid * a.F;
    const float f1 = fr[1], f4 = fr[4];
    for (int c = lane; c < a.F; c += 32) {
      float v = fr[c];
      if (c == 1) v = (f1 == -1.f) ? -1.f : f4;
      if (a.no_tct && c >= 1 && c <= 4) v = -1.f;
      out[c] = v;
    }
  }
  if (blockIdx.x == 0 && a.pooling != B200GAT_POOL_NONE) {
    if (t == 0) {
      src[pool_self + g] = pool_id;
      dst[pool_self + g] = pool_id;
      et[pool_self + g] = 0;
    }
    for (int c = t; c < a.F; c += kThreads) feat[pool_id * a.F + c] = 0.f;   // zero row (gnn_dgl.py:266-270)
  }
  if (blockIdx.x == 0 && a.node_mode) {
    for (int j = t; j < a.M; j += kThreads) {
      const int64_t mn = mach_node0 + j, e = mach_blk + (int64_t)g * a.M + j;
      src[e] = mn;
      dst[e] = mn;
      et[e] = 0;
      if (a.pooling == B200GAT_POOL_LEARN) {
        const int64_t q = mach_blk + MB + 2 * a.A + (int64_t)g * a.M + j;
        src[q] = mn;
        dst[q] = pool_id;
        et[q] = 5 + 2 * (int64_t)a.M;
      }
    }
    for (int i = t; i < a.M * a.F; i += kThreads) feat[mach_node0 * a.F + i] = 0.f;
  }
}

int fill(ObsArgs& a, const b200gat_jssp_obs* o) {
  B200_CHECK_ARG(o, "null observation descriptor");
  B200_CHECK_ARG(o->n_graphs >= 0 && o->max_nodes >= 0 && o->max_edges >= 0 && o->n_feat >= 0, "negative size");
  B200_CHECK_ARG(o->pooling >= B200GAT_POOL_NONE && o->pooling <= B200GAT_POOL_LEARNINV, "unknown pooling");
  B200_CHECK_ARG(o->n_nodes_total >= 0 && o->n_edges_graphs >= 0, "negative totals");
  a.B = o->n_graphs; a.max_nodes = o->max_nodes; a.F = o->n_feat; a.max_edges = o->max_edges;
  a.features = o->features; a.edge_index = o->edge_index; a.meta = o->meta;
  a.clique = o->clique_conflicts != 0; a.mid_in_edges = o->mid_in_edges != 0; a.bidir = o->bidir != 0;
  a.self_loops = o->self_loops != 0; a.no_tct = o->no_tct != 0; a.pooling = o->pooling;
  a.n_mid = o->n_machine_ids > 0 ? o->n_machine_ids : 0;
  a.N = o->n_nodes_total; a.Eg = o->n_edges_graphs;
  a.node_mode = o->node_conflicts != 0; a.M = a.node_mode ? o->max_n_machines : 0; a.A = a.node_mode ? o->n_affected : 0;
  B200_CHECK_ARG(!(a.clique && a.node_mode), "clique and node conflicts are exclusive");
  B200_CHECK_ARG(!a.node_mode || (a.M >= a.n_mid && a.A >= 0), "machine ids exceed max_n_machines");
  return B200GAT_OK;
}

}  // namespace
}  // namespace b200gat

using namespace b200gat;

extern "C" int64_t b200gat_jssp_batch_edges(const b200gat_jssp_obs* o) {
  if (!o) return -1;
  int64_t e = o->n_edges_graphs;
  if (o->pooling != B200GAT_POOL_NONE) e += o->n_graphs + o->n_nodes_total;
  if (o->pooling == B200GAT_POOL_LEARNINV) e += o->n_nodes_total;
  if (o->node_conflicts) {
    const int64_t MB = (int64_t)o->max_n_machines * o->n_graphs;
    e += MB + 2 * o->n_affected + (o->pooling == B200GAT_POOL_LEARN ? MB : 0);
  }
  return e;
}

extern "C" int b200gat_jssp_batch_build(const b200gat_jssp_obs* o, int64_t* src, int64_t* dst, int64_t* etype,
                                        float* feat, void* stream) {
  ObsArgs a;
  int rc = fill(a, o);
  if (rc) return rc;
  if (a.B == 0) return B200GAT_OK;
  B200_CHECK_ARG(a.features && a.meta && feat, "null pointer");
  B200_CHECK_ARG(a.F >= 7 || !(a.clique || a.node_mode), "conflicts need the machine id in feature column 6");
  B200_CHECK_ARG(a.F >= 5, "the feature fix-up reads columns 1 and 4");
  B200_CHECK_ARG(b200gat_jssp_batch_edges(o) == 0 || (src && dst && etype), "null edge output");
  B200_CHECK_ARG(a.max_edges == 0 || a.edge_index, "null edge_index");
  cudaStream_t st = (cudaStream_t)stream;
  ScopedTimer tm(B200GAT_TIMING_BATCH, st);
  if (a.max_edges > 0) {
    dim3 grid((unsigned)ceil_div(a.max_edges, kThreads), (unsigned)a.B);
    k_jssp_prec<<<grid, kThreads, 0, st>>>(a, src, dst, etype);
    B200_CHECK_LAUNCH("k_jssp_prec");
  }
  const size_t smem = sizeof(int) * ((size_t)((a.clique || a.node_mode) ? a.max_nodes + a.n_mid : 0) + kRowsPerCta + 1);
  if (smem > 200 * 1024) {
    set_error("b200gat_jssp_batch_build: %d nodes per graph + %d machine ids do not fit in shared memory",
              a.max_nodes, a.n_mid);
    return B200GAT_ERR_UNSUPPORTED;
  }
  static size_t granted = 48 * 1024;
  if (smem > granted) {
    if (cudaFuncSetAttribute(k_jssp_nodes, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
      set_error("k_jssp_nodes: cudaFuncSetAttribute(%zu) failed", smem);
      return B200GAT_ERR_CUDA;
    }
    granted = smem;
  }
  dim3 grid((unsigned)(a.max_nodes > 0 ? ceil_div(a.max_nodes, kRowsPerCta) : 1), (unsigned)a.B);
  k_jssp_nodes<<<grid, kThreads, smem, st>>>(a, src, dst, etype, feat);
  B200_CHECK_LAUNCH("k_jssp_nodes");
  return B200GAT_OK;
}
