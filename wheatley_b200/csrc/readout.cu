// Trunk readout: concat-all-layers + per-graph pooling + pad-to-max_n_nodes in one pass.
//
// The reference ends every trunk forward with torch.cat over the L+2 per-layer node tensors
// (psp/models/gnn_mp.py:419-425, jssp/models/gnn_dgl.py:401), a Python loop over the graphs that slices,
// pools (torch.max / matmul with 1/n) and F.pad's each graph (gnn_mp.py:429-470) and a torch.stack - three
// copies of the [N, F] tensor plus 3B small launches.  Here the sources are read where they lie:
//   k_readout_rows : node_features[b, r, :] = concat_i src_i[row_start[b] + r, :]   (r < n_b; zero rows after)
//                    graph_embedding[b, :]  = concat_i src_i[pool_row0 + b, :]      (learned pool node rows)
//   k_readout_pool : graph_embedding[b, c]  = max_r / mean_r node_features[b, r, c] (argmax = first maximal row,
//                    like torch.max(dim=0) on the CPU, which is what the reference's gradients follow)
//   k_readout_bwd  : every source's gradient, written completely (zeros for rows nothing reads)
#include <math.h>
#include "common.cuh"

namespace b200gat {

struct ReadoutSrc {
  const float* p[B200GAT_READOUT_MAX_SRC];
  float* g[B200GAT_READOUT_MAX_SRC];
  int64_t ld[B200GAT_READOUT_MAX_SRC];
  int32_t width[B200GAT_READOUT_MAX_SRC];
  int32_t off[B200GAT_READOUT_MAX_SRC + 1];
  int32_t n;
};

__global__ void __launch_bounds__(128) k_readout_rows(ReadoutSrc S, const int32_t* __restrict__ row_start, int B,
                                                      int max_n, int64_t pool_row0, int ld, float* __restrict__ nf,
                                                      float* __restrict__ ge) {
  const int F = S.off[S.n];
  int64_t row;          // source row, -1 = padding
  float* out;
  if ((int)blockIdx.x < B * max_n) {
    const int b = blockIdx.x / max_n, r = blockIdx.x - b * max_n;
    const int s0 = row_start[b], n_b = row_start[b + 1] - s0;
    row = r < n_b ? (int64_t)s0 + r : -1;
    out = nf + (int64_t)blockIdx.x * ld;
  } else {              // learned pooling: the pool node's row is the graph embedding
    const int b = blockIdx.x - B * max_n;
    row = pool_row0 + b;
    out = ge + (int64_t)b * ld;
  }
  if (row < 0) {
    for (int c = threadIdx.x; c < ld; c += 128) out[c] = 0.f;
    return;
  }
  if (threadIdx.x < ld - F) out[F + threadIdx.x] = 0.f;      // pad columns (ld - F < 4)
#pragma unroll 1
  for (int i = 0; i < S.n; ++i) {
    const float* __restrict__ src = S.p[i] + row * S.ld[i];
    float* __restrict__ o = out + S.off[i];
    for (int c = threadIdx.x; c < S.width[i]; c += 128) o[c] = __ldg(src + c);
  }
}

// block (32 columns, RY row lanes); grid (ceil(F / 32), B)
template <int RY>
__global__ void __launch_bounds__(32 * RY) k_readout_pool(const float* __restrict__ nf,
                                                          const int32_t* __restrict__ row_start, int max_n, int F,
                                                          int ld, int mode, float* __restrict__ ge,
                                                          int32_t* __restrict__ argmax) {
  __shared__ float sv[RY][33];
  __shared__ int si[RY][33];
  const int b = blockIdx.y, c = blockIdx.x * 32 + threadIdx.x, ty = threadIdx.y;
  const int n_b = row_start[b + 1] - row_start[b];
  const float* base = nf + (int64_t)b * max_n * ld;
  float best = -INFINITY, sum = 0.f;
  int bi = 0x7fffffff;
  if (c < F) {
#pragma unroll 4
    for (int r = ty; r < n_b; r += RY) {
      const float v = base[(int64_t)r * ld + c];
      sum += v;
      if (v > best || (bi == 0x7fffffff)) { best = v; bi = r; }      // strict >: first index inside a lane
    }
  }
  sv[ty][threadIdx.x] = mode == B200GAT_READOUT_MAX ? best : sum;
  si[ty][threadIdx.x] = bi;
  __syncthreads();
  if (ty == 0 && c < F) {
    if (mode == B200GAT_READOUT_MAX) {
      for (int k = 1; k < RY; ++k) {
        const float v = sv[k][threadIdx.x];
        const int i = si[k][threadIdx.x];
        if (i != 0x7fffffff && (bi == 0x7fffffff || v > best || (v == best && i < bi))) { best = v; bi = i; }
      }
      ge[(int64_t)b * ld + c] = best;
      argmax[(int64_t)b * F + c] = bi == 0x7fffffff ? -1 : bi;
    } else {
      // rows are summed lane-major then across lanes in lane order: fixed order, bit-reproducible
      for (int k = 1; k < RY; ++k) sum += sv[k][threadIdx.x];
      ge[(int64_t)b * ld + c] = n_b > 0 ? sum / (float)n_b : 0.f;
    }
  }
  if (ty == 0 && c >= F && c < ld) ge[(int64_t)b * ld + c] = 0.f;       // pad columns
}

__global__ void __launch_bounds__(128) k_readout_bwd(ReadoutSrc S, const int32_t* __restrict__ row_start, int B,
                                                     int max_n, int64_t n_task_rows, int64_t pool_row0, int mode,
                                                     const float* __restrict__ g_nf, int64_t ld_nf,
                                                     const float* __restrict__ g_ge, int64_t ld_ge,
                                                     const int32_t* __restrict__ argmax) {
  const int F = S.off[S.n];
  const int64_t row = blockIdx.x;
  __shared__ int s_b;
  const float* gn = nullptr;      // gradient row of node_features
  const float* gg = nullptr;      // gradient row of graph_embedding that reaches this source row
  const int32_t* am = nullptr;
  int local = 0;
  float scale = 1.f;
  if (row < n_task_rows) {
    if (threadIdx.x == 0) {       // graph of this row: last b with row_start[b] <= row
      int lo = 0, hi = B - 1;
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (row_start[mid] <= row) lo = mid; else hi = mid - 1;
      }
      s_b = lo;
    }
    __syncthreads();
    const int b = s_b;
    local = (int)(row - row_start[b]);
    if (g_nf) gn = g_nf + ((int64_t)b * max_n + local) * ld_nf;
    if (g_ge && mode == B200GAT_READOUT_MAX) { gg = g_ge + (int64_t)b * ld_ge; am = argmax + (int64_t)b * F; }
    if (g_ge && mode == B200GAT_READOUT_AVG) {
      gg = g_ge + (int64_t)b * ld_ge;
      scale = 1.f / (float)(row_start[b + 1] - row_start[b]);
    }
  } else if (mode == B200GAT_READOUT_ROWS && g_ge && row >= pool_row0 && row < pool_row0 + B) {
    gg = g_ge + (row - pool_row0) * ld_ge;
  }
#pragma unroll 1
  for (int i = 0; i < S.n; ++i) {
    if (!S.g[i]) continue;
    float* __restrict__ o = S.g[i] + row * S.width[i];
    const int off = S.off[i];
    for (int c = threadIdx.x; c < S.width[i]; c += 128) {
      float v = gn ? gn[off + c] : 0.f;
      if (gg) {
        if (am) { if (am[off + c] == local) v += gg[off + c]; }
        else v = fmaf(gg[off + c], scale, v);
      }
      o[c] = v;
    }
  }
}

static int fill_src(ReadoutSrc& S, const b200gat_readout* r, float* const* g_src) {
  B200_CHECK_ARG(r && r->n_src > 0 && r->n_src <= B200GAT_READOUT_MAX_SRC, "1 <= n_src <= B200GAT_READOUT_MAX_SRC");
  B200_CHECK_ARG(r->n_graphs > 0 && r->max_n_nodes > 0 && r->row_start, "n_graphs, max_n_nodes, row_start");
  S.n = r->n_src;
  S.off[0] = 0;
  for (int i = 0; i < r->n_src; ++i) {
    B200_CHECK_ARG(r->src[i] || g_src, "null source");
    B200_CHECK_ARG(r->width[i] > 0 && r->ld[i] >= r->width[i], "source width / leading dimension");
    S.p[i] = r->src[i];
    S.g[i] = g_src ? g_src[i] : nullptr;
    S.ld[i] = r->ld[i];
    S.width[i] = r->width[i];
    S.off[i + 1] = S.off[i] + r->width[i];
  }
  B200_CHECK_ARG(r->pooling >= B200GAT_READOUT_NONE && r->pooling <= B200GAT_READOUT_ROWS, "pooling mode");
  return B200GAT_OK;
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int b200gat_readout_fwd(const b200gat_readout* r, float* node_features, float* graph_embedding,
                                   int32_t* argmax, void* stream) {
  ReadoutSrc S;
  int rc = fill_src(S, r, nullptr);
  if (rc) return rc;
  B200_CHECK_ARG(node_features, "null node_features");
  B200_CHECK_ARG(r->pooling == B200GAT_READOUT_NONE || graph_embedding, "null graph_embedding");
  B200_CHECK_ARG(r->pooling != B200GAT_READOUT_MAX || argmax, "max pooling needs argmax");
  cudaStream_t st = (cudaStream_t)stream;
  const int F = S.off[S.n];
  const int ld = r->out_ld ? (int)r->out_ld : F;
  B200_CHECK_ARG(ld >= F && ld - F < 128, "out_ld must be F .. F + 127");
  const int64_t rows = (int64_t)r->n_graphs * r->max_n_nodes + (r->pooling == B200GAT_READOUT_ROWS ? r->n_graphs : 0);
  B200_CHECK_ARG(rows < (1ll << 31), "too many rows");
  k_readout_rows<<<(unsigned)rows, 128, 0, st>>>(S, r->row_start, r->n_graphs, r->max_n_nodes, r->pool_row0, ld,
                                                 node_features, graph_embedding);
  B200_CHECK_LAUNCH("k_readout_rows");
  if (r->pooling == B200GAT_READOUT_MAX || r->pooling == B200GAT_READOUT_AVG) {
    B200_CHECK_ARG(r->n_graphs <= 65535, "n_graphs <= 65535 for max / avg pooling");
    dim3 grid((unsigned)ceil_div(ld, 32), (unsigned)r->n_graphs);
    if (r->max_n_nodes >= 256)
      k_readout_pool<16><<<grid, dim3(32, 16), 0, st>>>(node_features, r->row_start, r->max_n_nodes, F, ld,
                                                        r->pooling, graph_embedding, argmax);
    else
      k_readout_pool<4><<<grid, dim3(32, 4), 0, st>>>(node_features, r->row_start, r->max_n_nodes, F, ld, r->pooling,
                                                       graph_embedding, argmax);
    B200_CHECK_LAUNCH("k_readout_pool");
  }
  return B200GAT_OK;
}

extern "C" int b200gat_readout_bwd(const b200gat_readout* r, const float* g_node_features, int64_t ld_g_nf,
                                   const float* g_graph_embedding, int64_t ld_g_ge, const int32_t* argmax,
                                   float* const* g_src, int64_t n_rows_total, void* stream) {
  ReadoutSrc S;
  B200_CHECK_ARG(g_src, "null g_src");
  int rc = fill_src(S, r, g_src);
  if (rc) return rc;
  B200_CHECK_ARG(n_rows_total >= r->n_task_rows && n_rows_total < (1ll << 31), "n_rows_total");
  B200_CHECK_ARG(r->pooling != B200GAT_READOUT_MAX || !g_graph_embedding || argmax, "max pooling needs argmax");
  B200_CHECK_ARG(r->pooling != B200GAT_READOUT_ROWS || r->pool_row0 + r->n_graphs <= n_rows_total, "pool rows");
  if (n_rows_total == 0) return B200GAT_OK;
  const int F = S.off[S.n];
  if (!ld_g_nf) ld_g_nf = F;
  if (!ld_g_ge) ld_g_ge = F;
  B200_CHECK_ARG(ld_g_nf >= F && ld_g_ge >= F, "gradient row strides must be >= F");
  k_readout_bwd<<<(unsigned)n_rows_total, 128, 0, (cudaStream_t)stream>>>(
      S, r->row_start, r->n_graphs, r->max_n_nodes, r->n_task_rows, r->pool_row0, r->pooling, g_node_features,
      ld_g_nf, g_graph_embedding, ld_g_ge, argmax);
  B200_CHECK_LAUNCH("k_readout_bwd");
  return B200GAT_OK;
}
