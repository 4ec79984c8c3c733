// Masked action head: one warp per row of [B, n] logits.
// forward : x = mask ? logit : fill ; logp = x - (max + log sum exp(x - max)) ;
//           entropy = -sum_mask exp(logp) * logp ; argmax = first index of max x
// backward: d/dlogit of  g_lp * logp[action] + g_ent * entropy
// (generic/agent.py:197-231, :238-254, :256-277)
#include <math.h>
#include "common.cuh"

namespace b200gat {

__global__ void __launch_bounds__(128) k_mask_head_fwd(const float* __restrict__ logits, int64_t ld,
                                                       const uint8_t* __restrict__ mask, int64_t ldm,
                                                       int B, int n, float fill, float* __restrict__ logp,
                                                       float* __restrict__ entropy,
                                                       int64_t* __restrict__ argmax) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= B) return;
  const float* x = logits + (int64_t)row * ld;
  const uint8_t* mk = mask ? mask + (int64_t)row * ldm : nullptr;
  float mx = -INFINITY;
  int am = 0x7fffffff;
  for (int j = lane; j < n; j += 32) {
    float v = (!mk || mk[j]) ? x[j] : fill;
    if (v > mx) { mx = v; am = j; }   // strict > keeps the first index inside a lane
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    float omx = __shfl_xor_sync(0xffffffffu, mx, o);
    int oam = __shfl_xor_sync(0xffffffffu, am, o);
    if (omx > mx || (omx == mx && oam < am)) { mx = omx; am = oam; }
  }
  float sum = 0.f;
  for (int j = lane; j < n; j += 32) {
    float v = (!mk || mk[j]) ? x[j] : fill;
    sum += expf(v - mx);
  }
  sum = warp_sum(sum);
  const float lse = mx + logf(sum);
  float ent = 0.f;
  for (int j = lane; j < n; j += 32) {
    bool on = (!mk || mk[j]);
    float lp = (on ? x[j] : fill) - lse;
    logp[(int64_t)row * n + j] = lp;
    if (on) ent -= expf(lp) * fmaxf(lp, -3.4028234663852886e38f);
  }
  ent = warp_sum(ent);
  if (lane == 0) {
    entropy[row] = ent;
    if (argmax) argmax[row] = n > 0 ? am : 0;
  }
}

__global__ void __launch_bounds__(128) k_mask_head_bwd(const float* __restrict__ logp,
                                                       const uint8_t* __restrict__ mask, int64_t ldm,
                                                       const int64_t* __restrict__ action,
                                                       const float* __restrict__ entropy,
                                                       const float* __restrict__ g_lp,
                                                       const float* __restrict__ g_ent, int B, int n,
                                                       float* __restrict__ g_logits) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= B) return;
  const uint8_t* mk = mask ? mask + (int64_t)row * ldm : nullptr;
  const float gl = g_lp ? g_lp[row] : 0.f, ge = g_ent ? g_ent[row] : 0.f;
  const float Hh = entropy[row];
  const int a = action ? (int)action[row] : -1;
  for (int j = lane; j < n; j += 32) {
    float out = 0.f;
    if (!mk || mk[j]) {
      float lp = logp[(int64_t)row * n + j];
      float p = expf(lp);
      out = ((j == a ? 1.f : 0.f) - p) * gl - p * (lp + Hh) * ge;
    }
    g_logits[(int64_t)row * n + j] = out;
  }
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int b200gat_mask_head_fwd(const float* logits, int64_t ld_logits, const uint8_t* mask,
                                     int64_t ld_mask, int32_t batch, int32_t n, float fill, float* logp,
                                     float* entropy, int64_t* argmax, void* stream) {
  B200_CHECK_ARG(batch >= 0 && n >= 0, "negative size");
  if (batch == 0) return B200GAT_OK;
  B200_CHECK_ARG(logits && logp && entropy, "null pointer");
  B200_CHECK_ARG(ld_logits >= n && (!mask || ld_mask >= n), "row stride smaller than n");
  ScopedTimer tm(B200GAT_TIMING_HEAD, (cudaStream_t)stream);
  k_mask_head_fwd<<<(unsigned)ceil_div((int64_t)batch * 32, 128), 128, 0, (cudaStream_t)stream>>>(
      logits, ld_logits, mask, ld_mask, batch, n, fill, logp, entropy, argmax);
  B200_CHECK_LAUNCH("k_mask_head_fwd");
  return B200GAT_OK;
}

extern "C" int b200gat_mask_head_bwd(const float* logp, const uint8_t* mask, int64_t ld_mask,
                                     const int64_t* action, const float* entropy, const float* g_logprob,
                                     const float* g_entropy, int32_t batch, int32_t n, float* g_logits,
                                     void* stream) {
  B200_CHECK_ARG(batch >= 0 && n >= 0, "negative size");
  if (batch == 0 || n == 0) return B200GAT_OK;
  B200_CHECK_ARG(logp && entropy && g_logits, "null pointer");
  ScopedTimer tm(B200GAT_TIMING_HEAD, (cudaStream_t)stream);
  k_mask_head_bwd<<<(unsigned)ceil_div((int64_t)batch * 32, 128), 128, 0, (cudaStream_t)stream>>>(
      logp, mask, ld_mask, action, entropy, g_logprob, g_entropy, batch, n, g_logits);
  B200_CHECK_LAUNCH("k_mask_head_bwd");
  return B200GAT_OK;
}
