// PPO update tail on the device: GAE and the clipped-PPO minibatch loss with its gradient.
//
// The reference computes GAE with a Python loop over the rollout steps on small CPU tensors
// (alg/ppo.py:276-299) and the loss with ~40 torch ops and eight .item() host syncs per minibatch
// (alg/ppo.py:582-710).  Here:
//   k_gae      : one thread per (env, reward component), scanning the T steps backwards in registers, with the
//                reference's operation order and no fused multiply-adds (bit-exact against the fp32 torch loop)
//   k_ppo_loss : ONE launch per minibatch: advantage normalisation (local, or from all-reduced moments when the
//                minibatch is sharded over ranks), ratio / clipping / value loss / entropy, the six
//                logged statistics, the scalar loss AND d loss / d (logprob, entropy, value) - the loss is a scalar,
//                so its gradient is produced in the same pass and the autograd backward is a scale
#include <math.h>
#include "common.cuh"

namespace b200gat {

__global__ void __launch_bounds__(128) k_gae(const float* __restrict__ rewards, const float* __restrict__ values,
                                             const float* __restrict__ dones, const float* __restrict__ next_value,
                                             const float* __restrict__ next_done, int T, int E, int R, float gamma,
                                             float gamma_lambda, float sigma, float* __restrict__ advantages,
                                             float* __restrict__ returns) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= E * R) return;
  const int e = i / R;
  float last = 0.f;
  for (int t = T - 1; t >= 0; --t) {
    const int64_t o = (int64_t)t * E * R + i;
    float nnt, nv;
    if (t == T - 1) { nnt = __fsub_rn(1.f, next_done[e]); nv = next_value[i]; }
    else { nnt = __fsub_rn(1.f, dones[(int64_t)(t + 1) * E + e]); nv = values[o + (int64_t)E * R]; }
    const float v = values[o];
    // delta = (rewards[t] + gamma * nextvalues * nextnonterminal - values[t]) / sigma        (:287-291)
    const float delta = __fdiv_rn(__fsub_rn(__fadd_rn(rewards[o], __fmul_rn(__fmul_rn(gamma, nv), nnt)), v), sigma);
    // advantages[t] = lastgaelam = delta + gamma * gae_lambda * nextnonterminal * lastgaelam  (:292-298)
    last = __fadd_rn(delta, __fmul_rn(__fmul_rn(gamma_lambda, nnt), last));
    advantages[o] = last;
    returns[o] = __fadd_rn(last, v);
  }
}

// {count, sum, sum of squares} of the advantages in double: what ranks all-reduce for the global normalisation
__global__ void __launch_bounds__(1024) k_adv_moments(const float* __restrict__ adv, int n, double* __restrict__ out) {
  __shared__ double s0[32], s1[32];
  double a = 0.0, b = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) { const double x = adv[i]; a += x; b += x * x; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
  if ((threadIdx.x & 31) == 0) { s0[threadIdx.x >> 5] = a; s1[threadIdx.x >> 5] = b; }
  __syncthreads();
  if (threadIdx.x == 0) {
    a = 0.0; b = 0.0;
    for (int w = 0; w < 32; ++w) { a += s0[w]; b += s1[w]; }
    out[0] = (double)n; out[1] = a; out[2] = b;
  }
}

struct PpoArgs {
  int n;
  const float *lp, *old_lp, *adv, *ent, *val, *ret, *old_val;
  float clip, ent_coef, vf_coef;
  int norm_adv, clip_vloss, critic, has_clip;
  const double* moments;
};

template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double (*sm)[32]) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
  }
  __syncthreads();                               // previous use of sm is over
  if ((threadIdx.x & 31) == 0)
#pragma unroll
    for (int k = 0; k < NV; ++k) sm[k][threadIdx.x >> 5] = v[k];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    double t = 0.0;
    for (int w = 0; w < 32; ++w) t += sm[k][w];  // same order in every thread: identical totals, fixed order
    v[k] = t;
  }
}

__device__ __forceinline__ float vloss_elem(int critic, float d) {
  if (critic == B200GAT_CRITIC_L2) return d * d;
  if (critic == B200GAT_CRITIC_L1) return fabsf(d);
  return fabsf(d) < 1.f ? 0.5f * d * d : fabsf(d) - 0.5f;            // smooth l1, beta = 1
}
__device__ __forceinline__ float vloss_grad(int critic, float d) {
  if (critic == B200GAT_CRITIC_L2) return 2.f * d;
  const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
  if (critic == B200GAT_CRITIC_L1) return sg;
  return fabsf(d) < 1.f ? d : sg;
}

__global__ void __launch_bounds__(1024) k_ppo_loss(PpoArgs A, float* __restrict__ loss, float* __restrict__ stats,
                                                   float* __restrict__ g_lp, float* __restrict__ g_ent,
                                                   float* __restrict__ g_val) {
  __shared__ double sm[6][32];
  const int n = A.n;
  // ---- advantage statistics (alg/ppo.py:602-605: (a - mean) / (std + 1e-8), unbiased std)
  double cnt = (double)n, mean = 0.0, stdv = 1.0;
  bool norm = false;
  if (A.moments) {
    cnt = A.moments[0];
    if (A.norm_adv && cnt > 1.0) {
      norm = true;
      mean = A.moments[1] / cnt;
      const double var = (A.moments[2] - A.moments[1] * mean) / (cnt - 1.0);
      stdv = sqrt(var > 0.0 ? var : 0.0);
    }
  } else if (A.norm_adv && n > 1) {
    norm = true;
    double s[1] = {0.0};
    for (int i = threadIdx.x; i < n; i += 1024) s[0] += (double)A.adv[i];
    block_sum<1>(s, sm);
    mean = s[0] / cnt;
    double q[1] = {0.0};
    for (int i = threadIdx.x; i < n; i += 1024) { const double d = (double)A.adv[i] - mean; q[0] += d * d; }
    block_sum<1>(q, sm);
    stdv = sqrt(q[0] / (cnt - 1.0));
  }
  const float meanf = (float)mean, inv_std = norm ? 1.f / ((float)stdv + 1e-8f) : 1.f;
  const float inv_cnt = (float)(1.0 / cnt);

  // ---- unclipped value loss mean first: with clip_vloss the reference takes max(MEAN unclipped, clipped_i)
  // (torch.max of the scalar mse_loss with the per-sample clipped loss, :665-690)
  double vu_mean = 0.0;
  if (A.clip_vloss) {
    double s[1] = {0.0};
    for (int i = threadIdx.x; i < n; i += 1024) s[0] += (double)vloss_elem(A.critic, A.val[i] - A.ret[i]);
    block_sum<1>(s, sm);
    vu_mean = s[0] / cnt;
  }
  const float vu = (float)vu_mean;

  double acc[6] = {0.0, 0.0, 0.0, 0.0, 0.0, 0.0};      // pg, v, entropy, kl, clipfrac, |ratio - 1|
  double w_mean[1] = {0.0};                             // clip_vloss: weight that flows to the unclipped mean
  for (int i = threadIdx.x; i < n; i += 1024) {
    const float logratio = A.lp[i] - A.old_lp[i];
    const float ratio = expf(logratio);
    const float a = norm ? (A.adv[i] - meanf) * inv_std : A.adv[i];
    const float pg1 = -a * ratio;
    float pg = pg1, dpg = -a;                            // d pg / d ratio
    if (A.has_clip) {
      const float lo = 1.f - A.clip, hi = 1.f + A.clip;
      const float rc = fminf(fmaxf(ratio, lo), hi);
      const float pg2 = -a * rc;
      const float d2 = (ratio >= lo && ratio <= hi) ? -a : 0.f;
      if (pg2 > pg1) { pg = pg2; dpg = d2; }
      else if (pg2 == pg1) dpg = 0.5f * (-a) + 0.5f * d2;          // torch.max splits the gradient on ties
      acc[4] += fabsf(ratio - 1.f) > A.clip ? 1.0 : 0.0;
    }
    g_lp[i] = dpg * ratio * inv_cnt;
    acc[0] += pg;
    acc[3] += (ratio - 1.f) - logratio;
    acc[5] += fabsf(ratio - 1.f);
    const float ent = A.ent ? A.ent[i] : 0.f;
    acc[2] += ent;
    if (g_ent) g_ent[i] = -A.ent_coef * inv_cnt;
    const float d = A.val[i] - A.ret[i];
    if (!A.clip_vloss) {
      acc[1] += vloss_elem(A.critic, d);
      g_val[i] = A.vf_coef * vloss_grad(A.critic, d) * inv_cnt;
    } else {
      const float ov = A.old_val[i];
      const float dv = A.val[i] - ov;
      const float vc = ov + fminf(fmaxf(dv, -A.clip), A.clip);
      const float dc = vc - A.ret[i];
      const float lc = vloss_elem(A.critic, dc);
      const float gin = (dv >= -A.clip && dv <= A.clip) ? vloss_grad(A.critic, dc) : 0.f;
      float wm = 0.f, gc = 0.f;
      if (vu > lc) wm = 1.f; else if (vu == lc) { wm = 0.5f; gc = 0.5f * gin; } else gc = gin;
      acc[1] += fmaxf(vu, lc);
      w_mean[0] += wm;
      g_val[i] = A.vf_coef * gc * inv_cnt;             // + the share through the unclipped mean, added below
    }
  }
  block_sum<6>(acc, sm);
  if (A.clip_vloss) {
    block_sum<1>(w_mean, sm);
    const float share = (float)(w_mean[0] / cnt) * inv_cnt * A.vf_coef;
    for (int i = threadIdx.x; i < n; i += 1024) g_val[i] += share * vloss_grad(A.critic, A.val[i] - A.ret[i]);
  }
  if (threadIdx.x == 0) {
    const float pg = (float)(acc[0] / cnt), v = (float)(acc[1] / cnt), en = (float)(acc[2] / cnt);
    loss[0] = pg - A.ent_coef * en + A.vf_coef * v;
    stats[0] = pg; stats[1] = v; stats[2] = en;
    stats[3] = (float)(acc[3] / cnt); stats[4] = (float)(acc[4] / cnt); stats[5] = (float)(acc[5] / cnt);
    stats[6] = meanf; stats[7] = norm ? (float)stdv : 0.f;
  }
}

}  // namespace b200gat

using namespace b200gat;

extern "C" int b200gat_gae(const float* rewards, const float* values, const float* dones, const float* next_value,
                           const float* next_done, int32_t n_steps, int32_t n_envs, int32_t reward_dim, float gamma,
                           float gae_lambda, float sigma, float* advantages, float* returns, void* stream) {
  B200_CHECK_ARG(n_steps >= 0 && n_envs >= 0 && reward_dim > 0, "sizes");
  if (n_steps == 0 || n_envs == 0) return B200GAT_OK;
  B200_CHECK_ARG(rewards && values && dones && next_value && next_done && advantages && returns, "null pointer");
  // gamma * gae_lambda is ONE Python float in the reference (double product, then cast by the tensor multiply)
  const float gl = (float)((double)gamma * (double)gae_lambda);
  const int n = n_envs * reward_dim;
  k_gae<<<(unsigned)ceil_div(n, 128), 128, 0, (cudaStream_t)stream>>>(rewards, values, dones, next_value, next_done,
                                                                      n_steps, n_envs, reward_dim, gamma, gl, sigma,
                                                                      advantages, returns);
  B200_CHECK_LAUNCH("k_gae");
  return B200GAT_OK;
}

extern "C" int b200gat_ppo_adv_moments(const float* advantages, int32_t n, double* moments, void* stream) {
  B200_CHECK_ARG(n >= 0 && moments && (advantages || n == 0), "advantages / moments");
  k_adv_moments<<<1, 1024, 0, (cudaStream_t)stream>>>(advantages, n, moments);
  B200_CHECK_LAUNCH("k_adv_moments");
  return B200GAT_OK;
}

extern "C" int b200gat_ppo_loss(const b200gat_ppo_loss_args* p, float* loss, float* stats, float* g_logprob,
                                float* g_entropy, float* g_value, void* stream) {
  B200_CHECK_ARG(p && p->n > 0, "empty minibatch");
  B200_CHECK_ARG(p->logprob && p->old_logprob && p->advantages && p->value && p->returns, "null input");
  B200_CHECK_ARG(loss && stats && g_logprob && g_value, "null output");
  B200_CHECK_ARG(p->critic_loss >= B200GAT_CRITIC_L2 && p->critic_loss <= B200GAT_CRITIC_SL1, "critic_loss");
  B200_CHECK_ARG(!p->clip_vloss || (p->old_values && p->clip_coef >= 0.f), "clip_vloss needs old_values and clip_coef");
  B200_CHECK_ARG(!p->clip_vloss || !p->moments, "clip_vloss is not defined for a sharded minibatch");
  B200_CHECK_ARG(p->clip_vloss || p->critic_loss != B200GAT_CRITIC_SL1, "sl1 exists for the clipped value loss only");
  PpoArgs A;
  A.n = p->n;
  A.lp = p->logprob; A.old_lp = p->old_logprob; A.adv = p->advantages; A.ent = p->entropy; A.val = p->value;
  A.ret = p->returns; A.old_val = p->old_values;
  A.has_clip = p->clip_coef >= 0.f; A.clip = p->clip_coef; A.ent_coef = p->ent_coef; A.vf_coef = p->vf_coef;
  A.norm_adv = p->norm_adv; A.clip_vloss = p->clip_vloss; A.critic = p->critic_loss; A.moments = p->moments;
  k_ppo_loss<<<1, 1024, 0, (cudaStream_t)stream>>>(A, loss, stats, g_logprob, g_entropy, g_value);
  B200_CHECK_LAUNCH("k_ppo_loss");
  return B200GAT_OK;
}
