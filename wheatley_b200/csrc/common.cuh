// Shared helpers for the b200gat kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/b200gat.h"

namespace b200gat {

void set_error(const char* fmt, ...);
void note_launch(int n);  // counts kernel launches (b200gat_launch_count)

// Optional per-kernel-class device timing (CUDA events on the launch stream), switched on by
// bench.py through b200gat_timing_enable(); a no-op otherwise.
struct ScopedTimer {
  ScopedTimer(int cls, cudaStream_t st);
  ~ScopedTimer();
  int cls_;
  cudaStream_t st_;
  cudaEvent_t e0_, e1_;
  bool on_, capturing_;
};

bool timing_enabled();     // b200gat_timing_enable(1) is in effect

// The weight / bias gradients of b200gat_mlp_bwd alone (mlp_tf32.cu), for a caller that has already run the input
// -gradient chain with gr->dW[] = NULL (so gr->dz[] hold every hidden layer's d loss / d z): used by trunk.cu to
// issue them on a side stream.
int mlp_bwd_dw(const b200gat_mlp* m, const b200gat_mlp_grads* gr, void* stream);

// Side stream for work that only produces parameter gradients (trunk.cu has the definition and the rationale).
struct AuxStream {
  cudaStream_t s = nullptr;
  cudaEvent_t m1 = nullptr, m2 = nullptr, a1 = nullptr, a2 = nullptr;   // recorded on main (m*) / on the side stream (a*)
  bool ok = false;
};
AuxStream* aux_stream_for(cudaStream_t main);                           // NULL: run everything on `main`
bool aux_fork(cudaStream_t main, cudaStream_t aux, cudaEvent_t ev);     // aux waits for what main has queued so far
bool aux_join(cudaStream_t main, cudaStream_t aux, cudaEvent_t ev);     // main waits for what aux has queued so far

#define B200_CHECK_ARG(cond, msg)                       \
  do {                                                  \
    if (!(cond)) {                                      \
      b200gat::set_error("invalid argument: %s", msg);  \
      return B200GAT_ERR_INVALID;                       \
    }                                                   \
  } while (0)

#define B200_CHECK_LAUNCH(name) B200_CHECK_LAUNCHES(name, 1)
#define B200_CHECK_LAUNCHES(name, n)                                                \
  do {                                                                              \
    b200gat::note_launch(n);                                                        \
    cudaError_t e__ = cudaGetLastError();                                           \
    if (e__ != cudaSuccess) {                                                       \
      b200gat::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__));   \
      return B200GAT_ERR_CUDA;                                                      \
    }                                                                               \
  } while (0)

constexpr int kNumSMs = 148;  // B200

// Programmatic dependent launch.  A kernel launched through launch_k() may become resident while the kernel before it
// on the stream is still running its LAST wave (every CTA of that kernel has executed pdl_sync()'s trigger or exited):
// its CTAs are scheduled, run pdl_sync() and block there until the predecessor has completed and its writes are
// visible.  The launch latency and the ramp-up of ~180 back-to-back dependent kernels per step come off the critical
// path; nothing is read or written before the wait, so results cannot change.  Every kernel launched through
// launch_k() MUST call pdl_sync() before its first global access.  B200GAT_PDL=0 launches them the ordinary way
// (the two instructions are no-ops then).
bool pdl_enabled();
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_sync() { pdl_trigger(); pdl_wait(); }
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                            Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__host__ __device__ inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float leaky(float z, float slope) { return z > 0.f ? z : z * slope; }

}  // namespace b200gat
