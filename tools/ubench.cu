// Pipe-throughput micro-benchmarks for the dense-clique kernel design.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/ubench tools/ubench.cu && gpurun_out/ubench
// Each kernel runs ITERS iterations of an unrolled body with independent chains in every thread and
// reports warp-instructions / clk / SM (from clock64 on one warp per CTA, averaged) for the
// "useful" instruction count of the body.  Development tool, not part of the library.
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdio.h>
#include <stdint.h>

constexpr int ITERS = 2048;
constexpr int NCH = 16;  // independent chains per thread

#define F2U(x) (*reinterpret_cast<unsigned long long*>(&(x)))

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, const float* in, long long* cyc) {
  unsigned smid;
  asm("mov.u32 %0, %%smid;" : "=r"(smid));
  float a[NCH], b[NCH], c[NCH];
#pragma unroll
  for (int i = 0; i < NCH; ++i) {
    a[i] = in[threadIdx.x + i * 256];
    b[i] = in[threadIdx.x + i * 256 + 1];
    c[i] = in[threadIdx.x + i * 256 + 2];
  }
  const float w = in[5], slope = in[6];
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
    if constexpr (MODE == 0) {  // FFMA
#pragma unroll
      for (int i = 0; i < NCH; ++i) c[i] = fmaf(a[i], b[i], c[i]);
    } else if constexpr (MODE == 1) {  // FADD
#pragma unroll
      for (int i = 0; i < NCH; ++i) c[i] = c[i] + a[i];
    } else if constexpr (MODE == 2) {  // FMNMX
#pragma unroll
      for (int i = 0; i < NCH; ++i) c[i] = fmaxf(c[i], a[i]), a[i] = fminf(a[i], b[i]);
    } else if constexpr (MODE == 3) {  // logit pattern, scalar: z=a+b ; r=max(z,0) ; c+=w*r   (3 instr)
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        float z = a[i] + b[(i + 5) & (NCH - 1)];
        float r = fmaxf(z, 0.f);
        c[i] = fmaf(w, r, c[i]);
        a[i] = r;
      }
    } else if constexpr (MODE == 4) {  // logit pattern, 4-instr leaky: z=a+b; t=z*slope; r=max(z,t); c+=w*r
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        float z = a[i] + b[(i + 5) & (NCH - 1)];
        float r = fmaxf(z, z * slope);
        c[i] = fmaf(w, r, c[i]);
        a[i] = r;
      }
    } else if constexpr (MODE == 5) {  // packed: FADD2, 2x FMNMX, FFMA2   (4 instr / 2 elements)
#pragma unroll
      for (int i = 0; i < NCH; i += 2) {
        float2 A = make_float2(a[i], a[i + 1]);
        float2 B = make_float2(b[(i + 5) & (NCH - 1)], b[(i + 6) & (NCH - 1)]);
        float2 Z, R, C = make_float2(c[i], c[i + 1]), W = make_float2(w, w);
        asm("add.rn.f32x2 %0, %1, %2;" : "=l"(F2U(Z)) : "l"(F2U(A)), "l"(F2U(B)));
        R.x = fmaxf(Z.x, 0.f);
        R.y = fmaxf(Z.y, 0.f);
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(F2U(C)) : "l"(F2U(W)), "l"(F2U(R)), "l"(F2U(C)));
        c[i] = C.x;
        c[i + 1] = C.y;
        a[i] = R.x;
        a[i + 1] = R.y;
      }
    } else if constexpr (MODE == 6) {  // FFMA2 only
#pragma unroll
      for (int i = 0; i < NCH; i += 2) {
        float2 A = make_float2(a[i], a[i + 1]), B = make_float2(b[i], b[i + 1]), C = make_float2(c[i], c[i + 1]);
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(F2U(C)) : "l"(F2U(A)), "l"(F2U(B)), "l"(F2U(C)));
        c[i] = C.x;
        c[i + 1] = C.y;
      }
    } else if constexpr (MODE == 7) {  // half2 logit pattern: HADD2, HMNMX2, HFMA2 (3 instr / 2 elements)
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        __half2 A = *reinterpret_cast<__half2*>(&a[i]);
        __half2 B = *reinterpret_cast<__half2*>(&b[(i + 5) & (NCH - 1)]);
        __half2 C = *reinterpret_cast<__half2*>(&c[i]);
        __half2 W = *reinterpret_cast<const __half2*>(&w);
        __half2 Z = __hadd2(A, B);
        __half2 R = __hmax2(Z, __float2half2_rn(0.f));
        C = __hfma2(W, R, C);
        c[i] = *reinterpret_cast<float*>(&C);
        a[i] = *reinterpret_cast<float*>(&R);
      }
    } else if constexpr (MODE == 8) {  // mma.sync m16n8k8 tf32, 4 independent accumulators
#pragma unroll
      for (int i = 0; i < NCH; i += 4) {
        unsigned A0 = __float_as_uint(a[i]), A1 = __float_as_uint(a[i + 1]), A2 = __float_as_uint(a[i + 2]),
                 A3 = __float_as_uint(a[i + 3]);
        unsigned B0 = __float_as_uint(b[i]), B1 = __float_as_uint(b[i + 1]);
        asm volatile(
            "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%0,%1,%2,%3};"
            : "+f"(c[i]), "+f"(c[i + 1]), "+f"(c[i + 2]), "+f"(c[i + 3])
            : "r"(A0), "r"(A1), "r"(A2), "r"(A3), "r"(B0), "r"(B1));
      }
    } else if constexpr (MODE == 9) {  // backward pattern: z=a+b; p=z>0; c+=p?g:0 ; d+=p?g:0 ; e+=p?g*z:0
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        float z = a[i] + b[(i + 5) & (NCH - 1)];
        float g = z > 0.f ? w : 0.f;
        c[i] += g;
        a[i] = fmaf(g, z, a[i]);
      }
    } else if constexpr (MODE == 10) {  // scalar logit pattern + one mma per 32 logit elements (mix)
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        float z = a[i] + b[(i + 5) & (NCH - 1)];
        float r = fmaxf(z, 0.f);
        c[i] = fmaf(w, r, c[i]);
        a[i] = r;
      }
      if ((it & 1) == 0) {
        unsigned A0 = __float_as_uint(a[0]), A1 = __float_as_uint(a[1]), A2 = __float_as_uint(a[2]),
                 A3 = __float_as_uint(a[3]);
        unsigned B0 = __float_as_uint(b[0]), B1 = __float_as_uint(b[1]);
        asm volatile(
            "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
            "{%0,%1,%2,%3};"
            : "+f"(a[4]), "+f"(a[5]), "+f"(a[6]), "+f"(a[7])
            : "r"(A0), "r"(A1), "r"(A2), "r"(A3), "r"(B0), "r"(B1));
      }
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NCH; ++i) s += a[i] + b[i] + c[i];
  out[blockIdx.x * 256 + threadIdx.x] = s;
  if (threadIdx.x == 0) { cyc[3 * blockIdx.x] = t0; cyc[3 * blockIdx.x + 1] = t1; cyc[3 * blockIdx.x + 2] = smid; }
}

template <int MODE>
void run(const char* name, double instr_per_iter_per_thread, int ctas_per_sm, float* out, float* in,
         long long* cyc) {
  const int grid = 148 * ctas_per_sm;
  k<MODE><<<grid, 256>>>(out, in, cyc);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(out, in, cyc);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  long long* h = new long long[3 * grid];
  cudaMemcpy(h, cyc, 3 * grid * sizeof(long long), cudaMemcpyDeviceToHost);
  // per SM: span between the first start and the last end of the CTAs it ran, and their count
  long long lo[256], hi[256]; int cnt[256];
  for (int i = 0; i < 256; ++i) { lo[i] = 1LL << 62; hi[i] = 0; cnt[i] = 0; }
  for (int i = 0; i < grid; ++i) {
    int s = (int)h[3 * i + 2] & 255;
    if (h[3 * i] < lo[s]) lo[s] = h[3 * i];
    if (h[3 * i + 1] > hi[s]) hi[s] = h[3 * i + 1];
    cnt[s]++;
  }
  double avg = 0, wipc = 0; int nsm = 0, maxc = 0;
  for (int s = 0; s < 256; ++s) if (cnt[s]) {
    double span = (double)(hi[s] - lo[s]);
    avg += span; nsm++; if (cnt[s] > maxc) maxc = cnt[s];
    wipc += 8.0 * cnt[s] * ITERS * instr_per_iter_per_thread / span;
  }
  avg /= nsm; wipc /= nsm;
  printf("[sms=%d maxcta/sm=%d] ", nsm, maxc);
  printf("%-44s ctas/SM=%d  %.3f ms  cyc=%.0f  warp-instr/clk/SM=%.2f  (elem/clk/SM=%.1f)\n", name, ctas_per_sm,
         ms, avg, wipc, wipc * 32);
  delete[] h;
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("CUDA error: %s\n", cudaGetErrorString(e));
}

int main() {
  float *out, *in;
  long long* cyc;
  cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaMalloc(&in, 256 * (NCH + 1) * 4 + 64);
  cudaMalloc(&cyc, 148 * 8 * 8 * 3);
  float* h = new float[256 * (NCH + 1) + 16];
  for (int i = 0; i < 256 * (NCH + 1) + 16; ++i) h[i] = 0.001f * (i % 97) - 0.04f;
  cudaMemcpy(in, h, (256 * (NCH + 1) + 16) * 4, cudaMemcpyHostToDevice);
  for (int c : {1, 2, 4}) {
    run<0>("FFMA", NCH, c, out, in, cyc);
    run<1>("FADD", NCH, c, out, in, cyc);
    run<2>("FMNMX x2", 2 * NCH, c, out, in, cyc);
    run<3>("logit3 (FADD,FMNMX,FFMA) as elements", NCH, c, out, in, cyc);
    run<4>("logit4 (FADD,FMUL,FMNMX,FFMA) as elements", NCH, c, out, in, cyc);
    run<5>("logit packed (FADD2,2xFMNMX,FFMA2) as elements", NCH, c, out, in, cyc);
    run<6>("FFMA2 (as FFMA2 instr)", NCH / 2, c, out, in, cyc);
    run<7>("half2 logit (HADD2,HMNMX2,HFMA2) as half2 elems", NCH, c, out, in, cyc);
    run<8>("mma.m16n8k8.tf32 (as mma instr)", NCH / 4, c, out, in, cyc);
    run<9>("bwd pattern (FADD,FSETP/FSEL,FADD,FFMA) elems", NCH, c, out, in, cyc);
    run<10>("logit3 + 1 mma / 2 iters, as elements", NCH, c, out, in, cyc);
  }
  return 0;
}
