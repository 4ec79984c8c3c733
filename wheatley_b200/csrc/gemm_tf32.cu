// Dense projections of the path (north_star item 4) as a hand-written tcgen05 GEMM:
//   D[m, n] = sum_k A(m, k) * B(n, k)      fp32 operands read as TF32 by the tensor core (kind::tf32),
//                                           fp32 accumulators in TMEM, fused epilogue in registers.
// One CTA = one 128 x BN output tile (x one K split), BN in {64, 128, 256}.  Warp roles (320 threads):
//   warps 0-7  epilogue (warp w: TMEM lanes 32 (w & 3).., column half w >> 2): tcgen05.ld of the accumulator rows (thread = row), transposed through shared memory so
//              that bias / GELU / GELU' / residual run on coalesced 128-byte row segments
//   warp  8    TMA producer: cp.async.bulk.tensor boxes (128-byte swizzle) into a 3-stage shared-memory ring
//   warp  9    MMA issuer: one thread issues tcgen05.mma (M=128, N=BN, K=8) x 4 per stage, tcgen05.commit
//              releases the stage / signals the epilogue
// Both operand majors are handled by descriptors, not by transposes:
//   K-major  operand: stored [mn rows, k cols]  (forward: X and W; dX: dY)
//   MN-major operand: stored [k rows, mn cols]  (dX: W read as W^T; dW: dY and X, reduction over the rows)
// so y = x W^T, dx = dy W and dW = dy^T x all run on this kernel with no copies.  "Parts" let an operand or
// the output live in several equally shaped matrices (the three EGAT projections Pi | Pj | V).
#include <cuda.h>
#include <math.h>
#include <stdlib.h>
#include <atomic>
#include "common.cuh"

namespace b200gat {
namespace {

std::atomic<int> g_gemm_x3{0};   // b200gat_gemm_precision(): 0 = TF32 operands, 1 = 3xTF32 split (fp32-level accuracy)

constexpr int BM = 128;      // rows of the output tile = TMEM lanes
constexpr int BK = 32;       // fp32 per stage row = 128 bytes = one swizzle span
constexpr int STAGES = 3;
constexpr int A_STAGE_BYTES = BM * BK * 4;

struct GemmArgs {
  int M, N;                   // output extent
  int kb_total, kb_per_split; // k-blocks (32 reduction indices each)
  int a_kb_per_part;          // K-major A split over parts along k
  int a_mt_per_part;          // MN-major A split over parts along m
  int n_tiles_per_part;       // output split over parts along n
  float* D;
  long long ldd, d_part_stride, d_split_stride;
  const float* bias;          // [N] added to every row, or null
  const float* R;             // [M, N] residual added after the activation-free epilogue, or null
  long long ldr;
  int epi;
  float* D2;                  // EPI_GELU: gelu(D) (D itself = pre-activation, may be null)
  const float* Zin;           // EPI_DGELU: D = acc * gelu'(Zin)
  long long ld2;
  unsigned mn_lbo, mn_sbo, mn_layout, mn_kadv;   // MN-major descriptor fields (bytes; see smem_desc)
  float* a_rowsum;            // [M] (+ split * rowsum_split_stride): sum_k A(m, k), or null
  long long rowsum_split_stride;
};

// ------------------------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
// Bounded wait: a protocol error traps (the launch fails) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0, spins = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (++spins > 20000000u) __trap();
  }
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* tm, uint32_t bar, int c0, int c1,
                                            int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// Shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address and offsets in 16-byte units,
// version 1 (Blackwell).
//   K-major : SWIZZLE_128B (layout type 2): rows of 128 B (32 fp32 along k), 8-row groups SBO = 1024 B apart;
//             LBO unused (1)
//   MN-major: 32-bit operands only exist as SWIZZLE_128B_BASE32B (layout type 1; the TMA map uses the matching
//             128B_ATOM_32B swizzle: 32-byte chunks permuted by row mod 4): atoms of 32 mn x 4 k (4 rows of
//             128 B); next 32 mn LBO bytes further, next 4 k SBO = 512 bytes further
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes,
                                              uint32_t layout_type) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout_type << 61;
  return d;
}

// Exact-form GELU (torch.nn.GELU default, erf) and its derivative with one exp and one reciprocal per element:
//   Phi(z) = 0.5 (1 + erf(z / sqrt 2)),  erf(x) = sign(x) (1 - (a1 t + ... + a5 t^5) exp(-x^2)),  t = 1 / (1 + p |x|)
// (Abramowitz-Stegun 7.1.26, |error| <= 1.5e-7, i.e. fp32 rounding level); exp(-x^2) = exp(-z^2 / 2) is also the
// Gaussian of the derivative: gelu'(z) = Phi(z) + z exp(-z^2 / 2) / sqrt(2 pi).
__device__ __forceinline__ void phi_and_gauss(float z, float& phi, float& gauss) {
  const float ax = fabsf(z) * 0.70710678118654752f;
  const float t = __fdividef(1.f, fmaf(0.3275911f, ax, 1.f));
  gauss = __expf(-0.5f * z * z);
  float poly = fmaf(1.061405429f, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float half_erfc = 0.5f * poly * t * gauss;          // 0.5 (1 - erf|x|)
  phi = z >= 0.f ? 1.f - half_erfc : half_erfc;
}
__device__ __forceinline__ float gelu_f(float z) {
  float phi, g;
  phi_and_gauss(z, phi, g);
  return z * phi;
}
__device__ __forceinline__ float dgelu_f(float z) {
  float phi, g;
  phi_and_gauss(z, phi, g);
  return fmaf(z * 0.39894228040143268f, g, phi);
}

// ------------------------------------------------------------------------------------------- the kernel
// X3 = 3xTF32 split arithmetic (fp32-level accuracy, for torch.backends.cuda.matmul.allow_tf32 == False): the tensor
// core reads only the upper 19 bits of an fp32 operand, so hi = x as it lies in shared memory and lo = x - trunc(x)
// (exact in fp32); a_hi*b_hi + a_hi*b_lo + a_lo*b_hi reproduces the fp32 product to ~2^-20 relative.  The lo tiles
// are written by the (otherwise idle) epilogue warps into a second copy of the stage - same swizzled layout, the
// split is element-wise - and the MMA warp issues three tcgen05.mma per k step.  Two stages instead of three.
template <int BN, bool A_MN, bool B_MN, bool X3>
__global__ void __launch_bounds__(320, ((BN == 256 || X3) ? 1 : 2)) k_gemm_tf32(const __grid_constant__ CUtensorMap tmA,
                                                   const __grid_constant__ CUtensorMap tmB, GemmArgs p) {
  constexpr int B_STAGE_BYTES = BN * BK * 4;
  constexpr int TILE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;          // what the TMA fills per stage
  constexpr int STAGE_BYTES = X3 ? 2 * TILE_BYTES : TILE_BYTES;      // X3: A | B | A_lo | B_lo
  constexpr int NST = X3 ? 2 : STAGES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // swizzle-128B tiles need 1024-byte alignment
  __shared__ __align__(8) uint64_t bars[3 * STAGES + 1];
  __shared__ uint32_t tmem_base_s;
  const uint32_t full0 = smem_u32(&bars[0]), empty0 = smem_u32(&bars[STAGES]), accum_bar = smem_u32(&bars[2 * STAGES]);
  const uint32_t lo0 = smem_u32(&bars[2 * STAGES + 1]);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int mt = blockIdx.x, nt = blockIdx.y;
  const int kb0 = blockIdx.z * p.kb_per_split;
  const int nkb = min(p.kb_per_split, p.kb_total - kb0);

  // Row sums of A (the bias gradient inside the weight-gradient GEMM, A = dy stored MN-major): epilogue warps
  // 0-3 are idle during the main loop, so they add up the staged A tiles in exact fp32 (thread = one of the 128
  // mn columns, k ascending: deterministic) before the stage is released - the stage's "empty" barrier then counts
  // the MMA commit plus these 128 readers.
  const bool rowsum = A_MN && p.a_rowsum != nullptr && nt == 0;
  pdl_trigger();                // the next kernel may be scheduled once every CTA of this one is running
  if (threadIdx.x == 0) {
    for (int s = 0; s < NST; ++s) {
      mbar_init(full0 + 8 * s, 1);
      mbar_init(empty0 + 8 * s, rowsum ? 129 : 1);
      mbar_init(lo0 + 8 * s, 256);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const uint32_t tmem_cols = (uint32_t)BN;          // power of two >= 32
  if (warp == 0) {      // TMEM: BN fp32 accumulator columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)),
                 "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;
  pdl_wait();                   // barriers and TMEM are set up; no global access before the predecessor has completed

  if (warp == 8) {
    // ------------------------------------------------------------------ TMA producer
    for (int i = 0; i < nkb; ++i) {
      const int s = i % NST;
      const uint32_t ph = (uint32_t)(i / NST) & 1u;
      if (lane == 0) {
        mbar_wait(empty0 + 8 * s, ph ^ 1u);
        const uint32_t fb = full0 + 8 * s;
        mbar_expect_tx(fb, TILE_BYTES);
        const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_STAGE_BYTES;
        const int kb = kb0 + i;
        if (!A_MN) {
          const int part = kb / p.a_kb_per_part, col = (kb - part * p.a_kb_per_part) * BK;
          tma_load_3d(sa, &tmA, fb, col, mt * BM, part);                    // box 32 k x 128 rows
        } else {
          const int part = mt / p.a_mt_per_part, col = (mt - part * p.a_mt_per_part) * BM;
#pragma unroll
          for (int j = 0; j < BM / 32; ++j) tma_load_3d(sa + j * 4096, &tmA, fb, col + 32 * j, kb * BK, part);
        }
        if (!B_MN) {
          tma_load_3d(sb, &tmB, fb, kb * BK, nt * BN, 0);                   // box 32 k x BN rows
        } else {
#pragma unroll
          for (int j = 0; j < BN / 32; ++j) tma_load_3d(sb + j * 4096, &tmB, fb, nt * BN + 32 * j, kb * BK, 0);
        }
      }
      __syncwarp();
    }
  } else if (warp == 9) {
    // ------------------------------------------------------------------ MMA issuer
    // instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = TF32, majors, N >> 3, M >> 4
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                           ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
    for (int i = 0; i < nkb; ++i) {
      const int s = i % NST;
      const uint32_t ph = (uint32_t)(i / NST) & 1u;
      if (lane == 0) {
        mbar_wait((X3 ? lo0 : full0) + 8 * s, ph);       // X3: the lo tiles are written (implies the stage is full)
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t sa = base + s * STAGE_BYTES, sb = sa + A_STAGE_BYTES;
#pragma unroll
        for (int k = 0; k < BK / 8; ++k) {
          const uint64_t ad = A_MN ? smem_desc(sa + k * p.mn_kadv, p.mn_lbo, p.mn_sbo, p.mn_layout)
                                   : smem_desc(sa + k * 32, 16, 1024, 2);
          const uint64_t bd = B_MN ? smem_desc(sb + k * p.mn_kadv, p.mn_lbo, p.mn_sbo, p.mn_layout)
                                   : smem_desc(sb + k * 32, 16, 1024, 2);
          if constexpr (X3) {
            const uint32_t la = sa + TILE_BYTES, lb = sb + TILE_BYTES;
            const uint64_t adl = A_MN ? smem_desc(la + k * p.mn_kadv, p.mn_lbo, p.mn_sbo, p.mn_layout)
                                      : smem_desc(la + k * 32, 16, 1024, 2);
            const uint64_t bdl = B_MN ? smem_desc(lb + k * p.mn_kadv, p.mn_lbo, p.mn_sbo, p.mn_layout)
                                      : smem_desc(lb + k * 32, 16, 1024, 2);
            umma_tf32(tmem_base, adl, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);     // small terms first
            umma_tf32(tmem_base, ad, bdl, idesc, 1u);
            umma_tf32(tmem_base, ad, bd, idesc, 1u);
          } else {
            umma_tf32(tmem_base, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
          }
        }
        umma_commit(empty0 + 8 * s);                 // frees the stage when these MMAs have read it
        if (i == nkb - 1) umma_commit(accum_bar);    // accumulator complete
      }
      __syncwarp();
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warp w owns TMEM lanes 32w..32w+31)
    // 32-column chunks: tcgen05.ld gives thread = row; the chunk is transposed through shared memory (the
    // pipeline's stage 0, idle once the accumulator is complete; rows padded to 36 floats: 16-byte accesses are
    // conflict free per quarter warp) so that every global access is a full 128-byte row segment; the
    // Zin / residual rows of a chunk are requested before the TMEM load is waited for.
    const int part = nt / p.n_tiles_per_part, col0 = (nt - part * p.n_tiles_per_part) * BN;
    const int ncol0 = nt * BN;                       // column in the un-parted output (bias / R / D2 / Zin index)
    const int lq = warp & 3, hf = warp >> 2;          // TMEM lane quarter, column half
    const int row_base = mt * BM + lq * 32;
    const int sub = lane >> 3, cq = (lane & 7) * 4;  // coalesced mapping: rows 4*it + sub, columns cq..cq+3
    float* stg = reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw))) + warp * (32 * 36);   // 8 x 4.5 KB <= the ring
    float* dbase = p.D ? p.D + (long long)blockIdx.z * p.d_split_stride + (long long)part * p.d_part_stride + col0
                       : nullptr;
    const bool dgelu = p.epi == B200GAT_EPI_DGELU, gelu = p.epi == B200GAT_EPI_GELU;
    bool waited = false;
    const bool do_rowsum = rowsum && warp < 4;
    if (X3 || do_rowsum) {
      // MN-major A stage: 4 boxes of 32 mn columns x 32 k rows (128 B per row, 32-byte chunks xor-ed with row & 3)
      const int t = threadIdx.x, c = t & 31;
      uint8_t* ring = smem_raw + (base - smem_u32(smem_raw));
      const uint8_t* a_tile = ring + (t >> 5) * 4096 + (c & 7) * 4;
      float sum = 0.f;
      for (int i = 0; i < nkb; ++i) {
        const int s2 = i % NST;
        mbar_wait(full0 + 8 * s2, (uint32_t)(i / NST) & 1u);
        if constexpr (X3) {
          // lo = x - trunc_tf32(x) for every element of the A and B tiles of this stage (element-wise: the lo tiles
          // inherit the swizzled layout), then make the generic-proxy writes visible to the tensor core's reads
          const float4* src = reinterpret_cast<const float4*>(ring + s2 * STAGE_BYTES);
          float4* dst = reinterpret_cast<float4*>(ring + s2 * STAGE_BYTES + TILE_BYTES);
#pragma unroll 4
          for (int idx = t; idx < TILE_BYTES / 16; idx += 256) {
            const float4 x = src[idx];
            float4 l;
            l.x = x.x - __uint_as_float(__float_as_uint(x.x) & 0xffffe000u);
            l.y = x.y - __uint_as_float(__float_as_uint(x.y) & 0xffffe000u);
            l.z = x.z - __uint_as_float(__float_as_uint(x.z) & 0xffffe000u);
            l.w = x.w - __uint_as_float(__float_as_uint(x.w) & 0xffffe000u);
            dst[idx] = l;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(lo0 + 8 * s2) : "memory");
        }
        if (do_rowsum) {
          const uint8_t* st_base = a_tile + s2 * STAGE_BYTES;
#pragma unroll 8
          for (int row = 0; row < BK; ++row)
            sum += *reinterpret_cast<const float*>(st_base + row * 128 + (((c >> 3) ^ (row & 3)) << 5));
          asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty0 + 8 * s2) : "memory");
        }
      }
      if (do_rowsum) {
        const int row = mt * BM + t;
        if (row < p.M) p.a_rowsum[(long long)blockIdx.z * p.rowsum_split_stride + row] = sum;
      }
    }
#pragma unroll 1
    for (int c = hf * (BN / 2); c < (hf + 1) * (BN / 2); c += 32) {
      if (ncol0 + c >= p.N) break;                   // ragged last column tile (N % 4 == 0): warp-uniform
      const bool col_ok = ncol0 + c + cq < p.N;
      float4 pre[8];                                 // Zin (DGELU) or residual rows of this chunk, never both
      const float* pre_src = dgelu ? p.Zin : p.R;
      const long long pre_ld = dgelu ? p.ld2 : p.ldr;
      if (pre_src) {
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int row = row_base + 4 * it + sub;
          pre[it] = (row < p.M && col_ok) ? __ldg(reinterpret_cast<const float4*>(pre_src + (long long)row * pre_ld + ncol0 + c + cq))
                              : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.bias && col_ok) b4 = __ldg(reinterpret_cast<const float4*>(p.bias + ncol0 + c + cq));
      if (!waited) {
        mbar_wait(accum_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        waited = true;
      }
      uint32_t r[32];
      tmem_ld32(tmem_base + ((uint32_t)(lq * 32) << 16) + (uint32_t)c, r);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<float4*>(stg + lane * 36 + 4 * q) =
            make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]), __uint_as_float(r[4 * q + 2]),
                        __uint_as_float(r[4 * q + 3]));
      __syncwarp();
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int rl = 4 * it + sub, row = row_base + rl;
        float4 v = *reinterpret_cast<const float4*>(stg + rl * 36 + cq);
        v.x += b4.x; v.y += b4.y; v.z += b4.z; v.w += b4.w;
        if (dgelu) {
          v.x *= dgelu_f(pre[it].x); v.y *= dgelu_f(pre[it].y); v.z *= dgelu_f(pre[it].z); v.w *= dgelu_f(pre[it].w);
        } else if (pre_src) {
          v.x += pre[it].x; v.y += pre[it].y; v.z += pre[it].z; v.w += pre[it].w;
        }
        if (row < p.M && col_ok) {
          if (dbase) *reinterpret_cast<float4*>(dbase + (long long)row * p.ldd + c + cq) = v;
          if (gelu)
            *reinterpret_cast<float4*>(p.D2 + (long long)row * p.ld2 + ncol0 + c + cq) =
                make_float4(gelu_f(v.x), gelu_f(v.y), gelu_f(v.z), gelu_f(v.w));
        }
      }
      __syncwarp();
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols) : "memory");
  }
}

// Deterministic split-K reduction: out[i] = sum_z ws[z * stride + i] for the n4 float4 of the output, and the same
// for the r4 float4 of the row-sum partials behind them.  Block = 8 z-lanes x 32 float4 columns: z-lane j sums the
// splits j, j+8, ... in order, the eight partials are added in lane order.
__global__ void __launch_bounds__(256) k_splitk_reduce(const float* __restrict__ ws, int splits, long long stride,
                                                       long long n4, float* __restrict__ out,
                                                       const float* __restrict__ ws_rs, long long rs_stride,
                                                       long long r4, float* __restrict__ out_rs) {
  pdl_sync();
  __shared__ float4 part[8][32];
  const int lane = threadIdx.x & 31, zl = threadIdx.x >> 5;
  const long long dblocks = (n4 + 31) / 32;
  long long i = (long long)blockIdx.x * 32 + lane;
  const float* src = ws;
  float* dst = out;
  long long st = stride, n = n4;
  if (blockIdx.x >= dblocks) {         // the blocks behind the output handle the row sums
    i -= dblocks * 32;
    src = ws_rs; dst = out_rs; st = rs_stride; n = r4;
  }
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n) {
#pragma unroll 4
    for (int z = zl; z < splits; z += 8) {
      const float4 b = reinterpret_cast<const float4*>(src + z * st)[i];
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
  }
  part[zl][lane] = a;
  __syncthreads();
  if (zl == 0 && i < n) {
#pragma unroll
    for (int j = 1; j < 8; ++j) {
      const float4 b = part[j][lane];
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    reinterpret_cast<float4*>(dst)[i] = a;
  }
}

// Column sums of X [M, N] (bias gradients): slab partials, then a fixed-order finish.
__global__ void k_colsum_partial(const float* __restrict__ X, long long M, int N, long long ld, int rows_per_slab,
                                 float* __restrict__ partial) {
  const long long r0 = (long long)blockIdx.x * rows_per_slab;
  const long long r1 = r0 + rows_per_slab < M ? r0 + rows_per_slab : M;
  for (int c = threadIdx.x; c < N; c += blockDim.x) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    long long r = r0;
    for (; r + 4 <= r1; r += 4) {
      a0 += X[r * ld + c];
      a1 += X[(r + 1) * ld + c];
      a2 += X[(r + 2) * ld + c];
      a3 += X[(r + 3) * ld + c];
    }
    for (; r < r1; ++r) a0 += X[r * ld + c];
    partial[(long long)blockIdx.x * N + c] = (a0 + a1) + (a2 + a3);
  }
}
__global__ void k_colsum_finish(const float* __restrict__ partial, int slabs, int N, float* __restrict__ out) {
  // one warp per column: lanes stride over the slabs, fixed-order shuffle tree
  const int c = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (c >= N) return;
  float a = 0.f;
  for (int s = lane; s < slabs; s += 32) a += partial[(long long)s * N + c];
  a = warp_sum(a);
  if (lane == 0) out[c] = a;
}

// ------------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// Development overrides of the MN-major descriptor fields (unset in production).
unsigned env_u(const char* name, unsigned dflt) {
  const char* e = getenv(name);
  return e && e[0] ? (unsigned)strtoul(e, nullptr, 0) : dflt;
}

// 3-D map over `parts` row-major matrices [rows, cols] (leading dimension ld, part_stride floats apart);
// box = box_cols x box_rows x 1, 128-byte swizzle (32-byte atoms for MN-major operands), out-of-bounds
// elements read as zero.
int make_map(CUtensorMap* tm, const float* ptr, long long cols, long long rows, long long ld, int parts,
             long long part_stride, int box_cols, int box_rows, bool mn_major) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from the driver");
    return B200GAT_ERR_CUDA;
  }
  cuuint64_t gdim[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)(parts > 0 ? parts : 1)};
  cuuint64_t gstr[2] = {(cuuint64_t)ld * 4, (cuuint64_t)(parts > 1 ? part_stride : rows * ld) * 4};
  cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = CUDA_SUCCESS;
  for (int attempt = 0; attempt < 2; ++attempt) {
    // autograd worker threads may not have the primary context current yet (CUDA_ERROR_INVALID_CONTEXT):
    // a runtime call binds it, then the encode is repeated
    if (attempt == 1) cudaFree(nullptr);
    r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(ptr), gdim, gstr, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                  mn_major ? (CUtensorMapSwizzle)env_u("B200GAT_MN_TMA_SWIZZLE", CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
                           : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_ERROR_INVALID_CONTEXT) break;
  }
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for a [%lld x %lld] ld %lld x %d map", (int)r, rows, cols, ld, parts);
    return B200GAT_ERR_CUDA;
  }
  return B200GAT_OK;
}

int pick_bn(long long n_per_part, bool a_mn) {
  static const unsigned force = env_u("B200GAT_GEMM_BN", 0);      // development override
  if (force == 64 && n_per_part % 64 == 0) return 64;
  if (force == 128 && n_per_part % 128 == 0) return 128;
  // 256-wide tiles read the weight operand once per 128 rows; weight gradients (split-K) keep 128 so that the
  // partials stay small.  In TF32 mode two 128-wide CTAs per SM overlap each other's load / MMA / epilogue phases
  // and win (8 x Taillard 100x20: 4.11 vs 4.20 ms per step, 30x20 x 64: 6.91 vs 7.22); in 3xTF32 mode the
  // epilogue warps also split the operands, and the wider tile (half as many splits per output) wins (4.94 vs 5.2).
  if (!a_mn && n_per_part % 256 == 0 && (force == 256 || g_gemm_x3.load(std::memory_order_relaxed))) return 256;
  if (n_per_part % 128 == 0) return 128;
  if (n_per_part % 4 == 0) return 64;       // ragged last tile: loads zero-filled by TMA, stores predicated
  return 0;
}

struct Shape {
  int bn, m_tiles, n_tiles, kb_total, splits, kb_per_split;
};

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// Validates the descriptor and derives the tiling; returns 0 when covered.
int plan_gemm(const b200gat_gemm* g, Shape* s, bool report) {
#define GEMM_REQ(cond, msg)                                     \
  do {                                                          \
    if (!(cond)) {                                              \
      if (report) set_error("b200gat_gemm_tf32: %s", msg);      \
      return B200GAT_ERR_UNSUPPORTED;                           \
    }                                                           \
  } while (0)
  GEMM_REQ(g->M > 0 && g->N > 0 && g->K > 0, "empty problem");
  GEMM_REQ(g->M < (1ll << 31) && g->N <= 65535ll * 64 && g->K < (1ll << 31), "extent too large");
  const int ap = g->a_parts > 1 ? g->a_parts : 1, dp = g->d_parts > 1 ? g->d_parts : 1;
  GEMM_REQ(g->N % dp == 0, "N not divisible by d_parts");
  s->bn = pick_bn(g->N / dp, g->a_major != 0);
  GEMM_REQ(s->bn != 0, "N must be a multiple of 4");
  GEMM_REQ(dp == 1 || (g->N / dp) % 64 == 0, "N per part must be a multiple of 64");
  GEMM_REQ(g->lda % 4 == 0 && g->ldb % 4 == 0 && g->ldd % 4 == 0, "leading dimensions must be multiples of 4 floats");
  if (g->a_rowsum)
    GEMM_REQ(g->d_parts <= 1 && g->a_major == 1 && g->a_parts <= 1,
             "row sums need an MN-major, un-parted A (A = dy of a weight-gradient GEMM) and an un-parted output");
  GEMM_REQ(aligned16(g->A) && aligned16(g->B) && aligned16(g->D) && aligned16(g->bias) && aligned16(g->residual) &&
               aligned16(g->D2) && aligned16(g->Zin),
           "pointers must be 16-byte aligned");
  GEMM_REQ(g->a_part_stride % 4 == 0 && g->d_part_stride % 4 == 0, "part strides must be multiples of 4 floats");
  if (ap > 1) {
    if (g->a_major == 0) GEMM_REQ(g->K % ap == 0 && (g->K / ap) % BK == 0, "K per A part must be a multiple of 32");
    else GEMM_REQ(g->M % ap == 0 && (g->M / ap) % BM == 0, "M per A part must be a multiple of 128");
  }
  if (dp > 1) GEMM_REQ(!g->residual && !g->D2 && !g->Zin, "parted outputs take no residual / activation tensors");
  GEMM_REQ(g->epilogue == B200GAT_EPI_NONE || g->epilogue == B200GAT_EPI_GELU || g->epilogue == B200GAT_EPI_DGELU,
           "unknown epilogue");
  if (g->epilogue == B200GAT_EPI_GELU) GEMM_REQ(g->D2 && g->ld2 % 4 == 0, "EPI_GELU needs D2");
  if (g->epilogue == B200GAT_EPI_DGELU)
    GEMM_REQ(g->Zin && g->ld2 % 4 == 0 && g->D && !g->residual, "EPI_DGELU needs Zin and D and takes no residual");
  if (g->epilogue == B200GAT_EPI_NONE) GEMM_REQ(g->D != nullptr, "null D");
  if (g->residual) GEMM_REQ(g->ldr % 4 == 0, "ldr must be a multiple of 4");
  s->m_tiles = (int)ceil_div(g->M, BM);
  s->n_tiles = (int)ceil_div(g->N, s->bn);
  s->kb_total = (int)ceil_div(g->K, BK);
  // split the reduction when the tile grid alone cannot fill the machine (weight gradients: K = number of nodes)
  int splits = 1;
  const int tiles = s->m_tiles * s->n_tiles;
  if (tiles < kNumSMs / 2 && s->kb_total >= 16 && g->workspace) {
    splits = (int)ceil_div(kNumSMs, tiles);
    const int max_by_k = s->kb_total / 4;
    if (splits > max_by_k) splits = max_by_k;
    if (splits < 1) splits = 1;
  }
  s->kb_per_split = (int)ceil_div(s->kb_total, splits);
  s->splits = (int)ceil_div(s->kb_total, s->kb_per_split);
  if (s->splits > 1) {
    GEMM_REQ(g->epilogue == B200GAT_EPI_NONE && !g->residual && !g->bias && dp == 1, "split-K takes a plain epilogue");
    GEMM_REQ(g->ldd == g->N, "split-K needs a contiguous output");
    GEMM_REQ(aligned16(g->workspace) && (g->M * g->N) % 4 == 0, "workspace alignment");
    if (g->a_rowsum) GEMM_REQ(g->M % 4 == 0 && aligned16(g->a_rowsum), "row sums under split-K need M % 4 == 0");
    GEMM_REQ(g->workspace_elems >= (long long)s->splits * (g->M * g->N + (g->a_rowsum ? g->M : 0)),
             "workspace too small");
  }
  return B200GAT_OK;
#undef GEMM_REQ
}

template <int BN, bool A_MN, bool B_MN, bool X3>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& a, dim3 grid, cudaStream_t st) {
  constexpr size_t smem = (X3 ? (size_t)2 * 2 : (size_t)STAGES) * (A_STAGE_BYTES + BN * BK * 4) + 1024 + 2048;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(k_gemm_tf32<BN, A_MN, B_MN, X3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
        cudaSuccess) {
      set_error("k_gemm_tf32: cudaFuncSetAttribute(%zu) failed", smem);
      return B200GAT_ERR_CUDA;
    }
    attr_set = true;
  }
  // 3xTF32 tiles fill an SM alone (two stages of A | B | A_lo | B_lo): a dependent launch cannot become resident early,
  // and measured 1 % slower with the attribute than without
  if (X3) k_gemm_tf32<BN, A_MN, B_MN, X3><<<grid, 320, smem, st>>>(ta, tb, a);
  else launch_k(k_gemm_tf32<BN, A_MN, B_MN, X3>, grid, dim3(320), smem, st, ta, tb, a);
  B200_CHECK_LAUNCH("k_gemm_tf32");
  return B200GAT_OK;
}

}  // namespace
}  // namespace b200gat

using namespace b200gat;

extern "C" void b200gat_gemm_precision(int mode) { g_gemm_x3.store(mode == B200GAT_GEMM_FP32X3 ? 1 : 0); }
extern "C" int b200gat_gemm_precision_get(void) {
  return g_gemm_x3.load() ? B200GAT_GEMM_FP32X3 : B200GAT_GEMM_TF32;
}

extern "C" int b200gat_gemm_supported(const b200gat_gemm* g) {
  if (!g) return 0;
  Shape s;
  return plan_gemm(g, &s, false) == B200GAT_OK ? 1 : 0;
}

extern "C" int64_t b200gat_gemm_workspace_elems(const b200gat_gemm* g) {
  if (!g) return 0;
  b200gat_gemm probe = *g;
  static float dummy_ws[4];
  probe.workspace = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(dummy_ws) + 15) & ~(uintptr_t)15);
  probe.workspace_elems = INT64_MAX;
  Shape s;
  if (plan_gemm(&probe, &s, false) != B200GAT_OK || s.splits <= 1) return 0;
  return (int64_t)s.splits * (g->M * g->N + (g->a_rowsum ? g->M : 0));
}

extern "C" int b200gat_gemm_tf32(const b200gat_gemm* g, void* stream) {
  B200_CHECK_ARG(g, "null descriptor");
  Shape s;
  int rc = plan_gemm(g, &s, true);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int ap = g->a_parts > 1 ? g->a_parts : 1, dp = g->d_parts > 1 ? g->d_parts : 1;
  const bool a_mn = g->a_major != 0, b_mn = g->b_major != 0;
  CUtensorMap ta, tb;
  if (!a_mn) rc = make_map(&ta, g->A, g->K / ap, g->M, g->lda, ap, g->a_part_stride, BK, BM, false);
  else rc = make_map(&ta, g->A, g->M / ap, g->K, g->lda, ap, g->a_part_stride, 32, BK, true);
  if (rc) return rc;
  if (!b_mn) rc = make_map(&tb, g->B, g->K, g->N, g->ldb, 1, 0, BK, s.bn, false);
  else rc = make_map(&tb, g->B, g->N, g->K, g->ldb, 1, 0, 32, BK, true);
  if (rc) return rc;

  GemmArgs a;
  a.M = (int)g->M; a.N = (int)g->N;
  a.kb_total = s.kb_total; a.kb_per_split = s.kb_per_split;
  a.a_kb_per_part = (!a_mn && ap > 1) ? (int)(g->K / ap / BK) : (1 << 30);
  a.a_mt_per_part = (a_mn && ap > 1) ? (int)(g->M / ap / BM) : (1 << 30);
  a.n_tiles_per_part = dp > 1 ? (int)(g->N / dp / s.bn) : (1 << 30);
  a.ldd = g->ldd; a.d_part_stride = g->d_part_stride; a.d_split_stride = 0;
  a.D = g->D;
  if (s.splits > 1) {
    a.D = g->workspace;
    a.d_split_stride = g->M * g->N;
  }
  a.bias = g->bias; a.R = g->residual; a.ldr = g->ldr;
  a.epi = g->epilogue; a.D2 = g->D2; a.Zin = g->Zin; a.ld2 = g->ld2;
  a.mn_lbo = env_u("B200GAT_MN_LBO", 4096);
  a.mn_sbo = env_u("B200GAT_MN_SBO", 512);
  a.mn_layout = env_u("B200GAT_MN_LAYOUT", 1);
  a.mn_kadv = env_u("B200GAT_MN_KADV", 1024);
  a.a_rowsum = g->a_rowsum; a.rowsum_split_stride = 0;
  if (s.splits > 1 && g->a_rowsum) {
    a.a_rowsum = g->workspace + (long long)s.splits * g->M * g->N;
    a.rowsum_split_stride = g->M;
  }
  dim3 grid((unsigned)s.m_tiles, (unsigned)s.n_tiles, (unsigned)s.splits);
  ScopedTimer tm(a_mn ? B200GAT_TIMING_GEMM_DW : (b_mn ? B200GAT_TIMING_GEMM_DX : B200GAT_TIMING_GEMM), st);
#define GEMM_LAUNCH2(BN_, X3_)                                                     \
  (a_mn ? (b_mn ? launch<BN_, true, true, X3_>(ta, tb, a, grid, st)                \
                : launch<BN_, true, false, X3_>(ta, tb, a, grid, st))              \
        : (b_mn ? launch<BN_, false, true, X3_>(ta, tb, a, grid, st)               \
                : launch<BN_, false, false, X3_>(ta, tb, a, grid, st)))
#define GEMM_LAUNCH(BN_) (g_gemm_x3.load(std::memory_order_relaxed) ? GEMM_LAUNCH2(BN_, true) : GEMM_LAUNCH2(BN_, false))
  rc = s.bn == 256 ? GEMM_LAUNCH(256) : (s.bn == 128 ? GEMM_LAUNCH(128) : GEMM_LAUNCH(64));
#undef GEMM_LAUNCH
#undef GEMM_LAUNCH2
  if (rc) return rc;
  if (s.splits > 1) {
    const long long n4 = g->M * g->N / 4, r4 = g->a_rowsum ? g->M / 4 : 0;
    const unsigned blocks = (unsigned)(ceil_div(n4, 32) + ceil_div(r4, 32));
    launch_k(k_splitk_reduce, dim3(blocks), dim3(256), 0, st, g->workspace, s.splits, g->M * g->N, n4, g->D, a.a_rowsum,
             g->M, r4, g->a_rowsum);
    B200_CHECK_LAUNCH("k_splitk_reduce");
  }
  return B200GAT_OK;
}

extern "C" int64_t b200gat_colsum_scratch_elems(int64_t M, int64_t N) {
  if (M <= 0 || N <= 0) return 0;
  const int64_t slabs = M < 2 * kNumSMs ? M : 2 * kNumSMs;
  return slabs * N;
}

extern "C" int b200gat_colsum(const float* X, int64_t M, int64_t N, int64_t ld, float* out, float* scratch,
                              void* stream) {
  B200_CHECK_ARG(N > 0 && out, "null output / empty width");
  cudaStream_t st = (cudaStream_t)stream;
  if (M <= 0) {
    if (cudaMemsetAsync(out, 0, (size_t)N * sizeof(float), st) != cudaSuccess) return B200GAT_ERR_CUDA;
    return B200GAT_OK;
  }
  B200_CHECK_ARG(X && scratch && ld >= N, "null input / scratch");
  const int64_t slabs_max = M < 2 * kNumSMs ? M : 2 * kNumSMs;
  const int rows = (int)ceil_div(M, slabs_max);
  const int slabs = (int)ceil_div(M, rows);
  ScopedTimer tm(B200GAT_TIMING_GEMM, st);
  k_colsum_partial<<<slabs, 256, 0, st>>>(X, M, (int)N, ld, rows, scratch);
  k_colsum_finish<<<(unsigned)ceil_div(N, 8), 256, 0, st>>>(scratch, slabs, (int)N, out);
  B200_CHECK_LAUNCHES("k_colsum", 2);
  return B200GAT_OK;
}

// ------------------------------------------------------------------------------------------- no-allocation variant
// b200gat_colsum for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc): the slab partials live in a file-scope device array instead of a caller-provided scratch buffer.
// 2 * 148 slabs x B200GAT_COLSUM_NOALLOC_MAX_N columns x 4 B = 1.16 MiB, held by every GPU that loads the library
// until the process exits.  Every call shares the array: calls that overlap on the GPU (two streams, two threads)
// race on it.  Nothing has to start at zero: k_colsum_partial writes every partial k_colsum_finish reads.
static __device__ __align__(256) float g_colsum_static[2 * kNumSMs * B200GAT_COLSUM_NOALLOC_MAX_N];

// Device address of the array on the CURRENT device, resolved once per device (in host code the symbol's name is a
// host-side stand-in, not a device pointer).
static int colsum_static_ptr(float** ptr) {
  constexpr int kMaxDev = 64;
  static std::atomic<float*> cache[kMaxDev];
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDev) {
    set_error("b200gat_colsum_noalloc: no current CUDA device (or device ordinal >= %d)", kMaxDev);
    return B200GAT_ERR_CUDA;
  }
  float* p = cache[dev].load(std::memory_order_acquire);
  if (!p) {
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static);
    if (e != cudaSuccess || !p) {
      set_error("b200gat_colsum_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return B200GAT_ERR_CUDA;
    }
    cache[dev].store(p, std::memory_order_release);
  }
  *ptr = p;
  return B200GAT_OK;
}

extern "C" int b200gat_colsum_noalloc(const float* X, int64_t M, int64_t N, int64_t ld, float* out, void* stream) {
  B200_CHECK_ARG(N > 0 && out, "null output / empty width");
  if (N > B200GAT_COLSUM_NOALLOC_MAX_N) {
    set_error("b200gat_colsum_noalloc: N = %lld exceeds the static scratch (B200GAT_COLSUM_NOALLOC_MAX_N = %d "
              "columns); use b200gat_colsum with a caller-provided scratch buffer",
              (long long)N, B200GAT_COLSUM_NOALLOC_MAX_N);
    return B200GAT_ERR_UNSUPPORTED;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (M <= 0) {
    if (cudaMemsetAsync(out, 0, (size_t)N * sizeof(float), st) != cudaSuccess) return B200GAT_ERR_CUDA;
    return B200GAT_OK;
  }
  B200_CHECK_ARG(X && ld >= N, "null input / leading dimension smaller than the width");
  float* scratch = nullptr;
  int rc = colsum_static_ptr(&scratch);
  if (rc) return rc;
  // same slab decomposition as b200gat_colsum: identical summation order, identical result
  const int64_t slabs_max = M < 2 * kNumSMs ? M : 2 * kNumSMs;
  const int rows = (int)ceil_div(M, slabs_max);
  const int slabs = (int)ceil_div(M, rows);
  k_colsum_partial<<<slabs, 256, 0, st>>>(X, M, (int)N, ld, rows, scratch);
  k_colsum_finish<<<(unsigned)ceil_div(N, 8), 256, 0, st>>>(scratch, slabs, (int)N, out);
  B200_CHECK_LAUNCHES("k_colsum(noalloc)", 2);
  return B200GAT_OK;
}
