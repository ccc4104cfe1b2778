// bf16 tensor-core path: the conv / fully-connected forward, weight-gradient and data-gradient of the
// Q-network as implicit GEMMs on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
// Same math and the same fused epilogues as gemm_core.cuh (reference: src/network.py:395-424 and the
// TF gradients of those ops); operands are bf16 (uint8 frames are exact in bf16), accumulation fp32.
//
// How a tile is computed
//   * 128 threads gather the operands (im2col of NHWC activations, transposed weights, or the
//     pool/ReLU-routed output gradient) as 16-byte chunks of 8 bf16 and store them into shared memory
//     in the UMMA canonical NO-SWIZZLE layout: 8x16-byte core matrices; the chunk of (mn-index i,
//     reduction-index r) lives at   K-major : (i/8)*1024 + (r/8)*128 + (i%8)*16      (forward, dgrad)
//                                   MN-major: (i/8)*1024 + (r/8)*128 + (r%8)*16      (wgrad: both operands)
//     so one 64-deep reduction stage is 1024 bytes per 8 mn-indices and UMMA_K=16 steps are 256 bytes apart.
//   * one thread issues 4 tcgen05.mma (M=128, N=tile width, K=16) per stage and commits them to an
//     mbarrier; two stages ping-pong so the gather of stage i+1 overlaps the MMAs of stage i.
//   * 4 warps read the accumulator with tcgen05.ld (warp w owns TMEM lanes 32w..32w+31, lane = tile row)
//     and run the epilogue: bias + ReLU + 2x2 max-pool across 4 neighbouring lanes (forward), fp32 partial
//     sums (wgrad), bf16 gradient (dgrad).
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

#include "gemm_core.cuh"

namespace dqn {
namespace umma {

constexpr int TM = 128;        // tile rows (UMMA M)
constexpr int TK = 64;         // reduction elements per stage (4 x UMMA_K)
constexpr int TT = 128;        // threads per CTA
constexpr uint32_t A_STAGE_BYTES = TM * TK * 2;   // 16 KB

// ------------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "DONE:\n\t"
      "}" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void fence_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {   // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {     // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// shared-memory matrix descriptor, no swizzle (cute::UMMA::SmemDescriptor: start[0,14) LBO[16,30) SBO[32,46) version[46,48)=1)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = (uint64_t)((saddr >> 4) & 0x3FFFu);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor for kind::f16 (cute::UMMA::InstrDescriptor): D=f32, A=B=bf16
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// 32 consecutive accumulator columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float v[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xFFFF0000u); }

// byte offset of the 16-byte chunk (mn-index i, reduction chunk rc of 8) in a K-major stage
__device__ __forceinline__ uint32_t kmajor_off(int i, int rc) { return (uint32_t)((i >> 3) * 1024 + rc * 128 + (i & 7) * 16); }
// byte offset of the 16-byte chunk (mn chunk g of 8, reduction index r) in an MN-major stage
__device__ __forceinline__ uint32_t mnmajor_off(int g, int r) { return (uint32_t)(g * 1024 + (r >> 3) * 128 + (r & 7) * 16); }

// issue the 4 MMAs of one stage (thread 0 only) and commit them to `bar`
__device__ __forceinline__ void issue_stage(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc, bool first,
                                            uint64_t* bar) {
  tc_fence_after();
#pragma unroll
  for (int j = 0; j < TK / 16; ++j) {
    const uint64_t ad = make_desc(a_addr + j * 256, 128, 1024);
    const uint64_t bd = make_desc(b_addr + j * 256, 128, 1024);
    mma_bf16(tmem_d, ad, bd, idesc, (first && j == 0) ? 0u : 1u);
  }
  mma_commit(bar);
}

// 8 consecutive input channels (bf16) or 2 pixels x 4 channels (uint8 conv1) of the input pixel that reduction
// chunk `kc8` (k = 8*kc8 .. 8*kc8+7 in (ky,kx,c) order) of output row `r` reads; zero outside the image.
template <class G, typename TIn>
__device__ __forceinline__ uint4 load_x_chunk(const TIn* __restrict__ x, const Row& r, int kc8) {
  uint4 v = make_uint4(0u, 0u, 0u, 0u);
  if (kc8 * 8 >= G::K || !r.valid) return v;
  if constexpr (sizeof(TIn) == 1) {
    static_assert(sizeof(TIn) != 1 || G::CIN == 4, "uint8 input is the 4-frame stack");
    const int t = kc8 * 2;                                // first of two taps (pixels)
    const int kx = t % G::KS, ky = t / G::KS;
    const int iy = r.oy * G::ST - G::PADH + ky, ix = r.ox * G::ST - G::PADW + kx;   // ix even, pair never straddles the border
    if ((unsigned)iy < (unsigned)G::IH && (unsigned)ix < (unsigned)G::IW) {
      const uint2 u = __ldg(reinterpret_cast<const uint2*>(x + ((size_t)(r.b * G::IH + iy) * G::IW + ix) * 4));
      v.x = pack_bf16((float)(u.x & 255u), (float)((u.x >> 8) & 255u));
      v.y = pack_bf16((float)((u.x >> 16) & 255u), (float)(u.x >> 24));
      v.z = pack_bf16((float)(u.y & 255u), (float)((u.y >> 8) & 255u));
      v.w = pack_bf16((float)((u.y >> 16) & 255u), (float)(u.y >> 24));
    }
  } else {
    constexpr int CC = G::CIN / 8;                        // chunks per tap
    const int c0 = (kc8 % CC) * 8;
    const int t = kc8 / CC;
    const int kx = t % G::KS, ky = t / G::KS;
    const int iy = r.oy * G::ST - G::PADH + ky, ix = r.ox * G::ST - G::PADW + kx;
    if ((unsigned)iy < (unsigned)G::IH && (unsigned)ix < (unsigned)G::IW)
      v = __ldg(reinterpret_cast<const uint4*>(x + ((size_t)(r.b * G::IH + iy) * G::IW + ix) * G::CIN + c0));
  }
  return v;
}

// 8 consecutive channels of the gradient w.r.t. the conv output, rebuilt from the gradient of the pooled
// activation (bf16), the pool arg-max and the ReLU gate; optionally adds them to 8 running column sums.
template <class G>
__device__ __forceinline__ uint4 load_dz_chunk(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* __restrict__ act,
                                               const uint8_t* __restrict__ arg, int pidx, int win, int n0, float* colsum) {
  const size_t off = (size_t)pidx * G::COUT + n0;
  uint4 d = *reinterpret_cast<const uint4*>(dact + off);
  const uint4 p = *reinterpret_cast<const uint4*>(act + off);
  uint32_t dw[4] = {d.x, d.y, d.z, d.w};
  const uint32_t pw[4] = {p.x, p.y, p.z, p.w};
  uint32_t aw[2] = {0u, 0u};
  if constexpr (G::POOL) {
    const uint2 a = *reinterpret_cast<const uint2*>(arg + off);
    aw[0] = a.x; aw[1] = a.y;
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    // bf16 > 0  <=>  sign bit clear and magnitude non-zero
    const bool lo_on = ((pw[q] & 0x7FFFu) != 0u) && !(pw[q] & 0x8000u);
    const bool hi_on = ((pw[q] & 0x7FFF0000u) != 0u) && !(pw[q] & 0x80000000u);
    bool lo = lo_on, hi = hi_on;
    if constexpr (G::POOL) {
      const uint32_t a2 = (aw[q >> 1] >> ((q & 1) * 16)) & 0xFFFFu;
      lo = lo && ((a2 & 255u) == (uint32_t)win);
      hi = hi && ((a2 >> 8) == (uint32_t)win);
    }
    dw[q] = (lo ? (dw[q] & 0xFFFFu) : 0u) | (hi ? (dw[q] & 0xFFFF0000u) : 0u);
    if (colsum) { colsum[2 * q] += bf16_lo(dw[q]); colsum[2 * q + 1] += bf16_hi(dw[q]); }
  }
  return make_uint4(dw[0], dw[1], dw[2], dw[3]);
}

// Common prologue: barriers + TMEM allocation.  Returns the TMEM base address.
__device__ __forceinline__ uint32_t setup(uint64_t* bars, uint32_t* tmem_slot, uint32_t ncols) {
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    __syncwarp();
    tmem_alloc(tmem_slot, ncols);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  return *tmem_slot;
}
__device__ __forceinline__ void teardown(uint32_t tmem_base, uint32_t ncols) {
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem_base, ncols);
}
__host__ __device__ constexpr uint32_t tmem_cols(int n) { return n <= 32 ? 32u : (n <= 64 ? 64u : (n <= 128 ? 128u : 256u)); }

// ------------------------------------------------------------------------------------------------ forward
struct FwdPassT {
  const void* x;               // uint8 [B,84,84,4] (conv1) or bf16 NHWC activations
  const __nv_bfloat16* wt;     // transposed weights [COUT][K]
  const float* bias;
  __nv_bfloat16* out;          // pooled activation, bf16 NHWC
  float* out_f32;              // optional fp32 copy (fc output feeds the fp32 head kernels)
  uint8_t* arg;                // pool arg-max
};
struct FwdArgsT {
  FwdPassT p[2];
  int B;
};

template <class G, typename TIn, int NTILE>
__global__ void __launch_bounds__(TT) umma_fwd_kernel(FwdArgsT a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr uint32_t B_STAGE = NTILE * TK * 2;
  __shared__ uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  const FwdPassT P = a.p[blockIdx.z];
  const TIn* __restrict__ x = static_cast<const TIn*>(P.x);
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * TM, n0 = blockIdx.y * NTILE;
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sA = smem_u32(smem), sB = sA + 2 * A_STAGE_BYTES;
  const Row row = decode_row<G>(m0 + tid, a.B);
  constexpr int T = (G::K + TK - 1) / TK;
  constexpr int BCH = NTILE * 8 / TT;                 // weight chunks per thread per stage
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  uint4 ra[8], rb[BCH];
  auto fetch = [&](int it) {
#pragma unroll
    for (int c = 0; c < 8; ++c) ra[c] = load_x_chunk<G, TIn>(x, row, it * 8 + c);
#pragma unroll
    for (int q = 0; q < BCH; ++q) {
      const int ch = tid + TT * q, n = ch >> 3, kc = ch & 7;
      const int k = it * TK + kc * 8;
      rb[q] = (k < G::K) ? __ldg(reinterpret_cast<const uint4*>(P.wt + (size_t)(n0 + n) * G::K + k)) : make_uint4(0u, 0u, 0u, 0u);
    }
  };
  fetch(0);
  for (int it = 0; it < T; ++it) {
    const int s = it & 1;
    if (it >= 2) mbar_wait(&bars[s], ((it >> 1) - 1) & 1);     // MMAs of iteration it-2 have drained stage s
    uint8_t* A = smem + s * A_STAGE_BYTES;
    uint8_t* Bm = smem + 2 * A_STAGE_BYTES + s * B_STAGE;
#pragma unroll
    for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(A + kmajor_off(tid, c)) = ra[c];
#pragma unroll
    for (int q = 0; q < BCH; ++q) {
      const int ch = tid + TT * q;
      *reinterpret_cast<uint4*>(Bm + kmajor_off(ch >> 3, ch & 7)) = rb[q];
    }
    fence_async_proxy();
    __syncthreads();
    if (tid == 0) issue_stage(tmem, sA + s * A_STAGE_BYTES, sB + s * B_STAGE, idesc, it == 0, &bars[s]);
    if (it + 1 < T) fetch(it + 1);
  }
  // the last commit covers every MMA issued before it
  mbar_wait(&bars[(T - 1) & 1], ((T - 1) >> 1) & 1);
  tc_fence_after();
  const int warp = tid >> 5, lane = tid & 31;
  const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
  const bool row_in = row.valid;
#pragma unroll 1
  for (int c0 = 0; c0 < NTILE; c0 += 32) {
    float v[32];
    tmem_ld32(trow + c0, v);
    const int n = n0 + c0;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + __ldg(P.bias + n + j), 0.f);
    if constexpr (G::POOL) {
      // rows 4q..4q+3 (= lanes) are one pool window; invalid rows never win; first maximum wins ties
      uint32_t idx[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float best = row_in ? v[j] : -1.f;
        uint32_t bi = (uint32_t)(lane & 3);
#pragma unroll
        for (int sft = 1; sft <= 2; sft <<= 1) {
          const float ov = __shfl_xor_sync(0xffffffffu, best, sft);
          const uint32_t oi = __shfl_xor_sync(0xffffffffu, bi, sft);
          if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        }
        v[j] = best;
        idx[j] = bi;
      }
      if ((lane & 3) == 0 && row.b < a.B) {
        const size_t off = (size_t)row.pidx * G::COUT + n;
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(P.out + off + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        if (P.arg) {
          uint32_t ak[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            ak[j] = idx[4 * j] | (idx[4 * j + 1] << 8) | (idx[4 * j + 2] << 16) | (idx[4 * j + 3] << 24);
          *reinterpret_cast<uint4*>(P.arg + off) = make_uint4(ak[0], ak[1], ak[2], ak[3]);
          *reinterpret_cast<uint4*>(P.arg + off + 16) = make_uint4(ak[4], ak[5], ak[6], ak[7]);
        }
      }
    } else {
      if (row_in) {
        const size_t off = (size_t)row.pidx * G::COUT + n;
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(P.out + off + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        if (P.out_f32) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(P.out_f32 + off + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
    }
  }
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ weight gradient
struct WgradArgsT {
  const void* x;                  // layer input: uint8 stacks (conv1) or bf16 activations
  const __nv_bfloat16* dact;      // gradient of the layer's pooled output
  const __nv_bfloat16* act;       // the layer's pooled output
  const uint8_t* arg;
  float* partial;                 // [nsplit][(K+1)*COUT] fp32
  int B;
  int rows_per_split;             // multiple of TK
};

template <class G, typename TIn, int NTILE>
__global__ void __launch_bounds__(TT) umma_wgrad_kernel(WgradArgsT a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr uint32_t B_STAGE = NTILE * TK * 2;
  __shared__ uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[NTILE];
  const TIn* __restrict__ x = static_cast<const TIn*>(a.x);
  const int tid = threadIdx.x;
  const int i0 = blockIdx.x * TM, n0 = blockIdx.y * NTILE;     // i = (ky,kx,c) index, rows of the output tile
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sA = smem_u32(smem), sB = sA + 2 * A_STAGE_BYTES;
  const int R = a.B * G::ROWS;
  const int r_begin = blockIdx.z * a.rows_per_split;
  const int r_end = min(R, r_begin + a.rows_per_split);
  const int T = (r_end - r_begin + TK - 1) / TK;
  const bool do_bias = blockIdx.x == 0;
  for (int j = tid; j < NTILE; j += TT) bias_sm[j] = 0.f;
  __syncthreads();
  constexpr uint32_t idesc = make_idesc(TM, NTILE, true, true);
  // A: 64 rows x 16 chunks (of 8 k) per stage -> 8 per thread; B: 64 rows x NTILE/8 chunks -> NTILE/16 per thread
  constexpr int BCH = NTILE / 16;
  const int am = tid & 63, ag0 = tid >> 6;            // A chunks g = ag0 + 2q of row am
  uint4 ra[8], rb[BCH];
  float colsum[BCH][8];
#pragma unroll
  for (int q = 0; q < BCH; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e) colsum[q][e] = 0.f;
  // B chunk assignment: chunk id = tid + TT*q ; row = id / (NTILE/8), g = id % (NTILE/8)  (g is the same for every q
  // when TT is a multiple of NTILE/8, so each thread owns fixed output columns for the bias sum)
  constexpr int GPR = NTILE / 8;
  const int bg = tid % GPR;
  auto fetch = [&](int it) {
    const int rbase = r_begin + it * TK;
    const Row ar = decode_row<G>(rbase + am, a.B);
    const bool arow_ok = (rbase + am) < r_end;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int g = ag0 + 2 * q;
      ra[q] = arow_ok ? load_x_chunk<G, TIn>(x, ar, (i0 >> 3) + g) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int q = 0; q < BCH; ++q) {
      const int id = tid + TT * q;
      const int rr = id / GPR;
      const Row br = decode_row<G>(rbase + rr, a.B);
      rb[q] = (br.valid && (rbase + rr) < r_end)
                  ? load_dz_chunk<G>(a.dact, a.act, a.arg, br.pidx, br.win, n0 + bg * 8, do_bias ? colsum[q] : nullptr)
                  : make_uint4(0u, 0u, 0u, 0u);
    }
  };
  if (T > 0) fetch(0);
  for (int it = 0; it < T; ++it) {
    const int s = it & 1;
    if (it >= 2) mbar_wait(&bars[s], ((it >> 1) - 1) & 1);
    uint8_t* A = smem + s * A_STAGE_BYTES;
    uint8_t* Bm = smem + 2 * A_STAGE_BYTES + s * B_STAGE;
#pragma unroll
    for (int q = 0; q < 8; ++q) *reinterpret_cast<uint4*>(A + mnmajor_off(ag0 + 2 * q, am)) = ra[q];
#pragma unroll
    for (int q = 0; q < BCH; ++q) {
      const int id = tid + TT * q;
      *reinterpret_cast<uint4*>(Bm + mnmajor_off(bg, id / GPR)) = rb[q];
    }
    fence_async_proxy();
    __syncthreads();
    if (tid == 0) issue_stage(tmem, sA + s * A_STAGE_BYTES, sB + s * B_STAGE, idesc, it == 0, &bars[s]);
    if (it + 1 < T) fetch(it + 1);
  }
  float* out = a.partial + (size_t)blockIdx.z * ((size_t)(G::K + 1) * G::COUT);
  const int warp = tid >> 5, lane = tid & 31;
  const int k = i0 + warp * 32 + lane;
  if (T > 0) {
    mbar_wait(&bars[(T - 1) & 1], ((T - 1) >> 1) & 1);
    tc_fence_after();
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
    for (int c0 = 0; c0 < NTILE; c0 += 32) {
      float v[32];
      tmem_ld32(trow + c0, v);
      if (k < G::K) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(out + (size_t)k * G::COUT + n0 + c0 + 4 * j) =
              make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
  } else if (k < G::K) {
    for (int c = 0; c < NTILE; ++c) out[(size_t)k * G::COUT + n0 + c] = 0.f;
  }
  if (do_bias) {
#pragma unroll
    for (int q = 0; q < BCH; ++q)
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(&bias_sm[bg * 8 + e], colsum[q][e]);
    __syncthreads();
    for (int j = tid; j < NTILE; j += TT) out[(size_t)G::K * G::COUT + n0 + j] = bias_sm[j];
  }
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ data gradient
struct DgradArgsT {
  const __nv_bfloat16* wd;        // weights permuted to [CIN][KS*KS*COUT]
  const __nv_bfloat16* dact;
  const __nv_bfloat16* act;
  const uint8_t* arg;
  __nv_bfloat16* dx;              // [B, IH, IW, CIN] gradient of the previous layer's pooled activation
  int B;
};

template <class G, int NTILE>
__global__ void __launch_bounds__(TT) umma_dgrad_kernel(DgradArgsT a) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr uint32_t B_STAGE = NTILE * TK * 2;
  constexpr int KD = G::KS * G::KS * G::COUT;
  constexpr int CC = G::COUT / 8;                     // chunks per filter tap
  __shared__ uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * TM, c0t = blockIdx.y * NTILE;
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sA = smem_u32(smem), sB = sA + 2 * A_STAGE_BYTES;
  const int M = a.B * G::IH * G::IW;
  const int m = m0 + tid;
  const int b = m / (G::IH * G::IW);
  const int rem = m - b * (G::IH * G::IW);
  const int iy = rem / G::IW, ix = rem - iy * G::IW;
  const bool mvalid = m < M;
  constexpr int T = KD / TK;
  static_assert(KD % TK == 0, "dgrad reduction must be a multiple of the stage depth");
  constexpr int BCH = NTILE * 8 / TT;
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  uint4 ra[8], rb[BCH];
  auto dz_at = [&](int kc8) -> uint4 {
    const int nq = kc8 % CC, t = kc8 / CC;
    const int kx = t % G::KS, ky = t / G::KS;
    const int ty = iy + G::PADH - ky, tx = ix + G::PADW - kx;
    if (!mvalid || ty < 0 || tx < 0 || (ty % G::ST) != 0 || (tx % G::ST) != 0) return make_uint4(0u, 0u, 0u, 0u);
    const int oy = ty / G::ST, ox = tx / G::ST;
    if (oy >= G::OH || ox >= G::OW) return make_uint4(0u, 0u, 0u, 0u);
    int pidx, win;
    if constexpr (G::POOL) {
      pidx = (b * G::PH + (oy >> 1)) * G::PW + (ox >> 1);
      win = (oy & 1) * 2 + (ox & 1);
    } else {
      pidx = (b * G::OH + oy) * G::OW + ox;
      win = 0;
    }
    return load_dz_chunk<G>(a.dact, a.act, a.arg, pidx, win, nq * 8, nullptr);
  };
  auto fetch = [&](int it) {
#pragma unroll
    for (int c = 0; c < 8; ++c) ra[c] = dz_at(it * 8 + c);
#pragma unroll
    for (int q = 0; q < BCH; ++q) {
      const int ch = tid + TT * q, n = ch >> 3, kc = ch & 7;
      rb[q] = __ldg(reinterpret_cast<const uint4*>(a.wd + (size_t)(c0t + n) * KD + it * TK + kc * 8));
    }
  };
  fetch(0);
  for (int it = 0; it < T; ++it) {
    const int s = it & 1;
    if (it >= 2) mbar_wait(&bars[s], ((it >> 1) - 1) & 1);
    uint8_t* A = smem + s * A_STAGE_BYTES;
    uint8_t* Bm = smem + 2 * A_STAGE_BYTES + s * B_STAGE;
#pragma unroll
    for (int c = 0; c < 8; ++c) *reinterpret_cast<uint4*>(A + kmajor_off(tid, c)) = ra[c];
#pragma unroll
    for (int q = 0; q < BCH; ++q) {
      const int ch = tid + TT * q;
      *reinterpret_cast<uint4*>(Bm + kmajor_off(ch >> 3, ch & 7)) = rb[q];
    }
    fence_async_proxy();
    __syncthreads();
    if (tid == 0) issue_stage(tmem, sA + s * A_STAGE_BYTES, sB + s * B_STAGE, idesc, it == 0, &bars[s]);
    if (it + 1 < T) fetch(it + 1);
  }
  mbar_wait(&bars[(T - 1) & 1], ((T - 1) >> 1) & 1);
  tc_fence_after();
  const int warp = tid >> 5;
  const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll 1
  for (int c0 = 0; c0 < NTILE; c0 += 32) {
    float v[32];
    tmem_ld32(trow + c0, v);
    if (mvalid) {
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(a.dx + (size_t)m * G::CIN + c0t + c0 + 8 * j) =
            make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    }
  }
  teardown(tmem, tmem_cols(NTILE));
}

}  // namespace umma
}  // namespace dqn
