// bf16 tensor-core path: the conv / fully-connected forward, weight-gradient and data-gradient of the
// Q-network as implicit GEMMs on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM).
// Same math and the same fused epilogues as gemm_core.cuh (reference: src/network.py:395-424 and the
// TF gradients of those ops); operands are bf16 (uint8 frames are exact in bf16), accumulation fp32.
//
// Two kernel families (DESIGN.md section 4 has the measurements behind the choices):
//
// RESIDENT-INPUT kernels (umma_fwd_res_kernel, umma_fwd_res_u8_kernel, umma_dgrad_res_kernel,
// umma_wgrad_res_u8_kernel) -- the conv layers of the training step:
//   * the tile's input (a contiguous NHWC block of a few samples, or a band of <= 40 uint8 frame rows) is fetched once
//     with a bulk async copy (cp.async.bulk, completion counted on an mbarrier); all weight stages are fetched the same
//     way, before the PDL wait;
//   * operand tiles are built shared->shared (16-byte copies into the SWIZZLE_128B K-major layout: row i at
//     (i/8)*1024 + (i%8)*128, chunk c at position c ^ (i%8); uint8 frames are widened to bf16 once per tile; pooled
//     gradients are routed to their arg-max position on the way) in big stages of several 64-deep k-blocks;
//   * one elected lane of warp 0 issues the stage's tcgen05.mma (M=128, N=tile width, K=16) back to back and commits
//     them to an mbarrier; the conv1 kernels walk over consecutive tiles per CTA.
// RING kernels (umma_fwd_kernel, umma_wgrad_kernel, umma_dgrad_kernel) -- fc layers, conv2/conv3 weight gradients, the
// Nature variant's conv1, and everything when DQN_NO_RESIDENT=1:
//   * 512 threads gather the operands (im2col of NHWC activations or the pool-routed output gradient) as 16-byte
//     chunks with cp.async into an N-slot ring several stages ahead; lanes run along the contiguous direction of global
//     memory.  Operands that need no arithmetic land directly in the swizzled layout; uint8 frames, pool-routed
//     gradients and the MN-major wgrad operands land in a padded raw ring and are rewritten shared->shared (no-swizzle
//     canonical layout: K-major chunk (i,r/8) at (i/8)*1024 + (r/8)*128 + (i%8)*16, MN-major chunk (i/8, r) at
//     (i/8)*1024 + (r/8)*128 + (r%8)*16).
// Both: weights live in global memory as pre-swizzled stage images (img_index) so that a stage's weight tile is one
// bulk copy; accumulators live in TMEM; warp w reads TMEM lanes 32*(w%4).. with tcgen05.ld (lane = tile row) and runs
// the epilogue: bias + ReLU + 2x2 max-pool across 4 neighbouring lanes (forward), fp32 partial sums (wgrad), ReLU-gated
// bf16 gradient (dgrad).
#pragma once
#include <cuda_bf16.h>
#include <cstdint>

#include "gemm_core.cuh"

namespace dqn {
namespace umma {

constexpr int TM = 128;        // tile rows (UMMA M)
constexpr int TK = 64;         // reduction elements per stage (4 x UMMA_K)
constexpr int TT = 512;        // threads per CTA: 16 warps gather operands, warp w reads TMEM lanes 32*(w%4)..
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
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
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
// One lane of a converged warp (the tensor-core instructions take their operands from the uniform datapath: issuing
// them under `if (tid == 0)` makes the compiler wrap every one of them in an elect loop).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ int warp_uniform_idx() { return __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0); }
// descriptor of the same operand `bytes` further on (bytes % 16 == 0; the 14-bit address field cannot carry: smem < 256 KB)
__device__ __forceinline__ uint64_t desc_advance(uint64_t d, uint32_t bytes) { return d + (uint64_t)(bytes >> 4); }
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

// 16 consecutive accumulator columns of this thread's TMEM lane; the caller issues tcgen05.wait::ld before using them
__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t r[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float v[16]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
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

// SWIZZLE_128B K-major descriptor (layout_type 2; 8-row groups 1024 bytes apart; tile base 1024-byte aligned)
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t saddr) {
  return make_desc(saddr, 16, 1024) | ((uint64_t)2 << 61);
}
// SWIZZLE_128B MN-major descriptor: the atom is 64 MN-elements (128 bytes, 16-byte chunks XOR-ed with k%8) x 8 k-rows 128
// bytes apart; LBO = distance between 64-element blocks along MN, SBO = distance between 8-row groups along K.  A
// [rows][64] tile stored as a SWIZZLE_128B K-major image (sw128_off) IS this layout with the roles of the axes swapped
// (rows = K): the forward A tiles double as the transposed operand of the weight gradient.
__device__ __forceinline__ uint64_t make_desc_sw128_mn(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return make_desc(saddr, lbo_bytes, sbo_bytes) | ((uint64_t)2 << 61);
}
// byte offset of chunk c (8 reduction elements) of row i in a SWIZZLE_128B K-major stage
__device__ __forceinline__ uint32_t sw128_off(int i, int c) { return (uint32_t)((i >> 3) * 1024 + (i & 7) * 128 + ((c ^ (i & 7)) << 4)); }

// issue the 4 MMAs of one stage and commit them to `bar`.  Called by ALL lanes of one converged warp (the branch
// leading here must be warp-uniform, see warp_uniform_idx); one elected lane issues.  *_SW: operand is in the swizzled
// K-major layout (UMMA_K steps are 32 bytes apart) instead of the no-swizzle one (256 bytes apart).
template <bool A_SW, bool B_SW>
__device__ __forceinline__ void issue_stage(uint32_t tmem_d, uint32_t a_addr, uint32_t b_addr, uint32_t idesc, bool first,
                                            uint64_t* bar) {
  tc_fence_after();
  if (elect_one()) {
    const uint64_t ad0 = A_SW ? make_desc_sw128(a_addr) : make_desc(a_addr, 128, 1024);
    const uint64_t bd0 = B_SW ? make_desc_sw128(b_addr) : make_desc(b_addr, 128, 1024);
#pragma unroll
    for (int j = 0; j < TK / 16; ++j)
      mma_bf16(tmem_d, desc_advance(ad0, j * (A_SW ? 32 : 256)), desc_advance(bd0, j * (B_SW ? 32 : 256)), idesc,
               (first && j == 0) ? 0u : 1u);
    mma_commit(bar);
  }
  __syncwarp();
}

// Weight operands live in global memory as "stage images": for a K-major operand matrix [rows][Kdim], stage s (64
// reduction elements) is one contiguous block of rows*128 bytes already in SWIZZLE_128B order, so the B tile of a
// stage (rows n0..n0+NTILE, n0 % 8 == 0) is ONE contiguous run of NTILE*128 bytes that a single thread fetches with a
// bulk async copy (TMA engine) straight into its ring slot; completion is counted on the slot's mbarrier.
__host__ __device__ inline size_t img_index(int row, int kk, int rows) {      // element (bf16) index of (row, kk)
  const int s = kk >> 6, c = (kk >> 3) & 7, e = kk & 7;
  return (size_t)s * rows * 64 + (size_t)((row >> 3) * 512 + (row & 7) * 64 + ((c ^ (row & 7)) << 3) + e);
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

constexpr int RS = 4;          // ring slots
// prefetch distance is 2 stages: slot (it+2)%4 was last read by the MMAs of stage it-2

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(valid ? 16 : 0) : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src, bool valid) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(valid ? 8 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Global address of the 8 input elements that reduction chunk `kc8` (k = 8*kc8.. in (ky,kx,c) order) of output
// row `r` reads: 8 bf16 channels of one pixel, or 2 pixels x 4 uint8 frames for conv1.  nullptr = zero (outside the
// image -- TF SAME padding --, beyond K, or an invalid row).
template <class G, typename TIn>
__device__ __forceinline__ const TIn* x_chunk_ptr(const TIn* __restrict__ x, const Row& r, int kc8) {
  if (kc8 * 8 >= G::K || !r.valid) return nullptr;
  int c0, t;
  if constexpr (sizeof(TIn) == 1) {
    static_assert(sizeof(TIn) != 1 || G::CIN == 4, "uint8 input is the 4-frame stack");
    c0 = 0;
    t = kc8 * 2;                       // first of two taps; the pair never straddles the border (ix is even)
  } else {
    constexpr int CC = G::CIN / 8;
    c0 = (kc8 % CC) * 8;
    t = kc8 / CC;
  }
  const int kx = t % G::KS, ky = t / G::KS;
  const int iy = r.oy * G::ST - G::PADH + ky, ix = r.ox * G::ST - G::PADW + kx;
  if ((unsigned)iy >= (unsigned)G::IH || (unsigned)ix >= (unsigned)G::IW) return nullptr;
  return x + ((size_t)(r.b * G::IH + iy) * G::IW + ix) * G::CIN + c0;
}

__device__ __forceinline__ uint4 cvt_u8x8(uint2 u) {     // 8 uint8 -> 8 bf16 (exact)
  uint4 v;
  v.x = pack_bf16((float)(u.x & 255u), (float)((u.x >> 8) & 255u));
  v.y = pack_bf16((float)((u.x >> 16) & 255u), (float)(u.x >> 24));
  v.z = pack_bf16((float)(u.y & 255u), (float)((u.y >> 8) & 255u));
  v.w = pack_bf16((float)((u.y >> 16) & 255u), (float)(u.y >> 24));
  return v;
}

// 8 uint8 -> 8 bf16 (exact) without the slow integer->float instruction: the byte goes into the low mantissa bits of
// 2^23 (0x4B000000 | n == 8388608 + n), one subtraction leaves n.
__device__ __forceinline__ uint4 widen_u8x8(uint2 u) {
  const float k = 8388608.f;
  uint4 v;
  v.x = pack_bf16(__uint_as_float(__byte_perm(u.x, 0x4B000000u, 0x7540)) - k, __uint_as_float(__byte_perm(u.x, 0x4B000000u, 0x7541)) - k);
  v.y = pack_bf16(__uint_as_float(__byte_perm(u.x, 0x4B000000u, 0x7542)) - k, __uint_as_float(__byte_perm(u.x, 0x4B000000u, 0x7543)) - k);
  v.z = pack_bf16(__uint_as_float(__byte_perm(u.y, 0x4B000000u, 0x7540)) - k, __uint_as_float(__byte_perm(u.y, 0x4B000000u, 0x7541)) - k);
  v.w = pack_bf16(__uint_as_float(__byte_perm(u.y, 0x4B000000u, 0x7542)) - k, __uint_as_float(__byte_perm(u.y, 0x4B000000u, 0x7543)) - k);
  return v;
}

// Pool routing of 8 (already ReLU-gated) gradient values: keep element e iff the window's arg-max is `win`.
__device__ __forceinline__ uint4 route_chunk(uint4 d, uint2 arg, int win) {
  uint32_t dw[4] = {d.x, d.y, d.z, d.w};
  const uint32_t aw[2] = {arg.x, arg.y};
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t a2 = (aw[q >> 1] >> ((q & 1) * 16)) & 0xFFFFu;
    const bool lo = (a2 & 255u) == (uint32_t)win, hi = (a2 >> 8) == (uint32_t)win;
    dw[q] = (lo ? (dw[q] & 0xFFFFu) : 0u) | (hi ? (dw[q] & 0xFFFF0000u) : 0u);
  }
  return make_uint4(dw[0], dw[1], dw[2], dw[3]);
}
__device__ __forceinline__ void add_chunk(float* cs, const uint4& d) {
  cs[0] += bf16_lo(d.x); cs[1] += bf16_hi(d.x); cs[2] += bf16_lo(d.y); cs[3] += bf16_hi(d.y);
  cs[4] += bf16_lo(d.z); cs[5] += bf16_hi(d.z); cs[6] += bf16_lo(d.w); cs[7] += bf16_hi(d.w);
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
// dynamic shared memory is re-based to a 1024-byte boundary inside the kernels (swizzled tiles need it)
constexpr uint32_t SMEM_SLACK = 1024;
__host__ __device__ constexpr uint32_t round1k_c(uint32_t v) { return (v + 1023u) & ~1023u; }
__device__ __forceinline__ uint8_t* align_smem(uint8_t* p) {
  const uint32_t a = smem_u32(p);
  return p + (((a + 1023u) & ~1023u) - a);
}

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
  int npass;            // umma_fwd_res_u8_kernel only: passes folded into blockIdx.x ...
  int tiles_per_cta;    // ... and consecutive tiles per CTA
};

// kernel parameters are selected field by field: indexing a.p[] with a run-time value makes the compiler copy the whole
// parameter block to local memory (112 bytes of stack, read back through L1 at the start of every thread)
__device__ __forceinline__ FwdPassT pick_pass(const FwdArgsT& a, int pass) {
  FwdPassT P;
  P.x = pass ? a.p[1].x : a.p[0].x;
  P.wt = pass ? a.p[1].wt : a.p[0].wt;
  P.bias = pass ? a.p[1].bias : a.p[0].bias;
  P.out = pass ? a.p[1].out : a.p[0].out;
  P.out_f32 = pass ? a.p[1].out_f32 : a.p[0].out_f32;
  P.arg = pass ? a.p[1].arg : a.p[0].arg;
  return P;
}

// Forward epilogue: TMEM -> bias + ReLU (+ 2x2 max-pool and arg-max) -> bf16 NHWC (+ optional fp32 copy).
template <class G, int NTILE>
__device__ __forceinline__ void fwd_epilogue(uint32_t tmem, const FwdPassT& P, int m0, int n0, int B, const float* bias_sm) {
  const int tid = threadIdx.x;
  const int arow = tid & (TM - 1);
  const int warp = tid >> 5, lane = tid & 31;
  const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);    // tile row = arow (same for warps w and w+4k)
  const Row row = decode_row<G>(m0 + arow, B);
  const bool row_in = row.valid;
#pragma unroll 1
  for (int c0 = (warp >> 2) * 32; c0 < NTILE; c0 += 128) {             // the 4 warps of a lane group split the columns
    float v[32];
    tmem_ld32(trow + c0, v);
    const int n = n0 + c0;
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + bias_sm[c0 + j], 0.f);
    if constexpr (G::POOL) {
      // Lanes 4q..4q+3 hold the 4 conv outputs of one pool window.  Values are >= 0 after ReLU, so their float bits
      // order like unsigned integers: pack (value with the 2 lowest mantissa bits replaced by 3-win) into one key and
      // take the maximum over the window -- largest value wins, ties go to the first window element, rows outside
      // the image (key 0) never beat a positive value.  The 2 dropped bits are far below bf16 resolution.
      uint32_t key[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        uint32_t kj = row_in ? ((__float_as_uint(v[j]) & ~3u) | (3u - (uint32_t)(lane & 3))) : 0u;
        kj = max(kj, __shfl_xor_sync(0xffffffffu, kj, 1));
        kj = max(kj, __shfl_xor_sync(0xffffffffu, kj, 2));
        key[j] = kj;
      }
      if ((lane & 3) == 0 && row.b < B) {
        const size_t off = (size_t)row.pidx * G::COUT + n;
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j)
          pk[j] = pack_bf16(__uint_as_float(key[2 * j] & ~3u), __uint_as_float(key[2 * j + 1] & ~3u));
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(P.out + off + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        if (P.arg) {
          uint32_t ak[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            ak[j] = (3u - (key[4 * j] & 3u)) | ((3u - (key[4 * j + 1] & 3u)) << 8) | ((3u - (key[4 * j + 2] & 3u)) << 16) |
                    ((3u - (key[4 * j + 3] & 3u)) << 24);
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
}

constexpr uint32_t FWD_RAW_PITCH = 80;                       // 64 uint8 of one tile row + pad (keeps 16-byte alignment)
constexpr uint32_t FWD_RAW_A = 128 * FWD_RAW_PITCH;          // 10 KB
// RSN ring slots, loads run PD stages ahead of the MMAs (PD <= RSN)
template <typename TIn, int NTILE, int RSN>
__host__ __device__ constexpr uint32_t fwd_smem_bytes() {
  return SMEM_SLACK + (sizeof(TIn) == 1 ? RSN * FWD_RAW_A + 2u * A_STAGE_BYTES : RSN * A_STAGE_BYTES) + RSN * (uint32_t)NTILE * 128u;
}

template <class G, typename TIn, int NTILE, int RSN, int PD>
__global__ void __launch_bounds__(TT) umma_fwd_kernel(FwdArgsT a) {
  static_assert(PD >= 1 && PD <= RSN, "prefetch distance");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  constexpr bool U8 = sizeof(TIn) == 1;
  constexpr uint32_t B_STAGE = NTILE * 128;
  constexpr uint32_t OFF_OPA = U8 ? RSN * FWD_RAW_A : 0;             // converted A tiles (uint8 input only), 2 slots
  constexpr uint32_t OFF_B = U8 ? RSN * FWD_RAW_A + 2 * A_STAGE_BYTES : RSN * A_STAGE_BYTES;
  __shared__ uint64_t bars[2];        // set-up only (TMEM allocation helper)
  __shared__ uint64_t done[RSN];      // the MMAs reading a ring slot have completed (tcgen05.commit)
  __shared__ uint64_t full[RSN];      // a ring slot is filled: one arrival per warp
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[NTILE];
  const FwdPassT P = pick_pass(a, (int)blockIdx.z);
  const TIn* __restrict__ x = static_cast<const TIn*>(P.x);
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * TM, n0 = blockIdx.y * NTILE;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < RSN; ++i) { mbar_init(&full[i], TT / 32 + 1); mbar_init(&done[i], 1); }   // 16 warps + the weight copy
  }
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sbase = smem_u32(smem);
  // gather mapping (lanes along the contiguous k direction): chunk id = tid + 512*q -> row id>>3, chunk id&7
  const Row grow0 = decode_row<G>(m0 + (tid >> 3), a.B), grow1 = decode_row<G>(m0 + 64 + (tid >> 3), a.B);
  const int gc = tid & 7;
  pdl_wait_then_trigger();                                           // everything above overlapped the previous kernel
  for (int j = tid; j < NTILE; j += TT) bias_sm[j] = __ldg(P.bias + n0 + j);
  // epilogue mapping: tile row tid&127
  const int arow = tid & (TM - 1);
  constexpr int T = G::K / TK;
  static_assert(G::K % TK == 0, "the reduction must be a whole number of stages");
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  // Address generation is hoisted out of the stage loop (the gather is instruction-issue bound otherwise): per thread
  // and row a base pointer at the window origin, per stage one (ky, kx, channel) offset shared by both rows.
  constexpr int EPC = 8;                               // elements per chunk (8 bf16, or 2 pixels x 4 uint8)
  const TIn* abase[2];
  int aiy0[2], aix0[2];
  bool avalid[2];
  uint32_t adst[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const Row& r = q ? grow1 : grow0;
    aiy0[q] = r.oy * G::ST - G::PADH;
    aix0[q] = r.ox * G::ST - G::PADW;
    avalid[q] = r.valid;
    abase[q] = x + ((ptrdiff_t)(r.b * G::IH + aiy0[q]) * G::IW + aix0[q]) * G::CIN;     // may point outside: only used when in bounds
    const int rr = (tid >> 3) + 64 * q;
    adst[q] = U8 ? (uint32_t)(rr * FWD_RAW_PITCH + gc * 8) : sw128_off(rr, gc);
  }
  const __nv_bfloat16* wimg = P.wt + (size_t)n0 * 64;       // stage image: COUT*64 elements per stage, rows n0.. contiguous
  auto issue = [&](int it) {
    if (it < T) {
      const int slot = it % RSN;
      // chunk (it*8 + gc) of the (ky, kx, c) reduction axis: which filter tap, which channel group
      const int e = (it * 8 + gc) * EPC;               // first reduction element of this chunk
      const int tap = e / G::CIN, c0 = e - tap * G::CIN;
      const int ky = tap / G::KS, kx = tap - ky * G::KS;
      const int delta = (ky * G::IW + kx) * G::CIN + c0;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const bool ok = avalid[q] && (unsigned)(aiy0[q] + ky) < (unsigned)G::IH && (unsigned)(aix0[q] + kx) < (unsigned)G::IW;
        const TIn* src = ok ? abase[q] + delta : x;
        if constexpr (U8) cp_async8(sbase + slot * FWD_RAW_A + adst[q], src, ok);
        else cp_async16(sbase + slot * A_STAGE_BYTES + adst[q], src, ok);
      }
      if (tid == 0) {                                    // the whole weight tile of the stage: one bulk copy
        mbar_expect_tx(&full[slot], B_STAGE);
        bulk_g2s(sbase + OFF_B + slot * B_STAGE, wimg + (size_t)it * (G::COUT * 64), B_STAGE, &full[slot]);
      }
    }
    cp_async_commit();
  };
#pragma unroll
  for (int s = 0; s < PD; ++s) issue(s);
  const int wu = warp_uniform_idx();
#pragma unroll 1
  for (int it = 0; it < T; ++it) {
    const int slot = it % RSN, par = it & 1;
    // stage it+PD goes into the slot stage it+PD-RSN used: the MMAs that read it must be done
    if (it + PD < T && it + PD >= RSN) mbar_wait(&done[(it + PD) % RSN], ((it + PD - RSN) / RSN) & 1);
    issue(it + PD);
    cp_async_wait<PD>();                                        // this thread's copies for stage `it` have landed
    uint32_t a_addr;
    if constexpr (U8) {
      // widen the two 8-byte chunks this thread fetched itself and store them swizzled (lanes run along the chunks
      // of a row, which the 128-byte swizzle keeps bank-conflict free); the tile being overwritten was read by the
      // MMAs of stage it-2
      if (it >= 2) mbar_wait(&done[(it - 2) % RSN], ((it - 2) / RSN) & 1);
      uint8_t* opa = smem + OFF_OPA + par * A_STAGE_BYTES;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int r = (tid >> 3) + 64 * q;
        const uint2 u = *reinterpret_cast<const uint2*>(smem + slot * FWD_RAW_A + r * FWD_RAW_PITCH + gc * 8);
        *reinterpret_cast<uint4*>(opa + sw128_off(r, gc)) = cvt_u8x8(u);
      }
      a_addr = sbase + OFF_OPA + par * A_STAGE_BYTES;
    } else {
      a_addr = sbase + slot * A_STAGE_BYTES;
    }
    // No CTA-wide barrier: every warp publishes "my part of this slot is written and visible to the tensor core"
    // on the slot's mbarrier and moves on to the next stage; only the MMA-issuing thread waits for all 16 warps.
    fence_async_proxy();
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(&full[slot]);
    if (wu == 0) {
      mbar_wait(&full[slot], (it / RSN) & 1);
      issue_stage<true, true>(tmem, a_addr, sbase + OFF_B + slot * B_STAGE, idesc, it == 0, &done[slot]);
    }
  }
  pdl_trigger();                                                // all loads issued: let the next kernel start its prologue
  // the last commit covers every MMA issued before it
  mbar_wait(&done[(T - 1) % RSN], ((T - 1) / RSN) & 1);
  tc_fence_after();
  fwd_epilogue<G, NTILE>(tmem, P, m0, n0, a.B, bias_sm);
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ forward, resident input
// bf16 conv layers whose tiles span only a few samples.  Two observations from cycle traces of the ring kernels:
// (1) the im2col gather from L2 re-reads every input pixel once per filter tap that touches it (conv2: 4x, conv3: 9x);
// (2) with N = 64 a 64-deep stage is only 128 cycles of tensor-core work, but every stage costs ~1000 cycles of fixed
//     latencies on the issuing warp (mbarrier try_waits, proxy fence, MMA issue + commit).
// Here the input feature maps of the tile's samples -- one contiguous NHWC block -- are fetched ONCE with a bulk async
// copy, ALL weight stages are fetched up front with bulk copies (before the PDL wait: they do not depend on the
// previous kernel), and the A operand is built shared->shared (one 16-byte load + store per chunk) in BIG stages of
// NK k-blocks (NK*64 reduction elements): one barrier round and one commit per 4*NK MMAs.
template <class G, int NK, int NA>
struct FwdRes {
  static constexpr int NS = (TM + G::ROWS - 2) / G::ROWS + 1;                       // samples a 128-row tile can touch
  static constexpr uint32_t SAMPLE_BYTES = G::IH * G::IW * G::CIN * 2;
  static constexpr uint32_t PATCH_BYTES = round1k_c(NS * SAMPLE_BYTES);
  static constexpr int T = G::K / TK;                                               // k-blocks
  static constexpr int NBS = T / NK;                                                // big stages
  static constexpr uint32_t A_SLOT = NK * A_STAGE_BYTES;
  static_assert(G::K % (TK * NK) == 0, "the reduction must be a whole number of big stages");
  static_assert(SAMPLE_BYTES % 16 == 0, "bulk copies move multiples of 16 bytes");
};
template <class G, int NTILE, int NK, int NA>
__host__ __device__ constexpr uint32_t fwd_res_smem_bytes() {
  using R = FwdRes<G, NK, NA>;
  return SMEM_SLACK + R::PATCH_BYTES + NA * R::A_SLOT + R::T * (uint32_t)NTILE * 128u;
}

template <class G, int NTILE, int NK, int NA>
__global__ void __launch_bounds__(TT) umma_fwd_res_kernel(FwdArgsT a) {
  using R = FwdRes<G, NK, NA>;
  constexpr int T = R::T, NBS = R::NBS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  constexpr uint32_t B_STAGE = NTILE * 128;
  constexpr uint32_t OFF_A = R::PATCH_BYTES, OFF_B = OFF_A + NA * R::A_SLOT;
  __shared__ uint64_t bars[2];
  __shared__ uint64_t afull[NA];      // an operand slot is built: one arrival per warp
  __shared__ uint64_t wfull[NBS];     // the weights of a big stage have landed (used once each)
  __shared__ uint64_t done[NBS];      // the MMAs of a big stage have completed (used once each)
  __shared__ uint64_t patchbar;       // the input feature maps have landed
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[NTILE];
  const FwdPassT P = pick_pass(a, (int)blockIdx.z);
  const __nv_bfloat16* __restrict__ x = static_cast<const __nv_bfloat16*>(P.x);
  const int tid = threadIdx.x, lane = tid & 31;
  const int m0 = blockIdx.x * TM, n0 = blockIdx.y * NTILE;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NA; ++i) mbar_init(&afull[i], TT / 32);
#pragma unroll
    for (int i = 0; i < NBS; ++i) { mbar_init(&wfull[i], 1); mbar_init(&done[i], 1); }
    mbar_init(&patchbar, 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sbase = smem_u32(smem);
  // weights were written by the previous step's optimizer, several launches back: fetch them before the PDL wait.
  // The NK k-block images of a big stage are contiguous in the stage-image layout when the tile spans all COUT rows.
  if (tid == 0) {
    const __nv_bfloat16* wimg = P.wt + (size_t)n0 * 64;
#pragma unroll
    for (int bs = 0; bs < NBS; ++bs) {
      mbar_expect_tx(&wfull[bs], NK * B_STAGE);
#pragma unroll
      for (int kb = 0; kb < NK; ++kb)
        bulk_g2s(sbase + OFF_B + (bs * NK + kb) * B_STAGE, wimg + (size_t)(bs * NK + kb) * (G::COUT * 64), B_STAGE, &wfull[bs]);
    }
  }
  for (int j = tid; j < NTILE; j += TT) bias_sm[j] = __ldg(P.bias + n0 + j);
  // rows of this thread in the build mapping: tile rows (tid>>3) and (tid>>3)+64, chunk gc of each k-block
  const int gc = tid & 7;
  const int b0 = m0 / G::ROWS;
  int b1 = (m0 + TM - 1) / G::ROWS;
  if (b1 > a.B - 1) b1 = a.B - 1;
  int poff[2], aiy0[2], aix0[2];
  bool avalid[2];
  uint32_t adst[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int rr = (tid >> 3) + 64 * q;
    const Row r = decode_row<G>(m0 + rr, a.B);
    aiy0[q] = r.oy * G::ST - G::PADH;
    aix0[q] = r.ox * G::ST - G::PADW;
    avalid[q] = r.valid;
    poff[q] = (((r.b - b0) * G::IH + aiy0[q]) * G::IW + aix0[q]) * G::CIN * 2;   // byte offset in the patch (may be < 0 at the border)
    adst[q] = OFF_A + sw128_off(rr, gc);
  }
  pdl_wait_then_trigger();
  if (tid == 0) {
    const uint32_t pbytes = (uint32_t)(b1 - b0 + 1) * R::SAMPLE_BYTES;
    mbar_expect_tx(&patchbar, pbytes);
    bulk_g2s(sbase, x + (size_t)b0 * (R::SAMPLE_BYTES / 2), pbytes, &patchbar);
  }
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  const int wu = warp_uniform_idx();
  mbar_wait(&patchbar, 0);
#pragma unroll 1
  for (int bs = 0; bs < NBS; ++bs) {
    const int as = bs % NA;
    if (bs >= NA) mbar_wait(&done[bs - NA], 0);        // the MMAs that read this operand slot are done
#pragma unroll
    for (int kb = 0; kb < NK; ++kb) {
      const int e = ((bs * NK + kb) * 8 + gc) * 8;     // first reduction element of this thread's chunk
      const int tap = e / G::CIN, c0 = e - tap * G::CIN;
      const int ky = tap / G::KS, kx = tap - ky * G::KS;
      const int delta = ((ky * G::IW + kx) * G::CIN + c0) * 2;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const bool ok = avalid[q] && (unsigned)(aiy0[q] + ky) < (unsigned)G::IH && (unsigned)(aix0[q] + kx) < (unsigned)G::IW;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (ok) v = *reinterpret_cast<const uint4*>(smem + poff[q] + delta);
        *reinterpret_cast<uint4*>(smem + as * R::A_SLOT + kb * A_STAGE_BYTES + adst[q]) = v;
      }
    }
    fence_async_proxy();
    __syncwarp();
    if (lane == 0) mbar_arrive(&afull[as]);
    if (wu == 0) {
      mbar_wait(&afull[as], (bs / NA) & 1);
      mbar_wait(&wfull[bs], 0);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t ad0 = make_desc_sw128(sbase + OFF_A + as * R::A_SLOT), bd0 = make_desc_sw128(sbase + OFF_B + bs * NK * B_STAGE);
#pragma unroll
        for (int kb = 0; kb < NK; ++kb) {
#pragma unroll
          for (int j = 0; j < TK / 16; ++j)
            mma_bf16(tmem, desc_advance(ad0, kb * A_STAGE_BYTES + j * 32), desc_advance(bd0, kb * B_STAGE + j * 32), idesc,
                     (bs | kb | j) ? 1u : 0u);
        }
        mma_commit(&done[bs]);
      }
      __syncwarp();
    }
  }
  pdl_trigger();
  mbar_wait(&done[NBS - 1], 0);
  tc_fence_after();
  fwd_epilogue<G, NTILE>(tmem, P, m0, n0, a.B, bias_sm);
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ forward, resident uint8 frames
// conv1 on the uint8 frame stacks (network.py:395-407 with the float cast of the feed, dqn_agent.py:186-189, fused in).
// A 128-row tile is 32 pool windows = at most 5 pool rows, i.e. a contiguous band of at most PATCH_ROWS input rows
// (336 bytes each) -- fetched once by one bulk copy and widened to bf16 ONCE (every pixel is used by 4 windows; the
// integer->float conversion instruction runs at 1/8 rate, so bytes are widened with a byte-permute into the mantissa of
// 2^23 and one subtraction instead).  The whole reduction (K = 8*8*4 = 256) is ONE big stage: every thread copies 8
// chunks of 2 pixels x 4 frames from the bf16 band into the swizzled operand tile, then 16 MMAs, then the pooled
// epilogue.  The uint8 band is staged in the tail of the operand area, which is only written after the widening.
template <class G>
struct FwdResU8 {
  static constexpr int PATCH_ROWS = 40;                                               // 5 pool rows: 8*5 + 4 rows minus the clipped border
  static constexpr uint32_t ROW_BYTES = G::IW * G::CIN;                               // 336
  static constexpr uint32_t RAW_BYTES = round1k_c(PATCH_ROWS * ROW_BYTES);            // uint8 band (staged inside the operand area)
  static constexpr uint32_t PATCH_BYTES = round1k_c(2 * PATCH_ROWS * ROW_BYTES);      // bf16 band
  static constexpr int T = G::K / TK;                                                 // 4 k-blocks
  static_assert(G::CIN == 4 && ROW_BYTES % 16 == 0 && TK == 2 * G::KS * G::CIN, "one k-block = two filter rows of 8 pixels x 4 frames");
};
// input-row band of the 128-row tile starting at row m0 (see umma_fwd_res_u8_kernel): global row indices b*IH + iy, inclusive
template <class G>
__device__ __forceinline__ void conv1_band(int m0, int B, int& grow_lo, int& grow_hi, int& w_first, int& w_last) {
  constexpr int WPS = G::PH * G::PW;
  w_first = m0 >> 2;
  w_last = (m0 + TM - 1) >> 2;
  if (w_last > B * WPS - 1) w_last = B * WPS - 1;
  const int bf = w_first / WPS, bl = w_last / WPS;
  const int pyf = (w_first - bf * WPS) / G::PW, pyl = (w_last - bl * WPS) / G::PW;
  int iy_lo = 2 * pyf * G::ST - G::PADH;
  if (iy_lo < 0) iy_lo = 0;
  int iy_hi = (2 * pyl + 1) * G::ST - G::PADH + G::KS - 1;
  if (iy_hi > G::IH - 1) iy_hi = G::IH - 1;
  grow_lo = bf * G::IH + iy_lo;
  grow_hi = bl * G::IH + iy_hi;
}

template <class G, int NTILE>
__host__ __device__ constexpr uint32_t fwd_res_u8_smem_bytes() {
  return SMEM_SLACK + FwdResU8<G>::PATCH_BYTES + FwdResU8<G>::T * (A_STAGE_BYTES + (uint32_t)NTILE * 128u);
}

template <class G, int NTILE>
__global__ void __launch_bounds__(TT, 2) umma_fwd_res_u8_kernel(FwdArgsT a) {      // 2 CTAs/SM: at most 64 registers
  using R = FwdResU8<G>;
  constexpr int T = R::T;
  static_assert(G::COUT == NTILE, "one tile spans all output channels");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  constexpr uint32_t B_STAGE = NTILE * 128;
  constexpr uint32_t OFF_A = R::PATCH_BYTES, OFF_B = OFF_A + T * A_STAGE_BYTES;
  constexpr uint32_t OFF_RAW = OFF_B - R::RAW_BYTES;                   // tail of the operand area (last k-block's tile)
  static_assert(R::RAW_BYTES <= A_STAGE_BYTES, "the uint8 band must fit in the last operand tile");
  __shared__ uint64_t bars[2];
  __shared__ uint64_t afull, wfull, done, patchbar;
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[NTILE];
  const int tid = threadIdx.x, lane = tid & 31;
  // A CTA walks over `tiles_per_cta` consecutive tiles of the combined (pass, 128-row tile) space: set-up, TMEM
  // allocation and (per pass) the weights are paid once, and the next tile's band is fetched under the epilogue.
  const int mtiles = (a.B * G::ROWS + TM - 1) / TM;
  const int total = mtiles * a.npass;
  const int g_begin = blockIdx.x * a.tiles_per_cta;
  const int g_end = min(total, g_begin + a.tiles_per_cta);
  if (tid == 0) {
    mbar_init(&afull, TT / 32);
    mbar_init(&wfull, 1);
    mbar_init(&done, 1);
    mbar_init(&patchbar, 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sbase = smem_u32(smem);
  // kernel parameters are selected field by field (indexing a.p[] with a run-time value would copy it to local memory)
  auto pass_of = [&](int pass) {
    FwdPassT P;
    P.x = pass ? a.p[1].x : a.p[0].x;
    P.wt = pass ? a.p[1].wt : a.p[0].wt;
    P.bias = pass ? a.p[1].bias : a.p[0].bias;
    P.out = pass ? a.p[1].out : a.p[0].out;
    P.out_f32 = nullptr;
    P.arg = pass ? a.p[1].arg : a.p[0].arg;
    return P;
  };
  auto load_weights = [&](int pass) {    // thread 0: all k-blocks of the pass's weight image
    const __nv_bfloat16* wimg = pass ? a.p[1].wt : a.p[0].wt;
    mbar_expect_tx(&wfull, T * B_STAGE);
#pragma unroll
    for (int kb = 0; kb < T; ++kb) bulk_g2s(sbase + OFF_B + kb * B_STAGE, wimg + (size_t)kb * (G::COUT * 64), B_STAGE, &wfull);
  };
  // per-tile state, prepared one tile ahead (under the previous tile's MMAs, or before the PDL wait for the first tile)
  int glo = 0, ghi = 0, m0 = 0, pass = 0;
  int poff[2], aiy0[2], aix0[2];
  bool avalid[2];
  const int gc = tid & 7;
  auto prep = [&](int g) {
    pass = g / mtiles;
    m0 = (g - pass * mtiles) * TM;
    int wf, wl;
    conv1_band<G>(m0, a.B, glo, ghi, wf, wl);
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const Row r = decode_row<G>(m0 + (tid >> 3) + 64 * q, a.B);
      aiy0[q] = r.oy * G::ST - G::PADH;
      aix0[q] = r.ox * G::ST - G::PADW;
      avalid[q] = r.valid;
      poff[q] = ((r.b * G::IH + aiy0[q] - glo) * G::IW + aix0[q]) * G::CIN;      // byte offset in the band (may be < 0 at the border)
    }
  };
  auto load_band = [&]() {               // thread 0: the input-row band of the prepared tile
    const int nrows = ghi - glo + 1;
    if (nrows > R::PATCH_ROWS) __trap();
    mbar_expect_tx(&patchbar, (uint32_t)nrows * R::ROW_BYTES);
    bulk_g2s(sbase + OFF_RAW, static_cast<const uint8_t*>(pass ? a.p[1].x : a.p[0].x) + (size_t)glo * R::ROW_BYTES,
             (uint32_t)nrows * R::ROW_BYTES, &patchbar);
  };
  int cur_pass = g_begin < g_end ? g_begin / mtiles : 0;
  int wphase = 0;
  if (g_begin < g_end) prep(g_begin);
  const uint32_t adst[2] = {OFF_A + sw128_off(tid >> 3, gc), OFF_A + sw128_off((tid >> 3) + 64, gc)};
  pdl_wait_then_trigger();
  // conv1 can be the first kernel after the previous update's optimizer (no gather in between when the caller passes
  // stacks directly), so unlike the deeper layers it reads weights and bias only after the wait; they arrive together
  // with the band
  if (tid == 0 && g_begin < g_end) {
    load_weights(cur_pass);
    load_band();
  }
  for (int j = tid; j < NTILE; j += TT) bias_sm[j] = __ldg((cur_pass ? a.p[1].bias : a.p[0].bias) + j);
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  const int kxo = 2 * (gc & 3), kyo = gc >> 2;                         // chunk gc of k-block kb: filter row 2*kb + gc/4, pixels 2*(gc%4), +1
  const int wu = warp_uniform_idx();
#pragma unroll 1
  for (int g = g_begin; g < g_end; ++g) {
    const int t = g - g_begin;
    mbar_wait(&patchbar, t & 1);
    {   // widen the band: 8 bytes -> 8 bf16 per step (all readers of the previous tile's band are past its MMAs)
      const int nchunks = (ghi - glo + 1) * (int)(R::ROW_BYTES / 8);
      for (int c = tid; c < nchunks; c += TT)
        *reinterpret_cast<uint4*>(smem + c * 16) = widen_u8x8(*reinterpret_cast<const uint2*>(smem + OFF_RAW + c * 8));
    }
    __syncthreads();
#pragma unroll
    for (int kb = 0; kb < T; ++kb) {
      const int ky = 2 * kb + kyo;
      const int delta = (ky * G::IW + kxo) * G::CIN;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const bool ok = avalid[q] && (unsigned)(aiy0[q] + ky) < (unsigned)G::IH && (unsigned)(aix0[q] + kxo) < (unsigned)G::IW;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (ok) v = *reinterpret_cast<const uint4*>(smem + 2 * (poff[q] + delta));
        *reinterpret_cast<uint4*>(smem + kb * A_STAGE_BYTES + adst[q]) = v;
      }
    }
    fence_async_proxy();
    __syncwarp();
    if (lane == 0) mbar_arrive(&afull);
    if (wu == 0) {
      mbar_wait(&afull, t & 1);           // every warp has built its part (and finished the previous tile's epilogue)
      mbar_wait(&wfull, wphase);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t ad0 = make_desc_sw128(sbase + OFF_A), bd0 = make_desc_sw128(sbase + OFF_B);
#pragma unroll
        for (int kb = 0; kb < T; ++kb) {
#pragma unroll
          for (int j = 0; j < TK / 16; ++j)
            mma_bf16(tmem, desc_advance(ad0, kb * A_STAGE_BYTES + j * 32), desc_advance(bd0, kb * B_STAGE + j * 32), idesc,
                     (kb | j) ? 1u : 0u);
        }
        mma_commit(&done);
      }
      __syncwarp();
    }
    if (g + 1 == g_end) pdl_trigger();
    // this tile's output coordinates; then prepare the next tile while the MMAs run
    const int e_pass = pass, e_m0 = m0;
    if (g + 1 < g_end) prep(g + 1);
    mbar_wait(&done, t & 1);
    tc_fence_after();
    // operands are free again: fetch the next tile's band (and, at the online/target boundary, the other weights)
    if (tid == 0 && g + 1 < g_end) {
      load_band();
      if (pass != e_pass) load_weights(pass);
    }
    fwd_epilogue<G, NTILE>(tmem, pass_of(e_pass), e_m0, 0, a.B, bias_sm);
    tc_fence_before();
    if (pass != e_pass) {                 // uniform over the CTA
      wphase ^= 1;
      __syncthreads();                    // everyone is done with the old bias
      for (int j = tid; j < NTILE; j += TT) bias_sm[j] = __ldg((pass ? a.p[1].bias : a.p[0].bias) + j);
      __syncthreads();
    }
  }
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ weight gradient
// dactg is the gradient w.r.t. the layer's pooled output, ALREADY multiplied by the ReLU gate (act > 0) by the
// kernel that produced it (td_head / umma_dgrad); only the pool routing (arg-max == window position) remains.
// Both operands are MN-major (the reduction runs over output pixels), so both go through a raw ring and are
// transposed into the no-swizzle MN-major layout shared->shared.
struct WgradArgsT {
  const void* x;                  // layer input: uint8 stacks (conv1) or bf16 activations
  const __nv_bfloat16* dactg;
  const uint8_t* arg;             // pool arg-max (POOL layers)
  float* partial;                 // [nsplit][(K+1)*COUT] fp32
  int B;
  int rows_per_split;             // multiple of TK
};

template <typename TIn>
__host__ __device__ constexpr uint32_t wg_a_pitch() { return sizeof(TIn) == 1 ? 144u : 272u; }   // 16 chunks of 8 / 16 bytes + pad
template <int NTILE>
__host__ __device__ constexpr uint32_t wg_bd_pitch() { return (uint32_t)NTILE * 2u + 16u; }       // gradient row + pad
template <int NTILE>
__host__ __device__ constexpr uint32_t wg_ba_pitch() { return (uint32_t)NTILE + 8u; }             // arg-max row + pad
__host__ __device__ constexpr uint32_t round1k(uint32_t v) { return (v + 1023u) & ~1023u; }
template <class G, typename TIn, int NTILE>
__host__ __device__ constexpr uint32_t wgrad_smem_bytes() {
  return SMEM_SLACK + RS * round1k(64u * wg_a_pitch<TIn>()) + RS * round1k(64u * wg_bd_pitch<NTILE>()) +
         (G::POOL ? RS * round1k(64u * wg_ba_pitch<NTILE>()) : 0u) + 2u * A_STAGE_BYTES + 2u * (uint32_t)NTILE * 128u;
}

template <class G, typename TIn, int NTILE>
__global__ void __launch_bounds__(TT) umma_wgrad_kernel(WgradArgsT a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  constexpr bool U8 = sizeof(TIn) == 1;
  constexpr uint32_t B_STAGE = NTILE * 128;
  constexpr uint32_t PA = wg_a_pitch<TIn>(), PBD = wg_bd_pitch<NTILE>(), PBA = wg_ba_pitch<NTILE>();
  constexpr uint32_t RAW_A = round1k(64 * PA), RAW_BD = round1k(64 * PBD), RAW_BA = G::POOL ? round1k(64 * PBA) : 0;
  constexpr uint32_t OFF_RBD = RS * RAW_A, OFF_RBA = OFF_RBD + RS * RAW_BD, OFF_OPA = OFF_RBA + RS * RAW_BA;
  constexpr uint32_t OFF_OPB = OFF_OPA + 2 * A_STAGE_BYTES;
  constexpr int ACB = U8 ? 8 : 16;                                   // bytes of one A chunk in the raw ring
  __shared__ uint64_t bars[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[NTILE];
  const TIn* __restrict__ x = static_cast<const TIn*>(a.x);
  const int tid = threadIdx.x;
  const int i0 = blockIdx.x * TM, n0 = blockIdx.y * NTILE;     // i = (ky,kx,c) index, rows of the output tile
  for (int j = tid; j < NTILE; j += TT) bias_sm[j] = 0.f;
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sbase = smem_u32(smem);
  pdl_wait_then_trigger();
  const int R = a.B * G::ROWS;
  const int r_begin = blockIdx.z * a.rows_per_split;
  const int r_end = min(R, r_begin + a.rows_per_split);
  const int T = (r_end - r_begin + TK - 1) / TK;
  const bool do_bias = blockIdx.x == 0;
  constexpr uint32_t idesc = make_idesc(TM, NTILE, true, true);
  // gather mapping (lanes along contiguous memory): A chunk id = tid + 512*q -> reduction row id>>4, k-chunk id&15;
  //                                                  B chunk id = tid + 512*q -> reduction row id/GPR, n-chunk id%GPR
  // transpose mapping (lanes along the reduction row, conflict-free stores): row tid&63, chunks (tid>>6)*CPT ..
  constexpr int GPR = NTILE / 8;
  constexpr int NCHB = 64 * GPR;
  constexpr int BQ = (NCHB + TT - 1) / TT;
  constexpr int CPT = GPR / 8 > 0 ? GPR / 8 : 1;                     // B chunks per transposing thread (GPR=4: half the threads idle)
  const int tr = tid & 63, tg = tid >> 6;
  float colsum[CPT][8];
#pragma unroll
  for (int q = 0; q < CPT; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e) colsum[q][e] = 0.f;
  auto issue = [&](int it) {
    if (it < T) {
      const int slot = it & (RS - 1);
      const int rbase = r_begin + it * TK;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int id = tid + TT * q, rr = id >> 4, g = id & 15;
        const Row ar = decode_row<G>(rbase + rr, a.B);
        const TIn* src = (rbase + rr) < r_end ? x_chunk_ptr<G, TIn>(x, ar, (i0 >> 3) + g) : nullptr;
        const uint32_t dst = sbase + slot * RAW_A + rr * PA + g * ACB;
        if constexpr (U8) cp_async8(dst, src ? src : x, src != nullptr);
        else cp_async16(dst, src ? src : x, src != nullptr);
      }
#pragma unroll
      for (int q = 0; q < BQ; ++q) {
        const int id = tid + TT * q, rr = id / GPR, g = id % GPR;
        if (id >= NCHB) break;
        const Row br = decode_row<G>(rbase + rr, a.B);
        const bool ok = br.valid && (rbase + rr) < r_end;
        const size_t off = (size_t)(ok ? br.pidx : 0) * G::COUT + n0 + g * 8;
        cp_async16(sbase + OFF_RBD + slot * RAW_BD + rr * PBD + g * 16, a.dactg + off, ok);
        if constexpr (G::POOL) cp_async8(sbase + OFF_RBA + slot * RAW_BA + rr * PBA + g * 8, a.arg + off, ok);
      }
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  const int wu = warp_uniform_idx();
  for (int it = 0; it < T; ++it) {
    const int slot = it & (RS - 1), par = it & 1;
    if (it >= 2) mbar_wait(&bars[par], ((it >> 1) - 1) & 1);
    issue(it + 2);
    cp_async_wait<2>();
    __syncthreads();
    uint8_t* opa = smem + OFF_OPA + par * A_STAGE_BYTES;
    uint8_t* opb = smem + OFF_OPB + par * B_STAGE;
    {   // A: k-chunks 2*tg, 2*tg+1 of reduction row tr
      const uint8_t* src = smem + slot * RAW_A + tr * PA + (2 * tg) * ACB;
      if constexpr (U8) {
        const uint4 u = *reinterpret_cast<const uint4*>(src);
        *reinterpret_cast<uint4*>(opa + mnmajor_off(2 * tg, tr)) = cvt_u8x8(make_uint2(u.x, u.y));
        *reinterpret_cast<uint4*>(opa + mnmajor_off(2 * tg + 1, tr)) = cvt_u8x8(make_uint2(u.z, u.w));
      } else {
        *reinterpret_cast<uint4*>(opa + mnmajor_off(2 * tg, tr)) = *reinterpret_cast<const uint4*>(src);
        *reinterpret_cast<uint4*>(opa + mnmajor_off(2 * tg + 1, tr)) = *reinterpret_cast<const uint4*>(src + 16);
      }
    }
    if (tg * CPT < GPR) {   // B: n-chunks tg*CPT .. of reduction row tr
      const int win = (r_begin + it * TK + tr) & 3;                     // pool-window position of this row
#pragma unroll
      for (int q = 0; q < CPT; ++q) {
        const int g = tg * CPT + q;
        uint4 z = *reinterpret_cast<const uint4*>(smem + OFF_RBD + slot * RAW_BD + tr * PBD + g * 16);
        if constexpr (G::POOL)
          z = route_chunk(z, *reinterpret_cast<const uint2*>(smem + OFF_RBA + slot * RAW_BA + tr * PBA + g * 8), win);
        *reinterpret_cast<uint4*>(opb + mnmajor_off(g, tr)) = z;
        if (do_bias) add_chunk(colsum[q], z);
      }
    }
    fence_async_proxy();
    __syncthreads();
    if (wu == 0)
      issue_stage<false, false>(tmem, sbase + OFF_OPA + par * A_STAGE_BYTES, sbase + OFF_OPB + par * B_STAGE, idesc, it == 0,
                                &bars[par]);
  }
  pdl_trigger();
  float* out = a.partial + (size_t)blockIdx.z * ((size_t)(G::K + 1) * G::COUT);
  const int warp = tid >> 5, lane = tid & 31;
  const int k = i0 + (warp & 3) * 32 + lane;
  if (T > 0) {
    mbar_wait(&bars[(T - 1) & 1], ((T - 1) >> 1) & 1);
    tc_fence_after();
    const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
#pragma unroll 1
    for (int c0 = (warp >> 2) * 32; c0 < NTILE; c0 += 128) {
      float v[32];
      tmem_ld32(trow + c0, v);
      if (k < G::K) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          *reinterpret_cast<float4*>(out + (size_t)k * G::COUT + n0 + c0 + 4 * j) =
              make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
  } else if (k < G::K && warp < 4) {
    for (int c = 0; c < NTILE; ++c) out[(size_t)k * G::COUT + n0 + c] = 0.f;
  }
  if (do_bias) {
    // lanes of a warp hold the same columns (same tg) for different reduction rows: reduce in the warp first
    if (tg * CPT < GPR) {
#pragma unroll
      for (int q = 0; q < CPT; ++q)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float v = colsum[q][e];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          if (lane == 0) atomicAdd(&bias_sm[(tg * CPT + q) * 8 + e], v);
        }
    }
    __syncthreads();
    for (int j = tid; j < NTILE; j += TT) out[(size_t)G::K * G::COUT + n0 + j] = bias_sm[j];
  }
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ weight gradient, resident uint8 frames
// conv1 weight gradient  dW[k][n] = sum_rows X[row][k] * dZ[row][n]  (k = (ky,kx,frame) 0..255, n = 0..31) with the rows
// in the same pool-window order as the forward pass.  A CTA walks over consecutive 128-row tiles (32 pool windows):
//   * per tile three bulk copies, issued one tile ahead: the band of uint8 input rows (as in umma_fwd_res_u8_kernel),
//     the 32 windows' gated gradients (2 KB) and arg-max bytes (1 KB);
//   * the band is widened to bf16 once; A = X^T is built with 16-byte shared->shared copies as four [128 rows][64 k]
//     SWIZZLE_128B images (the forward pass's A tiles, read MN-major here: 128-byte contiguous operand reads); B = dZ
//     with the pool routing applied, in the no-swizzle MN-major layout (chunk (n/8, row) at (n/8)*2048 + (row/8)*128 +
//     (row%8)*16);
//   * 16 MMAs per tile (2 halves of k x 8 steps of 16 rows) accumulate into TMEM over ALL tiles of the CTA; one fp32
//     partial [(256+1) x 32] per CTA at the end (summed in fixed order by the optimizer).
template <class G>
struct WgradResU8 {
  using F = FwdResU8<G>;
  static constexpr uint32_t BAND_BYTES = F::PATCH_BYTES;                 // bf16 band
  static constexpr uint32_t RAW_BYTES = F::RAW_BYTES;                    // uint8 band, double-buffered
  static constexpr uint32_t GRAD_BYTES = 32 * G::COUT * 2;               // 32 windows x COUT bf16
  static constexpr uint32_t ARG_BYTES = 32 * G::COUT;
  static constexpr uint32_t A_BYTES = G::K * TM * 2;                     // 256 x 128 bf16 = 64 KB
  static constexpr uint32_t B_BYTES = G::COUT * TM * 2;                  // 32 x 128 bf16 = 8 KB
  static constexpr uint32_t OFF_A = BAND_BYTES, OFF_B = OFF_A + A_BYTES, OFF_RAW = OFF_B + B_BYTES;
  static constexpr uint32_t OFF_GRAD = OFF_RAW + 2 * RAW_BYTES, OFF_ARG = OFF_GRAD + 2 * round1k_c(GRAD_BYTES);
  static constexpr uint32_t SMEM = SMEM_SLACK + OFF_ARG + 2 * round1k_c(ARG_BYTES);
  static_assert(G::K == 256 && G::COUT == 32 && G::POOL, "written for the reference conv1 (8x8x4 -> 32, pooled)");
};

template <class G>
__global__ void __launch_bounds__(TT) umma_wgrad_res_u8_kernel(WgradArgsT a) {
  using R = WgradResU8<G>;
  using F = FwdResU8<G>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t bars[2];
  __shared__ uint64_t ldbar[2];       // a tile's band + gradients + arg-max have landed
  __shared__ uint64_t done;           // the MMAs of a tile have completed
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_part[TT / 32][G::COUT];
  const uint8_t* __restrict__ x = static_cast<const uint8_t*>(a.x);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    mbar_init(&ldbar[0], 1);
    mbar_init(&ldbar[1], 1);
    mbar_init(&done, 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, 64);     // two 128 x 32 accumulators (k halves)
  const uint32_t sbase = smem_u32(smem);
  const int R_total = a.B * G::ROWS;
  const int r_begin = blockIdx.x * a.rows_per_split;                   // rows_per_split is a multiple of 128
  const int r_end = min(R_total, r_begin + a.rows_per_split);
  const int nt = (r_end - r_begin + TM - 1) / TM;
  auto issue_loads = [&](int t) {     // thread 0: the three bulk copies of tile t into buffer t&1
    const int m0 = r_begin + t * TM, buf = t & 1;
    int glo, ghi, wf, wl;
    conv1_band<G>(m0, a.B, glo, ghi, wf, wl);
    const int nrows = ghi - glo + 1, nw = wl - wf + 1;
    if (nrows > F::PATCH_ROWS) __trap();
    const uint32_t bb = (uint32_t)nrows * F::ROW_BYTES, gb = (uint32_t)nw * G::COUT * 2, ab = (uint32_t)nw * G::COUT;
    mbar_expect_tx(&ldbar[buf], bb + gb + ab);
    bulk_g2s(sbase + R::OFF_RAW + buf * R::RAW_BYTES, x + (size_t)glo * F::ROW_BYTES, bb, &ldbar[buf]);
    bulk_g2s(sbase + R::OFF_GRAD + buf * round1k_c(R::GRAD_BYTES), a.dactg + (size_t)wf * G::COUT, gb, &ldbar[buf]);
    bulk_g2s(sbase + R::OFF_ARG + buf * round1k_c(R::ARG_BYTES), a.arg + (size_t)wf * G::COUT, ab, &ldbar[buf]);
  };
  pdl_wait_then_trigger();
  if (tid == 0 && nt > 0) issue_loads(0);
  // build mapping: tile row r = tid & 127 (lanes along the rows: conflict-free stores), chunk column cq = tid >> 7
  const int r = tid & (TM - 1), cq = tid >> 7;
  const uint32_t rdst = (uint32_t)((r >> 3) * 128 + (r & 7) * 16);
  float colsum[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) colsum[e] = 0.f;
  constexpr uint32_t idesc = make_idesc(TM, G::COUT, true, true);
  const int wu = warp_uniform_idx();
#pragma unroll 1
  for (int t = 0; t < nt; ++t) {
    const int m0 = r_begin + t * TM, buf = t & 1;
    if (tid == 0 && t + 1 < nt) issue_loads(t + 1);                    // buffer (t+1)&1 was consumed by tile t-1 (barrier below)
    int glo, ghi, wf, wl;
    conv1_band<G>(m0, a.B, glo, ghi, wf, wl);
    const Row row = decode_row<G>(m0 + r, a.B);
    const int iy0 = row.oy * G::ST - G::PADH, ix0 = row.ox * G::ST - G::PADW;
    const int poff = ((row.b * G::IH + iy0 - glo) * G::IW + ix0) * G::CIN;        // uint8 offset in the band (may be < 0)
    mbar_wait(&ldbar[buf], (t >> 1) & 1);
    {
      const int nchunks = (ghi - glo + 1) * (int)(F::ROW_BYTES / 8);
      const uint8_t* raw = smem + R::OFF_RAW + buf * R::RAW_BYTES;
      for (int c = tid; c < nchunks; c += TT) *reinterpret_cast<uint4*>(smem + c * 16) = widen_u8x8(*reinterpret_cast<const uint2*>(raw + c * 8));
    }
    if (t > 0) mbar_wait(&done, (t - 1) & 1);                          // the previous tile's MMAs have read A and B
    __syncthreads();
    // A = X^T: chunk gg (k = 8*gg..8*gg+7: filter row gg/4, pixels 2*(gg%4), +1) of row r
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int gg = cq + 4 * q;
      const int ky = gg >> 2, kx = 2 * (gg & 3);
      const bool ok = row.valid && (unsigned)(iy0 + ky) < (unsigned)G::IH && (unsigned)(ix0 + kx) < (unsigned)G::IW;
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (ok) v = *reinterpret_cast<const uint4*>(smem + 2 * (poff + (ky * G::IW + kx) * G::CIN));
      *reinterpret_cast<uint4*>(smem + R::OFF_A + (gg >> 3) * A_STAGE_BYTES + sw128_off(r, gg & 7)) = v;
    }
    // B = dZ: channels 8*cq..8*cq+7 of row r, routed through the pool (gradient goes to the arg-max position only)
    {
      uint4 z = make_uint4(0u, 0u, 0u, 0u);
      if (row.valid) {
        const int wl_ = (m0 + r) >> 2;                                  // this row's pool window
        const uint4 d = *reinterpret_cast<const uint4*>(smem + R::OFF_GRAD + buf * round1k_c(R::GRAD_BYTES) + ((wl_ - wf) * G::COUT + cq * 8) * 2);
        const uint2 ag = *reinterpret_cast<const uint2*>(smem + R::OFF_ARG + buf * round1k_c(R::ARG_BYTES) + (wl_ - wf) * G::COUT + cq * 8);
        z = route_chunk(d, ag, r & 3);
        add_chunk(colsum, z);
      }
      *reinterpret_cast<uint4*>(smem + R::OFF_B + cq * 2048 + rdst) = z;
    }
    fence_async_proxy();
    __syncthreads();
    if (wu == 0) {
      tc_fence_after();
      if (elect_one()) {
        const uint64_t bd0 = make_desc(sbase + R::OFF_B, 128, 2048);
#pragma unroll
        for (int mt = 0; mt < 2; ++mt) {
          // k half mt = images 2*mt, 2*mt+1 (LBO = one image); 16 rows per MMA = two 8-row groups (SBO = 1024)
          const uint64_t ad0 = make_desc_sw128_mn(sbase + R::OFF_A + mt * 2 * A_STAGE_BYTES, A_STAGE_BYTES, 1024);
#pragma unroll
          for (int j = 0; j < TM / 16; ++j)
            mma_bf16(tmem + mt * 32, desc_advance(ad0, j * 2048), desc_advance(bd0, j * 256), idesc, (t | j) ? 1u : 0u);
        }
        mma_commit(&done);
      }
      __syncwarp();
    }
  }
  pdl_trigger();
  float* out = a.partial + (size_t)blockIdx.x * ((size_t)(G::K + 1) * G::COUT);
  if (nt > 0) {
    mbar_wait(&done, (nt - 1) & 1);
    tc_fence_after();
    if ((warp >> 2) < 2) {                                              // warps 0-3: k 0..127, warps 4-7: k 128..255
      const int mt = warp >> 2, k = mt * 128 + (warp & 3) * 32 + lane;
      float v[32];
      tmem_ld32(tmem + ((uint32_t)((warp & 3) * 32) << 16) + mt * 32, v);
#pragma unroll
      for (int j = 0; j < 8; ++j)
        *reinterpret_cast<float4*>(out + (size_t)k * G::COUT + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    }
  } else {
    for (int i = tid; i < G::K * G::COUT; i += TT) out[i] = 0.f;
  }
  // bias gradient = column sums of dZ: lanes hold different rows of the same 8 channels -> warp sum, then a fixed-order
  // sum over the 4 warps of each channel group
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    float v = colsum[e];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) bias_part[warp][e] = v;
  }
  __syncthreads();
  if (tid < G::COUT) {
    const int g = tid >> 3, e = tid & 7;                                // warps 4g..4g+3 hold channel group g
    out[(size_t)G::K * G::COUT + tid] = ((bias_part[4 * g][e] + bias_part[4 * g + 1][e]) + bias_part[4 * g + 2][e]) + bias_part[4 * g + 3][e];
  }
  teardown(tmem, 64);
}

// ------------------------------------------------------------------------------------------------ data gradient
struct DgradArgsT {
  const __nv_bfloat16* wd;        // weights permuted to [CIN][KS*KS*COUT]
  const __nv_bfloat16* dactg;     // ReLU-gated gradient of this layer's pooled output
  const uint8_t* arg;
  const __nv_bfloat16* act_prev;  // previous layer's pooled output: its ReLU gate is applied to dx here
  __nv_bfloat16* dx;              // [B, IH, IW, CIN] gated gradient of the previous layer's pooled output
  int B;
  int cls_tile[5];                // stride-2 layers: first tile of each input-pixel parity class (see below)
};

// Stride-2 convolutions: an input pixel only receives gradient through the filter taps whose parity matches its own
// ((iy + pad - ky) even), i.e. through (KS/2)^2 of the KS^2 taps.  Input pixels are therefore processed in 4 parity
// classes (iy&1, ix&1); every tile belongs to one class and reduces over that class's taps only -- 4x fewer stages
// than the dense formulation, whose other taps are all zero.
template <class G>
struct DgradClasses {
  static constexpr int NY0 = (G::IH + 1) / 2, NY1 = G::IH / 2, NX0 = (G::IW + 1) / 2, NX1 = G::IW / 2;
  __host__ __device__ static constexpr int ny(int c) { return (c >> 1) ? NY1 : NY0; }
  __host__ __device__ static constexpr int nx(int c) { return (c & 1) ? NX1 : NX0; }
  __host__ __device__ static constexpr int count(int c) { return ny(c) * nx(c); }
};

// Data-gradient epilogue: TMEM -> ReLU gate of the layer below -> bf16 gradient of its pooled output.
template <class G, int NTILE>
__device__ __forceinline__ void dgrad_epilogue(uint32_t tmem, const DgradArgsT& a, int m, bool mvalid, int c0t) {
  const int tid = threadIdx.x;
  const int warp = tid >> 5;
  const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
#pragma unroll 1
  for (int c0 = (warp >> 2) * 32; c0 < NTILE; c0 += 128) {
    float v[32];
    tmem_ld32(trow + c0, v);
    if (mvalid) {
      const size_t off = (size_t)m * G::CIN + c0t + c0;
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // ReLU gate of the layer below: zero the gradient where its (pooled) output is not positive
        const uint4 p = *reinterpret_cast<const uint4*>(a.act_prev + off + 8 * j);
        const uint32_t pw[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const bool lo = ((pw[q] & 0x7FFFu) != 0u) && !(pw[q] & 0x8000u);
          const bool hi = ((pw[q] & 0x7FFF0000u) != 0u) && !(pw[q] & 0x80000000u);
          pk[4 * j + q] = pack_bf16(lo ? v[8 * j + 2 * q] : 0.f, hi ? v[8 * j + 2 * q + 1] : 0.f);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(a.dx + off + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    }
  }
}

constexpr uint32_t DG_PD = 144, DG_PA = 72;                          // raw gradient / arg-max row pitches (8 chunks + pad)
constexpr uint32_t DG_RAW_D = 128 * DG_PD, DG_RAW_A = 128 * DG_PA, DG_RAW_W = 1024;   // 18 KB, 9 KB, 1 KB
template <class G, int NTILE>
__host__ __device__ constexpr uint32_t dgrad_smem_bytes() {
  return SMEM_SLACK + (G::POOL ? RS * (DG_RAW_D + DG_RAW_A + DG_RAW_W) + 2u * A_STAGE_BYTES : RS * A_STAGE_BYTES) +
         RS * (uint32_t)NTILE * 128u;
}

template <class G, int NTILE>
__global__ void __launch_bounds__(TT) umma_dgrad_kernel(DgradArgsT a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  constexpr uint32_t B_STAGE = NTILE * 128;
  constexpr int KD = G::KS * G::KS * G::COUT;
  constexpr int CC = G::COUT / 8;                     // chunks per filter tap
  // POOL: raw rings (gradient chunks, arg-max chunks, window byte per chunk) + 2 routed A tiles; else direct A ring
  constexpr uint32_t OFF_RG = RS * DG_RAW_D, OFF_RW = OFF_RG + RS * DG_RAW_A, OFF_OPA = OFF_RW + RS * DG_RAW_W;
  constexpr uint32_t OFF_B = G::POOL ? OFF_OPA + 2 * A_STAGE_BYTES : RS * A_STAGE_BYTES;
  __shared__ uint64_t bars[2];
  __shared__ uint64_t wfull[RS];      // the weight tile of a ring slot has landed (bulk copy)
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * TM, c0t = blockIdx.y * NTILE;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < RS; ++i) mbar_init(&wfull[i], 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sbase = smem_u32(smem);
  constexpr bool S2 = G::ST == 2;                     // parity-class formulation
  static_assert(!S2 || (G::KS % 2 == 0 && G::COUT == TK), "stride-2 path: even filter, one tap per stage");
  // which rows does this tile cover?  dense: m0 + r over all input pixels; S2: r-th pixel of parity class `cls`
  int cls = 0, row0 = m0;
  if constexpr (S2) {
    // (no run-time index into the parameter block: that would move it to local memory)
    const int bx = (int)blockIdx.x, t1 = a.cls_tile[1], t2 = a.cls_tile[2], t3 = a.cls_tile[3];
    cls = (bx >= t1) + (bx >= t2) + (bx >= t3);
    row0 = (bx - (cls == 0 ? a.cls_tile[0] : cls == 1 ? t1 : cls == 2 ? t2 : t3)) * TM;
  }
  const int cy = cls >> 1, cx = cls & 1;
  auto decode = [&](int r, int& b, int& iy, int& ix) -> bool {
    if constexpr (S2) {
      const int nxc = DgradClasses<G>::nx(cls), cnt = DgradClasses<G>::count(cls);
      b = r / cnt;
      const int rem = r - b * cnt;
      const int yy = rem / nxc;
      iy = 2 * yy + cy;
      ix = 2 * (rem - yy * nxc) + cx;
      return r < a.B * cnt;
    } else {
      b = r / (G::IH * G::IW);
      const int rem = r - b * (G::IH * G::IW);
      iy = rem / G::IW;
      ix = rem - iy * G::IW;
      return r < a.B * G::IH * G::IW;
    }
  };
  // gather mapping: chunk id = tid + 512*q -> tile row id>>3, reduction chunk id&7 (8 lanes = 128 contiguous bytes)
  const int gc = tid & 7;
  int gb[2], giy[2], gix[2];
  bool gvalid[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) gvalid[q] = decode(row0 + (tid >> 3) + 64 * q, gb[q], giy[q], gix[q]);
  // route / epilogue mapping: tile row tid&127
  const int arow = tid & (TM - 1), ac0 = (tid >> 7) * 2;
  int eb, eiy, eix;
  const bool mvalid = decode(row0 + arow, eb, eiy, eix);
  const int m = (eb * G::IH + eiy) * G::IW + eix;
  constexpr int T = S2 ? (G::KS / 2) * (G::KS / 2) : KD / TK;
  static_assert(KD % TK == 0, "dgrad reduction must be a multiple of the stage depth");
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  const __nv_bfloat16* wimg = a.wd + (size_t)c0t * 64;      // stage image: CIN*64 elements per stage, rows c0t.. contiguous
  pdl_wait_then_trigger();
  auto issue = [&](int it) {
    if (it < T) {
      const int slot = it & (RS - 1);
      int nq, t;
      if constexpr (S2) {        // stage = one of the class's taps: ky = ((cy+pad)&1) + 2a, kx = ((cx+pad)&1) + 2b
        const int ta = it / (G::KS / 2), tb = it % (G::KS / 2);
        t = (((cy + G::PADH) & 1) + 2 * ta) * G::KS + ((cx + G::PADW) & 1) + 2 * tb;
        nq = gc;
      } else {
        const int kc8 = it * 8 + gc;
        nq = kc8 % CC;
        t = kc8 / CC;
      }
      const int kx = t % G::KS, ky = t / G::KS;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int r = (tid >> 3) + 64 * q;
        const int ty = giy[q] + G::PADH - ky, tx = gix[q] + G::PADW - kx;
        bool ok = gvalid[q] && ty >= 0 && tx >= 0 && (ty % G::ST) == 0 && (tx % G::ST) == 0;
        const int oy = ty / G::ST, ox = tx / G::ST;
        ok = ok && oy < G::OH && ox < G::OW;
        int pidx = 0, win = 0;
        if (ok) {
          if constexpr (G::POOL) {
            pidx = (gb[q] * G::PH + (oy >> 1)) * G::PW + (ox >> 1);
            win = (oy & 1) * 2 + (ox & 1);
          } else {
            pidx = (gb[q] * G::OH + oy) * G::OW + ox;
          }
        }
        const size_t off = (size_t)pidx * G::COUT + nq * 8;
        if constexpr (G::POOL) {
          cp_async16(sbase + slot * DG_RAW_D + r * DG_PD + gc * 16, a.dactg + off, ok);
          cp_async8(sbase + OFF_RG + slot * DG_RAW_A + r * DG_PA + gc * 8, a.arg + off, ok);
          smem[OFF_RW + slot * DG_RAW_W + r * 8 + gc] = (uint8_t)win;
        } else {
          cp_async16(sbase + slot * A_STAGE_BYTES + sw128_off(r, gc), a.dactg + off, ok);
        }
      }
      if (tid == 0) {                                    // weight tile of the stage (S2: the stage is filter tap t)
        mbar_expect_tx(&wfull[slot], B_STAGE);
        bulk_g2s(sbase + OFF_B + slot * B_STAGE, wimg + (size_t)(S2 ? t : it) * (G::CIN * 64), B_STAGE, &wfull[slot]);
      }
    }
    cp_async_commit();
  };
  issue(0);
  issue(1);
  const int wu = warp_uniform_idx();
  for (int it = 0; it < T; ++it) {
    const int slot = it & (RS - 1), par = it & 1;
    if (it >= 2) mbar_wait(&bars[par], ((it >> 1) - 1) & 1);
    issue(it + 2);
    cp_async_wait<2>();
    uint32_t a_addr;
    if constexpr (G::POOL) {
      __syncthreads();
      uint8_t* opa = smem + OFF_OPA + par * A_STAGE_BYTES;
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = ac0 + cc;
        const uint4 d = *reinterpret_cast<const uint4*>(smem + slot * DG_RAW_D + arow * DG_PD + c * 16);
        const uint2 ag = *reinterpret_cast<const uint2*>(smem + OFF_RG + slot * DG_RAW_A + arow * DG_PA + c * 8);
        const int win = (int)smem[OFF_RW + slot * DG_RAW_W + arow * 8 + c];
        *reinterpret_cast<uint4*>(opa + kmajor_off(arow, c)) = route_chunk(d, ag, win);
      }
      a_addr = sbase + OFF_OPA + par * A_STAGE_BYTES;
    } else {
      a_addr = sbase + slot * A_STAGE_BYTES;
    }
    fence_async_proxy();
    __syncthreads();
    if (wu == 0) {
      mbar_wait(&wfull[slot], (it >> 2) & 1);
      issue_stage<!G::POOL, true>(tmem, a_addr, sbase + OFF_B + slot * B_STAGE, idesc, it == 0, &bars[par]);
    }
  }
  pdl_trigger();
  mbar_wait(&bars[(T - 1) & 1], ((T - 1) >> 1) & 1);
  tc_fence_after();
  dgrad_epilogue<G, NTILE>(tmem, a, m, mvalid, c0t);
  teardown(tmem, tmem_cols(NTILE));
}

// ------------------------------------------------------------------------------------------------ data gradient, resident
// Pooled conv layers (conv2: stride-2 parity classes, conv3: dense 3x3).  Same idea as umma_fwd_res_kernel: the
// (ReLU-gated) gradient of the pooled output and the pool arg-max of the tile's samples are two small contiguous
// blocks, fetched once with bulk copies; all weight stages are fetched before the PDL wait; every filter tap is one
// k-block (COUT == 64) whose A tile is built shared->shared with the pool routing applied on the way, NK taps per
// barrier round.
template <class G, bool S2>
struct DgradRes {
  static constexpr int RPS_MIN = S2 ? DgradClasses<G>::count(3) : G::IH * G::IW;    // fewest tile rows one sample contributes
  static constexpr int NS = (TM + RPS_MIN - 2) / RPS_MIN + 1;
  static constexpr uint32_t D_BYTES = G::PH * G::PW * G::COUT * 2;                   // gradient bytes per sample
  static constexpr uint32_t A_BYTES = G::PH * G::PW * G::COUT;                       // arg-max bytes per sample
  static constexpr uint32_t OFF_ARG = round1k_c(NS * D_BYTES);
  static constexpr uint32_t PATCH_BYTES = OFF_ARG + round1k_c(NS * A_BYTES);
  static constexpr int T = S2 ? (G::KS / 2) * (G::KS / 2) : G::KS * G::KS;           // taps = k-blocks
  static_assert(G::COUT == TK && G::POOL, "one filter tap per 64-deep k-block, pooled layers only");
  static_assert(D_BYTES % 16 == 0 && A_BYTES % 16 == 0, "bulk copies move multiples of 16 bytes");
};
template <class G, int NTILE, int NK, int NA>
__host__ __device__ constexpr uint32_t dgrad_res_smem_bytes() {
  using R = DgradRes<G, G::ST == 2>;
  return SMEM_SLACK + R::PATCH_BYTES + NA * NK * A_STAGE_BYTES + R::T * (uint32_t)NTILE * 128u;
}

template <class G, int NTILE, int NK, int NA>
__global__ void __launch_bounds__(TT) umma_dgrad_res_kernel(DgradArgsT a) {
  constexpr bool S2 = G::ST == 2;
  using R = DgradRes<G, S2>;
  constexpr int T = R::T, NBS = T / NK;
  static_assert(T % NK == 0, "taps must split into whole big stages");
  static_assert(!S2 || (G::KS % 2 == 0 && G::PADH == 1 && G::PADW == 1), "parity classes assume an even filter and pad 1");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  constexpr uint32_t B_STAGE = NTILE * 128, A_SLOT = NK * A_STAGE_BYTES;
  constexpr uint32_t OFF_A = R::PATCH_BYTES, OFF_B = OFF_A + NA * A_SLOT;
  __shared__ uint64_t bars[2];
  __shared__ uint64_t afull[NA];
  __shared__ uint64_t wfull[NBS];
  __shared__ uint64_t done[NBS];
  __shared__ uint64_t patchbar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, lane = tid & 31;
  const int c0t = blockIdx.y * NTILE;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < NA; ++i) mbar_init(&afull[i], TT / 32);
#pragma unroll
    for (int i = 0; i < NBS; ++i) { mbar_init(&wfull[i], 1); mbar_init(&done[i], 1); }
    mbar_init(&patchbar, 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, tmem_cols(NTILE));
  const uint32_t sbase = smem_u32(smem);
  // which rows does this tile cover?  dense: m0 + r over all input pixels; S2: r-th pixel of parity class `cls`
  int cls = 0, row0 = blockIdx.x * TM;
  if constexpr (S2) {
    // (no run-time index into the parameter block: that would move it to local memory)
    const int bx = (int)blockIdx.x, t1 = a.cls_tile[1], t2 = a.cls_tile[2], t3 = a.cls_tile[3];
    cls = (bx >= t1) + (bx >= t2) + (bx >= t3);
    row0 = (bx - (cls == 0 ? a.cls_tile[0] : cls == 1 ? t1 : cls == 2 ? t2 : t3)) * TM;
  }
  const int cy = cls >> 1, cx = cls & 1;
  // weights first (they do not depend on the previous kernel); S2: k-block i is the class's i-th tap
  auto tap_of = [&](int i) -> int {
    if constexpr (S2) return ((1 - cy) + 2 * (i / (G::KS / 2))) * G::KS + (1 - cx) + 2 * (i % (G::KS / 2));
    else return i;
  };
  if (tid == 0) {
    const __nv_bfloat16* wimg = a.wd + (size_t)c0t * 64;
#pragma unroll
    for (int bs = 0; bs < NBS; ++bs) {
      mbar_expect_tx(&wfull[bs], NK * B_STAGE);
#pragma unroll
      for (int kb = 0; kb < NK; ++kb)
        bulk_g2s(sbase + OFF_B + (bs * NK + kb) * B_STAGE, wimg + (size_t)tap_of(bs * NK + kb) * (G::CIN * 64), B_STAGE, &wfull[bs]);
    }
  }
  auto decode = [&](int r, int& b, int& iy, int& ix) -> bool {
    if constexpr (S2) {
      const int nxc = DgradClasses<G>::nx(cls), cnt = DgradClasses<G>::count(cls);
      b = r / cnt;
      const int rem = r - b * cnt;
      const int yy = rem / nxc;
      iy = 2 * yy + cy;
      ix = 2 * (rem - yy * nxc) + cx;
      return r < a.B * cnt;
    } else {
      b = r / (G::IH * G::IW);
      const int rem = r - b * (G::IH * G::IW);
      iy = rem / G::IW;
      ix = rem - iy * G::IW;
      return r < a.B * G::IH * G::IW;
    }
  };
  const int rps = S2 ? DgradClasses<G>::count(cls) : G::IH * G::IW;
  const int b0 = row0 / rps;
  int b1 = (row0 + TM - 1) / rps;
  if (b1 > a.B - 1) b1 = a.B - 1;
  // build mapping: tile rows (tid>>3) and (tid>>3)+64, channel chunk gc of the tap
  const int gc = tid & 7;
  int gb[2], giy[2], gix[2];
  bool gvalid[2];
  uint32_t adst[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int rr = (tid >> 3) + 64 * q;
    gvalid[q] = decode(row0 + rr, gb[q], giy[q], gix[q]);
    gb[q] -= b0;
    adst[q] = OFF_A + sw128_off(rr, gc);
  }
  // epilogue mapping: tile row tid&127
  int eb, eiy, eix;
  const bool mvalid = decode(row0 + (tid & (TM - 1)), eb, eiy, eix);
  const int m = (eb * G::IH + eiy) * G::IW + eix;
  pdl_wait_then_trigger();
  if (tid == 0) {
    const uint32_t nb = (uint32_t)(b1 - b0 + 1);
    mbar_expect_tx(&patchbar, nb * (R::D_BYTES + R::A_BYTES));
    bulk_g2s(sbase, a.dactg + (size_t)b0 * (R::D_BYTES / 2), nb * R::D_BYTES, &patchbar);
    bulk_g2s(sbase + R::OFF_ARG, a.arg + (size_t)b0 * R::A_BYTES, nb * R::A_BYTES, &patchbar);
  }
  constexpr uint32_t idesc = make_idesc(TM, NTILE, false, false);
  const int wu = warp_uniform_idx();
  mbar_wait(&patchbar, 0);
#pragma unroll 1
  for (int bs = 0; bs < NBS; ++bs) {
    const int as = bs % NA;
    if (bs >= NA) mbar_wait(&done[bs - NA], 0);
#pragma unroll
    for (int kb = 0; kb < NK; ++kb) {
      const int t = tap_of(bs * NK + kb);
      const int ky = t / G::KS, kx = t - ky * G::KS;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int ty = giy[q] + G::PADH - ky, tx = gix[q] + G::PADW - kx;
        bool ok = gvalid[q] && ty >= 0 && tx >= 0 && (ty % G::ST) == 0 && (tx % G::ST) == 0;
        const int oy = ty / G::ST, ox = tx / G::ST;
        ok = ok && oy < G::OH && ox < G::OW;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (ok) {
          const int pl = ((gb[q] * G::PH + (oy >> 1)) * G::PW + (ox >> 1)) * G::COUT + gc * 8;
          const uint4 d = *reinterpret_cast<const uint4*>(smem + pl * 2);
          const uint2 ag = *reinterpret_cast<const uint2*>(smem + R::OFF_ARG + pl);
          v = route_chunk(d, ag, (oy & 1) * 2 + (ox & 1));
        }
        *reinterpret_cast<uint4*>(smem + as * A_SLOT + kb * A_STAGE_BYTES + adst[q]) = v;
      }
    }
    fence_async_proxy();
    __syncwarp();
    if (lane == 0) mbar_arrive(&afull[as]);
    if (wu == 0) {
      mbar_wait(&afull[as], (bs / NA) & 1);
      mbar_wait(&wfull[bs], 0);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t ad0 = make_desc_sw128(sbase + OFF_A + as * A_SLOT), bd0 = make_desc_sw128(sbase + OFF_B + bs * NK * B_STAGE);
#pragma unroll
        for (int kb = 0; kb < NK; ++kb) {
#pragma unroll
          for (int j = 0; j < TK / 16; ++j)
            mma_bf16(tmem, desc_advance(ad0, kb * A_STAGE_BYTES + j * 32), desc_advance(bd0, kb * B_STAGE + j * 32), idesc,
                     (bs | kb | j) ? 1u : 0u);
        }
        mma_commit(&done[bs]);
      }
      __syncwarp();
    }
  }
  pdl_trigger();
  mbar_wait(&done[NBS - 1], 0);
  tc_fence_after();
  dgrad_epilogue<G, NTILE>(tmem, a, m, mvalid, c0t);
  teardown(tmem, tmem_cols(NTILE));
}

}  // namespace umma
}  // namespace dqn
