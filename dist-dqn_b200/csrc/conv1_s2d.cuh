// conv1 of the reference ConvNetwork (network.py:395-407: 8x8 stride-4 SAME convolution of the 84x84x4 frame stack,
// bias, ReLU, 2x2 max-pool) and its weight gradient on the tcgen05 tensor cores, in a space-to-depth formulation.
//
// Padded to 88x88 (TF SAME: 2 before, 2 after) and cut into 4x4-pixel super-pixels, the stack is a 22x22 image with
// 64 channels (4x4 pixels x 4 frames), and the 8x8/4 convolution is a 2x2/1 convolution over it.  One super-pixel =
// 64 bf16 = 128 bytes = one row of a SWIZZLE_128B operand tile, so NO im2col expansion is needed: with the super-pixels
// stored as four parity planes (Y&1, X&1) of 11x11 rows each, the A operand of filter tap (dy,dx) for the conv outputs
// at pool-window position (wy,wx) of ALL 121 pool windows of a sample is a run of 128 consecutive rows of ONE plane,
// starting (wy+dy, wx+dx)-dependent rows into it -- a descriptor start address, not a copy.  The M rows of the GEMM are
// the pool windows (121 of 128 used) and the four window positions are four accumulators in TMEM (4 x 32 columns), so
// the 2x2 max-pool of the epilogue is a per-thread maximum over four TMEM loads: no shuffles, no row decoding.
//
//   image row of super-pixel (Y, X):  (Y&1)*248 + (X&1)*124 + (Y>>1)*11 + (X>>1)        (512 rows x 128 B = 64 KB)
//   (the plane pitch 124 = 4 mod 8 makes the stores of two neighbouring super-pixels hit different swizzle phases)
//   channel of (yy, xx, f) inside a super-pixel: yy*16 + xx*4 + f;  weight reduction index k' = (dy*2+dx)*64 + channel
//
// Forward kernel: persistent, one CTA per SM, warp-specialised.
//   warp 0        bulk-copies the uint8 stack of the next sample into a 2-deep raw ring (from a stack array, or frame
//                 by frame straight out of the replay ring: the gather fused away)
//   warps 6..15   widen the bytes to bf16 and scatter them into the 2-deep image ring (16 source bytes -> two 16-byte
//                 chunks; every byte is read once and every image byte written once)
//   warp 1        issues the 64 MMAs (128x32x16) of a sample from one elected lane into one of two accumulator sets
//   warps 2..5    epilogue: bias, ReLU, max over the four window positions, arg-max byte, bf16 store
// so the epilogue of sample i, the MMAs of sample i+1 and the build of sample i+2 overlap.
//
// Weight-gradient kernel: the same image read MN-major (the reduction runs over the pool windows), the routed output
// gradient as B operand, accumulation in TMEM over all samples of the CTA, one partial per CTA.
#pragma once
#include "umma_core.cuh"

namespace dqn {
namespace umma {
namespace s2d {

constexpr int PW = 11, NWIN = 121;                     // pool windows per sample
constexpr int PLANE_X = 124, PLANE_Y = 248, IMG_ROWS = 512;
constexpr uint32_t IMG_BYTES = IMG_ROWS * 128;         // 64 KB
constexpr uint32_t FRAME_BYTES = 84 * 84;              // 7056
constexpr uint32_t STACK_BYTES = 4 * FRAME_BYTES;      // 28224
constexpr uint32_t RAW_BYTES = 28 * 1024;              // raw ring slot
constexpr int UNITS = 84 * 21;                         // 16-byte source units per stack (4 pixels x 4 frames)
constexpr uint32_t W_BYTES = 256 * 32 * 2;             // conv1 weight image, 16 KB
constexpr int NWARPS = 16;

__host__ __device__ constexpr int img_row(int Y, int X) { return (Y & 1) * PLANE_Y + (X & 1) * PLANE_X + (Y >> 1) * PW + (X >> 1); }
// reduction index of the permuted conv1 weight image for the reference index k = (ky*8 + kx)*4 + f
__host__ __device__ constexpr int kperm(int k) {
  return (((k >> 5) >> 2) * 2 + (((k >> 2) & 7) >> 2)) * 64 + ((k >> 5) & 3) * 16 + ((k >> 2) & 3) * 4 + (k & 3);
}
// ... and back: reference index of image channel c64 of tap (dy, dx)
__host__ __device__ constexpr int kref(int dy, int dx, int c64) {
  return ((4 * dy + (c64 >> 4)) * 8 + 4 * dx + ((c64 >> 2) & 3)) * 4 + (c64 & 3);
}

struct Src {
  const uint8_t* x[2];         // pass p: stacks [B][84][84][4] (null = take the frames from the ring)
  // replay ring (fused gather): sample b = logical index idx[b]; pass 0 reads the state stack, pass 1 the next-state stack
  const uint8_t* frames;
  const void* meta;
  const int64_t* ring_state;
  const int32_t* idx;
  int64_t capacity, frame_capacity;
  int32_t pitch;
  // scalars of the sampled transitions, written by pass 0 in ring mode (null = not wanted)
  int32_t* actions;
  double* rewards;
  uint8_t* terminals;
};

struct RingRec { int64_t state_pos, next_pos; double reward; int32_t action; uint8_t terminal; uint8_t pad[3]; };

// one elected thread: start the copy of sample b's stack into `dst`; completion (STACK_BYTES) is counted on `bar`
__device__ __forceinline__ void load_stack(const Src& s, int pass, int b, uint32_t dst, uint64_t* bar) {
  mbar_expect_tx(bar, STACK_BYTES);
  const uint8_t* x = pass ? s.x[1] : s.x[0];
  if (x) {
    bulk_g2s(dst, x + (size_t)b * STACK_BYTES, STACK_BYTES, bar);
    return;
  }
  int64_t slot = s.ring_state[0] + (int64_t)s.idx[b];
  if (slot >= s.capacity) slot -= s.capacity;
  const RingRec* r = static_cast<const RingRec*>(s.meta) + slot;
  const int64_t pos = pass ? r->next_pos : r->state_pos;
#pragma unroll
  for (int f = 0; f < 4; ++f) {
    const int64_t fs = (pos - 3 + f) % s.frame_capacity;
    bulk_g2s(dst + f * FRAME_BYTES, s.frames + fs * (int64_t)s.pitch, FRAME_BYTES, bar);
  }
  if (pass == 0) {
    if (s.actions) s.actions[b] = r->action;
    if (s.rewards) s.rewards[b] = r->reward;
    if (s.terminals) s.terminals[b] = r->terminal;
  }
}

// 4 pixels of frames a,b,c,d (one 32-bit word each) -> 16 bytes in HWC order (pixel-major, frame-minor)
__device__ __forceinline__ uint4 interleave4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  const uint32_t ab_lo = __byte_perm(a, b, 0x5140), ab_hi = __byte_perm(a, b, 0x7362);
  const uint32_t cd_lo = __byte_perm(c, d, 0x5140), cd_hi = __byte_perm(c, d, 0x7362);
  return make_uint4(__byte_perm(ab_lo, cd_lo, 0x5410), __byte_perm(ab_lo, cd_lo, 0x7632), __byte_perm(ab_hi, cd_hi, 0x5410),
                    __byte_perm(ab_hi, cd_hi, 0x7632));
}

// Builder threads (nb of them, this one is bt): widen the raw stack and scatter it into the image.  PLANAR: the raw
// slot holds 4 frames back to back (ring mode) instead of the HWC stack.
template <bool PLANAR>
__device__ __forceinline__ void build_image(const uint8_t* raw, uint8_t* img, int bt, int nb) {
#pragma unroll 2
  for (int u = bt; u < UNITS; u += nb) {
    uint4 v;
    if constexpr (PLANAR) {
      const uint32_t* p = reinterpret_cast<const uint32_t*>(raw) + u;
      v = interleave4(p[0], p[FRAME_BYTES / 4], p[2 * (FRAME_BYTES / 4)], p[3 * (FRAME_BYTES / 4)]);
    } else {
      v = reinterpret_cast<const uint4*>(raw)[u];
    }
    const int iy = u / 21, j = u - iy * 21;
    const int Y = (iy + 2) >> 2, yy = (iy + 2) & 3;
    const int rbase = (Y & 1) * PLANE_Y + (Y >> 1) * PW;
    const int ra = rbase + (j & 1) * PLANE_X + (j >> 1);               // super-pixel X = j:   pixels xx = 2,3 -> chunk 2*yy+1
    const int rb = rbase + ((j + 1) & 1) * PLANE_X + ((j + 1) >> 1);   // super-pixel X = j+1: pixels xx = 0,1 -> chunk 2*yy
    *reinterpret_cast<uint4*>(img + ra * 128 + (((2 * yy + 1) ^ (ra & 7)) << 4)) = widen_u8x8(make_uint2(v.x, v.y));
    *reinterpret_cast<uint4*>(img + rb * 128 + (((2 * yy) ^ (rb & 7)) << 4)) = widen_u8x8(make_uint2(v.z, v.w));
  }
}

// ------------------------------------------------------------------------------------------------ forward
struct FwdArgs {
  Src src;
  const __nv_bfloat16* wt[2];   // k-permuted weight images (kperm): 4 taps x [32 rows][64] SWIZZLE_128B
  const float* bias[2];
  __nv_bfloat16* out[2];        // pooled activation [B][121][32]
  uint8_t* arg[2];              // pool arg-max [B][121][32] (may be null)
  int B, npass;
};

constexpr uint32_t FWD_OFF_IMG = 2 * RAW_BYTES, FWD_OFF_W = FWD_OFF_IMG + 2 * IMG_BYTES;
constexpr uint32_t FWD_SMEM = SMEM_SLACK + FWD_OFF_W + W_BYTES;        // 205,824 bytes
constexpr int FWD_BUILD_WARP0 = 6, FWD_BUILD_WARPS = NWARPS - FWD_BUILD_WARP0;

template <bool PLANAR>
__global__ void __launch_bounds__(NWARPS * 32, 1) conv1_s2d_fwd_kernel(FwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t raw_full[2], raw_empty[2], img_full[2], img_empty[2], acc_full[2], acc_empty[2], wbar;
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[32];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = warp_uniform_idx();
  // tile = (pass, sample); a CTA stays on one pass so that it needs one set of weights
  const int nper = a.npass == 2 ? (int)gridDim.x >> 1 : (int)gridDim.x;          // CTAs per pass
  const int pass = a.npass == 2 ? (int)(blockIdx.x & 1) : 0;
  const int b0 = a.npass == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ntiles = b0 < a.B ? (a.B - b0 + nper - 1) / nper : 0;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      mbar_init(&raw_full[i], 1);
      mbar_init(&raw_empty[i], FWD_BUILD_WARPS);
      mbar_init(&img_full[i], FWD_BUILD_WARPS);
      mbar_init(&img_empty[i], 1);
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], 4);
    }
    mbar_init(&wbar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 256);                 // two sets of four 128x32 fp32 accumulators
  // the padding of the images (border super-pixels, rows between the planes) is never written again: zero it once
  for (int i = tid; i < (int)(2 * IMG_BYTES / 16); i += NWARPS * 32)
    reinterpret_cast<uint4*>(smem + FWD_OFF_IMG)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_async_proxy();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t sbase = smem_u32(smem);
  pdl_wait_then_trigger();                                    // everything above overlapped the previous kernel
  if (tid < 32) bias_sm[tid] = __ldg((pass ? a.bias[1] : a.bias[0]) + tid);
  __syncthreads();
  if (warp == 0) {
    // ---------------------------------------------------------------- producer
    if (elect_one()) {
      mbar_expect_tx(&wbar, W_BYTES);
      bulk_g2s(sbase + FWD_OFF_W, pass ? a.wt[1] : a.wt[0], W_BYTES, &wbar);
      for (int i = 0; i < ntiles; ++i) {
        const int rb = i & 1;
        mbar_wait(&raw_empty[rb], ((i >> 1) & 1) ^ 1);
        load_stack(a.src, pass, b0 + i * nper, sbase + rb * RAW_BYTES, &raw_full[rb]);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    constexpr uint32_t idesc = make_idesc(TM, 32, false, false);
    mbar_wait(&wbar, 0);
    for (int i = 0; i < ntiles; ++i) {
      const int ib = i & 1, ph = (i >> 1) & 1;
      mbar_wait(&img_full[ib], ph);
      mbar_wait(&acc_empty[ib], ph ^ 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t img = sbase + FWD_OFF_IMG + ib * IMG_BYTES;
        const uint64_t bd0 = make_desc_sw128(sbase + FWD_OFF_W);
#pragma unroll
        for (int w = 0; w < 4; ++w) {
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            const int s = (w >> 1) + (d >> 1), t = (w & 1) + (d & 1);
            const uint64_t ad0 = make_desc_sw128(img + (uint32_t)img_row(s, t) * 128u);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              mma_bf16(tmem + ib * 128 + w * 32, desc_advance(ad0, j * 32), desc_advance(bd0, d * 4096 + j * 32), idesc,
                       (d | j) ? 1u : 0u);
          }
        }
        mma_commit(&img_empty[ib]);
        mma_commit(&acc_full[ib]);
      }
      __syncwarp();
    }
  } else if (warp < FWD_BUILD_WARP0) {
    // ---------------------------------------------------------------- epilogue (TMEM lane quarter = warp & 3)
    const int q = warp & 3, m = 32 * q + lane;
    const int py = m / PW, px = m - py * PW;
    const bool valid = m < NWIN;
    __nv_bfloat16* outp = pass ? a.out[1] : a.out[0];
    uint8_t* argp = pass ? a.arg[1] : a.arg[0];
    for (int i = 0; i < ntiles; ++i) {
      const int ib = i & 1, ph = (i >> 1) & 1;
      const int b = b0 + i * nper;
      mbar_wait(&acc_full[ib], ph);
      tc_fence_after();
      uint32_t key[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) key[j] = 0u;
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        float v[32];
        tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + ib * 128 + w * 32, v);
        // window position (wy, wx) lies inside the 21x21 conv output unless it is the pool's padding row / column
        const bool in = valid && ((w >> 1) == 0 || py < PW - 1) && ((w & 1) == 0 || px < PW - 1);
        // values are >= 0 after ReLU, so their float bits order like unsigned integers: (value with the 2 lowest
        // mantissa bits replaced by 3-w) is a key whose maximum is the largest value, first window position on ties
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const uint32_t kj = in ? ((__float_as_uint(fmaxf(v[j] + bias_sm[j], 0.f)) & ~3u) | (3u - (uint32_t)w)) : 0u;
          key[j] = max(key[j], kj);
        }
      }
      // the accumulators are in registers: hand the TMEM set back before the global stores
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[ib]);
      if (valid) {
        const size_t off = ((size_t)b * NWIN + m) * 32;
        uint32_t pk[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) pk[j] = pack_bf16(__uint_as_float(key[2 * j] & ~3u), __uint_as_float(key[2 * j + 1] & ~3u));
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<uint4*>(outp + off + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        if (argp) {
          uint32_t ak[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            ak[j] = (3u - (key[4 * j] & 3u)) | ((3u - (key[4 * j + 1] & 3u)) << 8) | ((3u - (key[4 * j + 2] & 3u)) << 16) |
                    ((3u - (key[4 * j + 3] & 3u)) << 24);
          *reinterpret_cast<uint4*>(argp + off) = make_uint4(ak[0], ak[1], ak[2], ak[3]);
          *reinterpret_cast<uint4*>(argp + off + 16) = make_uint4(ak[4], ak[5], ak[6], ak[7]);
        }
      }
    }
  } else {
    // ---------------------------------------------------------------- builders
    const int bt = tid - FWD_BUILD_WARP0 * 32;
    for (int i = 0; i < ntiles; ++i) {
      const int rb = i & 1, ph = (i >> 1) & 1;
      mbar_wait(&raw_full[rb], ph);
      mbar_wait(&img_empty[rb], ph ^ 1);
      build_image<PLANAR>(smem + rb * RAW_BYTES, smem + FWD_OFF_IMG + rb * IMG_BYTES, bt, FWD_BUILD_WARPS * 32);
      fence_async_proxy();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&img_full[rb]);
        mbar_arrive(&raw_empty[rb]);
      }
    }
  }
  pdl_trigger();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

// ------------------------------------------------------------------------------------------------ weight gradient
// dW(dy,dx)[c64][n] = sum_samples sum_(wy,wx) sum_m image(wy+dy, wx+dx)[m][c64] * dZ_(wy,wx)[m][n], where dZ_w is the
// (ReLU-gated) gradient of the pooled output routed to window position w (arg-max == w), m runs over the pool windows.
// As MMAs: M = 128 = the two dy taps x 64 channels (two 64-element blocks of the MN-major A operand, LBO apart: both
// taps read plane rows a fixed distance apart), N = 32, reduction over m in 8 steps of 16.  For wy = 1 the row distance
// between tap dy = 0 and dy = 1 is negative, so there the blocks are taken in the order (dy = 1, dy = 0) and accumulated
// separately: four accumulators (wy, dx); the epilogue adds accumulator (1, dx) to (0, dx) with its halves swapped.
struct WgradArgs {
  Src src;                       // pass 0 only
  const __nv_bfloat16* dactg;    // [B][121][32] ReLU-gated gradient of the pooled conv1 output
  const uint8_t* arg;            // [B][121][32]
  float* partial;                // [gridDim.x][(256+1)*32] fp32, reference index order
  int B;
};

constexpr uint32_t WG_GRAD_BYTES = NWIN * 32 * 2, WG_ARG_BYTES = NWIN * 32;          // 7744, 3872 per sample
constexpr uint32_t WG_GA_SLOT = 12 * 1024;                                           // gradient (8 KB) + arg-max (4 KB) staging
constexpr uint32_t WG_OFF_IMG = 2 * RAW_BYTES, WG_OFF_DZ = WG_OFF_IMG + IMG_BYTES, WG_OFF_GA = WG_OFF_DZ + 4 * 8192;
constexpr uint32_t WG_SMEM = SMEM_SLACK + WG_OFF_GA + 2 * WG_GA_SLOT;                // 181,248 bytes
constexpr int WG_BUILD_WARP0 = 2, WG_BUILD_WARPS = NWARPS - WG_BUILD_WARP0, WG_BUILDERS = WG_BUILD_WARPS * 32;

template <bool PLANAR>
__global__ void __launch_bounds__(NWARPS * 32, 1) conv1_s2d_wgrad_kernel(WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t ld_full[2], ld_empty[2], op_full, op_empty;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = warp_uniform_idx();
  const int G = gridDim.x, b0 = blockIdx.x;
  const int ntiles = b0 < a.B ? (a.B - b0 + G - 1) / G : 0;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) { mbar_init(&ld_full[i], 2); mbar_init(&ld_empty[i], WG_BUILD_WARPS); }   // ld_full: two expect_tx arrivals (gradient + arg-max, stack)
    mbar_init(&op_full, WG_BUILD_WARPS);
    mbar_init(&op_empty, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 128);                 // four 128x32 fp32 accumulators
  // image padding and the unused dZ rows (m >= 121) stay zero for the whole kernel (the reduction runs over them)
  for (int i = tid; i < (int)((IMG_BYTES + 4 * 8192) / 16); i += NWARPS * 32)
    reinterpret_cast<uint4*>(smem + WG_OFF_IMG)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_async_proxy();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t sbase = smem_u32(smem);
  pdl_wait_then_trigger();
  float colsum[2][8];                                          // bias gradient: this thread's (m, channel chunk) items
#pragma unroll
  for (int q = 0; q < 2; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e) colsum[q][e] = 0.f;
  if (warp == 0) {
    // ---------------------------------------------------------------- producer: stack + gradient + arg-max of a sample
    if (elect_one()) {
      for (int i = 0; i < ntiles; ++i) {
        const int rb = i & 1, b = b0 + i * G;
        mbar_wait(&ld_empty[rb], ((i >> 1) & 1) ^ 1);
        mbar_expect_tx(&ld_full[rb], WG_GRAD_BYTES + WG_ARG_BYTES);
        bulk_g2s(sbase + WG_OFF_GA + rb * WG_GA_SLOT, a.dactg + (size_t)b * NWIN * 32, WG_GRAD_BYTES, &ld_full[rb]);
        bulk_g2s(sbase + WG_OFF_GA + rb * WG_GA_SLOT + 8192, a.arg + (size_t)b * NWIN * 32, WG_ARG_BYTES, &ld_full[rb]);
        load_stack(a.src, 0, b, sbase + rb * RAW_BYTES, &ld_full[rb]);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    constexpr uint32_t idesc = make_idesc(TM, 32, true, true);
    for (int i = 0; i < ntiles; ++i) {
      mbar_wait(&op_full, i & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t img = sbase + WG_OFF_IMG;
#pragma unroll
        for (int wy = 0; wy < 2; ++wy) {
#pragma unroll
          for (int dx = 0; dx < 2; ++dx) {
#pragma unroll
            for (int wx = 0; wx < 2; ++wx) {
              const int t = wx + dx;
              // block 0 / block 1 of the M = 128 rows: wy = 0 -> taps dy = 0, 1 (rows +248); wy = 1 -> taps dy = 1, 0 (rows +237)
              const int r0 = wy ? img_row(2, t) : img_row(0, t);
              const int r1 = img_row(1, t);
              const uint64_t ad0 = make_desc_sw128_mn(img + (uint32_t)r0 * 128u, (uint32_t)(r1 - r0) * 128u, 1024);
              const uint64_t bd0 = make_desc(sbase + WG_OFF_DZ + (wy * 2 + wx) * 8192, 128, 2048);
#pragma unroll
              for (int j = 0; j < 8; ++j)
                mma_bf16(tmem + (wy * 2 + dx) * 32, desc_advance(ad0, j * 2048), desc_advance(bd0, j * 256), idesc,
                         (i | wx | j) ? 1u : 0u);
            }
          }
        }
        mma_commit(&op_empty);
      }
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- builders: image + routed gradient tiles
    const int bt = tid - WG_BUILD_WARP0 * 32;
    for (int i = 0; i < ntiles; ++i) {
      const int rb = i & 1;
      mbar_wait(&ld_full[rb], (i >> 1) & 1);
      if (i > 0) mbar_wait(&op_empty, (i - 1) & 1);             // the previous sample's MMAs have read the operands
      build_image<PLANAR>(smem + rb * RAW_BYTES, smem + WG_OFF_IMG, bt, WG_BUILDERS);
      const uint8_t* ga = smem + WG_OFF_GA + rb * WG_GA_SLOT;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int item = bt + q * WG_BUILDERS;                  // item = (m, channel chunk cq): m = item >> 2
        if (item < NWIN * 4) {
          const int m = item >> 2, cq = item & 3;
          const uint4 d = *reinterpret_cast<const uint4*>(ga + (m * 32 + cq * 8) * 2);
          const uint2 ag = *reinterpret_cast<const uint2*>(ga + 8192 + m * 32 + cq * 8);
          add_chunk(colsum[q], d);
          const uint32_t off = (uint32_t)(cq * 2048 + (m >> 3) * 128 + (m & 7) * 16);
#pragma unroll
          for (int w = 0; w < 4; ++w) *reinterpret_cast<uint4*>(smem + WG_OFF_DZ + w * 8192 + off) = route_chunk(d, ag, w);
        }
      }
      fence_async_proxy();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&op_full);
        mbar_arrive(&ld_empty[rb]);
      }
    }
  }
  pdl_trigger();
  // ------------------------------------------------------------------ epilogue: one partial per CTA
  if (ntiles > 0) mbar_wait(&op_empty, (ntiles - 1) & 1);      // every MMA has completed (all threads wait: the operand area is reused)
  tc_fence_after();
  __syncthreads();
  float* out = a.partial + (size_t)blockIdx.x * (257 * 32);
  float* stage = reinterpret_cast<float*>(smem + WG_OFF_IMG);  // [2 dx][128 rows][32] fp32, 16-byte chunks XOR-ed with the row
  float* csum = reinterpret_cast<float*>(smem + WG_OFF_IMG + 32768);   // [4 cq][8 e][121 m] bias-gradient terms
  if (warp >= WG_BUILD_WARP0) {
    const int bt = tid - WG_BUILD_WARP0 * 32;
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int item = bt + q * WG_BUILDERS;
      if (item < NWIN * 4) {
#pragma unroll
        for (int e = 0; e < 8; ++e) csum[((item & 3) * 8 + e) * NWIN + (item >> 2)] = colsum[q][e];
      }
    }
  }
  const int q4 = warp & 3, dxe = (warp >> 2) & 1, row = 32 * q4 + lane;             // warps 0..7: accumulators (., dx = warp>>2)
  float acc0[32];
  if (warp < 8 && ntiles > 0) {
    float acc1[32];
    tmem_ld32(tmem + ((uint32_t)(32 * q4) << 16) + (2 + dxe) * 32, acc1);            // accumulator (wy = 1, dx): halves swapped
    const int prow = row ^ 64;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(stage + (dxe * 128 + prow) * 32 + ((j ^ (prow & 7)) << 2)) =
          make_float4(acc1[4 * j], acc1[4 * j + 1], acc1[4 * j + 2], acc1[4 * j + 3]);
    tmem_ld32(tmem + ((uint32_t)(32 * q4) << 16) + dxe * 32, acc0);                  // accumulator (wy = 0, dx)
  }
  __syncthreads();
  if (warp < 8) {
    const int dy = row >> 6, c64 = row & 63;
    float* dst = out + (size_t)kref(dy, dxe, c64) * 32;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ntiles > 0) {
        const float4 o = *reinterpret_cast<const float4*>(stage + (dxe * 128 + row) * 32 + ((j ^ (row & 7)) << 2));
        v = make_float4(acc0[4 * j] + o.x, acc0[4 * j + 1] + o.y, acc0[4 * j + 2] + o.z, acc0[4 * j + 3] + o.w);
      }
      *reinterpret_cast<float4*>(dst + 4 * j) = v;
    }
  } else if (warp == 8) {
    // bias gradient: channel n = lane, fixed-order sum over the 121 windows (and over the samples, in the registers above)
    float s = 0.f;
    if (ntiles > 0)
      for (int m = 0; m < NWIN; ++m) s += csum[lane * NWIN + m];
    out[256 * 32 + lane] = s;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

}  // namespace s2d
}  // namespace umma
}  // namespace dqn
