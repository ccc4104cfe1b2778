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
#include <cuda_fp16.h>

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
constexpr int NWARPS = 16;                           // weight-gradient kernel (the forward kernel has FWD_WARPS)

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
  // idx_cursor != null: idx is a [idx_rows][idx_k] table of pre-drawn index sets and the update uses row
  // (*idx_cursor % idx_rows); the optimizer advances the cursor at the end of the update (dqn_train_step_table)
  const int64_t* idx_cursor;
  int64_t idx_rows;
  int32_t idx_k;
  int64_t capacity, frame_capacity;
  int32_t pitch;
  // scalars of the sampled transitions, written by pass 0 in ring mode (null = not wanted)
  int32_t* actions;
  double* rewards;
  uint8_t* terminals;
};

struct RingRec { int64_t state_pos, next_pos; double reward; int32_t action; uint8_t terminal; uint8_t pad[3]; };

// Ring mode: where the 4 frames of sample b's stack (pass 0: state, pass 1: next state) start in the frame ring.  The
// look-up is a chain of dependent global loads (index -> record -> positions), so a whole warp resolves up to 64 tiles
// at once into a shared-memory table ahead of the copy loop.  Pass 0 also copies the transition's scalars.
__device__ __forceinline__ void resolve_stack(const Src& s, int pass, int b, int64_t off[4]) {
  const int32_t* ip = s.idx;
  if (s.idx_cursor) ip += (size_t)(*s.idx_cursor % s.idx_rows) * (size_t)s.idx_k;
  int64_t slot = s.ring_state[0] + (int64_t)ip[b];
  if (slot >= s.capacity) slot -= s.capacity;
  const RingRec* r = static_cast<const RingRec*>(s.meta) + slot;
  const int64_t pos = pass ? r->next_pos : r->state_pos;
#pragma unroll
  for (int f = 0; f < 4; ++f) off[f] = ((pos - 3 + f) % s.frame_capacity) * (int64_t)s.pitch;
  if (pass == 0) {
    if (s.actions) s.actions[b] = r->action;
    if (s.rewards) s.rewards[b] = r->reward;
    if (s.terminals) s.terminals[b] = r->terminal;
  }
}
constexpr int TAB_TILES = 64;

// one elected thread: start the copy of a stack into `dst`; completion (STACK_BYTES) is counted on `bar`
__device__ __forceinline__ void load_stack(const Src& s, int pass, int b, const int64_t* tab_entry, uint32_t dst, uint64_t* bar) {
  mbar_expect_tx(bar, STACK_BYTES);
  const uint8_t* x = pass ? s.x[1] : s.x[0];
  if (x) {
    bulk_g2s(dst, x + (size_t)b * STACK_BYTES, STACK_BYTES, bar);
    return;
  }
#pragma unroll
  for (int f = 0; f < 4; ++f) bulk_g2s(dst + f * FRAME_BYTES, s.frames + tab_entry[f], FRAME_BYTES, bar);
}

// The producer warp's loop: every lane resolves tiles of the next table round, then one elected lane issues the copies.
// slot_of(i) / wait_empty(i) are supplied by the kernel (raw ring slot and its "empty" wait of tile i).
template <class FDst, class FWait, class FBar, class FPost>
__device__ __forceinline__ void producer_loop(const Src& s, int pass, int b0, int bstep, int ntiles, int64_t (*tab)[4], FDst dst_of,
                                              FWait wait_empty, FBar bar_of, FPost after_issue) {
  const int lane = threadIdx.x & 31;
  const bool ring = (pass ? s.x[1] : s.x[0]) == nullptr;
  for (int base = 0; base < ntiles; base += TAB_TILES) {
    const int n = min(TAB_TILES, ntiles - base);
    if (ring) {
      for (int t = lane; t < n; t += 32) resolve_stack(s, pass, b0 + (base + t) * bstep, tab[t]);
    }
    __syncwarp();
    if (elect_one()) {
      for (int t = 0; t < n; ++t) {
        const int i = base + t;
        wait_empty(i);
        load_stack(s, pass, b0 + i * bstep, tab[t], dst_of(i), bar_of(i));
        after_issue(i);
      }
    }
    __syncwarp();
  }
}

// 4 pixels of frames a,b,c,d (one 32-bit word each) -> 16 bytes in HWC order (pixel-major, frame-minor)
__device__ __forceinline__ uint4 interleave4(uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  const uint32_t ab_lo = __byte_perm(a, b, 0x5140), ab_hi = __byte_perm(a, b, 0x7362);
  const uint32_t cd_lo = __byte_perm(c, d, 0x5140), cd_hi = __byte_perm(c, d, 0x7362);
  return make_uint4(__byte_perm(ab_lo, cd_lo, 0x5410), __byte_perm(ab_lo, cd_lo, 0x7632), __byte_perm(ab_hi, cd_hi, 0x5410),
                    __byte_perm(ab_hi, cd_hi, 0x7632));
}

// uint8 -> f16 without the (slow) integer conversion instruction: byte n next to the byte 0x64 is the f16 number
// 1024 + n (0x6400 is 1024.0 and its ulp is 1); one packed subtraction leaves n, exactly.  One byte permute and one
// HSUB2 per PAIR of values.
__device__ __forceinline__ uint32_t f16x2_from_bytes(uint32_t w, uint32_t sel) {
  const uint32_t biased = __byte_perm(w, 0x64646464u, sel);
  const __half2 d = __hsub2(*reinterpret_cast<const __half2*>(&biased), __half2half2(__ushort_as_half((unsigned short)0x6400)));
  return *reinterpret_cast<const uint32_t*>(&d);
}
__device__ __forceinline__ uint4 f16x8_from_u8(uint32_t w0, uint32_t w1) {        // 2 pixels x 4 frames, HWC order
  return make_uint4(f16x2_from_bytes(w0, 0x4140), f16x2_from_bytes(w0, 0x4342), f16x2_from_bytes(w1, 0x4140),
                    f16x2_from_bytes(w1, 0x4342));
}

constexpr int BUILD_THREADS = 336;      // 21 source columns x 16 row phases

// Builder thread bt (< BUILD_THREADS) owns source column j = bt % 21 (16 bytes: 4 pixels x 4 frames) of the rows
// r0 + 16k: the position inside the super-pixel (yy, chunk) and the plane are then fixed per thread and the image row
// advances by 22 per step, so the address arithmetic is a handful of instructions per 16 source bytes.  PLANAR: the raw
// slot holds 4 frames back to back (ring mode) instead of the HWC stack.  F16: the image holds IEEE halves (forward:
// f16 x f16 MMAs), else bf16 (weight gradient: the other operand is a bf16 gradient); exact either way.
template <bool PLANAR, bool F16>
__device__ __forceinline__ void build_image(const uint8_t* raw, uint8_t* img, int bt) {
  if (bt >= BUILD_THREADS) return;
  const int r0 = bt / 21, j = bt - r0 * 21;
  const int yy = (r0 + 2) & 3, Y0 = (r0 + 2) >> 2;
  const int rbase = (Y0 & 1) * PLANE_Y + (Y0 >> 1) * PW;
  const int ra0 = rbase + (j & 1) * PLANE_X + (j >> 1);               // super-pixel X = j:   pixels xx = 2,3 -> chunk 2*yy+1
  const int rb0 = rbase + ((j + 1) & 1) * PLANE_X + ((j + 1) >> 1);   // super-pixel X = j+1: pixels xx = 0,1 -> chunk 2*yy
  const int u0 = r0 * 21 + j;
  uint4 v[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    if (k < 5 || r0 < 4) {
      if constexpr (PLANAR) {
        const uint32_t* p = reinterpret_cast<const uint32_t*>(raw) + u0 + 336 * k;
        v[k] = make_uint4(p[0], p[FRAME_BYTES / 4], p[2 * (FRAME_BYTES / 4)], p[3 * (FRAME_BYTES / 4)]);
      } else {
        v[k] = reinterpret_cast<const uint4*>(raw)[u0 + 336 * k];
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    if (k < 5 || r0 < 4) {
      const int ra = ra0 + 22 * k, rb = rb0 + 22 * k;
      uint4 lo, hi;
      if constexpr (PLANAR && F16) {
        // v = 4 pixels of frames a, b, c, d: pair the frames first, then convert each pair
        const uint32_t ab01 = __byte_perm(v[k].x, v[k].y, 0x5140), ab23 = __byte_perm(v[k].x, v[k].y, 0x7362);
        const uint32_t cd01 = __byte_perm(v[k].z, v[k].w, 0x5140), cd23 = __byte_perm(v[k].z, v[k].w, 0x7362);
        lo = make_uint4(f16x2_from_bytes(ab01, 0x4140), f16x2_from_bytes(cd01, 0x4140), f16x2_from_bytes(ab01, 0x4342),
                        f16x2_from_bytes(cd01, 0x4342));
        hi = make_uint4(f16x2_from_bytes(ab23, 0x4140), f16x2_from_bytes(cd23, 0x4140), f16x2_from_bytes(ab23, 0x4342),
                        f16x2_from_bytes(cd23, 0x4342));
      } else {
        const uint4 w = PLANAR ? interleave4(v[k].x, v[k].y, v[k].z, v[k].w) : v[k];
        if constexpr (F16) {
          lo = f16x8_from_u8(w.x, w.y);
          hi = f16x8_from_u8(w.z, w.w);
        } else {
          lo = widen_u8x8(make_uint2(w.x, w.y));
          hi = widen_u8x8(make_uint2(w.z, w.w));
        }
      }
      *reinterpret_cast<uint4*>(img + ra * 128 + (((2 * yy + 1) ^ (ra & 7)) << 4)) = lo;
      *reinterpret_cast<uint4*>(img + rb * 128 + (((2 * yy) ^ (rb & 7)) << 4)) = hi;
    }
  }
}

// ------------------------------------------------------------------------------------------------ forward
struct FwdArgs {
  Src src;
  const __nv_bfloat16* wt[2];   // k-permuted F16 weight images (kperm): 4 taps x [32 rows][64] SWIZZLE_128B
  const float* bias[2];
  __nv_bfloat16* out[2];        // pooled activation [B][121][32]
  uint8_t* arg[2];              // pool arg-max [B][121][32] (may be null)
  int B, npass;
  unsigned long long* trace;    // debugging (DQN_S2D_TRACE=1): CTA 0 stamps clock64() per tile and pipeline event; else null
};
__device__ __forceinline__ void trace_stamp(unsigned long long* trace, int tile, int ev) {
  if (trace && blockIdx.x == 0 && tile < 60) trace[tile * 8 + ev] = (unsigned long long)clock64();
}
// per-CTA wall-clock stamps (ns): trace[480 + 4*cta + ev], ev 0 = entry, 1 = set-up done, 2 = past the PDL wait, 3 = exit
__device__ __forceinline__ void trace_cta(unsigned long long* trace, int ev) {
  if (trace && blockIdx.x < 160) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[480 + 4 * blockIdx.x + ev] = t;
  }
}

constexpr uint32_t FWD_OFF_IMG = 2 * RAW_BYTES, FWD_OFF_W = FWD_OFF_IMG + 2 * IMG_BYTES;
constexpr int FWD_W_BLOCKS = 8;                        // the four 4 KB tap blocks of the weight image, staged in two orders
constexpr uint32_t FWD_SMEM = SMEM_SLACK + FWD_OFF_W + FWD_W_BLOCKS * 4096;        // 222,208 bytes
constexpr int FWD_EPI_WARPS = 8;                       // two per TMEM lane quarter: 16 of the 32 channels each
constexpr int FWD_BUILD_WARP0 = 2 + FWD_EPI_WARPS, FWD_BUILD_WARPS = 11, FWD_WARPS = FWD_BUILD_WARP0 + FWD_BUILD_WARPS;   // 21 warps
static_assert(FWD_BUILD_WARPS * 32 >= BUILD_THREADS, "not enough builder threads");

// instruction descriptor with the operand formats chosen separately (0 = f16, 1 = bf16), fp32 accumulation
__host__ __device__ constexpr uint32_t make_idesc_ab(int M, int N, bool a_mn_major, bool b_mn_major, bool a_bf16, bool b_bf16) {
  return (1u << 4) | ((a_bf16 ? 1u : 0u) << 7) | ((b_bf16 ? 1u : 0u) << 10) | ((a_mn_major ? 1u : 0u) << 15) |
         ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
template <bool PLANAR>
__global__ void __launch_bounds__(FWD_WARPS * 32, 1) conv1_s2d_fwd_kernel(FwdArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t raw_full[2], raw_empty[2], img_full[2], img_empty[2], acc_full[2], acc_empty[2], wbar;
  __shared__ uint32_t tmem_slot;
  __shared__ float bias_sm[32];
  __shared__ int64_t tab[TAB_TILES][4];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = warp_uniform_idx();
  // tile = (pass, sample); a CTA stays on one pass so that it needs one set of weights
  const int nper = a.npass == 2 ? (int)gridDim.x >> 1 : (int)gridDim.x;          // CTAs per pass
  const int pass = a.npass == 2 ? (int)(blockIdx.x & 1) : 0;
  const int b0 = a.npass == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int ntiles = b0 < a.B ? (a.B - b0 + nper - 1) / nper : 0;
  if (tid == 0) trace_cta(a.trace, 0);
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      mbar_init(&raw_full[i], 1);
      mbar_init(&raw_empty[i], FWD_BUILD_WARPS);
      mbar_init(&img_full[i], FWD_BUILD_WARPS);
      mbar_init(&img_empty[i], 1);
      mbar_init(&acc_full[i], 1);
      mbar_init(&acc_empty[i], FWD_EPI_WARPS);
    }
    mbar_init(&wbar, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 256);                 // two sets of four 128x32 fp32 accumulators
  // the padding of the images (border super-pixels, rows between the planes) is never written again: zero it once
  for (int i = tid; i < (int)(2 * IMG_BYTES / 16); i += FWD_WARPS * 32)
    reinterpret_cast<uint4*>(smem + FWD_OFF_IMG)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_async_proxy();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t sbase = smem_u32(smem);
  if (tid == 0) trace_cta(a.trace, 1);
  pdl_wait();                                                 // everything above overlapped the previous kernel
  // the grid never exceeds one CTA per SM and a CTA owns its SM's shared memory, so the successor's CTAs can only
  // start where one of ours has exited: let it launch now, its launch latency and set-up hide under our tail
  pdl_trigger();
  if (tid == 0) trace_cta(a.trace, 2);
  if (warp == 0) {
    // ---------------------------------------------------------------- producer
    if (elect_one()) {
      // tap blocks (4 KB each, tap d = 2*dy + dx in the global image) in the orders of the merged B operands, see the MMA issuer
      const __nv_bfloat16* wsrc = pass ? a.wt[1] : a.wt[0];
      constexpr int order[FWD_W_BLOCKS] = {3, 2, 0, 1, 2, 3, 1, 0};
      mbar_expect_tx(&wbar, FWD_W_BLOCKS * 4096);
#pragma unroll
      for (int k = 0; k < FWD_W_BLOCKS; ++k) bulk_g2s(sbase + FWD_OFF_W + k * 4096, wsrc + order[k] * 2048, 4096, &wbar);
    }
    __syncwarp();
    producer_loop(
        a.src, pass, b0, nper, ntiles, tab, [&](int i) { return sbase + (i & 1) * RAW_BYTES; },
        [&](int i) { mbar_wait(&raw_empty[i & 1], ((i >> 1) & 1) ^ 1); trace_stamp(a.trace, i, 0); },
        [&](int i) { return &raw_full[i & 1]; }, [](int) {});
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer (f16 image x f16 weights: the two
    // operands of a kind::f16 MMA must share one format -- f16 x bf16 is an illegal instruction)
    // An MMA of this shape is bound by its shared-memory operand reads (4 KB image slice + 1 KB weights per 128x32x16 at
    // 128 B/clk = 40 cycles for 16 cycles of math), so MMAs that read the SAME image slice are merged into wider ones.  The
    // slice at super-pixel offset (Y, X) = (wy+dy, wx+dx) serves every (window position, tap) pair with that sum: the
    // centre (1,1) all four positions, an edge two, a corner one.  With the accumulators in the column order
    // [w00, w01, w11, w10] the centre is ONE N=128 MMA per k-step and three of the four edges are N=64 (the positions of
    // edge (1,0), w00 and w10, are not neighbours): 40 MMAs instead of 64, 224 KB instead of 320 KB of operand reads per
    // sample.  The weights are staged twice (FWD_W_BLOCKS), tap blocks in the orders the merged B operands need:
    //   blocks 0..3 = taps [11, 10, 00, 01] (centre; single taps), blocks 4..7 = taps [10, 11, 01, 00] (edge pairs).
    constexpr uint32_t idesc32 = make_idesc_ab(TM, 32, false, false, false, false);
    constexpr uint32_t idesc64 = make_idesc_ab(TM, 64, false, false, false, false);
    constexpr uint32_t idesc128 = make_idesc_ab(TM, 128, false, false, false, false);
    mbar_wait(&wbar, 0);
    for (int i = 0; i < ntiles; ++i) {
      const int ib = i & 1, ph = (i >> 1) & 1;
      mbar_wait(&img_full[ib], ph);
      mbar_wait(&acc_empty[ib], ph ^ 1);
      tc_fence_after();
      if (elect_one()) {
        trace_stamp(a.trace, i, 4);
        const uint32_t img = sbase + FWD_OFF_IMG + ib * IMG_BYTES;
        const uint64_t bd0 = make_desc_sw128(sbase + FWD_OFF_W);
        const uint32_t acc = tmem + ib * 128;
        // (image slice Y, X; first weight block; first accumulator column; N)
        auto run = [&](int Y, int X, int wblock, int col, uint32_t idesc, bool first) {
          const uint64_t ad0 = make_desc_sw128(img + (uint32_t)img_row(Y, X) * 128u);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            mma_bf16(acc + col, desc_advance(ad0, j * 32), desc_advance(bd0, wblock * 4096 + j * 32), idesc, (first && j == 0) ? 0u : 1u);
        };
        run(1, 1, 0, 0, idesc128, true);       // centre: w00<-11, w01<-10, w11<-00, w10<-01 (first write of every column)
        run(0, 1, 6, 0, idesc64, false);       // edge: w00<-01, w01<-00
        run(1, 2, 5, 32, idesc64, false);      // edge: w01<-11, w11<-01
        run(2, 1, 4, 64, idesc64, false);      // edge: w11<-10, w10<-11
        run(1, 0, 1, 0, idesc32, false);       // edge (1,0): w00<-10 ...
        run(1, 0, 2, 96, idesc32, false);      //             ... and w10<-00
        run(0, 0, 2, 0, idesc32, false);       // corners: w00<-00
        run(0, 2, 3, 32, idesc32, false);      //          w01<-01
        run(2, 0, 1, 96, idesc32, false);      //          w10<-10
        run(2, 2, 0, 64, idesc32, false);      //          w11<-11
        mma_commit(&img_empty[ib]);
        mma_commit(&acc_full[ib]);
        trace_stamp(a.trace, i, 5);
      }
      __syncwarp();
    }
  } else if (warp < FWD_BUILD_WARP0) {
    // ---------------------------------------------------------------- epilogue: TMEM lane quarter = warp & 3, channel
    // half h = (warp - 2) >> 2 (16 of the 32 output channels)
    if (warp == 2) bias_sm[lane] = __ldg((pass ? a.bias[1] : a.bias[0]) + lane);
    asm volatile("bar.sync 1, %0;" ::"n"(FWD_EPI_WARPS * 32) : "memory");            // the epilogue warps
    const int q = warp & 3, h = (warp - 2) >> 2, m = 32 * q + lane;
    const int py = m / PW, px = m - py * PW;
    const bool valid = m < NWIN;
    float bias[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) bias[j] = bias_sm[h * 16 + j];
    __nv_bfloat16* outp = pass ? a.out[1] : a.out[0];
    uint8_t* argp = pass ? a.arg[1] : a.arg[0];
    for (int i = 0; i < ntiles; ++i) {
      const int ib = i & 1, ph = (i >> 1) & 1;
      const int b = b0 + i * nper;
      mbar_wait(&acc_full[ib], ph);
      tc_fence_after();
      if (tid == 64) trace_stamp(a.trace, i, 6);
      // A value's float bits, taken as a SIGNED integer, order like the value for values >= 0 and put every negative
      // value below them; with the 2 lowest mantissa bits replaced by 3-w the maximum over the window positions is the
      // largest value (first position on ties) whenever the pooled ReLU output is positive, and negative otherwise.
      // Two window positions per round of TMEM loads: all four at once need more registers than a 21-warp CTA has per
      // thread (80), and the spills ran through the same L1/shared-memory path the MMAs and the builders saturate.
      int key[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) key[j] = (int)0x80000000;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        uint32_t r[2][16];
#pragma unroll
        for (int u = 0; u < 2; ++u) {        // accumulator column block of window position w: [w00, w01, w11, w10]
          const int w = 2 * half + u;
          tmem_ld16_nowait(tmem + ((uint32_t)(32 * q) << 16) + ib * 128 + (w ^ (w >> 1)) * 32 + h * 16, r[u]);
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (half == 1) {                     // the TMEM set goes back to the MMA warp
          if (tid == 64) trace_stamp(a.trace, i, 2);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[ib]);
        }
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int w = 2 * half + u;
          // window position (wy, wx) lies inside the 21x21 conv output unless it is the pool's padding row / column
          const bool in = ((w >> 1) == 0 || py < PW - 1) && ((w & 1) == 0 || px < PW - 1);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int kj = (__float_as_int(__uint_as_float(r[u][j]) + bias[j]) & ~3) | (3 - w);
            key[j] = in ? max(key[j], kj) : key[j];
          }
        }
      }
      if (valid) {
        const size_t off = ((size_t)b * NWIN + m) * 32 + h * 16;
        uint32_t pk[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
          pk[j] = pack_bf16(__int_as_float(max(key[2 * j], 0) & ~3), __int_as_float(max(key[2 * j + 1], 0) & ~3));
        *reinterpret_cast<uint4*>(outp + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        *reinterpret_cast<uint4*>(outp + off + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
        if (argp) {
          uint32_t ak[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            ak[j] = (3u - ((uint32_t)key[4 * j] & 3u)) | ((3u - ((uint32_t)key[4 * j + 1] & 3u)) << 8) |
                    ((3u - ((uint32_t)key[4 * j + 2] & 3u)) << 16) | ((3u - ((uint32_t)key[4 * j + 3] & 3u)) << 24);
          *reinterpret_cast<uint4*>(argp + off) = make_uint4(ak[0], ak[1], ak[2], ak[3]);
        }
      }
      if (tid == 64) trace_stamp(a.trace, i, 7);
    }
  } else {
    // ---------------------------------------------------------------- builders
    const int bt = tid - FWD_BUILD_WARP0 * 32;
    for (int i = 0; i < ntiles; ++i) {
      const int rb = i & 1, ph = (i >> 1) & 1;
      mbar_wait(&raw_full[rb], ph);
      if (bt == 0) trace_stamp(a.trace, i, 1);
      mbar_wait(&img_empty[rb], ph ^ 1);
      build_image<PLANAR, true>(smem + rb * RAW_BYTES, smem + FWD_OFF_IMG + rb * IMG_BYTES, bt);       // exact f16 image
      fence_async_proxy();
      __syncwarp();
      if (bt == 0) trace_stamp(a.trace, i, 3);
      if (lane == 0) {
        mbar_arrive(&img_full[rb]);
        mbar_arrive(&raw_empty[rb]);
      }
    }
  }
  pdl_trigger();
  // exit stamp: the latest warp's arrival (a stamp taken right AFTER a __syncthreads records the barrier's issue, not
  // its release: BAR.SYNC.DEFER_BLOCKING lets the warp run on until it touches barrier-protected state)
  if (a.trace && lane == 0 && blockIdx.x < 160) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    atomicMax(a.trace + 480 + 4 * blockIdx.x + 3, t);
  }
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
// separately: four accumulators (wy, dx) in the column order [(0,1), (0,0), (1,1), (1,0)]; the epilogue adds accumulator
// (1, dx) to (0, dx) with its halves swapped.
struct WgradArgs {
  Src src;                       // pass 0 only
  const __nv_bfloat16* dactg;    // [B][121][32] ReLU-gated gradient of the pooled conv1 output
  const uint8_t* arg;            // [B][121][32]
  float* partial;                // [gridDim.x][(256+1)*32] fp32, reference index order
  int B;
  unsigned long long* trace;     // debugging (DQN_S2D_TRACE=1): CTA 0 stamps clock64() per tile at trace[1136 + 8*tile + ev]
};
__device__ __forceinline__ void wg_stamp(unsigned long long* trace, int tile, int ev) {
  if (trace && blockIdx.x == 0 && tile < 8) trace[1136 + tile * 8 + ev] = (unsigned long long)clock64();
}

constexpr uint32_t WG_GRAD_BYTES = NWIN * 32 * 2, WG_ARG_BYTES = NWIN * 32;          // 7744, 3872 per sample
constexpr uint32_t WG_GA_SLOT = 12 * 1024;                                           // gradient (8 KB) + arg-max (4 KB) staging
constexpr uint32_t WG_DZ_SET = 4 * 8192;                                             // the four routed gradient tiles of a sample
constexpr uint32_t WG_OFF_IMG = 2 * RAW_BYTES, WG_OFF_DZ = WG_OFF_IMG + IMG_BYTES, WG_OFF_GA = WG_OFF_DZ + 2 * WG_DZ_SET;
constexpr uint32_t WG_SMEM = SMEM_SLACK + WG_OFF_GA + 2 * WG_GA_SLOT;                // 214,016 bytes
constexpr int WG_BUILD_WARP0 = 2, WG_BUILD_WARPS = NWARPS - WG_BUILD_WARP0, WG_BUILDERS = WG_BUILD_WARPS * 32;

template <bool PLANAR>
__global__ void __launch_bounds__(NWARPS * 32, 1) conv1_s2d_wgrad_kernel(WgradArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t stk_full[2], ga_full[2], ld_empty[2], op_full, op_empty;   // stack landed / gradient + arg-max landed / both consumed
  __shared__ uint32_t tmem_slot;
  __shared__ int64_t tab[TAB_TILES][4];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = warp_uniform_idx();
  const int G = gridDim.x, b0 = blockIdx.x;
  const int ntiles = b0 < a.B ? (a.B - b0 + G - 1) / G : 0;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 2; ++i) { mbar_init(&stk_full[i], 1); mbar_init(&ga_full[i], 1); mbar_init(&ld_empty[i], WG_BUILD_WARPS); }
    mbar_init(&op_full, WG_BUILD_WARPS);
    mbar_init(&op_empty, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_slot, 128);                 // four 128x32 fp32 accumulators
  // image padding and the unused dZ rows (m >= 121) stay zero for the whole kernel (the reduction runs over them)
  for (int i = tid; i < (int)((IMG_BYTES + 2 * WG_DZ_SET) / 16); i += NWARPS * 32)
    reinterpret_cast<uint4*>(smem + WG_OFF_IMG)[i] = make_uint4(0u, 0u, 0u, 0u);
  fence_async_proxy();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const uint32_t sbase = smem_u32(smem);
  // Only the output gradient and the arg-max bytes come from the predecessor in the stream (conv2's data gradient); the
  // frames were in place before the step's first kernel finished (every kernel of the chain waits for its predecessor
  // before it lets its successor launch).  So the stack of the first sample is fetched and its image built BEFORE the
  // PDL wait, in the shadow of the predecessor; only the producer thread waits, right before it touches dactg / arg.
  pdl_trigger();                                              // see the forward kernel
  float colsum[2][8];                                          // bias gradient: this thread's (m, channel chunk) items
#pragma unroll
  for (int q = 0; q < 2; ++q)
#pragma unroll
    for (int e = 0; e < 8; ++e) colsum[q][e] = 0.f;
  if (warp == 0) {
    // ---------------------------------------------------------------- producer: stack + gradient + arg-max of a sample
    producer_loop(
        a.src, 0, b0, G, ntiles, tab, [&](int i) { return sbase + (i & 1) * RAW_BYTES; },
        [&](int i) { mbar_wait(&ld_empty[i & 1], ((i >> 1) & 1) ^ 1); },
        [&](int i) { return &stk_full[i & 1]; },
        [&](int i) {
          const int rb = i & 1, b = b0 + i * G;
          if (i == 0) pdl_wait();
          mbar_expect_tx(&ga_full[rb], WG_GRAD_BYTES + WG_ARG_BYTES);
          bulk_g2s(sbase + WG_OFF_GA + rb * WG_GA_SLOT, a.dactg + (size_t)b * NWIN * 32, WG_GRAD_BYTES, &ga_full[rb]);
          bulk_g2s(sbase + WG_OFF_GA + rb * WG_GA_SLOT + 8192, a.arg + (size_t)b * NWIN * 32, WG_ARG_BYTES, &ga_full[rb]);
        });
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    // MMAs that read the same image slice are merged (see the forward kernel): for X = wx + dx = 1 the pairs (wx=0, dx=1)
    // and (wx=1, dx=0) share it, their B operands dZ(wy,0), dZ(wy,1) are neighbours in shared memory (one descriptor, N=64)
    // and so are their accumulators, which therefore sit in the order [(wy,dx=1), (wy,dx=0)]: 48 MMAs instead of 64.
    constexpr uint32_t idesc32 = make_idesc(TM, 32, true, true), idesc64 = make_idesc(TM, 64, true, true);
    for (int i = 0; i < ntiles; ++i) {
      mbar_wait(&op_full, i & 1);
      tc_fence_after();
      if (elect_one()) {
        wg_stamp(a.trace, i, 5);
        const uint32_t img = sbase + WG_OFF_IMG;
#pragma unroll
        for (int wy = 0; wy < 2; ++wy) {
          // (X; first dZ tile; first accumulator column block; N; first write of the columns)
          auto run = [&](int t, int wx, int col, uint32_t idesc, bool first) {
            // block 0 / block 1 of the M = 128 rows: wy = 0 -> taps dy = 0, 1 (rows +248); wy = 1 -> taps dy = 1, 0 (rows +237)
            const int r0 = wy ? img_row(2, t) : img_row(0, t);
            const int r1 = img_row(1, t);
            const uint64_t ad0 = make_desc_sw128_mn(img + (uint32_t)r0 * 128u, (uint32_t)(r1 - r0) * 128u, 1024);
            const uint64_t bd0 = make_desc(sbase + WG_OFF_DZ + (i & 1) * WG_DZ_SET + (wy * 2 + wx) * 8192, 128, 2048);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              mma_bf16(tmem + wy * 64 + col * 32, desc_advance(ad0, j * 2048), desc_advance(bd0, j * 256), idesc,
                       (first && i == 0 && j == 0) ? 0u : 1u);
          };
          run(1, 0, 0, idesc64, true);       // X = 1: (wx=0 -> dx=1, column block 0), (wx=1 -> dx=0, column block 1)
          run(0, 0, 1, idesc32, false);      // X = 0: wx=0, dx=0
          run(2, 1, 0, idesc32, false);      // X = 2: wx=1, dx=1
        }
        mma_commit(&op_empty);
        wg_stamp(a.trace, i, 6);
      }
      __syncwarp();
    }
  } else {
    // ---------------------------------------------------------------- builders: image + routed gradient tiles
    const int bt = tid - WG_BUILD_WARP0 * 32;
    // The routed gradient tiles are double-buffered: those of sample i are written while the MMAs of sample i-1 still
    // run (their set was last read by sample i-2, whose completion this thread saw before it built image i-1); only the
    // image, 64 KB, is single and waits for the MMAs.  Sample 0 goes the other way round: its image can be built before
    // the gradient has landed (before the predecessor kernel has finished, even).
    auto route = [&](int i) {
      const int rb = i & 1;
      mbar_wait(&ga_full[rb], (i >> 1) & 1);
      if (bt == 0) wg_stamp(a.trace, i, 3);
      const uint8_t* ga = smem + WG_OFF_GA + rb * WG_GA_SLOT;
      uint8_t* dz = smem + WG_OFF_DZ + rb * WG_DZ_SET;
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
          for (int w = 0; w < 4; ++w) *reinterpret_cast<uint4*>(dz + w * 8192 + off) = route_chunk(d, ag, w);
        }
      }
      if (bt == 0) wg_stamp(a.trace, i, 4);
    };
    for (int i = 0; i < ntiles; ++i) {
      const int rb = i & 1;
      if (i > 0) route(i);
      mbar_wait(&stk_full[rb], (i >> 1) & 1);
      if (bt == 0) wg_stamp(a.trace, i, 0);
      if (i > 0) mbar_wait(&op_empty, (i - 1) & 1);             // the previous sample's MMAs have read the image
      if (bt == 0) wg_stamp(a.trace, i, 1);
      build_image<PLANAR, false>(smem + rb * RAW_BYTES, smem + WG_OFF_IMG, bt);
      if (bt == 0) wg_stamp(a.trace, i, 2);
      if (i == 0) route(0);
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
  if (tid == 64) wg_stamp(a.trace, 0, 7);
  tc_fence_after();
  __syncthreads();
  if (tid == 64) wg_stamp(a.trace, 4, 7);
  float* out = a.partial + (size_t)blockIdx.x * (257 * 32);
  float* stage = reinterpret_cast<float*>(smem + WG_OFF_IMG);  // [2 dx][128 rows][32] fp32, 16-byte chunks XOR-ed with the row
  float* csum = reinterpret_cast<float*>(smem + WG_OFF_IMG + 32768);   // [builder warp][4 cq][8 e] bias-gradient partial sums
  if (warp >= WG_BUILD_WARP0) {
    // bias gradient: a thread's items are (m, channel chunk cq = lane & 3) for two m; sum over the 8 m of the warp with
    // shuffles (fixed tree), one partial per warp and channel; warp 8 adds the 14 partials in order
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float v = colsum[0][e] + colsum[1][e];
      v += __shfl_xor_sync(0xffffffffu, v, 4);
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if (lane < 4) csum[((warp - WG_BUILD_WARP0) * 4 + lane) * 8 + e] = v;
    }
  }
  const int q4 = warp & 3, dxe = (warp >> 2) & 1, row = 32 * q4 + lane;             // warps 0..7: accumulators (., dx = warp>>2)
  if (warp < 8 && ntiles > 0) {
    float acc1[32];
    tmem_ld32(tmem + ((uint32_t)(32 * q4) << 16) + (2 + (1 - dxe)) * 32, acc1);      // accumulator (wy = 1, dx): halves swapped
    const int prow = row ^ 64;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<float4*>(stage + (dxe * 128 + prow) * 32 + ((j ^ (prow & 7)) << 2)) =
          make_float4(acc1[4 * j], acc1[4 * j + 1], acc1[4 * j + 2], acc1[4 * j + 3]);
  }
  if (tid == 64) wg_stamp(a.trace, 2, 7);
  __syncthreads();
  if (warp < 8 && ntiles > 0) {
    // add accumulator (wy = 0, dx) to the row its partner left in the staging tile (in place: only this thread touches
    // this row now)
    float acc0[32];
    tmem_ld32(tmem + ((uint32_t)(32 * q4) << 16) + (1 - dxe) * 32, acc0);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4* p = reinterpret_cast<float4*>(stage + (dxe * 128 + row) * 32 + ((j ^ (row & 7)) << 2));
      const float4 o = *p;
      *p = make_float4(acc0[4 * j] + o.x, acc0[4 * j + 1] + o.y, acc0[4 * j + 2] + o.z, acc0[4 * j + 3] + o.w);
    }
  }
  if (tid == 64) wg_stamp(a.trace, 3, 7);
  __syncthreads();
  if (warp < 8) {
    // write-out: 8 lanes per row (16 bytes each), 4 rows per instruction; 4 consecutive channels are 4 consecutive rows of
    // the reference order, i.e. 512 contiguous bytes (instead of every thread writing its own row 16 bytes at a
    // time, 32 separate lines per instruction)
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int r = warp * 32 + it * 4 + (lane >> 3), ch = lane & 7;            // r = dx * 128 + dy * 64 + c64
      const int rdx = r >> 7, rdy = (r >> 6) & 1, c64 = r & 63;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (ntiles > 0) v = *reinterpret_cast<const float4*>(stage + r * 32 + ((ch ^ (r & 7)) << 2));
      *reinterpret_cast<float4*>(out + (size_t)kref(rdy, rdx, c64) * 32 + ch * 4) = v;
    }
  }
  if (warp == 8) {
    // bias gradient: channel n = lane, the builder warps' partial sums added in warp order
    float s = 0.f;
    if (ntiles > 0) {
#pragma unroll
      for (int w = 0; w < WG_BUILD_WARPS; ++w) s += csum[w * 32 + lane];
    }
    out[256 * 32 + lane] = s;
  }
  if (tid == 64) wg_stamp(a.trace, 1, 7);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 128);
}

}  // namespace s2d
}  // namespace umma
}  // namespace dqn
