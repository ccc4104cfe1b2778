// The middle of the training step of the reference ConvNetwork at small minibatches (B <= 32), as ONE launch of a
// thread-block cluster of 8 CTAs instead of three dependent launches (fc forward, TD head, fc data gradient):
//   a4 = relu(a3 w4 + b4) for both parameter sets (network.py:421-424)  ->  q = a4 w5 + b5 of both  ->  TD target, loss
//   term and dL/dq (dqn_agent.py:118-122, 224-253; network.py:141-157)  ->  dL/da4, ReLU-gated  ->  dL/da3 = (.) w4^T,
//   gated by conv3's ReLU.
// fc_fused.cuh does the same in one CTA and loses: that CTA has to pull all 384 KB of fc weights through one SM.  Here
// CTA r of the cluster owns hidden units [32r, 32r+32): it fetches 2 x 16 KB of forward weights and 16 KB of
// data-gradient weights (before the PDL wait), runs 32 + 16 tcgen05 MMAs (M = 128 sample rows of which <= 32 are used,
// N = 32), and the CTAs exchange only small vectors through distributed shared memory:
//   barrier 1: partial Q sums (this CTA's 32 hidden units)      -> CTA r sums them for samples 4r..4r+3 and does the TD step
//   barrier 2: dL/dq of those 4 samples                          -> every CTA builds its 32 columns of dL/da4
//   barrier 3: dL/da4 columns (through global memory: the fc weight gradient needs them there anyway)
//                                                                 -> every CTA reads all 256 columns as its A operand
#pragma once
#include "fc_fused.cuh"

namespace dqn {
namespace umma {
namespace fcc {

constexpr int CL = 8;                        // CTAs per cluster
constexpr int NS = FCF / CL;                 // 32 hidden units (fc outputs / data-gradient outputs) per CTA
constexpr int MAXB = 32;                     // sample rows used
constexpr uint32_t BLK = 4096;               // one 64-deep k-block of 32 rows (SWIZZLE_128B K-major)
constexpr uint32_t OFF_A = 0, OFF_W = 2 * 4 * BLK, OFF_WD = OFF_W + 2 * 4 * BLK, SMEM_END = OFF_WD + 4 * BLK;   // 80 KB
// the MMAs read 128 rows (16 KB) per k-block; rows >= 32 are whatever follows in shared memory (their accumulator rows
// are never looked at), so the allocation only has to reach 12 KB past the last A block, which the weights cover
constexpr uint32_t SMEM = SMEM_SLACK + SMEM_END;
constexpr int THREADS = 512;

__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// generic pointer to the same shared-memory variable in CTA `rank` of the cluster (plain loads through it are routed to
// that CTA's shared memory and can be in flight together, unlike a chain of volatile asm loads)
template <typename T>
__device__ __forceinline__ const T* dsmem_ptr(const T* local, uint32_t rank) {
  uint64_t out;
  asm volatile("mapa.u64 %0, %1, %2;" : "=l"(out) : "l"(reinterpret_cast<uint64_t>(local)), "r"(rank));
  return reinterpret_cast<const T*>(out);
}

// [rows][256] bf16 row-major in global memory -> four SWIZZLE_128B K-major k-blocks of 32 rows (rows >= nrows zero)
__device__ __forceinline__ void load_rows(uint32_t dst, const __nv_bfloat16* src, int nrows) {
  for (int ch = threadIdx.x; ch < MAXB * 32; ch += THREADS) {
    const int r = ch >> 5, cc = ch & 31;
    const bool ok = r < nrows;
    cp_async16(dst + (uint32_t)(cc >> 3) * BLK + sw128_off(r, cc & 7), src + (ok ? (size_t)r * FCF + cc * 8 : 0), ok);
  }
}

#define FCC_STAMP(i) do { if (a.trace && rank == 0 && tid == 0) a.trace[1080 + (i)] = (unsigned long long)clock64(); } while (0)
__global__ void __launch_bounds__(THREADS, 1) fc_td_cluster_kernel(FcFusedArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t bars[2];
  __shared__ uint64_t wbar, wdbar, fbar, dbar;   // forward weights / data-gradient weights landed; forward / data-gradient MMAs done
  __shared__ uint32_t tmem_slot;
  __shared__ float b4s[2][NS];
  __shared__ __align__(16) float w5s[2][NS * FC_AP];
  __shared__ float b5s[2][FC_AP];
  __shared__ __align__(16) float qpart[2][2][MAXB][FC_AP];   // [net][column half][sample][action]: partial head sums of this CTA
  __shared__ float g_own[4];                                 // dL/dq_selected of samples 4r..4r+3
  __shared__ float g_all[MAXB];
  __shared__ uint32_t gate[MAXB];                            // bit j: online a4[b][n0 + j] > 0
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = warp_uniform_idx();
  const uint32_t rank = cluster_rank();
  const int n0 = (int)rank * NS;
  if (tid == 0) {
    mbar_init(&wbar, 1);
    mbar_init(&wdbar, 1);
    mbar_init(&fbar, 1);
    mbar_init(&dbar, 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, 128);   // [0,32) online fc, [32,64) target fc, [64,96) data gradient
  const uint32_t sbase = smem_u32(smem);
  // parameters were last written by the previous step's optimizer, several launches back: fetched before the PDL wait.
  // Rows n0..n0+31 of a 64-deep stage image are one contiguous 4 KB run (umma::img_index).
  if (tid == 0) {
    mbar_expect_tx(&wbar, 2 * 4 * BLK);
#pragma unroll
    for (int net = 0; net < 2; ++net)
#pragma unroll
      for (int kb = 0; kb < 4; ++kb)
        bulk_g2s(sbase + OFF_W + (net * 4 + kb) * BLK, a.wt[net] + (size_t)kb * (FCF * 64) + (size_t)n0 * 64, BLK, &wbar);
    mbar_expect_tx(&wdbar, 4 * BLK);
#pragma unroll
    for (int kb = 0; kb < 4; ++kb) bulk_g2s(sbase + OFF_WD + kb * BLK, a.wd + (size_t)kb * (FCF * 64) + (size_t)n0 * 64, BLK, &wdbar);
  }
  for (int i = tid; i < 2 * NS; i += THREADS) b4s[i / NS][i % NS] = __ldg(a.b4[i / NS] + n0 + (i % NS));
  for (int i = tid; i < 2 * NS * FC_AP; i += THREADS) {
    const int net = i / (NS * FC_AP), j = (i / FC_AP) % NS, c = i % FC_AP;
    w5s[net][j * FC_AP + c] = c < a.A ? __ldg(a.w5[net] + (size_t)(n0 + j) * a.A + c) : 0.f;
  }
  if (tid < 2 * FC_AP) b5s[tid / FC_AP][tid % FC_AP] = (tid % FC_AP) < a.A ? __ldg(a.b5[tid / FC_AP] + (tid % FC_AP)) : 0.f;
  FCC_STAMP(0);
  pdl_wait();
  FCC_STAMP(1);
  pdl_trigger();                                  // 8 CTAs: the successor's CTAs fit beside us, let it run its set-up

  // ---------------------------------------------------------------- forward: both nets' activations -> A blocks
  load_rows(sbase + OFF_A, a.x[0], a.B);
  load_rows(sbase + OFF_A + 4 * BLK, a.x[1], a.B);
  cp_async_commit();
  cp_async_wait<0>();
  fence_async_proxy();
  __syncthreads();
  FCC_STAMP(2);
  constexpr uint32_t idesc = make_idesc(TM, NS, false, false);
  if (warp == 0) {
    mbar_wait(&wbar, 0);
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int net = 0; net < 2; ++net) {
#pragma unroll
        for (int kb = 0; kb < 4; ++kb) {
          const uint64_t ad0 = make_desc_sw128(sbase + OFF_A + (net * 4 + kb) * BLK);
          const uint64_t bd0 = make_desc_sw128(sbase + OFF_W + (net * 4 + kb) * BLK);
#pragma unroll
          for (int j = 0; j < TK / 16; ++j)
            mma_bf16(tmem + net * NS, desc_advance(ad0, j * 32), desc_advance(bd0, j * 32), idesc, (kb | j) ? 1u : 0u);
        }
      }
      mma_commit(&fbar);
    }
    __syncwarp();
  }
  // the TD step's inputs of samples 4r..4r+3 (warp 0, lane = (sample, action)): fetched now, used after barrier 1
  int td_act = 0;
  double td_reward = 0.0;
  bool td_terminal = false;
  if (warp == 0) {
    const int bb = 4 * (int)rank + (lane >> 3);
    if (bb < a.B) {
      td_act = a.actions[bb];
      td_reward = a.rewards[bb];
      td_terminal = a.terminals[bb] != 0;
    }
  }
  // ---------------------------------------------------------------- fc epilogue: warps 0, 4, 8, 12 own TMEM lanes 0..31
  // (= samples); warp w: net w >> 3, column half (w >> 2) & 1
  const int b = lane;                              // sample of this lane in the epilogue warps
  if ((warp & 3) == 0) {
    const int net = warp >> 3, half = (warp >> 2) & 1;
    mbar_wait(&fbar, 0);
    tc_fence_after();
    FCC_STAMP(3);
    float v[16];
    tmem_ld16(tmem + net * NS + half * 16, v);
    float qa[FC_AP];
#pragma unroll
    for (int c = 0; c < FC_AP; ++c) qa[c] = 0.f;
    uint32_t gbits = 0u;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int jj = half * 16 + j;
      v[j] = fmaxf(v[j] + b4s[net][jj], 0.f);
      gbits |= (v[j] > 0.f ? 1u : 0u) << jj;
      const float4 w0 = *reinterpret_cast<const float4*>(&w5s[net][jj * FC_AP]);
      const float4 w1 = *reinterpret_cast<const float4*>(&w5s[net][jj * FC_AP + 4]);
      qa[0] = fmaf(v[j], w0.x, qa[0]); qa[1] = fmaf(v[j], w0.y, qa[1]);
      qa[2] = fmaf(v[j], w0.z, qa[2]); qa[3] = fmaf(v[j], w0.w, qa[3]);
      qa[4] = fmaf(v[j], w1.x, qa[4]); qa[5] = fmaf(v[j], w1.y, qa[5]);
      qa[6] = fmaf(v[j], w1.z, qa[6]); qa[7] = fmaf(v[j], w1.w, qa[7]);
    }
    *reinterpret_cast<float4*>(&qpart[net][half][b][0]) = make_float4(qa[0], qa[1], qa[2], qa[3]);
    *reinterpret_cast<float4*>(&qpart[net][half][b][4]) = make_float4(qa[4], qa[5], qa[6], qa[7]);
    if (net == 0) {
      if (half == 0) gate[b] = gbits;             // half 1 ORs its bits in after the barrier below
      if (b < a.B) {
        float* dst = a.a4 + (size_t)b * FCF + n0 + half * 16;
#pragma unroll
        for (int j = 0; j < 4; ++j) *reinterpret_cast<float4*>(dst + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
      }
    }
    __syncwarp();
    asm volatile("bar.sync 2, 128;" ::: "memory");      // the four epilogue warps: gate[] of half 0 is written
    if (net == 0 && half == 1) gate[b] |= gbits;
  }
  tc_fence_before();
  __syncthreads();
  FCC_STAMP(4);
  cluster_sync_all();                               // barrier 1: every CTA's qpart is complete
  FCC_STAMP(5);

  // ---------------------------------------------------------------- Q values + TD step for samples 4r .. 4r+3 (warp 0)
  if (warp == 0) {
    const int s = lane >> 3, c = lane & 7;          // sample 4r+s, action c
    const int bb = 4 * (int)rank + s;
    const bool valid = bb < a.B;
    float p[CL][4];                                 // all 32 remote loads first (one DSMEM latency), then the sums
#pragma unroll
    for (uint32_t r = 0; r < (uint32_t)CL; ++r) {
      p[r][0] = *dsmem_ptr(&qpart[0][0][bb][c], r);
      p[r][1] = *dsmem_ptr(&qpart[0][1][bb][c], r);
      p[r][2] = *dsmem_ptr(&qpart[1][0][bb][c], r);
      p[r][3] = *dsmem_ptr(&qpart[1][1][bb][c], r);
    }
    float q_on = 0.f, q_tg = 0.f;
#pragma unroll
    for (int r = 0; r < CL; ++r) {                  // fixed order: CTA 0..7, column half 0, 1
      q_on += p[r][0];
      q_on += p[r][1];
      q_tg += p[r][2];
      q_tg += p[r][3];
    }
    q_on += b5s[0][c];
    q_tg += b5s[1][c];
    const int act = td_act;
    // over the 8 lanes of a sample: max_a Q_target (actions < A only, first maximum value), Q_online[action]
    float maxq = c < a.A ? q_tg : -INFINITY;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) maxq = fmaxf(maxq, __shfl_xor_sync(0xffffffffu, maxq, o));
    const float q_sel = __shfl_sync(0xffffffffu, q_on, (lane & ~7) | act);
    float se_v = 0.f, g_v = 0.f;
    if (valid) td_terms(q_sel, maxq, td_reward, td_terminal, a.gamma, a.B, a.loss_kind, &se_v, &g_v);
    if (c == 0) {
      g_own[s] = valid ? g_v : 0.f;
      if (valid) { a.se[bb] = se_v; a.g[bb] = g_v; }
    }
    if (valid && a.qout && c < a.A) a.qout[(size_t)bb * a.A + c] = q_on;
    if (rank == 0 && lane == 0) *a.step += 1;        // global_step (network.py:83, 198-202)
  }
  __syncthreads();
  FCC_STAMP(6);
  cluster_sync_all();                               // barrier 2: every CTA's g_own is complete
  FCC_STAMP(7);

  // ---------------------------------------------------------------- dL/da4 for this CTA's 32 hidden units, ReLU-gated
  if (tid < MAXB) g_all[tid] = *dsmem_ptr(&g_own[tid & 3], (uint32_t)(tid >> 2));
  __syncthreads();
  for (int i = tid; i < MAXB * NS / 2; i += THREADS) {      // pairs of hidden units
    const int bb = i / (NS / 2), j = (i % (NS / 2)) * 2;
    if (bb < a.B) {
      const int act = a.actions[bb];
      const float gg = g_all[bb];
      const uint32_t gt = gate[bb];
      const float d0 = (gt >> j) & 1u ? gg * w5s[0][j * FC_AP + act] : 0.f;
      const float d1 = (gt >> (j + 1)) & 1u ? gg * w5s[0][(j + 1) * FC_AP + act] : 0.f;
      *reinterpret_cast<uint32_t*>(a.dh_bf + (size_t)bb * FCF + n0 + j) = pack_bf16(d0, d1);
    }
  }
  __syncthreads();
  FCC_STAMP(8);
  cluster_sync_all();                               // barrier 3: all 256 columns of dL/da4 are in global memory
  FCC_STAMP(9);

  // ---------------------------------------------------------------- data gradient: dX[b, k] = sum_n dh[b, n] w4[k, n], k in this CTA's 32
  load_rows(sbase + OFF_A, a.dh_bf, a.B);
  cp_async_commit();
  cp_async_wait<0>();
  fence_async_proxy();
  __syncthreads();
  FCC_STAMP(10);
  if (warp == 0) {
    mbar_wait(&wdbar, 0);
    tc_fence_after();
    if (elect_one()) {
#pragma unroll
      for (int kb = 0; kb < 4; ++kb) {
        const uint64_t ad0 = make_desc_sw128(sbase + OFF_A + kb * BLK);
        const uint64_t bd0 = make_desc_sw128(sbase + OFF_WD + kb * BLK);
#pragma unroll
        for (int j = 0; j < TK / 16; ++j)
          mma_bf16(tmem + 2 * NS, desc_advance(ad0, j * 32), desc_advance(bd0, j * 32), idesc, (kb | j) ? 1u : 0u);
      }
      mma_commit(&dbar);
    }
    __syncwarp();
  }
  if ((warp & 3) == 0 && warp < 8) {                // warps 0 and 4: 16 columns each
    const int half = warp >> 2;
    mbar_wait(&dbar, 0);
    tc_fence_after();
    FCC_STAMP(11);
    float v[16];
    tmem_ld16(tmem + 2 * NS + half * 16, v);
    if (b < a.B) {
      const size_t off = (size_t)b * FCF + n0 + half * 16;
      uint32_t pk[8];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        // ReLU gate of conv3: zero the gradient where its pooled output is not positive
        const uint4 p = *reinterpret_cast<const uint4*>(a.x[0] + off + 8 * j);
        const uint32_t pw[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const bool lo = ((pw[q] & 0x7FFFu) != 0u) && !(pw[q] & 0x8000u);
          const bool hi = ((pw[q] & 0x7FFF0000u) != 0u) && !(pw[q] & 0x80000000u);
          pk[4 * j + q] = pack_bf16(lo ? v[8 * j + 2 * q] : 0.f, hi ? v[8 * j + 2 * q + 1] : 0.f);
        }
      }
      *reinterpret_cast<uint4*>(a.dx + off) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      *reinterpret_cast<uint4*>(a.dx + off + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
    }
  }
  FCC_STAMP(12);
  teardown(tmem, 128);
  FCC_STAMP(13);
}
#undef FCC_STAMP

}  // namespace fcc
}  // namespace umma
}  // namespace dqn
