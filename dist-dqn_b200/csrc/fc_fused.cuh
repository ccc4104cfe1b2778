// One kernel for the middle of the training step of the reference ConvNetwork (fc width 256, fc input 256):
//   fc forward of BOTH parameter sets (network.py:421-424 a4 = relu(a3 w4 + b4); online on the sampled states, target on
//   the next states)  ->  Q head of both (q = a4 w5 + b5)  ->  TD target, loss terms and dL/dq (dqn_agent.py:118-122,
//   224-253; network.py:141-157)  ->  gradient of the head w.r.t. a4, ReLU-gated  ->  fc data gradient
//   dL/da3 = (dL/da4 . relu') w4^T, gated by conv3's ReLU.
// It replaces three dependent launches (umma_fwd_kernel<F4>, td_head_kernel, umma_dgrad_kernel<F4>) on the critical
// path of a step; the three GEMMs run on the tensor cores (tcgen05.mma, M=128 samples, N=256), the head and the TD
// arithmetic in fp32/fp64 exactly like td_head_kernel.
//
// One CTA = 128 samples.  TMEM (all 512 columns): [0,256) online fc accumulator, [256,512) target fc accumulator,
// later re-used for the data gradient.  Shared memory: the 4-slot operand ring of the forward GEMMs (4 x (16 KB
// activations + 32 KB weights)); after the forward GEMMs the activation slots hold the gated head gradient (the A
// operand of the data-gradient GEMM, 4 k-blocks of 64) and the weight slots the four 64-deep stages of w4.
#pragma once
#include "umma_core.cuh"

namespace dqn {

__device__ __forceinline__ void td_terms(float q_sel, float maxq, double reward, bool terminal, double gamma, int B,
                                         int loss_kind, float* se, float* g) {
  // dqn_agent.py:242-246: python float64 arithmetic, rounded to float32 by the q_placeholder
  const float y = (float)(reward + gamma * (terminal ? 0.0 : (double)maxq));
  const float diff = q_sel - y;
  if (loss_kind == DQN_LOSS_HUBER) {
    const float ad = fabsf(diff);
    *se = ad <= 1.f ? 0.5f * diff * diff : ad - 0.5f;
    *g = fminf(fmaxf(diff, -1.f), 1.f) / (float)B;
  } else {                                   // network.py:150  reduce_mean(squared_difference)
    *se = diff * diff;
    *g = 2.f * diff / (float)B;
  }
}

namespace umma {

constexpr int FCF = 256;                                   // fc width == fc input width
constexpr int FC_AP = 8;                                   // head weights are padded to 8 actions per hidden unit
constexpr uint32_t FC_B_STAGE = FCF * 128;                 // 32 KB: 256 rows x 64 reduction elements
constexpr uint32_t FC_OFF_B = RS * A_STAGE_BYTES;          // weight slots start after the 4 activation slots
constexpr uint32_t FC_SMEM = SMEM_SLACK + RS * (A_STAGE_BYTES + FC_B_STAGE);

struct FcFusedArgs {
  const __nv_bfloat16* x[2];      // conv3 pooled output of {online(states), target(next states)}, [B,256] bf16
  const __nv_bfloat16* wt[2];     // fc weights transposed [n][k] bf16, {online, target}
  const float* b4[2];
  const float* w5[2];             // head weights fp32 [256][A]
  const float* b5[2];
  const __nv_bfloat16* wd;        // online fc weights [k][n] bf16 (data-gradient operand)
  const int32_t* actions;
  const double* rewards;
  const uint8_t* terminals;
  float* a4;                      // online fc output, fp32 [B,256] (input of the head weight gradient)
  float* qout;                    // optional [B,A]
  float* se;                      // per-sample loss term
  float* g;                       // per-sample dL/dq_selected
  __nv_bfloat16* dh_bf;           // [B,256] ReLU-gated dL/da4 (operand of the fc weight gradient)
  __nv_bfloat16* dx;              // [B,256] dL/da3 gated by conv3's ReLU
  long long* step;
  double gamma;
  int B, A, loss_kind;
  unsigned long long* trace;      // debugging (DQN_S2D_TRACE=1): CTA 0 of fc_td_cluster_kernel stamps clock64() per phase at trace[1080 + i]
};

__global__ void __launch_bounds__(TT) fc_td_fused_kernel(FcFusedArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = align_smem(smem_raw);
  __shared__ uint64_t bars[2];
  __shared__ uint64_t full[RS];       // a ring slot is filled: one arrival per warp
  __shared__ uint64_t done[RS];       // the MMAs reading a ring slot have completed
  __shared__ uint64_t wdbar[RS];      // a stage of w4 (data-gradient operand) has landed
  __shared__ uint64_t fin;            // all data-gradient MMAs have completed
  __shared__ uint32_t tmem_slot;
  __shared__ float b4s[2][FCF];
  __shared__ __align__(16) float w5s[2][FCF * FC_AP];
  __shared__ float b5s[2][FC_AP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int m0 = blockIdx.x * TM;
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < RS; ++i) { mbar_init(&full[i], TT / 32 + 1); mbar_init(&done[i], 1); mbar_init(&wdbar[i], 1); }
    mbar_init(&fin, 1);
  }
  const uint32_t tmem = setup(bars, &tmem_slot, 512);
  const uint32_t sbase = smem_u32(smem);
  // parameters were last written by the previous step's optimizer, several launches back: safe to read before the wait
  for (int i = tid; i < 2 * FCF; i += TT) b4s[i / FCF][i % FCF] = __ldg(a.b4[i / FCF] + (i % FCF));
  for (int i = tid; i < 2 * FCF * FC_AP; i += TT) {
    const int net = i / (FCF * FC_AP), j = (i / FC_AP) % FCF, c = i % FC_AP;
    w5s[net][j * FC_AP + c] = c < a.A ? __ldg(a.w5[net] + j * a.A + c) : 0.f;
  }
  if (tid < 2 * FC_AP) b5s[tid / FC_AP][tid % FC_AP] = (tid % FC_AP) < a.A ? __ldg(a.b5[tid / FC_AP] + (tid % FC_AP)) : 0.f;
  pdl_wait();

  // ---------------------------------------------------------------- forward GEMMs: 8 stages (net = j/4, k-block = j%4)
  const int gc = tid & 7;
  constexpr uint32_t idesc = make_idesc(TM, FCF, false, false);
  uint32_t adst[2];
  bool aok[2];
  size_t asrc[2];
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const int row = (tid >> 3) + 64 * q;
    aok[q] = m0 + row < a.B;
    asrc[q] = aok[q] ? (size_t)(m0 + row) * FCF + gc * 8 : 0;
    adst[q] = sw128_off(row, gc);
  }
  auto issue_fwd = [&](int j) {
    if (j < 8) {
      const int net = j >> 2, kb = j & 3, slot = j & 3;
#pragma unroll
      for (int q = 0; q < 2; ++q)
        cp_async16(sbase + slot * A_STAGE_BYTES + adst[q], a.x[net] + asrc[q] + (aok[q] ? kb * TK : 0), aok[q]);
      if (tid == 0) {                 // the stage's 256 x 64 weight tile is one contiguous image (umma::img_index)
        mbar_expect_tx(&full[slot], FC_B_STAGE);
        bulk_g2s(sbase + FC_OFF_B + slot * FC_B_STAGE, a.wt[net] + (size_t)kb * (FCF * 64), FC_B_STAGE, &full[slot]);
      }
    }
    cp_async_commit();
  };
  issue_fwd(0);
  issue_fwd(1);
  issue_fwd(2);
  const int wu = warp_uniform_idx();
#pragma unroll 1
  for (int it = 0; it < 8; ++it) {
    const int slot = it & 3;
    if (it >= 1 && it + 3 < 8) mbar_wait(&done[(it + 3) & 3], ((it - 1) >> 2) & 1);   // slot of stage it-1 is free again
    issue_fwd(it + 3);
    cp_async_wait<3>();
    fence_async_proxy();
    __syncwarp();
    if (lane == 0) mbar_arrive(&full[slot]);
    if (wu == 0) {
      mbar_wait(&full[slot], (it >> 2) & 1);
      issue_stage<true, true>(tmem + (uint32_t)(it >> 2) * FCF, sbase + slot * A_STAGE_BYTES, sbase + FC_OFF_B + slot * FC_B_STAGE,
                              idesc, (it & 3) == 0, &done[slot]);
    }
  }
  mbar_wait(&done[3], 1);             // stage 7's commit covers every MMA issued before it
  tc_fence_after();
  // the ring is free: start fetching w4 for the data-gradient GEMM (stages 0..2; slot 3 is used as scratch below)
  auto issue_wd = [&](int s) {
    if (tid == 0) {
      mbar_expect_tx(&wdbar[s], FC_B_STAGE);
      bulk_g2s(sbase + FC_OFF_B + s * FC_B_STAGE, a.wd + (size_t)s * (FCF * 64), FC_B_STAGE, &wdbar[s]);
    }
  };
  issue_wd(0);
  issue_wd(1);
  issue_wd(2);

  // ---------------------------------------------------------------- fc epilogue + Q head partial sums
  // thread = (sample row r of the tile, column group cg of 64 hidden units); warp w reads TMEM lanes 32*(w%4)..
  const int r = tid & (TM - 1), cg = tid >> 7;
  const int b = m0 + r;
  const bool valid = b < a.B;
  const uint32_t trow = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  int act = 0;
  double reward = 0.0;
  bool terminal = false;
  if (valid) {
    act = a.actions[b];
    reward = a.rewards[b];
    terminal = a.terminals[b] != 0;
  }
  float* qs = reinterpret_cast<float*>(smem + FC_OFF_B + 3 * FC_B_STAGE);     // [net][cg][row][8] partial head sums
  uint32_t pos[2] = {0u, 0u};                                                  // a4 > 0 for this thread's 64 units
#pragma unroll
  for (int net = 0; net < 2; ++net) {
    float qa[FC_AP];
#pragma unroll
    for (int c = 0; c < FC_AP; ++c) qa[c] = 0.f;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int c0 = cg * 64 + half * 32;
      float v[32];
      tmem_ld32(trow + (uint32_t)(net * FCF + c0), v);
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + b4s[net][c0 + j], 0.f);
      if (net == 0) {
        uint32_t m = 0u;
#pragma unroll
        for (int j = 0; j < 32; ++j) m |= (v[j] > 0.f ? 1u : 0u) << j;
        pos[half] = m;
        if (valid) {
          float* dst = a.a4 + (size_t)b * FCF + c0;
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<float4*>(dst + 4 * j) = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        }
      }
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float4 w0 = *reinterpret_cast<const float4*>(&w5s[net][(c0 + j) * FC_AP]);
        const float4 w1 = *reinterpret_cast<const float4*>(&w5s[net][(c0 + j) * FC_AP + 4]);
        qa[0] = fmaf(v[j], w0.x, qa[0]); qa[1] = fmaf(v[j], w0.y, qa[1]);
        qa[2] = fmaf(v[j], w0.z, qa[2]); qa[3] = fmaf(v[j], w0.w, qa[3]);
        qa[4] = fmaf(v[j], w1.x, qa[4]); qa[5] = fmaf(v[j], w1.y, qa[5]);
        qa[6] = fmaf(v[j], w1.z, qa[6]); qa[7] = fmaf(v[j], w1.w, qa[7]);
      }
    }
    float* dst = qs + (size_t)((net * 4 + cg) * TM + r) * FC_AP;
    *reinterpret_cast<float4*>(dst) = make_float4(qa[0], qa[1], qa[2], qa[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(qa[4], qa[5], qa[6], qa[7]);
  }
  tc_fence_before();
  __syncthreads();

  // ---------------------------------------------------------------- Q values, TD target, loss term, dL/dq (every thread of a row)
  float q_on[FC_AP], q_tg[FC_AP];
#pragma unroll
  for (int c = 0; c < FC_AP; ++c) {
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int k = 0; k < 4; ++k) {                          // fixed order: column groups 0..3
      s0 += qs[(size_t)((0 * 4 + k) * TM + r) * FC_AP + c];
      s1 += qs[(size_t)((1 * 4 + k) * TM + r) * FC_AP + c];
    }
    q_on[c] = s0 + b5s[0][c];
    q_tg[c] = s1 + b5s[1][c];
  }
  float maxq = q_tg[0], q_sel = q_on[0];
#pragma unroll
  for (int c = 1; c < FC_AP; ++c) {
    if (c < a.A && q_tg[c] > maxq) maxq = q_tg[c];
    if (c == act) q_sel = q_on[c];
  }
  float se_v, g_v;
  td_terms(q_sel, maxq, reward, terminal, a.gamma, a.B, a.loss_kind, &se_v, &g_v);
  if (cg == 0 && valid) {
    a.se[b] = se_v;
    a.g[b] = g_v;
    if (a.qout) {
#pragma unroll
      for (int c = 0; c < FC_AP; ++c)
        if (c < a.A) a.qout[(size_t)b * a.A + c] = q_on[c];
    }
  }
  if (blockIdx.x == 0 && tid == 0) *a.step += 1;            // global_step (network.py:83, 198-202)
  // head data gradient for this thread's 64 hidden units, ReLU-gated: A operand of the data-gradient GEMM (k-block cg)
  const float gg = valid ? g_v : 0.f;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int c0 = cg * 64 + half * 32;
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float d0 = (pos[half] >> (2 * j)) & 1u ? gg * w5s[0][(c0 + 2 * j) * FC_AP + act] : 0.f;
      const float d1 = (pos[half] >> (2 * j + 1)) & 1u ? gg * w5s[0][(c0 + 2 * j + 1) * FC_AP + act] : 0.f;
      pk[j] = pack_bf16(d0, d1);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint4 val = make_uint4(pk[4 * i], pk[4 * i + 1], pk[4 * i + 2], pk[4 * i + 3]);
      *reinterpret_cast<uint4*>(smem + cg * A_STAGE_BYTES + sw128_off(r, half * 4 + i)) = val;
      if (valid) *reinterpret_cast<uint4*>(a.dh_bf + (size_t)b * FCF + c0 + 8 * i) = val;
    }
  }
  fence_async_proxy();
  tc_fence_before();
  __syncthreads();

  // ---------------------------------------------------------------- data-gradient GEMM: dX[r,k] = sum_n dh[r,n] w4[k,n]
  issue_wd(3);                                   // the scratch in slot 3 is dead now
  tc_fence_after();
  if (wu == 0) {                                 // the A operand is complete (barrier above): only the weights arrive in stages
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      mbar_wait(&wdbar[s], 0);
      issue_stage<true, true>(tmem + FCF, sbase + s * A_STAGE_BYTES, sbase + FC_OFF_B + s * FC_B_STAGE, idesc, s == 0,
                              s == 3 ? &fin : &done[s]);
    }
  }
  pdl_trigger();
  mbar_wait(&fin, 0);
  tc_fence_after();
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int c0 = cg * 64 + half * 32;
    float v[32];
    tmem_ld32(trow + (uint32_t)(FCF + c0), v);
    if (valid) {
      const size_t off = (size_t)b * FCF + c0;
      uint32_t pk[16];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
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
#pragma unroll
      for (int j = 0; j < 4; ++j)
        *reinterpret_cast<uint4*>(a.dx + off + 8 * j) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
    }
  }
  teardown(tmem, 512);
}

}  // namespace umma
}  // namespace dqn
