// Q-network training step: everything Network.create_network builds (reference src/network.py:13-424)
// and DQNAgent runs through session.run (src/dqn_agent.py:108-143, 176-189, 215-253), as hand-written
// CUDA kernels over flat parameter buffers.
//
// Kernel inventory (fp32 path)
//   conv_fwd_kernel     conv/fc + bias + ReLU + 2x2 max-pool, online and target pass in one launch
//   td_head_kernel      Q head of both nets, max_a Q_target, TD target (float64), loss term, dL/dQ,
//                       gradient into the last hidden layer      (dqn_agent.py:118-125, 242-246; network.py:149-150)
//   head_wgrad_kernel   gradient of the head weights/bias
//   conv_wgrad_kernel   weight + bias gradients, split over the batch into scratch partial sums
//   conv_dgrad_kernel   gradient into the previous layer's activations
//   optimizer_kernel    adds the partial sums up, L2 term, and applies one of TF's 7 update rules (network.py:169-202)
//   mlp_*_kernel        the 4-20-20-A tanh net of SimpleNetwork as warp-level kernels (network.py:258-313)
#include <cuda_fp16.h>

#include "common.cuh"
#include "gemm_core.cuh"
#include "umma_core.cuh"
#include "fc_fused.cuh"
#include "conv1_s2d.cuh"
#include "fc_cluster.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <new>
#include <vector>

namespace dqn {

// ConvNetwork (network.py:318-333, 389-424) on 84x84x4: SAME pads (2,2),(1,2),(1,1); pools 21->11, 6->3, 3->2
using C1 = Geo<84, 84, 4, 8, 4, 2, 2, 21, 21, 32, true>;
using C2 = Geo<11, 11, 32, 4, 2, 1, 1, 6, 6, 64, true>;
using C3 = Geo<3, 3, 64, 3, 1, 1, 1, 3, 3, 64, true>;
using F4 = Geo<1, 1, 256, 1, 1, 0, 0, 1, 1, 256, false>;
// Nature-paper variant: VALID convs, no pooling, fc 512
using N1 = Geo<84, 84, 4, 8, 4, 0, 0, 20, 20, 32, false>;
using N2 = Geo<20, 20, 32, 4, 2, 0, 0, 9, 9, 64, false>;
using N3 = Geo<9, 9, 64, 3, 1, 0, 0, 7, 7, 64, false>;
using NF = Geo<1, 1, 3136, 1, 1, 0, 0, 1, 1, 512, false>;

constexpr int HID = 20;          // SimpleNetwork.HIDDEN1_SIZE == HIDDEN2_SIZE (network.py:255-256)
constexpr int MLP_MAX_IN = 16;
constexpr int MLP_MAX_A = 16;
constexpr int MLP_WARPS = 8;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One warp per sample; lane l owns hidden units l, l+32, ... (FPL per lane, kept in registers so every load of the
// head weights is independent of the reductions).
template <int FPL>
__global__ void __launch_bounds__(256)
td_head_kernel(const float* __restrict__ h_on, const float* __restrict__ h_tg, const float* __restrict__ w_on,
               const float* __restrict__ b_on, const float* __restrict__ w_tg, const float* __restrict__ b_tg,
               const int32_t* __restrict__ actions, const double* __restrict__ rewards,
               const uint8_t* __restrict__ terminals, int B, int A, double gamma, int loss_kind,
               float* __restrict__ qout, float* __restrict__ se, float* __restrict__ g, float* __restrict__ dh,
               __nv_bfloat16* __restrict__ dh_bf, long long* __restrict__ step) {
  constexpr int F = FPL * 32;
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  // head weights of both nets -> shared memory BEFORE the PDL wait (last written by the previous step's optimizer):
  // their L2 latency overlaps the tail of the fc forward kernel instead of sitting in the per-action loop
  extern __shared__ float td_sm[];                       // [2][F*A + A]
  float* ws_on = td_sm;
  float* ws_tg = td_sm + (F * A + A);
  pdl_trigger();
  for (int i = threadIdx.x; i < F * A; i += blockDim.x) { ws_on[i] = __ldg(w_on + i); ws_tg[i] = __ldg(w_tg + i); }
  if (threadIdx.x < A) { ws_on[F * A + threadIdx.x] = __ldg(b_on + threadIdx.x); ws_tg[F * A + threadIdx.x] = __ldg(b_tg + threadIdx.x); }
  pdl_wait();
  if (blockIdx.x == 0 && threadIdx.x == 0) *step += 1;      // global_step (network.py:83, 198-202)
  float x[FPL], xt[FPL];
  int act = 0;
  double reward = 0.0;
  bool terminal = false;
  if (b < B) {
#pragma unroll
    for (int i = 0; i < FPL; ++i) {
      x[i] = h_on[(size_t)b * F + lane + 32 * i];
      xt[i] = h_tg[(size_t)b * F + lane + 32 * i];
    }
    act = actions[b];
    reward = rewards[b];
    terminal = terminals[b] != 0;
  }
  __syncthreads();
  if (b >= B) return;
  float q_sel = 0.f, maxq = 0.f;
  for (int a = 0; a < A; ++a) {
    float s = 0.f, st = 0.f;
#pragma unroll
    for (int i = 0; i < FPL; ++i) {
      s = fmaf(x[i], ws_on[(lane + 32 * i) * A + a], s);
      st = fmaf(xt[i], ws_tg[(lane + 32 * i) * A + a], st);
    }
    s = warp_sum(s) + ws_on[F * A + a];
    st = warp_sum(st) + ws_tg[F * A + a];
    if (qout && lane == 0) qout[(size_t)b * A + a] = s;
    if (a == 0 || st > maxq) maxq = st;
    if (a == act) q_sel = s;
  }
  float e, gg;
  td_terms(q_sel, maxq, reward, terminal, gamma, B, loss_kind, &e, &gg);
  if (lane == 0) { se[b] = e; g[b] = gg; }
#pragma unroll
  for (int i = 0; i < FPL; ++i) {
    const int j = lane + 32 * i;
    const float d = gg * ws_on[j * A + act];
    dh[(size_t)b * F + j] = d;
    if (dh_bf) dh_bf[(size_t)b * F + j] = __float2bfloat16_rn(x[i] > 0.f ? d : 0.f);   // ReLU-gated for the tensor-core path
  }
}

// out[split][(F+1)*A]: rows 0..F-1 = dL/dw5 (data term), row F = dL/db5; blockIdx.y = batch split (partial sums are
// added up in split order by the optimizer, like the conv wgrad partials).
__global__ void __launch_bounds__(256)
head_wgrad_kernel(const float* __restrict__ h_on, const float* __restrict__ g, const int32_t* __restrict__ actions,
                  int Btot, int per_split, int F, int A, float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= (F + 1) * A) return;
  const int j = o / A, a = o - j * A;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};      // 4 independent chains; summed in a fixed order
  const int jj = j < F ? j : 0;
  int b = blockIdx.y * per_split;
  const int B = min(Btot, b + per_split);
  out += (size_t)blockIdx.y * (size_t)((F + 1) * A);
  for (; b + 16 <= B; b += 16) {            // 16 samples' loads in flight at once (the kernel is a chain of L2 latencies)
    float hv[16], gv[16];
    int av[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      hv[u] = j < F ? h_on[(size_t)(b + u) * F + jj] : 1.f;
      gv[u] = g[b + u];
      av[u] = actions[b + u];
    }
#pragma unroll
    for (int u = 0; u < 16; ++u) acc[u & 3] += (av[u] == a) ? hv[u] * gv[u] : 0.f;
  }
  for (; b + 4 <= B; b += 4) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const float hv = j < F ? h_on[(size_t)(b + u) * F + jj] : 1.f;
      acc[u] += (actions[b + u] == a) ? hv * g[b + u] : 0.f;
    }
  }
  for (; b < B; ++b) acc[0] += (actions[b] == a) ? (j < F ? h_on[(size_t)b * F + jj] : 1.f) * g[b] : 0.f;
  out[o] = (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// Predict path (dqn_agent.py:176-189, 159-160): q = h @ w + b, optional row-wise arg-max (first maximum).
__global__ void __launch_bounds__(256)
head_forward_kernel(const float* __restrict__ h, const float* __restrict__ w, const float* __restrict__ bias, int n,
                    int F, int A, float* __restrict__ q, int32_t* __restrict__ argmax) {
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= n) return;
  const float* x = h + (size_t)b * F;
  float best = 0.f;
  int bi = 0;
  for (int a = 0; a < A; ++a) {
    float s = 0.f;
    for (int j = lane; j < F; j += 32) s = fmaf(x[j], w[j * A + a], s);
    s = warp_sum(s) + bias[a];
    if (q && lane == 0) q[(size_t)b * A + a] = s;
    if (a == 0 || s > best) { best = s; bi = a; }
  }
  if (argmax && lane == 0) argmax[b] = bi;
}

// ------------------------------------------------------------------------------------------------
// Optimizer (network.py:159-203): one thread per parameter; the gradient is the fixed-order sum of
// the wgrad partials (or the all-reduced flat gradient), times grad_scale, plus reg_param * w on
// the regularised tensors (network.py:151-152).  Update rules are TF r0.9-era (see oracle/net_ref.py).
// ------------------------------------------------------------------------------------------------
struct Seg {
  long long off;          // first float of the segment [w | b] in the flat buffers
  int len;                // K*N + N
  int reg_len;            // leading floats under L2 (K*N or 0)
  const float* partial;   // [nsplit][stride]
  int nsplit;
  long long stride;
  // bf16 tensor-core mode: operand copies of the weight part, refreshed by the update itself
  __nv_bfloat16* wt;      // [N][K]            (forward)
  __nv_bfloat16* wd;      // [CIN][taps][N]    (dgrad), may be null
  int K, N, CIN;
  int kperm;              // conv1 in the space-to-depth formulation: the image's reduction index is s2d::kperm(k) ...
  int f16;                //   ... and its elements are IEEE halves (conv1's MMAs run f16 x f16, see conv1_s2d.cuh)
};
constexpr int MAX_SEGS = 6;
struct OptArgs {
  Seg seg[MAX_SEGS];
  int nseg;
  float* w;
  float* s0;
  float* s1;
  const long long* step;
  float lr, reg, momentum, decay, gscale;
  int opt;
  // data-parallel one-shot all-reduce over NVLink peer memory (world > 1): the gradient of every rank lives in a
  // double-buffered exchange block that all ranks have mapped; flags[p] = last step whose gradient rank p published
  int world;
  const float* peer_grad[8];             // exchange gradient buffers of all ranks ([2][P] floats each), rank order
  const unsigned long long* flags;       // this rank's flag arrays (written by the peers): slot s = [8s..8s+7], s = 0..3
  long long P;
  int peer_mask;                         // bit s set: segment s is summed over the ranks here; clear: already summed locally
  int flag_slot;                         // which flag array guards the peer-summed segments
  long long timeout_ns;                  // how long a rank waits for its peers' flags before it reports them missing
  unsigned long long* err;               // device word: last step whose wait timed out (0 = none); read by dqn_net_dp_status
  unsigned long long* trace;             // DQN_S2D_TRACE=1: globaltimer stamps of the exchange (tools/trace_dp.py), else null
  // optimizer only, conv nets: the blocks of the peer-summed segments first PUBLISH this rank's gradient of those segments
  // (sum of the weight-gradient partials -> exchange block, block counter, flags), then wait and sum over the ranks
  unsigned long long* const* pub_flags;  // null: a separate kernel published (MLP, NCCL mode)
  unsigned int* pub_counter;
  float* pub_grad;                       // this rank's exchange gradient buffers [2][P]
  int pub_total;                         // blocks that publish (those with work in the peer-summed segments)
  int rank, pub_sys_release;
  long long* idx_cursor;                 // dqn_train_step_table: cursor of the index table, advanced here (null otherwise)
};
__device__ __forceinline__ void dp_stamp(unsigned long long* trace, int slot, bool max_of_all = false) {
  if (!trace) return;
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  if (max_of_all) atomicMax(trace + 1136 + slot, t);
  else trace[1136 + slot] = t;
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// One thread waits until every rank has published its gradient of `step` (flags[r] >= step).  The wait is bounded by
// a.timeout_ns of wall-clock time (dqn_net_dp_set_timeout; ranks are driven by independent host code, so seconds of skew
// are normal): when it expires the step number is left in *a.err for the host (dqn_net_dp_status) and the kernel goes on
// with whatever the exchange buffers hold, instead of trapping and taking the CUDA context down.
// Called by the first warp of a block: lane r polls rank r's flag, so the system-scope acquire loads (0.6 us each,
// tools/trace_dp.py) overlap instead of queueing one rank after the other.
__device__ __forceinline__ void wait_peer_flags(const OptArgs& a, unsigned long long step) {
  const int r = threadIdx.x & 31;
  if (r < a.world) {
    const unsigned long long t0 = global_timer_ns();
    unsigned spins = 0;
    while (ld_acquire_sys(a.flags + 8 * a.flag_slot + r) < step) {
      if ((++spins & 1023u) == 0 && (long long)(global_timer_ns() - t0) > a.timeout_ns) {
        atomicMax(a.err, step);
        break;
      }
    }
  }
  __syncwarp();
}

// Sum over the ranks of element gi of the exchange buffers.  All loads are issued before the first add (a loop with
// a run-time trip count is a serial chain of NVLink round trips); the fixed tree is the same on every rank, so the
// ranks end up with bit-identical sums.
__device__ __forceinline__ float peer_sum1(const OptArgs& a, size_t gi) {
  float v[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) v[r] = r < a.world ? ld_relaxed_sys(a.peer_grad[r] + gi) : 0.f;
  return ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
}
__device__ __forceinline__ float4 peer_sum4(const OptArgs& a, size_t gi) {      // gi % 4 == 0
  float4 v[8];
#pragma unroll
  for (int r = 0; r < 8; ++r) v[r] = r < a.world ? ld_relaxed_sys_v4(a.peer_grad[r] + gi) : make_float4(0.f, 0.f, 0.f, 0.f);
  float4 s;
  s.x = ((v[0].x + v[1].x) + (v[2].x + v[3].x)) + ((v[4].x + v[5].x) + (v[6].x + v[7].x));
  s.y = ((v[0].y + v[1].y) + (v[2].y + v[3].y)) + ((v[4].y + v[5].y) + (v[6].y + v[7].y));
  s.z = ((v[0].z + v[1].z) + (v[2].z + v[3].z)) + ((v[4].z + v[5].z) + (v[6].z + v[7].z));
  s.w = ((v[0].w + v[1].w) + (v[2].w + v[3].w)) + ((v[4].w + v[5].w) + (v[6].w + v[7].w));
  return s;
}

// Sum of the weight-gradient partials of element i in a fixed order (deterministic).  Up to ~150 partials (one per CTA
// of the conv1 weight-gradient kernel) sit 33 KB apart: a plain loop is a chain of L2 round trips (measured 8.6 us for
// 61 partials), so 16 loads are kept in flight and summed as a fixed tree.
__device__ __forceinline__ float seg_grad(const Seg& s, int i) {
  const float* p = s.partial + i;
  const size_t st = (size_t)s.stride;
  if (s.nsplit <= 2) return s.nsplit == 2 ? p[0] + p[st] : p[0];
  float total = 0.f;
  int k = 0;
  for (; k + 16 <= s.nsplit; k += 16) {
    const float* q = p + (size_t)k * st;
    const float v0 = q[0], v1 = q[st], v2 = q[2 * st], v3 = q[3 * st], v4 = q[4 * st], v5 = q[5 * st], v6 = q[6 * st], v7 = q[7 * st];
    const float v8 = q[8 * st], v9 = q[9 * st], v10 = q[10 * st], v11 = q[11 * st], v12 = q[12 * st], v13 = q[13 * st],
                v14 = q[14 * st], v15 = q[15 * st];
    total += (((v0 + v1) + (v2 + v3)) + ((v4 + v5) + (v6 + v7))) + (((v8 + v9) + (v10 + v11)) + ((v12 + v13) + (v14 + v15)));
  }
  if (k < s.nsplit) {                       // tail: up to 15 partials, still loaded together
    const float* q = p + (size_t)k * st;
    const int n = s.nsplit - k;
    float v[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) v[u] = u < n ? q[(size_t)u * st] : 0.f;
    total += (((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]))) +
             (((v[8] + v[9]) + (v[10] + v[11])) + ((v[12] + v[13]) + (v[14] + v[15])));
  }
  return total;
}

__global__ void __launch_bounds__(256) optimizer_kernel(OptArgs a) {
  pdl_wait();
  // every reader of this update's row of the index table (the conv1 kernels) has finished: the next update takes the next row
  if (a.idx_cursor && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) *a.idx_cursor += 1;
  const Seg s = a.seg[blockIdx.y];
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  // bf16 mode, weight part of a conv / fc tensor: the block owns a tile of 8 reduction rows x 32 output columns instead
  // of 256 consecutive elements (K*N is a multiple of 256, N of 32), so that the refreshed [n][k] stage image is
  // written as 16-byte chunks of 8 consecutive k (through shared memory) instead of 2-byte writes 128 bytes apart
  const bool tiled = s.wt != nullptr && (int)(blockIdx.x * blockDim.x) < s.K * s.N;
  __shared__ uint16_t wtile[8][32 + 2];           // bf16 (or f16, s.f16) bit patterns
  if (tiled) {
    const int ntn = s.N >> 5, kt = blockIdx.x / ntn, nt = blockIdx.x - kt * ntn;
    i = (kt * 8 + (int)(threadIdx.x >> 5)) * s.N + nt * 32 + (int)(threadIdx.x & 31);
  }
  __shared__ float lr_t_s;
  if (a.opt == DQN_OPT_ADAM) {
    if (threadIdx.x == 0) {
      const double t = (double)*a.step;
      lr_t_s = (float)((double)a.lr * sqrt(1.0 - pow(0.999, t)) / (1.0 - pow(0.9, t)));
    }
    __syncthreads();
  }
  const long long step = *a.step;
  if (a.world > 1 && ((a.peer_mask >> blockIdx.y) & 1)) {
    const bool tr0 = threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0;
    if (a.pub_flags && (int)blockIdx.x < (s.len + 255) / 256) {
      // Publish first: this rank's gradient of the element (sum of the weight-gradient partials) goes to its exchange
      // block; the last block to get there raises the rank's flag on every peer (see reduce_grad_kernel).  Every block
      // that waits below has published before, and the few hundred blocks of these segments are the first of the grid
      // to be scheduled (blockIdx.y = 0, 1) and all fit on the GPU at once, so the wait cannot starve a publisher.
      if (tr0) dp_stamp(a.trace, 8);
      if (i < s.len) a.pub_grad[(size_t)(step & 1) * (size_t)a.P + (size_t)s.off + i] = seg_grad(s, i);
      __shared__ bool last;
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence();
        last = atomicAdd(a.pub_counter, 1u) == (unsigned)a.pub_total - 1u;
      }
      __syncthreads();
      if (last) {
        if (threadIdx.x == 0) { *a.pub_counter = 0; dp_stamp(a.trace, 9); }
        if ((int)threadIdx.x < a.world) {
          __threadfence();
          unsigned long long* f = a.pub_flags[threadIdx.x] + 8 * a.flag_slot + a.rank;
          if (a.pub_sys_release) asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(f), "l"((unsigned long long)step) : "memory");
          else asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(f), "l"((unsigned long long)step) : "memory");
          if (threadIdx.x == 0) dp_stamp(a.trace, 10);
        }
      }
    }
    // wait until every rank has published its gradient of this step (segments that were exchanged earlier in the step
    // do not wait: their blocks finish while the flags are still on their way)
    if (tr0) dp_stamp(a.trace, 48);
    if (threadIdx.x < 32) wait_peer_flags(a, (unsigned long long)step);
    if (tr0) dp_stamp(a.trace, 49);
    __syncthreads();
  }
  if (i >= s.len) return;
  const size_t p = (size_t)s.off + i;
  float w = a.w[p];
  float g;
  if (a.world > 1 && ((a.peer_mask >> blockIdx.y) & 1)) {
    const size_t gi = (size_t)(step & 1) * (size_t)a.P + p;
    g = peer_sum1(a, gi) * a.gscale;
  } else {
    g = seg_grad(s, i) * a.gscale;
  }
  if (i < s.reg_len) g = fmaf(a.reg, w, g);
  switch (a.opt) {
    case DQN_OPT_SGD:
      w -= a.lr * g;
      break;
    case DQN_OPT_MOMENTUM: {
      const float acc = a.momentum * a.s0[p] + g;
      a.s0[p] = acc;
      w -= a.lr * acc;
    } break;
    case DQN_OPT_RMSPROP: {
      float ms = a.s0[p];
      ms += (1.f - a.decay) * (g * g - ms);
      const float mom = a.lr * g / sqrtf(ms + 1e-10f);      // TF default momentum = 0
      a.s0[p] = ms;
      a.s1[p] = mom;
      w -= mom;
    } break;
    case DQN_OPT_ADAM: {
      float m = a.s0[p], v = a.s1[p];
      m += (1.f - 0.9f) * (g - m);
      v += (1.f - 0.999f) * (g * g - v);
      a.s0[p] = m;
      a.s1[p] = v;
      w -= lr_t_s * m / (sqrtf(v) + 1e-8f);
    } break;
    case DQN_OPT_ADAGRAD: {
      const float acc = a.s0[p] + g * g;
      a.s0[p] = acc;
      w -= a.lr * g / sqrtf(acc);
    } break;
    case DQN_OPT_ADADELTA: {
      float acc = a.s0[p], accu = a.s1[p];
      acc = 0.95f * acc + 0.05f * g * g;
      const float u = sqrtf(accu + 1e-8f) / sqrtf(acc + 1e-8f) * g;
      accu = 0.95f * accu + 0.05f * u * u;
      a.s0[p] = acc;
      a.s1[p] = accu;
      w -= a.lr * u;
    } break;
    case DQN_OPT_FTRL: {
      const float acc = a.s0[p];
      float lin = a.s1[p];
      const float nacc = acc + g * g;
      lin += g - (sqrtf(nacc) - sqrtf(acc)) / a.lr * w;
      w = -lin / (sqrtf(nacc) / a.lr);
      a.s0[p] = nacc;
      a.s1[p] = lin;
    } break;
  }
  a.w[p] = w;
  if (threadIdx.x == 0) dp_stamp(a.trace, 50, true);
  if (tiled) {
    const int k = i / s.N, n = i - k * s.N;
    const __nv_bfloat16 v = __float2bfloat16_rn(w);
    wtile[threadIdx.x >> 5][threadIdx.x & 31] = s.f16 ? __half_as_ushort(__float2half_rn(w)) : __bfloat16_as_ushort(v);
    __syncthreads();
    if (threadIdx.x < 32) {              // column n of the tile: its 8 consecutive k are one 16-byte chunk of the image
      __align__(16) uint16_t c8[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) c8[e] = wtile[e][threadIdx.x];
      // thread < 32: k = 8*kt (kperm keeps runs of 8 consecutive reduction indices together)
      *reinterpret_cast<uint4*>(s.wt + umma::img_index(n, s.kperm ? umma::s2d::kperm(k) : k, s.N)) = *reinterpret_cast<const uint4*>(c8);
    }
    if (s.wd) {
      const int t = k / s.CIN, c = k - t * s.CIN;
      s.wd[umma::img_index(c, t * s.N + n, s.CIN)] = v;
    }
  }
}

// Materialises the flat gradient (data-parallel path and parity checks).
struct PublishArgs {
  unsigned long long* const* peer_flags;   // null: nothing to publish
  unsigned int* done_counter;              // blocks finished so far (reset by the last one)
  int rank, world;
  int flag_slot;                           // flag array to raise: 0 = early segments, 1 = final segment(s)
  int segs;                                // segments this launch covers, 4 bits each: blockIdx.y -> (segs >> 4*y) & 15
  int sys_release;                         // 1: raise the flags with a system-scope release store (see reduce_grad_kernel)
};
static int dp_separate_publish() {     // DQN_DP_SEPARATE_PUBLISH=1: conv1 / conv2 are published by a kernel of their own (A/B switch)
  static const int v = [] { const char* e = getenv("DQN_DP_SEPARATE_PUBLISH"); return e && e[0] == '1' ? 1 : 0; }();
  return v;
}
static int dp_sys_release() {
  static const int v = [] { const char* e = getenv("DQN_DP_SYS_RELEASE"); return e && e[0] == '1' ? 1 : 0; }();
  return v;
}

__global__ void __launch_bounds__(256) reduce_grad_kernel(OptArgs a, float* __restrict__ grad, float reg, int by_parity,
                                                          PublishArgs pub) {
  pdl_wait();
  if (pub.peer_flags && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0) dp_stamp(a.trace, 8 * pub.flag_slot);
  const Seg s = a.seg[(pub.segs >> (4 * blockIdx.y)) & 15];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (by_parity) grad += (size_t)(*a.step & 1) * (size_t)a.P;     // exchange block: buffer of this step's parity
  if (i < s.len) {
    float g = seg_grad(s, i);
    if (i < s.reg_len) g = fmaf(reg, a.w[(size_t)s.off + i], g);
    grad[(size_t)s.off + i] = g;
  }
  if (pub.peer_flags) {
    // the last block to finish tells every rank (system-scope release) that this rank's gradient is complete
    __shared__ bool last;
    __syncthreads();                      // the block's stores happen-before thread 0's fence (one fence per block, not 256)
    if (threadIdx.x == 0) {
      __threadfence();
      const unsigned int total = gridDim.x * gridDim.y;
      last = atomicAdd(pub.done_counter, 1u) == total - 1;
    }
    __syncthreads();
    if (last) {
      if (threadIdx.x == 0) { *pub.done_counter = 0; dp_stamp(a.trace, 8 * pub.flag_slot + 1); }
      if ((int)threadIdx.x < pub.world) {
        // The gradient lives in THIS GPU's memory and the peers read it through this GPU's L2, so a gpu-scope fence
        // (acquire side of the counter + every block's stores ordered before the flag) is what the flag store needs;
        // a system-scope release here cost 2.5 us per publish (tools/trace_dp.py).  DQN_DP_SYS_RELEASE=1 restores it.
        __threadfence();
        const unsigned long long v = (unsigned long long)*a.step;
        if (pub.sys_release)
          asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(pub.peer_flags[threadIdx.x] + 8 * pub.flag_slot + pub.rank),
                       "l"(v)
                       : "memory");
        else
          asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(pub.peer_flags[threadIdx.x] + 8 * pub.flag_slot + pub.rank),
                     "l"(v)
                     : "memory");
        if (threadIdx.x == 0) dp_stamp(a.trace, 8 * pub.flag_slot + 2);
      }
    }
  }
}

// Early half of the data-parallel exchange: sums the already published segments of all ranks into the local flat
// gradient buffer while the data-gradient chain is still running; the optimizer then only has to exchange what the
// last kernel of the step produces.
__global__ void __launch_bounds__(256) peer_sum_kernel(OptArgs a, int segs, float* __restrict__ gsum) {
  pdl_wait();
  const Seg s = a.seg[(segs >> (4 * blockIdx.y)) & 15];
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;            // 4 consecutive parameters per thread: 16-byte peer loads
  const long long step = *a.step;
  const bool tr0 = threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0;
  if (tr0) dp_stamp(a.trace, 32 + 4 * a.flag_slot);
  if (threadIdx.x < 32) wait_peer_flags(a, (unsigned long long)step);
  if (tr0) dp_stamp(a.trace, 33 + 4 * a.flag_slot);
  __syncthreads();
  if (i >= s.len) return;
  const size_t p = (size_t)s.off + i;                                  // segments start at multiples of 4 floats
  const size_t gi = (size_t)(step & 1) * (size_t)a.P + p;              // P is a multiple of 4 as well
  if (i + 4 <= s.len) {
    *reinterpret_cast<float4*>(gsum + p) = peer_sum4(a, gi);
  } else {
    for (int e = 0; i + e < s.len; ++e) gsum[p + e] = peer_sum1(a, gi + e);
  }
  if (threadIdx.x == 0) dp_stamp(a.trace, 34 + 4 * a.flag_slot, true);
}

// loss = mean_b se_b + reg * sum_{reg tensors} ||w||^2 / 2   (network.py:150-152); one CTA.  The weights are not written by
// any kernel of the step before the optimizer, so the regularisation sum runs BEFORE the PDL wait, in the shadow of the
// kernels that produce the squared errors.
constexpr int LOSS_THREADS = 1024;
__global__ void __launch_bounds__(LOSS_THREADS) loss_kernel(OptArgs a, const float* __restrict__ se, int B, float* __restrict__ out) {
  __shared__ float red[LOSS_THREADS / 32];
  pdl_trigger();
  float r = 0.f;
  for (int k = 0; k < a.nseg; ++k) {
    const Seg sg = a.seg[k];
    const float* w = a.w + (size_t)sg.off;
    int i0 = 0;
    if ((sg.off & 3) == 0) {                       // 16-byte aligned: four weights per load
      const int n4 = sg.reg_len >> 2;
      for (int i = threadIdx.x; i < n4; i += LOSS_THREADS) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(w) + i);
        r = fmaf(v.x, v.x, r); r = fmaf(v.y, v.y, r); r = fmaf(v.z, v.z, r); r = fmaf(v.w, v.w, r);
      }
      i0 = n4 << 2;
    }
    for (int i = i0 + threadIdx.x; i < sg.reg_len; i += LOSS_THREADS) r = fmaf(w[i], w[i], r);
  }
  r *= 0.5f * a.reg;
  pdl_wait();
  float s = 0.f;
  for (int b = threadIdx.x; b < B; b += LOSS_THREADS) s += se[b];
  s = s / (float)B + r;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = red[threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) *out = s;
  }
}

// Publishes "my gradient of step *step is complete" to every rank's flag array (including my own).
__global__ void dp_publish_kernel(unsigned long long* const* peer_flags, int rank, int world, const long long* step) {
  const int r = threadIdx.x;
  if (r < world) {
    __threadfence_system();
    const unsigned long long v = (unsigned long long)*step;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(peer_flags[r] + rank), "l"(v) : "memory");
  }
}

__global__ void empty_kernel() {
  pdl_trigger();
  pdl_wait();
}

__global__ void fill_kernel(float* p, long long n, float v) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) p[i] = v;
}

// bf16 copies of one layer's weights for the tensor-core path: wt[n][k] (forward: B operand, K-major) and
// wd[c][(ky,kx,n)] (dgrad), both stored as swizzled stage images (umma::img_index) so that a stage's tile is one
// contiguous bulk copy.  w is the fp32 master in TF layout [k=(ky,kx,c)][n].
struct CvtLayer { const float* w; __nv_bfloat16* wt; __nv_bfloat16* wd; int K, N, CIN, kperm, f16; };
struct CvtArgs { CvtLayer l[4]; };
__global__ void __launch_bounds__(256) weights_to_bf16_kernel(CvtArgs a) {
  const CvtLayer L = a.l[blockIdx.y];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L.K * L.N) return;
  const int k = i / L.N, n = i - k * L.N;
  const __nv_bfloat16 v = __float2bfloat16_rn(L.w[i]);
  reinterpret_cast<uint16_t*>(L.wt)[umma::img_index(n, L.kperm ? umma::s2d::kperm(k) : k, L.N)] =
      L.f16 ? __half_as_ushort(__float2half_rn(L.w[i])) : __bfloat16_as_ushort(v);
  if (L.wd) {
    const int t = k / L.CIN, c = k - t * L.CIN;
    L.wd[umma::img_index(c, t * L.N + n, L.CIN)] = v;
  }
}

// ------------------------------------------------------------------------------------------------
// SimpleNetwork (network.py:258-313): lane j of a warp owns hidden unit j of both hidden layers.
// ------------------------------------------------------------------------------------------------
struct MlpOff { int w1, b1, w2, b2, w3, b3, total; };

template <typename TIn>
__device__ __forceinline__ void mlp_forward_lane(const float* __restrict__ w, const MlpOff& o, const TIn* __restrict__ xin,
                                                 int IN, int A, int lane, float xv[MLP_MAX_IN], float* h1o, float* h2o,
                                                 float* qo) {
  const int j = lane < HID ? lane : HID - 1;
#pragma unroll
  for (int i = 0; i < MLP_MAX_IN; ++i) xv[i] = i < IN ? (float)xin[i] : 0.f;   // float32 placeholder feed (network.py:61)
  float s = w[o.b1 + j];
#pragma unroll
  for (int i = 0; i < MLP_MAX_IN; ++i)
    if (i < IN) s = fmaf(xv[i], w[o.w1 + i * HID + j], s);
  const float h1 = lane < HID ? tanhf(s) : 0.f;
  s = w[o.b2 + j];
#pragma unroll
  for (int i = 0; i < HID; ++i) s = fmaf(__shfl_sync(0xffffffffu, h1, i), w[o.w2 + i * HID + j], s);
  const float h2 = lane < HID ? tanhf(s) : 0.f;
  const int a = lane < A ? lane : A - 1;
  s = w[o.b3 + a];
#pragma unroll
  for (int i = 0; i < HID; ++i) s = fmaf(__shfl_sync(0xffffffffu, h2, i), w[o.w3 + i * A + a], s);
  *h1o = h1; *h2o = h2; *qo = s;            // q valid on lanes < A
}

template <typename TIn>
__global__ void __launch_bounds__(256)
mlp_forward_kernel(const TIn* __restrict__ x, int n, int IN, int A, MlpOff o, const float* __restrict__ params,
                   float* __restrict__ q, int32_t* __restrict__ argmax) {
  extern __shared__ float sm[];
  for (int t = threadIdx.x; t < o.total; t += blockDim.x) sm[t] = params[t];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (b >= n) return;
  float xv[MLP_MAX_IN], h1, h2, qv;
  mlp_forward_lane<TIn>(sm, o, x + (size_t)b * IN, IN, A, lane, xv, &h1, &h2, &qv);
  if (q && lane < A) q[(size_t)b * A + lane] = qv;
  if (argmax) {
    float best = __shfl_sync(0xffffffffu, qv, 0);
    int bi = 0;
    for (int a = 1; a < A; ++a) {
      const float v = __shfl_sync(0xffffffffu, qv, a);
      if (v > best) { best = v; bi = a; }
    }
    if (lane == 0) argmax[b] = bi;
  }
}

// Whole backward of SimpleNetwork in ONE CTA: both forwards, TD target, loss term, all gradients
// (batch mean, data term) into the flat gradient buffer.
template <typename TIn>
__global__ void __launch_bounds__(MLP_WARPS * 32)
mlp_backward_kernel(const TIn* __restrict__ states, const TIn* __restrict__ next_states,
                    const int32_t* __restrict__ actions, const double* __restrict__ rewards,
                    const uint8_t* __restrict__ terminals, int B, int IN, int A, MlpOff o,
                    const float* __restrict__ online, const float* __restrict__ target, double gamma, int loss_kind,
                    float* __restrict__ grad, float* __restrict__ qout, float* __restrict__ se_out,
                    long long* __restrict__ step) {
  extern __shared__ float sm[];
  float* w = sm;
  float* wt = sm + o.total;
  float* accs = sm + 2 * o.total;                   // [MLP_WARPS][total]
  for (int t = threadIdx.x; t < o.total; t += blockDim.x) { w[t] = online[t]; wt[t] = target[t]; }
  for (int t = threadIdx.x; t < MLP_WARPS * o.total; t += blockDim.x) accs[t] = 0.f;
  if (threadIdx.x == 0) *step += 1;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int j = lane < HID ? lane : HID - 1;
  float dw1[MLP_MAX_IN], dw2[HID], dw3[MLP_MAX_A];
  float db1 = 0.f, db2 = 0.f, db3 = 0.f;
#pragma unroll
  for (int i = 0; i < MLP_MAX_IN; ++i) dw1[i] = 0.f;
#pragma unroll
  for (int i = 0; i < HID; ++i) dw2[i] = 0.f;
#pragma unroll
  for (int i = 0; i < MLP_MAX_A; ++i) dw3[i] = 0.f;
  for (int b = warp; b < B; b += MLP_WARPS) {
    float xv[MLP_MAX_IN], xn[MLP_MAX_IN], h1, h2, qv, t1, t2, tq;
    mlp_forward_lane<TIn>(w, o, states + (size_t)b * IN, IN, A, lane, xv, &h1, &h2, &qv);
    mlp_forward_lane<TIn>(wt, o, next_states + (size_t)b * IN, IN, A, lane, xn, &t1, &t2, &tq);
    float maxq = lane < A ? tq : -INFINITY;
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) maxq = fmaxf(maxq, __shfl_xor_sync(0xffffffffu, maxq, s));
    const int act = actions[b];
    const float q_sel = __shfl_sync(0xffffffffu, qv, act);
    float e, g;
    td_terms(q_sel, maxq, rewards[b], terminals[b] != 0, gamma, B, loss_kind, &e, &g);
    if (lane == 0) se_out[b] = e;
    if (qout && lane < A) qout[(size_t)b * A + lane] = qv;
    // output layer
#pragma unroll
    for (int a = 0; a < MLP_MAX_A; ++a)
      if (a == act) dw3[a] = fmaf(h2, g, dw3[a]);
    if (lane == act) db3 += g;
    const float d2 = lane < HID ? g * w[o.w3 + j * A + act] * (1.f - h2 * h2) : 0.f;
#pragma unroll
    for (int i = 0; i < HID; ++i) dw2[i] = fmaf(__shfl_sync(0xffffffffu, h1, i), d2, dw2[i]);
    db2 += d2;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < HID; ++k) s = fmaf(__shfl_sync(0xffffffffu, d2, k), w[o.w2 + j * HID + k], s);
    const float d1 = lane < HID ? s * (1.f - h1 * h1) : 0.f;
#pragma unroll
    for (int i = 0; i < MLP_MAX_IN; ++i) dw1[i] = fmaf(xv[i], d1, dw1[i]);
    db1 += d1;
  }
  float* my = accs + warp * o.total;
  if (lane < HID) {
#pragma unroll
    for (int i = 0; i < MLP_MAX_IN; ++i)
      if (i < IN) my[o.w1 + i * HID + lane] = dw1[i];
    my[o.b1 + lane] = db1;
#pragma unroll
    for (int i = 0; i < HID; ++i) my[o.w2 + i * HID + lane] = dw2[i];
    my[o.b2 + lane] = db2;
#pragma unroll
    for (int a = 0; a < MLP_MAX_A; ++a)
      if (a < A) my[o.w3 + lane * A + a] = dw3[a];
  }
  if (lane < A) my[o.b3 + lane] = db3;
  __syncthreads();
  for (int t = threadIdx.x; t < o.total; t += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < MLP_WARPS; ++k) s += accs[k * o.total + t];
    grad[t] = s;
  }
}

}  // namespace dqn

// ================================================================================================
// Host side
// ================================================================================================
using namespace dqn;

namespace {
struct Tensor { int64_t off, numel; int rank; int64_t shape[4]; };

struct GemmLayer {           // conv1..3 + fc of the conv architectures
  int K, N, rows, in_elems, out_elems;
  int maxsplit;
  int maxsplit_tc;        // cap of the reduction split of the tensor-core weight-gradient kernel
  float* partial = nullptr;
  int w_tensor;              // index of the weight tensor (bias = +1)
};
}  // namespace

struct dqn_net_s {
  int device = 0, arch = 0, A = 0, MB = 0, in_dtype = 0, in_dim = 0, F = 0;
  int num_sms = 148;
  int64_t P = 0;
  std::vector<Tensor> tensors;
  float *online = nullptr, *target = nullptr, *grad = nullptr, *slot0 = nullptr, *slot1 = nullptr;
  GemmLayer L[4];
  float* act[2][4] = {};         // [pass][layer] pooled activations; layer 3 = fc output
  uint8_t* arg[2][3] = {};
  float* dact[4] = {};
  float* head_partial = nullptr; // [16][(F+1)*A]
  int head_nsplit = 1;
  float *se = nullptr, *g = nullptr, *loss_tmp = nullptr, *q_tmp = nullptr;
  long long* step = nullptr;
  void *st_states = nullptr, *st_next = nullptr;
  int32_t* st_actions = nullptr;
  double* st_rewards = nullptr;
  uint8_t* st_term = nullptr;
  size_t state_bytes = 0;        // bytes of one state
  dqn_hyper_t hyper;
  MlpOff mo;
  std::vector<void*> allocs;
  // data-parallel exchange over peer memory (dqn_net_dp_*)
  void* dp_block = nullptr;                    // [256 B flags][2][P] floats, cudaMalloc'ed, IPC-shared
  float* dp_grad = nullptr;                    // = dp_block + 256
  int dp_rank = 0, dp_world = 1;
  void* dp_peer_block[8] = {};                 // mapped exchange blocks of all ranks (own = dp_block)
  unsigned long long** dp_peer_flags_dev = nullptr;   // device array of the peers' flag pointers
  unsigned int* dp_counter = nullptr;                 // finished-block counter of the publishing reduce kernel
  unsigned long long* dp_err = nullptr;               // last step whose flag wait timed out (0 = none)
  long long dp_timeout_ns = 120LL * 1000000000LL;     // dqn_net_dp_set_timeout
  // weight-gradient kernels run on a forked stream, concurrently with the data-gradient chain (captured into the
  // step's CUDA graph as a parallel branch)
  cudaStream_t side = nullptr, side2 = nullptr;     // side2: head weight gradient, early half of the data-parallel exchange
  cudaStream_t side3 = nullptr;                     // fc weight gradient + loss
  bool no_resident = false;                         // DQN_NO_RESIDENT=1: im2col gathers from L2 instead of a shared-memory resident input block
  bool no_fused_fc = true;                          // DQN_FUSED_FC=1: fc forward / TD head / fc dgrad as one launch (experimental)
  bool fc_cluster = true;                           // B <= 32: fc forward / TD head / fc dgrad as one 8-CTA cluster launch (DQN_NO_FC_CLUSTER=1: off)
  long long* idx_cursor = nullptr;                  // set for the duration of a dqn_train_step_table call
  bool conv1_s2d = false;                           // reference conv net, bf16: conv1 forward / weight gradient in the space-to-depth formulation
  bool fuse_gather = true;                          // ... reading the frames straight from the replay ring in dqn_train_step (DQN_NO_FUSED_GATHER=1: off)
  cudaEvent_t ev_fork[3] = {}, ev_join = nullptr, ev_x = nullptr, ev_join2 = nullptr, ev_head = nullptr, ev_loss = nullptr;
  // bf16 tensor-core mode (precision == 1; DQN_ARCH_CNN only)
  int precision = 0;
  __nv_bfloat16* act_bf[2][4] = {};
  __nv_bfloat16* dact_bf[4] = {};
  __nv_bfloat16* wt_bf[2] = {};       // per parameter set: the 4 layers' transposed weights back to back
  __nv_bfloat16* wd_bf = nullptr;     // online only: permuted weights for dgrad (layers 1..3)
  size_t wt_off[4] = {}, wd_off[4] = {}, wt_total = 0, wd_total = 0;
  // optional per-kernel timing (bench.py's roofline figures): one event pair per stage
  bool timing = false;
  bool timing_sticky = false;         // brackets captured into a graph: every replay refreshes the same stages
  cudaEvent_t ev0[DQN_NUM_STAGES] = {}, ev1[DQN_NUM_STAGES] = {};
  bool stage_used[DQN_NUM_STAGES] = {};
  double stage_ms[DQN_NUM_STAGES] = {};
  int stage_count[DQN_NUM_STAGES] = {};
};

namespace {
// Brackets one kernel launch: counts it and, in timing mode, records its event pair.  While the stream is being
// captured into a CUDA graph the records become external event-record NODES of the graph, so a replay of that graph
// stamps every kernel as it runs inside the graph (no host launch latency inside the bracket; the nodes also cut the
// programmatic-dependent-launch overlap between neighbours, so a bracket is one kernel's own duration).
static void stage_record(cudaEvent_t ev, cudaStream_t st) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(st, &cs);
  if (cs == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(ev, st, cudaEventRecordExternal);
  else cudaEventRecord(ev, st);
}
struct StageScope {
  dqn_net_s* h; int id; cudaStream_t st;
  StageScope(dqn_net_s* h_, int id_, cudaStream_t st_) : h(h_), id(id_), st(st_) {
    dqn::count_launch();
    if (h->timing) stage_record(h->ev0[id], st);
  }
  ~StageScope() {
    if (h->timing) { stage_record(h->ev1[id], st); h->stage_used[id] = true; }
  }
};
}  // namespace

static const char* kStageNames[DQN_NUM_STAGES] = {
    "gather", "fwd_conv1", "fwd_conv2", "fwd_conv3", "fwd_fc", "td_head", "head_wgrad", "loss",
    "wgrad_fc", "dgrad_fc", "wgrad_conv3", "dgrad_conv3", "wgrad_conv2", "dgrad_conv2", "wgrad_conv1",
    "reduce_grad", "optimizer", "mlp_backward", "mlp_forward", "head_forward", "fc_td_fused", "launch_floor"};

static int refresh_bf16_weights(dqn_net_s* h, int set, cudaStream_t st);

static int dev_alloc(dqn_net_s* h, void** p, size_t bytes) {
  DQN_CUDA(cudaMalloc(p, bytes ? bytes : 16));
  DQN_CUDA(cudaMemset(*p, 0, bytes ? bytes : 16));
  h->allocs.push_back(*p);
  return 0;
}
#define NET_ALLOC(ptr, bytes)                                         \
  do {                                                                \
    int rc__ = dev_alloc(h, reinterpret_cast<void**>(&(ptr)), bytes); \
    if (rc__) { dqn_net_destroy(h); return rc__; }                    \
  } while (0)

static void add_tensor(dqn_net_s* h, std::initializer_list<int64_t> shape) {
  Tensor t{};
  t.off = h->P;
  t.rank = (int)shape.size();
  t.numel = 1;
  int i = 0;
  for (int64_t s : shape) { t.shape[i++] = s; t.numel *= s; }
  h->P += round_up(t.numel, 4);
  h->tensors.push_back(t);
}

extern "C" int dqn_net_create(dqn_net_t* out, int32_t arch, const int32_t* in_shape, int32_t in_rank,
                              int32_t num_actions, int32_t max_batch, int32_t in_dtype, int32_t device) {
  DQN_REQUIRE(out && in_shape, DQN_ERR_INVALID, "dqn_net_create: NULL argument");
  *out = nullptr;
  DQN_REQUIRE(arch == DQN_ARCH_SIMPLE || arch == DQN_ARCH_CNN || arch == DQN_ARCH_NATURE, DQN_ERR_INVALID,
              "Unsupported network type %d", arch);                          // network.py:46-47
  DQN_REQUIRE(num_actions >= 1, DQN_ERR_INVALID, "num_actions must be positive");
  DQN_REQUIRE(max_batch >= 1 && max_batch <= 4096, DQN_ERR_INVALID, "max_batch must be in [1, 4096]");
  if (arch == DQN_ARCH_SIMPLE) {
    DQN_REQUIRE(in_rank == 1, DQN_ERR_INVALID, "SimpleNetwork expects 1-d input");         // network.py:265-266
    DQN_REQUIRE(in_shape[0] >= 1 && in_shape[0] <= MLP_MAX_IN, DQN_ERR_UNSUPPORTED,
                "SimpleNetwork kernel supports 1..%d inputs, got %d", MLP_MAX_IN, in_shape[0]);
    DQN_REQUIRE(num_actions <= MLP_MAX_A, DQN_ERR_UNSUPPORTED, "SimpleNetwork kernel supports up to %d actions", MLP_MAX_A);
    DQN_REQUIRE(in_dtype == DQN_F32 || in_dtype == DQN_F64, DQN_ERR_INVALID, "SimpleNetwork input must be f32 or f64");
  } else {
    DQN_REQUIRE(in_rank == 3, DQN_ERR_INVALID, "ConvNetwork expects 3-d input");           // network.py:338-339
    DQN_REQUIRE(in_shape[0] == 84 && in_shape[1] == 84 && in_shape[2] == 4, DQN_ERR_UNSUPPORTED,
                "conv kernels are specialised for 84x84x4 input (the reference's fc layer fixes this too, "
                "network.py:372-373); got %dx%dx%d", in_shape[0], in_shape[1], in_shape[2]);
    DQN_REQUIRE(in_dtype == DQN_U8 || in_dtype == DQN_F32, DQN_ERR_INVALID, "ConvNetwork input must be u8 or f32");
    DQN_REQUIRE(num_actions <= 64, DQN_ERR_UNSUPPORTED, "at most 64 actions");
  }
  int ndev = 0;
  DQN_CUDA(cudaGetDeviceCount(&ndev));
  DQN_REQUIRE(device >= 0 && device < ndev, DQN_ERR_INVALID, "no CUDA device %d (found %d); there is no CPU path", device, ndev);
  DeviceGuard guard(device);
  DQN_REQUIRE(guard.ok, DQN_ERR_CUDA, "cannot select device %d", device);
  dqn_net_s* h = new (std::nothrow) dqn_net_s();
  DQN_REQUIRE(h, DQN_ERR_INVALID, "out of host memory");
  h->device = device; h->arch = arch; h->A = num_actions; h->MB = max_batch; h->in_dtype = in_dtype;
  cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
  memset(&h->hyper, 0, sizeof(h->hyper));
  h->hyper.gamma = 0.9; h->hyper.lr = 0.001f; h->hyper.reg_param = 0.001f; h->hyper.momentum = 0.9f;
  h->hyper.rmsprop_decay = 0.95f; h->hyper.optimizer = DQN_OPT_SGD; h->hyper.grad_scale = 1.f;   // main.py:30-39, 55
  const int A = num_actions, MB = max_batch;
  if (arch == DQN_ARCH_SIMPLE) {
    const int IN = in_shape[0];
    h->in_dim = IN;
    h->state_bytes = (size_t)IN * (in_dtype == DQN_F64 ? 8 : 4);
    add_tensor(h, {IN, HID}); add_tensor(h, {HID}); add_tensor(h, {HID, HID}); add_tensor(h, {HID});
    add_tensor(h, {HID, A}); add_tensor(h, {A});
    h->mo = MlpOff{(int)h->tensors[0].off, (int)h->tensors[1].off, (int)h->tensors[2].off, (int)h->tensors[3].off,
                   (int)h->tensors[4].off, (int)h->tensors[5].off, (int)h->P};
  } else {
    h->state_bytes = (size_t)84 * 84 * 4 * (in_dtype == DQN_U8 ? 1 : 4);
    const bool nat = arch == DQN_ARCH_NATURE;
    add_tensor(h, {8, 8, 4, 32}); add_tensor(h, {32});
    add_tensor(h, {4, 4, 32, 64}); add_tensor(h, {64});
    add_tensor(h, {3, 3, 64, 64}); add_tensor(h, {64});
    h->F = nat ? 512 : 256;
    const int flat = nat ? 3136 : 256;
    add_tensor(h, {flat, h->F}); add_tensor(h, {h->F});
    add_tensor(h, {h->F, A}); add_tensor(h, {A});
    auto setL = [&](int i, int K, int N, int rows, int in_e, int out_e) {
      GemmLayer& l = h->L[i];
      l.K = K; l.N = N; l.rows = rows; l.in_elems = in_e; l.out_elems = out_e; l.w_tensor = 2 * i;
      const int base = (K / BM) * (N / BN);
      int ms = (3 * h->num_sms + base - 1) / base;
      l.maxsplit = ms < 1 ? 1 : (ms > h->num_sms ? h->num_sms : ms);
      // tensor-core weight gradients tile the output as 128 x 64 (fc: 128 x 128): one wave of CTAs over the reduction splits
      const int base_tc = ((K + 127) / 128) * (i == 3 ? (N + 127) / 128 : (N + 63) / 64);
      l.maxsplit_tc = h->num_sms / base_tc < 1 ? 1 : h->num_sms / base_tc;
      if (l.maxsplit < l.maxsplit_tc) l.maxsplit = l.maxsplit_tc;     // the partial-sum scratch is sized by maxsplit
      if (i == 0) l.maxsplit = l.maxsplit_tc = h->num_sms;      // conv1's resident weight-gradient kernel: one CTA per SM
    };
    if (!nat) {
      setL(0, C1::K, C1::COUT, C1::ROWS, 84 * 84 * 4, C1::OUT_PER_SAMPLE);
      setL(1, C2::K, C2::COUT, C2::ROWS, C1::OUT_PER_SAMPLE, C2::OUT_PER_SAMPLE);
      setL(2, C3::K, C3::COUT, C3::ROWS, C2::OUT_PER_SAMPLE, C3::OUT_PER_SAMPLE);
      setL(3, F4::K, F4::COUT, F4::ROWS, C3::OUT_PER_SAMPLE, F4::OUT_PER_SAMPLE);
    } else {
      setL(0, N1::K, N1::COUT, N1::ROWS, 84 * 84 * 4, N1::OUT_PER_SAMPLE);
      setL(1, N2::K, N2::COUT, N2::ROWS, N1::OUT_PER_SAMPLE, N2::OUT_PER_SAMPLE);
      setL(2, N3::K, N3::COUT, N3::ROWS, N2::OUT_PER_SAMPLE, N3::OUT_PER_SAMPLE);
      setL(3, NF::K, NF::COUT, NF::ROWS, N3::OUT_PER_SAMPLE, NF::OUT_PER_SAMPLE);
    }
  }
  const size_t PB = (size_t)h->P * sizeof(float);
  NET_ALLOC(h->online, PB); NET_ALLOC(h->target, PB); NET_ALLOC(h->grad, PB);
  NET_ALLOC(h->slot0, PB); NET_ALLOC(h->slot1, PB);
  NET_ALLOC(h->se, (size_t)MB * 4); NET_ALLOC(h->g, (size_t)MB * 4); NET_ALLOC(h->loss_tmp, 16);
  NET_ALLOC(h->q_tmp, (size_t)MB * A * 4); NET_ALLOC(h->step, 16);
  NET_ALLOC(h->st_states, (size_t)MB * h->state_bytes); NET_ALLOC(h->st_next, (size_t)MB * h->state_bytes);
  NET_ALLOC(h->st_actions, (size_t)MB * 4); NET_ALLOC(h->st_rewards, (size_t)MB * 8); NET_ALLOC(h->st_term, (size_t)MB);
  if (arch != DQN_ARCH_SIMPLE) {
    for (int p = 0; p < 2; ++p)
      for (int l = 0; l < 4; ++l) {
        NET_ALLOC(h->act[p][l], (size_t)MB * h->L[l].out_elems * 4);
        if (l < 3) NET_ALLOC(h->arg[p][l], (size_t)MB * h->L[l].out_elems);
      }
    for (int l = 0; l < 4; ++l) {
      NET_ALLOC(h->dact[l], (size_t)MB * h->L[l].out_elems * 4);
      NET_ALLOC(h->L[l].partial, (size_t)h->L[l].maxsplit * (h->L[l].K + 1) * h->L[l].N * 4);
    }
    NET_ALLOC(h->head_partial, (size_t)16 * (h->F + 1) * A * 4);
    cudaError_t e = cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->side2, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->side3, cudaStreamNonBlocking);
    { const char* nr = getenv("DQN_NO_RESIDENT"); h->no_resident = nr && nr[0] == '1'; }
    { const char* nf = getenv("DQN_FUSED_FC"); h->no_fused_fc = !(nf && nf[0] == '1'); }
    { const char* ng = getenv("DQN_NO_FUSED_GATHER"); h->fuse_gather = !(ng && ng[0] == '1'); }
    { const char* nc = getenv("DQN_NO_FC_CLUSTER"); h->fc_cluster = !(nc && nc[0] == '1'); }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_head, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_loss, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_x, cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_join2, cudaEventDisableTiming);
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&h->ev_fork[i], cudaEventDisableTiming);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming);
    if (e != cudaSuccess) {
      int rc = set_err(DQN_ERR_CUDA, "dqn_net_create: %s", cudaGetErrorString(e));
      dqn_net_destroy(h);
      return rc;
    }
  }
  *out = h;
  return 0;
}

extern "C" int dqn_net_destroy(dqn_net_t h) {
  if (!h) return 0;
  DeviceGuard g(h->device);
  cudaDeviceSynchronize();
  for (void* p : h->allocs) cudaFree(p);
  for (int r = 0; r < 8; ++r)
    if (h->dp_peer_block[r] && h->dp_peer_block[r] != h->dp_block) cudaIpcCloseMemHandle(h->dp_peer_block[r]);
  if (h->dp_block) cudaFree(h->dp_block);
  if (h->side) cudaStreamDestroy(h->side);
  if (h->side2) cudaStreamDestroy(h->side2);
  if (h->side3) cudaStreamDestroy(h->side3);
  if (h->ev_head) cudaEventDestroy(h->ev_head);
  if (h->ev_loss) cudaEventDestroy(h->ev_loss);
  if (h->ev_x) cudaEventDestroy(h->ev_x);
  if (h->ev_join2) cudaEventDestroy(h->ev_join2);
  for (int i = 0; i < 3; ++i) if (h->ev_fork[i]) cudaEventDestroy(h->ev_fork[i]);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  for (int i = 0; i < DQN_NUM_STAGES; ++i) {
    if (h->ev0[i]) cudaEventDestroy(h->ev0[i]);
    if (h->ev1[i]) cudaEventDestroy(h->ev1[i]);
  }
  delete h;
  return 0;
}

extern "C" int dqn_net_param_count(dqn_net_t h, int64_t* padded_floats, int32_t* num_tensors) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "NULL handle");
  if (padded_floats) *padded_floats = h->P;
  if (num_tensors) *num_tensors = (int32_t)h->tensors.size();
  return 0;
}

extern "C" int dqn_net_tensor_info(dqn_net_t h, int32_t i, int64_t* offset, int64_t* numel, int32_t* rank, int64_t shape[4]) {
  DQN_REQUIRE(h && i >= 0 && i < (int)h->tensors.size(), DQN_ERR_INVALID, "tensor index out of range");
  const Tensor& t = h->tensors[i];
  if (offset) *offset = t.off;
  if (numel) *numel = t.numel;
  if (rank) *rank = t.rank;
  if (shape) for (int k = 0; k < 4; ++k) shape[k] = k < t.rank ? t.shape[k] : 1;
  return 0;
}

static float* buf_of(dqn_net_s* h, int which) {
  switch (which) {
    case DQN_BUF_ONLINE: return h->online;
    case DQN_BUF_TARGET: return h->target;
    case DQN_BUF_GRAD: return h->grad;
    case DQN_BUF_SLOT0: return h->slot0;
    case DQN_BUF_SLOT1: return h->slot1;
  }
  return nullptr;
}

extern "C" int dqn_net_buffer(dqn_net_t h, int32_t which, float** d_ptr, int64_t* floats) {
  DQN_REQUIRE(h && buf_of(h, which), DQN_ERR_INVALID, "unknown buffer %d", which);
  if (d_ptr) *d_ptr = buf_of(h, which);
  if (floats) *floats = h->P;
  return 0;
}

extern "C" int dqn_net_upload(dqn_net_t h, int32_t which, const float* host, int64_t floats) {
  DQN_REQUIRE(h && host && buf_of(h, which), DQN_ERR_INVALID, "upload: bad argument");
  DQN_REQUIRE(floats == h->P, DQN_ERR_INVALID, "upload: expected %lld floats, got %lld", (long long)h->P, (long long)floats);
  DeviceGuard g(h->device);
  DQN_CUDA(cudaDeviceSynchronize());
  DQN_CUDA(cudaMemcpy(buf_of(h, which), host, (size_t)floats * 4, cudaMemcpyHostToDevice));
  if (h->precision == 1 && (which == DQN_BUF_ONLINE || which == DQN_BUF_TARGET)) {
    int rc = refresh_bf16_weights(h, which == DQN_BUF_ONLINE ? 0 : 1, nullptr);
    if (rc) return rc;
    DQN_CUDA(cudaDeviceSynchronize());
  }
  return 0;
}

extern "C" int dqn_net_download(dqn_net_t h, int32_t which, float* host, int64_t floats) {
  DQN_REQUIRE(h && host && buf_of(h, which), DQN_ERR_INVALID, "download: bad argument");
  DQN_REQUIRE(floats == h->P, DQN_ERR_INVALID, "download: expected %lld floats, got %lld", (long long)h->P, (long long)floats);
  DeviceGuard g(h->device);
  DQN_CUDA(cudaDeviceSynchronize());
  DQN_CUDA(cudaMemcpy(host, buf_of(h, which), (size_t)floats * 4, cudaMemcpyDeviceToHost));
  return 0;
}

extern "C" int dqn_net_set_hyper(dqn_net_t h, const dqn_hyper_t* hyper, int32_t reset_slots) {
  DQN_REQUIRE(h && hyper, DQN_ERR_INVALID, "set_hyper: NULL argument");
  DQN_REQUIRE(hyper->optimizer >= DQN_OPT_SGD && hyper->optimizer <= DQN_OPT_FTRL, DQN_ERR_INVALID,
              "Unsupported optimizer %d", hyper->optimizer);                  // network.py:179-180
  DQN_REQUIRE(hyper->loss == DQN_LOSS_MSE || hyper->loss == DQN_LOSS_HUBER, DQN_ERR_INVALID, "unknown loss %d", hyper->loss);
  h->hyper = *hyper;
  if (h->hyper.grad_scale == 0.f) h->hyper.grad_scale = 1.f;
  if (reset_slots) {
    DeviceGuard g(h->device);
    float v0 = 0.f;
    if (hyper->optimizer == DQN_OPT_RMSPROP) v0 = 1.f;                         // TF: rms slot starts at 1
    if (hyper->optimizer == DQN_OPT_ADAGRAD || hyper->optimizer == DQN_OPT_FTRL) v0 = 0.1f;   // initial_accumulator_value
    DQN_CUDA(cudaDeviceSynchronize());
    fill_kernel<<<64, 256>>>(h->slot0, h->P, v0);
    fill_kernel<<<64, 256>>>(h->slot1, h->P, 0.f);
    DQN_CUDA(cudaGetLastError());
    DQN_CUDA(cudaMemset(h->step, 0, 16));
    // the exchange flags count steps too (only meaningful if every rank resets at the same point, behind a barrier)
    if (h->dp_block) DQN_CUDA(cudaMemset(h->dp_block, 0, 256));
    DQN_CUDA(cudaDeviceSynchronize());
  }
  return 0;
}

extern "C" int dqn_net_set_global_step(dqn_net_t h, int64_t step) {
  DQN_REQUIRE(h && step >= 0, DQN_ERR_INVALID, "set_global_step: bad argument");
  DeviceGuard g(h->device);
  long long v = step;
  DQN_CUDA(cudaDeviceSynchronize());
  DQN_CUDA(cudaMemcpy(h->step, &v, sizeof(v), cudaMemcpyHostToDevice));
  if (h->dp_block) {
    // the exchange flags count steps too ("rank r has published step s"): realign this rank's flag arrays, otherwise a
    // smaller restored step would make every wait pass at once.  Every rank must do this at the same point, with no
    // dqn_net_dp_step in flight on any rank (Network.load puts a barrier on both sides).
    unsigned long long f[32];
    for (int i = 0; i < 32; ++i) f[i] = (unsigned long long)step;
    DQN_CUDA(cudaMemcpy(h->dp_block, f, sizeof(f), cudaMemcpyHostToDevice));
    if (h->dp_err) DQN_CUDA(cudaMemset(h->dp_err, 0, 8));
  }
  return 0;
}

extern "C" int dqn_net_global_step(dqn_net_t h, int64_t* step, dqn_stream_t stream) {
  DQN_REQUIRE(h && step, DQN_ERR_INVALID, "global_step: NULL argument");
  DeviceGuard g(h->device);
  long long v = 0;
  DQN_CUDA(cudaMemcpyAsync(&v, h->step, sizeof(v), cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
  DQN_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  *step = v;
  return 0;
}

// ---------------------------------------------------------------------------------------------- launch helpers
// Cluster size (split of the reduction) for a launch with `tiles` output tiles and `chunks` reduction chunks:
// enough CTAs to cover the SMs, at least 2 chunks per CTA, at most the portable cluster limit.
static int pick_split(int tiles, int chunks, int num_sms) {
  int s = (num_sms + tiles - 1) / tiles;
  if (s > chunks / 2) s = chunks / 2;
  if (s > MAX_CLUSTER) s = MAX_CLUSTER;
  return s < 1 ? 1 : s;
}

template <typename Kern, typename Args>
static void launch_clustered(Kern kern, dim3 grid, int split, const Args& args, cudaStream_t st) {
  launch_pdl(kern, grid, dim3(NT, 1, 1), 0, st, (unsigned)split, args);
}

template <class G, typename TIn>
static void launch_fwd(const FwdArgs& a, int npass, int num_sms, cudaStream_t st) {
  const int mt = (a.B * G::ROWS + BM - 1) / BM, nt = G::COUT / BN;
  const int split = pick_split(mt * nt * npass, G::K / BK, num_sms);
  dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)(npass * split));
  launch_clustered(conv_fwd_kernel<G, TIn>, grid, split, a, st);
}

template <class G, typename TIn>
static void launch_wgrad(dqn_net_s* h, int l, const void* x, int B, cudaStream_t st, int* nsplit_out) {
  const GemmLayer& L = h->L[l];
  const int R = B * G::ROWS;
  const int tiles = (R + BK - 1) / BK;
  int ns = L.maxsplit < tiles ? L.maxsplit : tiles;
  const int per = (tiles + ns - 1) / ns;
  ns = (tiles + per - 1) / per;
  WgradArgs a{x, h->dact[l], h->act[0][l], l < 3 ? h->arg[0][l] : nullptr, L.partial, B, per * BK};
  dim3 grid(G::K / BM, G::COUT / BN, ns);
  {
    StageScope sc(h, DQN_STAGE_WGRAD_FC + 2 * (3 - l), st);
    launch_pdl(conv_wgrad_kernel<G, TIn>, grid, dim3(NT), 0, st, 1, a);
  }
  *nsplit_out = ns;
}

template <class G>
static void launch_dgrad(dqn_net_s* h, int l, int B, cudaStream_t st) {
  DgradArgs a{h->online + h->tensors[2 * l].off, h->dact[l], h->act[0][l], l < 3 ? h->arg[0][l] : nullptr, h->dact[l - 1], B};
  const int mt = (B * G::IH * G::IW + BM - 1) / BM, nt = G::CIN / BN;
  const int split = pick_split(mt * nt, G::KS * G::KS * G::COUT / BK, h->num_sms);
  dim3 grid((unsigned)mt, (unsigned)nt, (unsigned)split);
  StageScope sc(h, DQN_STAGE_DGRAD_FC + 2 * (3 - l), st);
  launch_clustered(conv_dgrad_kernel<G>, grid, split, a, st);
}

static FwdPass make_pass(dqn_net_s* h, const float* params, int pass, int l, const void* x) {
  FwdPass p;
  p.x = x;
  p.w = params + h->tensors[2 * l].off;
  p.bias = params + h->tensors[2 * l + 1].off;
  p.out = h->act[pass][l];
  p.arg = l < 3 ? h->arg[pass][l] : nullptr;
  return p;
}

// Runs the trunk (conv1..3 + fc) for 1 or 2 passes. x[p]: input states of pass p, params[p]: flat params.
static int run_trunk(dqn_net_s* h, int npass, const void* const x[2], const float* const params[2], const int pass_id[2],
                     int B, cudaStream_t st) {
  const bool nat = h->arch == DQN_ARCH_NATURE;
  const bool u8 = h->in_dtype == DQN_U8;
  for (int l = 0; l < 4; ++l) {
    FwdArgs a;
    a.B = B;
    for (int p = 0; p < npass; ++p)
      a.p[p] = make_pass(h, params[p], pass_id[p], l, l == 0 ? x[p] : (const void*)h->act[pass_id[p]][l - 1]);
    if (npass == 1) a.p[1] = a.p[0];
    StageScope sc(h, DQN_STAGE_FWD_CONV1 + l, st);
    switch (l) {
      case 0:
        if (nat) { if (u8) launch_fwd<N1, uint8_t>(a, npass, h->num_sms, st); else launch_fwd<N1, float>(a, npass, h->num_sms, st); }
        else     { if (u8) launch_fwd<C1, uint8_t>(a, npass, h->num_sms, st); else launch_fwd<C1, float>(a, npass, h->num_sms, st); }
        break;
      case 1: if (nat) launch_fwd<N2, float>(a, npass, h->num_sms, st); else launch_fwd<C2, float>(a, npass, h->num_sms, st); break;
      case 2: if (nat) launch_fwd<N3, float>(a, npass, h->num_sms, st); else launch_fwd<C3, float>(a, npass, h->num_sms, st); break;
      case 3: if (nat) launch_fwd<NF, float>(a, npass, h->num_sms, st); else launch_fwd<F4, float>(a, npass, h->num_sms, st); break;
    }
  }
  DQN_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------------------------- bf16 tensor-core path
template <typename Kern>
static void set_smem(Kern kern, size_t bytes) {
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

template <class G, typename TIn, int NTILE, int RSN, int PD>
static void launch_fwd_tc_cfg(const umma::FwdArgsT& a, int npass, cudaStream_t st) {
  constexpr size_t smem = umma::fwd_smem_bytes<TIn, NTILE, RSN>();
  static_assert(smem <= 227 * 1024, "ring does not fit in shared memory");
  static bool once = (set_smem(umma::umma_fwd_kernel<G, TIn, NTILE, RSN, PD>, smem), true);
  (void)once;
  dim3 grid((unsigned)((a.B * G::ROWS + umma::TM - 1) / umma::TM), G::COUT / NTILE, npass);
  launch_pdl(umma::umma_fwd_kernel<G, TIn, NTILE, RSN, PD>, grid, dim3(umma::TT), smem, st, 1, a);
}
template <class G, int NTILE, int NK, int NA>
static void launch_fwd_res(const umma::FwdArgsT& a, int npass, cudaStream_t st) {
  constexpr size_t smem = umma::fwd_res_smem_bytes<G, NTILE, NK, NA>();
  static_assert(smem <= 227 * 1024, "patch + operand slots + weights do not fit in shared memory");
  static bool once = (set_smem(umma::umma_fwd_res_kernel<G, NTILE, NK, NA>, smem), true);
  (void)once;
  dim3 grid((unsigned)((a.B * G::ROWS + umma::TM - 1) / umma::TM), G::COUT / NTILE, npass);
  launch_pdl(umma::umma_fwd_res_kernel<G, NTILE, NK, NA>, grid, dim3(umma::TT), smem, st, 1, a);
}
template <class G, int NTILE>
static void launch_fwd_res_u8(umma::FwdArgsT a, int npass, int num_sms, cudaStream_t st) {
  constexpr size_t smem = umma::fwd_res_u8_smem_bytes<G, NTILE>();
  static bool once = (set_smem(umma::umma_fwd_res_u8_kernel<G, NTILE>, smem), true);
  (void)once;
  // one wave of CTAs (two fit per SM), each walking over consecutive tiles of the combined (pass, tile) space
  const int total = ((a.B * G::ROWS + umma::TM - 1) / umma::TM) * npass;
  const int per = (total + 2 * num_sms - 1) / (2 * num_sms);
  a.npass = npass;
  a.tiles_per_cta = per;
  dim3 grid((unsigned)((total + per - 1) / per), 1, 1);
  launch_pdl(umma::umma_fwd_res_u8_kernel<G, NTILE>, grid, dim3(umma::TT), smem, st, 1, a);
}
// DQN_S2D_TRACE=1: CTA 0 of the conv1 forward kernel stamps clock64() per (tile, pipeline event) here (dqn_debug_trace)
static __device__ unsigned long long g_s2d_trace[60 * 8 + 160 * 4 + 16 + 64];     // + 64 data-parallel exchange stamps
static unsigned long long* s2d_trace_ptr() {
  static unsigned long long* p = nullptr;
  static bool init = false;
  if (!init) {
    init = true;
    const char* e = getenv("DQN_S2D_TRACE");
    if (e && e[0] == '1') cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_s2d_trace);
  }
  return p;
}
extern "C" int dqn_debug_trace(unsigned long long* host, int32_t n) {
  DQN_REQUIRE(host && n > 0 && n <= 60 * 8 + 160 * 4 + 16 + 64, DQN_ERR_INVALID, "debug_trace: bad argument");
  DQN_CUDA(cudaDeviceSynchronize());
  DQN_CUDA(cudaMemcpyFromSymbol(host, g_s2d_trace, (size_t)n * 8));
  return 0;
}

// conv1 in the space-to-depth formulation (conv1_s2d.cuh): persistent, one CTA per SM
static void launch_conv1_s2d_fwd(dqn_net_s* h, const umma::s2d::Src& src, const int pass_id[2], int npass, int B, cudaStream_t st) {
  namespace S = umma::s2d;
  static bool once = (set_smem(S::conv1_s2d_fwd_kernel<false>, S::FWD_SMEM), set_smem(S::conv1_s2d_fwd_kernel<true>, S::FWD_SMEM), true);
  (void)once;
  S::FwdArgs a;
  a.src = src;
  a.B = B;
  a.npass = npass;
  a.trace = s2d_trace_ptr();
  for (int p = 0; p < 2; ++p) {
    const int id = pass_id[p < npass ? p : 0];
    const float* params = id == 0 ? h->online : h->target;
    a.wt[p] = h->wt_bf[id] + h->wt_off[0];
    a.bias[p] = params + h->tensors[1].off;
    a.out[p] = h->act_bf[id][0];
    a.arg[p] = h->arg[id][0];
  }
  int grid = npass == 2 ? (h->num_sms & ~1) : h->num_sms;
  if (grid > npass * B) grid = npass * B;
  if (src.x[0]) launch_pdl(S::conv1_s2d_fwd_kernel<false>, dim3(grid), dim3(S::FWD_WARPS * 32), S::FWD_SMEM, st, 1, a);
  else launch_pdl(S::conv1_s2d_fwd_kernel<true>, dim3(grid), dim3(S::FWD_WARPS * 32), S::FWD_SMEM, st, 1, a);
}

static void launch_conv1_s2d_wgrad(dqn_net_s* h, const umma::s2d::Src& src, int B, cudaStream_t st, int* nsplit_out) {
  namespace S = umma::s2d;
  static bool once = (set_smem(S::conv1_s2d_wgrad_kernel<false>, S::WG_SMEM), set_smem(S::conv1_s2d_wgrad_kernel<true>, S::WG_SMEM), true);
  (void)once;
  S::WgradArgs a;
  a.src = src;
  a.dactg = h->dact_bf[0];
  a.arg = h->arg[0][0];
  a.partial = h->L[0].partial;
  a.B = B;
  a.trace = h->dp_world > 1 ? nullptr : s2d_trace_ptr();      // (the stamps share their slots with the data-parallel exchange stamps)
  const int grid = B < h->num_sms ? B : h->num_sms;
  {
    StageScope sc(h, DQN_STAGE_WGRAD_CONV1, st);
    if (src.x[0]) launch_pdl(S::conv1_s2d_wgrad_kernel<false>, dim3(grid), dim3(S::NWARPS * 32), S::WG_SMEM, st, 1, a);
    else launch_pdl(S::conv1_s2d_wgrad_kernel<true>, dim3(grid), dim3(S::NWARPS * 32), S::WG_SMEM, st, 1, a);
  }
  *nsplit_out = grid;
}

// Ring depth: a grid of at most one CTA per SM is latency-bound (deep ring, loads far ahead of the MMAs); a grid of
// several waves wants two CTAs resident per SM (small ring).  (Splitting the reduction of a tile over a cluster with a
// DSMEM reduction was measured slower at B=32: the cluster launch and its two cluster barriers cost more than the
// shorter main loop saves.)
template <class G, typename TIn, int NTILE, int RS_DEEP, int PD_DEEP>
static void launch_fwd_tc(const umma::FwdArgsT& a, int npass, int num_sms, cudaStream_t st) {
  const long ctas = (long)((a.B * G::ROWS + umma::TM - 1) / umma::TM) * (G::COUT / NTILE) * npass;
  if (ctas <= 2L * num_sms) launch_fwd_tc_cfg<G, TIn, NTILE, RS_DEEP, PD_DEEP>(a, npass, st);
  else launch_fwd_tc_cfg<G, TIn, NTILE, 4, 2>(a, npass, st);
}

template <class G, typename TIn, int NTILE>
static void launch_wgrad_tc(dqn_net_s* h, int l, const void* x, int B, cudaStream_t st, int* nsplit_out) {
  constexpr size_t smem = umma::wgrad_smem_bytes<G, TIn, NTILE>();
  static bool once = (set_smem(umma::umma_wgrad_kernel<G, TIn, NTILE>, smem), true);
  (void)once;
  const GemmLayer& L = h->L[l];
  const int R = B * G::ROWS;
  const int tiles = (R + umma::TK - 1) / umma::TK;
  int ns = L.maxsplit_tc < tiles ? L.maxsplit_tc : tiles;
  const int per = (tiles + ns - 1) / ns;
  ns = (tiles + per - 1) / per;
  umma::WgradArgsT a{x, h->dact_bf[l], l < 3 ? h->arg[0][l] : nullptr, L.partial, B, per * umma::TK};
  dim3 grid((G::K + umma::TM - 1) / umma::TM, G::COUT / NTILE, ns);
  {
    StageScope sc(h, DQN_STAGE_WGRAD_FC + 2 * (3 - l), st);
    launch_pdl(umma::umma_wgrad_kernel<G, TIn, NTILE>, grid, dim3(umma::TT), smem, st, 1, a);
  }
  *nsplit_out = ns;
}

// conv1 weight gradient with the frames resident in shared memory: one CTA per run of consecutive 128-row tiles
template <class G>
static void launch_wgrad_res_u8(dqn_net_s* h, const void* x, int B, cudaStream_t st, int* nsplit_out) {
  constexpr size_t smem = umma::WgradResU8<G>::SMEM;
  static_assert(smem <= 227 * 1024, "does not fit in shared memory");
  static bool once = (set_smem(umma::umma_wgrad_res_u8_kernel<G>, smem), true);
  (void)once;
  const GemmLayer& L = h->L[0];
  const int tiles = (B * G::ROWS + umma::TM - 1) / umma::TM;
  int ns = L.maxsplit < tiles ? L.maxsplit : tiles;
  if (ns > h->num_sms) ns = h->num_sms;
  int per = (tiles + ns - 1) / ns;
  if (per < 2 && tiles > 32) per = 2;      // every partial is read back by the reduce / optimizer kernels: not too many of them
  ns = (tiles + per - 1) / per;
  umma::WgradArgsT a{x, h->dact_bf[0], h->arg[0][0], L.partial, B, per * umma::TM};
  {
    StageScope sc(h, DQN_STAGE_WGRAD_CONV1, st);
    launch_pdl(umma::umma_wgrad_res_u8_kernel<G>, dim3(ns), dim3(umma::TT), smem, st, 1, a);
  }
  *nsplit_out = ns;
}

template <class G, int NTILE>
static void launch_dgrad_tc(dqn_net_s* h, int l, int B, cudaStream_t st) {
  constexpr size_t smem = umma::dgrad_smem_bytes<G, NTILE>();
  static bool once = (set_smem(umma::umma_dgrad_kernel<G, NTILE>, smem), true);
  (void)once;
  umma::DgradArgsT a{h->wd_bf + h->wd_off[l], h->dact_bf[l], l < 3 ? h->arg[0][l] : nullptr, h->act_bf[0][l - 1],
                     h->dact_bf[l - 1], B, {0, 0, 0, 0, 0}};
  unsigned mtiles = (unsigned)((B * G::IH * G::IW + umma::TM - 1) / umma::TM);
  if constexpr (G::ST == 2) {          // tiles are grouped by input-pixel parity class
    int acc = 0;
    for (int c = 0; c < 4; ++c) {
      a.cls_tile[c] = acc;
      acc += (B * umma::DgradClasses<G>::count(c) + umma::TM - 1) / umma::TM;
    }
    a.cls_tile[4] = acc;
    mtiles = (unsigned)acc;
  }
  dim3 grid(mtiles, G::CIN / NTILE, 1);
  StageScope sc(h, DQN_STAGE_DGRAD_FC + 2 * (3 - l), st);
  launch_pdl(umma::umma_dgrad_kernel<G, NTILE>, grid, dim3(umma::TT), smem, st, 1, a);
}

template <class G, int NTILE, int NK, int NA>
static void launch_dgrad_res(dqn_net_s* h, int l, int B, cudaStream_t st) {
  constexpr size_t smem = umma::dgrad_res_smem_bytes<G, NTILE, NK, NA>();
  static_assert(smem <= 227 * 1024, "patch + operand slots + weights do not fit in shared memory");
  static bool once = (set_smem(umma::umma_dgrad_res_kernel<G, NTILE, NK, NA>, smem), true);
  (void)once;
  umma::DgradArgsT a{h->wd_bf + h->wd_off[l], h->dact_bf[l], h->arg[0][l], h->act_bf[0][l - 1], h->dact_bf[l - 1], B, {0, 0, 0, 0, 0}};
  unsigned mtiles = (unsigned)((B * G::IH * G::IW + umma::TM - 1) / umma::TM);
  if constexpr (G::ST == 2) {          // tiles are grouped by input-pixel parity class
    int acc = 0;
    for (int c = 0; c < 4; ++c) {
      a.cls_tile[c] = acc;
      acc += (B * umma::DgradClasses<G>::count(c) + umma::TM - 1) / umma::TM;
    }
    a.cls_tile[4] = acc;
    mtiles = (unsigned)acc;
  }
  dim3 grid(mtiles, G::CIN / NTILE, 1);
  StageScope sc(h, DQN_STAGE_DGRAD_FC + 2 * (3 - l), st);
  launch_pdl(umma::umma_dgrad_res_kernel<G, NTILE, NK, NA>, grid, dim3(umma::TT), smem, st, 1, a);
}

// fp32 master weights of one parameter set -> bf16 operand copies
static int refresh_bf16_weights(dqn_net_s* h, int set, cudaStream_t st) {
  CvtArgs c;
  const float* base = set == 0 ? h->online : h->target;
  int maxn = 0;
  for (int l = 0; l < 4; ++l) {
    const GemmLayer& L = h->L[l];
    c.l[l].w = base + h->tensors[2 * l].off;
    c.l[l].wt = h->wt_bf[set] + h->wt_off[l];
    c.l[l].wd = (set == 0 && l >= 1) ? h->wd_bf + h->wd_off[l] : nullptr;
    c.l[l].K = L.K; c.l[l].N = L.N;
    c.l[l].CIN = l == 0 ? 4 : (l == 1 ? 32 : (l == 2 ? 64 : L.K));
    c.l[l].kperm = c.l[l].f16 = (l == 0 && h->conv1_s2d) ? 1 : 0;
    maxn = L.K * L.N > maxn ? L.K * L.N : maxn;
  }
  dim3 grid((maxn + 255) / 256, 4);
  dqn::count_launch();
  weights_to_bf16_kernel<<<grid, 256, 0, st>>>(c);
  DQN_CUDA(cudaGetLastError());
  return 0;
}

static umma::s2d::Src stack_src(const void* x0, const void* x1) {
  umma::s2d::Src s;
  memset(&s, 0, sizeof(s));
  s.x[0] = static_cast<const uint8_t*>(x0);
  s.x[1] = static_cast<const uint8_t*>(x1);
  return s;
}

static int run_trunk_tc(dqn_net_s* h, int npass, const void* const x[2], const int pass_id[2], int B, cudaStream_t st,
                        int nlayers = 4, const umma::s2d::Src* ring = nullptr) {
  for (int l = 0; l < nlayers; ++l) {
    if (l == 0 && h->conv1_s2d) {
      StageScope sc(h, DQN_STAGE_FWD_CONV1, st);
      launch_conv1_s2d_fwd(h, ring ? *ring : stack_src(x[0], x[npass == 2 ? 1 : 0]), pass_id, npass, B, st);
      continue;
    }
    umma::FwdArgsT a;
    a.B = B;
    a.npass = npass;
    a.tiles_per_cta = 1;
    for (int p = 0; p < npass; ++p) {
      const int id = pass_id[p];
      const float* params = id == 0 ? h->online : h->target;
      a.p[p].x = l == 0 ? x[p] : (const void*)h->act_bf[id][l - 1];
      a.p[p].wt = h->wt_bf[id] + h->wt_off[l];
      a.p[p].bias = params + h->tensors[2 * l + 1].off;
      a.p[p].out = h->act_bf[id][l];
      a.p[p].out_f32 = l == 3 ? h->act[id][3] : nullptr;
      a.p[p].arg = l < 3 ? h->arg[id][l] : nullptr;
    }
    if (npass == 1) a.p[1] = a.p[0];
    StageScope sc(h, DQN_STAGE_FWD_CONV1 + l, st);
    if (h->arch == DQN_ARCH_NATURE) {
      // Nature variant (VALID convs, no pooling, fc 512): conv1 and fc through the ring kernel, conv2/conv3 resident
      switch (l) {
        case 0: launch_fwd_tc<N1, uint8_t, 32, 4, 4>(a, npass, h->num_sms, st); break;
        case 1: launch_fwd_res<N2, 64, 2, 2>(a, npass, st); break;
        case 2: launch_fwd_res<N3, 64, 3, 2>(a, npass, st); break;
        case 3:
          if (B <= 128) launch_fwd_tc<NF, __nv_bfloat16, 64, 4, 2>(a, npass, h->num_sms, st);
          else launch_fwd_tc<NF, __nv_bfloat16, 256, 4, 2>(a, npass, h->num_sms, st);
          break;
      }
      continue;
    }
    switch (l) {
      case 0:
        if (h->no_resident) launch_fwd_tc<C1, uint8_t, 32, 4, 4>(a, npass, h->num_sms, st);
        else launch_fwd_res_u8<C1, 32>(a, npass, h->num_sms, st);
        break;
      case 1:
        if (h->no_resident) launch_fwd_tc<C2, __nv_bfloat16, 64, 8, 4>(a, npass, h->num_sms, st);
        else launch_fwd_res<C2, 64, 2, 3>(a, npass, st);
        break;
      case 2:
        if (h->no_resident) launch_fwd_tc<C3, __nv_bfloat16, 64, 8, 4>(a, npass, h->num_sms, st);
        else launch_fwd_res<C3, 64, 3, 2>(a, npass, st);
        break;
      case 3:          // few rows: narrow tiles spread the weight traffic over more SMs (a lone CTA is bound by its SM's L2 port)
        // 32-column tiles at every batch size: the layer is tiny (256x256 weights), so its time is the latency of ONE
        // CTA's chain of loads; more, narrower CTAs shorten it (B=512: 10.4 -> 4.0 us against 256-column tiles)
        launch_fwd_tc<F4, __nv_bfloat16, 32, 4, 4>(a, npass, h->num_sms, st);
        break;
    }
  }
  DQN_CUDA(cudaGetLastError());
  return 0;
}

static void fill_segs(dqn_net_s* h, OptArgs* o, const int nsplit[4], bool from_flat_grad) {
  memset(o, 0, sizeof(*o));
  if (h->arch == DQN_ARCH_SIMPLE) {
    o->nseg = 3;
    for (int i = 0; i < 3; ++i) {
      const Tensor& w = h->tensors[2 * i];
      const Tensor& b = h->tensors[2 * i + 1];
      o->seg[i] = Seg{(long long)w.off, (int)(w.numel + b.numel), (int)w.numel, h->grad + w.off, 1, 0, nullptr, nullptr, 0, 0, 0, 0, 0};
    }
  } else {
    o->nseg = 5;
    for (int l = 0; l < 4; ++l) {
      const Tensor& w = h->tensors[2 * l];
      const GemmLayer& L = h->L[l];
      Seg s{(long long)w.off, L.K * L.N + L.N, l == 3 ? L.K * L.N : 0, L.partial, nsplit[l], (long long)(L.K + 1) * L.N,
            nullptr, nullptr, 0, 0, 0, 0, 0};
      if (from_flat_grad) { s.partial = h->grad + w.off; s.nsplit = 1; s.stride = 0; }
      if (h->precision == 1) {
        s.wt = h->wt_bf[0] + h->wt_off[l];
        s.wd = l >= 1 ? h->wd_bf + h->wd_off[l] : nullptr;
        s.K = L.K; s.N = L.N;
        s.CIN = l == 0 ? 4 : (l == 1 ? 32 : (l == 2 ? 64 : L.K));
        s.kperm = s.f16 = (l == 0 && h->conv1_s2d) ? 1 : 0;
      }
      o->seg[l] = s;
    }
    const Tensor& w5 = h->tensors[8];
    Seg s{(long long)w5.off, (h->F + 1) * h->A, h->F * h->A, h->head_partial, h->head_nsplit,
          (long long)(h->F + 1) * h->A, nullptr, nullptr, 0, 0, 0, 0, 0};
    if (from_flat_grad) { s.partial = h->grad + w5.off; s.nsplit = 1; s.stride = 0; }
    o->seg[4] = s;
  }
  o->w = h->online; o->s0 = h->slot0; o->s1 = h->slot1; o->step = h->step;
  o->world = 1; o->P = h->P;
  o->lr = h->hyper.lr; o->reg = h->hyper.reg_param; o->momentum = h->hyper.momentum; o->decay = h->hyper.rmsprop_decay;
  o->gscale = h->hyper.grad_scale; o->opt = h->hyper.optimizer;
  o->idx_cursor = h->idx_cursor;
}

// data-parallel fields of the optimizer / exchange kernels' arguments
static void set_dp_args(dqn_net_s* h, OptArgs* o, int flag_slot) {
  o->world = h->dp_world;
  for (int r = 0; r < h->dp_world; ++r)
    o->peer_grad[r] = reinterpret_cast<const float*>(static_cast<const char*>(h->dp_peer_block[r]) + 256);
  o->flags = static_cast<const unsigned long long*>(h->dp_block);
  o->flag_slot = flag_slot;
  o->P = h->P;
  o->timeout_ns = h->dp_timeout_ns;
  o->err = h->dp_err;
  o->trace = s2d_trace_ptr();
}

static int max_seg_len(const OptArgs& o) {
  int m = 0;
  for (int i = 0; i < o.nseg; ++i) m = o.seg[i].len > m ? o.seg[i].len : m;
  return m;
}

extern "C" int dqn_net_forward(dqn_net_t h, int32_t which, const void* d_x, int32_t n, float* d_q, int32_t* d_argmax,
                               dqn_stream_t stream) {
  DQN_REQUIRE(h && d_x, DQN_ERR_INVALID, "forward: NULL argument");
  DQN_REQUIRE(which == DQN_BUF_ONLINE || which == DQN_BUF_TARGET, DQN_ERR_INVALID, "forward: which must be online or target");
  DQN_REQUIRE(n >= 1 && n <= h->MB, DQN_ERR_INVALID, "forward: n=%d outside [1, max_batch=%d]", n, h->MB);
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float* params = which == DQN_BUF_ONLINE ? h->online : h->target;
  if (h->arch == DQN_ARCH_SIMPLE) {
    const int blocks = (n + 7) / 8;
    const size_t smem = (size_t)h->mo.total * 4;
    StageScope sc(h, DQN_STAGE_MLP_FORWARD, st);
    if (h->in_dtype == DQN_F64)
      mlp_forward_kernel<double><<<blocks, 256, smem, st>>>(static_cast<const double*>(d_x), n, h->in_dim, h->A, h->mo, params, d_q, d_argmax);
    else
      mlp_forward_kernel<float><<<blocks, 256, smem, st>>>(static_cast<const float*>(d_x), n, h->in_dim, h->A, h->mo, params, d_q, d_argmax);
  } else {
    const int pass = which == DQN_BUF_ONLINE ? 0 : 1;
    const void* xs[2] = {d_x, d_x};
    const float* ps[2] = {params, params};
    const int ids[2] = {pass, pass};
    int rc = h->precision == 1 ? run_trunk_tc(h, 1, xs, ids, n, st) : run_trunk(h, 1, xs, ps, ids, n, st);
    if (rc) return rc;
    StageScope sc(h, DQN_STAGE_HEAD_FORWARD, st);
    head_forward_kernel<<<(n + 7) / 8, 256, 0, st>>>(h->act[pass][3], params + h->tensors[8].off, params + h->tensors[9].off,
                                                     n, h->F, h->A, d_q, d_argmax);
  }
  DQN_CUDA(cudaGetLastError());
  return 0;
}

// Forward of both nets + TD + full backward.  mode 0: fused optimizer; 1: materialise flat gradient.
// ring != null (conv1_s2d only): the conv1 kernels read the frames of the sampled transitions from the replay ring
// themselves and the forward kernel writes actions / rewards / terminals (d_states / d_next are not used).
static int run_step(dqn_net_s* h, const void* d_states, const void* d_next, const int32_t* d_actions,
                    const double* d_rewards, const uint8_t* d_term, int k, float* d_loss, float* d_q, int mode,
                    int include_reg, cudaStream_t st, const umma::s2d::Src* ring = nullptr) {
  DQN_REQUIRE((ring || (d_states && d_next)) && d_actions && d_rewards && d_term, DQN_ERR_INVALID, "train: NULL batch pointer");
  DQN_REQUIRE(k >= 1 && k <= h->MB, DQN_ERR_INVALID, "train: minibatch %d outside [1, max_batch=%d]", k, h->MB);
  int nsplit[4] = {1, 1, 1, 1};
  OptArgs oa;
  if (h->arch == DQN_ARCH_SIMPLE) {
    const size_t smem = (size_t)(2 + MLP_WARPS) * h->mo.total * 4;
    {
    StageScope sc(h, DQN_STAGE_MLP_BACKWARD, st);
    if (h->in_dtype == DQN_F64)
      mlp_backward_kernel<double><<<1, MLP_WARPS * 32, smem, st>>>(
          static_cast<const double*>(d_states), static_cast<const double*>(d_next), d_actions, d_rewards, d_term, k,
          h->in_dim, h->A, h->mo, h->online, h->target, h->hyper.gamma, h->hyper.loss, h->grad, d_q, h->se, h->step);
    else
      mlp_backward_kernel<float><<<1, MLP_WARPS * 32, smem, st>>>(
          static_cast<const float*>(d_states), static_cast<const float*>(d_next), d_actions, d_rewards, d_term, k,
          h->in_dim, h->A, h->mo, h->online, h->target, h->hyper.gamma, h->hyper.loss, h->grad, d_q, h->se, h->step);
    }
    DQN_CUDA(cudaGetLastError());
    fill_segs(h, &oa, nsplit, true);
    if (d_loss) {
      StageScope sc(h, DQN_STAGE_LOSS, st);
      loss_kernel<<<1, LOSS_THREADS, 0, st>>>(oa, h->se, k, d_loss);
    }
    if (mode == 1 && include_reg) {
      // grad already holds the data term; add reg*w in place through the reduce kernel (partial == grad)
      dim3 grid((max_seg_len(oa) + 255) / 256, oa.nseg);
      StageScope sc(h, DQN_STAGE_REDUCE_GRAD, st);
      reduce_grad_kernel<<<grid, 256, 0, st>>>(oa, h->grad, h->hyper.reg_param, 0, PublishArgs{nullptr, nullptr, 0, 1, 0, 0x43210, 0});
    }
    if (mode == 2) {
      dim3 grid((max_seg_len(oa) + 255) / 256, oa.nseg);
      StageScope sc(h, DQN_STAGE_REDUCE_GRAD, st);
      reduce_grad_kernel<<<grid, 256, 0, st>>>(oa, h->dp_grad, 0.f, 1,      // copy into this step's exchange buffer + publish
                                               PublishArgs{h->dp_peer_flags_dev, h->dp_counter + 1, h->dp_rank, h->dp_world, 1, 0x43210, dp_sys_release()});
    }
  } else {
    const bool nat = h->arch == DQN_ARCH_NATURE;
    const bool u8 = h->in_dtype == DQN_U8;
    const void* xs[2] = {d_states, d_next};
    const float* ps[2] = {h->online, h->target};
    const int ids[2] = {0, 1};
    // reference architecture in tensor-core mode: fc forward (both nets) + head + TD + fc data gradient are one kernel
    const bool fc_ok = h->precision == 1 && !nat && h->F == umma::FCF && h->A <= umma::FC_AP;
    const bool fused_cl = fc_ok && h->fc_cluster && h->no_fused_fc && k <= umma::fcc::MAXB;      // one cluster launch (small minibatches)
    const bool fused = fused_cl || (fc_ok && !h->no_fused_fc);
    int rc = h->precision == 1 ? run_trunk_tc(h, 2, xs, ids, k, st, fused ? 3 : 4, ring) : run_trunk(h, 2, xs, ps, ids, k, st);
    if (rc) return rc;
    const float* w5 = h->online + h->tensors[8].off;
    cudaStream_t sd = h->side;
    cudaStream_t sh = h->side3;                        // fc weight gradient, loss
    if (fused) {
      static bool once = (set_smem(umma::fc_td_fused_kernel, umma::FC_SMEM), true);
      (void)once;
      umma::FcFusedArgs fa;
      fa.x[0] = h->act_bf[0][2]; fa.x[1] = h->act_bf[1][2];
      fa.wt[0] = h->wt_bf[0] + h->wt_off[3]; fa.wt[1] = h->wt_bf[1] + h->wt_off[3];
      fa.b4[0] = h->online + h->tensors[7].off; fa.b4[1] = h->target + h->tensors[7].off;
      fa.w5[0] = w5; fa.w5[1] = h->target + h->tensors[8].off;
      fa.b5[0] = h->online + h->tensors[9].off; fa.b5[1] = h->target + h->tensors[9].off;
      fa.wd = h->wd_bf + h->wd_off[3];
      fa.actions = d_actions; fa.rewards = d_rewards; fa.terminals = d_term;
      fa.a4 = h->act[0][3]; fa.qout = d_q; fa.se = h->se; fa.g = h->g;
      fa.dh_bf = h->dact_bf[3]; fa.dx = h->dact_bf[2];
      fa.step = h->step; fa.gamma = h->hyper.gamma; fa.B = k; fa.A = h->A; fa.loss_kind = h->hyper.loss;
      fa.trace = s2d_trace_ptr();
      StageScope sc(h, DQN_STAGE_FC_TD_FUSED, st);
      if (fused_cl) {
        static bool once_cl = (set_smem(umma::fcc::fc_td_cluster_kernel, umma::fcc::SMEM), true);
        (void)once_cl;
        launch_pdl(umma::fcc::fc_td_cluster_kernel, dim3(1, 1, umma::fcc::CL), dim3(umma::fcc::THREADS), umma::fcc::SMEM, st, umma::fcc::CL, fa);
      } else {
        launch_pdl(umma::fc_td_fused_kernel, dim3((k + umma::TM - 1) / umma::TM), dim3(umma::TT), umma::FC_SMEM, st, 1, fa);
      }
    } else {
    StageScope sc(h, DQN_STAGE_TD_HEAD, st);
    if (h->F == 256)
      launch_pdl(td_head_kernel<8>, dim3((k + 7) / 8), dim3(256), (size_t)2 * (256 * h->A + h->A) * 4, st, 1, h->act[0][3], h->act[1][3], w5, h->online + h->tensors[9].off,
                                                     h->target + h->tensors[8].off, h->target + h->tensors[9].off,
                                                     d_actions, d_rewards, d_term, k, h->A, h->hyper.gamma, h->hyper.loss,
                                                     d_q, h->se, h->g, h->dact[3],
                                                     h->precision == 1 ? h->dact_bf[3] : nullptr, h->step);
    else {
      static bool once = (set_smem(td_head_kernel<16>, 160 * 1024), true);     // 2 x 512 x A floats can exceed the 48 KB default
      (void)once;
      launch_pdl(td_head_kernel<16>, dim3((k + 7) / 8), dim3(256), (size_t)2 * (512 * h->A + h->A) * 4, st, 1, h->act[0][3], h->act[1][3], w5, h->online + h->tensors[9].off,
                                                      h->target + h->tensors[8].off, h->target + h->tensors[9].off,
                                                      d_actions, d_rewards, d_term, k, h->A, h->hyper.gamma, h->hyper.loss,
                                                      d_q, h->se, h->g, h->dact[3],
                                                      h->precision == 1 ? h->dact_bf[3] : nullptr, h->step);
    }
    }
    // fork: head + fc weight gradients on `sh`, conv3/conv2 weight gradients on `sd`, data-gradient chain + conv1 wgrad on `st`
    DQN_CUDA(cudaEventRecord(h->ev_fork[0], st));
    DQN_CUDA(cudaStreamWaitEvent(sd, h->ev_fork[0], 0));
    DQN_CUDA(cudaStreamWaitEvent(sh, h->ev_fork[0], 0));
    DQN_CUDA(cudaStreamWaitEvent(h->side2, h->ev_fork[0], 0));
    {
    // the head weight gradient gets a branch of its own (in data-parallel mode the early exchange follows it there)
    StageScope sc(h, DQN_STAGE_HEAD_WGRAD, h->side2);
    h->head_nsplit = (k + 31) / 32 > 16 ? 16 : (k + 31) / 32;
    const int head_per = (k + h->head_nsplit - 1) / h->head_nsplit;
    h->head_nsplit = (k + head_per - 1) / head_per;
    launch_pdl(head_wgrad_kernel, dim3(((h->F + 1) * h->A + 255) / 256, h->head_nsplit), dim3(256), 0, h->side2, 1, h->act[0][3],
               h->g, d_actions, k, head_per, h->F, h->A, h->head_partial);
    }
    if (mode != 2) DQN_CUDA(cudaEventRecord(h->ev_join2, h->side2));
    // data-parallel: the gradients of conv3, fc and the head (segments 2..4, 72% of the parameters) are published and summed
    // over the ranks into the local flat gradient buffer in the shadow of the rest of the backward pass, in two parts
    // that run side by side: fc (the largest tensor) on its own stream right behind its weight-gradient kernel (flag
    // slot 2), conv3 + head on a third stream (flag slot 0).  At 8 ranks one part pulls 1-2 MB over NVLink; as a single
    // exchange behind the LAST of the three weight gradients it ended after the conv1 weight gradient (tools/trace_dp.py).
    auto exchange_part = [&](cudaStream_t s, int segs, int nsegs, int slot) -> int {
      OptArgs ea;
      fill_segs(h, &ea, nsplit, false);
      set_dp_args(h, &ea, slot);
      int len = 0;
      for (int y = 0; y < nsegs; ++y) len = std::max(len, ea.seg[(segs >> (4 * y)) & 15].len);
      dqn::count_launch();
      launch_pdl(reduce_grad_kernel, dim3((len + 255) / 256, nsegs), dim3(256), 0, s, 1, ea, h->dp_grad, 0.f, 1,
                 PublishArgs{h->dp_peer_flags_dev, h->dp_counter + slot, h->dp_rank, h->dp_world, slot, segs,
                             dp_sys_release()});
      dqn::count_launch();
      launch_pdl(peer_sum_kernel, dim3((len / 4 + 256) / 256, nsegs), dim3(256), 0, s, 1, ea, segs, h->grad);   // 4 parameters per thread
      return 0;
    };
    auto early_exchange_fc = [&]() -> int {          // on `sh`, right behind the fc weight gradient
      if (mode != 2) return 0;
      return exchange_part(sh, 0x3, 1, 2);
    };
    auto early_exchange = [&]() -> int {             // conv3 + head, on the stream of the head weight gradient
      if (mode != 2) return 0;
      cudaStream_t s2 = h->side2;
      DQN_CUDA(cudaEventRecord(h->ev_x, sd));
      DQN_CUDA(cudaStreamWaitEvent(s2, h->ev_x, 0));
      if (exchange_part(s2, 0x42, 2, 0)) return DQN_ERR_CUDA;
      DQN_CUDA(cudaEventRecord(h->ev_join2, s2));
      return 0;
    };
    auto fork_after_dgrad = [&](int i) -> int {      // dact of the next layer down is ready on `st`
      DQN_CUDA(cudaEventRecord(h->ev_fork[i], st));
      DQN_CUDA(cudaStreamWaitEvent(sd, h->ev_fork[i], 0));
      return 0;
    };
    if (h->precision == 1 && nat) {
      launch_wgrad_tc<NF, __nv_bfloat16, 128>(h, 3, h->act_bf[0][2], k, sh, &nsplit[3]);
      if (early_exchange_fc()) return DQN_ERR_CUDA;
      DQN_CUDA(cudaEventRecord(h->ev_head, sh));
      launch_dgrad_tc<NF, 64>(h, 3, k, st);
      if (fork_after_dgrad(1)) return DQN_ERR_CUDA;
      launch_wgrad_tc<N3, __nv_bfloat16, 64>(h, 2, h->act_bf[0][1], k, sd, &nsplit[2]); launch_dgrad_tc<N3, 64>(h, 2, k, st);
      if (early_exchange()) return DQN_ERR_CUDA;
      if (fork_after_dgrad(2)) return DQN_ERR_CUDA;
      launch_wgrad_tc<N2, __nv_bfloat16, 64>(h, 1, h->act_bf[0][0], k, sd, &nsplit[1]); launch_dgrad_tc<N2, 32>(h, 1, k, st);
      launch_wgrad_tc<N1, uint8_t, 32>(h, 0, d_states, k, st, &nsplit[0]);
    } else if (h->precision == 1) {
      launch_wgrad_tc<F4, __nv_bfloat16, 128>(h, 3, h->act_bf[0][2], k, sh, &nsplit[3]);
      if (early_exchange_fc()) return DQN_ERR_CUDA;
      DQN_CUDA(cudaEventRecord(h->ev_head, sh));
      if (!fused) {                  // the fused kernel already produced the fc data gradient
        launch_dgrad_tc<F4, 32>(h, 3, k, st);          // narrow tiles, see the fc forward
        if (fork_after_dgrad(1)) return DQN_ERR_CUDA;
      }
      launch_wgrad_tc<C3, __nv_bfloat16, 64>(h, 2, h->act_bf[0][1], k, sd, &nsplit[2]); if (h->no_resident) launch_dgrad_tc<C3, 64>(h, 2, k, st); else launch_dgrad_res<C3, 64, 3, 2>(h, 2, k, st);
      if (early_exchange()) return DQN_ERR_CUDA;
      if (fork_after_dgrad(2)) return DQN_ERR_CUDA;
      launch_wgrad_tc<C2, __nv_bfloat16, 64>(h, 1, h->act_bf[0][0], k, sd, &nsplit[1]); if (h->no_resident) launch_dgrad_tc<C2, 32>(h, 1, k, st); else launch_dgrad_res<C2, 32, 4, 1>(h, 1, k, st);
      if (h->conv1_s2d) launch_conv1_s2d_wgrad(h, ring ? *ring : stack_src(d_states, nullptr), k, st, &nsplit[0]);
      else if (h->no_resident) launch_wgrad_tc<C1, uint8_t, 32>(h, 0, d_states, k, st, &nsplit[0]);
      else launch_wgrad_res_u8<C1>(h, d_states, k, st, &nsplit[0]);
    } else if (nat) {
      launch_wgrad<NF, float>(h, 3, h->act[0][2], k, sh, &nsplit[3]); launch_dgrad<NF>(h, 3, k, st);
      if (early_exchange_fc()) return DQN_ERR_CUDA;
      DQN_CUDA(cudaEventRecord(h->ev_head, sh));
      if (fork_after_dgrad(1)) return DQN_ERR_CUDA;
      launch_wgrad<N3, float>(h, 2, h->act[0][1], k, sd, &nsplit[2]); launch_dgrad<N3>(h, 2, k, st);
      if (early_exchange()) return DQN_ERR_CUDA;
      if (fork_after_dgrad(2)) return DQN_ERR_CUDA;
      launch_wgrad<N2, float>(h, 1, h->act[0][0], k, sd, &nsplit[1]); launch_dgrad<N2>(h, 1, k, st);
      if (u8) launch_wgrad<N1, uint8_t>(h, 0, d_states, k, st, &nsplit[0]);
      else launch_wgrad<N1, float>(h, 0, d_states, k, st, &nsplit[0]);
    } else {
      launch_wgrad<F4, float>(h, 3, h->act[0][2], k, sh, &nsplit[3]); launch_dgrad<F4>(h, 3, k, st);
      if (early_exchange_fc()) return DQN_ERR_CUDA;
      DQN_CUDA(cudaEventRecord(h->ev_head, sh));
      if (fork_after_dgrad(1)) return DQN_ERR_CUDA;
      launch_wgrad<C3, float>(h, 2, h->act[0][1], k, sd, &nsplit[2]); launch_dgrad<C3>(h, 2, k, st);
      if (early_exchange()) return DQN_ERR_CUDA;
      if (fork_after_dgrad(2)) return DQN_ERR_CUDA;
      launch_wgrad<C2, float>(h, 1, h->act[0][0], k, sd, &nsplit[1]); launch_dgrad<C2>(h, 1, k, st);
      if (u8) launch_wgrad<C1, uint8_t>(h, 0, d_states, k, st, &nsplit[0]);
      else launch_wgrad<C1, float>(h, 0, d_states, k, st, &nsplit[0]);
    }
    fill_segs(h, &oa, nsplit, false);
    if (d_loss) {
      // last on its stream, so that no weight-gradient kernel queues behind it; the optimizer must not overwrite the
      // weights before it has read them (ev_loss)
      {
        StageScope sc(h, DQN_STAGE_LOSS, sh);
        launch_pdl(loss_kernel, dim3(1), dim3(LOSS_THREADS), 0, sh, 1, oa, h->se, k, d_loss);
      }
      DQN_CUDA(cudaEventRecord(h->ev_loss, sh));
      DQN_CUDA(cudaStreamWaitEvent(st, h->ev_loss, 0));
    }
    // join
    DQN_CUDA(cudaEventRecord(h->ev_join, sd));
    DQN_CUDA(cudaStreamWaitEvent(st, h->ev_join, 0));
    DQN_CUDA(cudaStreamWaitEvent(st, h->ev_head, 0));
    DQN_CUDA(cudaGetLastError());
    if (mode == 2 && dp_separate_publish()) {
      // last part of the exchange: conv1 and conv2 (segments 0, 1) are published here and summed by the optimizer.  (conv2
      // as a part of its own behind its weight gradient was measured: that chain of two kernels ends after this one.)
      set_dp_args(h, &oa, 1);
      dim3 g01((std::max(oa.seg[0].len, oa.seg[1].len) + 255) / 256, 2);
      {
        StageScope sc(h, DQN_STAGE_REDUCE_GRAD, st);
        launch_pdl(reduce_grad_kernel, g01, dim3(256), 0, st, 1, oa, h->dp_grad, 0.f, 1,
                   PublishArgs{h->dp_peer_flags_dev, h->dp_counter + 1, h->dp_rank, h->dp_world, 1, 0x10, dp_sys_release()});
      }
    }
    DQN_CUDA(cudaStreamWaitEvent(st, h->ev_join2, 0));
    if (mode == 1) {
      dim3 grid((max_seg_len(oa) + 255) / 256, oa.nseg);
      StageScope sc(h, DQN_STAGE_REDUCE_GRAD, st);
      launch_pdl(reduce_grad_kernel, grid, dim3(256), 0, st, 1, oa, h->grad, include_reg ? h->hyper.reg_param : 0.f, 0,
                 PublishArgs{nullptr, nullptr, 0, 1, 0, 0x43210, 0});
    }
  }
  if (mode == 0) {
    dim3 grid((max_seg_len(oa) + 255) / 256, oa.nseg);
    {
      StageScope sc(h, DQN_STAGE_OPTIMIZER, st);
      launch_pdl(optimizer_kernel, grid, dim3(256), 0, st, 1, oa);
    }
  }
  if (mode == 2) {
    // the reduce kernel above published this rank's gradient; now sum all ranks' gradients inside the optimizer
    // (peer loads over NVLink)
    const int one[4] = {1, 1, 1, 1};
    OptArgs part = oa;                         // conv nets: segments with the weight-gradient partials of this step
    fill_segs(h, &oa, one, true);
    set_dp_args(h, &oa, 1);
    // conv nets: only segments 0, 1 (conv1, conv2) are still to be summed over the ranks; the MLP exchanges everything here
    oa.peer_mask = h->arch == DQN_ARCH_SIMPLE ? 0x7 : 0x3;
    if (h->arch != DQN_ARCH_SIMPLE && !dp_separate_publish()) {
      // the optimizer's own blocks publish conv1 / conv2 (no separate kernel on the critical path)
      for (int l = 0; l < 2; ++l) {
        oa.seg[l].partial = part.seg[l].partial; oa.seg[l].nsplit = part.seg[l].nsplit; oa.seg[l].stride = part.seg[l].stride;
      }
      oa.pub_flags = h->dp_peer_flags_dev;
      oa.pub_counter = h->dp_counter + 1;
      oa.pub_grad = h->dp_grad;
      oa.pub_total = (oa.seg[0].len + 255) / 256 + (oa.seg[1].len + 255) / 256;
      oa.rank = h->dp_rank;
      oa.pub_sys_release = dp_sys_release();
    }
    dim3 grid((max_seg_len(oa) + 255) / 256, oa.nseg);
    StageScope sc(h, DQN_STAGE_OPTIMIZER, st);
    launch_pdl(optimizer_kernel, grid, dim3(256), 0, st, 1, oa);
  }
  DQN_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dqn_net_train_step(dqn_net_t h, const void* d_states, const void* d_next_states, const int32_t* d_actions,
                                  const double* d_rewards, const uint8_t* d_terminals, int32_t k, float* d_loss,
                                  float* d_q, dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "train_step: NULL handle");
  DeviceGuard g(h->device);
  return run_step(h, d_states, d_next_states, d_actions, d_rewards, d_terminals, k, d_loss, d_q, 0, 0,
                  static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_net_backward(dqn_net_t h, const void* d_states, const void* d_next_states, const int32_t* d_actions,
                                const double* d_rewards, const uint8_t* d_terminals, int32_t k, float* d_loss, float* d_q,
                                int32_t include_reg, dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "backward: NULL handle");
  DeviceGuard g(h->device);
  return run_step(h, d_states, d_next_states, d_actions, d_rewards, d_terminals, k, d_loss, d_q, 1, include_reg,
                  static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_net_apply(dqn_net_t h, dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "apply: NULL handle");
  DeviceGuard g(h->device);
  const int nsplit[4] = {1, 1, 1, 1};
  OptArgs oa;
  fill_segs(h, &oa, nsplit, true);
  dim3 grid((max_seg_len(oa) + 255) / 256, oa.nseg);
  {
    StageScope sc(h, DQN_STAGE_OPTIMIZER, static_cast<cudaStream_t>(stream));
    optimizer_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(oa);
  }
  DQN_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dqn_net_target_sync(dqn_net_t h, dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "target_sync: NULL handle");
  DeviceGuard g(h->device);
  DQN_CUDA(cudaMemcpyAsync(h->target, h->online, (size_t)h->P * 4, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  if (h->precision == 1)
    DQN_CUDA(cudaMemcpyAsync(h->wt_bf[1], h->wt_bf[0], h->wt_total * 2, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
  return 0;
}

// Sample-to-update.  Reference conv net in bf16 mode: no gather launch -- the conv1 kernels fetch the frames of the
// sampled transitions from the ring themselves (4 bulk copies per stack) and the forward kernel copies the scalars.
// Everything else: gather into the net's staging buffers, then the step.
struct IndexTableRef { const int32_t* table; int64_t rows; int64_t* cursor; int32_t* scratch; };
int dqn_index_table_next_(const int32_t* d_table, int64_t rows, int32_t k, int64_t* d_cursor, int32_t* d_idx, cudaStream_t st);

static int step_from_ring(dqn_net_s* net, dqn_replay_t replay, const int32_t* d_idx, int k, float* d_loss, int mode, cudaStream_t st,
                          const IndexTableRef* tab = nullptr) {
  RingView rv;
  dqn_replay_view_(replay, &rv);
  const bool fused = net->conv1_s2d && net->fuse_gather && rv.frames_per_state == 4 && rv.elem_bytes == 1 &&
                     rv.frame_bytes == (int)umma::s2d::FRAME_BYTES;
  if (tab && !fused) {
    // kernels that cannot read the table themselves: one small kernel takes the row (and advances the cursor)
    int rc = dqn_index_table_next_(tab->table, tab->rows, k, tab->cursor, tab->scratch, st);
    if (rc) return rc;
    d_idx = tab->scratch;
    tab = nullptr;
  }
  if (fused) {
    int rc = dqn_replay_flush(replay, st);                  // frames / records staged on the host go to the ring first
    if (rc) return rc;
    umma::s2d::Src src;
    memset(&src, 0, sizeof(src));
    src.frames = rv.frames; src.meta = rv.meta; src.ring_state = rv.state; src.idx = d_idx;
    src.capacity = rv.capacity; src.frame_capacity = rv.frame_capacity; src.pitch = rv.pitch;
    src.actions = net->st_actions; src.rewards = net->st_rewards; src.terminals = net->st_term;
    if (tab) {
      // the conv1 kernels read row (*cursor % rows) of the table themselves; the optimizer advances the cursor
      src.idx = tab->table; src.idx_cursor = tab->cursor; src.idx_rows = tab->rows; src.idx_k = k;
      net->idx_cursor = reinterpret_cast<long long*>(tab->cursor);
    }
    rc = run_step(net, nullptr, nullptr, net->st_actions, net->st_rewards, net->st_term, k, d_loss, nullptr, mode, 0, st, &src);
    net->idx_cursor = nullptr;
    return rc;
  }
  if (net->timing) stage_record(net->ev0[DQN_STAGE_GATHER], st);
  int rc = dqn_replay_gather(replay, d_idx, k, net->st_states, net->st_next, net->st_actions, net->st_rewards, net->st_term, st);
  if (net->timing) { stage_record(net->ev1[DQN_STAGE_GATHER], st); net->stage_used[DQN_STAGE_GATHER] = true; }
  if (rc) return rc;
  return run_step(net, net->st_states, net->st_next, net->st_actions, net->st_rewards, net->st_term, k, d_loss, nullptr, mode, 0, st);
}

extern "C" int dqn_net_refresh_operands(dqn_net_t h, int32_t which, dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "refresh_operands: NULL handle");
  DQN_REQUIRE(which == DQN_BUF_ONLINE || which == DQN_BUF_TARGET, DQN_ERR_INVALID, "refresh_operands: which must be online or target");
  if (h->precision != 1) return 0;                      // the fp32 kernels read the master weights themselves
  DeviceGuard g(h->device);
  return refresh_bf16_weights(h, which == DQN_BUF_ONLINE ? 0 : 1, static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_train_step(dqn_net_t net, dqn_replay_t replay, const int32_t* d_logical_idx, int32_t k, float* d_loss,
                              dqn_stream_t stream) {
  DQN_REQUIRE(net && replay && d_logical_idx, DQN_ERR_INVALID, "dqn_train_step: NULL argument");
  DQN_REQUIRE(k >= 1 && k <= net->MB, DQN_ERR_INVALID, "dqn_train_step: minibatch %d outside [1, max_batch=%d]", k, net->MB);
  DQN_REQUIRE(dqn_replay_device_(replay) == net->device, DQN_ERR_INVALID, "replay and net live on different devices");
  RingView rv;
  dqn_replay_view_(replay, &rv);
  DQN_REQUIRE((size_t)rv.frame_bytes * rv.frames_per_state == net->state_bytes, DQN_ERR_INVALID,
              "replay stacks are %d bytes but the net expects %zu", rv.frame_bytes * rv.frames_per_state, net->state_bytes);
  DeviceGuard g(net->device);
  return step_from_ring(net, replay, d_logical_idx, k, d_loss, 0, static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_train_step_table(dqn_net_t net, dqn_replay_t replay, const int32_t* d_table, int64_t rows, int32_t k,
                                    int64_t* d_cursor, int32_t* d_idx_scratch, float* d_loss, int32_t data_parallel,
                                    dqn_stream_t stream) {
  DQN_REQUIRE(net && replay && d_table && d_cursor && d_idx_scratch && rows > 0, DQN_ERR_INVALID, "dqn_train_step_table: bad argument");
  DQN_REQUIRE(k >= 1 && k <= net->MB, DQN_ERR_INVALID, "dqn_train_step_table: minibatch %d outside [1, max_batch=%d]", k, net->MB);
  DQN_REQUIRE(dqn_replay_device_(replay) == net->device, DQN_ERR_INVALID, "replay and net live on different devices");
  DQN_REQUIRE(!data_parallel || (net->dp_world >= 2 && net->dp_peer_flags_dev), DQN_ERR_STATE,
              "dqn_train_step_table: call dqn_net_dp_connect first");
  RingView rv;
  dqn_replay_view_(replay, &rv);
  DQN_REQUIRE((size_t)rv.frame_bytes * rv.frames_per_state == net->state_bytes, DQN_ERR_INVALID,
              "replay stacks are %d bytes but the net expects %zu", rv.frame_bytes * rv.frames_per_state, net->state_bytes);
  DeviceGuard g(net->device);
  const IndexTableRef tab{d_table, rows, d_cursor, d_idx_scratch};
  return step_from_ring(net, replay, nullptr, k, d_loss, data_parallel ? 2 : 0, static_cast<cudaStream_t>(stream), &tab);
}

// ---------------------------------------------------------------------------------------------- peer-memory data parallelism
static int dp_alloc(dqn_net_s* h) {
  if (h->dp_block) return 0;
  const size_t bytes = 256 + 2 * (size_t)h->P * sizeof(float);
  DQN_CUDA(cudaMalloc(&h->dp_block, bytes));            // its own allocation: IPC handles cover whole allocations
  DQN_CUDA(cudaMemset(h->dp_block, 0, bytes));
  h->dp_grad = reinterpret_cast<float*>(static_cast<char*>(h->dp_block) + 256);
  DQN_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->dp_peer_flags_dev), 8 * sizeof(void*)));
  h->allocs.push_back(h->dp_peer_flags_dev);
  DQN_CUDA(cudaMalloc(reinterpret_cast<void**>(&h->dp_counter), 32));
  DQN_CUDA(cudaMemset(h->dp_counter, 0, 32));
  h->allocs.push_back(h->dp_counter);
  h->dp_err = reinterpret_cast<unsigned long long*>(h->dp_counter + 4);       // bytes 16..23 of the same allocation
  { const char* to = getenv("DQN_DP_TIMEOUT_S"); if (to && atof(to) > 0) h->dp_timeout_ns = (long long)(atof(to) * 1e9); }
  return 0;
}

extern "C" int dqn_net_dp_handle(dqn_net_t h, void* out_handle64) {
  DQN_REQUIRE(h && out_handle64, DQN_ERR_INVALID, "dp_handle: NULL argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  DeviceGuard g(h->device);
  int rc = dp_alloc(h);
  if (rc) return rc;
  cudaIpcMemHandle_t hd;
  DQN_CUDA(cudaIpcGetMemHandle(&hd, h->dp_block));
  memcpy(out_handle64, &hd, 64);
  return 0;
}

extern "C" int dqn_net_dp_connect(dqn_net_t h, int32_t rank, int32_t world, const void* handles) {
  DQN_REQUIRE(h && handles, DQN_ERR_INVALID, "dp_connect: NULL argument");
  DQN_REQUIRE(world >= 2 && world <= 8 && rank >= 0 && rank < world, DQN_ERR_INVALID,
              "dp_connect: world must be 2..8 (one NVLink domain) and 0 <= rank < world");
  DeviceGuard g(h->device);
  int rc = dp_alloc(h);
  if (rc) return rc;
  unsigned long long* flags[8] = {};
  for (int r = 0; r < world; ++r) {
    if (r == rank) {
      h->dp_peer_block[r] = h->dp_block;
    } else if (!h->dp_peer_block[r]) {
      cudaIpcMemHandle_t hd;
      memcpy(&hd, static_cast<const char*>(handles) + 64 * r, 64);
      DQN_CUDA(cudaIpcOpenMemHandle(&h->dp_peer_block[r], hd, cudaIpcMemLazyEnablePeerAccess));
    }
    flags[r] = static_cast<unsigned long long*>(h->dp_peer_block[r]);
  }
  DQN_CUDA(cudaMemcpy(h->dp_peer_flags_dev, flags, sizeof(flags), cudaMemcpyHostToDevice));
  h->dp_rank = rank;
  h->dp_world = world;
  return 0;
}

extern "C" int dqn_net_dp_step(dqn_net_t h, const void* d_states, const void* d_next_states, const int32_t* d_actions,
                               const double* d_rewards, const uint8_t* d_terminals, int32_t k, float* d_loss,
                               dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "dp_step: NULL handle");
  DQN_REQUIRE(h->dp_world >= 2 && h->dp_peer_flags_dev, DQN_ERR_STATE, "dp_step: call dqn_net_dp_connect first");
  DeviceGuard g(h->device);
  return run_step(h, d_states, d_next_states, d_actions, d_rewards, d_terminals, k, d_loss, nullptr, 2, 0,
                  static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_train_step_dp(dqn_net_t net, dqn_replay_t replay, const int32_t* d_logical_idx, int32_t k, float* d_loss,
                                 dqn_stream_t stream) {
  DQN_REQUIRE(net && replay && d_logical_idx, DQN_ERR_INVALID, "dqn_train_step_dp: NULL argument");
  DQN_REQUIRE(k >= 1 && k <= net->MB, DQN_ERR_INVALID, "dqn_train_step_dp: minibatch %d outside [1, max_batch=%d]", k, net->MB);
  DQN_REQUIRE(net->dp_world >= 2 && net->dp_peer_flags_dev, DQN_ERR_STATE, "dqn_train_step_dp: call dqn_net_dp_connect first");
  DQN_REQUIRE(dqn_replay_device_(replay) == net->device, DQN_ERR_INVALID, "replay and net live on different devices");
  DeviceGuard g(net->device);
  return step_from_ring(net, replay, d_logical_idx, k, d_loss, 2, static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_net_dp_set_timeout(dqn_net_t h, double seconds) {
  DQN_REQUIRE(h && seconds > 0, DQN_ERR_INVALID, "dp_set_timeout: bad argument");
  h->dp_timeout_ns = (long long)(seconds * 1e9);
  return 0;
}

extern "C" int dqn_net_dp_status(dqn_net_t h, int64_t* timed_out_step, dqn_stream_t stream) {
  DQN_REQUIRE(h && timed_out_step, DQN_ERR_INVALID, "dp_status: NULL argument");
  *timed_out_step = 0;
  if (!h->dp_err) return 0;
  DeviceGuard g(h->device);
  unsigned long long v = 0;
  DQN_CUDA(cudaMemcpyAsync(&v, h->dp_err, sizeof(v), cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
  DQN_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  *timed_out_step = (int64_t)v;
  return 0;
}

// ---------------------------------------------------------------------------------------------- precision mode
extern "C" int dqn_net_set_precision(dqn_net_t h, int32_t precision) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "set_precision: NULL handle");
  DQN_REQUIRE(precision == 0 || precision == 1, DQN_ERR_INVALID, "precision must be 0 (fp32) or 1 (bf16 tensor cores)");
  if (precision == 1) {
    DQN_REQUIRE((h->arch == DQN_ARCH_CNN || h->arch == DQN_ARCH_NATURE) && h->in_dtype == DQN_U8, DQN_ERR_UNSUPPORTED,
                "the bf16 tensor-core path covers the conv nets (DQN_ARCH_CNN, DQN_ARCH_NATURE) fed uint8 frames");
    DeviceGuard g(h->device);
    if (!h->wt_bf[0]) {
      size_t wt = 0, wd = 0;
      for (int l = 0; l < 4; ++l) {
        h->wt_off[l] = wt; wt += (size_t)h->L[l].K * h->L[l].N;
        h->wd_off[l] = wd; if (l >= 1) wd += (size_t)h->L[l].K * h->L[l].N;
      }
      h->wt_total = wt; h->wd_total = wd;
      NET_ALLOC(h->wt_bf[0], wt * 2); NET_ALLOC(h->wt_bf[1], wt * 2); NET_ALLOC(h->wd_bf, wd * 2);
      for (int p = 0; p < 2; ++p)
        for (int l = 0; l < 4; ++l) NET_ALLOC(h->act_bf[p][l], (size_t)h->MB * h->L[l].out_elems * 2);
      for (int l = 0; l < 4; ++l) NET_ALLOC(h->dact_bf[l], (size_t)h->MB * h->L[l].out_elems * 2);
    }
    h->precision = 1;
    h->conv1_s2d = h->arch == DQN_ARCH_CNN && !h->no_resident;
    int rc = refresh_bf16_weights(h, 0, nullptr);
    if (!rc) rc = refresh_bf16_weights(h, 1, nullptr);
    if (rc) return rc;
    DQN_CUDA(cudaDeviceSynchronize());
  } else {
    h->precision = 0;
    h->conv1_s2d = false;
  }
  return 0;
}

// ---------------------------------------------------------------------------------------------- instrumentation
extern "C" int64_t dqn_launch_count(void) { return dqn::launch_count(); }

extern "C" const char* dqn_stage_name(int32_t stage) {
  return (stage >= 0 && stage < DQN_NUM_STAGES) ? kStageNames[stage] : "";
}

extern "C" int dqn_net_set_timing(dqn_net_t h, int32_t enable) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "set_timing: NULL handle");
  DeviceGuard g(h->device);
  if (enable && !h->ev0[0]) {
    for (int i = 0; i < DQN_NUM_STAGES; ++i) {
      DQN_CUDA(cudaEventCreate(&h->ev0[i]));
      DQN_CUDA(cudaEventCreate(&h->ev1[i]));
    }
  }
  h->timing = enable != 0;
  h->timing_sticky = enable == 3;
  if (enable != 3) for (int i = 0; i < DQN_NUM_STAGES; ++i) h->stage_used[i] = false;
  if (enable == 2) for (int i = 0; i < DQN_NUM_STAGES; ++i) { h->stage_ms[i] = 0.0; h->stage_count[i] = 0; }
  return 0;
}

extern "C" int dqn_net_timing_collect(dqn_net_t h) {
  DQN_REQUIRE(h && h->timing, DQN_ERR_STATE, "timing_collect: timing is not enabled");
  DeviceGuard g(h->device);
  DQN_CUDA(cudaDeviceSynchronize());
  for (int i = 0; i < DQN_NUM_STAGES; ++i) {
    if (!h->stage_used[i]) continue;
    float ms = 0.f;
    DQN_CUDA(cudaEventElapsedTime(&ms, h->ev0[i], h->ev1[i]));
    h->stage_ms[i] += ms;
    h->stage_count[i] += 1;
    if (!h->timing_sticky) h->stage_used[i] = false;
  }
  return 0;
}

// Enqueues the timing bracket around an EMPTY kernel as the pseudo-stage "launch_floor" (works eagerly and under stream
// capture, like the real brackets): what a bracket costs by itself, in the same mode as the step's brackets.
extern "C" int dqn_net_timing_floor_enqueue(dqn_net_t h, dqn_stream_t stream) {
  DQN_REQUIRE(h && h->timing, DQN_ERR_STATE, "timing_floor_enqueue: timing is not enabled");
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  {
    StageScope sc(h, DQN_STAGE_LAUNCH_FLOOR, st);
    launch_pdl(empty_kernel, dim3(1), dim3(32), 0, st, 1);
  }
  DQN_CUDA(cudaGetLastError());
  return 0;
}

// What the event bracket of the timing hooks measures around a kernel that does nothing: event record, launch of an
// empty kernel the same way the step's kernels are launched, event record.  bench.py subtracts it from the brackets.
extern "C" int dqn_net_timing_floor(dqn_net_t h, dqn_stream_t stream, double* us) {
  DQN_REQUIRE(h && us, DQN_ERR_INVALID, "timing_floor: NULL argument");
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaEvent_t e0, e1;
  DQN_CUDA(cudaEventCreate(&e0));
  DQN_CUDA(cudaEventCreate(&e1));
  double total = 0.0;
  const int n = 20;
  for (int i = 0; i < n + 2; ++i) {
    DQN_CUDA(cudaEventRecord(e0, st));
    launch_pdl(empty_kernel, dim3(1), dim3(32), 0, st, 1);
    DQN_CUDA(cudaEventRecord(e1, st));
    DQN_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    DQN_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (i >= 2) total += ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *us = total / n * 1e3;
  return 0;
}

extern "C" int dqn_net_timing_read(dqn_net_t h, int32_t stage, double* total_ms, int32_t* count) {
  DQN_REQUIRE(h && stage >= 0 && stage < DQN_NUM_STAGES, DQN_ERR_INVALID, "timing_read: bad argument");
  if (total_ms) *total_ms = h->stage_ms[stage];
  if (count) *count = h->stage_count[stage];
  return 0;
}

// ================================================================================================
// Forward pass that allocates nothing
// ================================================================================================
// dqn_net_forward keeps the activations between its kernels in buffers that dqn_net_create obtained
// from cudaMalloc.  Some hosts of this library are not allowed to allocate device memory at all: an
// actor process that was handed a fixed carve-out of the GPU by its parent, a real-time control loop
// that must not enter the allocator (cudaMalloc may take a device-wide lock and synchronise), or a
// sandboxed plugin whose allocation calls are intercepted.  For them the variant below runs the same
// kernels for the reference conv net (DQN_ARCH_CNN, uint8 84x84x4 states) with the inter-kernel
// activations in a fixed array that is part of the library image itself.
//
// The array costs kNoAllocFloats * 4 bytes (about 1.27 MB) on every GPU that loads the library,
// from load until the process exits, whether or not the entry point is ever called.
namespace {
constexpr int kNoAllocMaxStates = 64;                                   // largest n per call
constexpr int kNoAllocPerState = C1::OUT_PER_SAMPLE + C2::OUT_PER_SAMPLE + C3::OUT_PER_SAMPLE + F4::OUT_PER_SAMPLE;
constexpr int kNoAllocFloats = kNoAllocMaxStates * kNoAllocPerState;    // 64 * 4960 floats
constexpr int kNoAllocMaxDevices = 64;
}  // namespace

// In host code this name is only a handle for cudaGetSymbolAddress; it is NOT a device pointer.
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocFloats];

// Device address of the array on each device, looked up once per device with that device current.
static float* g_noalloc_ptr[kNoAllocMaxDevices] = {};

extern "C" int dqn_cnn_forward_noalloc(const float* d_params, int32_t num_actions, const uint8_t* d_states, int32_t n,
                                       float* d_q, int32_t* d_argmax, dqn_stream_t stream) {
  DQN_REQUIRE(d_params && d_states, DQN_ERR_INVALID, "dqn_cnn_forward_noalloc: NULL params or states");
  DQN_REQUIRE(d_q || d_argmax, DQN_ERR_INVALID, "dqn_cnn_forward_noalloc: pass d_q, d_argmax or both");
  DQN_REQUIRE(num_actions >= 1 && num_actions <= 64, DQN_ERR_INVALID, "dqn_cnn_forward_noalloc: num_actions must be in [1, 64]");
  DQN_REQUIRE(n >= 1, DQN_ERR_INVALID, "dqn_cnn_forward_noalloc: n must be positive");
  DQN_REQUIRE(n <= kNoAllocMaxStates, DQN_ERR_INVALID,
              "dqn_cnn_forward_noalloc: n=%d exceeds the built-in scratch, which holds at most %d states per call; "
              "split the batch or use dqn_net_forward", n, kNoAllocMaxStates);
  DQN_REQUIRE(((uintptr_t)d_params % 16) == 0 && ((uintptr_t)d_states % 4) == 0, DQN_ERR_INVALID,
              "dqn_cnn_forward_noalloc: d_params must be 16-byte and d_states 4-byte aligned");
  // every argument check is above this line: a refused call has touched nothing
  int dev = -1;
  DQN_CUDA(cudaGetDevice(&dev));
  DQN_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, DQN_ERR_UNSUPPORTED, "device index %d not supported", dev);
  if (!g_noalloc_ptr[dev]) {
    // one-time set-up for this device (the current one): resolve the array's device address.
    float* p = nullptr;
    DQN_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch));
    DQN_REQUIRE(p != nullptr, DQN_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[dev] = p;
  }
  int num_sms = 148;
  DQN_CUDA(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* base = g_noalloc_ptr[dev];
  float* act[4];
  act[0] = base;
  act[1] = act[0] + (size_t)kNoAllocMaxStates * C1::OUT_PER_SAMPLE;
  act[2] = act[1] + (size_t)kNoAllocMaxStates * C2::OUT_PER_SAMPLE;
  act[3] = act[2] + (size_t)kNoAllocMaxStates * C3::OUT_PER_SAMPLE;
  // parameter tensors back to back in the reference's tuple order (network.py:387); every tensor but the
  // last has a multiple of 4 elements, so this equals the layout of the handle-based flat buffers
  const int64_t sizes[10] = {8 * 8 * 4 * 32, 32, 4 * 4 * 32 * 64, 64, 3 * 3 * 64 * 64, 64, 256 * 256, 256,
                             256 * (int64_t)num_actions, num_actions};
  const float* t[10];
  int64_t off = 0;
  for (int i = 0; i < 10; ++i) { t[i] = d_params + off; off += sizes[i]; }
  for (int l = 0; l < 4; ++l) {
    FwdArgs a;
    a.B = n;
    a.p[0].x = l == 0 ? static_cast<const void*>(d_states) : static_cast<const void*>(act[l - 1]);
    a.p[0].w = t[2 * l];
    a.p[0].bias = t[2 * l + 1];
    a.p[0].out = act[l];
    a.p[0].arg = nullptr;
    a.p[1] = a.p[0];
    dqn::count_launch();
    switch (l) {
      case 0: launch_fwd<C1, uint8_t>(a, 1, num_sms, st); break;
      case 1: launch_fwd<C2, float>(a, 1, num_sms, st); break;
      case 2: launch_fwd<C3, float>(a, 1, num_sms, st); break;
      case 3: launch_fwd<F4, float>(a, 1, num_sms, st); break;
    }
  }
  dqn::count_launch();
  head_forward_kernel<<<(n + 7) / 8, 256, 0, st>>>(act[3], t[8], t[9], n, 256, num_actions, d_q, d_argmax);
  DQN_CUDA(cudaGetLastError());
  return 0;
}
