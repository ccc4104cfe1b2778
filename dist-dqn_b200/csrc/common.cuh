// Shared helpers of libdqn_b200: error reporting across the C ABI, checked CUDA calls.
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>

#include "../../include/dqn_b200.h"

namespace dqn {

// Thread-local message behind dqn_last_error().
char* err_buf();
int set_err(int code, const char* fmt, ...);

#define DQN_CUDA(expr)                                                                              \
  do {                                                                                              \
    cudaError_t e__ = (expr);                                                                       \
    if (e__ != cudaSuccess)                                                                         \
      return ::dqn::set_err(DQN_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__),  \
                            __FILE__, __LINE__);                                                    \
  } while (0)

#define DQN_REQUIRE(cond, code, ...)                              \
  do {                                                            \
    if (!(cond)) return ::dqn::set_err(code, __VA_ARGS__);        \
  } while (0)

struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
    cur = dev;
  }
  ~DeviceGuard() { if (prev >= 0 && prev != cur) cudaSetDevice(prev); }
  int cur = -1;
};

static inline int64_t round_up(int64_t v, int64_t m) { return (v + m - 1) / m * m; }

// Process-wide count of kernels this library has launched (bench.py's gpu_launches evidence).
void count_launch();
int64_t launch_count();

// Programmatic dependent launch (PDL): a kernel launched with launch_pdl may start while its predecessor in the
// stream is still running; it must call pdl_wait() before touching any global memory (the wait returns once the
// predecessor has completed and flushed), and calls pdl_trigger() early so that its own successor can be launched
// and run its prologue (barrier init, TMEM allocation, index arithmetic) in the shadow of this kernel.  Both are
// no-ops for kernels launched the ordinary way.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// Wait for the predecessor, then -- if this launch is a single wave (every CTA is already resident, so early dependents
// cannot take shared memory / TMEM away from CTAs that still have to start) -- let the successor launch right away: it
// runs its set-up and its weight prefetch in the shadow of this kernel's main loop and blocks in its own wait.
// Multi-wave launches trigger late (after their main loop), as before.
__device__ __forceinline__ void pdl_wait_then_trigger() {
  pdl_wait();
  if (gridDim.x * gridDim.y * gridDim.z <= 128u) pdl_trigger();
}

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                     unsigned cluster_z, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  int na = 0;
  at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[na].val.programmaticStreamSerializationAllowed = 1;
  ++na;
  if (cluster_z > 1) {
    at[na].id = cudaLaunchAttributeClusterDimension;
    at[na].val.clusterDim.x = 1;
    at[na].val.clusterDim.y = 1;
    at[na].val.clusterDim.z = cluster_z;
    ++na;
  }
  cfg.attrs = at;
  cfg.numAttrs = na;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// Ring view passed to kernels that read the replay ring (replay.cu, and the fused step).
struct RingView {
  const uint8_t* frames;   // frame_capacity x pitch bytes
  const void* meta;        // capacity x 32-byte records
  const int64_t* state;    // device {head, size}
  int64_t capacity;
  int64_t frame_capacity;
  int32_t pitch;           // bytes between frames (frame_bytes rounded up to 16)
  int32_t frame_bytes;
  int32_t elem_bytes;
  int32_t frames_per_state;
};

}  // namespace dqn

// Stage ids of the per-kernel timing hooks (dqn_net_set_timing); names in net.cu.
enum {
  DQN_STAGE_GATHER = 0, DQN_STAGE_FWD_CONV1, DQN_STAGE_FWD_CONV2, DQN_STAGE_FWD_CONV3, DQN_STAGE_FWD_FC,
  DQN_STAGE_TD_HEAD, DQN_STAGE_HEAD_WGRAD, DQN_STAGE_LOSS, DQN_STAGE_WGRAD_FC, DQN_STAGE_DGRAD_FC,
  DQN_STAGE_WGRAD_CONV3, DQN_STAGE_DGRAD_CONV3, DQN_STAGE_WGRAD_CONV2, DQN_STAGE_DGRAD_CONV2, DQN_STAGE_WGRAD_CONV1,
  DQN_STAGE_REDUCE_GRAD, DQN_STAGE_OPTIMIZER, DQN_STAGE_MLP_BACKWARD, DQN_STAGE_MLP_FORWARD, DQN_STAGE_HEAD_FORWARD,
  DQN_STAGE_FC_TD_FUSED, DQN_STAGE_LAUNCH_FLOOR,
  DQN_NUM_STAGES
};

// Internal cross-file hooks (not part of the public ABI).
int dqn_replay_view_(dqn_replay_t h, dqn::RingView* out);
int dqn_replay_device_(dqn_replay_t h);
