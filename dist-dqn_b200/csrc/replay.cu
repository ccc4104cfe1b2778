// Device-resident replay ring + the sample gather.
//
// Replaces ReplayMemory (reference src/replay_memory.py:18-53) and the per-transition frame
// stacks FrameBuffer.get_state() materialises (src/frame_buffer.py:40-51): the reference stores
// every 84x84 frame four times (once per stack it appears in, ~28 GB of host RAM at 1M
// transitions) and samples by O(n) deque indexing; here every frame lives ONCE in an HBM byte
// ring and a sample is one gather kernel that rebuilds the HWC stacks (channel = time).
//
// HBM layout
//   frames : frame_capacity x pitch bytes   (pitch = frame_bytes rounded up to 16)
//   meta   : capacity x 32 B  {int64 state_pos, int64 next_pos, double reward, int32 action, u8 terminal}
//   state  : {int64 head, int64 size}        logical index i -> record slot (head + i) % capacity
// A stack that "ends at position p" is frames p-(S-1) .. p (ring positions, slot = p % frame_capacity).
#include "common.cuh"

#include <atomic>
#include <cstring>
#include <new>
#include <vector>

namespace dqn {

static thread_local char g_err[512] = "";
char* err_buf() { return g_err; }
int set_err(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

static std::atomic<long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
int64_t launch_count() { return g_launches.load(std::memory_order_relaxed); }

struct __align__(16) TransitionRec {
  int64_t state_pos;
  int64_t next_pos;
  double reward;
  int32_t action;
  uint8_t terminal;
  uint8_t pad[3];
};
static_assert(sizeof(TransitionRec) == 32, "record must be 32 bytes");

// ---------------------------------------------------------------------------------------------
// Gather kernels
// ---------------------------------------------------------------------------------------------

// 4x4 byte transpose: in a,b,c,d = 4 pixels of frames 0..3; out[p] = {a[p], b[p], c[p], d[p]}.
__device__ __forceinline__ void transpose4x4(uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t out[4]) {
  uint32_t ab_lo = __byte_perm(a, b, 0x5140);  // a0 b0 a1 b1
  uint32_t ab_hi = __byte_perm(a, b, 0x7362);  // a2 b2 a3 b3
  uint32_t cd_lo = __byte_perm(c, d, 0x5140);
  uint32_t cd_hi = __byte_perm(c, d, 0x7362);
  out[0] = __byte_perm(ab_lo, cd_lo, 0x5410);  // a0 b0 c0 d0
  out[1] = __byte_perm(ab_lo, cd_lo, 0x7632);  // a1 b1 c1 d1
  out[2] = __byte_perm(ab_hi, cd_hi, 0x5410);
  out[3] = __byte_perm(ab_hi, cd_hi, 0x7632);
}

__device__ __forceinline__ void store_stack16(uint4* dst, const uint4& f0, const uint4& f1, const uint4& f2,
                                              const uint4& f3) {
  // 16 pixels x 4 frames -> 64 contiguous bytes of the HWC stack
  uint32_t o[4];
  uint4 v;
  transpose4x4(f0.x, f1.x, f2.x, f3.x, o); v = make_uint4(o[0], o[1], o[2], o[3]); dst[0] = v;
  transpose4x4(f0.y, f1.y, f2.y, f3.y, o); v = make_uint4(o[0], o[1], o[2], o[3]); dst[1] = v;
  transpose4x4(f0.z, f1.z, f2.z, f3.z, o); v = make_uint4(o[0], o[1], o[2], o[3]); dst[2] = v;
  transpose4x4(f0.w, f1.w, f2.w, f3.w, o); v = make_uint4(o[0], o[1], o[2], o[3]); dst[3] = v;
}

__device__ __forceinline__ const uint4* frame_chunk(const uint8_t* frames, int64_t pos, int64_t fcap, int32_t pitch,
                                                    int32_t chunk) {
  int64_t slot = pos % fcap;
  return reinterpret_cast<const uint4*>(frames + slot * (int64_t)pitch) + chunk;
}

// Fast path: uint8 frames, 4 frames per state, frame_bytes % 16 == 0.
// One thread = one 16-pixel chunk of one sampled transition: 5 (or 8) 16-byte loads, 8 16-byte stores.
__global__ void __launch_bounds__(256)
gather_stack4_u8_kernel(RingView rv, const int32_t* __restrict__ idx, int32_t k, int32_t chunks,
                        uint8_t* __restrict__ states, uint8_t* __restrict__ next_states,
                        int32_t* __restrict__ actions, double* __restrict__ rewards,
                        uint8_t* __restrict__ terminals) {
  pdl_wait();
  if (gridDim.x <= 1024u) pdl_trigger();      // single wave: the consumer may launch and run its set-up now (it waits for our completion)
  const int64_t total = (int64_t)k * chunks;
  const int64_t head = rv.state[0];
  const TransitionRec* meta = static_cast<const TransitionRec*>(rv.meta);
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int32_t tr = (int32_t)(t / chunks);
    const int32_t ch = (int32_t)(t - (int64_t)tr * chunks);
    int64_t slot = head + (int64_t)__ldg(idx + tr);
    if (slot >= rv.capacity) slot -= rv.capacity;
    const TransitionRec* m = meta + slot;
    const int64_t ps = m->state_pos, pn = m->next_pos;
    if (ch == 0) {
      if (actions) actions[tr] = m->action;
      if (rewards) rewards[tr] = m->reward;
      if (terminals) terminals[tr] = m->terminal;
    }
    const int64_t out_off = ((int64_t)tr * chunks + ch) * 4;  // in uint4 units (64 B per chunk)
    if (pn == ps + 1) {
      // common case: the two stacks overlap in 3 frames -> 5 unique frames
      uint4 f0 = __ldg(frame_chunk(rv.frames, ps - 3, rv.frame_capacity, rv.pitch, ch));
      uint4 f1 = __ldg(frame_chunk(rv.frames, ps - 2, rv.frame_capacity, rv.pitch, ch));
      uint4 f2 = __ldg(frame_chunk(rv.frames, ps - 1, rv.frame_capacity, rv.pitch, ch));
      uint4 f3 = __ldg(frame_chunk(rv.frames, ps, rv.frame_capacity, rv.pitch, ch));
      uint4 f4 = __ldg(frame_chunk(rv.frames, pn, rv.frame_capacity, rv.pitch, ch));
      if (states) store_stack16(reinterpret_cast<uint4*>(states) + out_off, f0, f1, f2, f3);
      if (next_states) store_stack16(reinterpret_cast<uint4*>(next_states) + out_off, f1, f2, f3, f4);
    } else {
      if (states) {
        uint4 f0 = __ldg(frame_chunk(rv.frames, ps - 3, rv.frame_capacity, rv.pitch, ch));
        uint4 f1 = __ldg(frame_chunk(rv.frames, ps - 2, rv.frame_capacity, rv.pitch, ch));
        uint4 f2 = __ldg(frame_chunk(rv.frames, ps - 1, rv.frame_capacity, rv.pitch, ch));
        uint4 f3 = __ldg(frame_chunk(rv.frames, ps, rv.frame_capacity, rv.pitch, ch));
        store_stack16(reinterpret_cast<uint4*>(states) + out_off, f0, f1, f2, f3);
      }
      if (next_states) {
        uint4 f0 = __ldg(frame_chunk(rv.frames, pn - 3, rv.frame_capacity, rv.pitch, ch));
        uint4 f1 = __ldg(frame_chunk(rv.frames, pn - 2, rv.frame_capacity, rv.pitch, ch));
        uint4 f2 = __ldg(frame_chunk(rv.frames, pn - 1, rv.frame_capacity, rv.pitch, ch));
        uint4 f3 = __ldg(frame_chunk(rv.frames, pn, rv.frame_capacity, rv.pitch, ch));
        store_stack16(reinterpret_cast<uint4*>(next_states) + out_off, f0, f1, f2, f3);
      }
    }
  }
}

// Generic path: any frames_per_state, element size EB in {1,4,8}.  One thread = one frame element of
// one sampled transition; writes its S copies (one per stacked frame) for both stacks.
template <typename ET>
__global__ void __launch_bounds__(256)
gather_generic_kernel(RingView rv, const int32_t* __restrict__ idx, int32_t k, int32_t elems, ET* __restrict__ states,
                      ET* __restrict__ next_states, int32_t* __restrict__ actions, double* __restrict__ rewards,
                      uint8_t* __restrict__ terminals) {
  const int64_t total = (int64_t)k * elems;
  const int64_t head = rv.state[0];
  const int S = rv.frames_per_state;
  const TransitionRec* meta = static_cast<const TransitionRec*>(rv.meta);
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int32_t tr = (int32_t)(t / elems);
    const int32_t e = (int32_t)(t - (int64_t)tr * elems);
    int64_t slot = head + (int64_t)idx[tr];
    if (slot >= rv.capacity) slot -= rv.capacity;
    const TransitionRec* m = meta + slot;
    if (e == 0) {
      if (actions) actions[tr] = m->action;
      if (rewards) rewards[tr] = m->reward;
      if (terminals) terminals[tr] = m->terminal;
    }
    const int64_t o = ((int64_t)tr * elems + e) * S;
    for (int c = 0; c < S; ++c) {
      if (states) {
        int64_t fs = (m->state_pos - (S - 1) + c) % rv.frame_capacity;
        states[o + c] = reinterpret_cast<const ET*>(rv.frames + fs * (int64_t)rv.pitch)[e];
      }
      if (next_states) {
        int64_t fs = (m->next_pos - (S - 1) + c) % rv.frame_capacity;
        next_states[o + c] = reinterpret_cast<const ET*>(rv.frames + fs * (int64_t)rv.pitch)[e];
      }
    }
  }
}

// Graph-embedded ingest: this kernel reads a pinned staging buffer of the host DIRECTLY (zero-copy over PCIe: header +
// up to kGraphIngest records, up to kGraphIngest frames -- copy-engine memcpy nodes cost several microseconds each
// inside a graph) and scatters it into the ring.  Block j moves frame j; block 0 also writes the records, the ring's
// {head, size} and copies the sample indices of the update from pinned memory to the device.
struct IngestHdr {
  int32_t nframes, nrecs;
  int64_t first_frame_pos, first_rec, head, size;
  int64_t pad[3];
};
static_assert(sizeof(IngestHdr) == 64, "header is one 64-byte block");

__global__ void __launch_bounds__(256)
ingest_kernel(const uint8_t* __restrict__ meta_blk, const uint8_t* __restrict__ frames_blk, uint8_t* __restrict__ d_frames,
              TransitionRec* __restrict__ d_meta, int64_t* __restrict__ d_state, int64_t fcap, int64_t cap, int32_t pitch,
              const int32_t* __restrict__ host_idx, int32_t* __restrict__ d_idx, int32_t k) {
  pdl_trigger();                                  // the gather may launch now; it waits for this grid's completion itself
  // grid (ceil(pitch/16/256), kGraphIngest): block (x, j) moves 256 16-byte chunks of frame slot j.  Every PCIe read is
  // issued before the header is examined (a stale frame slot is harmless to read), so a block pays one host round
  // trip, not two.
  const int j = blockIdx.y, i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool first = blockIdx.x == 0 && blockIdx.y == 0;
  uint4 v = make_uint4(0u, 0u, 0u, 0u), rec = make_uint4(0u, 0u, 0u, 0u);
  int32_t myidx = 0;
  if (i < pitch / 16) v = reinterpret_cast<const uint4*>(frames_blk + (size_t)j * pitch)[i];
  if (first && (int)threadIdx.x < 16) rec = reinterpret_cast<const uint4*>(meta_blk + sizeof(IngestHdr))[threadIdx.x];   // 8 records
  if (first && host_idx)
    for (int q = threadIdx.x; q < k; q += blockDim.x) { myidx = host_idx[q]; d_idx[q] = myidx; }
  const IngestHdr hd = *reinterpret_cast<const IngestHdr*>(meta_blk);
  if (j < hd.nframes && i < pitch / 16)
    reinterpret_cast<uint4*>(d_frames + ((hd.first_frame_pos + j) % fcap) * (int64_t)pitch)[i] = v;
  if (first && (hd.nframes > 0 || hd.nrecs > 0)) {
    if ((int)threadIdx.x < hd.nrecs * 2)           // a record is two 16-byte halves
      reinterpret_cast<uint4*>(d_meta + (hd.first_rec + (threadIdx.x >> 1)) % cap)[threadIdx.x & 1] = rec;
    if (threadIdx.x == 0) { d_state[0] = hd.head; d_state[1] = hd.size; }
  }
}

// Lock-step actors (dqn_replay_add_batch): copy op list -> frame ring.  src >= 0: frame index in the caller's batch;
// src < 0: ring slot -src-1 (re-storing the tail of a stack at the start of an environment's lane).
struct LaneOp { int64_t src, dst; };
__global__ void __launch_bounds__(256)
lane_scatter_kernel(const LaneOp* __restrict__ ops, const uint8_t* __restrict__ batch, int64_t batch_pitch,
                    uint8_t* __restrict__ frames, int32_t pitch) {
  const LaneOp op = ops[blockIdx.y];
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= pitch / 16) return;
  const uint8_t* src = op.src >= 0 ? batch + op.src * batch_pitch : frames + (-op.src - 1) * (int64_t)pitch;
  // a batch frame has frame_bytes valid bytes; the ring's pitch rounds that up to 16: the tail chunk may read a few
  // bytes past the frame inside the caller's (larger) allocation only when batch_pitch == pitch, which the host enforces
  reinterpret_cast<uint4*>(frames + op.dst * (int64_t)pitch)[i] = reinterpret_cast<const uint4*>(src)[i];
}

// Stacks ending at explicit ring positions (the actors' current states): out[n][elems][S], same layout as the gather.
__global__ void __launch_bounds__(256)
stack_at_positions_kernel(RingView rv, const int64_t* __restrict__ pos, int32_t n, int32_t elems, uint8_t* __restrict__ out) {
  const int S = rv.frames_per_state;
  const int64_t total = (int64_t)n * elems;
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
    const int32_t e = (int32_t)(t / elems), i = (int32_t)(t - (int64_t)e * elems);
    for (int c = 0; c < S; ++c) {
      const int64_t fs = (pos[e] - (S - 1) + c) % rv.frame_capacity;
      out[((int64_t)e * elems + i) * S + c] = rv.frames[fs * (int64_t)rv.pitch + i];
    }
  }
}

}  // namespace dqn

// ---------------------------------------------------------------------------------------------
// Host side of the ring
// ---------------------------------------------------------------------------------------------
using namespace dqn;

namespace {
constexpr int kStageFrames = 64;       // frames per pinned staging buffer
constexpr int kStageRecs = 1024;       // transition records per pinned staging buffer
constexpr int kStages = 3;             // pinned staging buffers (the host may run one update ahead of the GPU)
constexpr int kGraphIngest = 8;        // frames / records a step graph ingests per replay (dqn_replay_seal)
constexpr size_t kIngestMetaBytes = sizeof(IngestHdr) + kGraphIngest * sizeof(TransitionRec);

struct Stage {
  uint8_t* frames = nullptr;   // pinned, kStageFrames x pitch
  TransitionRec* recs = nullptr;   // pinned
  int64_t* ring_state = nullptr;   // pinned {head,size}
  uint8_t* ing = nullptr;          // pinned: IngestHdr + kGraphIngest records (graph-embedded ingest)
  int64_t first_frame_pos = 0;
  int nframes = 0;
  int64_t first_rec = 0;   // total_adds index of recs[0]
  int nrecs = 0;
  cudaEvent_t done = nullptr;
  bool in_flight = false;
};
}  // namespace

struct dqn_replay_s {
  int device = 0;
  int64_t cap = 0, fcap = 0;
  int32_t frame_bytes = 0, pitch = 0, elem_bytes = 1, S = 1;
  uint8_t* d_frames = nullptr;
  TransitionRec* d_meta = nullptr;
  int64_t* d_state = nullptr;
  int64_t total_adds = 0;     // transitions ever added
  int64_t total_frames = 0;   // frames ever appended (= next frame position)
  std::vector<int64_t> need;  // per record slot: oldest frame position the transition references
  int64_t last_need = 0;
  Stage stage[kStages];
  int cur = 0;
  cudaStream_t last_stream = nullptr;
  int sticky = 0;
  int num_sms = 148;
  // lock-step actor lanes (dqn_replay_set_lanes / dqn_replay_add_batch): environment e owns the frame slots
  // [base, base + cap) and writes them in order, so the frames of its stacks stay consecutive although E environments
  // append in turn.  `written` counts the slots consumed so far (virtual, never wraps), `tail` the oldest still needed.
  struct Lane { int64_t base, cap, written, tail, last_pos, state_vlo; bool has_state; };   // state_vlo: virtual index of the oldest frame of the current state stack
  std::vector<Lane> lanes;
  std::vector<int32_t> rec_lane;     // per record slot: lane of the transition ...
  std::vector<int64_t> rec_vlo;      // ... and the virtual index of the oldest frame it references
  dqn::LaneOp* ops_host[2] = {nullptr, nullptr};   // pinned
  dqn::LaneOp* ops_dev[2] = {nullptr, nullptr};
  int64_t* pos_host = nullptr;       // pinned: current-state positions for dqn_replay_current_states
  int64_t* pos_dev = nullptr;
  cudaEvent_t ops_done[2] = {nullptr, nullptr};
  int ops_turn = 0, ops_cap = 0;
};

static int64_t live_size(const dqn_replay_s* h) { return h->total_adds < h->cap ? h->total_adds : h->cap; }
static int64_t oldest_needed(const dqn_replay_s* h) {
  int64_t sz = live_size(h);
  if (sz == 0) return h->total_frames;
  int64_t head = (h->total_adds - sz) % h->cap;
  return h->need[head];
}

static int wait_stage(Stage& s) {
  if (s.in_flight) {
    DQN_CUDA(cudaEventSynchronize(s.done));
    s.in_flight = false;
  }
  return 0;
}

static int flush_impl(dqn_replay_s* h, cudaStream_t stream) {
  Stage& s = h->stage[h->cur];
  if (s.nframes == 0 && s.nrecs == 0) return 0;
  // frames: contiguous in position space, at most two ranges in slot space
  int64_t p = s.first_frame_pos;
  int left = s.nframes;
  const uint8_t* src = s.frames;
  while (left > 0) {
    int64_t slot = p % h->fcap;
    int n = (int)((h->fcap - slot) < left ? (h->fcap - slot) : left);
    DQN_CUDA(cudaMemcpyAsync(h->d_frames + slot * (int64_t)h->pitch, src, (size_t)n * h->pitch,
                             cudaMemcpyHostToDevice, stream));
    src += (size_t)n * h->pitch; p += n; left -= n;
  }
  int64_t r = s.first_rec;
  left = s.nrecs;
  const TransitionRec* rsrc = s.recs;
  while (left > 0) {
    int64_t slot = r % h->cap;
    int n = (int)((h->cap - slot) < left ? (h->cap - slot) : left);
    DQN_CUDA(cudaMemcpyAsync(h->d_meta + slot, rsrc, (size_t)n * sizeof(TransitionRec), cudaMemcpyHostToDevice, stream));
    rsrc += n; r += n; left -= n;
  }
  int64_t sz = live_size(h);
  s.ring_state[0] = (h->total_adds - sz) % h->cap;
  s.ring_state[1] = sz;
  DQN_CUDA(cudaMemcpyAsync(h->d_state, s.ring_state, 2 * sizeof(int64_t), cudaMemcpyHostToDevice, stream));
  DQN_CUDA(cudaEventRecord(s.done, stream));
  s.in_flight = true;
  s.nframes = s.nrecs = 0;
  h->cur = (h->cur + 1) % kStages;
  h->last_stream = stream;
  int rc = wait_stage(h->stage[h->cur]);   // the buffer we are about to fill must be idle
  if (rc) return rc;
  return 0;
}

extern "C" const char* dqn_last_error(void) { return dqn::err_buf(); }
extern "C" int32_t dqn_version(void) { return 100; }

extern "C" int dqn_replay_create(dqn_replay_t* out, int64_t capacity, int64_t frame_capacity, int32_t frame_bytes,
                                 int32_t elem_bytes, int32_t frames_per_state, int32_t device) {
  DQN_REQUIRE(out != nullptr, DQN_ERR_INVALID, "dqn_replay_create: out is NULL");
  *out = nullptr;
  DQN_REQUIRE(capacity > 0 && capacity < (int64_t)1 << 31, DQN_ERR_INVALID, "capacity must be in [1, 2^31)");
  DQN_REQUIRE(frames_per_state > 0, DQN_ERR_INVALID, "Frames per state should be greater than 0");
  DQN_REQUIRE(elem_bytes == 1 || elem_bytes == 4 || elem_bytes == 8, DQN_ERR_INVALID, "elem_bytes must be 1, 4 or 8");
  DQN_REQUIRE(frame_bytes > 0 && frame_bytes % elem_bytes == 0, DQN_ERR_INVALID,
              "frame_bytes must be a positive multiple of elem_bytes");
  if (frame_capacity <= 0) frame_capacity = 2 * capacity + 2 * (int64_t)frames_per_state;
  DQN_REQUIRE(frame_capacity >= 2 * (int64_t)frames_per_state + 1, DQN_ERR_INVALID, "frame_capacity too small");
  int ndev = 0;
  DQN_CUDA(cudaGetDeviceCount(&ndev));
  DQN_REQUIRE(device >= 0 && device < ndev, DQN_ERR_INVALID, "no CUDA device %d (found %d); there is no CPU path",
              device, ndev);
  DeviceGuard g(device);
  DQN_REQUIRE(g.ok, DQN_ERR_CUDA, "cannot select device %d", device);
  dqn_replay_s* h = new (std::nothrow) dqn_replay_s();
  DQN_REQUIRE(h != nullptr, DQN_ERR_INVALID, "out of host memory");
  h->device = device; h->cap = capacity; h->fcap = frame_capacity; h->frame_bytes = frame_bytes;
  h->pitch = (int32_t)round_up(frame_bytes, 16); h->elem_bytes = elem_bytes; h->S = frames_per_state;
  h->need.assign((size_t)capacity, 0);
  cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
  cudaError_t e = cudaMalloc(&h->d_frames, (size_t)h->fcap * h->pitch);
  if (e == cudaSuccess) e = cudaMalloc(&h->d_meta, (size_t)h->cap * sizeof(TransitionRec));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_state, 2 * sizeof(int64_t));
  if (e == cudaSuccess) e = cudaMemset(h->d_state, 0, 2 * sizeof(int64_t));

  for (int i = 0; i < kStages && e == cudaSuccess; ++i) {
    e = cudaMallocHost(&h->stage[i].frames, (size_t)kStageFrames * h->pitch);
    if (e == cudaSuccess) e = cudaMallocHost(&h->stage[i].recs, kStageRecs * sizeof(TransitionRec));
    if (e == cudaSuccess) e = cudaMallocHost(&h->stage[i].ring_state, 2 * sizeof(int64_t));
    if (e == cudaSuccess) e = cudaMallocHost(&h->stage[i].ing, kIngestMetaBytes);
    if (e == cudaSuccess) memset(h->stage[i].ing, 0, kIngestMetaBytes);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->stage[i].done, cudaEventDisableTiming);
  }
  if (e != cudaSuccess) {
    int rc = set_err(DQN_ERR_CUDA, "dqn_replay_create: %s (ring of %lld frames x %d B)", cudaGetErrorString(e),
                     (long long)h->fcap, h->pitch);
    dqn_replay_destroy(h);
    return rc;
  }
  *out = h;
  return 0;
}

extern "C" int dqn_replay_destroy(dqn_replay_t h) {
  if (!h) return 0;
  DeviceGuard g(h->device);
  cudaDeviceSynchronize();
  cudaFree(h->d_frames); cudaFree(h->d_meta); cudaFree(h->d_state);

  for (int i = 0; i < 2; ++i) {
    if (h->ops_host[i]) cudaFreeHost(h->ops_host[i]);
    if (h->ops_dev[i]) cudaFree(h->ops_dev[i]);
    if (h->ops_done[i]) cudaEventDestroy(h->ops_done[i]);
  }
  if (h->pos_host) cudaFreeHost(h->pos_host);
  if (h->pos_dev) cudaFree(h->pos_dev);
  for (int i = 0; i < kStages; ++i) {
    if (h->stage[i].frames) cudaFreeHost(h->stage[i].frames);
    if (h->stage[i].recs) cudaFreeHost(h->stage[i].recs);
    if (h->stage[i].ring_state) cudaFreeHost(h->stage[i].ring_state);
    if (h->stage[i].ing) cudaFreeHost(h->stage[i].ing);
    if (h->stage[i].done) cudaEventDestroy(h->stage[i].done);
  }
  delete h;
  return 0;
}

static int reserve_frames(dqn_replay_s* h, int32_t n) {
  DQN_REQUIRE(n > 0, DQN_ERR_INVALID, "append_frames: n must be positive");
  DQN_REQUIRE(h->lanes.empty(), DQN_ERR_STATE, "append_frames: this replay memory is in lock-step actor mode (dqn_replay_set_lanes)");
  int64_t need = oldest_needed(h);
  DQN_REQUIRE(h->total_frames + n - need <= h->fcap, DQN_ERR_OVERFLOW,
              "frame ring overflow: %lld frames would be live but frame_capacity is %lld; create the replay "
              "with a larger frame_capacity",
              (long long)(h->total_frames + n - need), (long long)h->fcap);
  return 0;
}

extern "C" int dqn_replay_append_frames(dqn_replay_t h, const void* host_frames, int32_t n, int64_t* first_pos) {
  DQN_REQUIRE(h && host_frames, DQN_ERR_INVALID, "append_frames: NULL argument");
  int rc = reserve_frames(h, n);
  if (rc) return rc;
  DeviceGuard g(h->device);
  if (first_pos) *first_pos = h->total_frames;
  const uint8_t* src = static_cast<const uint8_t*>(host_frames);
  for (int i = 0; i < n; ++i) {
    Stage* s = &h->stage[h->cur];
    if (s->nframes == kStageFrames) {
      // staging full: push it out now.  Synchronous with respect to the last stream used, so the
      // ordering against kernels already enqueued there is preserved.
      rc = flush_impl(h, h->last_stream);
      if (rc) return rc;
      s = &h->stage[h->cur];
    }
    if (s->nframes == 0) s->first_frame_pos = h->total_frames;
    memcpy(s->frames + (size_t)s->nframes * h->pitch, src + (size_t)i * h->frame_bytes, (size_t)h->frame_bytes);
    s->nframes++;
    h->total_frames++;
  }
  return 0;
}

extern "C" int dqn_replay_append_frames_device(dqn_replay_t h, const void* d_frames, int32_t n, int64_t* first_pos,
                                               dqn_stream_t stream) {
  DQN_REQUIRE(h && d_frames, DQN_ERR_INVALID, "append_frames_device: NULL argument");
  int rc = reserve_frames(h, n);
  if (rc) return rc;
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // keep position order: anything staged on the host goes first
  rc = flush_impl(h, st);
  if (rc) return rc;
  if (first_pos) *first_pos = h->total_frames;
  int64_t p = h->total_frames;
  int left = n;
  const uint8_t* src = static_cast<const uint8_t*>(d_frames);
  while (left > 0) {
    int64_t slot = p % h->fcap;
    int m = (int)((h->fcap - slot) < left ? (h->fcap - slot) : left);
    if (h->pitch == h->frame_bytes) {
      DQN_CUDA(cudaMemcpyAsync(h->d_frames + slot * (int64_t)h->pitch, src, (size_t)m * h->pitch,
                               cudaMemcpyDeviceToDevice, st));
    } else {
      DQN_CUDA(cudaMemcpy2DAsync(h->d_frames + slot * (int64_t)h->pitch, h->pitch, src, h->frame_bytes,
                                 h->frame_bytes, m, cudaMemcpyDeviceToDevice, st));
    }
    src += (size_t)m * h->frame_bytes; p += m; left -= m;
  }
  h->total_frames += n;
  h->last_stream = st;
  return 0;
}

extern "C" int dqn_replay_add(dqn_replay_t h, int64_t state_pos, int64_t next_pos, int32_t action, double reward,
                              uint8_t terminal) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "add: NULL handle");
  DQN_REQUIRE(h->lanes.empty(), DQN_ERR_STATE, "add: this replay memory is in lock-step actor mode (dqn_replay_set_lanes); use dqn_replay_add_batch");
  const int64_t S = h->S;
  int64_t lo = (state_pos < next_pos ? state_pos : next_pos) - (S - 1);
  int64_t hi = state_pos > next_pos ? state_pos : next_pos;
  DQN_REQUIRE(lo >= 0 && hi < h->total_frames, DQN_ERR_INVALID,
              "add: stack positions (%lld, %lld) outside the appended frames [0, %lld)", (long long)state_pos,
              (long long)next_pos, (long long)h->total_frames);
  DQN_REQUIRE(h->total_frames - lo <= h->fcap, DQN_ERR_INVALID, "add: frames of this transition were already overwritten");
  DQN_REQUIRE(lo >= h->last_need || h->total_adds == 0, DQN_ERR_INVALID,
              "add: a transition may not reference frames older than the previous transition's oldest frame");
  DeviceGuard g(h->device);
  Stage* s = &h->stage[h->cur];
  if (s->nrecs == kStageRecs) {
    int rc = flush_impl(h, h->last_stream);
    if (rc) return rc;
    s = &h->stage[h->cur];
  }
  if (s->nrecs == 0) s->first_rec = h->total_adds;
  TransitionRec& r = s->recs[s->nrecs++];
  r.state_pos = state_pos; r.next_pos = next_pos; r.reward = reward; r.action = action; r.terminal = terminal ? 1 : 0;
  r.pad[0] = r.pad[1] = r.pad[2] = 0;
  h->need[(size_t)(h->total_adds % h->cap)] = lo;
  h->last_need = lo;
  h->total_adds++;
  return 0;
}

// Stages one transition record (the part of dqn_replay_add after its position checks).
static int stage_record(dqn_replay_s* h, int64_t state_pos, int64_t next_pos, int32_t action, double reward, uint8_t terminal) {
  Stage* s = &h->stage[h->cur];
  if (s->nrecs == kStageRecs) {
    int rc = flush_impl(h, h->last_stream);
    if (rc) return rc;
    s = &h->stage[h->cur];
  }
  if (s->nrecs == 0) s->first_rec = h->total_adds;
  TransitionRec& r = s->recs[s->nrecs++];
  r.state_pos = state_pos; r.next_pos = next_pos; r.reward = reward; r.action = action; r.terminal = terminal ? 1 : 0;
  r.pad[0] = r.pad[1] = r.pad[2] = 0;
  return 0;
}

extern "C" int dqn_replay_set_lanes(dqn_replay_t h, int32_t num_envs) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "set_lanes: NULL handle");
  DQN_REQUIRE(num_envs >= 1, DQN_ERR_INVALID, "set_lanes: num_envs must be positive");
  DQN_REQUIRE(h->total_adds == 0 && h->total_frames == 0, DQN_ERR_STATE, "set_lanes: the replay memory must be empty");
  DQN_REQUIRE(h->pitch == h->frame_bytes, DQN_ERR_UNSUPPORTED, "set_lanes: frame_bytes must be a multiple of 16");
  const int64_t cap = h->fcap / num_envs;
  DQN_REQUIRE(cap >= 3 * (int64_t)h->S + 2, DQN_ERR_INVALID,
              "set_lanes: %d environments leave %lld frame slots each; need at least %d (raise frame_capacity)", num_envs,
              (long long)cap, 3 * h->S + 2);
  DeviceGuard g(h->device);
  h->lanes.assign((size_t)num_envs, dqn_replay_s::Lane{0, cap, 0, 0, -1, 0, false});
  for (int e = 0; e < num_envs; ++e) h->lanes[(size_t)e].base = (int64_t)e * cap;
  h->rec_lane.assign((size_t)h->cap, -1);
  h->rec_vlo.assign((size_t)h->cap, 0);
  h->ops_cap = num_envs * (2 * h->S);
  for (int i = 0; i < 2; ++i) {
    DQN_CUDA(cudaMallocHost(&h->ops_host[i], (size_t)h->ops_cap * sizeof(LaneOp)));
    DQN_CUDA(cudaMalloc(&h->ops_dev[i], (size_t)h->ops_cap * sizeof(LaneOp)));
    DQN_CUDA(cudaEventCreateWithFlags(&h->ops_done[i], cudaEventDisableTiming));
  }
  DQN_CUDA(cudaMallocHost(&h->pos_host, (size_t)num_envs * sizeof(int64_t)));
  DQN_CUDA(cudaMalloc(&h->pos_dev, (size_t)num_envs * sizeof(int64_t)));
  return 0;
}

// Reserves n consecutive slots in a lane (never straddling the lane's end) and returns the slot of the first one.
static int lane_reserve(dqn_replay_s* h, dqn_replay_s::Lane& L, int n, int64_t* first_slot) {
  int64_t off = L.written % L.cap;
  int64_t skip = off + n > L.cap ? L.cap - off : 0;          // not enough room before the end: start over at the lane's base
  DQN_REQUIRE(L.written + skip + n - L.tail <= L.cap, DQN_ERR_OVERFLOW,
              "frame ring overflow in an actor lane: %lld frames would be live but the lane holds %lld; create the replay "
              "with a larger frame_capacity", (long long)(L.written + skip + n - L.tail), (long long)L.cap);
  L.written += skip;
  *first_slot = L.base + L.written % L.cap;
  L.written += n;
  return 0;
}

extern "C" int dqn_replay_add_batch(dqn_replay_t h, int32_t num_envs, const void* d_frames, const int32_t* actions,
                                    const double* rewards, const uint8_t* terminals, const uint8_t* episode_start,
                                    dqn_stream_t stream) {
  DQN_REQUIRE(h && d_frames && episode_start, DQN_ERR_INVALID, "add_batch: NULL argument");
  DQN_REQUIRE(!h->lanes.empty() && num_envs == (int)h->lanes.size(), DQN_ERR_STATE,
              "add_batch: call dqn_replay_set_lanes(h, %d) on the empty replay memory first", num_envs);
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int S = h->S;
  const int turn = h->ops_turn;
  DQN_CUDA(cudaEventSynchronize(h->ops_done[turn]));          // the op list about to be overwritten has been consumed
  LaneOp* ops = h->ops_host[turn];
  int nops = 0;
  for (int e = 0; e < num_envs; ++e) {
    dqn_replay_s::Lane& L = h->lanes[(size_t)e];
    int64_t slot = 0;
    if (episode_start[e]) {
      // FrameBuffer.clear() + append(env.reset()): the window is the first frame S times (frame_buffer.py:35-36); no transition yet
      int rc = lane_reserve(h, L, S, &slot);
      if (rc) return rc;
      for (int c = 0; c < S; ++c) ops[nops++] = LaneOp{e, slot + c};
      L.last_pos = slot + S - 1;
      L.state_vlo = L.written - S;
      L.has_state = true;
      continue;
    }
    DQN_REQUIRE(L.has_state, DQN_ERR_STATE, "add_batch: environment %d has no current state; its first frame must come with episode_start", e);
    DQN_REQUIRE(actions && rewards && terminals, DQN_ERR_INVALID, "add_batch: NULL actions / rewards / terminals");
    const int64_t spos = L.last_pos, svlo = L.state_vlo;
    int64_t npos;
    if ((L.written % L.cap) != 0 && L.base + L.written % L.cap == spos + 1) {
      int rc = lane_reserve(h, L, 1, &slot);                  // the usual case: one new frame right behind the stack
      if (rc) return rc;
      ops[nops++] = LaneOp{e, slot};
      npos = slot;
      L.state_vlo = svlo + 1;
    } else {
      // the lane wrapped: the next-state stack is re-stored in full at its start (S-1 old frames + the new one)
      int rc = lane_reserve(h, L, S, &slot);
      if (rc) return rc;
      for (int c = 0; c < S - 1; ++c) ops[nops++] = LaneOp{-(spos - (S - 2) + c) - 1, slot + c};
      ops[nops++] = LaneOp{e, slot + S - 1};
      npos = slot + S - 1;
      L.state_vlo = L.written - S;
    }
    // record; the record slot being overwritten frees its lane's frames up to its own oldest one
    const size_t rslot = (size_t)(h->total_adds % h->cap);
    if (h->total_adds >= h->cap && h->rec_lane[rslot] >= 0) {
      dqn_replay_s::Lane& old = h->lanes[(size_t)h->rec_lane[rslot]];
      if (h->rec_vlo[rslot] > old.tail) old.tail = h->rec_vlo[rslot];
    }
    int rc = stage_record(h, spos, npos, actions[e], rewards[e], terminals[e]);
    if (rc) return rc;
    h->rec_lane[rslot] = e;
    h->rec_vlo[rslot] = svlo;                                  // oldest frame this transition references
    h->total_adds++;
    L.last_pos = npos;
  }
  DQN_CUDA(cudaMemcpyAsync(h->ops_dev[turn], ops, (size_t)nops * sizeof(LaneOp), cudaMemcpyHostToDevice, st));
  if (nops > 0) {
    count_launch();
    lane_scatter_kernel<<<dim3((h->pitch / 16 + 255) / 256, nops), 256, 0, st>>>(h->ops_dev[turn], static_cast<const uint8_t*>(d_frames),
                                                                               h->frame_bytes, h->d_frames, h->pitch);
    DQN_CUDA(cudaGetLastError());
  }
  DQN_CUDA(cudaEventRecord(h->ops_done[turn], st));
  h->ops_turn ^= 1;
  h->last_stream = st;
  return 0;
}

extern "C" int dqn_replay_current_states(dqn_replay_t h, void* d_states, dqn_stream_t stream) {
  DQN_REQUIRE(h && d_states, DQN_ERR_INVALID, "current_states: NULL argument");
  DQN_REQUIRE(!h->lanes.empty(), DQN_ERR_STATE, "current_states: no actor lanes (dqn_replay_set_lanes)");
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int E = (int)h->lanes.size();
  DQN_CUDA(cudaStreamSynchronize(st));                        // pos_host is reused by every call
  for (int e = 0; e < E; ++e) {
    DQN_REQUIRE(h->lanes[(size_t)e].has_state, DQN_ERR_STATE, "current_states: environment %d has no current state", e);
    h->pos_host[e] = h->lanes[(size_t)e].last_pos;
  }
  DQN_CUDA(cudaMemcpyAsync(h->pos_dev, h->pos_host, (size_t)E * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  RingView rv;
  dqn_replay_view_(h, &rv);
  const int elems = h->frame_bytes;
  DQN_REQUIRE(h->elem_bytes == 1, DQN_ERR_UNSUPPORTED, "current_states: uint8 frames only");
  int64_t blocks = ((int64_t)E * elems + 255) / 256;
  if (blocks > (int64_t)h->num_sms * 32) blocks = (int64_t)h->num_sms * 32;
  count_launch();
  stack_at_positions_kernel<<<(unsigned)blocks, 256, 0, st>>>(rv, h->pos_dev, E, elems, static_cast<uint8_t*>(d_states));
  DQN_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dqn_replay_push(dqn_replay_t h, const void* host_frame, int32_t action, double reward, uint8_t terminal) {
  DQN_REQUIRE(h && host_frame, DQN_ERR_INVALID, "push: NULL argument");
  DQN_REQUIRE(h->total_frames >= h->S, DQN_ERR_STATE, "push: the ring does not hold a complete state stack yet");
  int64_t pos = 0;
  int rc = dqn_replay_append_frames(h, host_frame, 1, &pos);
  if (rc) return rc;
  return dqn_replay_add(h, pos - 1, pos, action, reward, terminal);
}

extern "C" int dqn_replay_flush(dqn_replay_t h, dqn_stream_t stream) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "flush: NULL handle");
  DeviceGuard g(h->device);
  return flush_impl(h, static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_replay_seal(dqn_replay_t h, int32_t* stage, dqn_stream_t stream) {
  DQN_REQUIRE(h && stage, DQN_ERR_INVALID, "seal: NULL argument");
  DeviceGuard g(h->device);
  const int i = h->cur;
  Stage& s = h->stage[i];
  IngestHdr* hd = reinterpret_cast<IngestHdr*>(s.ing);
  *stage = i;
  if (s.nframes > kGraphIngest || s.nrecs > kGraphIngest) {
    // too much for the fixed-size copy of the graph: flush now (eager copies), the graph's ingest sees an empty header
    hd->nframes = 0; hd->nrecs = 0;
    return flush_impl(h, stream ? static_cast<cudaStream_t>(stream) : h->last_stream);   // NULL: the stream of the last gather / flush
  }
  hd->nframes = s.nframes; hd->nrecs = s.nrecs;
  hd->first_frame_pos = s.first_frame_pos; hd->first_rec = s.first_rec;
  const int64_t sz = live_size(h);
  hd->head = (h->total_adds - sz) % h->cap; hd->size = sz;
  if (s.nrecs) memcpy(s.ing + sizeof(IngestHdr), s.recs, (size_t)s.nrecs * sizeof(TransitionRec));
  s.nframes = s.nrecs = 0;
  s.in_flight = false;                  // completion of the graph that reads this buffer is the caller's to track
  h->cur = (h->cur + 1) % kStages;
  if (stream) h->last_stream = static_cast<cudaStream_t>(stream);
  return wait_stage(h->stage[h->cur]);  // the buffer we are about to fill may still be the source of an eager flush
}

extern "C" int dqn_replay_ingest_staged(dqn_replay_t h, int32_t stage, const int32_t* pinned_idx, int32_t* d_idx, int32_t k,
                                        dqn_stream_t stream) {
  DQN_REQUIRE(h && stage >= 0 && stage < kStages, DQN_ERR_INVALID, "ingest_staged: bad argument");
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  Stage& s = h->stage[stage];
  DQN_REQUIRE((pinned_idx == nullptr) == (d_idx == nullptr) && k >= 0, DQN_ERR_INVALID, "ingest_staged: index pointers");
  dqn::count_launch();
  static_assert(kGraphIngest * 2 <= 16, "block (0,0) moves the records with its first 16 threads");
  ingest_kernel<<<dim3((h->pitch / 16 + 255) / 256, kGraphIngest), 256, 0, st>>>(s.ing, s.frames, h->d_frames, h->d_meta, h->d_state,
                                                                              h->fcap, h->cap, h->pitch, pinned_idx, d_idx, k);
  DQN_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int dqn_replay_size(dqn_replay_t h, int64_t* size, int64_t* capacity) {
  DQN_REQUIRE(h, DQN_ERR_INVALID, "size: NULL handle");
  if (size) *size = live_size(h);
  if (capacity) *capacity = h->cap;
  return 0;
}

extern "C" int dqn_replay_next_frame_pos(dqn_replay_t h, int64_t* pos) {
  DQN_REQUIRE(h && pos, DQN_ERR_INVALID, "next_frame_pos: NULL argument");
  *pos = h->total_frames;
  return 0;
}

int dqn_replay_view_(dqn_replay_t h, dqn::RingView* out) {
  out->frames = h->d_frames; out->meta = h->d_meta; out->state = h->d_state; out->capacity = h->cap;
  out->frame_capacity = h->fcap; out->pitch = h->pitch; out->frame_bytes = h->frame_bytes;
  out->elem_bytes = h->elem_bytes; out->frames_per_state = h->S;
  return 0;
}
int dqn_replay_device_(dqn_replay_t h) { return h->device; }

// row (*cursor % rows) of a device-resident [rows][k] index table -> idx; *cursor += 1
__global__ void __launch_bounds__(256) index_table_next_kernel(const int32_t* __restrict__ table, long long rows, int k,
                                                               long long* __restrict__ cursor, int32_t* __restrict__ idx) {
  pdl_trigger();
  pdl_wait();
  const long long cur = *cursor;
  const int32_t* row = table + (size_t)(cur % rows) * (size_t)k;
  for (int i = threadIdx.x; i < k; i += blockDim.x) idx[i] = row[i];
  __syncthreads();                          // every thread has read the cursor
  if (threadIdx.x == 0) *cursor = cur + 1;
}

int dqn_index_table_next_(const int32_t* d_table, int64_t rows, int32_t k, int64_t* d_cursor, int32_t* d_idx, cudaStream_t st) {
  DQN_REQUIRE(d_table && d_cursor && d_idx && rows > 0 && k > 0, DQN_ERR_INVALID, "index_table_next: bad argument");
  dqn::count_launch();
  DQN_CUDA(launch_pdl(index_table_next_kernel, dim3(1), dim3(256), 0, st, 1, d_table, (long long)rows, (int)k,
                      reinterpret_cast<long long*>(d_cursor), d_idx));
  return 0;
}
extern "C" int dqn_index_table_next(const int32_t* d_table, int64_t rows, int32_t k, int64_t* d_cursor, int32_t* d_idx,
                                    dqn_stream_t stream) {
  return dqn_index_table_next_(d_table, rows, k, d_cursor, d_idx, static_cast<cudaStream_t>(stream));
}

extern "C" int dqn_replay_gather(dqn_replay_t h, const int32_t* d_logical_idx, int32_t k, void* d_states,
                                 void* d_next_states, int32_t* d_actions, double* d_rewards, uint8_t* d_terminals,
                                 dqn_stream_t stream) {
  DQN_REQUIRE(h && d_logical_idx, DQN_ERR_INVALID, "gather: NULL argument");
  DQN_REQUIRE(k > 0, DQN_ERR_INVALID, "gather: k must be positive");
  DQN_REQUIRE(live_size(h) >= 1, DQN_ERR_STATE, "gather: replay memory is empty");
  DeviceGuard g(h->device);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = flush_impl(h, st);
  if (rc) return rc;
  h->last_stream = st;
  RingView rv;
  dqn_replay_view_(h, &rv);
  const int threads = 256;
  if (h->S == 4 && h->elem_bytes == 1 && h->frame_bytes % 16 == 0 &&
      ((uintptr_t)d_states % 16 == 0) && ((uintptr_t)d_next_states % 16 == 0)) {
    const int chunks = h->frame_bytes / 16;
    int64_t total = (int64_t)k * chunks;
    int64_t blocks = (total + threads - 1) / threads;
    int64_t maxb = (int64_t)h->num_sms * 32;
    if (blocks > maxb) blocks = maxb;
    launch_pdl(gather_stack4_u8_kernel, dim3((unsigned)blocks), dim3(threads), 0, st, 1, rv, d_logical_idx, k, chunks,
               static_cast<uint8_t*>(d_states), static_cast<uint8_t*>(d_next_states), d_actions, d_rewards, d_terminals);
  } else {
    const int elems = h->frame_bytes / h->elem_bytes;
    int64_t total = (int64_t)k * elems;
    int64_t blocks = (total + threads - 1) / threads;
    int64_t maxb = (int64_t)h->num_sms * 32;
    if (blocks > maxb) blocks = maxb;
    if (h->elem_bytes == 1)
      gather_generic_kernel<uint8_t><<<(unsigned)blocks, threads, 0, st>>>(
          rv, d_logical_idx, k, elems, static_cast<uint8_t*>(d_states), static_cast<uint8_t*>(d_next_states),
          d_actions, d_rewards, d_terminals);
    else if (h->elem_bytes == 4)
      gather_generic_kernel<uint32_t><<<(unsigned)blocks, threads, 0, st>>>(
          rv, d_logical_idx, k, elems, static_cast<uint32_t*>(d_states), static_cast<uint32_t*>(d_next_states),
          d_actions, d_rewards, d_terminals);
    else
      gather_generic_kernel<uint64_t><<<(unsigned)blocks, threads, 0, st>>>(
          rv, d_logical_idx, k, elems, static_cast<uint64_t*>(d_states), static_cast<uint64_t*>(d_next_states),
          d_actions, d_rewards, d_terminals);
  }
  count_launch();
  DQN_CUDA(cudaGetLastError());
  return 0;
}
