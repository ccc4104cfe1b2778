/*
 * dqn_b200.h -- C ABI of libdqn_b200.so: the DQN minibatch-update hot path of hbfs/dist-dqn,
 * rebuilt as hand-written CUDA for sm_100a (B200).
 *
 * The reference has no FFI of its own: its boundary is a Python class surface.
 * Every entry point below therefore cites the reference Python interface (file:line under the
 * reference's src/) whose work it takes over; the Python shells in dist-dqn_b200/ keep those
 * class/method names and call down to this ABI through ctypes.
 *
 * Conventions
 *  - every function returns 0 on success and a negative dqn_status on failure; the message is
 *    available (thread-local) from dqn_last_error().  Nothing throws across the boundary.
 *  - "d_" pointers are device pointers owned by the caller (e.g. torch tensors' data_ptr()); they
 *    must stay alive until the stream reaches the work enqueued by the call.
 *  - all work is enqueued on the caller's stream (a cudaStream_t passed as void*; NULL = the legacy
 *    default stream).  No hidden synchronisation except where stated.
 *  - a handle belongs to one GPU and is not thread-safe (cf. replay_memory.py:11-17).
 *  - there is no CPU fallback anywhere: without a CUDA device every compute call fails.
 */
#ifndef DQN_B200_H_
#define DQN_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct dqn_replay_s* dqn_replay_t;
typedef struct dqn_net_s* dqn_net_t;
typedef struct dqn_preproc_s* dqn_preproc_t;
typedef void* dqn_stream_t; /* cudaStream_t */

enum dqn_status {
  DQN_OK = 0,
  DQN_ERR_INVALID = -1,  /* bad argument */
  DQN_ERR_CUDA = -2,     /* a CUDA runtime call failed (sticky on the handle) */
  DQN_ERR_STATE = -3,    /* call not valid in the current state (e.g. sampling more than size()) */
  DQN_ERR_OVERFLOW = -4, /* frame ring would overwrite frames still referenced by live transitions */
  DQN_ERR_UNSUPPORTED = -5
};

enum dqn_arch {
  DQN_ARCH_SIMPLE = 0, /* network.py:258-313  SimpleNetwork: in -> 20 tanh -> 20 tanh -> A */
  DQN_ARCH_CNN = 1,    /* network.py:317-424  ConvNetwork: 3 x (conv SAME + relu + 2x2 max-pool), fc256, A */
  DQN_ARCH_NATURE = 2  /* Mnih 2015 variant (VALID convs, no pooling, fc512); north_star sweep, not in the reference */
};

enum dqn_dtype { DQN_U8 = 0, DQN_F32 = 1, DQN_F64 = 2 };

enum dqn_optimizer { /* network.py:169-177 */
  DQN_OPT_SGD = 0, DQN_OPT_MOMENTUM = 1, DQN_OPT_RMSPROP = 2, DQN_OPT_ADAM = 3,
  DQN_OPT_ADAGRAD = 4, DQN_OPT_ADADELTA = 5, DQN_OPT_FTRL = 6
};

enum dqn_loss { DQN_LOSS_MSE = 0 /* network.py:150 */, DQN_LOSS_HUBER = 1 /* north_star variant, delta=1 */ };

enum dqn_buffer { /* flat fp32 buffers behind a net handle, all in the reference's tuple order */
  DQN_BUF_ONLINE = 0, DQN_BUF_TARGET = 1, DQN_BUF_GRAD = 2, DQN_BUF_SLOT0 = 3, DQN_BUF_SLOT1 = 4
};

/* Hyper-parameters of one update; mirrors the argparse namespace the reference passes around as
 * `config` (main.py:28-39, 42-69): lr, reg_param, optimizer, momentum, rmsprop_decay, reward_discount. */
typedef struct dqn_hyper {
  double gamma;        /* --reward_discount; used in float64 like dqn_agent.py:246 */
  float lr;            /* --lr */
  float reg_param;     /* --reg_param (L2 on the fully-connected weights, network.py:151-152, 311, 407) */
  float momentum;      /* --momentum (MomentumOptimizer only) */
  float rmsprop_decay; /* --rmsprop_decay */
  int32_t optimizer;   /* enum dqn_optimizer */
  int32_t loss;        /* enum dqn_loss */
  float grad_scale;    /* multiplies the data gradient before the update: 1/world_size after an
                          all-reduce(sum) across data-parallel ranks (SyncReplicas mean, network.py:187-192) */
  int32_t reserved;
} dqn_hyper_t;

const char* dqn_last_error(void);
int32_t dqn_version(void);

/* ------------------------------------------------------------------------------------------------
 * Replay ring -- replaces ReplayMemory (replay_memory.py:18-53) and the frame stacking that
 * FrameBuffer.get_state() materialises per transition (frame_buffer.py:40-51).
 *
 * HBM layout: ONE copy of every frame in a byte ring (frame_capacity x frame_bytes) plus a
 * 32-byte record per transition {state_pos, next_pos, reward f64, action i32, terminal u8};
 * state_pos/next_pos are the ring positions of the NEWEST frame of each stack, whose
 * frames_per_state frames are consecutive.  Logical index 0 is the oldest live transition and
 * the oldest is evicted on every add when full, exactly like deque(maxlen) (replay_memory.py:20-23).
 * ------------------------------------------------------------------------------------------------ */

/* ReplayMemory.__init__ (replay_memory.py:19-20).  frame_bytes: bytes of one preprocessed frame
 * (84*84 for Atari, 8*obs_dim for float64 vector observations), must be a multiple of
 * elem_bytes; elem_bytes in {1,4,8} is the size of one frame element (stack interleave unit);
 * frame_capacity 0 = 2*capacity + 2*frames_per_state. */
int dqn_replay_create(dqn_replay_t* out, int64_t capacity, int64_t frame_capacity, int32_t frame_bytes,
                      int32_t elem_bytes, int32_t frames_per_state, int32_t device);
int dqn_replay_destroy(dqn_replay_t h);

/* Appends n consecutive frames from HOST memory to the frame ring (staged in pinned memory, sent
 * by the next flush); *first_pos receives the ring position of the first one.  Together with
 * dqn_replay_add this is ReplayMemory.add (replay_memory.py:22-23): the Python shell appends only
 * the frames it has not stored yet (normally one per env step). */
int dqn_replay_append_frames(dqn_replay_t h, const void* host_frames, int32_t n, int64_t* first_pos);
/* Same, from DEVICE memory (device-side actors; enqueued on stream immediately). */
int dqn_replay_append_frames_device(dqn_replay_t h, const void* d_frames, int32_t n, int64_t* first_pos,
                                    dqn_stream_t stream);
/* Records one transition whose state / next_state stacks end at the given frame positions. */
int dqn_replay_add(dqn_replay_t h, int64_t state_pos, int64_t next_pos, int32_t action, double reward,
                   uint8_t terminal);
/* The common case of ReplayMemory.add in one call (dqn_agent.py:96-100: the transition's state is the previous
 * transition's next_state): appends ONE host frame and records the transition from the stack that ended at the
 * previous newest frame to the stack ending at the new one. */
int dqn_replay_push(dqn_replay_t h, const void* host_frame, int32_t action, double reward, uint8_t terminal);
/* Sends staged frames + records to the device (async H2D on stream). Called by gather itself. */
int dqn_replay_flush(dqn_replay_t h, dqn_stream_t stream);

/* Graph-embedded ingest (no reference counterpart: it removes the per-update host calls of the H2D path that
 * ReplayMemory.add feeds, replay_memory.py:22-23).
 * dqn_replay_seal is host-only work: it publishes the header of the current pinned staging buffer (frame / record
 * counts and positions, ring head and size), switches staging to the next of the 3 buffers (round robin) and returns
 * the index (0..2) of the sealed one in *stage.  A buffer holding more than 8 frames or 8 records is flushed eagerly on `stream` instead and
 * its header reads empty.  Before the next dqn_replay_push the caller must know that the last
 * dqn_replay_ingest_staged of the NEXT buffer ((*stage + 1) % 3) has completed (its pinned memory is refilled from now
 * on).
 * dqn_replay_ingest_staged enqueues ONE kernel with fixed addresses (capturable in a CUDA graph) that reads staging
 * buffer `stage` straight from pinned host memory and scatters its frames / records into the ring; if pinned_idx is
 * given (pinned host memory, k int32) it also copies the update's sample indices to d_idx. */
int dqn_replay_seal(dqn_replay_t h, int32_t* stage, dqn_stream_t stream);
int dqn_replay_ingest_staged(dqn_replay_t h, int32_t stage, const int32_t* pinned_idx, int32_t* d_idx, int32_t k,
                             dqn_stream_t stream);
/* Lock-step actors on the device: the per-step actor work of the reference -- FrameBuffer.append +
 * get_state + ReplayMemory.add for every environment (dqn_agent.py:96-100, frame_buffer.py:27-51) -- for E environments
 * that step together, with no per-transition host call and no frame passing through the host.
 *   dqn_replay_set_lanes(h, E): once, on the empty ring.  Environment e gets its own run of frame slots, so that the
 *     frames of its stacks stay consecutive although the E environments append in turn.  Afterwards the ring only takes
 *     dqn_replay_add_batch (dqn_replay_add / append_frames / push are refused).
 *   dqn_replay_add_batch: one preprocessed frame per environment, d_frames = device [E][frame_bytes], in environment order.
 *     episode_start[e] != 0: the frame is env.reset()'s observation -- FrameBuffer.clear() + append, i.e. the window becomes
 *       that frame frames_per_state times (frame_buffer.py:35-36); no transition is recorded (actions/rewards/terminals[e]
 *       are ignored).
 *     otherwise: the frame is env.step(action[e])'s observation: it shifts environment e's window and the transition
 *       (previous window, action, reward, new window, terminal) is recorded exactly as ReplayMemory.add would, in
 *       environment order (logical indices and eviction follow deque(maxlen) over that order).
 *     actions / rewards / terminals / episode_start are HOST arrays of E entries.
 *   dqn_replay_current_states: the E current windows as stacks [E][frame elems][frames_per_state] (what
 *     FrameBuffer.get_state() returns per environment), e.g. as the input of dqn_net_forward for batched acting. */
int dqn_replay_set_lanes(dqn_replay_t h, int32_t num_envs);
int dqn_replay_add_batch(dqn_replay_t h, int32_t num_envs, const void* d_frames, const int32_t* actions,
                         const double* rewards, const uint8_t* terminals, const uint8_t* episode_start,
                         dqn_stream_t stream);
int dqn_replay_current_states(dqn_replay_t h, void* d_states, dqn_stream_t stream);
/* ReplayMemory.size / capacity (replay_memory.py:25-29); host-side counters, no sync. */
int dqn_replay_size(dqn_replay_t h, int64_t* size, int64_t* capacity);
/* Frame-ring position that the next appended frame will get. */
int dqn_replay_next_frame_pos(dqn_replay_t h, int64_t* pos);

/* The gather behind ReplayMemory.get_minibatch + get_states/get_next_states + the feed conversion
 * (replay_memory.py:31-53, dqn_agent.py:186-189, 249-253): for each logical index (0 = oldest,
 * the values random.sample(range(size), k) yields) rebuilds state and next_state stacks
 * [k][frame elems][frames_per_state] (HWC, channel = time, oldest -> newest, frame_buffer.py:51)
 * and copies action / reward / terminal.  Any of the five outputs may be NULL.  Bit-exact. */
int dqn_replay_gather(dqn_replay_t h, const int32_t* d_logical_idx, int32_t k, void* d_states,
                      void* d_next_states, int32_t* d_actions, double* d_rewards, uint8_t* d_terminals,
                      dqn_stream_t stream);

/* Pre-drawn sample indices resident on the device: a host that draws the index sets of the next `rows` updates at
 * once (random.sample per update, dqn_agent.py:110, uploaded as one [rows][k] int32 table) replays a captured update
 * without touching the indices again.  ONE kernel: copies row (*d_cursor % rows) to d_idx and advances *d_cursor
 * (int64, caller-owned device memory, start it at 0).  Capturable; the kernel that follows it on the stream may be
 * launched with programmatic dependent launch. */
int dqn_index_table_next(const int32_t* d_table, int64_t rows, int32_t k, int64_t* d_cursor, int32_t* d_idx,
                         dqn_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Q-network -- replaces Network.create_network and the graph it builds (network.py:13-116), the
 * session.run calls of DQNAgent (dqn_agent.py:132-138, 189, 222) and the host TD-target loop
 * (dqn_agent.py:118-125, 224-253).
 *
 * Parameters live in ONE flat fp32 buffer per set (online, target, gradient, two optimizer slots),
 * tensors in the reference's tuple order (network.py:296, 387) and TF layouts (HWIO conv filters,
 * [in,out] matrices), each tensor starting at a multiple of 4 floats.
 * ------------------------------------------------------------------------------------------------ */

/* Network.create_network (network.py:35-57).  in_shape: {dim} for SIMPLE, {84,84,4} for CNN/NATURE.
 * in_dtype: element type of the states the net is fed (DQN_U8 frames, DQN_F32, or DQN_F64 vectors).
 * Weights start at zero: the shell draws truncated_normal(0.01) (network.py:269, 341) and uploads. */
int dqn_net_create(dqn_net_t* out, int32_t arch, const int32_t* in_shape, int32_t in_rank,
                   int32_t num_actions, int32_t max_batch, int32_t in_dtype, int32_t device);
int dqn_net_destroy(dqn_net_t h);
int dqn_net_param_count(dqn_net_t h, int64_t* padded_floats, int32_t* num_tensors);
/* Tensor i of the reference's params tuple: offset (floats) into a flat buffer, element count, shape. */
int dqn_net_tensor_info(dqn_net_t h, int32_t i, int64_t* offset, int64_t* numel, int32_t* rank, int64_t shape[4]);
/* Device pointer of a flat buffer (enum dqn_buffer). */
int dqn_net_buffer(dqn_net_t h, int32_t which, float** d_ptr, int64_t* floats);
/* Synchronous host<->device copies of a whole flat buffer (init / checkpoint / tests). */
int dqn_net_upload(dqn_net_t h, int32_t which, const float* host, int64_t floats);
int dqn_net_download(dqn_net_t h, int32_t which, float* host, int64_t floats);
/* (Re)initialises optimizer slots the way TF does for h->optimizer (RMSProp ms=1, Adagrad/Ftrl 0.1)
 * and resets the update counter (global_step, network.py:83). */
int dqn_net_set_hyper(dqn_net_t h, const dqn_hyper_t* hyper, int32_t reset_slots);
int dqn_net_global_step(dqn_net_t h, int64_t* step, dqn_stream_t stream); /* synchronises the stream */
/* Restores the update counter when resuming from a checkpoint (the Supervisor restored global_step, main.py:136-143).
 * On a connected data-parallel net it also realigns this rank's exchange flags to `step`: every rank must call it at
 * the same point with no dqn_net_dp_step in flight anywhere (barrier before and after). */
int dqn_net_set_global_step(dqn_net_t h, int64_t step);
/* Arithmetic of the conv / fully-connected GEMMs (north_star's two modes; no reference counterpart, TF ran fp32):
 * 0 = fp32 on CUDA cores, the parity mode (<= 1e-4 vs the oracle);
 * 1 = bf16 operands on the tcgen05 tensor cores with fp32 accumulation, fp32 master weights, fp32 optimizer
 *     (<= 2e-2 vs the oracle).  DQN_ARCH_CNN and DQN_ARCH_NATURE with uint8 input.  Allocates its bf16 buffers on
 *     first use. */
int dqn_net_set_precision(dqn_net_t h, int32_t precision);

/* DQNAgent._predict_q_values (dqn_agent.py:176-189): forward of n states through the online
 * (which = DQN_BUF_ONLINE) or target (DQN_BUF_TARGET) parameters -> d_q [n][A]; if d_argmax is not
 * NULL also the greedy action per row (first maximum, dqn_agent.py:159-160). n <= max_batch. */
int dqn_net_forward(dqn_net_t h, int32_t which, const void* d_x, int32_t n, float* d_q, int32_t* d_argmax,
                    dqn_stream_t stream);

/* One minibatch update (DQNAgent._train_minibatch, dqn_agent.py:108-143, minus sampling and the
 * target-sync cadence): target forward on next_states, y = r + gamma*max_a Q_target (0 bootstrap for
 * terminals, float64 then rounded to float32), online forward, loss (network.py:149-152), backward,
 * optimizer update (network.py:198-202), global_step += 1.  d_loss (device float) and d_q
 * (device [k][A], the online Q-values before the update) may be NULL. */
int dqn_net_train_step(dqn_net_t h, const void* d_states, const void* d_next_states, const int32_t* d_actions,
                       const double* d_rewards, const uint8_t* d_terminals, int32_t k, float* d_loss, float* d_q,
                       dqn_stream_t stream);
/* The same step split for data parallelism: backward leaves the LOCAL mean gradient of the data
 * term (no L2 term) in the DQN_BUF_GRAD buffer (include_reg != 0 adds reg_param*w, for parity
 * checks); the caller all-reduces that buffer; apply runs the optimizer on grad_scale * grad
 * (+ reg_param*w) and bumps global_step. */
int dqn_net_backward(dqn_net_t h, const void* d_states, const void* d_next_states, const int32_t* d_actions,
                     const double* d_rewards, const uint8_t* d_terminals, int32_t k, float* d_loss, float* d_q,
                     int32_t include_reg, dqn_stream_t stream);
int dqn_net_apply(dqn_net_t h, dqn_stream_t stream);
/* Data-parallel update with the gradient exchange fused into the optimizer kernel (replaces the reference's
 * parameter-server round trip, main.py:117-120, with the mean semantics of SyncReplicasOptimizer, network.py:187-192):
 * every rank publishes its local mean gradient in a double-buffered exchange block that all ranks of the NVLink
 * domain have mapped (CUDA IPC), raises a per-rank flag with a system-scope release store, and the optimizer kernel
 * of every rank waits for all flags and sums the N gradients with peer loads in rank order (bit-identical weights on
 * all ranks) -- no separate all-reduce launch.  Set hyper.grad_scale = 1/world.  2 <= world <= 8, one process per GPU.
 *   dqn_net_dp_handle : 64-byte IPC handle of this rank's exchange block (all-gather it across ranks)
 *   dqn_net_dp_connect: maps the peers' blocks; handles = world x 64 bytes in rank order
 *   dqn_net_dp_step   : dqn_net_train_step across ranks; every rank must call it the same number of times
 * Skew between ranks: a rank's kernels wait for the peers' flags for at most the time set with
 * dqn_net_dp_set_timeout (default 120 s, or the environment variable DQN_DP_TIMEOUT_S read by dqn_net_dp_handle /
 * dqn_net_dp_connect).  When the wait expires nothing traps: the step number is recorded, the update goes on with
 * whatever the exchange buffers hold, and dqn_net_dp_status returns that step (0 = every wait so far was answered) so
 * that the host can stop or fall back to dqn_net_backward + all-reduce + dqn_net_apply. */
int dqn_net_dp_handle(dqn_net_t h, void* out_handle64);
int dqn_net_dp_connect(dqn_net_t h, int32_t rank, int32_t world, const void* handles);
int dqn_net_dp_set_timeout(dqn_net_t h, double seconds);
int dqn_net_dp_status(dqn_net_t h, int64_t* timed_out_step, dqn_stream_t stream); /* synchronises the stream */
int dqn_net_dp_step(dqn_net_t h, const void* d_states, const void* d_next_states, const int32_t* d_actions,
                    const double* d_rewards, const uint8_t* d_terminals, int32_t k, float* d_loss, dqn_stream_t stream);
/* dqn_train_step (gather + update) across ranks. */
int dqn_train_step_dp(dqn_net_t net, dqn_replay_t replay, const int32_t* d_logical_idx, int32_t k, float* d_loss,
                      dqn_stream_t stream);

/* DQNAgent._update_target_network's assigns (dqn_agent.py:215-222, network.py:251-252): one D2D copy. */
int dqn_net_target_sync(dqn_net_t h, dqn_stream_t stream);
/* bf16 mode keeps pre-swizzled bf16 operand copies of the weights beside the fp32 master buffers; dqn_net_upload, the
 * optimizer and dqn_net_target_sync keep them current.  Anything else that writes a DQN_BUF_ONLINE / DQN_BUF_TARGET
 * buffer in place through dqn_net_buffer -- the start-up broadcast of rank 0's parameters (the reference's chief
 * initialises and the others wait, main.py:133-143) -- must call this afterwards (which = the buffer written).
 * A no-op in fp32 mode. */
int dqn_net_refresh_operands(dqn_net_t h, int32_t which, dqn_stream_t stream);

/* Sample-to-update in one call: gather from the ring (d_logical_idx as in dqn_replay_gather) into the
 * net's own staging buffers, then dqn_net_train_step.  This is the unit bench.py times. */
int dqn_train_step(dqn_net_t net, dqn_replay_t replay, const int32_t* d_logical_idx, int32_t k, float* d_loss,
                   dqn_stream_t stream);

/* dqn_train_step / dqn_train_step_dp (data_parallel != 0) on row (*d_cursor % rows) of a device-resident [rows][k] table
 * of pre-drawn index sets (see dqn_index_table_next); the update advances *d_cursor itself, so a captured update replays
 * with no per-update work on the indices at all.  The conv1 kernels of the reference architecture in 16-bit mode read
 * the row directly; every other configuration first copies it to d_idx_scratch [k] with one small kernel. */
int dqn_train_step_table(dqn_net_t net, dqn_replay_t replay, const int32_t* d_table, int64_t rows, int32_t k,
                         int64_t* d_cursor, int32_t* d_idx_scratch, float* d_loss, int32_t data_parallel,
                         dqn_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Frame preprocessing on the device -- utils.resize_image (utils.py:39-45: cv2.cvtColor(RGB2GRAY) then
 * cv2.resize, bilinear), bit-exact with OpenCV's fixed-point uint8 paths.  d_rgb: n frames uint8
 * [n][src_h][src_w][3]; d_out: uint8 [n][dst_h][dst_w], ready for dqn_replay_append_frames_device.
 * ------------------------------------------------------------------------------------------------ */
int dqn_preproc_create(dqn_preproc_t* out, int32_t src_h, int32_t src_w, int32_t dst_h, int32_t dst_w, int32_t device);
int dqn_preproc_destroy(dqn_preproc_t h);
int dqn_preproc_run(dqn_preproc_t h, const uint8_t* d_rgb, int32_t n, uint8_t* d_out, dqn_stream_t stream);

/* Forward pass for callers that may not allocate device memory.  No counterpart in the reference.
 *
 * Same result as dqn_net_forward on a DQN_ARCH_CNN net fed uint8 states, but it needs no handle and
 * calls no cudaMalloc/cudaFree (nor any other allocating routine): the activations that flow
 * between its kernels live in a fixed array compiled into the library.  Meant for processes that
 * were given a fixed slice of the GPU and must not touch the allocator -- embedded actors, control
 * loops with hard latency bounds, sandboxed plugins.
 *
 *   d_params   the ten parameter tensors back to back, reference order (w1,b1,...,w5,b5; network.py:387),
 *              TF layouts, fp32, 16-byte aligned: 143,776 + 257*num_actions floats
 *   d_states   n stacks, uint8 [n][84][84][4]
 *   d_q        out, [n][num_actions] Q-values (may be NULL if d_argmax is given)
 *   d_argmax   out, [n] greedy actions, first maximum (may be NULL)
 *
 * Limits and rules
 *   - at most 64 states per call; a larger n is refused with DQN_ERR_INVALID and a message saying so.
 *   - the array occupies about 1.27 MB on every GPU that loads the library, until the process exits.
 *   - the first call on each device looks the array's address up (cudaGetSymbolAddress on the current
 *     device); make that first call outside of any stream capture, and from one host thread at a time
 *     (two first calls on the same device must not race).
 *   - every call on a device uses the same array, and the array keeps whatever the previous call left
 *     in it (every element a call reads is written earlier in the same call, so nothing needs
 *     clearing).  Two calls whose kernels can overlap on the GPU -- different streams without an
 *     ordering between them, or different host threads -- race on it; order them yourself. */
int dqn_cnn_forward_noalloc(const float* d_params, int32_t num_actions, const uint8_t* d_states, int32_t n,
                            float* d_q, int32_t* d_argmax, dqn_stream_t stream);

/* ------------------------------------------------------------------------------------------------
 * Instrumentation (no reference counterpart: the reference's only throughput signal is the
 * Supervisor's global_step/sec log line, main.py:136-143).
 * ------------------------------------------------------------------------------------------------ */
/* Kernels launched by this library in this process so far (eager launches; a replayed CUDA graph
 * launches the kernels that were counted when it was captured). */
int64_t dqn_launch_count(void);
/* enable=1: every kernel of the following eager steps is bracketed by a CUDA event pair on the
 * caller's stream; enable=2 additionally zeroes the accumulated totals; 0 switches it off.
 * If a step is CAPTURED into a CUDA graph while timing is on, the brackets become event-record nodes of that graph
 * and every replay stamps the kernels as they run inside the graph; enable=3 (set after the capture, before the
 * replays) keeps the set of bracketed stages from the capture so that dqn_net_timing_collect can be called after
 * every replay. */
int dqn_net_set_timing(dqn_net_t h, int32_t enable);
/* Synchronises the device and adds the last step's per-kernel durations to the totals. */
int dqn_net_timing_collect(dqn_net_t h);
int dqn_net_timing_read(dqn_net_t h, int32_t stage, double* total_ms, int32_t* count);
/* The same event bracket around an empty kernel (microseconds, mean of 20): the floor to subtract from the above. */
int dqn_net_timing_floor(dqn_net_t h, dqn_stream_t stream, double* us);
/* The same bracket as a pseudo-stage "launch_floor" of the step being timed (eager or captured): enqueue it after the
 * step and read it back with dqn_net_timing_read like any other stage. */
int dqn_net_timing_floor_enqueue(dqn_net_t h, dqn_stream_t stream);
const char* dqn_stage_name(int32_t stage); /* "" beyond the last stage */

#ifdef __cplusplus
}
#endif
#endif /* DQN_B200_H_ */
