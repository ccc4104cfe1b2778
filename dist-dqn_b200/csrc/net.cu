// TEMPORARY stub so the ring can be exercised on the GPU first; replaced by the real kernels.
#include "common.cuh"
using namespace dqn;
#define STUB(name, ...) extern "C" int name(__VA_ARGS__) { return set_err(DQN_ERR_UNSUPPORTED, #name ": not built yet"); }
STUB(dqn_net_create, dqn_net_t*, int32_t, const int32_t*, int32_t, int32_t, int32_t, int32_t, int32_t)
STUB(dqn_net_destroy, dqn_net_t)
STUB(dqn_net_param_count, dqn_net_t, int64_t*, int32_t*)
STUB(dqn_net_tensor_info, dqn_net_t, int32_t, int64_t*, int64_t*, int32_t*, int64_t*)
STUB(dqn_net_buffer, dqn_net_t, int32_t, float**, int64_t*)
STUB(dqn_net_upload, dqn_net_t, int32_t, const float*, int64_t)
STUB(dqn_net_download, dqn_net_t, int32_t, float*, int64_t)
STUB(dqn_net_set_hyper, dqn_net_t, const dqn_hyper_t*, int32_t)
STUB(dqn_net_global_step, dqn_net_t, int64_t*, dqn_stream_t)
STUB(dqn_net_forward, dqn_net_t, int32_t, const void*, int32_t, float*, int32_t*, dqn_stream_t)
STUB(dqn_net_train_step, dqn_net_t, const void*, const void*, const int32_t*, const double*, const uint8_t*, int32_t, float*, float*, dqn_stream_t)
STUB(dqn_net_backward, dqn_net_t, const void*, const void*, const int32_t*, const double*, const uint8_t*, int32_t, float*, float*, int32_t, dqn_stream_t)
STUB(dqn_net_apply, dqn_net_t, dqn_stream_t)
STUB(dqn_net_target_sync, dqn_net_t, dqn_stream_t)
STUB(dqn_train_step, dqn_net_t, dqn_replay_t, const int32_t*, int32_t, float*, dqn_stream_t)
