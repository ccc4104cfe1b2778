// Frame preprocessing on the device: RGB -> gray -> bilinear resize, bit-exact with what the reference's
// utils.resize_image does through OpenCV (reference src/utils.py:39-45: cv2.cvtColor(RGB2GRAY) + cv2.resize,
// INTER_LINEAR).  OpenCV's uint8 paths are fixed point; the same integer arithmetic is used here
// (see oracle/preprocess_ref.py for the restatement and its pin against cv2):
//   gray = (R*9798 + G*19235 + B*3735 + 2^14) >> 15
//   11-bit tap weights, horizontal pass in int32, dst = (((b0*(S0>>4))>>16) + ((b1*(S1>>4))>>16) + 2) >> 2
// One thread per output pixel; the frames it produces go straight into the replay ring
// (dqn_replay_append_frames_device), so raw frames never visit the host.
#include "common.cuh"

#include <cmath>
#include <new>
#include <vector>

using namespace dqn;

namespace {
struct Tap { int idx; short w0, w1; };

__device__ __forceinline__ int gray_at(const uint8_t* __restrict__ img, int W, int y, int x) {
  const uint8_t* p = img + ((size_t)y * W + x) * 3;
  return (p[0] * 9798 + p[1] * 19235 + p[2] * 3735 + 16384) >> 15;
}

__global__ void __launch_bounds__(256)
preprocess_kernel(const uint8_t* __restrict__ rgb, int n, int H, int W, int OH, int OW, const Tap* __restrict__ xt,
                  const Tap* __restrict__ yt, uint8_t* __restrict__ out) {
  const long long total = (long long)n * OH * OW;
  for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    const int ox = (int)(t % OW);
    const int oy = (int)((t / OW) % OH);
    const int f = (int)(t / ((long long)OW * OH));
    const uint8_t* img = rgb + (size_t)f * H * W * 3;
    const Tap tx = xt[ox], ty = yt[oy];
    const int x0 = tx.idx, x1 = min(tx.idx + 1, W - 1);
    const int y0 = ty.idx, y1 = min(ty.idx + 1, H - 1);
    const int s0 = gray_at(img, W, y0, x0) * tx.w0 + gray_at(img, W, y0, x1) * tx.w1;
    const int s1 = gray_at(img, W, y1, x0) * tx.w0 + gray_at(img, W, y1, x1) * tx.w1;
    int v = (((ty.w0 * (s0 >> 4)) >> 16) + ((ty.w1 * (s1 >> 4)) >> 16) + 2) >> 2;
    v = v < 0 ? 0 : (v > 255 ? 255 : v);
    out[t] = (uint8_t)v;
  }
}

void make_taps(int dn, int sn, std::vector<Tap>& taps) {
  taps.resize(dn);
  const double scale = (double)sn / (double)dn;
  for (int d = 0; d < dn; ++d) {
    double f = (d + 0.5) * scale - 0.5;
    int s = (int)std::floor(f);
    f -= s;
    if (s < 0) { s = 0; f = 0.0; }
    if (s >= sn - 1) { s = sn - 1; f = 0.0; }
    const float ff = (float)f;
    taps[d].idx = s;
    taps[d].w0 = (short)lrintf((1.f - ff) * 2048.f);      // cvRound: round half to even
    taps[d].w1 = (short)lrintf(ff * 2048.f);
  }
}
}  // namespace

struct dqn_preproc_s {
  int device, H, W, OH, OW, num_sms;
  Tap *d_xt, *d_yt;
};

extern "C" int dqn_preproc_create(dqn_preproc_t* out, int32_t src_h, int32_t src_w, int32_t dst_h, int32_t dst_w,
                                  int32_t device) {
  DQN_REQUIRE(out, DQN_ERR_INVALID, "dqn_preproc_create: out is NULL");
  *out = nullptr;
  DQN_REQUIRE(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, DQN_ERR_INVALID, "image sizes must be positive");
  int ndev = 0;
  DQN_CUDA(cudaGetDeviceCount(&ndev));
  DQN_REQUIRE(device >= 0 && device < ndev, DQN_ERR_INVALID, "no CUDA device %d (found %d); there is no CPU path", device, ndev);
  DeviceGuard g(device);
  dqn_preproc_s* h = new (std::nothrow) dqn_preproc_s();
  DQN_REQUIRE(h, DQN_ERR_INVALID, "out of host memory");
  h->device = device; h->H = src_h; h->W = src_w; h->OH = dst_h; h->OW = dst_w;
  h->d_xt = h->d_yt = nullptr; h->num_sms = 148;
  cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, device);
  std::vector<Tap> xt, yt;
  make_taps(dst_w, src_w, xt);
  make_taps(dst_h, src_h, yt);
  cudaError_t e = cudaMalloc(&h->d_xt, xt.size() * sizeof(Tap));
  if (e == cudaSuccess) e = cudaMalloc(&h->d_yt, yt.size() * sizeof(Tap));
  if (e == cudaSuccess) e = cudaMemcpy(h->d_xt, xt.data(), xt.size() * sizeof(Tap), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(h->d_yt, yt.data(), yt.size() * sizeof(Tap), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    int rc = set_err(DQN_ERR_CUDA, "dqn_preproc_create: %s", cudaGetErrorString(e));
    dqn_preproc_destroy(h);
    return rc;
  }
  *out = h;
  return 0;
}

extern "C" int dqn_preproc_destroy(dqn_preproc_t h) {
  if (!h) return 0;
  DeviceGuard g(h->device);
  cudaFree(h->d_xt);
  cudaFree(h->d_yt);
  delete h;
  return 0;
}

extern "C" int dqn_preproc_run(dqn_preproc_t h, const uint8_t* d_rgb, int32_t n, uint8_t* d_out, dqn_stream_t stream) {
  DQN_REQUIRE(h && d_rgb && d_out, DQN_ERR_INVALID, "dqn_preproc_run: NULL argument");
  DQN_REQUIRE(n >= 1, DQN_ERR_INVALID, "dqn_preproc_run: n must be positive");
  DeviceGuard g(h->device);
  const long long total = (long long)n * h->OH * h->OW;
  long long blocks = (total + 255) / 256;
  const long long maxb = (long long)h->num_sms * 16;
  if (blocks > maxb) blocks = maxb;
  count_launch();
  preprocess_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(d_rgb, n, h->H, h->W, h->OH, h->OW,
                                                                                     h->d_xt, h->d_yt, d_out);
  DQN_CUDA(cudaGetLastError());
  return 0;
}
