// fp32 implicit-GEMM building blocks shared by the conv / fully-connected forward, weight-gradient
// and data-gradient kernels (reference: the TF ops at src/network.py:303-308, 395-424 and their
// gradients, which TF dispatched to cuDNN/cuBLAS).
//
// All three are the same 64x32 register-tiled product  C[i,j] = sum_r A[r][i] * B[r][j]  with
// different gather functions filling the shared-memory tiles:
//   forward : i = output pixel (pool-window order), j = out channel, r = (ky,kx,c)
//   wgrad   : i = (ky,kx,c),                        j = out channel, r = output pixel
//   dgrad   : i = input pixel,                      j = in channel,  r = (ky,kx,n)
// Layer geometry is a compile-time struct so every index decomposition is a multiply-shift.
//
// Small batches (B=32) leave the deeper layers with a handful of output tiles and a long serial
// reduction, so forward and dgrad split the reduction over a THREAD-BLOCK CLUSTER: the S CTAs of a
// cluster each take 1/S of the chunks of one output tile, park their partial accumulators in their
// own shared memory, and rank 0 adds them up through distributed shared memory before the fused
// epilogue (deterministic order, no global scratch, no atomics).
#pragma once
#include <cstdint>
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include "common.cuh"

namespace dqn {

namespace cg = cooperative_groups;

constexpr int BM = 64;   // tile rows    (i)
constexpr int BN = 32;   // tile columns (j)
constexpr int BK = 32;   // reduction chunk (r)
constexpr int NT = 128;  // threads per CTA; each owns a 4x4 micro-tile
constexpr int MAX_CLUSTER = 8;
static_assert(BK * BM >= BM * BN, "the A tile is reused as the cluster-reduction buffer");

template <int IH_, int IW_, int CIN_, int KS_, int ST_, int PADH_, int PADW_, int OH_, int OW_, int COUT_, bool POOL_>
struct Geo {
  static constexpr int IH = IH_, IW = IW_, CIN = CIN_, KS = KS_, ST = ST_, PADH = PADH_, PADW = PADW_;
  static constexpr int OH = OH_, OW = OW_, COUT = COUT_;
  static constexpr bool POOL = POOL_;
  static constexpr int PH = POOL_ ? (OH_ + 1) / 2 : OH_;   // TF SAME 2x2/2 pool: ceil(OH/2), pad after only
  static constexpr int PW = POOL_ ? (OW_ + 1) / 2 : OW_;
  static constexpr int K = KS_ * KS_ * CIN_;
  // rows of the implicit GEMM per sample; with pooling the 4 conv outputs of a pool window are
  // 4 consecutive rows (some of them outside the image) so the pool is a register max in the epilogue
  static constexpr int ROWS = POOL_ ? PH * PW * 4 : OH_ * OW_;
  static constexpr int OUT_PER_SAMPLE = PH * PW * COUT_;
  static_assert(CIN_ % 4 == 0 && COUT_ % BN == 0 && K % BM == 0 && (KS_ * KS_ * COUT_) % BK == 0 && COUT_ % BK == 0,
                "geometry not tileable");
};

struct Row {
  int b, oy, ox;  // sample, conv-output coordinates
  int pidx;       // index of the (pooled) activation pixel this row contributes to
  int win;        // position inside the 2x2 pool window (dy*2+dx), 0 without pooling
  bool valid;
};

template <class G>
__device__ __forceinline__ Row decode_row(int m, int B) {
  Row r;
  if constexpr (G::POOL) {
    const int p = m >> 2;
    r.win = m & 3;
    r.b = p / (G::PH * G::PW);
    const int rem = p - r.b * (G::PH * G::PW);
    const int py = rem / G::PW, px = rem - py * G::PW;
    r.oy = py * 2 + (r.win >> 1);
    r.ox = px * 2 + (r.win & 1);
    r.pidx = p;
    r.valid = (r.b < B) && (r.oy < G::OH) && (r.ox < G::OW);
  } else {
    r.b = m / (G::OH * G::OW);
    const int rem = m - r.b * (G::OH * G::OW);
    r.oy = rem / G::OW;
    r.ox = rem - r.oy * G::OW;
    r.pidx = m;
    r.win = 0;
    r.valid = r.b < B;
  }
  return r;
}

// 4 consecutive input channels of the input pixel that filter tap `kquad` of output row `r` reads
// (zero outside the image: TF SAME padding, network.py:417).  uint8 inputs are converted to float
// here, unscaled, exactly like the reference's float32 placeholder feed (network.py:61).
template <class G, typename TIn>
__device__ __forceinline__ float4 load_x_quad(const TIn* __restrict__ x, const Row& r, int kquad) {
  constexpr int CQ = G::CIN / 4;
  const int c0 = (kquad % CQ) * 4;
  const int t = kquad / CQ;
  const int kx = t % G::KS, ky = t / G::KS;
  const int iy = r.oy * G::ST - G::PADH + ky;
  const int ix = r.ox * G::ST - G::PADW + kx;
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (r.valid && (unsigned)iy < (unsigned)G::IH && (unsigned)ix < (unsigned)G::IW) {
    const size_t off = ((size_t)(r.b * G::IH + iy) * G::IW + ix) * G::CIN + c0;
    if constexpr (sizeof(TIn) == 1) {
      const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(x + off));
      v = make_float4((float)(u & 255u), (float)((u >> 8) & 255u), (float)((u >> 16) & 255u), (float)(u >> 24));
    } else {
      v = __ldg(reinterpret_cast<const float4*>(x + off));
    }
  }
  return v;
}

// Gradient w.r.t. the conv output (pre-ReLU) for 4 consecutive channels, rebuilt on the fly from
// the gradient of the pooled activation: routed to the arg-max of the pool window (first maximum,
// TF MaxPoolGrad) and gated by ReLU (act > 0).  Nothing pre-pool is ever stored.
template <class G>
__device__ __forceinline__ float4 load_dz_quad(const float* __restrict__ dact, const float* __restrict__ act,
                                               const uint8_t* __restrict__ arg, int pidx, int win, int n0) {
  const size_t off = (size_t)pidx * G::COUT + n0;
  float4 d = *reinterpret_cast<const float4*>(dact + off);
  const float4 p = *reinterpret_cast<const float4*>(act + off);
  if constexpr (G::POOL) {
    const uchar4 a = *reinterpret_cast<const uchar4*>(arg + off);
    d.x = (a.x == win && p.x > 0.f) ? d.x : 0.f;
    d.y = (a.y == win && p.y > 0.f) ? d.y : 0.f;
    d.z = (a.z == win && p.z > 0.f) ? d.z : 0.f;
    d.w = (a.w == win && p.w > 0.f) ? d.w : 0.f;
  } else {
    d.x = p.x > 0.f ? d.x : 0.f;
    d.y = p.y > 0.f ? d.y : 0.f;
    d.z = p.z > 0.f ? d.z : 0.f;
    d.w = p.w > 0.f ? d.w : 0.f;
  }
  return d;
}

__device__ __forceinline__ void tile_fma(const float (*As)[BM], const float (*Bs)[BN], int ti, int tj,
                                         float acc[4][4]) {
#pragma unroll
  for (int kk = 0; kk < BK; ++kk) {
    const float4 a = *reinterpret_cast<const float4*>(&As[kk][ti * 4]);
    const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tj * 4]);
    const float av[4] = {a.x, a.y, a.z, a.w};
    const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
  }
}

// Split-K epilogue: ranks > 0 park their 4x4 accumulators in their shared memory, rank 0 adds them
// (in rank order) through DSMEM.  Returns true on the CTA that owns the final sum.
__device__ __forceinline__ bool cluster_reduce(float* red, float acc[4][4], int tid) {
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned S = cluster.num_blocks();
  if (S == 1) return true;
  const unsigned rank = cluster.block_rank();
  float4* mine = reinterpret_cast<float4*>(red) + tid * 4;
  if (rank != 0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) mine[i] = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
  cluster.sync();
  if (rank == 0) {
    for (unsigned r = 1; r < S; ++r) {
      const float4* rem = reinterpret_cast<const float4*>(cluster.map_shared_rank(red, r)) + tid * 4;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 v = rem[i];
        acc[i][0] += v.x; acc[i][1] += v.y; acc[i][2] += v.z; acc[i][3] += v.w;
      }
    }
  }
  cluster.sync();          // nobody leaves while its shared memory may still be read
  return rank == 0;
}

// ------------------------------------------------------------------------------------------------
// Forward: out = pool(relu(conv(x, w) + bias)); blockIdx.z selects the pass (online net on the
// states / target net on the next states) so both forwards of a training step share one launch.
// ------------------------------------------------------------------------------------------------
struct FwdPass {
  const void* x;
  const float* w;
  const float* bias;
  float* out;
  uint8_t* arg;
};
struct FwdArgs {
  FwdPass p[2];
  int B;
};

template <class G, typename TIn>
__global__ void __launch_bounds__(NT) conv_fwd_kernel(FwdArgs a) {
  __shared__ __align__(16) float As[BK][BM];
  __shared__ __align__(16) float Bs[BK][BN];
  const unsigned S = cg::this_cluster().num_blocks();          // cluster = (1,1,S): split of the reduction
  const unsigned rank = cg::this_cluster().block_rank();
  // selected field by field: a run-time index into the parameter block would move it to local memory
  const bool p1 = blockIdx.z / S != 0;
  FwdPass P;
  P.x = p1 ? a.p[1].x : a.p[0].x;
  P.w = p1 ? a.p[1].w : a.p[0].w;
  P.bias = p1 ? a.p[1].bias : a.p[0].bias;
  P.out = p1 ? a.p[1].out : a.p[0].out;
  P.arg = p1 ? a.p[1].arg : a.p[0].arg;
  const TIn* __restrict__ x = static_cast<const TIn*>(P.x);
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int am = tid % BM, akq = tid / BM;          // this thread gathers quads akq, akq+2, akq+4, akq+6 of row am
  const Row row = decode_row<G>(m0 + am, a.B);
  const int bkk = tid >> 3, bnq = tid & 7;           // weight rows bkk and bkk+16
  const int ti = tid >> 3, tj = tid & 7;
  pdl_wait();
  constexpr int T = G::K / BK;
  const int it0 = (int)(T * rank / S), it1 = (int)(T * (rank + 1) / S);
  float acc[4][4] = {};
  float4 ra[4], rb[2];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) ra[q] = load_x_quad<G, TIn>(x, row, (k0 >> 2) + akq + 2 * q);
#pragma unroll
    for (int q = 0; q < 2; ++q)
      rb[q] = __ldg(reinterpret_cast<const float4*>(P.w + (size_t)(k0 + bkk + 16 * q) * G::COUT + n0 + bnq * 4));
  };
  if (it0 < it1) fetch(it0 * BK);
  for (int it = it0; it < it1; ++it) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int kk = (akq + 2 * q) * 4;
      As[kk + 0][am] = ra[q].x; As[kk + 1][am] = ra[q].y; As[kk + 2][am] = ra[q].z; As[kk + 3][am] = ra[q].w;
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) *reinterpret_cast<float4*>(&Bs[bkk + 16 * q][bnq * 4]) = rb[q];
    __syncthreads();
    if (it + 1 < it1) fetch((it + 1) * BK);
    tile_fma(As, Bs, ti, tj, acc);
    __syncthreads();
  }
  pdl_trigger();                                     // main loop done: the next kernel may start its prologue
  if (!cluster_reduce(&As[0][0], acc, tid)) return;
  const int n = n0 + tj * 4;
  const float4 bias = __ldg(reinterpret_cast<const float4*>(P.bias + n));
  const float bv[4] = {bias.x, bias.y, bias.z, bias.w};
  if constexpr (G::POOL) {
    const int mrow = m0 + ti * 4;                   // the 4 rows of this thread are one pool window
    const Row r0 = decode_row<G>(mrow, a.B);
    if (r0.b < a.B) {
      float best[4];
      unsigned char bi[4] = {0, 0, 0, 0};
#pragma unroll
      for (int j = 0; j < 4; ++j) best[j] = fmaxf(acc[0][j] + bv[j], 0.f);   // window element 0 is always inside
#pragma unroll
      for (int i = 1; i < 4; ++i) {
        const bool inside = (r0.oy + (i >> 1) < G::OH) && (r0.ox + (i & 1) < G::OW);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float v = fmaxf(acc[i][j] + bv[j], 0.f);
          if (inside && v > best[j]) { best[j] = v; bi[j] = (unsigned char)i; }
        }
      }
      const size_t off = (size_t)r0.pidx * G::COUT + n;
      *reinterpret_cast<float4*>(P.out + off) = make_float4(best[0], best[1], best[2], best[3]);
      if (P.arg) *reinterpret_cast<uchar4*>(P.arg + off) = make_uchar4(bi[0], bi[1], bi[2], bi[3]);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + ti * 4 + i;
      if (m < a.B * G::ROWS) {
        float4 v;
        v.x = fmaxf(acc[i][0] + bv[0], 0.f); v.y = fmaxf(acc[i][1] + bv[1], 0.f);
        v.z = fmaxf(acc[i][2] + bv[2], 0.f); v.w = fmaxf(acc[i][3] + bv[3], 0.f);
        *reinterpret_cast<float4*>(P.out + (size_t)m * G::COUT + n) = v;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Weight gradient: partial[split][k][n] = sum over this split's output pixels of x_patch[k] * dz[n];
// row K of each partial is the bias gradient.  The partial sums live in scratch memory and are
// added up (in a fixed order) by the optimizer / reduce kernel.
// ------------------------------------------------------------------------------------------------
struct WgradArgs {
  const void* x;        // layer input  [B, IH, IW, CIN]
  const float* dact;    // gradient of the layer's (pooled) output
  const float* act;     // the layer's (pooled) output
  const uint8_t* arg;   // pool arg-max (POOL only)
  float* partial;       // [nsplit][(K+1)*COUT]
  int B;
  int rows_per_split;   // multiple of BK
};

template <class G, typename TIn>
__global__ void __launch_bounds__(NT) conv_wgrad_kernel(WgradArgs a) {
  __shared__ __align__(16) float As[BK][BM];
  __shared__ __align__(16) float Bs[BK][BN];
  const TIn* __restrict__ x = static_cast<const TIn*>(a.x);
  const int tid = threadIdx.x;
  const int i0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int R = a.B * G::ROWS;
  const int r_begin = blockIdx.z * a.rows_per_split;
  const int r_end = min(R, r_begin + a.rows_per_split);
  pdl_wait();
  const int aiq = tid & 15, ar = tid >> 4;          // gathers k-quad aiq of rows ar, ar+8, ar+16, ar+24
  const int kquad = (i0 >> 2) + aiq;
  const int br = tid >> 3, bnq = tid & 7;           // dz quad bnq of rows br and br+16
  const int ti = tid >> 3, tj = tid & 7;
  float acc[4][4] = {};
  float bsum = 0.f;                                  // bias gradient, threads 0..31 of the first row-tile
  float4 ra[4], rb[2];
  auto fetch = [&](int r0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const Row qa = decode_row<G>(r0 + ar + 8 * q, a.B);
      ra[q] = load_x_quad<G, TIn>(x, qa, kquad);
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const Row qb = decode_row<G>(r0 + br + 16 * q, a.B);
      rb[q] = qb.valid ? load_dz_quad<G>(a.dact, a.act, a.arg, qb.pidx, qb.win, n0 + bnq * 4)
                       : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  if (r_begin < r_end) fetch(r_begin);
  for (int r0 = r_begin; r0 < r_end; r0 += BK) {
#pragma unroll
    for (int q = 0; q < 4; ++q) *reinterpret_cast<float4*>(&As[ar + 8 * q][aiq * 4]) = ra[q];
#pragma unroll
    for (int q = 0; q < 2; ++q) *reinterpret_cast<float4*>(&Bs[br + 16 * q][bnq * 4]) = rb[q];
    __syncthreads();
    if (r0 + BK < r_end) fetch(r0 + BK);
    tile_fma(As, Bs, ti, tj, acc);
    if (blockIdx.x == 0 && tid < BN) {
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) bsum += Bs[kk][tid];
    }
    __syncthreads();
  }
  pdl_trigger();
  float* out = a.partial + (size_t)blockIdx.z * ((size_t)(G::K + 1) * G::COUT);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int k = i0 + ti * 4 + i;
    *reinterpret_cast<float4*>(out + (size_t)k * G::COUT + n0 + tj * 4) =
        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
  if (blockIdx.x == 0 && tid < BN) out[(size_t)G::K * G::COUT + n0 + tid] = bsum;
}

// ------------------------------------------------------------------------------------------------
// Data gradient: dx[b,iy,ix,c] = sum_{ky,kx,n} dz[b,oy,ox,n] * w[ky,kx,c,n], (oy,ox) = ((iy,ix)+pad-(ky,kx))/stride.
// dx is the gradient of the previous layer's pooled activation.
// ------------------------------------------------------------------------------------------------
struct DgradArgs {
  const float* w;       // [KS,KS,CIN,COUT]
  const float* dact;
  const float* act;
  const uint8_t* arg;
  float* dx;            // [B, IH, IW, CIN]
  int B;
};

template <class G>
__global__ void __launch_bounds__(NT) conv_dgrad_kernel(DgradArgs a) {
  __shared__ __align__(16) float As[BK][BM];
  __shared__ __align__(16) float Bs[BK][BN];
  constexpr int KD = G::KS * G::KS * G::COUT;       // reduction length
  constexpr int NQ = G::COUT / 4;
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, c0 = blockIdx.y * BN;
  const int M = a.B * G::IH * G::IW;
  const int am = tid % BM, akq = tid / BM;
  const int m = m0 + am;
  const int b = m / (G::IH * G::IW);
  const int rem = m - b * (G::IH * G::IW);
  const int iy = rem / G::IW, ix = rem - iy * G::IW;
  const bool mvalid = m < M;
  pdl_wait();
  const int brq = tid & 7, bc = tid >> 3;           // w quad (4 consecutive n) brq for input channels c0+bc, c0+bc+16
  const int ti = tid >> 3, tj = tid & 7;
  const unsigned S = cg::this_cluster().num_blocks();
  const unsigned rank = cg::this_cluster().block_rank();
  constexpr int T = KD / BK;
  const int it0 = (int)(T * rank / S), it1 = (int)(T * (rank + 1) / S);
  float acc[4][4] = {};
  float4 ra[4], rb[2];
  auto dz_at = [&](int kq) -> float4 {
    const int nq = kq % NQ;
    const int t = kq / NQ;
    const int kx = t % G::KS, ky = t / G::KS;
    const int ty = iy + G::PADH - ky, tx = ix + G::PADW - kx;
    if (!mvalid || ty < 0 || tx < 0 || (ty % G::ST) != 0 || (tx % G::ST) != 0) return make_float4(0.f, 0.f, 0.f, 0.f);
    const int oy = ty / G::ST, ox = tx / G::ST;
    if (oy >= G::OH || ox >= G::OW) return make_float4(0.f, 0.f, 0.f, 0.f);
    int pidx, win;
    if constexpr (G::POOL) {
      pidx = (b * G::PH + (oy >> 1)) * G::PW + (ox >> 1);
      win = (oy & 1) * 2 + (ox & 1);
    } else {
      pidx = (b * G::OH + oy) * G::OW + ox;
      win = 0;
    }
    return load_dz_quad<G>(a.dact, a.act, a.arg, pidx, win, nq * 4);
  };
  auto fetch = [&](int k0) {
#pragma unroll
    for (int q = 0; q < 4; ++q) ra[q] = dz_at((k0 >> 2) + akq + 2 * q);
    const int kq = (k0 >> 2) + brq;                  // all 8 quads of a chunk share (ky,kx): COUT % BK == 0
    const int nq = kq % NQ, t = kq / NQ;
#pragma unroll
    for (int q = 0; q < 2; ++q)
      rb[q] = __ldg(reinterpret_cast<const float4*>(a.w + ((size_t)t * G::CIN + c0 + bc + 16 * q) * G::COUT + nq * 4));
  };
  if (it0 < it1) fetch(it0 * BK);
  for (int it = it0; it < it1; ++it) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int kk = (akq + 2 * q) * 4;
      As[kk + 0][am] = ra[q].x; As[kk + 1][am] = ra[q].y; As[kk + 2][am] = ra[q].z; As[kk + 3][am] = ra[q].w;
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      Bs[brq * 4 + 0][bc + 16 * q] = rb[q].x; Bs[brq * 4 + 1][bc + 16 * q] = rb[q].y;
      Bs[brq * 4 + 2][bc + 16 * q] = rb[q].z; Bs[brq * 4 + 3][bc + 16 * q] = rb[q].w;
    }
    __syncthreads();
    if (it + 1 < it1) fetch((it + 1) * BK);
    tile_fma(As, Bs, ti, tj, acc);
    __syncthreads();
  }
  pdl_trigger();                                     // main loop done: the next kernel may start its prologue
  if (!cluster_reduce(&As[0][0], acc, tid)) return;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int mm = m0 + ti * 4 + i;
    if (mm < M)
      *reinterpret_cast<float4*>(a.dx + (size_t)mm * G::CIN + c0 + tj * 4) =
          make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
  }
}

}  // namespace dqn
