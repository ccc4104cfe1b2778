// How long does a gpu-scope release (red.release / __threadfence) take after a CTA has written its outputs?
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long gt() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__global__ void k(float4* out, unsigned int* flag, unsigned long long* stamps, int nstore, int mode) {
  extern __shared__ unsigned char sm[];
  const int tid = threadIdx.x;
  unsigned long long t0 = gt();
  for (int i = 0; i < nstore; ++i) out[(size_t)blockIdx.x * 65536 + i * blockDim.x + tid] = make_float4(tid, i, 0, 0);
  __syncthreads();
  unsigned long long t1 = gt();
  if (tid == 0) {
    if (mode == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(flag) : "memory");
    else if (mode == 1) { __threadfence(); atomicAdd(flag, 1u); }
    else if (mode == 2) asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(flag) : "memory");
    else if (mode == 3) { asm volatile("fence.acq_rel.gpu;" ::: "memory"); asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(flag) : "memory"); }
    unsigned long long t2 = gt();
    stamps[blockIdx.x * 3 + 0] = t0; stamps[blockIdx.x * 3 + 1] = t1; stamps[blockIdx.x * 3 + 2] = t2;
  }
}
int main() {
  float4* out; unsigned int* flag; unsigned long long* st;
  cudaMalloc(&out, (size_t)148 * 65536 * 16); cudaMalloc(&flag, 64); cudaMalloc(&st, 148 * 3 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  unsigned long long h[148 * 3];
  for (int smem : {0, 200 * 1024}) for (int nstore : {0, 1, 4, 16}) for (int mode = 0; mode < 4; ++mode) {
    double worst = 0, mean = 0;
    for (int rep = 0; rep < 5; ++rep) {
      k<<<64, 512, smem>>>(out, flag, st, nstore, mode);
      cudaDeviceSynchronize();
      cudaMemcpy(h, st, sizeof(h), cudaMemcpyDeviceToHost);
      double w = 0, m = 0;
      for (int b = 0; b < 64; ++b) { double d = (double)(h[b * 3 + 2] - h[b * 3 + 1]); w = d > w ? d : w; m += d / 64; }
      if (rep) { worst += w / 4; mean += m / 4; }
    }
    printf("smem %3d KB, %2d x 8 KB stores per CTA, mode %d (%s): release takes mean %.0f ns, worst CTA %.0f ns\n", smem / 1024, nstore, mode,
           mode == 0 ? "red.release" : mode == 1 ? "__threadfence+atomicAdd" : mode == 2 ? "red.relaxed (no ordering)" : "fence.acq_rel+red.relaxed", mean, worst);
  }
  printf("err %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
