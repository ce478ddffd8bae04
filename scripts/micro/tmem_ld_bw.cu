// Microbenchmark: tcgen05.ld throughput (32x32b.x32) and MUFU.EX2 throughput per SM on B200.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_ld_bw scripts/micro/tmem_ld_bw.cu && /tmp/tmem_ld_bw
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../video_depth_anything_b200/csrc/ptx.cuh"

__global__ void tmem_ld_kernel(int iters, float* sink, long long* cycles) {
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::tc_fence_before(); __syncthreads(); ptx::tc_fence_after();
  const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
  float acc = 0.f;
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
    float v[32];
    ptx::tmem_ld32(base + ((i * 32) & 255) + (warp >> 2) * 32 % 256, v);
    ptx::tmem_ld_wait();
    acc += v[i & 31];
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  __syncthreads();
  if (warp == 0) ptx::tmem_dealloc(slot, 512);
}

__global__ void mufu_kernel(int iters, float* sink, long long* cycles) {
  float x[8];
  for (int k = 0; k < 8; ++k) x[k] = -0.001f * (threadIdx.x + k);
  __syncthreads();
  long long t0 = clock64();
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[k]));
  }
  long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  float s = 0; for (int k = 0; k < 8; ++k) s += x[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  float* sink; long long* cyc; cudaMalloc(&sink, 1 << 22); cudaMalloc(&cyc, 4096);
  long long h[1];
  for (int warps : {4, 8, 16}) {
    const int iters = 4096;
    tmem_ld_kernel<<<1, warps * 32>>>(iters, sink, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    double bytes = (double)warps * iters * 32 * 32 * 4;
    printf("tcgen05.ld 32x32b.x32: %2d warps: %lld cycles, %.1f B/clk/SM (%s)\n", warps, h[0], bytes / h[0], cudaGetErrorString(e));
  }
  for (int warps : {4, 8, 16, 32}) {
    const int iters = 2048;
    mufu_kernel<<<1, warps * 32>>>(iters, sink, cyc);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, cyc, 8, cudaMemcpyDeviceToHost);
    double n = (double)warps * 32 * iters * 8;
    printf("MUFU.EX2: %2d warps: %lld cycles, %.2f ex2/clk/SM (%s)\n", warps, h[0], n / h[0], cudaGetErrorString(e));
  }
  return 0;
}
