// Which pipe does F2FP.BF16.F32.PACK_AB (cvt.rn.bf16x2.f32) use, and does it share it with MUFU.EX2?
// Per-SM throughput of (a) ex2 only, (b) cvt-pack only, (c) both interleaved.  nvcc -arch=sm_100a -O3 xu_pipe.cu -o xu_pipe.bin
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pk(float a, float b) { uint32_t r; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a)); return r; }
template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
  float v[16];
  uint32_t u[16];
  for (int i = 0; i < 16; ++i) { v[i] = threadIdx.x * 1e-3f + i; u[i] = i; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0 || MODE == 2) v[i] = ex2(v[i]);
      if (MODE == 1 || MODE == 2) u[i] = pk(v[i], __uint_as_float(u[i]));
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 16; ++i) s += v[i] + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE>
void run(const char* name, int threads) {
  float* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  k<MODE><<<148, threads>>>(out, iters, cyc);
  k<MODE><<<148, threads>>>(out, iters, cyc);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  double ops = (double)iters * 16 * threads;  // thread-level ops of EACH kind per SM
  printf("%-10s warps=%2d : %9lld clk  %6.2f thread-ops/clk/SM (per kind)\n", name, threads / 32, h, ops / h);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int th : {128, 256, 512}) { run<0>("ex2", th); run<1>("cvt.pack", th); run<2>("ex2+cvt", th); }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
