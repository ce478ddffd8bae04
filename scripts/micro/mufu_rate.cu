// Throughput of the MUFU operations on one SM (results per clk): ex2, rcp, rsqrt, lg2, tanh -- 16 resident warps, 8 independent
// chains per thread.  nvcc -arch=sm_100a -O3 -o mufu_rate mufu_rate.cu && ./mufu_rate
#include <cstdio>
#include <cuda_runtime.h>

template <int OP>
__device__ __forceinline__ float mufu(float x) {
  float y;
  if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  if (OP == 1) asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  if (OP == 2) asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  if (OP == 3) asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  if (OP == 4) asm volatile("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int OP>
__global__ void k(float* out, long long* clk, int iters) {
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = 1.0f + 0.001f * (threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = mufu<OP>(v[i]) + 1.5f;  // the add keeps rcp(rcp(x)) from being folded away
  }
  __syncthreads();
  const long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name) {
  const int threads = 512, iters = 2000;
  float* out;
  long long* clk;
  cudaMalloc(&out, 148 * threads * sizeof(float));
  cudaMalloc(&clk, 148 * sizeof(long long));
  k<OP><<<148, threads>>>(out, clk, iters);
  k<OP><<<148, threads>>>(out, clk, iters);
  long long h[148];
  cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-6s %6.2f results / clk / SM  (%lld clk for %d x 8 x %d)\n", name, (double)threads * 8 * iters / h[0], h[0], iters, threads);
  cudaFree(out);
  cudaFree(clk);
}

int main() {
  run<0>("ex2");
  run<1>("rcp");
  run<2>("rsqrt");
  run<3>("lg2");
  run<4>("tanh");
  return 0;
}
