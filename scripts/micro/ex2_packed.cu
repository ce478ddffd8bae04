// Throughput of the packed half-precision exponentials against the fp32 one (softmax of the flash attention):
//   ex2.approx.ftz.f32 (1 result per lane and instruction), ex2.approx.f16x2 and ex2.approx.ftz.bf16x2 (2 results).
// If the packed forms issue at the fp32 rate they halve the MUFU time per score.  Also: cvt.rn.f16x2.f32 (pack) rate.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 ex2_packed.cu -o ex2_packed.bin
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t ex2h2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t ex2b2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.ftz.bf16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ uint32_t pkh(float a, float b) { uint32_t r; asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a)); return r; }
template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
  float v[16];
  uint32_t u[16];
  for (int i = 0; i < 16; ++i) { v[i] = -(threadIdx.x & 7) * 0.25f - i * 0.1f; u[i] = 0xB800B400u + i; }
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0) v[i] = ex2f(v[i]);
      if (MODE == 1) u[i] = ex2h2(u[i]);
      if (MODE == 2) u[i] = ex2b2(u[i]);
      if (MODE == 3) u[i] = pkh(v[i], __uint_as_float(u[i]));
      if (MODE == 4) { u[i] = pkh(v[i], v[(i + 1) & 15]); u[i] = ex2h2(u[i]); v[i] = __uint_as_float(u[i]); }
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 16; ++i) s += v[i] + __uint_as_float(u[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE>
void run(const char* name, int threads, int per_instr) {
  float* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  k<MODE><<<148, threads>>>(out, iters, cyc);
  k<MODE><<<148, threads>>>(out, iters, cyc);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  double instr = (double)iters * 16 * threads;
  printf("%-22s warps=%2d : %9lld clk  %6.2f lane-instr/clk/SM = %6.2f results/clk/SM\n", name, threads / 32, h, instr / h, instr * per_instr / h);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int th : {256, 512}) {
    run<0>("ex2.f32", th, 1); run<1>("ex2.f16x2", th, 2); run<2>("ex2.ftz.bf16x2", th, 2); run<3>("cvt.rn.f16x2.f32", th, 2);
    run<4>("cvt.f16x2 + ex2.f16x2", th, 2);
  }
  // accuracy of ex2.f16x2 on arguments in [-16, 0]
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
