// Issue rate of fma.rn.f32 (3 register operands) vs fma.rn.f32x2 (sm_100 packed pair) per SM.
// nvcc -arch=sm_100a -O3 ffma2.cu -o ffma2.bin
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* out, int iters, long long* cyc) {
  float v[16], w[16];
  for (int i = 0; i < 16; ++i) { v[i] = threadIdx.x * 1e-3f + i; w[i] = 1.0f + i * 1e-3f; }
  const float c = out[0];
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(v[i]) : "f"(w[i]), "f"(c));
    } else {
#pragma unroll
      for (int i = 0; i < 16; i += 2) {
        asm volatile(
            "{ .reg .b64 a, b, cc;\n\t"
            "mov.b64 a, {%0, %1};\n\t"
            "mov.b64 b, {%2, %3};\n\t"
            "mov.b64 cc, {%4, %4};\n\t"
            "fma.rn.f32x2 a, b, cc, a;\n\t"
            "mov.b64 {%0, %1}, a; }"
            : "+f"(v[i]), "+f"(v[i + 1])
            : "f"(w[i]), "f"(w[i + 1]), "f"(c));
      }
    }
  }
  long long t1 = clock64();
  float s = 0;
  for (int i = 0; i < 16; ++i) s += v[i];
  out[1 + blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE>
void run(const char* name, int threads) {
  float* out; long long* cyc; long long h;
  cudaMalloc(&out, (148 * 1024 + 1) * 4); cudaMemset(out, 0, 4); cudaMalloc(&cyc, 8);
  const int iters = 4096;
  k<MODE><<<148, threads>>>(out, iters, cyc);
  k<MODE><<<148, threads>>>(out, iters, cyc);
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-12s warps=%2d : %9lld clk  %6.1f fp32 FMA/clk/SM\n", name, threads / 32, h, (double)iters * 16 * threads / h);
  cudaFree(out); cudaFree(cyc);
}
int main() {
  for (int th : {128, 256, 512, 1024}) { run<0>("fma.f32", th); run<1>("fma.f32x2", th); }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
