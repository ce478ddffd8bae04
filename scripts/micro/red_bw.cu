// Throughput of the residual-stream update forms on one [rows, 1024] fp32 matrix (180 MB, larger than L2):
// red.global.add.v4.f32 vs ld.v4 + add + st.v4 vs plain st.v4.  nvcc -arch=sm_100a -O3 red_bw.cu -o red_bw.bin
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(float* p, size_t n4) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n4; i += stride) {
    float4* q = reinterpret_cast<float4*>(p) + i;
    if (MODE == 0) {
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(q), "f"(1.f), "f"(2.f), "f"(3.f), "f"(4.f) : "memory");
    } else if (MODE == 1) {
      float4 v = *q;
      v.x += 1.f; v.y += 2.f; v.z += 3.f; v.w += 4.f;
      *q = v;
    } else if (MODE == 2) {
      *q = make_float4(1.f, 2.f, 3.f, 4.f);
    } else {
      asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p + 4 * i), "f"(1.f) : "memory");
      asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p + 4 * i + 1), "f"(1.f) : "memory");
      asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p + 4 * i + 2), "f"(1.f) : "memory");
      asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p + 4 * i + 3), "f"(1.f) : "memory");
    }
  }
}
template <int MODE>
void run(const char* name, float* p, size_t n4, int blocks, int threads) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  k<MODE><<<blocks, threads>>>(p, n4);
  cudaEventRecord(a);
  for (int i = 0; i < 5; ++i) k<MODE><<<blocks, threads>>>(p, n4);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  printf("%-28s blocks=%4d threads=%4d : %8.1f us  %7.1f GB/s (payload)\n", name, blocks, threads, ms / 5 * 1e3, n4 * 16.0 / (ms / 5 * 1e-3) / 1e9);
}
int main() {
  size_t n4 = (size_t)43840 * 1024 / 4;
  float* p; cudaMalloc(&p, n4 * 16); cudaMemset(p, 0, n4 * 16);
  for (int th : {256, 1024}) for (int bl : {148, 148 * 4, 148 * 16}) {
    run<0>("red.v4.f32", p, n4, bl, th);
    run<1>("ld.v4+st.v4", p, n4, bl, th);
    run<2>("st.v4", p, n4, bl, th);
    run<3>("red.f32 x4", p, n4, bl, th);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
