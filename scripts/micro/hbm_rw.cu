// HBM streaming rates for buffers far larger than L2: write-only, read-only, copy (1.5 GB each).
// nvcc -arch=sm_100a -O3 hbm_rw.cu -o hbm_rw.bin
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void k(const float4* __restrict__ a, float4* __restrict__ b, size_t n, float* sink) {
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (; i < n; i += stride) {
    if (MODE == 0) b[i] = make_float4(1.f, 2.f, 3.f, 4.f);
    if (MODE == 1) { float4 v = a[i]; acc += v.x + v.y + v.z + v.w; }
    if (MODE == 2) b[i] = a[i];
  }
  if (MODE == 1 && acc == 123.456f) *sink = acc;
}
template <int MODE>
void run(const char* name, const float4* a, float4* b, size_t n, float* sink, double bytes) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  k<MODE><<<148 * 16, 512>>>(a, b, n, sink);
  cudaEventRecord(e0);
  for (int i = 0; i < 5; ++i) k<MODE><<<148 * 16, 512>>>(a, b, n, sink);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("%-10s %8.1f us  %7.1f GB/s\n", name, ms / 5 * 1e3, bytes / (ms / 5 * 1e-3) / 1e9);
}
int main() {
  size_t n = (size_t)1536 << 20 >> 4;  // 1.5 GB of float4
  float4 *a, *b; float* sink;
  cudaMalloc(&a, n * 16); cudaMalloc(&b, n * 16); cudaMalloc(&sink, 4);
  cudaMemset(a, 0, n * 16);
  run<0>("write", a, b, n, sink, n * 16.0);
  run<1>("read", a, b, n, sink, n * 16.0);
  run<2>("copy", a, b, n, sink, n * 32.0);
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
}
