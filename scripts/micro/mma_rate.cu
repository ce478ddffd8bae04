// Issue / execution interval of tcgen05.mma.cta_group::1.kind::f16 (M = 128, K = 16) as a function of N: is there a
// per-instruction floor that makes narrow MMAs cost more than their FLOPs?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../video_depth_anything_b200/csrc mma_rate.cu -o mma_rate.bin
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "ptx.cuh"

template <int N, bool TF32>
__global__ void __launch_bounds__(128, 1) k(long long* out, int reps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < (128 * 128 + 256 * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (threadIdx.x < 32) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = slot;
  if (threadIdx.x == 0) {
    ptx::fence_proxy_async();
    const uint64_t da = ptx::make_sw128_desc(ptx::smem_u32(smem), 16, 1024);
    const uint64_t db = ptx::make_sw128_desc(ptx::smem_u32(smem + 128 * 128), 16, 1024);
    constexpr uint32_t idesc = TF32 ? ptx::make_idesc_tf32(128, N) : ptx::make_idesc_bf16(128, N, 0, 0);
    uint32_t ph = 0;
    for (int warm = 0; warm < 2; ++warm) {
      const long long t0 = clock64();
      for (int r = 0; r < reps; ++r) {
        if (TF32) ptx::mma_tf32_ss(tmem + (r & 1) * 256, da + 2 * (r & 3), db + 2 * (r & 3), idesc, 1);
        else ptx::mma_f16_ss(tmem + (r & 1) * 256, da + 2 * (r & 3), db + 2 * (r & 3), idesc, 1);
      }
      const long long t1 = clock64();
      ptx::tc_commit(&bar);
      ptx::mbar_wait(&bar, ph);
      ph ^= 1;
      const long long t2 = clock64();
      if (blockIdx.x == 0 && warm == 1) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tmem, 512); }
}

template <int N, bool TF32 = false>
void run(int reps) {
  long long *d, h[2];
  cudaMalloc(&d, 16);
  const int smem = 128 * 128 + 256 * 128 + 1024;
  cudaFuncSetAttribute(k<N, TF32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  k<N, TF32><<<148, 128, smem>>>(d, reps);
  cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("%s N=%3d: %d MMAs issued in %lld clk, complete after %lld clk -> %.1f clk per MMA (FLOP bound %d)  %s\n", TF32 ? "tf32 K=8 " : "bf16 K=16", N, reps, h[0], h[1],
         (double)h[1] / reps, N / 2, cudaGetErrorString(cudaGetLastError()));
  cudaFree(d);
}
int main() {
  run<32>(512); run<64>(512); run<128>(512); run<256>(512);
  run<32, true>(512); run<64, true>(512); run<128, true>(512);
  return 0;
}
