// How many bytes per clock can the SMs pull out of L2 with TMA, and does multicast across a cluster raise it?
// The encoder GEMMs (256x256 CTA-pair tiles) sit exactly at ~6100 B/clk of L2 -> SM traffic chip-wide (ncu:
// l1tex__m_xbar2l1tex_read_bytes / cycles): is that a limit of the L2 slices (multicast would lift it) or of what an SM
// can take in (only larger tiles would)?  Every CTA receives 64 KB per iteration into a double-buffered 128 KB of
// shared memory, from an L2-resident 48 MB matrix:
//   mode 0  unicast, every CTA its own tiles
//   mode 1  unicast, all CTAs of a cluster the SAME tiles (does L2 merge them?)
//   mode 2  multicast: each CTA of the cluster issues 1/CS of the tiles to all CS CTAs
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../video_depth_anything_b200/csrc tma_l2_bw.cu -o tma_l2_bw.bin -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include <cuda.h>
#include "ptx.cuh"

constexpr int TILE = 128 * 128;          // 128 rows x 64 bf16 = 16 KB
constexpr int TILES = 4;                 // per CTA and iteration
constexpr int GROUP = TILES * TILE;      // 64 KB
constexpr int ROWS = 48 * 1024 * 1024 / 128;  // 48 MB matrix of 64 bf16 columns

__device__ __forceinline__ void tma_load_2d_mc(void* dst, const CUtensorMap* m, int x, int y, uint64_t* bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;" ::
          "r"(ptx::smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(ptx::smem_u32(bar)), "r"(x), "r"(y), "h"(mask)
      : "memory");
}

template <int CS>
__global__ void __launch_bounds__(128, 1) k(const __grid_constant__ CUtensorMap tm, int mode, int iters, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar[2];
  const uint32_t rank = CS > 1 ? ptx::cluster_ctarank() : 0;
  const int cluster = blockIdx.x / CS;
  if (threadIdx.x == 0) {
    ptx::mbar_init(&bar[0], 1);
    ptx::mbar_init(&bar[1], 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  if (CS > 1) ptx::cluster_sync();
  long long t0 = 0;
  const bool lead = threadIdx.x == 0;
  if (lead) t0 = clock64();
  for (int i = 0; i < iters; ++i) {   // every thread walks the loop: barrier.cluster is .aligned
    const int g = i & 1;
    if (lead) {
      if (i >= 2) ptx::mbar_wait(&bar[g], ((i - 2) >> 1) & 1);   // the loads of iteration i-2 into this group have landed
      ptx::mbar_arrive_expect_tx(&bar[g], GROUP);
    }
    if (CS > 1 && mode == 2) {
      __syncwarp();
      ptx::cluster_sync();                                        // every peer has armed its barrier of this group
    }
    if (lead) {
      uint8_t* dst = smem + g * GROUP;
      if (mode == 2) {
        for (int t = rank; t < TILES; t += CS) {                 // my share of the cluster's tiles, to everybody
          const int row = ((cluster * 977 + i * 131 + t * 17) * 128) % (ROWS - 128);
          tma_load_2d_mc(dst + t * TILE, &tm, 0, row, &bar[g], (uint16_t)((1u << CS) - 1));
        }
      } else {
        const int who = mode == 1 ? cluster : blockIdx.x;
        for (int t = 0; t < TILES; ++t) {
          const int row = ((who * 977 + i * 131 + t * 17) * 128) % (ROWS - 128);
          ptx::tma_load_2d(dst + t * TILE, &tm, 0, row, &bar[g]);
        }
      }
    }
    __syncwarp();
  }
  if (lead) {
    for (int i = max(0, iters - 2); i < iters; ++i) ptx::mbar_wait(&bar[i & 1], (i >> 1) & 1);
    out[blockIdx.x] = clock64() - t0;
  }
  __syncthreads();
  if (CS > 1) ptx::cluster_sync();
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int CS>
void run(const CUtensorMap& tm, int mode, int iters, long long* d_out) {
  cudaFuncSetAttribute(k<CS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * GROUP);
  cudaLaunchConfig_t cfg{};
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = 2 * GROUP;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = CS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  int maxc = 0;
  cfg.gridDim = dim3(148 / CS * CS);
  if (CS > 8) cudaFuncSetAttribute(k<CS>, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  cudaError_t e = cudaOccupancyMaxActiveClusters(&maxc, k<CS>, &cfg);
  int ctas = maxc * CS < 148 ? maxc * CS : 148 / CS * CS;
  if (ctas <= 0) { printf("cluster %d: no active clusters (%s)\n", CS, cudaGetErrorString(e)); return; }
  cfg.gridDim = dim3(ctas);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaEventRecord(e0);
    cudaLaunchKernelEx(&cfg, k<CS>, tm, mode, iters, d_out);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
  }
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  long long h[148];
  cudaMemcpy(h, d_out, sizeof(long long) * ctas, cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < ctas; ++i) mx = h[i] > mx ? h[i] : mx;
  const double bytes_per_cta = (double)iters * GROUP;
  printf("cluster %d (max active clusters %d, %d CTAs) mode %d: %.1f B/clk received per SM, %.0f B/clk chip-wide, %.2f TB/s received "
         "(%.3f ms)  %s\n", CS, maxc, ctas, mode, bytes_per_cta / mx, bytes_per_cta / mx * ctas, bytes_per_cta * ctas / (ms * 1e-3) / 1e12, ms,
         cudaGetErrorString(cudaGetLastError()));
}

int main() {
  void* buf;
  cudaMalloc(&buf, (size_t)ROWS * 128);
  cudaMemset(buf, 1, (size_t)ROWS * 128);
  long long* d_out;
  cudaMalloc(&d_out, 148 * sizeof(long long));
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  CUtensorMap tm;
  cuuint64_t dims[2] = {64, (cuuint64_t)ROWS}, strides[1] = {128};
  cuuint32_t box[2] = {64, 128}, estr[2] = {1, 1};
  CUresult r = ((EncodeTiledFn)p)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { printf("tensor map failed %d\n", (int)r); return 1; }
  const int iters = 2000;
  run<1>(tm, 0, iters, d_out);
  run<2>(tm, 0, iters, d_out); run<2>(tm, 1, iters, d_out); run<2>(tm, 2, iters, d_out);
  run<4>(tm, 0, iters, d_out); run<4>(tm, 1, iters, d_out); run<4>(tm, 2, iters, d_out);
  run<8>(tm, 0, iters, d_out); run<8>(tm, 1, iters, d_out); run<8>(tm, 2, iters, d_out);
  return 0;
}
