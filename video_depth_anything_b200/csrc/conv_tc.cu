// 3x3 convolution (stride 1, pad 1) over a padded channels-last map as a tcgen05 implicit GEMM with TAP SHARING,
// (DPT `output_conv1`, dpt_temporal.py:91 / dpt.py:117; the fusion-block and layer_rn convolutions, blocks.py:20-33,68-91).
//
// Why a second kernel: in gemm_tc.cu every (tap, k-block) step loads its own 128-row A tile, so a 128 x 128 output
// tile moves 32 KB into the SM per 256 MMA cycles -- twice what an SM can take in (~64 B/clk): `output_conv1` ran at
// 729 TFLOP/s with 26 GB through L1 for 2.1 GB of DRAM traffic (profiles/ncu_gemm_r01f.txt).  Two observations fix it:
//  * M runs over the padded pixel grid, so the A rows of tap (dy, dx) are the rows of tap (dy, 0) shifted by dx: ONE
//    load of rows [m0 + dy W2 - 1, m0 + dy W2 + 257) serves the three dx taps; the MMA reads it through descriptors
//    whose start address is advanced by dx rows (128 B) inside the 128B-swizzled tile (the swizzle is a function of the
//    shared-memory address bits, so a row-shifted start reads the rows TMA wrote);
//  * two 128-row M tiles (two TMEM accumulators) share every B (weight) tile.
// Per (dy, k-block) group: A 33 KB + 3 B tiles of 16 KB for 6 x 4 MMAs of 128x128x16 = 1536 tensor cycles, i.e.
// 53 B/clk instead of 128.  A and B travel through separate rings (A slots live three times as long).
//
// Roles as in gemm_tc.cu: warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 epilogue (epilogue_tc.cuh), persistent
// over 256-row tiles, double-buffered accumulators (2 stages x 2 M tiles x 128 columns = all 512 TMEM columns).
#include "epilogue_tc.cuh"
#include "ptx.cuh"
#include "tmap.h"

namespace {

constexpr int BM = 128, BN = 128, BK = 64;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
constexpr int A_BOX_ROWS = 88;                       // three boxes = 264 rows >= 2 * 128 + 2 (TMA boxes are <= 256 rows)
constexpr int A_BOX_BYTES = A_BOX_ROWS * BK * 2;     // 11 KB, a multiple of the 1024 B swizzle period
constexpr int A_BYTES = 3 * A_BOX_BYTES;             // 33 KB
constexpr int B_BYTES = BN * BK * 2;                 // 16 KB
constexpr int NA = 3, NB = 5;                        // ring depths
constexpr int EPI_BYTES = NUM_EPI_WARPS * 4096 + 2 * BN * 4;
constexpr int OFF_B = NA * A_BYTES;
constexpr int OFF_EPI = OFF_B + NB * B_BYTES;
constexpr int SMEM_BYTES = OFF_EPI + EPI_BYTES + 256;
constexpr int TMEM_COLS = 512;
static_assert(A_BOX_BYTES % 1024 == 0 && SMEM_BYTES <= 232448, "shared memory layout");

struct ConvArgs {
  int M, N, K;   // M over the padded grid, K = input channels (per tap)
  int W2;        // padded row length
  int tiles_m;   // 256-row tiles
  int tiles_n;   // 128-column tiles (n fastest: the CTAs sharing A rows run together and hit L2)
};

template <int EK>
__global__ void __launch_bounds__(NUM_THREADS, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ConvArgs a,
                  const Epi e) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* epi_smem = smem + OFF_EPI;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + EPI_BYTES);
  uint64_t* afull = bars;                 // [NA] TMA -> MMA
  uint64_t* aempty = afull + NA;          // [NA] MMA -> TMA
  uint64_t* bfull = aempty + NA;          // [NB]
  uint64_t* bempty = bfull + NB;          // [NB]
  uint64_t* tfull = bempty + NB;          // [2]  MMA -> epilogue
  uint64_t* tempty = tfull + 2;           // [2]  epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  static_assert((2 * NA + 2 * NB + 4) * 8 + 8 <= 256, "barrier area");

  const int warp = (int)ptx::uniform_u32(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int kchunks = a.K / BK;
  const int ngroups = 3 * kchunks;  // (k-block, dy)
  const int num_tiles = a.tiles_m * a.tiles_n;

  if (warp == 0 && lane == 0) {
    if (ptx::smem_u32(smem) & 1023) {
      printf("vda conv3x3_tc: dynamic shared memory is not 1024B aligned\n");
      __trap();
    }
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < NA; ++s) { ptx::mbar_init(&afull[s], 1); ptx::mbar_init(&aempty[s], 1); }
    for (int s = 0; s < NB; ++s) { ptx::mbar_init(&bfull[s], 1); ptx::mbar_init(&bempty[s], 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(&tfull[s], 1); ptx::mbar_init(&tempty[s], NUM_EPI_WARPS); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = ptx::uniform_u32(*tmem_slot);

  // role loops walked by the whole warp, issuing instructions under ptx::elect_one() (uniform-register operands, ptx.cuh)
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    {
      int as = 0, bs = 0;
      uint32_t aph = 0, bph = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / a.tiles_n) * 2 * BM, n0 = (tile % a.tiles_n) * BN;
        for (int g = 0; g < ngroups; ++g) {
          const int kc = g / 3, dy = g % 3;
          ptx::mbar_wait(&aempty[as], aph ^ 1);
          uint8_t* sa = smem + as * A_BYTES;
          const int r0 = m0 + (dy - 1) * a.W2 - 1;  // rows before / past the map are zero-filled by TMA
          if (ptx::elect_one()) {
            ptx::mbar_arrive_expect_tx(&afull[as], A_BYTES);
#pragma unroll
            for (int i = 0; i < 3; ++i) ptx::tma_load_2d(sa + i * A_BOX_BYTES, &tmA, kc * BK, r0 + i * A_BOX_ROWS, &afull[as]);
          }
          __syncwarp();
          if (++as == NA) { as = 0; aph ^= 1; }
          for (int dx = 0; dx < 3; ++dx) {
            ptx::mbar_wait(&bempty[bs], bph ^ 1);
            if (ptx::elect_one()) {
              ptx::mbar_arrive_expect_tx(&bfull[bs], B_BYTES);
              ptx::tma_load_2d(smem + OFF_B + bs * B_BYTES, &tmB, (dy * 3 + dx) * a.K + kc * BK, n0, &bfull[bs]);
            }
            __syncwarp();
            if (++bs == NB) { bs = 0; bph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(BM, BN, 0, 0);
      int as = 0, bs = 0, acc = 0;
      uint32_t aph = 0, bph = 0, acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        ptx::mbar_wait(&tempty[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        for (int g = 0; g < ngroups; ++g) {
          ptx::mbar_wait(&afull[as], aph);
          const uint32_t sa = ptx::smem_u32(smem + as * A_BYTES);
          for (int dx = 0; dx < 3; ++dx) {
            ptx::mbar_wait(&bfull[bs], bph);
            ptx::tc_fence_after();
            const uint64_t db = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_B + bs * B_BYTES), 16, 1024);
            if (ptx::elect_one()) {
#pragma unroll
              for (int i = 0; i < 2; ++i) {
                // rows [128 i + dx, +128) of the 264-row A slot: the start address moves by whole 128 B rows
                const uint64_t da = ptx::make_sw128_desc(sa + (i * BM + dx) * 128, 16, 1024);
                const uint32_t d_tmem = tmem_base + (acc * 2 + i) * BN;
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) ptx::mma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (g | dx | k) != 0);
              }
              ptx::tc_commit(&bempty[bs]);
              if (dx == 2) {
                ptx::tc_commit(&aempty[as]);
                if (g == ngroups - 1) ptx::tc_commit(&tfull[acc]);
              }
            }
            __syncwarp();
            if (++bs == NB) { bs = 0; bph ^= 1; }
          }
          if (++as == NA) { as = 0; aph ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps: one 64-column chunk per M tile each
    const int ew = warp - 2;
    const int q = warp & 3;
    const int half = ew >> 2;
    const int et = threadIdx.x - 64;
    uint8_t* stage = epi_smem + ew * 4096;
    float* sbias_all = reinterpret_cast<float*>(epi_smem + NUM_EPI_WARPS * 4096);
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile / a.tiles_n) * 2 * BM, n0 = (tile % a.tiles_n) * BN;
      float* sbias = sbias_all + acc * BN;
      for (int i = et; i < BN; i += NUM_EPI_WARPS * 32) sbias[i] = (e.bias != nullptr && n0 + i < a.N) ? e.bias[n0 + i] : 0.f;
      asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
      ptx::mbar_wait(&tfull[acc], acc_phase);
      ptx::tc_fence_after();
      // N <= 64: one 64-column chunk per M tile -- the two warp sets take one M tile each instead of column halves
      const bool narrow = a.N <= 64;
      const int c = narrow ? 0 : half * 64;
#pragma unroll 1
      for (int i = narrow ? half : 0; i < (narrow ? half + 1 : 2); ++i) {
        const int m = m0 + i * BM + q * 32 + lane;
        const uint32_t trow = tmem_base + (acc * 2 + i) * BN + ((uint32_t)(q * 32) << 16);
        if (n0 + c >= a.N) break;
        if (epi_fast_ok(e, n0 + c)) {
          const RowIdx ri = make_row_idx(e, m);
          epi_chunk_tc<EK, false>(e, ri, n0, c, trow + c, stage, sbias, lane);
        } else {
          const RowCtx r = make_row_ctx(e, m);
          for (int cc = c; cc < c + 64 && n0 + cc < a.N; cc += 32) {
            float v[32], g[32];
            ptx::tmem_ld32(trow + cc, v);
            ptx::tmem_ld_wait();
            epi_chunk<bf16, true>(e, r, n0 + cc, v, g);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

template <int EK>
int launch_conv(const vda_gemm_desc* d, cudaStream_t st) {
  CUtensorMap tmA, tmB;
  int rc = vda_make_tmap_2d(&tmA, d->A, (uint64_t)d->K, (uint64_t)d->rowsA, (uint64_t)d->lda * 2, BK, A_BOX_ROWS, 2);
  if (rc) return rc;
  rc = vda_make_tmap_2d(&tmB, d->B, (uint64_t)9 * d->K, (uint64_t)d->N, (uint64_t)d->ldb * 2, BK, BN, 2);
  if (rc) return rc;
  ConvArgs a;
  a.M = d->M; a.N = d->N; a.K = d->K; a.W2 = d->gw + 2;
  a.tiles_m = (d->M + 2 * BM - 1) / (2 * BM);
  a.tiles_n = (d->N + BN - 1) / BN;
  const int tiles = a.tiles_m * a.tiles_n;
  const int grid = tiles < vda_sm_count() ? tiles : vda_sm_count();
  Epi e = make_epi(d);
  auto kern = conv3x3_tc_kernel<EK>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (ce != cudaSuccess) {
      vda_set_error("conv3x3_tc: cudaFuncSetAttribute(%d) failed: %s", SMEM_BYTES, cudaGetErrorString(ce));
      return (int)ce;
    }
    attr_set = true;
  }
  kern<<<grid, NUM_THREADS, SMEM_BYTES, st>>>(tmA, tmB, a, e);
  VDA_CUDA_LAUNCH_CHECK("conv3x3_tc_kernel");
  return 0;
}

}  // namespace

// taps == 9, ksplit == 1, bf16 x bf16, N > 64, K % 64 == 0: checked by the caller (vda_gemm_tc).
int vda_conv3x3_tc_shared(const vda_gemm_desc* d, cudaStream_t st) {
  const int kind = epi_kind_of(d);
  if (kind == EK_STORE) return launch_conv<EK_STORE>(d, st);
  if (kind == EK_RES) return launch_conv<EK_RES>(d, st);
  return launch_conv<EK_GENERIC>(d, st);
}
