// CTA-pair (cta_group::2) variant of the tcgen05 GEMM / implicit-GEMM conv in gemm_tc.cu, for the large contractions.
//
// Why: with one CTA per 128 x 256 tile every k-block moves 48 KB into the SM for 512 MMA cycles, and an SM ingests
// about 64 B/clk from L2 -- the mainloop is fill-bound at ~2/3 of the tensor rate (profiles/ncu_gemm_r01d.txt).
// A pair of CTAs on the two SMs of a TPC computes a 256 x 256 tile with ONE tcgen05.mma.cta_group::2 per k-step:
// each CTA loads its own 128 rows of A and only HALF of the B tile (the tensor cores read the other half from the
// peer's shared memory), i.e. 32 KB per k-block per SM, which balances the 512 MMA cycles.
//
// Roles per CTA (same as gemm_tc.cu): warp 0 TMA producer, warp 1 MMA issuer (leader CTA only), warps 2..9 epilogue
// (epilogue_tc.cuh, unchanged: each CTA drains its own 128 x 256 accumulator half from its own TMEM).
// Cross-CTA signalling: both producers' TMA bytes complete on the LEADER's full barrier; the MMA commit multicasts to
// both CTAs' empty / accumulator-full barriers; the peer's epilogue warps arrive remotely on the leader's
// accumulator-empty barrier.
#include "epilogue_tc.cuh"
#include "ptx.cuh"
#include "tmap.h"

#ifdef TC2_TRACE  // epilogue timeline of CTA 0, epilogue warp 0 (scripts/gemm_trace.py); debugging aid
__device__ long long g_tc2_trace[64 * 4];
extern "C" int vda_debug_tc2_trace(long long* out) { return (int)cudaMemcpyFromSymbol(out, g_tc2_trace, sizeof(g_tc2_trace)); }
#define TC2_T(ev) do { if (blockIdx.x == 0 && ew == 0 && lane == 0 && tcount < 64) g_tc2_trace[tcount * 4 + (ev)] = clock64(); } while (0)
#else
#define TC2_T(ev)
#endif
namespace {

constexpr int BM = 128, BK = 64, BN = 256;  // per-CTA accumulator: 128 x 256; pair tile 256 x 256
constexpr int A_BYTES = BM * BK * 2;          // 16 KB: this CTA's rows of A
constexpr int B_BYTES = (BN / 2) * BK * 2;    // 16 KB: this CTA's half of the B tile
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int TMEM_COLS = 2 * BN;
// Eight epilogue warps and six stages for every kind.  (The GELU kind of mlp.fc1 ran 16 warps and 5 stages while its
// epilogue sat behind a cluster-scope release per tile and the producer behind divisions; with both gone, 8 + 6 is 2 %
// faster there too: 269 against 275 us.)
template <int EK> struct PairCfg {
  static constexpr int NUM_EPI_WARPS = 8;
  static constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
  static constexpr int STAGES = 6;
  static constexpr int EPI_BYTES = NUM_EPI_WARPS * 4096 + 2 * BN * 4;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 256;
  static_assert(SMEM_BYTES <= 232448, "shared memory budget exceeded");
};

struct Tc2Args {
  int M, N, K;
  int taps, ksplit, m_padded, gb, gh, gw;
  int tiles_m, tiles_n;  // pair tiles (256 x 256)
};


template <int EK>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(PairCfg<EK>::NUM_THREADS, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Tc2Args a,
                const Epi e) {
  constexpr int NUM_EPI_WARPS = PairCfg<EK>::NUM_EPI_WARPS, STAGES = PairCfg<EK>::STAGES, EPI_BYTES = PairCfg<EK>::EPI_BYTES;
  constexpr int NGROUPS = NUM_EPI_WARPS / 4;  // column groups: warps of one group cover the four TMEM lane quadrants
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* epi_smem = smem + STAGES * STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + EPI_BYTES);
  uint64_t* full = bars;                    // [STAGES]  TMA (both CTAs) -> MMA; only the leader's are used
  uint64_t* empty = bars + STAGES;          // [STAGES]  MMA -> TMA, one set per CTA
  uint64_t* tfull = bars + 2 * STAGES;      // [2]       MMA -> epilogue, one set per CTA
  uint64_t* tempty = bars + 2 * STAGES + 2; // [2]       epilogue (both CTAs) -> MMA; only the leader's are used
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = (int)ptx::uniform_u32(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();  // 0 = leader
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int kb_per_tap = (a.K + BK - 1) / BK;
  const int num_kb = a.ksplit * a.taps * kb_per_tap;
  const int num_tiles = a.tiles_m * a.tiles_n;

  if (warp == 0 && lane == 0) {
    if (ptx::smem_u32(smem) & 1023) {
      printf("vda gemm_tc2: dynamic shared memory is not 1024B aligned\n");
      __trap();
    }
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], 2 * NUM_EPI_WARPS);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc_pair(tmem_slot, TMEM_COLS);
    ptx::tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // the peer's barriers are initialised before anything signals them
  ptx::tc_fence_after();
  const uint32_t tmem_base = ptx::uniform_u32(*tmem_slot);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs): the warp walks the loop,
    // one elected lane issues
    {
      int stage = 0;
      uint32_t phase = 0;
      // This one thread has to hand a k-block to the TMA unit every 512 clk (4 MMAs) for the tensor pipe to stay busy:
      // the loop carries its coordinates incrementally (no division per k-block: k-block -> (split, tap, column) is a
      // pair of counters) and keeps the leader's barrier address as one add.
      const uint32_t lead_full0 = ptx::mapa(ptx::smem_u32(&full[0]), 0);
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        const int m0 = (tile / a.tiles_n) * (2 * BM) + (int)rank * BM;
        const int n0 = (tile % a.tiles_n) * BN + (int)rank * (BN / 2);
        int kin = 0, tap = 0, bcol = 0;  // k-block inside the tap; tap; first B column of (split, tap) = st * K
        int off = vda_tap_off(a.taps, 0, a.m_padded, a.gb, a.gh, a.gw);
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          uint8_t* sb = sa + A_BYTES;
          const uint32_t lead_full = lead_full0 + stage * 8;
          if (ptx::elect_one()) {
            if (rank == 0) ptx::mbar_arrive_expect_tx(&full[stage], 2 * STAGE_BYTES);
            ptx::tma_load_2d_pair(sa, &tmA, kin * BK, m0 + off, lead_full);
            ptx::tma_load_2d_pair(sb, &tmB, bcol + kin * BK, n0, lead_full);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
          if (++kin == kb_per_tap) {
            kin = 0;
            bcol += a.K;
            if (++tap == a.taps) tap = 0;
            off = vda_tap_off(a.taps, tap, a.m_padded, a.gb, a.gh, a.gw);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only): warp-uniform loop,
    // one elected lane issues
    if (rank == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = pair; tile < num_tiles; tile += num_pairs) {
        ptx::mbar_wait(&tempty[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t sb = sa + A_BYTES;
          const uint64_t da = ptx::make_sw128_desc(sa, 16, 1024);
          const uint64_t db = ptx::make_sw128_desc(sb, 16, 1024);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k)
              ptx::mma_f16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            ptx::tc_commit_pair(&empty[stage], 3);
            if (kb == num_kb - 1) ptx::tc_commit_pair(&tfull[acc], 3);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs, own accumulator half)
    const int ew = warp - 2;
    const int q = warp & 3;
    const int grp = ew >> 2;
    const int et = threadIdx.x - 64;
    const bool geglu = (EK == EK_GENERIC) ? (e.act == VDA_ACT_GEGLU) : (EK == EK_GEGLU);
    const int cw = geglu ? 128 : 64;
    uint8_t* stage = epi_smem + ew * 4096;
    float* sbias_all = reinterpret_cast<float*>(epi_smem + NUM_EPI_WARPS * 4096);
    int acc = 0;
    uint32_t acc_phase = 0;
    [[maybe_unused]] int tcount = 0;
    // One bias column per epilogue thread (BN = 256 = 8 warps x 32), fetched one tile AHEAD: the global load's round trip
    // sits under the previous tile's epilogue instead of between the tiles -- the GELU kind is epilogue-bound (8.4 k clk
    // of epilogue per 8.0 k clk of MMAs per tile, scripts/gemm_trace.py), so what happens between two tiles is on its
    // critical path.
    static_assert(NUM_EPI_WARPS * 32 == BN, "one bias column per epilogue thread");
    auto bias_of = [&](int tile_) {
      const int n = (tile_ % a.tiles_n) * BN + et;
      return (tile_ < num_tiles && e.bias != nullptr && n < a.N) ? epi_bias(e, n) : 0.f;
    };
    float bias_next = bias_of(pair);
    for (int tile = pair; tile < num_tiles; tile += num_pairs, ++tcount) {
      const int m0 = (tile / a.tiles_n) * (2 * BM) + (int)rank * BM, n0 = (tile % a.tiles_n) * BN;
      float* sbias = sbias_all + acc * BN;
      sbias[et] = bias_next;
      // (reading the bias straight from global memory instead, to drop this CTA-wide barrier, made the GELU epilogue
      // slower: 9.2 k against 8.3 k clk per tile -- generic 16-byte loads in place of shared-memory ones)
      asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
      bias_next = bias_of(tile + num_pairs);
      TC2_T(0);
      ptx::mbar_wait(&tfull[acc], acc_phase);
      ptx::tc_fence_after();
      TC2_T(1);
      const int m = m0 + q * 32 + lane;
      const RowIdx ri = make_row_idx(e, m);
      const uint32_t trow = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
      for (int c = grp * cw; c < BN; c += NGROUPS * cw) {
        if (n0 + c >= a.N) break;
        if (c + cw <= BN && epi_fast_ok(e, n0 + c)) {
          if constexpr (EK == EK_GENERIC) {
            if (geglu) epi_chunk_tc<EK_GENERIC, true>(e, ri, n0, c, trow + c, stage, sbias, lane);
            else epi_chunk_tc<EK_GENERIC, false>(e, ri, n0, c, trow + c, stage, sbias, lane);
          } else {
            epi_chunk_tc<EK, EK == EK_GEGLU>(e, ri, n0, c, trow + c, stage, sbias, lane);
          }
        } else {
          const RowCtx r = make_row_ctx(e, m);
          const int sw = geglu ? 64 : 32;
          for (int cc = c; cc < c + cw && cc < BN && n0 + cc < a.N; cc += sw) {
            float v[32], g[32];
            ptx::tmem_ld32(trow + cc, v);
            if (geglu) ptx::tmem_ld32(trow + cc + 32, g);
            ptx::tmem_ld_wait();
            epi_chunk<bf16, true>(e, r, n0 + cc, v, g);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      TC2_T(2);
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tempty[acc]), 0));
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // the peer has finished reading this CTA's shared memory and signalling its barriers
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, TMEM_COLS);
  }
}

template <int EK>
int launch_tc2(const vda_gemm_desc* d, cudaStream_t st) {
  CUtensorMap tmA, tmB;
  int rc = vda_make_tmap_2d(&tmA, d->A, (uint64_t)d->K, (uint64_t)d->rowsA, (uint64_t)d->lda * 2, BK, BM, 2);
  if (rc) return rc;
  rc = vda_make_tmap_2d(&tmB, d->B, (uint64_t)d->ksplit * d->taps * d->K, (uint64_t)d->N, (uint64_t)d->ldb * 2, BK, BN / 2, 2);
  if (rc) return rc;
  Tc2Args a;
  a.M = d->M; a.N = d->N; a.K = d->K; a.taps = d->taps; a.ksplit = d->ksplit;
  a.m_padded = d->m_padded; a.gb = d->gb; a.gh = d->gh; a.gw = d->gw;
  a.tiles_m = (d->M + 2 * BM - 1) / (2 * BM);
  a.tiles_n = (d->N + BN - 1) / BN;
  const int tiles = a.tiles_m * a.tiles_n;
  const int max_pairs = vda_sm_count() / 2;
  const int grid = 2 * (tiles < max_pairs ? tiles : max_pairs);
  Epi e = make_epi(d);
  auto kern = gemm_tc2_kernel<EK>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PairCfg<EK>::SMEM_BYTES);
    if (ce != cudaSuccess) {
      vda_set_error("gemm_tc2: cudaFuncSetAttribute(%d) failed: %s", PairCfg<EK>::SMEM_BYTES, cudaGetErrorString(ce));
      return (int)ce;
    }
    attr_set = true;
  }
  kern<<<grid, PairCfg<EK>::NUM_THREADS, PairCfg<EK>::SMEM_BYTES, st>>>(tmA, tmB, a, e);
  VDA_CUDA_LAUNCH_CHECK("gemm_tc2_kernel");
  return 0;
}

}  // namespace

// bf16 x bf16, round_bf16, N >= 256: checked by the caller (vda_gemm_tc).
int vda_gemm_tc_pair(const vda_gemm_desc* d, cudaStream_t st) {
  switch (epi_kind_of(d)) {
    case EK_STORE: return launch_tc2<EK_STORE>(d, st);
    case EK_GELU: return launch_tc2<EK_GELU>(d, st);
    case EK_RESID: return launch_tc2<EK_RESID>(d, st);
    case EK_GEGLU: return launch_tc2<EK_GEGLU>(d, st);
    case EK_RES: return launch_tc2<EK_RES>(d, st);
    default: return launch_tc2<EK_GENERIC>(d, st);
  }
}
