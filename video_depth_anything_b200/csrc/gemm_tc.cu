// tcgen05 / TMEM / TMA GEMM with fused epilogues (bf16 operands, fp32 accumulation in tensor memory).
//
// Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer (+ TMEM owner), warps 2..9 = epilogue.
// Tile 128 x BN x 64, UMMA 128 x BN x 16 (cta_group::1), STAGES-deep smem ring, two TMEM accumulator stages so the
// epilogue of tile i overlaps the main loop of tile i+1.  A 3x3 convolution is the same kernel: with padded
// channels-last activations each tap is a constant row offset of the A tensor map (K loop = 9 taps x K/64).
#include <stdlib.h>

#include "epilogue_tc.cuh"
#include "ptx.cuh"
#include "tmap.h"

namespace {

constexpr int BM = 128, BK = 64;
constexpr int NUM_EPI_WARPS = 8;
constexpr int NUM_THREADS = 64 + NUM_EPI_WARPS * 32;
constexpr int SMEM_LIMIT = 232448;  // 227 KB opt-in maximum per CTA

template <int BN>
struct TcCfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (196608 / STAGE_BYTES) > 8 ? 8 : (196608 / STAGE_BYTES);
  static constexpr int TMEM_COLS = (2 * BN < 32) ? 32 : 2 * BN;  // power of two for BN in {32,64,128,256}
  static constexpr int EPI_BYTES = NUM_EPI_WARPS * 4096 /*staging tiles*/ + 2 * BN * 4 /*bias tiles*/;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + 256 /*barriers*/;
};

struct TcArgs {
  int M, N, K;         // K per tap
  int taps, ksplit;    // K loop = ksplit * taps * ceil(K/64)
  int m_padded, gb, gh, gw;  // tap row offsets (vda_tap_off)
  int tiles_m, tiles_n;
};

template <int BN, typename TO, bool RB, bool TF32, int EK>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcArgs a,
               const Epi e) {
  using C = TcCfg<BN>;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* epi_smem = smem + C::STAGES * C::STAGE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(epi_smem + C::EPI_BYTES);
  uint64_t* full = bars;                     // [STAGES]   TMA -> MMA
  uint64_t* empty = bars + C::STAGES;        // [STAGES]   MMA -> TMA
  uint64_t* tfull = bars + 2 * C::STAGES;    // [2]        MMA -> epilogue
  uint64_t* tempty = bars + 2 * C::STAGES + 2;  // [2]     epilogue -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * C::STAGES + 4);

  const int warp = (int)ptx::uniform_u32(threadIdx.x >> 5), lane = threadIdx.x & 31;
  constexpr int BKE = TF32 ? 32 : 64;  // elements per 128-byte K block
  const int kb_per_tap = (a.K + BKE - 1) / BKE;
  const int num_kb = a.ksplit * a.taps * kb_per_tap;
  const int num_tiles = a.tiles_m * a.tiles_n;

  if (warp == 0 && lane == 0) {
    if (ptx::smem_u32(smem) & 1023) {
      printf("vda gemm_tc: dynamic shared memory is not 1024B aligned\n");
      __trap();
    }
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < C::STAGES; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], BN <= 64 ? NUM_EPI_WARPS / 2 : NUM_EPI_WARPS);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = ptx::uniform_u32(*tmem_slot);

  // The two role loops below are walked by their whole warp with the issuing instructions under ptx::elect_one(): the
  // operands stay in uniform registers (ptx.cuh), which is worth 5 % of the kernel over a loop under `if (lane == 0)`.
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / a.tiles_n) * BM, n0 = (tile % a.tiles_n) * BN;
        // coordinates carried incrementally: this one thread feeds the TMA unit a k-block per 4 MMAs, no division per step
        int kin = 0, tap = 0, bcol = 0;  // k-block inside the tap; tap; first B column of (split, tap) = st * K
        int off = vda_tap_off(a.taps, 0, a.m_padded, a.gb, a.gh, a.gw);
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&empty[stage], phase ^ 1);
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          uint8_t* sb = sa + C::A_BYTES;
          if (ptx::elect_one()) {
            ptx::mbar_arrive_expect_tx(&full[stage], C::STAGE_BYTES);
            ptx::tma_load_2d(sa, &tmA, kin * BKE, m0 + off, &full[stage]);
            ptx::tma_load_2d(sb, &tmB, bcol + kin * BKE, n0, &full[stage]);
          }
          __syncwarp();
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
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
    // ------------------------------------------------------------------ MMA issuer
    {
      constexpr uint32_t idesc = TF32 ? ptx::make_idesc_tf32(BM, BN) : ptx::make_idesc_bf16(BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        ptx::mbar_wait(&tempty[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full[stage], phase);
          ptx::tc_fence_after();
          const uint32_t sa = ptx::smem_u32(smem + stage * C::STAGE_BYTES);
          const uint32_t sb = sa + C::A_BYTES;
          const uint64_t da = ptx::make_sw128_desc(sa, 16, 1024);
          const uint64_t db = ptx::make_sw128_desc(sb, 16, 1024);
          if (ptx::elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              // advance 32 B (16 bf16 / 8 tf32) inside the 128 B swizzle row: +2 in the (addr >> 4) field
              if constexpr (TF32) ptx::mma_tf32_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
              else ptx::mma_f16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
            }
            ptx::tc_commit(&empty[stage]);
            if (kb == num_kb - 1) ptx::tc_commit(&tfull[acc]);
          }
          __syncwarp();
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int ew = warp - 2;                   // 0..7
    const int q = warp & 3;                    // TMEM lane quadrant this warp may access
    const int half = ew >> 2;                  // which interleaved set of column chunks
    const int et = threadIdx.x - 64;           // 0..255
    const bool geglu = (EK == EK_GENERIC) ? (e.act == VDA_ACT_GEGLU) : (EK == EK_GEGLU);
    const int cw = geglu ? 128 : 64;           // accumulator columns per chunk (64 output columns)
    uint8_t* stage = epi_smem + ew * 4096;
    float* sbias_all = reinterpret_cast<float*>(epi_smem + NUM_EPI_WARPS * 4096);
    // Narrow tiles (BN <= 64) are one 64-column chunk: interleaving chunks over the two warp sets would idle one of them,
    // and a tile's epilogue is a single latency chain per scheduler (~5000 clk for a K = 64 1x1 convolution whose
    // mainloop is four MMAs).  The sets take ALTERNATE tiles instead (set == accumulator stage), two epilogues in flight.
    constexpr bool SPLIT = BN <= 64;
    constexpr int SET_THREADS = SPLIT ? NUM_EPI_WARPS * 16 : NUM_EPI_WARPS * 32;
    const int eti = SPLIT ? (et & (SET_THREADS - 1)) : et;
    int acc = SPLIT ? half : 0;
    uint32_t acc_phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      if (SPLIT && (it & 1) != half) continue;
      const int m0 = (tile / a.tiles_n) * BM, n0 = (tile % a.tiles_n) * BN;
      // bias values of this tile's columns -> smem (double-buffered with the accumulator stage: one barrier per tile)
      float* sbias = sbias_all + acc * BN;
      for (int i = eti; i < BN; i += SET_THREADS) {
        const int n = n0 + i;
        sbias[i] = (e.bias != nullptr && n < a.N) ? epi_bias(e, n) : 0.f;
      }
      if constexpr (SPLIT) asm volatile("bar.sync %0, %1;" ::"r"(2 + half), "n"(SET_THREADS) : "memory");
      else asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI_WARPS * 32) : "memory");
      ptx::mbar_wait(&tfull[acc], acc_phase);
      ptx::tc_fence_after();
      const int m = m0 + q * 32 + lane;
      const RowIdx ri = make_row_idx(e, m);
      const uint32_t trow = tmem_base + acc * BN + ((uint32_t)(q * 32) << 16);
      for (int c = SPLIT ? 0 : half * cw; c < BN; c += SPLIT ? cw : 2 * cw) {
        if (n0 + c >= a.N) break;
        if (c + cw <= BN && epi_fast_ok(e, n0 + c)) {
          if constexpr (EK == EK_GENERIC) {
            if (geglu) epi_chunk_tc<EK_GENERIC, true>(e, ri, n0, c, trow + c, stage, sbias, lane);
            else epi_chunk_tc<EK_GENERIC, false>(e, ri, n0, c, trow + c, stage, sbias, lane);
          } else {
            epi_chunk_tc<EK, EK == EK_GEGLU>(e, ri, n0, c, trow + c, stage, sbias, lane);
          }
        } else {
          // ragged chunk (N tail, narrow conv-transpose blocks, the fused fp32 head): one row per thread
          const RowCtx r = make_row_ctx(e, m);
          const int sw = geglu ? 64 : 32;
          for (int cc = c; cc < c + cw && cc < BN && n0 + cc < a.N; cc += sw) {
            float v[32], g[32];
            ptx::tmem_ld32(trow + cc, v);
            if (geglu) ptx::tmem_ld32(trow + cc + 32, g);
            ptx::tmem_ld_wait();
            epi_chunk<TO, RB>(e, r, n0 + cc, v, g);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty[acc]);
      if constexpr (SPLIT) acc_phase ^= 1;
      else if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

template <int BN, typename TO, bool RB, bool TF32, int EK = EK_GENERIC>
int launch_tc(const vda_gemm_desc* d, cudaStream_t st) {
  using C = TcCfg<BN>;
  constexpr int EB = TF32 ? 4 : 2, BKE = 128 / EB;
  CUtensorMap tmA, tmB;
  int rc = vda_make_tmap_2d(&tmA, d->A, (uint64_t)d->K, (uint64_t)d->rowsA, (uint64_t)d->lda * EB, BKE, BM, EB);
  if (rc) return rc;
  rc = vda_make_tmap_2d(&tmB, d->B, (uint64_t)d->ksplit * d->taps * d->K, (uint64_t)d->N, (uint64_t)d->ldb * EB, BKE, BN, EB);
  if (rc) return rc;
  TcArgs a;
  a.M = d->M; a.N = d->N; a.K = d->K; a.taps = d->taps; a.ksplit = d->ksplit;
  a.m_padded = d->m_padded; a.gb = d->gb; a.gh = d->gh; a.gw = d->gw;
  a.tiles_m = (d->M + BM - 1) / BM;
  a.tiles_n = (d->N + BN - 1) / BN;
  const int tiles = a.tiles_m * a.tiles_n;
  const int grid = tiles < vda_sm_count() ? tiles : vda_sm_count();
  Epi e = make_epi(d);
  static_assert(C::SMEM_BYTES <= SMEM_LIMIT, "shared memory budget exceeded");
  auto kern = gemm_tc_kernel<BN, TO, RB, TF32, EK>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES);
    if (ce != cudaSuccess) {
      vda_set_error("gemm_tc: cudaFuncSetAttribute(%d) failed: %s", C::SMEM_BYTES, cudaGetErrorString(ce));
      return (int)ce;
    }
    attr_set = true;
  }
  kern<<<grid, NUM_THREADS, C::SMEM_BYTES, st>>>(tmA, tmB, a, e);
  VDA_CUDA_LAUNCH_CHECK("gemm_tc_kernel");
  return 0;
}

template <int BN>
int launch_kind(const vda_gemm_desc* d, cudaStream_t st) {
  switch (epi_kind_of(d)) {
    case EK_STORE: return launch_tc<BN, bf16, true, false, EK_STORE>(d, st);
    case EK_GELU: return launch_tc<BN, bf16, true, false, EK_GELU>(d, st);
    case EK_RESID: return launch_tc<BN, bf16, true, false, EK_RESID>(d, st);
    case EK_GEGLU: return launch_tc<BN, bf16, true, false, EK_GEGLU>(d, st);
    case EK_RES: return launch_tc<BN, bf16, true, false, EK_RES>(d, st);
    default: return launch_tc<BN, bf16, true, false, EK_GENERIC>(d, st);
  }
}

}  // namespace

int vda_gemm_tc_pair(const vda_gemm_desc* d, cudaStream_t st);       // gemm_tc2.cu
int vda_conv3x3_tc_shared(const vda_gemm_desc* d, cudaStream_t st);  // conv_tc.cu

// VDA_NO_PAIR=1 keeps every GEMM on the single-CTA kernel, VDA_NO_CONVSHARE=1 keeps the narrow 3x3 convolutions on the
// per-tap kernel (A/B measurements).
static bool env_off(const char* name) {
  const char* s = getenv(name);
  return s && s[0] == '1';
}
static bool pair_enabled() {
  static int v = -1;
  if (v < 0) v = env_off("VDA_NO_PAIR") ? 0 : 1;
  return v == 1;
}
// smallest N handled by the tap-sharing conv kernel (its weight tile is 128 wide: narrower convs compute zero columns,
// still far cheaper than nine A loads per tile and an epilogue per 128 x N tile); VDA_CONVSHARE_MIN_N overrides
static int convshare_min_n() {
  static int v = -1;
  if (v < 0) {
    const char* s = getenv("VDA_CONVSHARE_MIN_N");
    v = s ? atoi(s) : 32;
  }
  return v;
}
static bool convshare_enabled() {
  static int v = -1;
  if (v < 0) v = env_off("VDA_NO_CONVSHARE") ? 0 : 1;
  return v == 1;
}

int vda_gemm_tc(const vda_gemm_desc* d, cudaStream_t st) {
  const bool tf32 = d->a_dtype == VDA_F32 && d->b_dtype == VDA_F32;
  VDA_CHECK_ARG(tf32 || (d->a_dtype == VDA_BF16 && d->b_dtype == VDA_BF16), "vda_gemm(tc): operands must be bf16/bf16 or f32/f32 (tf32)");
  VDA_CHECK_ARG(d->out_dtype == VDA_BF16 && d->round_bf16, "vda_gemm(tc): the tensor-core path is the bf16 build");
  const int al = tf32 ? 4 : 8;
  VDA_CHECK_ARG((d->lda % al) == 0 && (d->ldb % al) == 0, "vda_gemm(tc): lda/ldb must be multiples of 16 bytes (TMA)");
  VDA_CHECK_ARG(((uintptr_t)d->A % 16) == 0 && ((uintptr_t)d->B % 16) == 0, "vda_gemm(tc): operands must be 16B aligned");
  VDA_CHECK_ARG(d->act != VDA_ACT_GEGLU || d->N % 64 == 0, "vda_gemm(tc): GEGLU needs N %% 64 == 0");
  VDA_CHECK_ARG(d->head_w == nullptr || d->N == 32, "vda_gemm(tc): fused head needs N == 32");
  VDA_CHECK_ARG(d->act != VDA_ACT_GEGLU || (d->res == nullptr && d->res2 == nullptr), "vda_gemm(tc): GEGLU has no skip inputs");
  if (tf32) {
    // the fp32 island of the bf16 build (dpt_temporal.py:95-97): fp32 operands through kind::tf32, as cuDNN does by default
    VDA_CHECK_ARG(d->N <= 32, "vda_gemm(tc, tf32): only the N <= 32 output head is instantiated");
    return launch_tc<32, bf16, true, true>(d, st);
  }
  // 3x3 convolutions: tap-sharing kernel (one A load per three taps, two M tiles per weight tile)
  if (d->taps == 9 && d->m_padded == 1 && d->ksplit == 1 && d->N >= convshare_min_n() && d->K % 64 == 0 && d->head_w == nullptr &&
      d->M >= 2048 && convshare_enabled())
    return vda_conv3x3_tc_shared(d, st);
  int bn;
  if (d->N <= 32) bn = 32;
  else if (d->N <= 64) bn = 64;
  else if (d->N <= 128 || (d->N % 256 != 0 && d->N % 128 == 0 && d->N < 1024)) bn = 128;
  else bn = 256;
  // small problems: prefer more tiles over wide tiles so every SM gets work
  if (bn == 256 && (int64_t)((d->M + 127) / 128) * ((d->N + 255) / 256) < vda_sm_count()) bn = 128;
  // large contractions: CTA pairs (256 x 256 tile per pair, half the B traffic per SM)
  // (short contractions are epilogue / HBM bound and run better without the pair's lock-step)
  if (bn == 256 && pair_enabled() && (int64_t)d->K * d->taps * d->ksplit >= 512 &&
      (int64_t)((d->M + 255) / 256) * ((d->N + 255) / 256) >= vda_sm_count() / 2)
    return vda_gemm_tc_pair(d, st);
  switch (bn) {
    case 32: return launch_tc<32, bf16, true, false>(d, st);
    case 64: return launch_tc<64, bf16, true, false>(d, st);
    case 128: return launch_kind<128>(d, st);
    default: return launch_kind<256>(d, st);
  }
}
