// tcgen05 flash attention for the DINOv2 encoder (dinov2_layers/attention.py:49-59), head_dim 64, non-causal.
//
// One CTA = 128 query rows (one tile) of one (frame, head); TWO CTAs are resident per SM (256 TMEM columns, 112 KB of
// shared memory and 11 warps each), so one CTA's prologue (barrier set-up, TMEM allocation, the first Q / K round trip)
// and epilogue (merge, store, teardown) run under the other's main loop -- with one 256-row CTA per SM that per-CTA
// overhead was a quarter of the kernel.  The work of a CTA is cut into TWO independent streams (key half hh): stream hh
// runs an online softmax of the tile's rows over the keys [128 j + 64 hh, +64) of every key block j, with its own S and
// O accumulators in TMEM, its own running max / sum and its own MMA issuer, and the halves are merged once, at the end
// (a two-way split-KV).
//   warp 0      : TMA producer (the Q tile once, then K_j / V_j tiles of 128 keys through 2-stage rings)
//   warps 1-2   : MMA issuer of stream w-1 (one thread each): S = Q K_j[hh]^T (128x64x64), O += P V_j[hh] (128x64x64)
//   warps 3-10  : softmax, four warps per stream (one per TMEM lane quadrant); a thread owns one query row x 64 keys
// Why streams: a softmax warp's iteration is a CHAIN of waits (S ready -> TMEM read -> max -> 64 exponentials -> PV
// done -> P store), and the exponentials (MUFU.EX2, 4 lanes/clk/scheduler: scripts/micro/xu_pipe.cu) are the busiest
// unit.  With one issuer and an exchange of the row max between the key halves (the previous version) the sixteen
// warps ran in lock step -- all four warps of a scheduler queued on the MUFU together, then left it idle together
// (per-phase clock64 trace: 1750 clk in the exponentials with ~3.4 warps contending, 1750 clk outside them).
// Independent streams have no per-block exchange and no in-order issue coupling, so they drift apart and one stream's
// exponentials run under the others' waits.  What bounds the loop now is instruction issue (the exponential phase is
// ~290 instructions per warp and four softmax warps share a scheduler: replacing MUFU.EX2 by an FMA changes nothing)
// and tcgen05.mma's ~48 clk floor per instruction at N = 64 (scripts/micro/mma_rate.cu).
// A softmax thread reads its 64 scores out of TMEM ONCE, rounding to bf16 and packing as it goes (32 registers),
// releases the S tile immediately (the next S MMA overlaps the exponentials), then: online max / sum in fp32, P
// written to shared memory as the bf16 A operand of the PV MMA, O accumulated in TMEM across key blocks (rescaled
// lazily, only when the running max moves by more than 2^8).
// Tried and dropped: evaluating half of the exponentials with a Cody-Waite + cubic polynomial on the FMA pipe (the
// FlashAttention-4 trick): 12 % slower, the FMA pipe and the issue slots are nearly as loaded as the MUFU.
// Numerics follow autocast: scores are rounded to bf16 (q@k^T is a bf16 tensor; the 2^-3 scale commutes with the
// rounding), softmax runs in fp32, probabilities are rounded to bf16 before P@V, the output is rounded to bf16.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"

#ifdef FA_TRACE  // per-phase clock64 timeline of the four softmax warps of one scheduler (scripts/fa_trace.py); debugging aid
__device__ long long g_fa_trace[4 * 16 * 8];  // [stream][block j][event]
#define FA_T(ev) do { if (blockIdx.x == 4 && blockIdx.y == 5 && blockIdx.z == 7 && q == 2 && lane == 0 && j < 16) \
                        g_fa_trace[(sid * 16 + j) * 8 + (ev)] = clock64(); } while (0)
extern "C" int vda_debug_fa_trace(long long* out) { return (int)cudaMemcpyFromSymbol(out, g_fa_trace, sizeof(g_fa_trace)); }
#else
#define FA_T(ev)
#endif
namespace {

constexpr int BQ = 128, BKV = 128, HD = 64, KH = 64;  // KH: keys per stream and block
constexpr int NS = 2;                  // K / V ring depth
constexpr int NSTREAM = 2;
constexpr int TILE_BYTES = 128 * 128;  // 128 rows x 64 bf16
constexpr int OFF_Q = 0;                         // 1 tile; after the last S MMA it holds the (max, sum) exchange
constexpr int OFF_K = TILE_BYTES;                // NS tiles
constexpr int OFF_V = OFF_K + NS * TILE_BYTES;   // NS tiles
constexpr int OFF_P = OFF_V + NS * TILE_BYTES;   // 2 streams x (128 rows x 64 keys)
constexpr int OFF_BAR = OFF_P + NSTREAM * TILE_BYTES;
constexpr int OFF_X = OFF_Q;
constexpr int FA_SMEM = OFF_BAR + 256;           // 112.25 KB: two CTAs per SM (228 KB, 1 KB reserved per CTA)
constexpr int FA_WARPS = 1 + NSTREAM + 4 * NSTREAM;
constexpr int FA_THREADS = FA_WARPS * 32;
constexpr uint32_t TM_S = 0, TM_O = 128, TM_COLS = 256;  // S of stream s at TM_S + 64 s, O at TM_O + 64 s

// barrier indices
constexpr int B_QFULL = 0;               // 1
constexpr int B_KFULL = 1;               // NS
constexpr int B_KEMPTY = B_KFULL + NS;   // NS
constexpr int B_VFULL = B_KEMPTY + NS;   // NS
constexpr int B_VEMPTY = B_VFULL + NS;   // NS
constexpr int B_SFULL = B_VEMPTY + NS;         // per stream
constexpr int B_SFREE = B_SFULL + NSTREAM;
constexpr int B_PFULL = B_SFREE + NSTREAM;
constexpr int B_PVDONE = B_PFULL + NSTREAM;
constexpr int B_QFREE = B_PVDONE + NSTREAM;    // 1: every S MMA has read Q (its tile then carries the final exchange)
constexpr int NBARS = B_QFREE + 1;
static_assert(NBARS * 8 + 8 <= 256, "barrier area");
static_assert(2 * (FA_SMEM + 1024) <= 233472, "two CTAs per SM");

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

__global__ void __launch_bounds__(FA_THREADS, 2)
flash_attn_tc_kernel(const __grid_constant__ CUtensorMap tm, bf16* __restrict__ o, int ntok, int heads) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + NBARS);
  const int warp = (int)ptx::uniform_u32(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, f = blockIdx.z;
  const int D = heads * HD;
  const int nblk = (ntok + BKV - 1) / BKV;
  const int row0 = f * ntok;                    // first row of this frame in the [frames*ntok, 3D] matrix

  if (threadIdx.x == 0) {
    if (ptx::smem_u32(smem) & 1023) {
      printf("vda flash_attn: dynamic shared memory is not 1024B aligned\n");
      __trap();
    }
    ptx::prefetch_tmap(&tm);
    ptx::mbar_init(&bars[B_QFULL], 1);
    for (int s = 0; s < NS; ++s) {
      ptx::mbar_init(&bars[B_KFULL + s], 1);
      ptx::mbar_init(&bars[B_KEMPTY + s], NSTREAM);  // one commit per stream
      ptx::mbar_init(&bars[B_VFULL + s], 1);
      ptx::mbar_init(&bars[B_VEMPTY + s], NSTREAM);
    }
    ptx::mbar_init(&bars[B_QFREE], NSTREAM);
    for (int s = 0; s < NSTREAM; ++s) {
      ptx::mbar_init(&bars[B_SFULL + s], 1);
      ptx::mbar_init(&bars[B_SFREE + s], 4);
      ptx::mbar_init(&bars[B_PFULL + s], 4);
      ptx::mbar_init(&bars[B_PVDONE + s], 1);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 1) {
    ptx::tmem_alloc(tmem_slot, TM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = ptx::uniform_u32(*tmem_slot);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer.  This loop and the MMA issuers'
    // are walked by the whole warp, the issuing instructions sit under ptx::elect_one(): the operands stay in uniform
    // registers instead of going through an ELECT + R2UR waterfall per instruction (ptx.cuh) -- here that latency sits
    // on the softmax warps' dependent chain.
    {
      if (ptx::elect_one()) {
        ptx::mbar_arrive_expect_tx(&bars[B_QFULL], TILE_BYTES);
        ptx::tma_load_2d(smem + OFF_Q, &tm, h * HD, row0 + q0, &bars[B_QFULL]);
      }
      __syncwarp();
      for (int j = 0; j < nblk; ++j) {
        const int s = j % NS;
        const uint32_t ph = (j / NS) & 1;
        ptx::mbar_wait(&bars[B_KEMPTY + s], ph ^ 1);
        if (ptx::elect_one()) {
          ptx::mbar_arrive_expect_tx(&bars[B_KFULL + s], TILE_BYTES);
          ptx::tma_load_2d(smem + OFF_K + s * TILE_BYTES, &tm, D + h * HD, row0 + j * BKV, &bars[B_KFULL + s]);
        }
        __syncwarp();
        ptx::mbar_wait(&bars[B_VEMPTY + s], ph ^ 1);
        if (ptx::elect_one()) {
          ptx::mbar_arrive_expect_tx(&bars[B_VFULL + s], TILE_BYTES);
          ptx::tma_load_2d(smem + OFF_V + s * TILE_BYTES, &tm, 2 * D + h * HD, row0 + j * BKV, &bars[B_VFULL + s]);
        }
        __syncwarp();
      }
    }
  } else if (warp <= NSTREAM) {
    // ------------------------------------------------------------------ MMA issuer of stream warp-1
    const int sid = warp - 1, hh = sid;
    {
      constexpr uint32_t idesc_s = ptx::make_idesc_bf16(BQ, KH, 0, 0);
      constexpr uint32_t idesc_pv = ptx::make_idesc_bf16(BQ, HD, 0, 1);  // B = V is MN-major (d contiguous)
      const uint64_t dq = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_Q), 16, 1024);
      const uint64_t dp = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_P + sid * TILE_BYTES), 16, 1024);
      const uint32_t ds_tmem = tmem + TM_S + sid * KH, do_tmem = tmem + TM_O + sid * HD;
      auto issue_s = [&](int j) {  // S(j) of this stream: keys [64 hh, +64) of block j are rows 64 hh.. of the K tile
        const int s = j % NS;
        ptx::mbar_wait(&bars[B_KFULL + s], (j / NS) & 1);
        if (j > 0) ptx::mbar_wait(&bars[B_SFREE + sid], (j - 1) & 1);  // the softmax warps hold S(j-1) in registers
        ptx::tc_fence_after();
        const uint64_t dk = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_K + s * TILE_BYTES + hh * (KH * 128)), 16, 1024);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < HD / 16; ++k) ptx::mma_f16_ss(ds_tmem, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
          ptx::tc_commit(&bars[B_SFULL + sid]);
          ptx::tc_commit(&bars[B_KEMPTY + s]);
          if (j == nblk - 1) ptx::tc_commit(&bars[B_QFREE]);
        }
        __syncwarp();
      };
      ptx::mbar_wait(&bars[B_QFULL], 0);
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) issue_s(j + 1);
        const int s = j % NS;
        ptx::mbar_wait(&bars[B_VFULL + s], (j / NS) & 1);
        ptx::mbar_wait(&bars[B_PFULL + sid], j & 1);  // P(j) is in smem, PV(j-1) has been read out of TMEM if rescaled
        ptx::tc_fence_after();
        const uint32_t sv = ptx::smem_u32(smem + OFF_V + s * TILE_BYTES) + hh * (KH * 128);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < KH / 16; ++k) {
            const uint64_t dv = ptx::make_sw128_desc(sv + k * 2048, 1024, 1024);
            ptx::mma_f16_ss(do_tmem, dp + 2 * k, dv, idesc_pv, (j | k) != 0);  // O accumulates in TMEM
          }
          ptx::tc_commit(&bars[B_PVDONE + sid]);
          ptx::tc_commit(&bars[B_VEMPTY + s]);
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax warps: stream (warp-3)/4, TMEM lane quadrant warp%4
    const int sid = (warp - 1 - NSTREAM) >> 2, hh = sid;
    {
      const int q = warp & 3;       // TMEM lane quadrant this warp may access
      const int r = q * 32 + lane;  // query row inside the tile == TMEM lane
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
      const uint32_t ts = trow + TM_S + sid * KH, to = trow + TM_O + sid * HD;
      const float C2 = 0.125f * 1.4426950408889634f;  // scale * log2(e)
      float m = -INFINITY, l = 0.f;  // m: reference max (RAW, unscaled bf16 scores) of the exponentials; may be stale
      uint8_t* prow = smem + OFF_P + sid * TILE_BYTES + r * 128;  // this thread's 64 keys = one swizzle atom row
      // The barrier unit answers a phase test after ~200 clk even when the phase completed long ago, and this loop is a
      // chain of dependent steps: the tests of "S(j+1) ready" and "PV(j-1) done" are issued early and consumed later.
      uint32_t s_ok = 0;
      for (int j = 0; j < nblk; ++j) {
        FA_T(0);
        if (!s_ok) ptx::mbar_wait(&bars[B_SFULL + sid], j & 1);
        ptx::tc_fence_after();
        FA_T(1);
        // ---- 64 scores of the row -> registers, rounded to bf16 and packed
        uint32_t sp[32];
#pragma unroll
        for (int c0 = 0; c0 < KH; c0 += 32) {
          float sv[32];
          ptx::tmem_ld32(ts + c0, sv);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; i += 2) sp[(c0 + i) >> 1] = pack_bf16x2(sv[i], sv[i + 1]);
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bars[B_SFREE + sid]);  // the next S MMA of this stream may start
        FA_T(2);
        const int kvalid = ntok - j * BKV - hh * KH;  // keys >= kvalid (in this thread's 64) are outside the frame
        if (kvalid < KH) {  // one select per packed pair, and one more pass only when the boundary splits a pair
          const int inside = (kvalid + 1) >> 1;  // pairs with at least one key of the frame
#pragma unroll
          for (int i = 0; i < 32; ++i) sp[i] = (i < inside) ? sp[i] : 0xFF80FF80u;
          if (kvalid & 1) {
            const int split = kvalid >> 1;
#pragma unroll
            for (int i = 0; i < 32; ++i) sp[i] = (i == split) ? ((sp[i] & 0xFFFFu) | 0xFF800000u) : sp[i];
          }
        }
        uint32_t mxa = 0xFF80FF80u, mxb = 0xFF80FF80u;  // (-inf, -inf); two chains
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          mxa = bf16x2_max(mxa, sp[i]);
          mxb = bf16x2_max(mxb, sp[i + 1]);
        }
        mxa = bf16x2_max(mxa, mxb);
        const float bm = fmaxf(bf16lo_to_f32(mxa), bf16hi_to_f32(mxa));  // block max over this stream's 64 keys
        // Lazy rescale: the reference max only moves when the block max exceeds it by more than 2^8 in the exponent
        // (probabilities then stay <= 256, harmless in bf16 / fp32); O and l are rescaled only on those blocks.
        // (bm = -inf when all 64 keys are outside the frame: nothing moves, every probability is 0.)
        const bool need = bm > -INFINITY && (bm - m) * C2 > 8.0f;  // true on the first block with a key in the frame
        const float alpha = need ? ex2_approx((m - bm) * C2) : 1.0f;
        if (need) m = bm;
        const float mb = (m > -INFINITY) ? m * C2 : 0.f;
        const uint32_t pv_ok = (j > 0) ? ptx::mbar_test(&bars[B_PVDONE + sid], (j - 1) & 1) : 1u;
        FA_T(3);
        // ---- probabilities -> bf16 (in place).  The four softmax warps of a scheduler keep its MUFU pipe ~95 % busy
        // (4 warps x 64 exp2 x 8 clk = 2048 of the ~2150 clk a block takes) and its issue port ~80 %; scale-and-shift and
        // row sum run as packed fp32 pairs (FFMA2 / FADD2: one issue slot for two).  Moving 2 / 4 / 6 of the 32 pairs' exp2
        // onto the FMA pipe (Cody-Waite + cubic, +5 issue slots per element) measured 376 / 390 / 435 us against 376:
        // both resources are at their balance point, so the MUFU stays the exponential's unit.
        uint64_t ps = pk2(0.f, 0.f);  // (sum over the even keys, sum over the odd keys), each in key order
        const uint64_t c2 = pk2(C2, C2), nmb = pk2(-mb, -mb);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const uint64_t x = fma2(pk2(bf16lo_to_f32(sp[i]), bf16hi_to_f32(sp[i])), c2, nmb);
          float x0, x1;
          upk2(x, x0, x1);
          const float p0 = ex2_approx(x0), p1 = ex2_approx(x1);
          ps = add2(ps, pk2(p0, p1));
          sp[i] = pack_bf16x2(p0, p1);
        }
        float psum0, psum1;
        upk2(ps, psum0, psum1);
        l = fmaf(l, alpha, psum0 + psum1);  // row sum over this stream's keys; the halves are merged at the end
        FA_T(4);
        // ---- PV(j-1) must be complete before P (smem) is overwritten or O (TMEM) is rescaled
        if (j > 0) {
          if (!pv_ok) ptx::mbar_wait(&bars[B_PVDONE + sid], (j - 1) & 1);
          ptx::tc_fence_after();
          if (__any_sync(0xffffffffu, need)) {
#pragma unroll 1
            for (int c0 = 0; c0 < HD; c0 += 32) {
              float ov[32];
              ptx::tmem_ld32(to + c0, ov);
              ptx::tmem_ld_wait();
#pragma unroll
              for (int i = 0; i < 32; ++i) ov[i] *= alpha;
              ptx::tmem_st32(to + c0, ov);
            }
            ptx::tmem_st_wait();
          }
        }
        FA_T(5);
        s_ok = (j + 1 < nblk) ? ptx::mbar_test(&bars[B_SFULL + sid], (j + 1) & 1) : 0u;
        // ---- P -> swizzled smem (A operand of the PV MMA)
#pragma unroll
        for (int g = 0; g < 8; ++g)
          *reinterpret_cast<uint4*>(prow + ((g ^ (r & 7)) << 4)) = make_uint4(sp[4 * g], sp[4 * g + 1], sp[4 * g + 2], sp[4 * g + 3]);
        ptx::fence_proxy_async();  // generic-proxy P writes -> visible to the tensor core (async proxy)
        ptx::tc_fence_before();    // TMEM accesses of this iteration are complete
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bars[B_PFULL + sid]);
        FA_T(6);
      }
      ptx::mbar_wait(&bars[B_PVDONE + sid], (nblk - 1) & 1);
      ptx::tc_fence_after();
      // ---- merge the two key halves of the tile: exchange (m, l) with the partner stream's warp of the same quadrant;
      // after the barrier the partner has seen its own last PV complete, so both O accumulators are final
      ptx::mbar_wait(&bars[B_QFREE], 0);  // both streams' last S MMAs have read Q: its tile is free for the exchange
      float2* xs = reinterpret_cast<float2*>(smem + OFF_X);
      xs[hh * 128 + r] = make_float2(m, l);
      ptx::tc_fence_before();
      asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
      ptx::tc_fence_after();
      const float2 pml = xs[(hh ^ 1) * 128 + r];
      const float mm = fmaxf(m, pml.x);  // finite: block 0 of key half 0 always holds a key of the frame
      const float a_own = ex2_approx((m - mm) * C2), a_par = ex2_approx((pml.x - mm) * C2);
      const float inv = 1.f / fmaf(l, a_own, pml.y * a_par);
      const float c_own = a_own * inv, c_par = a_par * inv;
      // this warp writes output columns [32 hh, +32) of its 32 rows: own and partner accumulator, normalised
      float oacc[32];
      {
        float opar[32];
        ptx::tmem_ld32(to + hh * 32, oacc);
        ptx::tmem_ld32(trow + TM_O + (sid ^ 1) * HD + hh * 32, opar);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) oacc[i] = fmaf(oacc[i], c_own, opar[i] * c_par);
      }
      // ---- round and store through this warp's (now idle) P rows: 32 rows x 64 B, row-contiguous stores
      uint8_t* wst = smem + OFF_P + sid * TILE_BYTES + q * 32 * 128;  // written only by this warp's threads
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 u;
        u.x = pack_bf16x2(oacc[8 * g], oacc[8 * g + 1]);
        u.y = pack_bf16x2(oacc[8 * g + 2], oacc[8 * g + 3]);
        u.z = pack_bf16x2(oacc[8 * g + 4], oacc[8 * g + 5]);
        u.w = pack_bf16x2(oacc[8 * g + 6], oacc[8 * g + 7]);
        *reinterpret_cast<uint4*>(wst + lane * 64 + ((g ^ ((lane >> 1) & 3)) << 4)) = u;
      }
      __syncwarp();
      const int g = lane & 3;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rr = 8 * it + (lane >> 2);
        const int qrow = q0 + q * 32 + rr;
        const uint4 u = *reinterpret_cast<const uint4*>(wst + rr * 64 + ((g ^ ((rr >> 1) & 3)) << 4));
        if (qrow < ntok) *reinterpret_cast<uint4*>(o + ((long long)row0 + qrow) * D + h * HD + hh * 32 + 8 * g) = u;
      }
    }
    ptx::tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, TM_COLS);
  }
}

}  // namespace

int vda_attention_tc(const void* qkv, void* o, int frames, int ntok, int heads, cudaStream_t st) {
  const int D = heads * HD;
  CUtensorMap tm;
  int rc = vda_make_tmap_2d(&tm, qkv, (uint64_t)3 * D, (uint64_t)frames * ntok, (uint64_t)3 * D * 2, 64, 128);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t ce = cudaFuncSetAttribute(flash_attn_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FA_SMEM);
    if (ce != cudaSuccess) {
      vda_set_error("flash_attn_tc: cudaFuncSetAttribute(%d) failed: %s", FA_SMEM, cudaGetErrorString(ce));
      return (int)ce;
    }
    attr_set = true;
  }
  dim3 grid((ntok + BQ - 1) / BQ, heads, frames);
  flash_attn_tc_kernel<<<grid, FA_THREADS, FA_SMEM, st>>>(tm, reinterpret_cast<bf16*>(o), ntok, heads);
  VDA_CUDA_LAUNCH_CHECK("flash_attn_tc_kernel");
  return 0;
}
