// tcgen05 flash attention for the DINOv2 encoder (dinov2_layers/attention.py:49-59), head_dim 64, non-causal.
//
// One CTA = up to 256 query rows (two 128-row tiles) of one (frame, head); one CTA per SM.  The two tiles ping-pong:
// while softmax group 0 works on its S tile the tensor core computes the other tile's S / PV.
//   warp 0     : TMA producer (Q tiles once, then K_j / V_j tiles of 128 keys through 3-stage rings)
//   warp 1     : MMA issuer;  S_t = Q_t K_j^T (128x128x64) and PV_t = P_t V_j (128x64x128), accumulators in TMEM
//   warps 2-9  : softmax group 0 (tile 0): one query row x 64 keys per thread (TMEM lane == row; two warps per lane
//                quadrant split the 128 keys of a block and exchange the block max through shared memory)
//   warps 10-17: softmax group 1 (tile 1)
// A softmax thread reads its S row out of TMEM ONCE, rounding to bf16 and packing as it goes (64 registers), releases
// the TMEM tile immediately (the next S MMA overlaps the exponentials), then: online max / sum in fp32, P written to
// shared memory as the bf16 A operand of the PV MMA, O accumulated in TMEM across key blocks (rescaled lazily, only
// when the running max moves by more than 2^8).
// Tried and dropped: evaluating half of the exponentials with a Cody-Waite + cubic polynomial on the FMA pipe (the
// FlashAttention-4 trick) made the kernel 12 % slower -- it is instruction-issue bound (IPC 0.44 per scheduler with
// ~7 instructions per score), not MUFU bound (XU pipe 44 %).
// Numerics follow autocast: scores are rounded to bf16 (q@k^T is a bf16 tensor; the 2^-3 scale commutes with the
// rounding), softmax runs in fp32, probabilities are rounded to bf16 before P@V, the output is rounded to bf16.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"

namespace {

constexpr int BQ = 128, BKV = 128, HD = 64;
constexpr int NS = 3;                  // K / V ring depth
constexpr int TILE_BYTES = 128 * 128;  // 128 rows x 64 bf16
constexpr int OFF_Q = 0;                         // 2 tiles
constexpr int OFF_K = 2 * TILE_BYTES;            // NS tiles
constexpr int OFF_V = OFF_K + NS * TILE_BYTES;   // NS tiles
constexpr int OFF_P = OFF_V + NS * TILE_BYTES;   // 2 x (2 atoms of 128 rows x 64 keys)
constexpr int OFF_BAR = OFF_P + 4 * TILE_BYTES;
constexpr int OFF_X = OFF_BAR + 256;             // block-max / row-sum exchange between partner warps: 6 KB
constexpr int FA_SMEM = OFF_X + 6 * 1024;
constexpr int FA_THREADS = 64 + 512;
constexpr uint32_t TM_S = 0, TM_O = 256, TM_COLS = 512;  // S_t at TM_S + 128 t, O_t at TM_O + 64 t

// barrier indices
constexpr int B_QFULL = 0;               // 1
constexpr int B_KFULL = 1;               // NS
constexpr int B_KEMPTY = B_KFULL + NS;   // NS
constexpr int B_VFULL = B_KEMPTY + NS;   // NS
constexpr int B_VEMPTY = B_VFULL + NS;   // NS
constexpr int B_SFULL = B_VEMPTY + NS;   // 2
constexpr int B_SFREE = B_SFULL + 2;     // 2
constexpr int B_PFULL = B_SFREE + 2;     // 2
constexpr int B_PVDONE = B_PFULL + 2;    // 2
constexpr int NBARS = B_PVDONE + 2;
static_assert(NBARS * 8 + 8 <= 256, "barrier area");
static_assert(FA_SMEM <= 232448, "shared memory budget");

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

__global__ void __launch_bounds__(FA_THREADS, 1)
flash_attn_tc_kernel(const __grid_constant__ CUtensorMap tm, bf16* __restrict__ o, int ntok, int heads) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + NBARS);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * 2 * BQ, h = blockIdx.y, f = blockIdx.z;
  const int D = heads * HD;
  const int nblk = (ntok + BKV - 1) / BKV;
  const int ntiles = (q0 + BQ < ntok) ? 2 : 1;  // the last CTA of a frame may own a single tile
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
      ptx::mbar_init(&bars[B_KEMPTY + s], 1);
      ptx::mbar_init(&bars[B_VFULL + s], 1);
      ptx::mbar_init(&bars[B_VEMPTY + s], 1);
    }
    for (int t = 0; t < 2; ++t) {
      ptx::mbar_init(&bars[B_SFULL + t], 1);
      ptx::mbar_init(&bars[B_SFREE + t], 8);
      ptx::mbar_init(&bars[B_PFULL + t], 8);
      ptx::mbar_init(&bars[B_PVDONE + t], 1);
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
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      ptx::mbar_arrive_expect_tx(&bars[B_QFULL], ntiles * TILE_BYTES);
      for (int t = 0; t < ntiles; ++t)
        ptx::tma_load_2d(smem + OFF_Q + t * TILE_BYTES, &tm, h * HD, row0 + q0 + t * BQ, &bars[B_QFULL]);
      for (int j = 0; j < nblk; ++j) {
        const int s = j % NS;
        const uint32_t ph = (j / NS) & 1;
        ptx::mbar_wait(&bars[B_KEMPTY + s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&bars[B_KFULL + s], TILE_BYTES);
        ptx::tma_load_2d(smem + OFF_K + s * TILE_BYTES, &tm, D + h * HD, row0 + j * BKV, &bars[B_KFULL + s]);
        ptx::mbar_wait(&bars[B_VEMPTY + s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&bars[B_VFULL + s], TILE_BYTES);
        ptx::tma_load_2d(smem + OFF_V + s * TILE_BYTES, &tm, 2 * D + h * HD, row0 + j * BKV, &bars[B_VFULL + s]);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_s = ptx::make_idesc_bf16(BQ, BKV, 0, 0);
      constexpr uint32_t idesc_pv = ptx::make_idesc_bf16(BQ, HD, 0, 1);  // B = V is MN-major (d contiguous)
      auto issue_s = [&](int j) {  // S_t(j) for every tile; frees the K stage when done
        const int s = j % NS;
        ptx::mbar_wait(&bars[B_KFULL + s], (j / NS) & 1);
        const uint64_t dk = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_K + s * TILE_BYTES), 16, 1024);
        for (int t = 0; t < ntiles; ++t) {
          if (j > 0) ptx::mbar_wait(&bars[B_SFREE + t], (j - 1) & 1);  // softmax t holds S_t(j-1) in registers
          ptx::tc_fence_after();
          const uint64_t dq = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_Q + t * TILE_BYTES), 16, 1024);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            ptx::mma_f16_ss(tmem + TM_S + t * BKV, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
          ptx::tc_commit(&bars[B_SFULL + t]);
        }
        ptx::tc_commit(&bars[B_KEMPTY + s]);
      };
      ptx::mbar_wait(&bars[B_QFULL], 0);
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) issue_s(j + 1);
        const int s = j % NS;
        ptx::mbar_wait(&bars[B_VFULL + s], (j / NS) & 1);
        const uint32_t sv = ptx::smem_u32(smem + OFF_V + s * TILE_BYTES);
        for (int t = 0; t < ntiles; ++t) {
          ptx::mbar_wait(&bars[B_PFULL + t], j & 1);  // P_t(j) is in smem, PV_t(j-1) has been read out of TMEM
          ptx::tc_fence_after();
          const uint32_t sp = ptx::smem_u32(smem + OFF_P + t * 2 * TILE_BYTES);
#pragma unroll
          for (int k = 0; k < BKV / 16; ++k) {
            const uint64_t dp = ptx::make_sw128_desc(sp + (k >> 2) * TILE_BYTES, 16, 1024) + 2 * (k & 3);
            const uint64_t dv = ptx::make_sw128_desc(sv + k * 2048, 1024, 1024);
            ptx::mma_f16_ss(tmem + TM_O + t * HD, dp, dv, idesc_pv, (j | k) != 0);  // O accumulates in TMEM
          }
          ptx::tc_commit(&bars[B_PVDONE + t]);
        }
        ptx::tc_commit(&bars[B_VEMPTY + s]);
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax groups
    // 8 warps per tile: two warps share each TMEM lane quadrant (warp ids 4 apart) and split the 128 keys of a block,
    // so a thread owns one query row x 64 keys.  Four warps per scheduler instead of two hide the MUFU / TMEM latencies.
    const int sw = warp - 2;        // 0..15
    const int t = sw >> 3;          // tile / group
    if (t < ntiles) {
      const int q = warp & 3;       // TMEM lane quadrant of this warp
      const int hc = (sw >> 2) & 1; // which 64 keys of the block
      const int r = q * 32 + lane;  // query row inside the tile == TMEM lane
      const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
      const uint32_t ts = trow + TM_S + t * BKV + hc * 64, to = trow + TM_O + t * HD + hc * 32;
      const float C2 = 0.125f * 1.4426950408889634f;  // scale * log2(e)
      float m = -INFINITY, l = 0.f;  // m: reference max (RAW, unscaled bf16 scores) of the exponentials; may be stale
      uint8_t* prow = smem + OFF_P + (t * 2 + hc) * TILE_BYTES + r * 128;  // this thread's 64 keys = one swizzle atom row
      float* xch = reinterpret_cast<float*>(smem + OFF_X);                 // [2 parities][2 tiles][2 halves][128 rows]
      const int pair_bar = 1 + t * 4 + q;                                  // named barrier of the two partner warps
      for (int j = 0; j < nblk; ++j) {
        ptx::mbar_wait(&bars[B_SFULL + t], j & 1);
        ptx::tc_fence_after();
        // ---- 64 scores of the row -> registers, rounded to bf16 and packed
        uint32_t sp[32];
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 32) {
          float sv[32];
          ptx::tmem_ld32(ts + c0, sv);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; i += 2) sp[(c0 + i) >> 1] = pack_bf16x2(sv[i], sv[i + 1]);
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bars[B_SFREE + t]);  // the next S MMA of this tile may start
        const int kvalid = ntok - j * BKV - hc * 64;  // keys >= kvalid (in this thread's 64) are outside the frame
        if (kvalid < 64) {
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            if (2 * i >= kvalid) sp[i] = 0xFF80FF80u;
            else if (2 * i + 1 >= kvalid) sp[i] = (sp[i] & 0xFFFFu) | 0xFF800000u;
          }
        }
        uint32_t mxa = 0xFF80FF80u, mxb = 0xFF80FF80u;  // (-inf, -inf); two chains
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          mxa = bf16x2_max(mxa, sp[i]);
          mxb = bf16x2_max(mxb, sp[i + 1]);
        }
        mxa = bf16x2_max(mxa, mxb);
        float bm = fmaxf(bf16lo_to_f32(mxa), bf16hi_to_f32(mxa));
        // block max of the whole row: exchange with the partner warp (double-buffered by block parity: one barrier)
        float* xm = xch + (((j & 1) * 2 + t) * 2) * 128;
        xm[hc * 128 + r] = bm;
        asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
        bm = fmaxf(bm, xm[(hc ^ 1) * 128 + r]);
        // Lazy rescale: the reference max only moves when the block max exceeds it by more than 2^8 in the exponent
        // (probabilities then stay <= 256, harmless in bf16 / fp32); O and l are rescaled only on those blocks.
        const bool need = (bm - m) * C2 > 8.0f;  // always true on the first block (m = -inf); identical in both partners
        const float alpha = need ? ex2_approx((m - bm) * C2) : 1.0f;
        if (need) m = bm;
        const float mb = m * C2;
        // ---- probabilities -> bf16 (in place)
        float psum0 = 0.f, psum1 = 0.f;
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float p0 = ex2_approx(fmaf(bf16lo_to_f32(sp[i]), C2, -mb));
          const float p1 = ex2_approx(fmaf(bf16hi_to_f32(sp[i]), C2, -mb));
          psum0 += p0;
          psum1 += p1;
          sp[i] = pack_bf16x2(p0, p1);
        }
        l = fmaf(l, alpha, psum0 + psum1);  // partial row sum (this thread's keys); the halves are added at the end
        // ---- PV_{j-1} must be complete before P (smem) is overwritten or O (TMEM) is rescaled
        if (j > 0) {
          ptx::mbar_wait(&bars[B_PVDONE + t], (j - 1) & 1);
          ptx::tc_fence_after();
          if (__any_sync(0xffffffffu, need)) {  // this thread rescales its 32 of the row's 64 output columns
            float ov[32];
            ptx::tmem_ld32(to, ov);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) ov[i] *= alpha;
            ptx::tmem_st32(to, ov);
            ptx::tmem_st_wait();
          }
        }
        // ---- P -> swizzled smem (A operand of the PV MMA)
#pragma unroll
        for (int g = 0; g < 8; ++g)
          *reinterpret_cast<uint4*>(prow + ((g ^ (r & 7)) << 4)) = make_uint4(sp[4 * g], sp[4 * g + 1], sp[4 * g + 2], sp[4 * g + 3]);
        ptx::fence_proxy_async();  // generic-proxy P writes -> visible to the tensor core (async proxy)
        ptx::tc_fence_before();    // TMEM accesses of this iteration are complete
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bars[B_PFULL + t]);
      }
      ptx::mbar_wait(&bars[B_PVDONE + t], (nblk - 1) & 1);
      ptx::tc_fence_after();
      float oacc[32];  // this thread's 32 output columns
      ptx::tmem_ld32(to, oacc);
      ptx::tmem_ld_wait();
      float* xs = xch + 4 * 2 * 128 + (t * 2) * 128;  // row sums of the two halves
      xs[hc * 128 + r] = l;
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      const float inv = 1.f / (l + xs[(hc ^ 1) * 128 + r]);
      // ---- normalise, round, and store through this warp's (now idle) P rows: 32 rows x 64 B, row-contiguous stores
      uint8_t* wst = smem + OFF_P + (t * 2 + hc) * TILE_BYTES + q * 32 * 128;  // written only by this warp's threads
#pragma unroll
      for (int g = 0; g < 4; ++g) {
        uint4 u;
        u.x = pack_bf16x2(oacc[8 * g] * inv, oacc[8 * g + 1] * inv);
        u.y = pack_bf16x2(oacc[8 * g + 2] * inv, oacc[8 * g + 3] * inv);
        u.z = pack_bf16x2(oacc[8 * g + 4] * inv, oacc[8 * g + 5] * inv);
        u.w = pack_bf16x2(oacc[8 * g + 6] * inv, oacc[8 * g + 7] * inv);
        *reinterpret_cast<uint4*>(wst + lane * 64 + ((g ^ ((lane >> 1) & 3)) << 4)) = u;
      }
      __syncwarp();
      const int g = lane & 3;
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int rr = 8 * it + (lane >> 2);
        const int qrow = q0 + t * BQ + q * 32 + rr;
        const uint4 u = *reinterpret_cast<const uint4*>(wst + rr * 64 + ((g ^ ((rr >> 1) & 3)) << 4));
        if (qrow < ntok) *reinterpret_cast<uint4*>(o + ((long long)row0 + qrow) * D + h * HD + hc * 32 + 8 * g) = u;
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
  dim3 grid((ntok + 2 * BQ - 1) / (2 * BQ), heads, frames);
  flash_attn_tc_kernel<<<grid, FA_THREADS, FA_SMEM, st>>>(tm, reinterpret_cast<bf16*>(o), ntok, heads);
  VDA_CUDA_LAUNCH_CHECK("flash_attn_tc_kernel");
  return 0;
}
