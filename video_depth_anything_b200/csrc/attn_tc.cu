// tcgen05 flash attention for the DINOv2 encoder (dinov2_layers/attention.py:49-59), head_dim 64, non-causal.
//
// One CTA = 128 query rows of one (frame, head); two CTAs per SM so one CTA's MMAs overlap the other's softmax.
//   warp 0   : TMA producer (Q once, then K_j / V_j tiles of 128 keys through 2-stage rings)
//   warp 1   : MMA issuer;  S = Q K_j^T (128x128x64) and PV = P_j V_j (128x64x128), accumulators in TMEM
//   warps 2-5: softmax, one query row per thread (TMEM lane == row): S from TMEM, online max/sum in fp32,
//              P written to shared memory as the bf16 A operand of the PV MMA, O kept in registers.
// Numerics follow autocast: scores are rounded to bf16 (q@k^T is a bf16 tensor), softmax runs in fp32,
// probabilities are rounded to bf16 before P@V, the output is rounded to bf16.
#include "common.cuh"
#include "ptx.cuh"
#include "tmap.h"

namespace {

constexpr int BQ = 128, BKV = 128, HD = 64;
constexpr int TILE_BYTES = 128 * 128;  // 128 rows x 64 bf16
constexpr int OFF_Q = 0, OFF_K = TILE_BYTES, OFF_V = 3 * TILE_BYTES, OFF_P = 5 * TILE_BYTES, OFF_BAR = 7 * TILE_BYTES;
constexpr int FA_SMEM = OFF_BAR + 128;
constexpr int FA_THREADS = 192;
constexpr uint32_t TM_S = 0, TM_O = 128, TM_COLS = 256;

enum { B_QFULL = 0, B_KFULL = 1, B_KEMPTY = 3, B_VFULL = 5, B_VEMPTY = 7, B_SFULL = 9, B_SFREE = 10, B_PFULL = 11, B_PVDONE = 12 };

__global__ void __launch_bounds__(FA_THREADS, 2)
flash_attn_tc_kernel(const __grid_constant__ CUtensorMap tm, bf16* __restrict__ o, int ntok, int heads) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 13);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q0 = blockIdx.x * BQ, h = blockIdx.y, f = blockIdx.z;
  const int D = heads * HD;
  const int nblk = (ntok + BKV - 1) / BKV;
  const int row0 = f * ntok;  // first row of this frame in the [frames*ntok, 3D] matrix

  if (threadIdx.x == 0) {
    if (ptx::smem_u32(smem) & 1023) {
      printf("vda flash_attn: dynamic shared memory is not 1024B aligned\n");
      __trap();
    }
    ptx::prefetch_tmap(&tm);
    ptx::mbar_init(&bars[B_QFULL], 1);
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&bars[B_KFULL + s], 1);
      ptx::mbar_init(&bars[B_KEMPTY + s], 1);
      ptx::mbar_init(&bars[B_VFULL + s], 1);
      ptx::mbar_init(&bars[B_VEMPTY + s], 1);
    }
    ptx::mbar_init(&bars[B_SFULL], 1);
    ptx::mbar_init(&bars[B_SFREE], 4);
    ptx::mbar_init(&bars[B_PFULL], 4);
    ptx::mbar_init(&bars[B_PVDONE], 1);
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
    if (lane == 0) {
      ptx::mbar_arrive_expect_tx(&bars[B_QFULL], TILE_BYTES);
      ptx::tma_load_2d(smem + OFF_Q, &tm, h * HD, row0 + q0, &bars[B_QFULL]);
      for (int j = 0; j < nblk; ++j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        ptx::mbar_wait(&bars[B_KEMPTY + s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&bars[B_KFULL + s], TILE_BYTES);
        ptx::tma_load_2d(smem + OFF_K + s * TILE_BYTES, &tm, D + h * HD, row0 + j * BKV, &bars[B_KFULL + s]);
        ptx::mbar_wait(&bars[B_VEMPTY + s], ph ^ 1);
        ptx::mbar_arrive_expect_tx(&bars[B_VFULL + s], TILE_BYTES);
        ptx::tma_load_2d(smem + OFF_V + s * TILE_BYTES, &tm, 2 * D + h * HD, row0 + j * BKV, &bars[B_VFULL + s]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = ptx::make_idesc_bf16(BQ, BKV, 0, 0);
      constexpr uint32_t idesc_pv = ptx::make_idesc_bf16(BQ, HD, 0, 1);  // B = V is MN-major (d contiguous)
      const uint32_t sq = ptx::smem_u32(smem + OFF_Q), sp = ptx::smem_u32(smem + OFF_P);
      const uint64_t dq = ptx::make_sw128_desc(sq, 16, 1024);
      auto issue_s = [&](int j) {
        const int s = j & 1;
        ptx::mbar_wait(&bars[B_KFULL + s], (j >> 1) & 1);
        ptx::tc_fence_after();
        const uint64_t dk = ptx::make_sw128_desc(ptx::smem_u32(smem + OFF_K + s * TILE_BYTES), 16, 1024);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k) ptx::mma_f16_ss(tmem + TM_S, dq + 2 * k, dk + 2 * k, idesc_s, k != 0);
        ptx::tc_commit(&bars[B_KEMPTY + s]);
        ptx::tc_commit(&bars[B_SFULL]);
      };
      ptx::mbar_wait(&bars[B_QFULL], 0);
      issue_s(0);
      for (int j = 0; j < nblk; ++j) {
        if (j + 1 < nblk) {
          ptx::mbar_wait(&bars[B_SFREE], j & 1);  // softmax has read S_j out of TMEM
          issue_s(j + 1);
        }
        const int s = j & 1;
        ptx::mbar_wait(&bars[B_VFULL + s], (j >> 1) & 1);
        ptx::mbar_wait(&bars[B_PFULL], j & 1);  // P_j is in smem, PV_{j-1} has been read out of TMEM
        ptx::tc_fence_after();
        const uint32_t sv = ptx::smem_u32(smem + OFF_V + s * TILE_BYTES);
#pragma unroll
        for (int k = 0; k < BKV / 16; ++k) {
          const uint64_t dp = ptx::make_sw128_desc(sp + (k >> 2) * TILE_BYTES, 16, 1024) + 2 * (k & 3);
          const uint64_t dv = ptx::make_sw128_desc(sv + k * 2048, 1024, 1024);
          ptx::mma_f16_ss(tmem + TM_O, dp, dv, idesc_pv, k != 0);
        }
        ptx::tc_commit(&bars[B_VEMPTY + s]);
        ptx::tc_commit(&bars[B_PVDONE]);
      }
    }
  } else {
    const int q = warp & 3;
    const int r = q * 32 + lane;  // query row inside the tile == TMEM lane
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const float LOG2E = 1.4426950408889634f;
    float oacc[HD];
#pragma unroll
    for (int c = 0; c < HD; ++c) oacc[c] = 0.f;
    float m = -INFINITY, l = 0.f, alpha_prev = 0.f;
    uint8_t* prow = smem + OFF_P + r * 128;
    for (int j = 0; j < nblk; ++j) {
      ptx::mbar_wait(&bars[B_SFULL], j & 1);
      ptx::tc_fence_after();
      const int kvalid = ntok - j * BKV;  // keys >= kvalid are outside the frame
      // pass 1: block max of the bf16-rounded scaled scores
      float bm = -INFINITY;
#pragma unroll
      for (int c0 = 0; c0 < BKV; c0 += 32) {
        float sv[32];
        ptx::tmem_ld32(trow + TM_S + c0, sv);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float s = round_bf16(sv[i] * 0.125f);
          if (c0 + i < kvalid) bm = fmaxf(bm, s);
        }
      }
      const float m_new = fmaxf(m, bm);
      const float alpha = exp2f((m - m_new) * LOG2E);
      const float mb = m_new * LOG2E;
      // fold PV_{j-1} (relative to the previous max) into the register accumulator before P / O_part are overwritten
      if (j > 0) {
        ptx::mbar_wait(&bars[B_PVDONE], (j - 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int c0 = 0; c0 < HD; c0 += 32) {
          float pv[32];
          ptx::tmem_ld32(trow + TM_O + c0, pv);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) oacc[c0 + i] = fmaf(oacc[c0 + i], alpha_prev, pv[i]);
        }
      }
      // pass 2: probabilities -> bf16 -> swizzled smem (A operand of the PV MMA)
      float psum = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < BKV; c0 += 32) {
        float sv[32];
        ptx::tmem_ld32(trow + TM_S + c0, sv);
        ptx::tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float s0 = round_bf16(sv[i] * 0.125f), s1 = round_bf16(sv[i + 1] * 0.125f);
          float p0 = (c0 + i < kvalid) ? exp2f(fmaf(s0, LOG2E, -mb)) : 0.f;
          float p1 = (c0 + i + 1 < kvalid) ? exp2f(fmaf(s1, LOG2E, -mb)) : 0.f;
          psum += p0 + p1;
          pk[i >> 1] = pack_bf16x2(p0, p1);
        }
        uint8_t* half = prow + (c0 >> 6) * TILE_BYTES;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int chunk = ((c0 & 63) >> 3) + g;
          *reinterpret_cast<uint4*>(half + ((chunk ^ (r & 7)) << 4)) =
              make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
        }
      }
      l = fmaf(l, alpha, psum);
      m = m_new;
      alpha_prev = alpha;
      ptx::fence_proxy_async();  // generic-proxy P writes -> visible to the tensor core (async proxy)
      ptx::tc_fence_before();    // all TMEM reads of S_j and PV_{j-1} are complete
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(&bars[B_SFREE]);
        ptx::mbar_arrive(&bars[B_PFULL]);
      }
    }
    ptx::mbar_wait(&bars[B_PVDONE], (nblk - 1) & 1);
    ptx::tc_fence_after();
#pragma unroll
    for (int c0 = 0; c0 < HD; c0 += 32) {
      float pv[32];
      ptx::tmem_ld32(trow + TM_O + c0, pv);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i) oacc[c0 + i] = fmaf(oacc[c0 + i], alpha_prev, pv[i]);
    }
    if (q0 + r < ntok) {
      const float inv = 1.f / l;
      bf16* op = o + ((long long)row0 + q0 + r) * D + h * HD;
#pragma unroll
      for (int c = 0; c < HD; c += 8) {
        float t[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) t[i] = oacc[c + i] * inv;
        Vec8<bf16>::store(op + c, t);
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
