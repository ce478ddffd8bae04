// Coalesced epilogue of the tcgen05 GEMM (bf16 build).  One call handles a chunk of 64 OUTPUT columns for the 32 rows
// (TMEM lanes) owned by one epilogue warp:
//   phase 1 (thread == row): tcgen05.ld the accumulators, add the bias (smem tile, broadcast reads), apply the
//            activation, round to bf16 (every value is a bf16 number at this point under autocast, so staging in
//            bf16 is lossless) and park the 64 values in a 128B-swizzled per-warp smem tile;
//   phase 2 (8 lanes == one row segment of 64 columns, 4 rows per instruction): read the tile back transposed and do
//            all global traffic -- residual stream read-modify-write, skip-connection reads, output stores -- as
//            16/32-byte per-lane accesses that cover whole 128-byte lines.
// The semantic twin is epi_chunk() in epilogue.cuh (used by the SIMT build and as the fallback for ragged chunks);
// any change to rounding points must be made in both.
#pragma once
#include "epilogue.cuh"
#include "ptx.cuh"

// Epilogue kinds: the common shapes are compiled as separate kernels so that their inner loops carry no run-time
// dispatch (the fully generic body is ~20k SASS instructions and thrashes the instruction cache).
enum EpiKind {
  EK_STORE = 0,    // out = R(acc + bias)                                  (qkv, 1x1 projections, plain convs)
  EK_GELU = 1,     // out = R(gelu(R(acc + bias)))                         (mlp.fc1)
  EK_RESID = 2,    // resid += gamma * R(acc + bias), nothing else stored  (attn.proj, mlp.fc2)
  EK_GEGLU = 3,    // out = R(value * R(gelu(gate)))                       (motion-module feed-forward)
  EK_GENERIC = 4,  // everything, decided at run time
  EK_RES = 5       // out = R(R(acc + bias) + res)                         (motion-module out / ff / proj_out skips)
};

static inline int epi_kind_of(const vda_gemm_desc* d) {
  const bool plain = d->res == nullptr && d->res2 == nullptr && d->out_relu == nullptr && d->head_w == nullptr;
  if (plain && d->resid == nullptr && d->out != nullptr) {
    if (d->act == VDA_ACT_NONE) return EK_STORE;
    if (d->act == VDA_ACT_GELU) return EK_GELU;
    if (d->act == VDA_ACT_GEGLU) return EK_GEGLU;
  }
  if (plain && d->resid != nullptr && !d->resid_assign && d->out == nullptr && d->act == VDA_ACT_NONE) return EK_RESID;
  if (d->res != nullptr && d->res2 == nullptr && d->out_relu == nullptr && d->head_w == nullptr && d->resid == nullptr &&
      d->out != nullptr && d->act == VDA_ACT_NONE && d->out_map != VDA_MAP_CONVT)
    return EK_RES;
  return EK_GENERIC;
}

struct RowIdx {
  int out_row, res_row, resid_row;  // -1 = this row produces no output
};

__device__ __forceinline__ RowIdx make_row_idx(const Epi& e, int m) {
  RowCtx r = make_row_ctx(e, m);
  RowIdx x;
  x.res_row = (int)r.res_row;
  x.resid_row = (int)r.resid_row;
  if (e.out_map == VDA_MAP_CONVT)  // padded row of (b, y*k, x*k); the (ky, kx) of the column block is added per chunk
    x.out_row = (int)(((long long)r.b * (e.gh * e.ct_k + 2) + r.y * e.ct_k + 1) * (e.gw * e.ct_k + 2) + r.x * e.ct_k + 1);
  else
    x.out_row = (int)r.out_row;
  if (!r.valid) x.out_row = -1;
  return x;
}

// Is the chunk of 64 output columns starting at accumulator column n (global) eligible for the coalesced path?
__device__ __forceinline__ bool epi_fast_ok(const Epi& e, int n) {
  const bool geglu = e.act == VDA_ACT_GEGLU;
  const int ncols = geglu ? (e.N >> 1) : e.N;
  const int c0 = geglu ? (n >> 1) : n;
  if (e.head_w != nullptr || c0 >= ncols) return false;
  // a ragged last chunk (N % 64 != 0, e.g. the 32-channel output_conv1 of vits) stays on this path when it is a plain
  // store: lanes whose 8 columns lie beyond N are masked (the accumulator columns there are zero: OOB weight rows)
  if (c0 + 64 > ncols && ((ncols & 7) || geglu || e.res != nullptr || e.res2 != nullptr || e.resid != nullptr)) return false;
  if ((e.ldo & 7) || (e.ldres & 7) || (e.ldresid & 3)) return false;
  if (e.out_map == VDA_MAP_CONVT && (e.ct_stride & 63)) return false;
  return true;
}

__device__ __forceinline__ uint32_t relu_bf16x2(uint32_t a) {
  uint32_t r;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(0u));
  return r;
}

// stage: this warp's 4 KB tile (1024B aligned); sbias: bias values of the CTA tile's accumulator columns, indexed by
// the tile-local accumulator column; tcol: TMEM address of (this warp's lane quadrant, tile-local accumulator column c).
template <int EK, bool GEGLU>
__device__ __forceinline__ void epi_chunk_tc(const Epi& e, const RowIdx& ri, int n0, int c, uint32_t tcol, uint8_t* stage,
                                             const float* sbias, int lane) {
  static_assert((EK == EK_GEGLU) == GEGLU || EK == EK_GENERIC, "GEGLU flag must match the kind");
  constexpr bool RESK = EK == EK_RES;
  constexpr bool GEN = EK == EK_GENERIC;
  const int act = GEN ? e.act : (EK == EK_GELU ? (int)VDA_ACT_GELU : (int)VDA_ACT_NONE);
  const bool has_resid = GEN ? (e.resid != nullptr) : (EK == EK_RESID);
  const bool has_out = GEN ? (e.out != nullptr) : (EK != EK_RESID);
  // skip connections / relu copy only exist on the generic, non-GEGLU kind
  const bool has_res = GEN && !GEGLU && e.res != nullptr, has_res2 = GEN && !GEGLU && e.res2 != nullptr;
  const bool has_relu = GEN && e.out_relu != nullptr;
  const bool assign = GEN && e.resid_assign;
  uint32_t pk[32];
  // phase-2 coordinates of this lane: 8 output columns, rows 4 t + lane / 8
  const int j = lane & 7;
  const int col = (GEGLU ? ((n0 + c) >> 1) : (n0 + c)) + 8 * j;  // first of this lane's 8 output columns
  int ocol = col;
  int row_add = 0;
  bool col_ok = GEGLU || col < e.N;  // false: this lane's 8 columns are padding (ragged last chunk / conv-transpose)
  if (e.out_map == VDA_MAP_CONVT) {
    const int kk = col / e.ct_stride;
    ocol = col - kk * e.ct_stride;
    col_ok = col_ok && ocol < e.ct_cout;
    row_add = (kk / e.ct_k) * (e.gw * e.ct_k + 2) + (kk % e.ct_k);
  }
  // EK_RES: the skip operand in phase-2 coordinates (8 lanes cover the 128 B of a row segment), requested before the
  // accumulator read so the round trip hides behind phase 1.  Only this lean kind does it: with the run-time flags of
  // the generic kind in the same function ptxas spilled the prefetched values, which serialised the loads.
  uint4 rv[RESK ? 8 : 1];
  if constexpr (RESK) {
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int rr = 4 * t + (lane >> 3);
      const int orow = __shfl_sync(0xffffffffu, ri.out_row, rr);
      const int rrow = __shfl_sync(0xffffffffu, ri.res_row, rr);
      rv[t] = make_uint4(0, 0, 0, 0);
      if (orow >= 0) rv[t] = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.res) + (long long)rrow * e.ldres + col);
    }
  }
  // ---------------------------------------------------------------- phase 1
  constexpr bool LEAN = (EK == EK_STORE || EK == EK_GELU) && !GEGLU;
  if constexpr (LEAN) {
    // Both 32-column halves are requested from TMEM before the first wait: one round trip per chunk instead of two (the
    // GELU kind's epilogue is its GEMM's critical path; registers are not short here).
    float v[2][32];
    ptx::tmem_ld32(tcol, v[0]);
    ptx::tmem_ld32(tcol + 32, v[1]);
    ptx::tmem_ld_wait();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const float* bv = sbias + c + h * 32;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bv + i);
        v[h][i] += b4.x; v[h][i + 1] += b4.y; v[h][i + 2] += b4.z; v[h][i + 3] += b4.w;
      }
      if constexpr (EK == EK_GELU) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {  // fc1's output is a bf16 tensor: round (paired convert), then GELU
          const uint32_t r = pack_bf16x2(v[h][i], v[h][i + 1]);
          gelu_fast2(bf16lo_to_f32(r), bf16hi_to_f32(r), v[h][i], v[h][i + 1]);
        }
      }
#pragma unroll
      for (int i = 0; i < 32; i += 2) pk[h * 16 + (i >> 1)] = pack_bf16x2(v[h][i], v[h][i + 1]);
    }
  }
#pragma unroll
  for (int h = 0; h < (LEAN ? 0 : 2); ++h) {
    float v[32];
    if constexpr (GEGLU) {
      float g[32];
      ptx::tmem_ld32(tcol + h * 64, v);
      ptx::tmem_ld32(tcol + h * 64 + 32, g);
      ptx::tmem_ld_wait();
      const float* bv = sbias + c + h * 64;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bv + i), g4 = *reinterpret_cast<const float4*>(bv + 32 + i);
        const float bb[4] = {b4.x, b4.y, b4.z, b4.w}, gb[4] = {g4.x, g4.y, g4.z, g4.w};
        // roundings to bf16 go through the paired convert (one F2FP per two values) instead of one F2F each
#pragma unroll
        for (int k = 0; k < 4; k += 2) {
          const uint32_t vr = pack_bf16x2(v[i + k] + bb[k], v[i + k + 1] + bb[k + 1]);
          const uint32_t gr = pack_bf16x2(g[i + k] + gb[k], g[i + k + 1] + gb[k + 1]);
          float g0, g1;
          gelu_fast2(bf16lo_to_f32(gr), bf16hi_to_f32(gr), g0, g1);
          const uint32_t ar = pack_bf16x2(g0, g1);
          v[i + k] = bf16lo_to_f32(vr) * bf16lo_to_f32(ar);
          v[i + k + 1] = bf16hi_to_f32(vr) * bf16hi_to_f32(ar);
        }
      }
    } else {
      ptx::tmem_ld32(tcol + h * 32, v);
      // Skip connections (GENERIC kind): this lane's row, the 32 columns of this half = 64 contiguous bytes = two full
      // sectors per lane, requested before the accumulator wait so the latency hides behind it.  (Fetching them
      // row-contiguously for phase 2 instead needed 32-64 more live registers: ptxas spilled the prefetched values,
      // which serialised the loads -- the motion-module out-projections ran at 200 us against 85 us for a plain store.)
      uint4 r1[(GEN && !GEGLU) ? 4 : 1], r2[(GEN && !GEGLU) ? 4 : 1];
      if constexpr (GEN && !GEGLU) {
        if (has_res && ri.out_row >= 0) {
          const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.res) + (long long)ri.res_row * e.ldres + n0 + c + h * 32);
#pragma unroll
          for (int k = 0; k < 4; ++k) r1[k] = rp[k];
        }
        if (has_res2 && ri.out_row >= 0) {
          const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.res2) + (long long)ri.res_row * e.ldres + n0 + c + h * 32);
#pragma unroll
          for (int k = 0; k < 4; ++k) r2[k] = rp[k];
        }
      }
      ptx::tmem_ld_wait();
      const float* bv = sbias + c + h * 32;
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bv + i);
        v[i] += b4.x; v[i + 1] += b4.y; v[i + 2] += b4.z; v[i + 3] += b4.w;
      }
      // the activation switch stays outside the element loops: one compact loop runs, the others are never fetched
      if (act == VDA_ACT_GELU) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {  // fc1's output is a bf16 tensor: round (paired convert), then GELU
          const uint32_t r = pack_bf16x2(v[i], v[i + 1]);
          gelu_fast2(bf16lo_to_f32(r), bf16hi_to_f32(r), v[i], v[i + 1]);
        }
      } else if (act == VDA_ACT_RELU) {
#pragma unroll
        for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
      }
      if constexpr (GEN && !GEGLU) {
        if (has_res || has_res2) {  // R(R(R(conv) + x) + x2): every bf16 add of the reference rounds (blocks.py:91,155)
          const uint32_t* w1 = reinterpret_cast<const uint32_t*>(r1);
          const uint32_t* w2 = reinterpret_cast<const uint32_t*>(r2);
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            uint32_t u = pack_bf16x2(v[i], v[i + 1]);
            if (has_res) u = pack_bf16x2(bf16lo_to_f32(u) + bf16lo_to_f32(w1[i >> 1]), bf16hi_to_f32(u) + bf16hi_to_f32(w1[i >> 1]));
            if (has_res2) u = pack_bf16x2(bf16lo_to_f32(u) + bf16lo_to_f32(w2[i >> 1]), bf16hi_to_f32(u) + bf16hi_to_f32(w2[i >> 1]));
            pk[h * 16 + (i >> 1)] = u;
          }
          continue;
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 32; i += 2) pk[h * 16 + (i >> 1)] = pack_bf16x2(v[i], v[i + 1]);
  }
  {
    uint8_t* rowp = stage + lane * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4)) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
  }
  __syncwarp();
  // ---------------------------------------------------------------- phase 2
  if constexpr (EK == EK_RESID) {
    // resid += gamma * v as fire-and-forget vector reductions (the L2 performs the single fp32 add: the same bits as
    // load-add-store, denormals aside: REDG flushes them).  A lane takes FOUR columns, so the sixteen lanes of a row
    // cover its 256 B contiguously and every 32 B sector is reduced by one request: with eight columns per lane and two
    // reductions each, every sector was hit twice, the L2 reduction rate (not HBM) bounded the kernel and `proj`
    // (K = 1024) could not hide it (123 us against 79 us with a plain store; 104 us now).  A prefetched
    // load - add - store of the same tile was slower still (136 us).
    const int cj = lane & 15;
    const int col4 = n0 + c + 4 * cj;
    const float4 g4 = *reinterpret_cast<const float4*>(e.gamma + col4);
#pragma unroll
    for (int t = 0; t < 16; ++t) {
      const int rr = 2 * t + (lane >> 4);
      const uint2 p = *reinterpret_cast<const uint2*>(stage + rr * 128 + (((cj >> 1) ^ (rr & 7)) << 4) + ((cj & 1) << 3));
      const int orow = __shfl_sync(0xffffffffu, ri.out_row, rr);
      const int drow = __shfl_sync(0xffffffffu, ri.resid_row, rr);
      if (orow < 0) continue;
      float* rp = e.resid + (long long)drow * e.ldresid + col4;
      asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(rp), "f"(__fmul_rn(bf16lo_to_f32(p.x), g4.x)),
                   "f"(__fmul_rn(bf16hi_to_f32(p.x), g4.y)), "f"(__fmul_rn(bf16lo_to_f32(p.y), g4.z)),
                   "f"(__fmul_rn(bf16hi_to_f32(p.y), g4.w))
                   : "memory");
    }
    __syncwarp();  // the tile is rewritten by the next chunk
    return;
  }
  if constexpr (EK == EK_STORE || EK == EK_GELU || EK == EK_GEGLU) {
    // Nothing to add: the staged bf16 is the output.  All eight row-index shuffles and all eight 16-byte shared-memory
    // loads go out before the first store, and the stores are predicated instead of branched around: the loop form
    // (shuffle -> compare -> branch -> load -> address -> store, eight times in a row) was a chain of ~100 clk per
    // iteration and 37 % of the GELU kind's epilogue, which is that GEMM's critical path (profiles/ncu_gemm_gelu8_r02c.txt).
    uint4 pp[8];
    int orow[8];
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int rr = 4 * t + (lane >> 3);
      pp[t] = *reinterpret_cast<const uint4*>(stage + rr * 128 + ((j ^ (rr & 7)) << 4));
      orow[t] = __shfl_sync(0xffffffffu, ri.out_row, rr);
    }
    bf16* obase = reinterpret_cast<bf16*>(e.out) + ocol;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      if (orow[t] >= 0 && col_ok) *reinterpret_cast<uint4*>(obase + (long long)(orow[t] + row_add) * e.ldo) = pp[t];
    }
    __syncwarp();  // the tile is rewritten by the next chunk
    return;
  }
  float gam[8];
  if (has_resid && !assign) {
    const float4 g0 = *reinterpret_cast<const float4*>(e.gamma + col), g1 = *reinterpret_cast<const float4*>(e.gamma + col + 4);
    gam[0] = g0.x; gam[1] = g0.y; gam[2] = g0.z; gam[3] = g0.w; gam[4] = g1.x; gam[5] = g1.y; gam[6] = g1.z; gam[7] = g1.w;
  }
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const int rr = 4 * t + (lane >> 3);
    const uint4 p = *reinterpret_cast<const uint4*>(stage + rr * 128 + ((j ^ (rr & 7)) << 4));
    const int orow = __shfl_sync(0xffffffffu, ri.out_row, rr);
    const int drow = __shfl_sync(0xffffffffu, ri.resid_row, rr);
    if (orow < 0 || !col_ok) continue;
    if constexpr (RESK) {  // bf16 add of the reference: R(R(acc + bias) + skip)
      const uint32_t* w = reinterpret_cast<const uint32_t*>(&rv[t]);
      const uint32_t* q = reinterpret_cast<const uint32_t*>(&p);
      uint4 o;
      uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
      for (int k = 0; k < 4; ++k)
        ow[k] = pack_bf16x2(bf16lo_to_f32(q[k]) + bf16lo_to_f32(w[k]), bf16hi_to_f32(q[k]) + bf16hi_to_f32(w[k]));
      *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out) + (long long)(orow + row_add) * e.ldo + ocol) = o;
      continue;
    }
    if constexpr (EK == EK_STORE || EK == EK_GELU || EK == EK_GEGLU) {  // nothing to add: the staged bf16 is the output
      *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out) + (long long)(orow + row_add) * e.ldo + ocol) = p;
      continue;
    }
    if (!has_resid) {  // skip connections were added in phase 1: the staged bf16 is the output (+ its ReLU copy)
      const long long oo = (long long)(orow + row_add) * e.ldo + ocol;
      *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out) + oo) = p;
      if (has_relu) *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out_relu) + oo) = make_uint4(relu_bf16x2(p.x), relu_bf16x2(p.y), relu_bf16x2(p.z), relu_bf16x2(p.w));
      continue;
    }
    float v[8] = {bf16lo_to_f32(p.x), bf16hi_to_f32(p.x), bf16lo_to_f32(p.y), bf16hi_to_f32(p.y),
                  bf16lo_to_f32(p.z), bf16hi_to_f32(p.z), bf16lo_to_f32(p.w), bf16hi_to_f32(p.w)};
    if (has_resid) {
      float* rp = e.resid + (long long)drow * e.ldresid + col;
      float4 x0, x1;
      if (assign) {
        const float* pp = e.pos + (long long)(drow % e.pos_rows) * e.ldresid + col;
        x0 = *reinterpret_cast<const float4*>(pp);
        x1 = *reinterpret_cast<const float4*>(pp + 4);
        x0.x = __fadd_rn(v[0], x0.x); x0.y = __fadd_rn(v[1], x0.y); x0.z = __fadd_rn(v[2], x0.z); x0.w = __fadd_rn(v[3], x0.w);
        x1.x = __fadd_rn(v[4], x1.x); x1.y = __fadd_rn(v[5], x1.y); x1.z = __fadd_rn(v[6], x1.z); x1.w = __fadd_rn(v[7], x1.w);
      } else {
        // resid += gamma * v as a fire-and-forget vector reduction: the L2 performs the single fp32 add (round to
        // nearest, one adder per element: the same bits as load-add-store, denormals aside: REDG flushes them) and the warp never waits
        // for the old value to travel to the SM.
        x0.x = __fmul_rn(v[0], gam[0]); x0.y = __fmul_rn(v[1], gam[1]); x0.z = __fmul_rn(v[2], gam[2]); x0.w = __fmul_rn(v[3], gam[3]);
        x1.x = __fmul_rn(v[4], gam[4]); x1.y = __fmul_rn(v[5], gam[5]); x1.z = __fmul_rn(v[6], gam[6]); x1.w = __fmul_rn(v[7], gam[7]);
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(rp), "f"(x0.x), "f"(x0.y), "f"(x0.z), "f"(x0.w) : "memory");
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(rp + 4), "f"(x1.x), "f"(x1.y), "f"(x1.z), "f"(x1.w) : "memory");
      }
      if (assign) {
        *reinterpret_cast<float4*>(rp) = x0;
        *reinterpret_cast<float4*>(rp + 4) = x1;
      }
      if (!has_out) continue;
    }
    const long long oo = (long long)(orow + row_add) * e.ldo + ocol;
    Vec8<bf16>::store(reinterpret_cast<bf16*>(e.out) + oo, v);
    if (has_relu) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
      Vec8<bf16>::store(reinterpret_cast<bf16*>(e.out_relu) + oo, v);
    }
  }
  __syncwarp();  // the tile is rewritten by the next chunk
}
