// SIMT attention kernels.
//  * encoder_attn_simt: exact two-pass softmax attention, one query row per thread, head_dim 64.  It is the
//    arithmetic of the fp32 build and the on-device cross-check of the tcgen05 flash kernel (attn_tc.cu).
//  * temporal_attn: the motion-module attention over the frame axis (T <= 64 keys, one problem per spatial
//    token and head): K/V of one (token, head) live in shared memory, one query per lane, scores in registers.
#include "common.cuh"

namespace {

constexpr int HD = 64;   // DINOv2 head dim
constexpr int KT = 64;   // keys per shared-memory tile

template <typename T, bool RB>
__global__ void __launch_bounds__(128) encoder_attn_simt(const T* __restrict__ qkv, T* __restrict__ o, int ntok, int heads) {
  __shared__ float Ks[KT][HD];
  __shared__ float Vs[KT][HD];
  const int f = blockIdx.z, h = blockIdx.y;
  const int qi = blockIdx.x * 128 + threadIdx.x;
  const int ld = 3 * heads * HD;
  const T* base = qkv + (long long)f * ntok * ld;
  const bool active = qi < ntok;
  float q[HD];
  if (active) {
    const T* qp = base + (long long)qi * ld + h * HD;
#pragma unroll
    for (int c = 0; c < HD; c += 8) {
      float t[8];
      Vec8<T>::load(qp + c, t);
#pragma unroll
      for (int i = 0; i < 8; ++i) q[c + i] = t[i] * 0.125f;  // attention.py:53 (exact power of two)
    }
  } else {
#pragma unroll
    for (int c = 0; c < HD; ++c) q[c] = 0.f;
  }
  // three sweeps over the keys: row max, softmax denominator, then P @ V with P = softmax rounded like autocast does
  float mx = -INFINITY, sum = 0.f;
  float acc[HD];
#pragma unroll
  for (int c = 0; c < HD; ++c) acc[c] = 0.f;
  for (int pass = 0; pass < 3; ++pass) {
    for (int k0 = 0; k0 < ntok; k0 += KT) {
      __syncthreads();
      for (int i = threadIdx.x; i < KT * (HD / 8); i += 128) {
        int r = i / (HD / 8), c = (i % (HD / 8)) * 8;
        float tk[8] = {}, tv[8] = {};
        if (k0 + r < ntok) {
          const T* kp = base + (long long)(k0 + r) * ld + heads * HD + h * HD + c;
          Vec8<T>::load(kp, tk);
          if (pass == 2) Vec8<T>::load(kp + heads * HD, tv);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) { Ks[r][c + j] = tk[j]; Vs[r][c + j] = tv[j]; }
      }
      __syncthreads();
      const int kn = min(KT, ntok - k0);
      for (int j = 0; j < kn; ++j) {
        float s = 0.f;
#pragma unroll
        for (int c = 0; c < HD; ++c) s = fmaf(q[c], Ks[j][c], s);
        s = R<RB>(s);  // q @ k^T is a bf16 tensor under autocast
        if (pass == 0) mx = fmaxf(mx, s);
        else if (pass == 1) sum += expf(s - mx);
        else {
          float p = R<RB>(expf(s - mx) / sum);  // softmax in fp32, probabilities cast to bf16 for attn @ v
#pragma unroll
          for (int c = 0; c < HD; ++c) acc[c] = fmaf(p, Vs[j][c], acc[c]);
        }
      }
    }
  }
  if (active) {
    T* op = o + ((long long)f * ntok + qi) * (heads * HD) + h * HD;
#pragma unroll
    for (int c = 0; c < HD; c += 8) {
      float t[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) t[i] = R<RB>(acc[c + i]);
      Vec8<T>::store(op + c, t);
    }
  }
}

// ---------------------------------------------------------------- temporal attention
template <typename T, int TT, bool RB>
__global__ void __launch_bounds__(256) temporal_attn_kernel(const T* __restrict__ qkv, T* __restrict__ o, int hw, int C,
                                                           int heads, int hpc, float scale, int Tn) {
  // TT = compile-time capacity (registers, shared memory), Tn <= TT = the frames that exist: any --inference_length /
  // forward T runs on the next capacity up; key columns j >= Tn score -inf, query rows t >= Tn are idle.
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int d = C / heads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.y * hpc + warp;
  const int b = blockIdx.x / hw, p = blockIdx.x % hw;
  T* Ks = reinterpret_cast<T*>(smem_raw) + (size_t)warp * 2 * TT * d;
  T* Vs = Ks + TT * d;
  const long long ld = 3LL * C;
  auto row_ptr = [&](int t) { return qkv + (((long long)b * Tn + t) * hw + p) * ld + h * d; };
  // stage K and V of this (token, head): Tn rows of d contiguous elements
  const int vpr = d >> 3;
  for (int i = lane; i < Tn * vpr; i += 32) {
    int t = i / vpr, c = (i % vpr) * 8;
    const T* rp = row_ptr(t);
    if constexpr (sizeof(T) == 2) {
      *reinterpret_cast<uint4*>(Ks + t * d + c) = *reinterpret_cast<const uint4*>(rp + C + c);
      *reinterpret_cast<uint4*>(Vs + t * d + c) = *reinterpret_cast<const uint4*>(rp + 2 * C + c);
    } else {
      const float4* k4 = reinterpret_cast<const float4*>(rp + C + c);
      const float4* v4 = reinterpret_cast<const float4*>(rp + 2 * C + c);
      float4* ks4 = reinterpret_cast<float4*>(Ks + t * d + c);
      float4* vs4 = reinterpret_cast<float4*>(Vs + t * d + c);
      ks4[0] = k4[0]; ks4[1] = k4[1];
      vs4[0] = v4[0]; vs4[1] = v4[1];
    }
  }
  __syncwarp();
  constexpr int NQ = (TT + 31) / 32;
#pragma unroll
  for (int qq = 0; qq < NQ; ++qq) {
    const int t = qq * 32 + lane;
    const bool active = t < Tn;
    const T* qp = row_ptr(active ? t : 0);
    float s[TT];
#pragma unroll
    for (int j = 0; j < TT; ++j) s[j] = 0.f;
    for (int c = 0; c < d; c += 8) {
      float q8[8];
      Vec8<T>::load(qp + c, q8);
#pragma unroll
      for (int j = 0; j < TT; ++j) {
        if (j < Tn) {
          float k8[8];
          Vec8<T>::load(Ks + j * d + c, k8);
#pragma unroll
          for (int i = 0; i < 8; ++i) s[j] = fmaf(q8[i], k8[i], s[j]);
        }
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < TT; ++j) {
      s[j] = j < Tn ? R<RB>(scale * s[j]) : -INFINITY;  // baddbmm(alpha=scale) output dtype
      mx = fmaxf(mx, s[j]);
    }
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < TT; ++j) {
      s[j] = expf(s[j] - mx);
      sum += s[j];
    }
#pragma unroll
    for (int j = 0; j < TT; ++j) s[j] = R<RB>(s[j] / sum);
    T* op = o + (((long long)b * Tn + (active ? t : 0)) * hw + p) * C + h * d;
    for (int c = 0; c < d; c += 8) {
      float a8[8] = {};
#pragma unroll
      for (int j = 0; j < TT; ++j) {
        if (j < Tn) {
          float v8[8];
          Vec8<T>::load(Vs + j * d + c, v8);
#pragma unroll
          for (int i = 0; i < 8; ++i) a8[i] = fmaf(s[j], v8[i], a8[i]);
        }
      }
      if (active) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a8[i] = R<RB>(a8[i]);
        Vec8<T>::store(op + c, a8);
      }
    }
  }
}


// ---------------------------------------------------------------- temporal attention, bf16 build, warp-level MMA
// One warp per (token, head): Q, K, V (T x d bf16 each) are staged in shared memory with row-contiguous 16-byte loads,
// S = Q K^T and O = P V run on mma.sync.m16n8k16 (the problem is 32 x 32 x d -- far too small for a tcgen05 tile, and
// on fp32 FMAs it is issue-bound at ~4x the memory time), the softmax works on the accumulator fragments with quad
// shuffles, P never leaves registers (accumulator layout == A-operand layout), O goes back through shared memory so
// that global stores are row-contiguous.  Rounding points as in the SIMT kernel: scores -> bf16, softmax fp32,
// probabilities -> bf16, output -> bf16 (motion_module/attention.py:187-207 under autocast).
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ float exp2f_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void mma16816(float* c, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// SPLIT warps share one (b, p, h) problem: they stage q / k / v together and each takes MT / SPLIT query tiles.  The
// problem is a latency chain (ldmatrix -> mma -> shuffles -> exp -> mma) and the d = 128 staging tiles allow only two
// blocks per SM, so two warps per problem double the chains in flight for the same shared memory.
template <int TT, int SPLIT>
__global__ void __launch_bounds__(128 * SPLIT) temporal_attn_mma_kernel(const bf16* __restrict__ qkv, bf16* __restrict__ o,
                                                                        int hw, int C, int heads, int d, float scale,
                                                                        int items) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  constexpr int NT = TT / 8;   // score n-tiles (keys)
  constexpr int MT = TT / 16;  // m-tiles (queries)
  static_assert(MT % SPLIT == 0, "query tiles per warp");
  const int lane = threadIdx.x & 31;
  const int part = (threadIdx.x >> 5) % SPLIT;   // which share of the query tiles
  const int warp = (threadIdx.x >> 5) / SPLIT;   // problem slot inside the block
  const int item = blockIdx.x * (blockDim.x / (32 * SPLIT)) + warp;  // (b, p, h)
  if (item >= items) return;  // both warps of a problem leave together; the barrier below is per problem
  const int h = item % heads;
  const int bp = item / heads;
  const int b = bp / hw, p = bp % hw;
  const int dp = (d + 15) & ~15;  // head dim rounded up to the MMA's k = 16: the extra columns are staged as zeros
  const int pitch = 2 * dp + 16;  // bytes per staged row: 16 B of padding keeps ldmatrix conflict-free
  uint8_t* Qs = smem_raw + (size_t)warp * 3 * TT * pitch;
  uint8_t* Ks = Qs + TT * pitch;
  uint8_t* Vs = Ks + TT * pitch;
  const long long ld = 3LL * C;
  const int vpr = d >> 3;  // 16-byte vectors per row
  const uint32_t qs = (uint32_t)__cvta_generic_to_shared(Qs), ks = (uint32_t)__cvta_generic_to_shared(Ks),
                 vs = (uint32_t)__cvta_generic_to_shared(Vs);
  // The kernel is HBM bound (q, k, v in, o out) and a warp's rows are short (2 d bytes): every 16-byte piece goes global ->
  // shared as its own cp.async, ALL of them in flight before the first wait, so a warp keeps 3 T d 2 bytes outstanding
  // instead of three loads (few warps are resident: the staging tiles of d = 128 take 104 KB per block).
  for (int i = part * 32 + lane; i < TT * vpr; i += 32 * SPLIT) {
    const int t = i / vpr, c = i - t * vpr;
    const bf16* rp = qkv + (((long long)b * TT + t) * hw + p) * ld + h * d + c * 8;
    const uint32_t off = t * pitch + c * 16;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(qs + off), "l"(rp) : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(ks + off), "l"(rp + C) : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(vs + off), "l"(rp + 2 * C) : "memory");
  }
  if (dp != d) {  // d = 8 or 24 (vits): one zero 16-byte piece per row completes the last k-step / output tile
    for (int t = part * 32 + lane; t < TT; t += 32 * SPLIT) {
      const uint32_t off = t * pitch + vpr * 16;
      *reinterpret_cast<uint4*>(Qs + off) = make_uint4(0, 0, 0, 0);
      *reinterpret_cast<uint4*>(Ks + off) = make_uint4(0, 0, 0, 0);
      *reinterpret_cast<uint4*>(Vs + off) = make_uint4(0, 0, 0, 0);
    }
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if constexpr (SPLIT == 1) __syncwarp();
  else asm volatile("bar.sync %0, %1;" ::"r"(1 + warp), "n"(32 * SPLIT) : "memory");
  const int g = lane >> 2, tq = lane & 3;
  // ldmatrix lane -> row / column-half assignments
  const int a_row = lane & 15, a_col = (lane >> 4) * 16;                       // A: 16 rows x (2 x 8 cols)
  const int b_row = (lane & 7) + ((lane >> 4) << 3), b_col = ((lane >> 3) & 1) * 16;  // B (two n-tiles): [n][k]
  const int v_row = (lane & 7) + (((lane >> 3) & 1) << 3), v_col = (lane >> 4) * 16;   // V^T (two n-tiles): [k][n]
#pragma unroll
  for (int mi = 0; mi < MT / SPLIT; ++mi) {
    const int mt = part * (MT / SPLIT) + mi;
    float sc[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) sc[n][0] = sc[n][1] = sc[n][2] = sc[n][3] = 0.f;
    for (int k0 = 0; k0 < dp; k0 += 16) {
      uint32_t a[4];
      ldsm_x4(qs + (mt * 16 + a_row) * pitch + k0 * 2 + a_col, a[0], a[1], a[2], a[3]);
#pragma unroll
      for (int n2 = 0; n2 < NT / 2; ++n2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(ks + (n2 * 16 + b_row) * pitch + k0 * 2 + b_col, b0, b1, b2, b3);
        mma16816(sc[2 * n2], a, b0, b1);
        mma16816(sc[2 * n2 + 1], a, b2, b3);
      }
    }
    // rows g and g + 8 of this m-tile: scores -> bf16, softmax in fp32 over the quad, probabilities -> bf16
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < NT; ++n) {
#pragma unroll
      for (int i = 0; i < 4; ++i) sc[n][i] = round_bf16(scale * sc[n][i]);
      mx0 = fmaxf(mx0, fmaxf(sc[n][0], sc[n][1]));
      mx1 = fmaxf(mx1, fmaxf(sc[n][2], sc[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      // ex2.approx on (x - max) log2 e: 2^-22 relative, far inside the bf16 rounding of the probabilities that follows;
      // libm expf and the fp32 divisions below were most of this kernel's instructions at head dim 32 (SM bound, not HBM)
      sc[n][0] = exp2f_approx((sc[n][0] - mx0) * 1.4426950408889634f);
      sc[n][1] = exp2f_approx((sc[n][1] - mx0) * 1.4426950408889634f);
      sc[n][2] = exp2f_approx((sc[n][2] - mx1) * 1.4426950408889634f);
      sc[n][3] = exp2f_approx((sc[n][3] - mx1) * 1.4426950408889634f);
      s0 += sc[n][0] + sc[n][1];
      s1 += sc[n][2] + sc[n][3];
    }
    s0 += __shfl_xor_sync(0xffffffffu, s0, 1);
    s0 += __shfl_xor_sync(0xffffffffu, s0, 2);
    s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
    s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
    const float r0 = __frcp_rn(s0), r1 = __frcp_rn(s1);
    uint32_t pa[NT / 2][4];  // P as A fragments: k-step n2 covers keys 16 n2 .. 16 n2 + 15
#pragma unroll
    for (int n2 = 0; n2 < NT / 2; ++n2) {
      pa[n2][0] = pack_bf16x2(sc[2 * n2][0] * r0, sc[2 * n2][1] * r0);
      pa[n2][1] = pack_bf16x2(sc[2 * n2][2] * r1, sc[2 * n2][3] * r1);
      pa[n2][2] = pack_bf16x2(sc[2 * n2 + 1][0] * r0, sc[2 * n2 + 1][1] * r0);
      pa[n2][3] = pack_bf16x2(sc[2 * n2 + 1][2] * r1, sc[2 * n2 + 1][3] * r1);
    }
    // O = P V, 16 output columns at a time; written into the (already consumed) Q rows of this m-tile
    for (int n0 = 0; n0 < dp; n0 += 16) {
      float oc[2][4] = {};
#pragma unroll
      for (int n2 = 0; n2 < NT / 2; ++n2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(vs + (n2 * 16 + v_row) * pitch + n0 * 2 + v_col, b0, b1, b2, b3);
        mma16816(oc[0], pa[n2], b0, b1);
        mma16816(oc[1], pa[n2], b2, b3);
      }
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        *reinterpret_cast<uint32_t*>(Qs + (mt * 16 + g) * pitch + (n0 + j * 8 + tq * 2) * 2) = pack_bf16x2(oc[j][0], oc[j][1]);
        *reinterpret_cast<uint32_t*>(Qs + (mt * 16 + g + 8) * pitch + (n0 + j * 8 + tq * 2) * 2) = pack_bf16x2(oc[j][2], oc[j][3]);
      }
    }
  }
  __syncwarp();
  // each warp stores the rows of its own query tiles
  for (int i = lane; i < (TT / SPLIT) * vpr; i += 32) {
    const int t = part * (TT / SPLIT) + i / vpr, c = i % vpr;
    *reinterpret_cast<uint4*>(o + (((long long)b * TT + t) * hw + p) * C + h * d + c * 8) =
        *reinterpret_cast<const uint4*>(Qs + t * pitch + c * 16);
  }
}

template <int TT>
int launch_temporal_mma(const void* qkv, void* o, int batch, int hw, int C, int heads, cudaStream_t st) {
  const int d = C / heads;
  const int pitch = 2 * ((d + 15) & ~15) + 16;
  const int wpb = 4;
  const size_t smem = (size_t)wpb * 3 * TT * pitch;
  constexpr int SPLIT = TT >= 32 ? 2 : 1;
  auto kern = temporal_attn_mma_kernel<TT, SPLIT>;
  if (smem > 48 * 1024) {
    cudaError_t ce = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (ce != cudaSuccess) {
      vda_set_error("temporal_attn_mma: smem attribute %zu failed: %s", smem, cudaGetErrorString(ce));
      return (int)ce;
    }
  }
  const int items = batch * hw * heads;
  kern<<<(items + wpb - 1) / wpb, wpb * 32 * SPLIT, smem, st>>>((const bf16*)qkv, (bf16*)o, hw, C, heads, d, 1.0f / sqrtf((float)d), items);
  VDA_CUDA_LAUNCH_CHECK("temporal_attn_mma_kernel");
  return 0;
}

template <typename T, int TT, bool RB>
int launch_temporal(const void* qkv, void* o, int batch, int hw, int C, int heads, cudaStream_t st, int Tn) {
  const int d = C / heads;
  int hpc = heads;
  while (hpc > 1 && (size_t)hpc * 2 * TT * d * sizeof(T) > 96 * 1024) hpc >>= 1;
  const size_t smem = (size_t)hpc * 2 * TT * d * sizeof(T);
  auto kern = temporal_attn_kernel<T, TT, RB>;
  if (smem > 48 * 1024) {
    cudaError_t ce = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (ce != cudaSuccess) {
      vda_set_error("temporal_attn: smem attribute %zu failed: %s", smem, cudaGetErrorString(ce));
      return (int)ce;
    }
  }
  dim3 grid(batch * hw, heads / hpc);
  kern<<<grid, hpc * 32, smem, st>>>((const T*)qkv, (T*)o, hw, C, heads, hpc, 1.0f / sqrtf((float)d), Tn);
  VDA_CUDA_LAUNCH_CHECK("temporal_attn_kernel");
  return 0;
}

}  // namespace

int vda_attention_tc(const void* qkv, void* o, int frames, int ntok, int heads, cudaStream_t st);

extern "C" int vda_attention(const void* qkv, void* o, int32_t dtype, int32_t frames, int32_t ntok, int32_t heads,
                             int32_t round_bf16, int32_t use_tc, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  VDA_CHECK_ARG(frames > 0 && ntok > 0 && heads > 0, "vda_attention: bad shape");
  if (use_tc) {
    VDA_CHECK_ARG(dtype == VDA_BF16 && round_bf16, "vda_attention(tc): bf16 build only");
    return vda_attention_tc(qkv, o, frames, ntok, heads, st);
  }
  dim3 grid((ntok + 127) / 128, heads, frames);
  if (dtype == VDA_BF16 && round_bf16) encoder_attn_simt<bf16, true><<<grid, 128, 0, st>>>((const bf16*)qkv, (bf16*)o, ntok, heads);
  else if (dtype == VDA_F32 && !round_bf16) encoder_attn_simt<float, false><<<grid, 128, 0, st>>>((const float*)qkv, (float*)o, ntok, heads);
  else VDA_CHECK_ARG(false, "vda_attention: unsupported dtype/rounding combination");
  VDA_CUDA_LAUNCH_CHECK("encoder_attn_simt");
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// Rotary position embedding of the temporal attention (pe='rope': motion_module/attention.py:403-429 applied at
// motion_module.py:290-293).  The reference rotates the full C-wide q and k rows BEFORE the head split: consecutive
// channel pairs (2i, 2i+1) are one complex number multiplied by exp(i t theta_i), theta_i = 10000^(-2i/C), t = frame
// index, in fp32, then cast back to the activation type.  q and k live in columns [0, 2C) of the fused qkv matrix and
// are rotated in place; tab[t, 2i .. 2i+1] = (cos, sin)(t theta_i) is the caller's copy of `freqs_cis`.
template <typename T, bool RB>
__global__ void rope_qk_kernel(T* __restrict__ qkv, const float* __restrict__ tab, long long rows, int hw, int Tn, int C) {
  const int vpr = (2 * C) >> 3;  // 8-element vectors per row over the q and k columns
  const long long total = rows * vpr;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / vpr;
    const int c = (int)(i - r * vpr) << 3;   // column in [0, 2C)
    const int cc = c >= C ? c - C : c;       // channel inside q or k
    const int t = (int)((r / hw) % Tn);
    T* p = qkv + r * (3LL * C) + c;
    float v[8], w[8];
    Vec8<T>::load(p, v);
    const float4 a = *reinterpret_cast<const float4*>(tab + (long long)t * C + cc);
    const float4 b = *reinterpret_cast<const float4*>(tab + (long long)t * C + cc + 4);
    const float cs[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
    for (int k = 0; k < 8; k += 2) {
      w[k] = R<RB>(v[k] * cs[k] - v[k + 1] * cs[k + 1]);
      w[k + 1] = R<RB>(v[k] * cs[k + 1] + v[k + 1] * cs[k]);
    }
    Vec8<T>::store(p, w);
  }
}

extern "C" int vda_rope_qk(void* qkv, int32_t dtype, const float* tab, int64_t rows, int32_t hw, int32_t T, int32_t C,
                           int32_t round_bf16, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  VDA_CHECK_ARG(C % 8 == 0 && hw > 0 && T > 0, "vda_rope_qk: C must be a multiple of 8");
  const bool bf = dtype == VDA_BF16 && round_bf16, f32 = dtype == VDA_F32 && !round_bf16;
  VDA_CHECK_ARG(bf || f32, "vda_rope_qk: unsupported dtype/rounding combination");
  if (rows == 0) return 0;
  const long long total = rows * ((2LL * C) >> 3);
  const int threads = 256;
  long long want = (total + threads - 1) / threads;
  const long long cap = (long long)vda_sm_count() * 16;
  const int blocks = (int)(want < cap ? want : cap);
  if (bf) rope_qk_kernel<bf16, true><<<blocks, threads, 0, st>>>(reinterpret_cast<bf16*>(qkv), tab, rows, hw, T, C);
  else rope_qk_kernel<float, false><<<blocks, threads, 0, st>>>(reinterpret_cast<float*>(qkv), tab, rows, hw, T, C);
  VDA_CUDA_LAUNCH_CHECK("rope_qk_kernel");
  return 0;
}

extern "C" int vda_temporal_attention(const void* qkv, void* o, int32_t dtype, int32_t batch, int32_t T, int32_t hw,
                                      int32_t C, int32_t heads, int32_t round_bf16, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  VDA_CHECK_ARG(C % heads == 0 && (C / heads) % 8 == 0, "vda_temporal_attention: head dim must be a multiple of 8");
  VDA_CHECK_ARG(heads <= 8 && (heads & (heads - 1)) == 0, "vda_temporal_attention: heads must be 1,2,4,8");
  const bool bf = dtype == VDA_BF16 && round_bf16, f32 = dtype == VDA_F32 && !round_bf16;
  VDA_CHECK_ARG(bf || f32, "vda_temporal_attention: unsupported dtype/rounding combination");
  const int dd = C / heads;
  if (bf && dd % 8 == 0 && dd <= 128 && (T == 16 || T == 32 || T == 64)) {
    if (T == 16) return launch_temporal_mma<16>(qkv, o, batch, hw, C, heads, st);
    if (T == 32) return launch_temporal_mma<32>(qkv, o, batch, hw, C, heads, st);
    return launch_temporal_mma<64>(qkv, o, batch, hw, C, heads, st);
  }
#define TA(TT)                                                                              \
  return bf ? launch_temporal<bf16, TT, true>(qkv, o, batch, hw, C, heads, st, T)           \
            : launch_temporal<float, TT, false>(qkv, o, batch, hw, C, heads, st, T)
  // any T in [1, 64] (the reference accepts any --inference_length, run.py:35): the SIMT kernel at the next capacity up
  VDA_CHECK_ARG(T >= 1 && T <= 64, "vda_temporal_attention: T=%d not in [1, 64]", T);
  if (T <= 4) TA(4);
  if (T <= 8) TA(8);
  if (T <= 16) TA(16);
  if (T <= 32) TA(32);
  TA(64);
#undef TA
}
