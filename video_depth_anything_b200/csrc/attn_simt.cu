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
                                                           int heads, int hpc, float scale) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const int d = C / heads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.y * hpc + warp;
  const int b = blockIdx.x / hw, p = blockIdx.x % hw;
  T* Ks = reinterpret_cast<T*>(smem_raw) + (size_t)warp * 2 * TT * d;
  T* Vs = Ks + TT * d;
  const long long ld = 3LL * C;
  auto row_ptr = [&](int t) { return qkv + (((long long)b * TT + t) * hw + p) * ld + h * d; };
  // stage K and V of this (token, head): TT rows of d contiguous elements
  const int vpr = d >> 3;
  for (int i = lane; i < TT * vpr; i += 32) {
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
    const bool active = t < TT;
    const T* qp = row_ptr(active ? t : 0);
    float s[TT];
#pragma unroll
    for (int j = 0; j < TT; ++j) s[j] = 0.f;
    for (int c = 0; c < d; c += 8) {
      float q8[8];
      Vec8<T>::load(qp + c, q8);
#pragma unroll
      for (int j = 0; j < TT; ++j) {
        float k8[8];
        Vec8<T>::load(Ks + j * d + c, k8);
#pragma unroll
        for (int i = 0; i < 8; ++i) s[j] = fmaf(q8[i], k8[i], s[j]);
      }
    }
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < TT; ++j) {
      s[j] = R<RB>(scale * s[j]);  // baddbmm(alpha=scale) output dtype
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
    T* op = o + (((long long)b * TT + (active ? t : 0)) * hw + p) * C + h * d;
    for (int c = 0; c < d; c += 8) {
      float a8[8] = {};
#pragma unroll
      for (int j = 0; j < TT; ++j) {
        float v8[8];
        Vec8<T>::load(Vs + j * d + c, v8);
#pragma unroll
        for (int i = 0; i < 8; ++i) a8[i] = fmaf(s[j], v8[i], a8[i]);
      }
      if (active) {
#pragma unroll
        for (int i = 0; i < 8; ++i) a8[i] = R<RB>(a8[i]);
        Vec8<T>::store(op + c, a8);
      }
    }
  }
}

template <typename T, int TT, bool RB>
int launch_temporal(const void* qkv, void* o, int batch, int hw, int C, int heads, cudaStream_t st) {
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
  kern<<<grid, hpc * 32, smem, st>>>((const T*)qkv, (T*)o, hw, C, heads, hpc, 1.0f / sqrtf((float)d));
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

extern "C" int vda_temporal_attention(const void* qkv, void* o, int32_t dtype, int32_t batch, int32_t T, int32_t hw,
                                      int32_t C, int32_t heads, int32_t round_bf16, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  VDA_CHECK_ARG(C % heads == 0 && (C / heads) % 8 == 0, "vda_temporal_attention: head dim must be a multiple of 8");
  VDA_CHECK_ARG(heads <= 8 && (heads & (heads - 1)) == 0, "vda_temporal_attention: heads must be 1,2,4,8");
  const bool bf = dtype == VDA_BF16 && round_bf16, f32 = dtype == VDA_F32 && !round_bf16;
  VDA_CHECK_ARG(bf || f32, "vda_temporal_attention: unsupported dtype/rounding combination");
#define TA(TT)                                                                              \
  return bf ? launch_temporal<bf16, TT, true>(qkv, o, batch, hw, C, heads, st)              \
            : launch_temporal<float, TT, false>(qkv, o, batch, hw, C, heads, st)
  switch (T) {
    case 1: TA(1);
    case 2: TA(2);
    case 3: TA(3);
    case 4: TA(4);
    case 8: TA(8);
    case 16: TA(16);
    case 32: TA(32);
    case 64: TA(64);
    default: VDA_CHECK_ARG(false, "vda_temporal_attention: T=%d not in {1,2,3,4,8,16,32,64}", T);
  }
#undef TA
}
