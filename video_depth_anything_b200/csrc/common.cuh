// Shared device/host helpers for libvda_sm100a.  sm_100a only.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/vda.h"

typedef __nv_bfloat16 bf16;

// ---------------------------------------------------------------- error plumbing
void vda_set_error(const char* fmt, ...);
#define VDA_CHECK_ARG(cond, ...)        \
  do {                                  \
    if (!(cond)) {                      \
      vda_set_error(__VA_ARGS__);       \
      return -1;                        \
    }                                   \
  } while (0)
#define VDA_CUDA_LAUNCH_CHECK(name)                                             \
  do {                                                                          \
    cudaError_t e__ = cudaGetLastError();                                       \
    if (e__ != cudaSuccess) {                                                   \
      vda_set_error("%s launch failed: %s", name, cudaGetErrorString(e__));     \
      return (int)e__;                                                          \
    }                                                                           \
  } while (0)

// ---------------------------------------------------------------- numeric helpers
// Round an fp32 value to the nearest bf16 (RNE) and come back: the "rounding point" of autocast.
__device__ __forceinline__ float round_bf16(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }
template <bool RB>
__device__ __forceinline__ float R(float v) {
  if constexpr (RB) return round_bf16(v);
  else return v;
}
// packed fp32 pairs (FMUL2 / FFMA2: one issue slot for two IEEE fp32 operations) for the interpolation kernels
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) {
  uint64_t r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}

// exact-erf GELU, the nn.GELU() / F.gelu default (dinov2.py:61, motion_module/attention.py:378)
__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

// GELU for the bf16 build's GEMM epilogue: x * Phi(x) with erfc from Abramowitz-Stegun 7.1.26 (|erf error| <= 1.5e-7,
// two MUFU ops + 9 FMA-pipe ops instead of erff's ~40 instructions).  The negative side is formed without
// cancellation (Phi = erfc/2).  The result is rounded to bf16 (2^-9 relative) right after, so this is far inside the
// rounding noise; the fp32 build keeps gelu_erf.
// 13 instructions: the 0.5 is folded into the polynomial, the reciprocal is a bare MUFU.RCP (its argument is >= 1),
// and x Phi(x) = max(x, 0) - |x| erfc(|x| / sqrt 2) / 2 needs no sign select.
__device__ __forceinline__ float gelu_fast(float x) {
  const float ax = fabsf(x);
  float t, e;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(0.3275911f * 0.70710678118654752440f, ax, 1.0f)));
  float p = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);
  p = fmaf(t, p, 0.5f * 1.421413741f);
  p = fmaf(t, p, 0.5f * -0.284496736f);
  p = fmaf(t, p, 0.5f * 0.254829592f);
  const float y = x * 0.84932180028801904272f;  // sqrt(log2(e) / 2): 2^(-y^2) = exp(-x^2 / 2)
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-y * y));
  const float h = (t * p) * e;  // erfc(|x| / sqrt 2) / 2
  return fmaf(-ax, h, fmaxf(x, 0.f));
}

// Two GELUs at once with the FMA-pipe work as packed fp32 pairs (FMUL2 / FFMA2): the same operations in the same order
// as gelu_fast on each element (bit-identical results), ~8 instead of 13 issue slots per element.
__device__ __forceinline__ void gelu_fast2(float x0, float x1, float& o0, float& o1) {
  const float a0 = fabsf(x0), a1 = fabsf(x1);
  const uint64_t ax = pk2(a0, a1), x = pk2(x0, x1);
  float u0, u1, t0, t1, z0, z1, e0, e1;
  constexpr float K = 0.3275911f * 0.70710678118654752440f;
  upk2(fma2(pk2(K, K), ax, pk2(1.0f, 1.0f)), u0, u1);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(u0));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(u1));
  const uint64_t t = pk2(t0, t1);
  uint64_t p = fma2(t, pk2(0.5f * 1.061405429f, 0.5f * 1.061405429f), pk2(0.5f * -1.453152027f, 0.5f * -1.453152027f));
  p = fma2(t, p, pk2(0.5f * 1.421413741f, 0.5f * 1.421413741f));
  p = fma2(t, p, pk2(0.5f * -0.284496736f, 0.5f * -0.284496736f));
  p = fma2(t, p, pk2(0.5f * 0.254829592f, 0.5f * 0.254829592f));
  const uint64_t y = mul2(x, pk2(0.84932180028801904272f, 0.84932180028801904272f));
  upk2(mul2(y, y), z0, z1);
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e0) : "f"(-z0));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e1) : "f"(-z1));
  const uint64_t h = mul2(mul2(t, p), pk2(e0, e1));
  upk2(fma2(pk2(-a0, -a1), h, pk2(fmaxf(x0, 0.f), fmaxf(x1, 0.f))), o0, o1);
}

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 p = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&p);
}
__device__ __forceinline__ float bf16lo_to_f32(uint32_t p) { return __uint_as_float(p << 16); }
__device__ __forceinline__ float bf16hi_to_f32(uint32_t p) { return __uint_as_float(p & 0xffff0000u); }

// load / store `n` (<=8, multiple handled by caller) consecutive elements as fp32
template <typename T> struct Vec8;
template <> struct Vec8<float> {
  static __device__ __forceinline__ void load(const float* p, float* v) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
  }
};
template <> struct Vec8<bf16> {
  static __device__ __forceinline__ void load(const bf16* p, float* v) {
    uint4 a = *reinterpret_cast<const uint4*>(p);
    v[0] = bf16lo_to_f32(a.x); v[1] = bf16hi_to_f32(a.x); v[2] = bf16lo_to_f32(a.y); v[3] = bf16hi_to_f32(a.y);
    v[4] = bf16lo_to_f32(a.z); v[5] = bf16hi_to_f32(a.z); v[6] = bf16lo_to_f32(a.w); v[7] = bf16hi_to_f32(a.w);
  }
  static __device__ __forceinline__ void store(bf16* p, const float* v) {
    uint4 a;
    a.x = pack_bf16x2(v[0], v[1]); a.y = pack_bf16x2(v[2], v[3]);
    a.z = pack_bf16x2(v[4], v[5]); a.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = a;
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

static inline int vda_sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}
