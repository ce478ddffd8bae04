// Egress side of the path: the depth visualisation of `save_video(..., is_depths=True)`
// (utils/dc_utils.py:70-86) on the device, so that what crosses PCIe is the uint8 picture (1 or 3 B per pixel) instead of
// the fp32 depth stack (4 B per pixel) and the host does no per-pixel work before the encoder.
//   d_min, d_max = frames.min(), frames.max()                         -> vda_depth_minmax   (one HBM read, 4 B/pixel)
//   depth_norm   = ((depth - d_min) / (d_max - d_min) * 255).astype(np.uint8)
//   picture      = depth_norm (grayscale) | (colormap[depth_norm] * 255).astype(np.uint8)      -> vda_depth_vis
// numpy evaluates the normalisation in float32 (float32 array, float32 scalars, a Python int): subtract, IEEE divide,
// multiply, truncate.  The kernels use the same three round-to-nearest fp32 operations, so the uint8 picture is
// bit-identical to the reference's.  HBM bound: 4 B read + 1 (grey) or 3 (colour) B written per pixel.
#include "common.cuh"

namespace {

constexpr int MM_BLOCKS_MAX = 1024;

__device__ __forceinline__ void warp_minmax(float& mn, float& mx) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
}

__device__ __forceinline__ void block_minmax(float& mn, float& mx) {
  __shared__ float smn[8], smx[8];
  warp_minmax(mn, mx);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { smn[w] = mn; smx[w] = mx; }
  __syncthreads();
  if (w == 0) {
    mn = l < (int)(blockDim.x >> 5) ? smn[l] : INFINITY;
    mx = l < (int)(blockDim.x >> 5) ? smx[l] : -INFINITY;
    warp_minmax(mn, mx);
  }
}

// partial[2b], partial[2b+1] = min, max of block b's grid-stride share (float4 loads on the aligned body)
__global__ void __launch_bounds__(256) minmax_partial_kernel(const float* __restrict__ x, long long n,
                                                            float* __restrict__ partial, int vec) {
  float mn = INFINITY, mx = -INFINITY;
  const long long n4 = vec ? n >> 2 : 0;  // vec = 0: the pointer is not 16-byte aligned (odd frame sizes), scalar loads
  const float4* x4 = reinterpret_cast<const float4*>(x);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = x4[i];
    mn = fminf(fminf(mn, v.x), fminf(v.y, fminf(v.z, v.w)));
    mx = fmaxf(fmaxf(mx, v.x), fmaxf(v.y, fmaxf(v.z, v.w)));
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    mn = fminf(mn, x[i]);
    mx = fmaxf(mx, x[i]);
  }
  block_minmax(mn, mx);
  if (threadIdx.x == 0) { partial[2 * blockIdx.x] = mn; partial[2 * blockIdx.x + 1] = mx; }
}

__global__ void __launch_bounds__(256) minmax_final_kernel(const float* __restrict__ partial, int blocks,
                                                          float* __restrict__ out) {
  float mn = INFINITY, mx = -INFINITY;
  for (int i = threadIdx.x; i < blocks; i += blockDim.x) { mn = fminf(mn, partial[2 * i]); mx = fmaxf(mx, partial[2 * i + 1]); }
  block_minmax(mn, mx);
  if (threadIdx.x == 0) { out[0] = mn; out[1] = mx; }
}

// ((d - mn) / (mx - mn) * 255).astype(uint8): three rn fp32 operations, then truncation; NaN (mx == mn) -> 0
__device__ __forceinline__ uint32_t norm_u8(float d, float mn, float range) {
  const float v = __fmul_rn(__fdiv_rn(__fsub_rn(d, mn), range), 255.f);
  return v >= 0.f ? (uint32_t)min((int)v, 255) : 0u;   // the comparison is false for NaN
}

// four pixels per thread: one float4 in, one uchar4 (grey) or three 32-bit words (RGB through the 256-entry table) out
template <bool COLOUR>
__global__ void __launch_bounds__(256) depth_vis_kernel(const float* __restrict__ d, long long n,
                                                       const float* __restrict__ minmax,
                                                       const uint8_t* __restrict__ lut, uint8_t* __restrict__ out, int vec) {
  __shared__ uint32_t slut[256];
  if (COLOUR) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x)
      slut[i] = (uint32_t)lut[3 * i] | ((uint32_t)lut[3 * i + 1] << 8) | ((uint32_t)lut[3 * i + 2] << 16);
    __syncthreads();
  }
  const float mn = minmax[0], range = __fsub_rn(minmax[1], minmax[0]);
  const long long n4 = vec ? n >> 2 : 0;
  // two float4 loads in flight per thread (the loop body is ~80 instructions of IEEE division behind each load)
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < n4; i0 += 2 * stride) {
    float4 v[2];
    v[0] = reinterpret_cast<const float4*>(d)[i0];
    const bool two = i0 + stride < n4;
    if (two) v[1] = reinterpret_cast<const float4*>(d)[i0 + stride];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (u == 1 && !two) break;
      const long long i = i0 + u * stride;
      const uint32_t a = norm_u8(v[u].x, mn, range), b = norm_u8(v[u].y, mn, range), c = norm_u8(v[u].z, mn, range),
                     e = norm_u8(v[u].w, mn, range);
      if (!COLOUR) {
        reinterpret_cast<uint32_t*>(out)[i] = a | (b << 8) | (c << 16) | (e << 24);
      } else {
        const uint32_t A = slut[a], B = slut[b], Cc = slut[c], E = slut[e];  // 0x00BBGGRR each
        uint32_t* o = reinterpret_cast<uint32_t*>(out) + 3 * i;             // 12 bytes: R G B R | G B R G | B R G B
        o[0] = A | (B << 24);
        o[1] = (B >> 8) | (Cc << 16);
        o[2] = (Cc >> 16) | (E << 8);
      }
    }
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
      const uint32_t a = norm_u8(d[i], mn, range);
      if (!COLOUR) out[i] = (uint8_t)a;
      else { out[3 * i] = lut[3 * a]; out[3 * i + 1] = lut[3 * a + 1]; out[3 * i + 2] = lut[3 * a + 2]; }
    }
}

}  // namespace

extern "C" int64_t vda_depth_minmax_workspace_floats(void) { return 2 * MM_BLOCKS_MAX; }

extern "C" int vda_depth_minmax(const float* depth, int64_t count, float* partial, float* minmax, vda_stream_t stream) {
  VDA_CHECK_ARG(count > 0, "vda_depth_minmax: empty input (numpy raises on min() of an empty array)");
  const int vec = (uintptr_t)depth % 16 == 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  long long blocks = (count / 4 + 256 * 8 - 1) / (256 * 8);
  if (blocks > 4LL * vda_sm_count()) blocks = 4LL * vda_sm_count();
  if (blocks > MM_BLOCKS_MAX) blocks = MM_BLOCKS_MAX;
  if (blocks < 1) blocks = 1;
  minmax_partial_kernel<<<(unsigned)blocks, 256, 0, s>>>(depth, count, partial, vec);
  VDA_CUDA_LAUNCH_CHECK("minmax_partial_kernel");
  minmax_final_kernel<<<1, 256, 0, s>>>(partial, (int)blocks, minmax);
  VDA_CUDA_LAUNCH_CHECK("minmax_final_kernel");
  return 0;
}

extern "C" int vda_depth_vis(const float* depth, int64_t count, const float* minmax, const uint8_t* lut, uint8_t* out,
                             vda_stream_t stream) {
  if (count <= 0) return 0;
  const int vec = ((uintptr_t)depth % 16 == 0) && ((uintptr_t)out % 4 == 0);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  long long blocks = (count / 4 + 256 * 4 - 1) / (256 * 4);
  if (blocks > 8LL * vda_sm_count()) blocks = 8LL * vda_sm_count();
  if (blocks < 1) blocks = 1;
  if (lut) depth_vis_kernel<true><<<(unsigned)blocks, 256, 0, s>>>(depth, count, minmax, lut, out, vec);
  else depth_vis_kernel<false><<<(unsigned)blocks, 256, 0, s>>>(depth, count, minmax, nullptr, out, vec);
  VDA_CUDA_LAUNCH_CHECK("depth_vis_kernel");
  return 0;
}
