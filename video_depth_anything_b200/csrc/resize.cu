// Memory-bound layout / resampling kernels of the forward pass: bilinear (align_corners=True) on channels-last
// maps, patch-embed im2col, class-token rows, bicubic resampling of the position grid.
#include "common.cuh"

namespace {

// ATen upsample_bilinear2d semantics (align_corners=True): src = dst * (in-1)/(out-1), lambda in fp32,
// value = h0*(w0*a + w1*b) + h1*(w0*c + w1*d), computed in fp32 and rounded once to the storage type.
template <typename T, typename TO>
__global__ void __launch_bounds__(256) bilinear_kernel(const T* __restrict__ in, TO* __restrict__ out, int ih, int iw,
                                                      int ipad, int oh, int ow, int opad, int C, float rh, float rw,
                                                      int rows) {
  // grid = (x-chunks, oh / rows, frames): the row quantities are block-uniform and no 64-bit division is needed.
  // A thread walks `rows` consecutive output rows of its column and keeps the two horizontally interpolated input rows
  // (w0*a + w1*b and w0*c + w1*d, unpacked) in registers: the upper one is re-used while the source row index does not
  // move and the lower one becomes the upper one when it moves by one, so a 2x upsample loads and interpolates one new
  // input row per two output rows.  One output row per block re-read its input rows from L2 (the 148 -> 296 upsample,
  // 1.4 GB out, sat at the L2 bandwidth); four fresh corner loads per output row then sat at the issue rate (158
  // instructions per eight channels, profiles/ncu_bilinear_r01m.txt).
  const int nv = C >> 3;
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= ow * nv) return;
  const int ox = idx / nv, cv = idx - ox * nv;
  const int f = blockIdx.z;
  const float w1r = rw * ox;
  const int w1 = (int)w1r;
  const int w1p = (w1 < iw - 1) ? 1 : 0;
  const float w1l = w1r - w1, w0l = 1.f - w1l;
  const int IW = iw + 2 * ipad, IH = ih + 2 * ipad;
  const int OW = ow + 2 * opad, OH = oh + 2 * opad;
  const T* col = in + (((long long)f * IH + ipad) * IW + w1 + ipad) * C + cv * 8;
  TO* ocol = out + (((long long)f * OH + opad) * OW + ox + opad) * C + cv * 8;
  const long long irow = (long long)IW * C, orow = (long long)OW * C;
  const int step = w1p * C;
  // channel pairs as packed fp32 (FMUL2 / FFMA2: half the issue slots of the sixteen multiply-adds per row)
  uint64_t tp[4], bt[4];
  float a[8], b[8], o[8];
  const uint64_t w0l2 = pk2(w0l, w0l), w1l2 = pk2(w1l, w1l);
  int top = -2, bot = -2;  // input rows whose horizontal interpolation sits in tp and bt
  const int oy0 = blockIdx.y * rows, oy1 = min(oy0 + rows, oh);
  for (int oy = oy0; oy < oy1; ++oy) {
    const float h1r = rh * oy;
    const int h1 = (int)h1r;
    const int h1b = (h1 < ih - 1) ? h1 + 1 : h1;
    const float h1l = h1r - h1, h0l = 1.f - h1l;
    if (h1 != top) {
      if (h1 == bot) {
#pragma unroll
        for (int i = 0; i < 4; ++i) tp[i] = bt[i];
      } else {
        const T* p = col + h1 * irow;
        Vec8<T>::load(p, a);
        Vec8<T>::load(p + step, b);
#pragma unroll
        for (int i = 0; i < 4; ++i) tp[i] = fma2(w1l2, pk2(b[2 * i], b[2 * i + 1]), mul2(w0l2, pk2(a[2 * i], a[2 * i + 1])));
      }
      top = h1;
    }
    if (h1b != bot) {
      if (h1b == top) {
#pragma unroll
        for (int i = 0; i < 4; ++i) bt[i] = tp[i];
      } else {
        const T* p = col + h1b * irow;
        Vec8<T>::load(p, a);
        Vec8<T>::load(p + step, b);
#pragma unroll
        for (int i = 0; i < 4; ++i) bt[i] = fma2(w1l2, pk2(b[2 * i], b[2 * i + 1]), mul2(w0l2, pk2(a[2 * i], a[2 * i + 1])));
      }
      bot = h1b;
    }
    const uint64_t h0l2 = pk2(h0l, h0l), h1l2 = pk2(h1l, h1l);
#pragma unroll
    for (int i = 0; i < 4; ++i) upk2(fma2(h1l2, bt[i], mul2(h0l2, tp[i])), o[2 * i], o[2 * i + 1]);
    Vec8<TO>::store(ocol + oy * orow, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) im2col_patch_kernel(const float* __restrict__ x, const float* const* __restrict__ fptr,
                                                          T* __restrict__ A, int H, int W, int kpad, long long total) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int k = (int)(idx % kpad);
  const long long m = idx / kpad;
  const int pw = W / 14, ph = H / 14;
  const int px = (int)(m % pw);
  const int py = (int)((m / pw) % ph);
  const int f = (int)(m / ((long long)pw * ph));
  float v = 0.f;
  if (k < 588) {
    const int c = k / 196, r = k % 196, i = r / 14, j = r % 14;
    const float* img = fptr ? fptr[f] : x + (long long)f * 3 * H * W;  // per-frame pointers: the driver's window gather
    v = img[((long long)c * H + py * 14 + i) * W + px * 14 + j];
  }
  A[idx] = from_f32<T>(v);
}

// bf16 fast path: one thread per 14-pixel patch row (56 contiguous input bytes as seven float2, 28 output bytes as seven
// bf16 pairs), 42 such rows per patch plus one thread for the zero tail [588, kpad); consecutive threads write
// consecutive pieces of the same A row.  The element-per-thread form above moved 2 bytes per store instruction.
__global__ void __launch_bounds__(256) im2col_patch_rows_kernel(const float* __restrict__ x, const float* const* __restrict__ fptr,
                                                               bf16* __restrict__ A, int H, int W, int kpad, long long tasks) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= tasks) return;
  const int r = (int)(idx % 43);   // 0..41: (channel, patch row); 42: the padding tail
  const long long m = idx / 43;
  bf16* out = A + m * kpad;
  if (r == 42) {
    uint32_t* o = reinterpret_cast<uint32_t*>(out + 588);
    for (int i = 0; i < (kpad - 588) / 2; ++i) o[i] = 0u;
    return;
  }
  const int pw = W / 14, ph = H / 14;
  const int px = (int)(m % pw);
  const int py = (int)((m / pw) % ph);
  const int f = (int)(m / ((long long)pw * ph));
  const int c = r / 14, i = r - c * 14;
  const float* img = fptr ? fptr[f] : x + (long long)f * 3 * H * W;
  const float2* in = reinterpret_cast<const float2*>(img + ((long long)c * H + py * 14 + i) * W + px * 14);
  uint32_t* o = reinterpret_cast<uint32_t*>(out + r * 14);
  float2 v[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) v[j] = in[j];
#pragma unroll
  for (int j = 0; j < 7; ++j) o[j] = pack_bf16x2(v[j].x, v[j].y);
}

__global__ void cls_rows_kernel(float* __restrict__ resid, const float* __restrict__ cls, const float* __restrict__ pos,
                                int frames, int ntok, int D) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= frames * D) return;
  const int f = idx / D, c = idx % D;
  resid[(long long)f * ntok * D + c] = __fadd_rn(cls[c], pos[c]);
}

// ATen upsample_bicubic2d (align_corners=False, scale_factor given, A = -0.75): src = (dst + 0.5)/scale - 0.5,
// taps floor(src)-1..+2 with clamped indices.
__device__ __forceinline__ float cubic1(float x, float A) { return ((A + 2) * x - (A + 3)) * x * x + 1; }
__device__ __forceinline__ float cubic2(float x, float A) { return ((A * x - 5 * A) * x + 8 * A) * x - 4 * A; }
__device__ __forceinline__ void cubic_coeffs(float t, float* c) {
  const float A = -0.75f;
  c[0] = cubic2(t + 1.f, A);
  c[1] = cubic1(t, A);
  c[2] = cubic1(1.f - t, A);
  c[3] = cubic2(2.f - t, A);
}
__global__ void pos_bicubic_kernel(const float* __restrict__ pos, float* __restrict__ out, int D, int src, int oh, int ow,
                                   float inv_sh, float inv_sw) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)oh * ow * D) return;
  const int c = (int)(idx % D);
  const int ox = (int)((idx / D) % ow), oy = (int)(idx / ((long long)D * ow));
  const float sy = inv_sh * (oy + 0.5f) - 0.5f, sx = inv_sw * (ox + 0.5f) - 0.5f;
  const int iy = (int)floorf(sy), ix = (int)floorf(sx);
  float cy[4], cx[4];
  cubic_coeffs(sy - iy, cy);
  cubic_coeffs(sx - ix, cx);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int yy = min(max(iy - 1 + i, 0), src - 1);
    float row = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int xx = min(max(ix - 1 + j, 0), src - 1);
      row += pos[((long long)(1 + yy * src + xx)) * D + c] * cx[j];
    }
    acc += row * cy[i];
  }
  out[(long long)(1 + oy * ow + ox) * D + c] = acc;
  if (idx < D) out[c] = pos[c];  // class-token position row is copied through
}

}  // namespace

extern "C" int vda_bilinear(const void* in, int32_t in_dtype, void* out, int32_t out_dtype, int32_t frames, int32_t ih,
                            int32_t iw, int32_t in_padded, int32_t oh, int32_t ow, int32_t out_padded, int32_t C,
                            vda_stream_t stream) {
  VDA_CHECK_ARG(C % 8 == 0, "vda_bilinear: C=%d must be a multiple of 8", C);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const float rh = oh > 1 ? (float)(ih - 1) / (float)(oh - 1) : 0.f;
  const float rw = ow > 1 ? (float)(iw - 1) / (float)(ow - 1) : 0.f;
  VDA_CHECK_ARG(oh <= 65535 && frames <= 65535, "vda_bilinear: oh / frames exceed the grid limits");
  static const int rows_env = [] { const char* e = getenv("VDA_BILINEAR_ROWS"); return e ? atoi(e) : 0; }();
  const int rows = rows_env > 0 ? rows_env : 16;  // output rows per thread (4: 437 us, 8: 362, 16: 344, 37: 338 on 148 -> 296)
  const dim3 grid((unsigned)((ow * (C / 8) + 255) / 256), (unsigned)((oh + rows - 1) / rows), (unsigned)frames);
  const int ip = in_padded ? 1 : 0, op = out_padded ? 1 : 0;
#define BL(TI, TO) bilinear_kernel<TI, TO><<<grid, 256, 0, st>>>((const TI*)in, (TO*)out, ih, iw, ip, oh, ow, op, C, rh, rw, rows)
  if (in_dtype == VDA_BF16 && out_dtype == VDA_BF16) BL(bf16, bf16);
  else if (in_dtype == VDA_BF16 && out_dtype == VDA_F32) BL(bf16, float);
  else if (in_dtype == VDA_F32 && out_dtype == VDA_F32) BL(float, float);
  else VDA_CHECK_ARG(false, "vda_bilinear: unsupported dtypes %d -> %d", in_dtype, out_dtype);
#undef BL
  VDA_CUDA_LAUNCH_CHECK("bilinear_kernel");
  return 0;
}

static int im2col_patch_launch(const float* x, const float* const* fptr, bool aligned8, void* A, int32_t a_dtype, int32_t frames,
                               int32_t H, int32_t W, int32_t kpad, cudaStream_t st) {
  const long long total = (long long)frames * (H / 14) * (W / 14) * kpad;
  const unsigned grid = (unsigned)((total + 255) / 256);
  if (a_dtype == VDA_BF16 && W % 2 == 0 && kpad % 2 == 0 && kpad >= 588 && aligned8 && ((uintptr_t)A % 4) == 0) {
    const long long tasks = (long long)frames * (H / 14) * (W / 14) * 43;
    im2col_patch_rows_kernel<<<(unsigned)((tasks + 255) / 256), 256, 0, st>>>(x, fptr, (bf16*)A, H, W, kpad, tasks);
  } else if (a_dtype == VDA_BF16) im2col_patch_kernel<bf16><<<grid, 256, 0, st>>>(x, fptr, (bf16*)A, H, W, kpad, total);
  else im2col_patch_kernel<float><<<grid, 256, 0, st>>>(x, fptr, (float*)A, H, W, kpad, total);
  VDA_CUDA_LAUNCH_CHECK("im2col_patch_kernel");
  return 0;
}

extern "C" int vda_im2col_patch(const float* x, void* A, int32_t a_dtype, int32_t frames, int32_t H, int32_t W,
                                int32_t kpad, vda_stream_t stream) {
  VDA_CHECK_ARG(H % 14 == 0 && W % 14 == 0, "vda_im2col_patch: H, W must be multiples of 14 (patch_embed.py:73-74)");
  VDA_CHECK_ARG(kpad >= 588, "vda_im2col_patch: kpad < 588");
  return im2col_patch_launch(x, nullptr, ((uintptr_t)x % 8) == 0, A, a_dtype, frames, H, W, kpad,
                             reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int vda_im2col_patch_frames(const float* const* frame_ptrs, void* A, int32_t a_dtype, int32_t frames, int32_t H,
                                       int32_t W, int32_t kpad, vda_stream_t stream) {
  VDA_CHECK_ARG(H % 14 == 0 && W % 14 == 0, "vda_im2col_patch_frames: H, W must be multiples of 14 (patch_embed.py:73-74)");
  VDA_CHECK_ARG(kpad >= 588 && frame_ptrs != nullptr, "vda_im2col_patch_frames: kpad < 588 or null pointer table");
  // the pointer table lives on the device (it changes per window under a replayed CUDA graph), so alignment of the frames
  // is the caller's contract: 8-byte aligned planar [3, H, W] fp32 images
  return im2col_patch_launch(nullptr, frame_ptrs, true, A, a_dtype, frames, H, W, kpad, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int vda_cls_rows(float* resid, const float* cls, const float* pos, int32_t frames, int32_t ntok, int32_t D,
                            vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  cls_rows_kernel<<<(frames * D + 255) / 256, 256, 0, st>>>(resid, cls, pos, frames, ntok, D);
  VDA_CUDA_LAUNCH_CHECK("cls_rows_kernel");
  return 0;
}

extern "C" int vda_pos_embed_bicubic(const float* pos, float* out, int32_t D, int32_t src, int32_t oh, int32_t ow,
                                     double scale_h, double scale_w, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)oh * ow * D;
  pos_bicubic_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(pos, out, D, src, oh, ow, (float)(1.0 / scale_h),
                                                                      (float)(1.0 / scale_w));
  VDA_CUDA_LAUNCH_CHECK("pos_bicubic_kernel");
  return 0;
}
