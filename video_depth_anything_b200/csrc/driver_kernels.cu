// Kernels of the sliding-window driver (video_depth.py:329-417): frame pre-processing, the resize of the depth maps
// back to frame size, the scale/shift least squares, scale+clamp and the 8-frame cross-fade.  All HBM-bound.
#include "common.cuh"

namespace {

// cv2.resize(INTER_CUBIC) on float32 HWC (transform.py:113): src = (dst+0.5)*scale-0.5, taps floor-1..+2 with replicate
// border, coefficients with A=-0.75 and c3 = 1-c0-c1-c2, horizontal pass then vertical pass.
__device__ __forceinline__ void cv_cubic_coeffs(float x, float* c) {
  const float A = -0.75f;
  c[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
  c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
  c[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
  c[3] = 1.f - c[0] - c[1] - c[2];
}

__global__ void __launch_bounds__(256) preprocess_kernel(const uint8_t* __restrict__ frames, float* __restrict__ out, int fh,
                                                        int fw, int oh, int ow, double sy, double sx, long long total) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int x = (int)(idx % ow);
  const int y = (int)((idx / ow) % oh);
  const int f = (int)(idx / ((long long)ow * oh));
  const uint8_t* img = frames + (long long)f * fh * fw * 3;
  float v[3];
  if (oh == fh && ow == fw) {
#pragma unroll
    for (int c = 0; c < 3; ++c) v[c] = (float)img[((long long)y * fw + x) * 3 + c] / 255.0f;
  } else {
    float fy = (float)((y + 0.5) * sy - 0.5), fx = (float)((x + 0.5) * sx - 0.5);
    const int iy = (int)floorf(fy), ix = (int)floorf(fx);
    fy -= iy;
    fx -= ix;
    float cy[4], cx[4];
    cv_cubic_coeffs(fy, cy);
    cv_cubic_coeffs(fx, cx);
    v[0] = v[1] = v[2] = 0.f;
    float rows[4][3];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int yy = min(max(iy - 1 + i, 0), fh - 1);
      float p[4][3];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int xx = min(max(ix - 1 + j, 0), fw - 1);
        const uint8_t* px = img + ((long long)yy * fw + xx) * 3;
#pragma unroll
        for (int c = 0; c < 3; ++c) p[j][c] = (float)px[c] / 255.0f;
      }
#pragma unroll
      for (int c = 0; c < 3; ++c) rows[i][c] = p[0][c] * cx[0] + p[1][c] * cx[1] + p[2][c] * cx[2] + p[3][c] * cx[3];
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) v[c] = rows[0][c] * cy[0] + rows[1][c] * cy[1] + rows[2][c] * cy[2] + rows[3][c] * cy[3];
  }
  // NormalizeImage promotes to float64 (python-list mean/std), PrepareForNet casts back to f32 (transform.py:133-148)
  const double mean[3] = {0.485, 0.456, 0.406}, stdv[3] = {0.229, 0.224, 0.225};
#pragma unroll
  for (int c = 0; c < 3; ++c)
    out[(((long long)f * 3 + c) * oh + y) * ow + x] = (float)(((double)v[c] - mean[c]) / stdv[c]);
}

template <typename T>
__global__ void __launch_bounds__(256) resize_depth_kernel(const T* __restrict__ in, float* __restrict__ out, int ih, int iw,
                                                          int oh, int ow, float rh, float rw, long long total) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int ox = (int)(idx % ow);
  const int oy = (int)((idx / ow) % oh);
  const int f = (int)(idx / ((long long)ow * oh));
  const float h1r = rh * oy;
  const int h1 = (int)h1r, h1p = (h1 < ih - 1) ? 1 : 0;
  const float h1l = h1r - h1, h0l = 1.f - h1l;
  const float w1r = rw * ox;
  const int w1 = (int)w1r, w1p = (w1 < iw - 1) ? 1 : 0;
  const float w1l = w1r - w1, w0l = 1.f - w1l;
  const T* p = in + ((long long)f * ih + h1) * iw + w1;
  const float a = to_f32<T>(p[0]), b = to_f32<T>(p[w1p]), c = to_f32<T>(p[(long long)h1p * iw]),
              d = to_f32<T>(p[(long long)h1p * iw + w1p]);
  out[idx] = h0l * (w0l * a + w1l * b) + h1l * (w0l * c + w1l * d);
}

__global__ void __launch_bounds__(256) ls_sums_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                     long long count, double* __restrict__ sums) {
  double s[4] = {0, 0, 0, 0};  // sum p*p, sum p, sum p*t, sum t
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
    const double p = pred[i], t = tgt[i];
    s[0] += p * p; s[1] += p; s[2] += p * t; s[3] += t;
  }
  __shared__ double sh[4][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
    if (lane == 0) sh[k][warp] = s[k];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0;
    for (int w = 0; w < 8; ++w) t += sh[threadIdx.x][w];
    atomicAdd(&sums[threadIdx.x], t);
  }
  if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(&sums[4], (double)count);
}

__global__ void ls_solve_kernel(const double* __restrict__ sums, float* __restrict__ st) {
  const double a00 = sums[0], a01 = sums[1], b0 = sums[2], b1 = sums[3], a11 = sums[4];
  const double det = a00 * a11 - a01 * a01;
  double x0 = 1.0, x1 = 0.0;
  if (det != 0.0) {
    x0 = (a11 * b0 - a01 * b1) / det;
    x1 = (-a01 * b0 + a00 * b1) / det;
  }
  st[0] = (float)x0;
  st[1] = (float)x1;
}

__global__ void __launch_bounds__(256) scale_clamp_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                         const float* __restrict__ st, long long count) {
  const float s = st[0], t = st[1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x)
    out[i] = fmaxf(__fadd_rn(__fmul_rn(in[i], s), t), 0.f);
}

// ---------------------------------------------------------------- evaluation aligner + metrics
// utils/align.py:151-160,202-226: least squares of [p, 1] c = 1/gt over the pixels that are valid and have gt != 0
// (the reference inverts the masked ground truth, which masks its zeros), scale = 1/c0, shift = -c1/c0.
template <int NS, typename F>
__device__ __forceinline__ void block_sums_to(double* s, double* out, F) {
  __shared__ double sh[NS][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < NS; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
    if (lane == 0) sh[k][warp] = s[k];
  }
  __syncthreads();
  if (threadIdx.x < NS) {
    double t = 0;
    for (int w = 0; w < 8; ++w) t += sh[threadIdx.x][w];
    atomicAdd(&out[threadIdx.x], t);
  }
}

__global__ void __launch_bounds__(256) eval_fit_sums_kernel(const float* __restrict__ pred, const float* __restrict__ gt,
                                                           const uint8_t* __restrict__ valid, long long count,
                                                           double* __restrict__ sums) {
  double s[5] = {0, 0, 0, 0, 0};  // sum p*p, sum p, sum p*g, sum g, n
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
    const float g0 = gt[i];
    if (valid[i] && g0 != 0.f) {
      const double p = pred[i], g = 1.0 / (double)g0;  // `1 / masked float32 array` is a float64 array in numpy
      s[0] += p * p; s[1] += p; s[2] += p * g; s[3] += g; s[4] += 1.0;
    }
  }
  block_sums_to<5>(s, sums, 0);
}

__global__ void eval_fit_solve_kernel(const double* __restrict__ sums, double* __restrict__ coef) {
  const double a00 = sums[0], a01 = sums[1], b0 = sums[2], b1 = sums[3], a11 = sums[4];
  const double det = a00 * a11 - a01 * a01;
  double c0 = 0.0, c1 = 0.0;
  if (det != 0.0) {
    c0 = (a11 * b0 - a01 * b1) / det;
    c1 = (-a01 * b0 + a00 * b1) / det;
  }
  // float64 throughout, like the reference's lstsq on the float64 target (utils/align.py:155-159)
  if (fabs(c0) <= 0.0) { coef[0] = (double)INFINITY; coef[1] = 0.0; }   // :156-157
  else { coef[0] = 1.0 / c0; coef[1] = -c1 / c0; }
}

// utils/align.py:228-236 + utils/metrics.py:92-192: aligned metric depth of every pixel and the seven error sums over
// the valid ones.  sums: n, |d-g|, |d-g|/g, (d-g)/g, (d-g)^2, and the outlier counts max(d/g, g/d) > 1.25, 1.25^2, 1.25^3.
__global__ void __launch_bounds__(256) eval_apply_kernel(const float* __restrict__ pred, const float* __restrict__ gt,
                                                        const uint8_t* __restrict__ valid, const double* __restrict__ coef,
                                                        double max_depth, float* __restrict__ aligned, long long count,
                                                        double* __restrict__ sums) {
  // the reference's aligned map is a float64 array ((float32 - float64 scalar) / float64 scalar): the arithmetic below
  // is double, the stored map is its float32 rounding
  const double scale = coef[0], shift = coef[1];
  const double t1 = 1.25, t2 = 1.25 * 1.25, t3 = 1.25 * 1.25 * 1.25;
  double s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x) {
    double a = ((double)pred[i] - shift) / scale;
    a = fmin(fmax(a, 0.0), 1.0);
    if (a == 0.0) a = 1e-4;
    const double d = fmin(fmax(1.0 / a, 0.0), max_depth);
    if (aligned) aligned[i] = (float)d;
    if (valid[i]) {
      const double g = gt[i];
      const double diff = d - g;
      const double r = fmax(d / g, g / d);
      s[0] += 1.0;
      s[1] += fabs(diff);
      s[2] += fabs(diff) / g;
      s[3] += diff / g;
      s[4] += diff * diff;
      s[5] += r > t1 ? 1.0 : 0.0;
      s[6] += r > t2 ? 1.0 : 0.0;
      s[7] += r > t3 ? 1.0 : 0.0;
    }
  }
  block_sums_to<8>(s, sums, 0);
}

// streaming mode (video_depth.py:313): cur_prediction * scale + shift, NOT clamped
__global__ void __launch_bounds__(256) scale_apply_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                         const float* __restrict__ st, long long count) {
  const float s = st[0], t = st[1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (long long)gridDim.x * blockDim.x)
    out[i] = __fadd_rn(__fmul_rn(in[i], s), t);
}

// Frame gather / scatter for the streaming mode's feature ring (video_depth.py:228-231,244-247,281-296):
// dst frame dst_idx[i] <- src frame src_idx[i]; frames are `vecs` 16-byte vectors; NULL index arrays mean i.
__global__ void __launch_bounds__(256) copy_frames_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                                         const int* __restrict__ src_idx, const int* __restrict__ dst_idx,
                                                         long long vecs) {
  const int f = blockIdx.y;
  const uint4* a = src + (long long)(src_idx ? src_idx[f] : f) * vecs;
  uint4* b = dst + (long long)(dst_idx ? dst_idx[f] : f) * vecs;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < vecs; i += (long long)gridDim.x * blockDim.x)
    b[i] = a[i];
}

__global__ void __launch_bounds__(256) blend_kernel(float* __restrict__ inout, const float* __restrict__ post,
                                                   const float* __restrict__ st, int n, long long frame_elems) {
  const float s = st[0], t = st[1];
  const int f = blockIdx.y;
  // python-float weights (utils/util.py:69-70): 0, i*(1/(n-1)), ..., 1, applied to f32 arrays (stay f32)
  const double step = 1.0 / (double)(n - 1);
  const double wd = (f == 0) ? 0.0 : (f == n - 1 ? 1.0 : f * step);
  const float w1 = (float)wd, w0 = (float)(1.0 - wd);
  float* a = inout + (long long)f * frame_elems;
  const float* b = post + (long long)f * frame_elems;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < frame_elems;
       i += (long long)gridDim.x * blockDim.x) {
    const float pv = fmaxf(__fadd_rn(__fmul_rn(b[i], s), t), 0.f);
    a[i] = __fadd_rn(__fmul_rn(a[i], w0), __fmul_rn(pv, w1));
  }
}


// Deterministic form of the least-squares sums (the window-sharded driver evaluates the (scale, shift) recurrence
// redundantly on every rank and needs bit-identical results everywhere, and equal to a single-rank run): block c sums
// elements [c*LS_CHUNK, (c+1)*LS_CHUNK) in a fixed order (per-thread strided sums, shuffle tree, warp order) into
// partial[c][0..3]; one block then adds the partials in a fixed order (thread t takes c = t, t+256, ...; shuffle tree;
// warp order) and solves.  No atomics, and no dependence on the grid size or the SM count.
constexpr int LS_CHUNK = 4096;

__global__ void __launch_bounds__(256) ls_partial_kernel(const float* __restrict__ pred, const float* __restrict__ tgt,
                                                        long long count, double* __restrict__ partial) {
  double s[4] = {0, 0, 0, 0};  // sum p*p, sum p, sum p*t, sum t
  const long long base = (long long)blockIdx.x * LS_CHUNK;
  const long long end = min(count, base + LS_CHUNK);
  for (long long i = base + threadIdx.x; i < end; i += 256) {
    const double p = pred[i], t = tgt[i];
    s[0] += p * p; s[1] += p; s[2] += p * t; s[3] += t;
  }
  __shared__ double sh[4][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
    if (lane == 0) sh[k][warp] = s[k];
  }
  __syncthreads();
  if (threadIdx.x < 4) {
    double t = 0;
    for (int w = 0; w < 8; ++w) t += sh[threadIdx.x][w];
    partial[(long long)blockIdx.x * 4 + threadIdx.x] = t;
  }
}

__global__ void __launch_bounds__(256) ls_final_kernel(const double* __restrict__ partial, int chunks, long long count,
                                                      float* __restrict__ st) {
  double s[4] = {0, 0, 0, 0};
  for (int c = threadIdx.x; c < chunks; c += 256) {
#pragma unroll
    for (int k = 0; k < 4; ++k) s[k] += partial[(long long)c * 4 + k];
  }
  __shared__ double sh[4][8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
    if (lane == 0) sh[k][warp] = s[k];
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t[4];
    for (int k = 0; k < 4; ++k) {
      t[k] = 0;
      for (int w = 0; w < 8; ++w) t[k] += sh[k][w];
    }
    const double a00 = t[0], a01 = t[1], b0 = t[2], b1 = t[3], a11 = (double)count;
    const double det = a00 * a11 - a01 * a01;
    double x0 = 1.0, x1 = 0.0;
    if (det != 0.0) {
      x0 = (a11 * b0 - a01 * b1) / det;
      x1 = (-a01 * b0 + a00 * b1) / det;
    }
    st[0] = (float)x0;
    st[1] = (float)x1;
  }
}

// Cross-fade of two RAW tails (window-sharded driver): the previous window's last frames arrive un-aligned from their
// owner, so both sides are aligned here.  Same fp32 operations, in the same order, as scale_clamp_kernel followed by
// blend_kernel (bit-identical to the single-buffer form).  st_pre == NULL: `pre` is used as it is (window 0's frames are
// never scaled or clamped, video_depth.py:382-387).
__global__ void __launch_bounds__(256) blend_raw_kernel(const float* __restrict__ pre, const float* __restrict__ st_pre,
                                                       const float* __restrict__ post, const float* __restrict__ st_post,
                                                       float* __restrict__ out, int n, long long frame_elems) {
  const float s1 = st_post[0], t1 = st_post[1];
  const bool has0 = st_pre != nullptr;
  const float s0 = has0 ? st_pre[0] : 1.f, t0 = has0 ? st_pre[1] : 0.f;
  const int f = blockIdx.y;
  const double step = 1.0 / (double)(n - 1);
  const double wd = (f == 0) ? 0.0 : (f == n - 1 ? 1.0 : f * step);
  const float w1 = (float)wd, w0 = (float)(1.0 - wd);
  const float* a = pre + (long long)f * frame_elems;
  const float* b = post + (long long)f * frame_elems;
  float* o = out + (long long)f * frame_elems;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < frame_elems;
       i += (long long)gridDim.x * blockDim.x) {
    const float av = has0 ? fmaxf(__fadd_rn(__fmul_rn(a[i], s0), t0), 0.f) : a[i];
    const float pv = fmaxf(__fadd_rn(__fmul_rn(b[i], s1), t1), 0.f);
    o[i] = __fadd_rn(__fmul_rn(av, w0), __fmul_rn(pv, w1));
  }
}

}  // namespace

extern "C" int vda_preprocess_u8(const uint8_t* frames, float* out, int32_t n, int32_t fh, int32_t fw, int32_t oh,
                                 int32_t ow, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const long long total = (long long)n * oh * ow;
  // cv2: inv_scale = dsize/ssize (double), scale = 1/inv_scale
  const double sx = 1.0 / ((double)ow / (double)fw), sy = 1.0 / ((double)oh / (double)fh);
  preprocess_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(frames, out, fh, fw, oh, ow, sy, sx, total);
  VDA_CUDA_LAUNCH_CHECK("preprocess_kernel");
  return 0;
}

extern "C" int vda_resize_depth(const void* in, int32_t dtype, float* out, int32_t n, int32_t ih, int32_t iw, int32_t oh,
                                int32_t ow, vda_stream_t stream) {
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const float rh = oh > 1 ? (float)(ih - 1) / (float)(oh - 1) : 0.f;
  const float rw = ow > 1 ? (float)(iw - 1) / (float)(ow - 1) : 0.f;
  const long long total = (long long)n * oh * ow;
  const unsigned grid = (unsigned)((total + 255) / 256);
  if (dtype == VDA_BF16) resize_depth_kernel<bf16><<<grid, 256, 0, st>>>((const bf16*)in, out, ih, iw, oh, ow, rh, rw, total);
  else resize_depth_kernel<float><<<grid, 256, 0, st>>>((const float*)in, out, ih, iw, oh, ow, rh, rw, total);
  VDA_CUDA_LAUNCH_CHECK("resize_depth_kernel");
  return 0;
}

extern "C" int vda_scale_shift(const float* pred, const float* target, int64_t count, double* scratch, float* st,
                               int32_t phase, vda_stream_t stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (phase & 1) {
    cudaError_t ce = cudaMemsetAsync(scratch, 0, 5 * sizeof(double), s);
    if (ce != cudaSuccess) {
      vda_set_error("vda_scale_shift: memset failed: %s", cudaGetErrorString(ce));
      return (int)ce;
    }
  }
  int blocks = (int)((count + 256 * 8 - 1) / (256 * 8));
  if (blocks > 4 * vda_sm_count()) blocks = 4 * vda_sm_count();
  if (blocks < 1) blocks = 1;
  ls_sums_kernel<<<blocks, 256, 0, s>>>(pred, target, count, scratch);
  if (phase & 2) ls_solve_kernel<<<1, 1, 0, s>>>(scratch, st);
  VDA_CUDA_LAUNCH_CHECK("scale_shift kernels");
  return 0;
}

extern "C" int vda_scale_clamp(const float* in, float* out, const float* st, int64_t count, vda_stream_t stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int blocks = (int)((count + 256 * 4 - 1) / (256 * 4));
  if (blocks > 8 * vda_sm_count()) blocks = 8 * vda_sm_count();
  if (blocks < 1) blocks = 1;
  scale_clamp_kernel<<<blocks, 256, 0, s>>>(in, out, st, count);
  VDA_CUDA_LAUNCH_CHECK("scale_clamp_kernel");
  return 0;
}

extern "C" int vda_eval_fit(const float* pred, const float* gt, const uint8_t* valid, int64_t count, double* scratch,
                            double* coef, int32_t phase, vda_stream_t stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (phase & 1) {
    cudaError_t ce = cudaMemsetAsync(scratch, 0, 5 * sizeof(double), s);
    if (ce != cudaSuccess) {
      vda_set_error("vda_eval_fit: memset failed: %s", cudaGetErrorString(ce));
      return (int)ce;
    }
  }
  int blocks = (int)((count + 256 * 8 - 1) / (256 * 8));
  if (blocks > 4 * vda_sm_count()) blocks = 4 * vda_sm_count();
  if (blocks < 1) blocks = 1;
  if (count > 0) eval_fit_sums_kernel<<<blocks, 256, 0, s>>>(pred, gt, valid, count, scratch);
  if (phase & 2) eval_fit_solve_kernel<<<1, 1, 0, s>>>(scratch, coef);
  VDA_CUDA_LAUNCH_CHECK("eval_fit kernels");
  return 0;
}

extern "C" int vda_eval_apply(const float* pred, const float* gt, const uint8_t* valid, const double* coef, double max_depth,
                              float* aligned, int64_t count, double* sums, vda_stream_t stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  cudaError_t ce = cudaMemsetAsync(sums, 0, 8 * sizeof(double), s);
  if (ce != cudaSuccess) {
    vda_set_error("vda_eval_apply: memset failed: %s", cudaGetErrorString(ce));
    return (int)ce;
  }
  if (count <= 0) return 0;
  int blocks = (int)((count + 256 * 8 - 1) / (256 * 8));
  if (blocks > 4 * vda_sm_count()) blocks = 4 * vda_sm_count();
  if (blocks < 1) blocks = 1;
  eval_apply_kernel<<<blocks, 256, 0, s>>>(pred, gt, valid, coef, max_depth, aligned, count, sums);
  VDA_CUDA_LAUNCH_CHECK("eval_apply_kernel");
  return 0;
}

extern "C" int vda_scale_apply(const float* in, float* out, const float* st, int64_t count, vda_stream_t stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int blocks = (int)((count + 256 * 4 - 1) / (256 * 4));
  if (blocks > 8 * vda_sm_count()) blocks = 8 * vda_sm_count();
  if (blocks < 1) blocks = 1;
  scale_apply_kernel<<<blocks, 256, 0, s>>>(in, out, st, count);
  VDA_CUDA_LAUNCH_CHECK("scale_apply_kernel");
  return 0;
}

extern "C" int vda_copy_frames(const void* src, void* dst, const int32_t* src_idx, const int32_t* dst_idx, int32_t n,
                               int64_t frame_bytes, vda_stream_t stream) {
  VDA_CHECK_ARG(frame_bytes % 16 == 0, "vda_copy_frames: frame_bytes must be a multiple of 16");
  VDA_CHECK_ARG(((uintptr_t)src | (uintptr_t)dst) % 16 == 0, "vda_copy_frames: pointers must be 16-byte aligned");
  if (n <= 0 || frame_bytes == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long vecs = frame_bytes / 16;
  long long bx = (vecs + 256 * 4 - 1) / (256 * 4);
  const long long cap = (8LL * vda_sm_count() + n - 1) / n;
  if (bx > cap) bx = cap;
  if (bx < 1) bx = 1;
  copy_frames_kernel<<<dim3((unsigned)bx, (unsigned)n), 256, 0, s>>>(reinterpret_cast<const uint4*>(src),
                                                                   reinterpret_cast<uint4*>(dst), src_idx, dst_idx, vecs);
  VDA_CUDA_LAUNCH_CHECK("copy_frames_kernel");
  return 0;
}

extern "C" int vda_blend(float* inout, const float* post, const float* st, int32_t n, int64_t frame_elems,
                         vda_stream_t stream) {
  VDA_CHECK_ARG(n >= 2, "vda_blend: n must be >= 2");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int bx = (int)((frame_elems + 256 * 4 - 1) / (256 * 4));
  if (bx > 2 * vda_sm_count()) bx = 2 * vda_sm_count();
  if (bx < 1) bx = 1;
  blend_kernel<<<dim3(bx, n), 256, 0, s>>>(inout, post, st, n, frame_elems);
  VDA_CUDA_LAUNCH_CHECK("blend_kernel");
  return 0;
}

extern "C" int64_t vda_scale_shift_det_workspace_doubles(int64_t count) {
  return count <= 0 ? 4 : ((count + LS_CHUNK - 1) / LS_CHUNK) * 4;
}

extern "C" int vda_scale_shift_det(const float* pred, const float* target, int64_t count, double* partial, float* st,
                                   vda_stream_t stream) {
  VDA_CHECK_ARG(count > 0 && count <= (int64_t)LS_CHUNK * 0x7fffffff, "vda_scale_shift_det: bad count %lld", (long long)count);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long chunks = (count + LS_CHUNK - 1) / LS_CHUNK;
  ls_partial_kernel<<<(unsigned)chunks, 256, 0, s>>>(pred, target, count, partial);
  ls_final_kernel<<<1, 256, 0, s>>>(partial, (int)chunks, count, st);
  VDA_CUDA_LAUNCH_CHECK("scale_shift_det kernels");
  return 0;
}

extern "C" int vda_blend_raw(const float* pre, const float* st_pre, const float* post, const float* st_post, float* out,
                             int32_t n, int64_t frame_elems, vda_stream_t stream) {
  VDA_CHECK_ARG(n >= 2, "vda_blend_raw: n must be >= 2");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int bx = (int)((frame_elems + 256 * 4 - 1) / (256 * 4));
  if (bx > 2 * vda_sm_count()) bx = 2 * vda_sm_count();
  if (bx < 1) bx = 1;
  blend_raw_kernel<<<dim3(bx, n), 256, 0, s>>>(pre, st_pre, post, st_post, out, n, frame_elems);
  VDA_CUDA_LAUNCH_CHECK("blend_raw_kernel");
  return 0;
}

extern "C" int vda_copy_device(const void* src, void* dst, int64_t bytes, vda_stream_t stream) {
  if (bytes <= 0) return 0;
  cudaError_t e = cudaMemcpyAsync(dst, src, (size_t)bytes, cudaMemcpyDeviceToDevice, reinterpret_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) {
    vda_set_error("vda_copy_device: cudaMemcpyAsync failed: %s", cudaGetErrorString(e));
    return (int)e;
  }
  return 0;
}

// Page-lock / release a HOST range (the node-wide result buffer of the window-sharded driver is shared memory mapped by
// every rank: each process registers its mapping so that its device->host copies are asynchronous DMA).  A failed
// registration is reported and cleared from the runtime's sticky last-error slot, so the caller can fall back to staged
// copies without poisoning the next CUDA call.
extern "C" int vda_host_register(void* host_ptr, int64_t bytes) {
  cudaError_t e = cudaHostRegister(host_ptr, (size_t)bytes, cudaHostRegisterPortable);
  if (e != cudaSuccess) {
    vda_set_error("vda_host_register(%p, %lld): %s", host_ptr, (long long)bytes, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return (int)e;
  }
  return 0;
}

extern "C" int vda_host_unregister(void* host_ptr) {
  cudaError_t e = cudaHostUnregister(host_ptr);
  if (e != cudaSuccess) {
    vda_set_error("vda_host_unregister(%p): %s", host_ptr, cudaGetErrorString(e));
    (void)cudaGetLastError();
    return (int)e;
  }
  return 0;
}
