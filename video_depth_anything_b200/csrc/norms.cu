// LayerNorm (one warp per token row, row held in registers) and GroupNorm over channels-last tokens.
// HBM-bound: LayerNorm moves 4 B in + 2 B out per element in the bf16 build.
#include "common.cuh"

namespace {

// ---------------------------------------------------------------- LayerNorm
// Lane `l` owns the column groups [ (l + 32 j) * V, +V ), V = 16 bytes of input (4 fp32 / 8 bf16), so every warp-wide
// load, parameter load and store is one contiguous, fully used segment.
template <typename T, int V> struct VecIO;
template <> struct VecIO<float, 4> {
  static __device__ __forceinline__ void load(const float* p, float* v) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  }
  static __device__ __forceinline__ void store(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};
template <> struct VecIO<float, 8> {
  static __device__ __forceinline__ void load(const float* p, float* v) { Vec8<float>::load(p, v); }
  static __device__ __forceinline__ void store(float* p, const float* v) { Vec8<float>::store(p, v); }
};
template <> struct VecIO<bf16, 8> {
  static __device__ __forceinline__ void load(const bf16* p, float* v) { Vec8<bf16>::load(p, v); }
  static __device__ __forceinline__ void store(bf16* p, const float* v) { Vec8<bf16>::store(p, v); }
};
template <> struct VecIO<bf16, 4> {
  static __device__ __forceinline__ void store(bf16* p, const float* v) {
    *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]));
  }
};

// LPR lanes per row (32, 16 or 8): narrow rows (D <= 512 / 256) put two / four rows in a warp, so every lane still has
// several 16-byte loads in flight -- with a whole warp on a 384-wide fp32 row the kernel sat at half the HBM rate.
template <typename TI, typename TO, bool RB, int LPR>
__global__ void __launch_bounds__(256) layernorm_kernel(const TI* __restrict__ x, TO* __restrict__ y,
                                                       const float* __restrict__ w, const float* __restrict__ b, float eps,
                                                       long long rows, int D, int group_rows, int skip_first,
                                                       const float* __restrict__ pe, int pe_group, int pe_T) {
  constexpr int V = 16 / (int)sizeof(TI);    // elements per 16-byte input vector
  constexpr int MAXV = 1024 / (32 * V);      // vectors per lane for D <= 1024 * LPR / 32
  constexpr int RPW = 32 / LPR;              // rows per warp
  const int lane = threadIdx.x & (LPR - 1);
  const int sub = (threadIdx.x & 31) / LPR;
  // Blocks walk the rows from the LAST to the first: the GEMM epilogue that produced x (the fp32 residual stream, larger
  // than L2) wrote its rows in ascending order, so the tail of x is what L2 still holds when this kernel starts, and the
  // head of y -- written last here -- is what the next GEMM (ascending again) reads first.
  const long long blk = (long long)gridDim.x - 1 - blockIdx.x;
  const long long row = (blk * (blockDim.x >> 5) + (threadIdx.x >> 5)) * RPW + sub;
  const bool live = row < rows;  // dead rows still take part in the shuffles below
  long long orow = row;
  bool emit = live;
  if (skip_first && live) {
    long long g = row / group_rows;
    int i = (int)(row % group_rows);
    if (i == 0) emit = false;
    orow = g * (group_rows - 1) + i - 1;
  }
  const int nv = D / V;
  float v[MAXV][V];
  const TI* xr = x + row * D;
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + LPR * j;
    if (c < nv && emit) {
      VecIO<TI, V>::load(xr + c * V, v[j]);
#pragma unroll
      for (int i = 0; i < V; ++i) s += v[j][i];
    }
  }
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mean = s / (float)D;
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + LPR * j;
    if (c < nv && emit) {
#pragma unroll
      for (int i = 0; i < V; ++i) {
        const float d = v[j][i] - mean;
        ss = fmaf(d, d, ss);
      }
    }
  }
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  if (!emit) return;
  const float rstd = rsqrtf(ss / (float)D + eps);
  const float* per = pe ? pe + (long long)((row / pe_group) % pe_T) * D : nullptr;
  TO* yr = y + orow * D;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int c = lane + LPR * j;
    if (c < nv) {
      float wv[V], bv[V], pv[V], o[V];
      VecIO<float, V>::load(w + c * V, wv);
      VecIO<float, V>::load(b + c * V, bv);
      if (per) VecIO<float, V>::load(per + c * V, pv);
#pragma unroll
      for (int i = 0; i < V; ++i) {
        float t = wv[i] * (rstd * (v[j][i] - mean)) + bv[i];
        if (per) t = t + pv[i];
        o[i] = R<RB>(t);
      }
      VecIO<TO, V>::store(yr + c * V, o);
    }
  }
}

// ---------------------------------------------------------------- GroupNorm
// Deterministic three-step reduction (no atomics): per-block partial sums -> per (frame, group) mean / rstd -> apply.
constexpr int GN_MAX_BLOCKS = 32;

template <typename TI>
__global__ void __launch_bounds__(256) gn_stats_kernel(const TI* __restrict__ x, double* __restrict__ partial, int hw, int C,
                                                      int groups, int rows_per_block) {
  __shared__ float sh_s[256 * 8];
  __shared__ float sh_q[256 * 8];
  const int f = blockIdx.y;
  const int nv = C >> 3;
  const int cv = threadIdx.x % nv, rsub = threadIdx.x / nv, rstep = blockDim.x / nv;
  const int cpg = C / groups;
  float s[8] = {}, q[8] = {};
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(hw, r0 + rows_per_block);
  if (rsub < rstep) {
    const TI* xp = x + (long long)f * hw * C + cv * 8;
    int r = r0 + rsub;
    for (; r + 3 * rstep < r1; r += 4 * rstep) {  // four independent 16 B loads in flight per thread
      float v[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) Vec8<TI>::load(xp + (long long)(r + u * rstep) * C, v[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          s[i] += v[u][i];
          q[i] = fmaf(v[u][i], v[u][i], q[i]);
        }
    }
    for (; r < r1; r += rstep) {
      float v[8];
      Vec8<TI>::load(xp + (long long)r * C, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s[i] += v[i];
        q[i] = fmaf(v[i], v[i], q[i]);
      }
    }
  }
  // sh[rsub][channel]; rows of inactive threads stay zero
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sh_s[threadIdx.x * 8 + i] = 0.f;
    sh_q[threadIdx.x * 8 + i] = 0.f;
  }
  __syncthreads();
  if (rsub < rstep) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sh_s[rsub * C + cv * 8 + i] = s[i];
      sh_q[rsub * C + cv * 8 + i] = q[i];
    }
  }
  __syncthreads();
  if (threadIdx.x < groups) {
    const int g = threadIdx.x;
    double ts = 0, tq = 0;
    for (int rs = 0; rs < rstep; ++rs)
      for (int c = g * cpg; c < (g + 1) * cpg; ++c) {
        ts += (double)sh_s[rs * C + c];
        tq += (double)sh_q[rs * C + c];
      }
    double* out = partial + (((long long)f * GN_MAX_BLOCKS + blockIdx.x) * groups + g) * 2;
    out[0] = ts;
    out[1] = tq;
  }
}

__global__ void gn_finalize_kernel(const double* __restrict__ partial, double* __restrict__ stats, float2* __restrict__ fstats,
                                   int nblocks, int groups, double cnt, float eps, int total) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // (frame, group)
  if (i >= total) return;
  const int f = i / groups, g = i % groups;
  double ts = 0, tq = 0;
  for (int b = 0; b < nblocks; ++b) {
    const double* p = partial + (((long long)f * GN_MAX_BLOCKS + b) * groups + g) * 2;
    ts += p[0];
    tq += p[1];
  }
  const double mean = ts / cnt;
  double var = tq / cnt - mean * mean;
  if (var < 0) var = 0;
  stats[2 * i] = mean;
  stats[2 * i + 1] = (double)rsqrtf((float)var + eps);
  fstats[i] = make_float2((float)mean, rsqrtf((float)var + eps));  // what the apply pass reads (no fp64 there)
}

// y = a*x + b' with a = rstd*w, b' = -a*mean + bias (the fused-parameter form ATen uses).
template <typename TI, typename TO, bool RB>
__global__ void __launch_bounds__(256) gn_apply_kernel(const TI* __restrict__ x, TO* __restrict__ y,
                                                      const float2* __restrict__ fstats, const float* __restrict__ w,
                                                      const float* __restrict__ b, int hw, int C, int groups,
                                                      long long total_vec) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_vec) return;
  const int nv = C >> 3;
  const int cv = (int)(idx % nv);
  const long long row = idx / nv;
  const int f = (int)(row / hw);
  const int cpg = C / groups;
  float v[8], o[8], wv[8], bv[8];
  Vec8<TI>::load(x + row * C + cv * 8, v);
  Vec8<float>::load(w + cv * 8, wv);
  Vec8<float>::load(b + cv * 8, bv);
  if ((cpg & 7) == 0) {  // the 8 channels of this vector share a group
    const float2 ms = fstats[(long long)f * groups + (cv * 8) / cpg];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float a = ms.y * wv[i];
      o[i] = R<RB>(a * v[i] + (-a * ms.x + bv[i]));
    }
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float2 ms = fstats[(long long)f * groups + (cv * 8 + i) / cpg];
      const float a = ms.y * wv[i];
      o[i] = R<RB>(a * v[i] + (-a * ms.x + bv[i]));
    }
  }
  Vec8<TO>::store(y + row * C + cv * 8, o);
}

}  // namespace

extern "C" int vda_layernorm(const void* x, int32_t x_dtype, void* y, int32_t y_dtype, const float* w, const float* b,
                             float eps, int64_t rows, int32_t D, int32_t group_rows, int32_t skip_first, const float* pe,
                             int32_t pe_group, int32_t pe_T, vda_stream_t stream) {
  VDA_CHECK_ARG(D % 8 == 0 && D <= 1024, "vda_layernorm: D=%d must be a multiple of 8 and <= 1024", D);
  VDA_CHECK_ARG(!skip_first || group_rows > 1, "vda_layernorm: skip_first needs group_rows");
  VDA_CHECK_ARG(!pe || (pe_group > 0 && pe_T > 0), "vda_layernorm: pe needs pe_group, pe_T");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int wpb = 8;
  const int lpr = D <= 256 ? 8 : (D <= 512 ? 16 : 32);
  const long long rpb = (long long)wpb * (32 / lpr);
  const unsigned grid = (unsigned)((rows + rpb - 1) / rpb);
#define LN_(TI, TO, RB, LPR)                                                                                          \
  layernorm_kernel<TI, TO, RB, LPR><<<grid, wpb * 32, 0, st>>>((const TI*)x, (TO*)y, w, b, eps, rows, D, group_rows, \
                                                               skip_first, pe, pe_group, pe_T)
#define LN(TI, TO, RB)                     \
  do {                                     \
    if (lpr == 8) LN_(TI, TO, RB, 8);      \
    else if (lpr == 16) LN_(TI, TO, RB, 16); \
    else LN_(TI, TO, RB, 32);              \
  } while (0)
  if (x_dtype == VDA_F32 && y_dtype == VDA_BF16) LN(float, bf16, true);
  else if (x_dtype == VDA_BF16 && y_dtype == VDA_BF16) LN(bf16, bf16, true);
  else if (x_dtype == VDA_F32 && y_dtype == VDA_F32) LN(float, float, false);
  else VDA_CHECK_ARG(false, "vda_layernorm: unsupported dtypes %d -> %d", x_dtype, y_dtype);
#undef LN_
#undef LN
  VDA_CUDA_LAUNCH_CHECK("layernorm_kernel");
  return 0;
}

extern "C" int vda_groupnorm_tokens(const void* x, int32_t x_dtype, void* y, int32_t y_dtype, const float* w,
                                    const float* b, float eps, int32_t frames, int32_t hw, int32_t C, int32_t groups,
                                    double* stats, vda_stream_t stream) {
  VDA_CHECK_ARG(C % 8 == 0 && C % groups == 0 && groups <= 64 && C / 8 <= 256, "vda_groupnorm_tokens: bad C=%d groups=%d",
                C, groups);
  VDA_CHECK_ARG(x_dtype == y_dtype, "vda_groupnorm_tokens: x/y dtypes must match");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int blocks_per_frame = (hw + 31) / 32;  // enough blocks to fill the machine: 32 frames x up to 32 blocks
  if (blocks_per_frame > GN_MAX_BLOCKS) blocks_per_frame = GN_MAX_BLOCKS;
  const int rpb = (hw + blocks_per_frame - 1) / blocks_per_frame;
  blocks_per_frame = (hw + rpb - 1) / rpb;
  dim3 g1(blocks_per_frame, frames);
  double* partial = stats + (size_t)frames * groups * 2;  // [frames][GN_MAX_BLOCKS][groups][2]
  // float (mean, rstd) pairs for the apply pass (no fp64 conversions there) follow the partial sums: the caller sizes
  // `stats` as frames * groups * 2 * (GN_MAX_BLOCKS + 2) doubles (include/vda.h).
  float2* fstats = reinterpret_cast<float2*>(stats + (size_t)frames * groups * 2 * (GN_MAX_BLOCKS + 1));
  const long long total_vec = (long long)frames * hw * (C / 8);
  const unsigned g2 = (unsigned)((total_vec + 255) / 256);
  const double cnt = (double)hw * (C / groups);
  const int fg = frames * groups;
  if (x_dtype == VDA_BF16) {
    gn_stats_kernel<bf16><<<g1, 256, 0, st>>>((const bf16*)x, partial, hw, C, groups, rpb);
    gn_finalize_kernel<<<(fg + 127) / 128, 128, 0, st>>>(partial, stats, fstats, blocks_per_frame, groups, cnt, eps, fg);
    gn_apply_kernel<bf16, bf16, true><<<g2, 256, 0, st>>>((const bf16*)x, (bf16*)y, fstats, w, b, hw, C, groups, total_vec);
  } else {
    gn_stats_kernel<float><<<g1, 256, 0, st>>>((const float*)x, partial, hw, C, groups, rpb);
    gn_finalize_kernel<<<(fg + 127) / 128, 128, 0, st>>>(partial, stats, fstats, blocks_per_frame, groups, cnt, eps, fg);
    gn_apply_kernel<float, float, false><<<g2, 256, 0, st>>>((const float*)x, (float*)y, fstats, w, b, hw, C, groups, total_vec);
  }
  VDA_CUDA_LAUNCH_CHECK("groupnorm kernels");
  return 0;
}

// ---------------------------------------------------------------- allocation-free GroupNorm entry point
// Same kernels as vda_groupnorm_tokens, with the statistics scratch in a file-scope device array instead of a caller
// buffer, for callers that cannot allocate (CUDA graph capture, serving loops).  The array is part of the module image:
// every GPU that loads the library holds VDA_GN_STATIC_MAX_FG * 2 * (GN_MAX_BLOCKS + 2) doubles (1.06 MiB) until exit.
// In host code `g_gn_static_stats` is only a handle for the symbol, never a device address: the address is looked up
// once per device with cudaGetSymbolAddress while that device is current.  All calls on one device share the array (it
// keeps the previous call's contents; every slot a call reads is written by that call first, so nothing needs zeroing),
// hence two calls that overlap on the GPU (different streams) race.
constexpr int VDA_GN_STATIC_MAX_FG = 2048;  // frames * groups, e.g. 64 frames x 32 groups
static __device__ __align__(256) double g_gn_static_stats[(size_t)VDA_GN_STATIC_MAX_FG * 2 * (GN_MAX_BLOCKS + 2)];

static int gn_static_scratch(double** out) {
  constexpr int MAX_DEV = 64;
  static double* ptrs[MAX_DEV] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
    vda_set_error("vda_groupnorm_tokens_noalloc: cudaGetDevice failed or device %d out of range: %s", dev,
                  cudaGetErrorString(e));
    return e != cudaSuccess ? (int)e : -1;
  }
  double* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
  if (!p) {  // first call on this device (it is the current one); a concurrent first call resolves the same address
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_static_stats);
    if (e != cudaSuccess || !p) {
      vda_set_error("vda_groupnorm_tokens_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -1;
    }
    __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return 0;
}

extern "C" int vda_groupnorm_tokens_noalloc(const void* x, int32_t x_dtype, void* y, int32_t y_dtype, const float* w,
                                            const float* b, float eps, int32_t frames, int32_t hw, int32_t C,
                                            int32_t groups, vda_stream_t stream) {
  VDA_CHECK_ARG(frames > 0 && groups > 0 && (long long)frames * groups <= VDA_GN_STATIC_MAX_FG,
                "vda_groupnorm_tokens_noalloc: frames * groups = %lld exceeds the static scratch (max %d); use "
                "vda_groupnorm_tokens with a caller buffer", (long long)frames * groups, VDA_GN_STATIC_MAX_FG);
  double* stats = nullptr;
  const int rc = gn_static_scratch(&stats);
  if (rc != 0) return rc;
  return vda_groupnorm_tokens(x, x_dtype, y, y_dtype, w, b, eps, frames, hw, C, groups, stats, stream);
}

extern "C" int32_t vda_groupnorm_noalloc_max_frames_x_groups(void) { return VDA_GN_STATIC_MAX_FG; }
