// The fused GEMM epilogue, shared by the tcgen05 kernel (from TMEM registers) and the SIMT path
// (from an fp32 accumulator scratch).  One call handles 32 consecutive accumulator columns of one row.
// See include/vda.h (vda_gemm_desc) for the meaning of every field and the reference lines it replaces.
#pragma once
#include "common.cuh"

struct Epi {
  const float* bias;
  const float* gamma;
  const float* pos;
  const float* head_w;
  void* out;
  void* out_relu;
  const void* res;
  const void* res2;
  float* resid;
  void* head_out;
  long long ldo, ldres, ldresid;
  int M, N;  // accumulator rows / columns
  int gb, gh, gw, m_padded;
  int act, out_map, res_map, resid_map, resid_assign, pos_rows, ct_k, ct_cout, ct_stride;
  float head_b;
};

static inline Epi make_epi(const vda_gemm_desc* d) {
  Epi e;
  e.bias = d->bias; e.gamma = d->gamma; e.pos = d->pos; e.head_w = d->head_w;
  e.out = d->out; e.out_relu = d->out_relu; e.res = d->res; e.res2 = d->res2; e.resid = d->resid;
  e.head_out = d->head_out;
  e.ldo = d->ldo; e.ldres = d->ldres; e.ldresid = d->ldresid;
  e.M = d->M; e.N = d->N;
  e.gb = d->gb; e.gh = d->gh; e.gw = d->gw; e.m_padded = d->m_padded;
  e.act = d->act; e.out_map = d->out_map; e.res_map = d->res_map; e.resid_map = d->resid_map;
  e.resid_assign = d->resid_assign; e.pos_rows = d->pos_rows; e.ct_k = d->ct_k; e.ct_cout = d->ct_cout; e.ct_stride = d->ct_stride ? d->ct_stride : d->ct_cout;
  e.head_b = d->head_b;
  return e;
}

// bias of accumulator column n (conv-transpose: per output channel; padding columns have none)
__device__ __forceinline__ float epi_bias(const Epi& e, int n) {
  if (!e.ct_k) return e.bias[n];
  const int co = n % e.ct_stride;
  return co < e.ct_cout ? e.bias[co] : 0.f;
}

// Row offset of tap t of a 3x3 convolution in the A operand (include/vda.h: taps, m_padded).
__host__ __device__ __forceinline__ int vda_tap_off(int taps, int t, int m_padded, int gb, int gh, int gw) {
  if (taps != 9) return 0;
  const int dy = t / 3, dx = t - dy * 3;
  if (m_padded == 2) return ((dy & 1) * 2 + (dx & 1)) * gb * gh * gw + (dy >> 1) * gw + (dx >> 1);
  return (dy - 1) * (gw + 2) + (dx - 1);
}

struct RowCtx {
  long long out_row, res_row, resid_row;
  int b, y, x, m;
  bool valid;
};

__device__ __forceinline__ long long map_row(int kind, const Epi& e, int m, int b, int y, int x, bool& ok) {
  switch (kind) {
    case VDA_MAP_SAME: return m;
    case VDA_MAP_DENSE: return ((long long)b * e.gh + y) * e.gw + x;
    case VDA_MAP_PADDED: return ((long long)b * (e.gh + 2) + y + 1) * (e.gw + 2) + x + 1;
    case VDA_MAP_STRIDE2: {
      if ((y | x) & 1) ok = false;
      int gh2 = (e.gh + 1) >> 1, gw2 = (e.gw + 1) >> 1;
      return ((long long)b * gh2 + (y >> 1)) * gw2 + (x >> 1);
    }
    case VDA_MAP_TOKENS: return (long long)m + m / (e.gh * e.gw) + 1;
    case VDA_MAP_S2D: {
      const int Y = y + 1, X = x + 1, hp = ((e.gh + 1) >> 1) + 1, wp = ((e.gw + 1) >> 1) + 1;
      return ((long long)(((Y & 1) * 2 + (X & 1)) * e.gb + b) * hp + (Y >> 1)) * wp + (X >> 1);
    }
    case VDA_MAP_CROP: {
      if (y >= e.gh - 1 || x >= e.gw - 1) ok = false;
      return ((long long)b * (e.gh - 1) + y) * (e.gw - 1) + x;
    }
    default: return m;  // VDA_MAP_CONVT resolved per column
  }
}

__device__ __forceinline__ RowCtx make_row_ctx(const Epi& e, int m) {
  RowCtx r;
  r.m = m; r.b = 0; r.y = 0; r.x = 0;
  r.valid = m < e.M;
  if (e.gh > 0) {
    if (e.m_padded == 1) {
      int W2 = e.gw + 2, H2 = e.gh + 2;
      int xp = m % W2, t = m / W2;
      int yp = t % H2;
      r.b = t / H2;
      r.valid = r.valid && xp >= 1 && xp <= e.gw && yp >= 1 && yp <= e.gh;
      r.y = yp - 1; r.x = xp - 1;
    } else {
      r.x = m % e.gw;
      int t = m / e.gw;
      r.y = t % e.gh;
      r.b = t / e.gh;
    }
  }
  bool ok = true;
  r.out_row = map_row(e.out_map, e, m, r.b, r.y, r.x, ok);
  r.valid = r.valid && ok;
  bool ok2 = true;
  r.res_row = map_row(e.res_map, e, m, r.b, r.y, r.x, ok2);
  r.resid_row = map_row(e.resid_map, e, m, r.b, r.y, r.x, ok2);
  return r;
}

// acc: 32 accumulator columns [n0, n0+32) of row r.m.  GEGLU: acc = 32 value columns, accg = the 32 gate columns
// (accumulator columns n0+32..n0+63) and the output columns are n0/2 .. n0/2+31.
template <typename TO, bool RB>
__device__ __forceinline__ void epi_chunk(const Epi& e, const RowCtx& r, int n0, const float* acc, const float* accg) {
  if (!r.valid) return;
  const int ncols = (e.act == VDA_ACT_GEGLU) ? (e.N >> 1) : e.N;  // output columns
  const int c0 = (e.act == VDA_ACT_GEGLU) ? (n0 >> 1) : n0;       // first output column
  float v[32];
  // ---- bias, rounding, activation
  if (e.act == VDA_ACT_GEGLU) {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float bv = e.bias ? e.bias[n0 + i] : 0.f, bg = e.bias ? e.bias[n0 + 32 + i] : 0.f;
      float val = R<RB>(acc[i] + bv), gate = R<RB>(accg[i] + bg);
      v[i] = R<RB>(val * R<RB>(gelu_erf(gate)));
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      int n = n0 + i;
      float bvv = 0.f;
      if (e.bias && n < e.N) bvv = epi_bias(e, n);
      float t = acc[i] + bvv;
      if (e.head_w == nullptr) t = R<RB>(t);  // the fused tail is an fp32 island
      if (e.act == VDA_ACT_GELU) t = R<RB>(gelu_erf(t));
      else if (e.act == VDA_ACT_RELU) t = fmaxf(t, 0.f);
      v[i] = t;
    }
  }
  // ---- fused ReLU -> 1x1 (N -> 1) -> ReLU tail (N == 32)
  if (e.head_w != nullptr) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) s = fmaf(fmaxf(v[i], 0.f), e.head_w[i], s);
    s = fmaxf(s + e.head_b, 0.f);
    reinterpret_cast<float*>(e.head_out)[r.out_row] = s;  // fp32 island: never rounded
    return;
  }
  // ---- fp32 residual stream
  if (e.resid != nullptr) {
    float* rp = e.resid + r.resid_row * e.ldresid + c0;
    if (e.resid_assign) {
      const float* pp = e.pos + (long long)(r.resid_row % e.pos_rows) * e.ldresid + c0;
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c0 + i < ncols) rp[i] = __fadd_rn(v[i], pp[i]);
    } else if (c0 + 32 <= ncols) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        float4 x = *reinterpret_cast<float4*>(rp + i);
        float4 g = *reinterpret_cast<const float4*>(e.gamma + c0 + i);
        x.x = __fadd_rn(x.x, __fmul_rn(v[i], g.x));
        x.y = __fadd_rn(x.y, __fmul_rn(v[i + 1], g.y));
        x.z = __fadd_rn(x.z, __fmul_rn(v[i + 2], g.z));
        x.w = __fadd_rn(x.w, __fmul_rn(v[i + 3], g.w));
        *reinterpret_cast<float4*>(rp + i) = x;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (c0 + i < ncols) rp[i] = __fadd_rn(rp[i], __fmul_rn(v[i], e.gamma[c0 + i]));
    }
    if (e.out == nullptr) return;
  }
  // ---- skip connections + stores
  long long orow = r.out_row;
  int oc0 = c0, owidth = ncols;
  if (e.out_map == VDA_MAP_CONVT) {
    // fast path requires the 32-chunk to sit inside one (ky,kx) block
    if (e.ct_stride % 32 != 0) {
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        int n = c0 + i;
        if (n >= ncols) continue;
        int kk = n / e.ct_stride, co = n % e.ct_stride, ky = kk / e.ct_k, kx = kk % e.ct_k;
        if (co >= e.ct_cout) continue;
        long long row = ((long long)r.b * (e.gh * e.ct_k + 2) + r.y * e.ct_k + ky + 1) * (e.gw * e.ct_k + 2) +
                        r.x * e.ct_k + kx + 1;
        reinterpret_cast<TO*>(e.out)[row * e.ldo + co] = from_f32<TO>(v[i]);
      }
      return;
    }
    int kk = c0 / e.ct_stride, ky = kk / e.ct_k, kx = kk % e.ct_k;
    oc0 = c0 % e.ct_stride;
    owidth = e.ct_cout;
    if (oc0 >= owidth) return;  // a chunk of padding columns
    orow = ((long long)r.b * (e.gh * e.ct_k + 2) + r.y * e.ct_k + ky + 1) * (e.gw * e.ct_k + 2) + r.x * e.ct_k + kx + 1;
    if (c0 >= ncols) return;
  }
  const bool full = (oc0 + 32 <= owidth) && ((e.ldo & 7) == 0) && ((e.ldres & 7) == 0);
  TO* op = reinterpret_cast<TO*>(e.out) + orow * e.ldo + oc0;
  if (full) {
    if (e.res) {
      const TO* rp = reinterpret_cast<const TO*>(e.res) + r.res_row * e.ldres + oc0;
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float t[8];
        Vec8<TO>::load(rp + i, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i + j] = R<RB>(v[i + j] + t[j]);
      }
    }
    if (e.res2) {
      const TO* rp = reinterpret_cast<const TO*>(e.res2) + r.res_row * e.ldres + oc0;
#pragma unroll
      for (int i = 0; i < 32; i += 8) {
        float t[8];
        Vec8<TO>::load(rp + i, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i + j] = R<RB>(v[i + j] + t[j]);
      }
    }
#pragma unroll
    for (int i = 0; i < 32; i += 8) Vec8<TO>::store(op + i, v + i);
    if (e.out_relu) {
      TO* op2 = reinterpret_cast<TO*>(e.out_relu) + orow * e.ldo + oc0;
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
#pragma unroll
      for (int i = 0; i < 32; i += 8) Vec8<TO>::store(op2 + i, v + i);
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if (oc0 + i >= owidth) continue;
      float t = v[i];
      if (e.res) t = R<RB>(t + to_f32<TO>(reinterpret_cast<const TO*>(e.res)[r.res_row * e.ldres + oc0 + i]));
      if (e.res2) t = R<RB>(t + to_f32<TO>(reinterpret_cast<const TO*>(e.res2)[r.res_row * e.ldres + oc0 + i]));
      op[i] = from_f32<TO>(t);
      if (e.out_relu) reinterpret_cast<TO*>(e.out_relu)[orow * e.ldo + oc0 + i] = from_f32<TO>(fmaxf(t, 0.f));
    }
  }
}
