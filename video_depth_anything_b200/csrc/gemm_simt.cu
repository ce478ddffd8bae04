// SIMT fp32-accumulate GEMM / 3x3-conv-as-GEMM: the arithmetic of the fp32 build (rel-err <= 1e-3 gate) and the
// on-device cross-check for the tcgen05 kernel.  Two launches: accumulate into an fp32 scratch, then run the same
// epilogue code (epilogue.cuh) the tensor-core kernel runs from TMEM.
#include "epilogue.cuh"

namespace {

constexpr int TM = 64, TN = 64, TK = 16;

template <typename TA, typename TB>
__global__ void __launch_bounds__(256) gemm_simt_kernel(const TA* __restrict__ A, const TB* __restrict__ B,
                                                       float* __restrict__ C, int M, int N, int K, int taps, int ksplit,
                                                       long long lda, long long ldb, long long rowsA, int m_padded,
                                                       int gb, int gh, int gw) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  float acc[4][4] = {};
  const int lr = threadIdx.x >> 2;        // 0..63 : tile row loaded by this thread
  const int lk = (threadIdx.x & 3) * 4;   // 0,4,8,12 : first k loaded
  for (int s = 0; s < ksplit; ++s) {
    for (int t = 0; t < taps; ++t) {
      const int off = vda_tap_off(taps, t, m_padded, gb, gh, gw);
      const long long arow = (long long)m0 + lr + off;
      const bool arow_ok = (m0 + lr < M) && arow >= 0 && arow < rowsA;
      const long long bcol0 = (long long)(s * taps + t) * K;
      for (int k0 = 0; k0 < K; k0 += TK) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          int k = k0 + lk + i;
          float a = 0.f, b = 0.f;
          if (arow_ok && k < K) a = to_f32<TA>(A[arow * lda + k]);
          if (n0 + lr < N && k < K) b = to_f32<TB>(B[(long long)(n0 + lr) * ldb + bcol0 + k]);
          As[lk + i][lr] = a;
          Bs[lk + i][lr] = b;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < TK; ++k) {
          float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
          float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
          float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int n = n0 + tx * 4 + j;
      if (n < N) C[(long long)m * N + n] = acc[i][j];
    }
  }
}

template <typename TO, bool RB>
__global__ void __launch_bounds__(128) epilogue_kernel(Epi e, const float* __restrict__ C, int chunks_per_row, int cw) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  int m = (int)(idx / chunks_per_row);
  if (m >= e.M) return;
  int n0 = (int)(idx % chunks_per_row) * cw;
  float acc[32], accg[32];
  const float* row = C + (long long)m * e.N;
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    acc[i] = (n0 + i < e.N) ? row[n0 + i] : 0.f;
    accg[i] = (cw == 64 && n0 + 32 + i < e.N) ? row[n0 + 32 + i] : 0.f;
  }
  RowCtx r = make_row_ctx(e, m);
  epi_chunk<TO, RB>(e, r, n0, acc, accg);
}

}  // namespace

int vda_gemm_simt(const vda_gemm_desc* d, cudaStream_t st) {
  VDA_CHECK_ARG(d->workspace && d->workspace_bytes >= (int64_t)d->M * d->N * 4,
                "vda_gemm(simt): workspace of %lld bytes needed", (long long)d->M * d->N * 4);
  float* C = reinterpret_cast<float*>(d->workspace);
  dim3 grid((d->N + TN - 1) / TN, (d->M + TM - 1) / TM);
  const int W2 = d->gw + 2;
#define LAUNCH(TA, TB)                                                                                        \
  gemm_simt_kernel<TA, TB><<<grid, 256, 0, st>>>((const TA*)d->A, (const TB*)d->B, C, d->M, d->N, d->K, d->taps, \
                                                 d->ksplit, d->lda, d->ldb, d->rowsA, d->m_padded, d->gb, d->gh, d->gw)
  if (d->a_dtype == VDA_F32 && d->b_dtype == VDA_F32) LAUNCH(float, float);
  else if (d->a_dtype == VDA_BF16 && d->b_dtype == VDA_BF16) LAUNCH(bf16, bf16);
  else if (d->a_dtype == VDA_BF16 && d->b_dtype == VDA_F32) LAUNCH(bf16, float);
  else VDA_CHECK_ARG(false, "vda_gemm(simt): unsupported operand dtypes %d/%d", d->a_dtype, d->b_dtype);
#undef LAUNCH
  VDA_CUDA_LAUNCH_CHECK("gemm_simt_kernel");
  Epi e = make_epi(d);
  const int cw = (d->act == VDA_ACT_GEGLU) ? 64 : 32;
  const int cpr = (d->N + cw - 1) / cw;
  long long total = (long long)d->M * cpr;
  int blocks = (int)((total + 127) / 128);
  const bool rb = d->round_bf16 != 0;
  if (d->out_dtype == VDA_BF16) {
    if (rb) epilogue_kernel<bf16, true><<<blocks, 128, 0, st>>>(e, C, cpr, cw);
    else epilogue_kernel<bf16, false><<<blocks, 128, 0, st>>>(e, C, cpr, cw);
  } else {
    if (rb) epilogue_kernel<float, true><<<blocks, 128, 0, st>>>(e, C, cpr, cw);
    else epilogue_kernel<float, false><<<blocks, 128, 0, st>>>(e, C, cpr, cw);
  }
  VDA_CUDA_LAUNCH_CHECK("epilogue_kernel");
  return 0;
}
