// C-ABI glue: error text, device check, tensor maps, vda_gemm dispatch.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"
#include "tmap.h"

static thread_local char g_err[512] = "";

void vda_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char* vda_last_error(void) { return g_err; }
extern "C" int vda_version(void) { return VDA_VERSION; }

extern "C" int vda_device_check(int dev, int* sm_count) {
  cudaDeviceProp p;
  cudaError_t e = cudaGetDeviceProperties(&p, dev);
  if (e != cudaSuccess) {
    vda_set_error("cudaGetDeviceProperties(%d): %s", dev, cudaGetErrorString(e));
    return (int)e;
  }
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (p.major != 10) {
    vda_set_error("libvda_sm100a needs an sm_100 device, found sm_%d%d (%s)", p.major, p.minor, p.name);
    return -2;
  }
  return 0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

int vda_make_tmap_2d(CUtensorMap* out, const void* base, uint64_t cols, uint64_t rows, uint64_t pitch_bytes,
                     uint32_t box_cols, uint32_t box_rows, int elem_bytes) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    vda_set_error("cuTensorMapEncodeTiled is unavailable (driver too old?)");
    return -3;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    vda_set_error("cuTensorMapEncodeTiled failed (%d): base=%p cols=%llu rows=%llu pitch=%llu box=%ux%u", (int)r, base,
                  (unsigned long long)cols, (unsigned long long)rows, (unsigned long long)pitch_bytes, box_cols, box_rows);
    return -4;
  }
  return 0;
}

int vda_gemm_simt(const vda_gemm_desc* d, cudaStream_t st);
int vda_gemm_tc(const vda_gemm_desc* d, cudaStream_t st);

extern "C" int vda_gemm(const vda_gemm_desc* d, vda_stream_t stream) {
  VDA_CHECK_ARG(d != nullptr && d->A && d->B, "vda_gemm: null operand");
  VDA_CHECK_ARG(d->M > 0 && d->N > 0 && d->K > 0, "vda_gemm: bad shape %d x %d x %d", d->M, d->N, d->K);
  VDA_CHECK_ARG(d->taps == 1 || d->taps == 9, "vda_gemm: taps must be 1 or 9");
  VDA_CHECK_ARG(d->ksplit == 1 || d->ksplit == 2, "vda_gemm: ksplit must be 1 or 2");
  VDA_CHECK_ARG(d->taps == 1 || (d->m_padded && d->gh > 0 && d->gw > 0), "vda_gemm: 3x3 taps need a padded M space");
  VDA_CHECK_ARG(d->m_padded != 2 || (d->taps == 9 && d->gb > 0), "vda_gemm: m_padded = 2 is the stride-2 3x3 form (taps = 9, gb set)");
  VDA_CHECK_ARG(d->out || d->resid || d->head_out, "vda_gemm: no output");
  const bool needs_geo = d->out_map != VDA_MAP_SAME || d->res_map != VDA_MAP_SAME || d->resid_map != VDA_MAP_SAME;
  VDA_CHECK_ARG(!needs_geo || (d->gh > 0 && d->gw > 0), "vda_gemm: row maps need gh, gw");
  const int ct_stride = d->ct_stride ? d->ct_stride : d->ct_cout;
  VDA_CHECK_ARG(d->out_map != VDA_MAP_CONVT || (d->ct_k > 0 && d->ct_cout > 0 && ct_stride >= d->ct_cout && d->ct_cout % 8 == 0 &&
                                                 d->N == d->ct_k * d->ct_k * ct_stride),
                "vda_gemm: conv-transpose needs N == ct_k^2 * ct_stride, ct_stride >= ct_cout, ct_cout %% 8 == 0");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  return d->use_tc ? vda_gemm_tc(d, st) : vda_gemm_simt(d, st);
}
