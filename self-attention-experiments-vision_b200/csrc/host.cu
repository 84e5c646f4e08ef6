#include "host.h"

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <mutex>

namespace sav {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}
void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int check_launch(const char* what) {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    cudaGetLastError();
    return set_error(SAV_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
  }
  return SAV_OK;
}

typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_fn_t get_encode() {
  static encode_fn_t fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<encode_fn_t>(p);
  });
  return fn;
}

int make_tmap_nd(CUtensorMap* out, const void* base, int elem_bytes, int rank, const uint64_t* dims,
                 const uint64_t* strides_bytes, const uint32_t* box) {
  encode_fn_t enc = get_encode();
  if (!enc) return set_error(SAV_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) {
      gstr[i - 1] = strides_bytes[i];
      if (strides_bytes[i] % 16 != 0)
        return set_error(SAV_ERR_ALIGN, "tensor-map stride %llu not a multiple of 16 B",
                         (unsigned long long)strides_bytes[i]);
    }
  }
  if (!aligned16(base)) return set_error(SAV_ERR_ALIGN, "tensor-map base not 16 B aligned");
  CUresult r = enc(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                   (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(SAV_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return SAV_OK;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims,
                   const uint64_t* strides_bytes, const uint32_t* box) {
  return make_tmap_nd(out, base, 2, rank, dims, strides_bytes, box);
}

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, int64_t cols, int64_t rows,
                 int64_t ld, uint32_t box_cols, uint32_t box_rows) {
  encode_fn_t enc = get_encode();
  if (!enc) return set_error(SAV_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
  if (!aligned16(base) || (ld * elem_bytes) % 16)
    return set_error(SAV_ERR_ALIGN, "tensor-map base / row pitch not 16 B aligned");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * elem_bytes};
  cuuint32_t bx[2] = {box_cols, box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(out, elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16,
                   2, const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(SAV_ERR_CUDA, "cuTensorMapEncodeTiled (2d) failed (%d)", (int)r);
  return SAV_OK;
}

static void* g_debug_buffer = nullptr;
void* sav_debug_buffer() { return g_debug_buffer; }
void set_debug_buffer(void* p) { g_debug_buffer = p; }

}  // namespace sav

extern "C" {
/* development aid: device buffer (>= 6*512 int64) receiving a kernel phase timeline; NULL disables */
void sav_set_debug_buffer(void* p) { sav::set_debug_buffer(p); }
int sav_abi_version(void) { return SAV_ABI_VERSION; }
const char* sav_last_error(void) { return sav::g_err; }
int64_t sav_launch_count(void) { return sav::g_launches.load(); }
int sav_device_check(void) {
  int dev = 0;
  cudaDeviceProp p;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&p, dev) != cudaSuccess)
    return sav::set_error(SAV_ERR_CUDA, "no CUDA device");
  if (p.major != 10) return sav::set_error(SAV_ERR_ARCH, "device is sm_%d%d, need sm_100", p.major, p.minor);
  return SAV_OK;
}
}
