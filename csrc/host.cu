// Host-side runtime of the library: error reporting, device queries, TMA descriptor
// construction through the driver entry point, and a trivial smoke kernel.
#include <stdarg.h>
#include <string.h>

#include <mutex>
#include <unordered_map>

#include "../include/b200gpt.h"
#include "common.cuh"

thread_local char g_b200_errmsg[512] = {0};
__device__ int g_b200_device_error = 0;

int b200_set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_b200_errmsg, sizeof(g_b200_errmsg), fmt, ap);
  va_end(ap);
  return code;
}

int* b200_error_word() {
  static int* cached[64] = {nullptr};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  if (cached[dev] == nullptr) {
    int* p = nullptr;
    if (cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_b200_device_error) != cudaSuccess)
      return nullptr;
    cached[dev] = p;
  }
  return cached[dev];
}

int b200_num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                    const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                    const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

int b200_make_tmap_2d(CUtensorMap* out, const void* base, uint64_t inner, uint64_t outer,
                      uint64_t ld_elems, uint32_t box_inner, uint32_t box_outer, int elem_bytes) {
  PFN_encodeTiled enc = get_encode();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not found");
  B200_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "TMA element size %d not supported", elem_bytes);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16B aligned", base);
  B200_REQUIRE((ld_elems * elem_bytes) % 16 == 0, "TMA row stride %llu elems (x%d B) not a multiple of 16 B",
               (unsigned long long)ld_elems, elem_bytes);
  B200_REQUIRE(box_inner * elem_bytes <= 128 && box_outer <= 256, "TMA box too large");
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {ld_elems * elem_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                   const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS,
               "cuTensorMapEncodeTiled(2d) failed: %d (inner=%llu outer=%llu ld=%llu box=%ux%u esz=%d)",
               (int)r, (unsigned long long)inner, (unsigned long long)outer,
               (unsigned long long)ld_elems, box_inner, box_outer, elem_bytes);
  return 0;
}

int b200_make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2,
                      uint64_t stride1_elems, uint64_t stride2_elems, uint32_t b0, uint32_t b1,
                      uint32_t b2) {
  PFN_encodeTiled enc = get_encode();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not found");
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16B aligned", base);
  B200_REQUIRE((stride1_elems * 2) % 16 == 0 && (stride2_elems * 2) % 16 == 0,
               "TMA strides must be multiples of 16 bytes");
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {stride1_elems * 2, stride2_elems * 2};
  cuuint32_t box[3] = {b0, b1, b2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(3d) failed: %d", (int)r);
  return 0;
}

int b200_make_tmap_nd(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int elem_bytes, int swizzle128) {
  PFN_encodeTiled enc = get_encode();
  B200_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point not found");
  B200_REQUIRE(rank >= 1 && rank <= 5, "TMA rank %d not supported", rank);
  B200_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "TMA element size %d not supported", elem_bytes);
  B200_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base %p not 16B aligned", base);
  cuuint64_t d[5], st[4];
  cuuint32_t bx[5], estr[5];
  for (int i = 0; i < rank; ++i) { d[i] = dims[i]; bx[i] = box[i]; estr[i] = 1; }
  for (int i = 0; i + 1 < rank; ++i) {
    B200_REQUIRE(strides_bytes[i] % 16 == 0, "TMA stride %d (%llu B) not a multiple of 16 B", i, (unsigned long long)strides_bytes[i]);
    st[i] = strides_bytes[i];
  }
  CUresult r = enc(out, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank,
                   const_cast<void*>(base), d, st, bx, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(rank %d) failed: %d", rank, (int)r);
  return 0;
}

__global__ void smoke_kernel(int* out) {
  if (threadIdx.x == 0) atomicAdd(out, 0);
}

extern "C" {

const char* b200_last_error() { return g_b200_errmsg; }

int b200_device_error(int* out, int clear) {
  int v = 0;
  B200_CHECK_CUDA(cudaMemcpyFromSymbol(&v, g_b200_device_error, sizeof(int)));
  if (clear && v != 0) {
    int z = 0;
    B200_CHECK_CUDA(cudaMemcpyToSymbol(g_b200_device_error, &z, sizeof(int)));
  }
  *out = v;
  return 0;
}

int b200_smoke(void* stream) {
  int* p = b200_error_word();
  B200_REQUIRE(p != nullptr, "no CUDA device");
  smoke_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(p);
  B200_LAUNCH_CHECK();
  return 0;
}

}  // extern "C"
