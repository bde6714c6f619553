#include "host.h"

#include <stdarg.h>
#include <string.h>

#include <mutex>

namespace stb {

static thread_local char g_err[1024] = {0};

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

static int g_sm_count[64] = {0};
static int g_smem_optin[64] = {0};

int sm_count() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (g_sm_count[dev] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    g_sm_count[dev] = v > 0 ? v : 148;
  }
  return g_sm_count[dev];
}

int max_smem_optin() {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 227 * 1024;
  if (g_smem_optin[dev] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    g_smem_optin[dev] = v > 0 ? v : 227 * 1024;
  }
  return g_smem_optin[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess) {
      fn = reinterpret_cast<EncodeTiledFn>(p);
    }
  });
  return fn;
}

// cuTensorMapEncodeTiled is a pure driver call: it needs a current context, which a thread that has
// not yet made any runtime call (e.g. an autograd worker on device 0) does not have.
static void ensure_context() {
  static thread_local bool done = false;
  if (!done) {
    cudaFree(0);
    done = true;
  }
}

static CUtensorMapDataType dtype_of(int elem_bytes) {
  return elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                         : (elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                            : CU_TENSOR_MAP_DATA_TYPE_UINT8);
}

int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t dim0, uint64_t dim1,
                 uint64_t stride1_bytes, uint32_t box0, uint32_t box1, int swizzle_128b) {
  ensure_context();
  EncodeTiledFn enc = get_encode();
  STB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[2] = {dim0, dim1};
  cuuint64_t strides[1] = {stride1_bytes};
  cuuint32_t box[2] = {box0, box1};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, dtype_of(elem_bytes), 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle_128b ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  STB_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeTiled(2d) failed: %d (base=%p dims=%llu,%llu stride=%llu box=%u,%u)",
              (int)r, base, (unsigned long long)dim0, (unsigned long long)dim1,
              (unsigned long long)stride1_bytes, box0, box1);
  return 0;
}

int make_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t dim0, uint64_t dim1,
                 uint64_t dim2, uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0,
                 uint32_t box1, uint32_t box2, int swizzle_128b) {
  ensure_context();
  EncodeTiledFn enc = get_encode();
  STB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[3] = {dim0, dim1, dim2};
  cuuint64_t strides[2] = {stride1_bytes, stride2_bytes};
  cuuint32_t box[3] = {box0, box1, box2};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(out, dtype_of(elem_bytes), 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE,
                   swizzle_128b ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  STB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(3d) failed: %d", (int)r);
  return 0;
}

int make_tmap_bshd(CUtensorMap* out, const void* base, int B, int S, int H, int D, long sb, long ss,
                   long sh, uint32_t box_d, uint32_t box_s) {
  ensure_context();
  EncodeTiledFn enc = get_encode();
  STB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  // dims ordered fastest-first: D, H, S, B
  cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)H, (cuuint64_t)S, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)sh * 2, (cuuint64_t)ss * 2, (cuuint64_t)sb * 2};
  cuuint32_t box[4] = {box_d, 1, box_s, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  STB_REQUIRE(r == CUDA_SUCCESS,
              "cuTensorMapEncodeTiled(bshd) failed: %d (B=%d S=%d H=%d D=%d strides=%ld,%ld,%ld)", (int)r,
              B, S, H, D, sb, ss, sh);
  return 0;
}

int make_tmap_bshd_f32(CUtensorMap* out, const void* base, int B, int S, int H, int D, long sb, long ss, long sh,
                       uint32_t box_d, uint32_t box_s) {
  ensure_context();
  EncodeTiledFn enc = get_encode();
  STB_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)H, (cuuint64_t)S, (cuuint64_t)B};
  cuuint64_t strides[3] = {(cuuint64_t)sh * 4, (cuuint64_t)ss * 4, (cuuint64_t)sb * 4};
  cuuint32_t box[4] = {box_d, 1, box_s, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  STB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(bshd f32) failed: %d (B=%d S=%d H=%d D=%d strides=%ld,%ld,%ld)",
              (int)r, B, S, H, D, sb, ss, sh);
  return 0;
}

}  // namespace stb

extern "C" const char* stb_last_error() { return stb::g_err; }

extern "C" int stb_abi_version() { return 1; }
