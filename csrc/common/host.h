// Host-side helpers shared by all translation units of libscaletorch_b200.so:
// error reporting through a thread-local string (the C ABI returns int codes),
// cached device properties and TMA tensor-map construction via the driver
// entry point (no link-time dependency on libcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace stb {

void set_last_error(const char* fmt, ...);
int sm_count();           // SM count of the current device (cached per device)
int max_smem_optin();     // opt-in shared memory per block of the current device

// Builds a 2-D bf16/fp32 tiled tensor map. dim0 is the contiguous dimension.
// Returns 0 on success.
int make_tmap_2d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t dim0, uint64_t dim1,
                 uint64_t stride1_bytes, uint32_t box0, uint32_t box1, int swizzle_128b);
int make_tmap_3d(CUtensorMap* out, const void* base, int elem_bytes, uint64_t dim0, uint64_t dim1,
                 uint64_t dim2, uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box0,
                 uint32_t box1, uint32_t box2, int swizzle_128b);

// [B, S, H, D] bf16 tensor with element strides (sb, ss, sh), D contiguous; box = {box_d, 1, box_s, 1}
int make_tmap_bshd(CUtensorMap* out, const void* base, int B, int S, int H, int D, long sb, long ss,
                   long sh, uint32_t box_d, uint32_t box_s);

// same for an fp32 tensor, not swizzled (shared-memory tile is dense [box_s][box_d]); used for TMA reduce-add tiles
int make_tmap_bshd_f32(CUtensorMap* out, const void* base, int B, int S, int H, int D, long sb, long ss, long sh,
                       uint32_t box_d, uint32_t box_s);

// CTA-pair GEMM (csrc/gemm_pair.cu): entry, its work model, and the k-split it would choose.
struct GemmFused;   // csrc/common/gemm_fused.cuh
namespace pair {
int run(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int a_mn, int b_mn,
        int out_fp32, int accumulate, int bn, int splits, cudaStream_t stream, const GemmFused* fx);
float pick_config(int M, int N, int K, int out_fp32, int n_pairs, int* bn_out, int* splits_out);
float pair_cost(int M, int N, int K, int out_fp32, int bn, int splits, int n_pairs);
}  // namespace pair

#define STB_CHECK_CUDA(expr)                                                              \
  do {                                                                                    \
    cudaError_t _e = (expr);                                                              \
    if (_e != cudaSuccess) {                                                              \
      stb::set_last_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                   \
                          cudaGetErrorString(_e));                                        \
      return -1;                                                                          \
    }                                                                                     \
  } while (0)

#define STB_REQUIRE(cond, ...)            \
  do {                                    \
    if (!(cond)) {                        \
      stb::set_last_error(__VA_ARGS__);   \
      return -2;                          \
    }                                     \
  } while (0)

#define STB_CHECK_LAUNCH() STB_CHECK_CUDA(cudaGetLastError())

}  // namespace stb
