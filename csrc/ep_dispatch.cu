// Expert-parallel token dispatch / combine over NVLink peer memory (no NCCL, no host synchronisation).
//
//   stb_ep_exchange_counts  every rank stores its per-expert pair counts [E] into row `rank` of a [world, E] table in
//                           every peer's symmetric window, device barrier, then copies the table out.
//   stb_ep_push             dispatch: one warp per (token, k) pair copies the token row (optionally scaled: the
//                           backward of the combine pushes gate-weighted output gradients) straight into the row
//                           `dst_row[p]` of rank `dst_rank[p]`'s receive region, which is already in the 128-row-padded
//                           expert-major layout the grouped GEMM (csrc/grouped_gemm.cu) consumes -- the receiver
//                           needs no re-sort, no split sizes and no padding pass.
//   stb_ep_pull             combine: one warp per token loads its k expert-output rows out of the owners' regions,
//                           applies the gate weights, sums in fp32 and writes the [T, H] result (optionally keeping
//                           the raw rows for the gate-weight gradient).  With weights == nullptr it is the backward
//                           of the dispatch (sum of the k input-gradient rows).
//
// Both data kernels open and close with an on-device cross-GPU barrier (epochs `epoch`, `epoch + 1` of `slot`), so
// regions can be recycled layer after layer with stream ordering alone.
//
// Reference counterpart: scaletorch/parallel/expert_parallel/ep_comms.py:30-171 -- five all_to_all_single calls per
// layer (tokens, expert ids, gate weights, outputs, ...) with two .tolist() host syncs for the split sizes and no
// gradient through the exchange.
#include "common/host.h"
#include "common/ptx.cuh"
#include "common/symm.cuh"

namespace stb {

__global__ void __launch_bounds__(256)
ep_counts_kernel(char* const* bufs, uint32_t* const* flag_pages, long table_off, const int* __restrict__ counts, int E,
                 int* __restrict__ out, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  for (int peer = 0; peer < world; ++peer) {
    int* tbl = reinterpret_cast<int*>(bufs[peer] + table_off) + long(rank) * E;
    for (int e = threadIdx.x; e < E; e += blockDim.x) tbl[e] = counts[e];
  }
  __syncthreads();
  if (threadIdx.x < 32) cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  __syncthreads();
  const volatile int* mine = reinterpret_cast<const volatile int*>(bufs[rank] + table_off);
  for (int i = threadIdx.x; i < world * E; i += blockDim.x) out[i] = mine[i];
  if (threadIdx.x == 0) symm_epoch_store(flag_pages, rank, slot, epoch_arg, epoch + 1);
}

__global__ void __launch_bounds__(256)
ep_push_kernel(char* const* bufs, uint32_t* const* flag_pages, long region_off, const __nv_bfloat16* __restrict__ src,
               const int* __restrict__ dst_rank, const int* __restrict__ dst_row, const float* __restrict__ scale,
               long P, int k, int H, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);   // every destination region is zeroed / free
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const int vec = H >> 3;                                           // uint4 (8 bf16) per row
  for (long p = long(blockIdx.x) * warps + (threadIdx.x >> 5); p < P; p += long(gridDim.x) * warps) {
    const int row = dst_row[p];
    if (row < 0) continue;                                          // over capacity: pair dropped
    const uint4* s = reinterpret_cast<const uint4*>(src + (p / k) * long(H));
    uint4* d = reinterpret_cast<uint4*>(bufs[dst_rank[p]] + region_off) + long(row) * vec;
    if (scale == nullptr) {
      for (int c = lane; c < vec; c += 32) d[c] = __ldg(s + c);
    } else {
      const float w = scale[p];
      for (int c = lane; c < vec; c += 32) {
        uint4 v = __ldg(s + c);
        uint32_t* u = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float2 f = unpack_bf16x2(u[i]);
          u[i] = pack_bf16x2(f.x * w, f.y * w);
        }
        d[c] = v;
      }
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);   // every rank's rows have landed
}

template <int K>
__global__ void __launch_bounds__(256)
ep_pull_kernel(char* const* bufs, uint32_t* const* flag_pages, long region_off, __nv_bfloat16* __restrict__ out,
               __nv_bfloat16* __restrict__ rows_out, const int* __restrict__ src_rank, const int* __restrict__ src_row,
               const float* __restrict__ weights, long T, int k_rt, int H, int rank, int world, int slot,
               uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);   // every rank's expert outputs are complete
  const int k = K > 0 ? K : k_rt;
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  const int vec = H >> 3;
  for (long t = long(blockIdx.x) * warps + (threadIdx.x >> 5); t < T; t += long(gridDim.x) * warps) {
    for (int c = lane; c < vec; c += 32) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      auto add_row = [&](long p, const uint4& v) {
        if (rows_out != nullptr) reinterpret_cast<uint4*>(rows_out + p * long(H))[c] = v;
        const float w = weights != nullptr ? weights[p] : 1.f;
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float2 f = unpack_bf16x2(u[i]);
          acc[2 * i] = fmaf(w, f.x, acc[2 * i]);
          acc[2 * i + 1] = fmaf(w, f.y, acc[2 * i + 1]);
        }
      };
      auto load_row = [&](long p) {
        const int row = src_row[p];
        uint4 v = make_uint4(0, 0, 0, 0);
        if (row >= 0) v = *(reinterpret_cast<const uint4*>(bufs[src_rank[p]] + region_off) + long(row) * vec + c);
        return v;
      };
      if constexpr (K > 0) {
        uint4 v[K];                                                  // K peer loads in flight per lane
#pragma unroll
        for (int j = 0; j < K; ++j) v[j] = load_row(t * K + j);
#pragma unroll
        for (int j = 0; j < K; ++j) add_row(t * K + j, v[j]);
      } else {
        for (int j = 0; j < k; ++j) add_row(t * k + j, load_row(t * k + j));
      }
      uint4 o;
      o.x = pack_bf16x2(acc[0], acc[1]);
      o.y = pack_bf16x2(acc[2], acc[3]);
      o.z = pack_bf16x2(acc[4], acc[5]);
      o.w = pack_bf16x2(acc[6], acc[7]);
      reinterpret_cast<uint4*>(out + t * long(H))[c] = o;
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);   // owners may overwrite their regions
}

static int ep_grid(int sm_budget) {
  const int sms = sm_count();
  return (sm_budget <= 0 || sm_budget > sms) ? sms : sm_budget;
}

}  // namespace stb

using namespace stb;

// Consumes epoch `epoch` of `slot`.  `table_off` is a byte offset into every rank's symmetric window ([world, E] int32).
extern "C" int stb_ep_exchange_counts(const void* bufs_dev, const void* flag_pages_dev, long table_off,
                                      const int* counts, int E, int* out, int rank, int world, int slot,
                                      unsigned epoch, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "ep_exchange_counts: world %d / slot %d out of range", world, slot);
  ep_counts_kernel<<<1, 256, 0, stream>>>((char* const*)bufs_dev, (uint32_t* const*)flag_pages_dev, table_off, counts, E,
                                          out, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes epochs `epoch`, `epoch + 1`.  src [P / k, H] bf16 (pair p reads row p / k); dst_rank / dst_row [P]
// (dst_row < 0: dropped); scale [P] fp32 or nullptr.
extern "C" int stb_ep_push(const void* bufs_dev, const void* flag_pages_dev, long region_off, const void* src,
                           const int* dst_rank, const int* dst_row, const float* scale, long P, int k, int H, int rank,
                           int world, int slot, unsigned epoch, int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "ep_push: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(H % 8 == 0 && region_off % 16 == 0 && k > 0, "ep_push: H %% 8, 16-byte aligned region, k > 0 required");
  ep_push_kernel<<<ep_grid(sm_budget), 256, 0, stream>>>((char* const*)bufs_dev, (uint32_t* const*)flag_pages_dev,
                                                         region_off, (const __nv_bfloat16*)src, dst_rank, dst_row, scale,
                                                         P, k, H, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes epochs `epoch`, `epoch + 1`.  out [T, H] bf16; rows_out [T * k, H] bf16 or nullptr; src_rank / src_row
// [T * k]; weights [T * k] fp32 or nullptr (= 1).
extern "C" int stb_ep_pull(const void* bufs_dev, const void* flag_pages_dev, long region_off, void* out, void* rows_out,
                           const int* src_rank, const int* src_row, const float* weights, long T, int k, int H, int rank,
                           int world, int slot, unsigned epoch, int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "ep_pull: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(H % 8 == 0 && region_off % 16 == 0 && k > 0, "ep_pull: H %% 8, 16-byte aligned region, k > 0 required");
  const int grid = ep_grid(sm_budget);
  auto args = [&](auto kern) {
    kern<<<grid, 256, 0, stream>>>((char* const*)bufs_dev, (uint32_t* const*)flag_pages_dev, region_off,
                                   (__nv_bfloat16*)out, (__nv_bfloat16*)rows_out, src_rank, src_row, weights, T, k, H,
                                   rank, world, slot, epoch);
  };
  if (k == 8) args(ep_pull_kernel<8>);
  else if (k == 4) args(ep_pull_kernel<4>);
  else if (k == 2) args(ep_pull_kernel<2>);
  else args(ep_pull_kernel<0>);
  STB_CHECK_LAUNCH();
  return 0;
}
