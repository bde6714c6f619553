// Pieces shared by the single-CTA (csrc/gemm_tcgen05.cu) and the CTA-pair (csrc/gemm_pair.cu) GEMM kernels when a
// collective is fused into them: per-destination output tensor maps (GEMM -> reduce-scatter), the host-side option
// block, and the copy role of the all-gather -> GEMM kernel (a few CTAs of the grid pull the peers' shards over NVLink
// and raise row-block flags the TMA producers wait on).
#pragma once
#include "host.h"
#include "ptx.cuh"

namespace stb {

constexpr int kMaxDest = 8;
struct CMaps {
  CUtensorMap m[kMaxDest];
};

constexpr int kAgCopyThreads = 256;   // block size of both GEMM kernels

struct GemmFused {
  int m_rotate = 0;
  int n_dest = 0;                 // > 0: GEMM->reduce-scatter into c_dest[d] (each [M / n_dest, N], ld = ldc)
  void* c_dest[kMaxDest] = {nullptr};
  const uint32_t* a_ready = nullptr;
  uint32_t a_epoch = 0;
  int a_flag_rows = 0;
  int n_copy = 0;
  const uint8_t* const* ag_bufs = nullptr;
  uint32_t* const* ag_flag_pages = nullptr;
  long ag_offset_bytes = 0;
  uint8_t* ag_dst = nullptr;
  int ag_rank = 0, ag_world = 1, ag_slot = 0;
};


// ---- copy role of the all-gather->GEMM kernel --------------------------------------------------------
// Flag-page layout and barrier protocol are shared with csrc/comm.cu (kMaxWorld = 16, kMaxSlots = 8).
STB_DEVICE void ag_cross_gpu_barrier(uint32_t* const* pages, int rank, int world, int slot, uint32_t epoch) {
  const int t = threadIdx.x;
  if (t < world) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pages[t] + slot * 16 + rank), "r"(epoch) : "memory");
    const uint32_t* mine = pages[rank] + slot * 16 + t;
    uint32_t v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    } while (int32_t(v - epoch) < 0);
  }
}

template <class Params>
STB_DEVICE void ag_copy_role(const Params& p, const uint32_t a_epoch) {
  uint32_t* go = p.ag_flag_pages[p.ag_rank] + 8 * 16 + p.ag_slot;
  uint32_t* done = p.ag_flag_pages[p.ag_rank] + 8 * 16 + 8 + p.ag_slot;
  // 1. every rank has written its shard (cross-GPU barrier by copy CTA 0, released locally to the others)
  if (blockIdx.x == 0) {
    if (threadIdx.x < 32) ag_cross_gpu_barrier(p.ag_flag_pages, p.ag_rank, p.ag_world, p.ag_slot, a_epoch);
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(go), "r"(a_epoch) : "memory");
  } else {
    if (threadIdx.x == 0) {
      uint32_t v;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(go) : "memory");
      } while (int32_t(v - a_epoch) < 0);
    }
    __syncthreads();
  }
  // 2. pull row blocks: own shard first, then the peers in ring order (the order the compute CTAs consume)
  const long rows_per_rank = p.M / p.ag_world;
  const long row_bytes = long(p.K) * 2;
  const long blocks_per_rank = (rows_per_rank + p.a_flag_rows - 1) / p.a_flag_rows;
  const long total = blocks_per_rank * p.ag_world;
  const long vec_per_row = row_bytes >> 4;
  for (long blk = blockIdx.x; blk < total; blk += p.n_copy) {
    const int src = int((p.ag_rank + blk / blocks_per_rank) % p.ag_world);
    const long rb = blk % blocks_per_rank;
    const long r0 = rb * p.a_flag_rows, r1 = min(rows_per_rank, r0 + long(p.a_flag_rows));
    const uint4* s = reinterpret_cast<const uint4*>(p.ag_bufs[src] + p.ag_offset_bytes + r0 * row_bytes);
    uint4* d = reinterpret_cast<uint4*>(p.ag_dst + (long(src) * rows_per_rank + r0) * row_bytes);
    const long nv = (r1 - r0) * vec_per_row;
    long i = threadIdx.x;
    for (; i + 7 * kAgCopyThreads < nv; i += 8 * kAgCopyThreads) {   // 8 x 16 B in flight per thread
      uint4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = ld_peer_v4(s + i + u * kAgCopyThreads);
#pragma unroll
      for (int u = 0; u < 8; ++u) d[i + u * kAgCopyThreads] = v[u];
    }
    for (; i < nv; i += kAgCopyThreads) d[i] = ld_peer_v4(s + i);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.ag_ready + long(src) * blocks_per_rank + rb),
                   "r"(a_epoch)
                   : "memory");
  }
  // 3. tell the peers their shard buffers may be reused: all copy CTAs done -> cross-GPU barrier
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    atomicAdd(done, 1u);
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x == 0) {
      uint32_t v;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(done) : "memory");
      } while (v < uint32_t(p.n_copy));
      *done = 0;
    }
    __syncthreads();
    if (threadIdx.x < 32) ag_cross_gpu_barrier(p.ag_flag_pages, p.ag_rank, p.ag_world, p.ag_slot, a_epoch + 1);
    if (p.a_epoch == 0 && threadIdx.x == 0) {
      // device-driven epochs: advance the slot counter once every CTA of the grid has read it
      uint32_t* checkin = p.ag_flag_pages[p.ag_rank] + 8 * 16 + 24 + p.ag_slot;
      uint32_t v;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(checkin) : "memory");
      } while (v < gridDim.x);
      *checkin = 0;
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.ag_flag_pages[p.ag_rank] + 8 * 16 + 16 + p.ag_slot),
                   "r"(a_epoch + 2)
                   : "memory");
    }
  }
}


}  // namespace stb
