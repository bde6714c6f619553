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
constexpr int kCpChunk = 32768;       // bytes per bulk copy of the all-gather copy role
constexpr int kCpStages = 6;          // chunks in shared memory per copy CTA (6 x 32 KB + barriers < the GEMM's own smem)
constexpr int kCpAhead = 3;           // loads issued ahead of the store that drains them

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
  // 2. pull row blocks: own shard first, then the peers in ring order (the order the compute CTAs consume).
  // One thread per copy CTA drives the TMA engine: bulk loads peer -> shared memory (kCpStages x kCpChunk bytes in
  // flight per CTA, enough to cover the ~2 us NVLink round trip with a handful of CTAs), bulk stores shared -> local
  // A_full.  The loads are plain cp.async.bulk (no multicast); the rest of the CTA idles, so the copy role costs the GEMM
  // two to four SMs instead of sixteen.
  const long rows_per_rank = p.M / p.ag_world;
  const long row_bytes = long(p.K) * 2;
  const long blocks_per_rank = (rows_per_rank + p.a_flag_rows - 1) / p.a_flag_rows;
  const long total = blocks_per_rank * p.ag_world;
  if (threadIdx.x == 0) {
    extern __shared__ uint8_t ag_smem_raw[];
    uint8_t* buf = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ag_smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bar = reinterpret_cast<uint64_t*>(buf + kCpStages * kCpChunk);
    for (int s_ = 0; s_ < kCpStages; ++s_) mbar_init(&bar[s_], 1);
    fence_barrier_init();
    long issued = 0, stored = 0;          // chunk counters over the whole role (stage = counter % kCpStages)
    for (long blk = blockIdx.x; blk < total; blk += p.n_copy) {
      const int src = int((p.ag_rank + blk / blocks_per_rank) % p.ag_world);
      const long rb = blk % blocks_per_rank;
      const long r0 = rb * p.a_flag_rows, r1 = min(rows_per_rank, r0 + long(p.a_flag_rows));
      const uint8_t* s = p.ag_bufs[src] + p.ag_offset_bytes + r0 * row_bytes;
      uint8_t* d = p.ag_dst + (long(src) * rows_per_rank + r0) * row_bytes;
      const long bytes = (r1 - r0) * row_bytes;
      const long n_chunks = (bytes + kCpChunk - 1) / kCpChunk;
      const long first = issued;
      auto store_chunk = [&](long c) {     // c: chunk index inside this block
        const long g = first + c;
        const int st = int(g % kCpStages);
        mbar_wait(&bar[st], uint32_t((g / kCpStages) & 1));
        const long off = c * kCpChunk;
        const uint32_t n = uint32_t(min(long(kCpChunk), bytes - off));
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(d + off),
                     "r"(smem_u32(buf + st * kCpChunk)), "r"(n)
                     : "memory");
        tma_store_commit();
        ++stored;
      };
      for (long c = 0; c < n_chunks; ++c) {
        const int st = int(issued % kCpStages);
        // the store that last used this stage must have finished reading shared memory: at most kCpAhead - 1 younger
        // stores may still be pending
        if (issued >= kCpStages) tma_store_wait_read<kCpStages - kCpAhead - 1>();
        const long off = c * kCpChunk;
        const uint32_t n = uint32_t(min(long(kCpChunk), bytes - off));
        mbar_arrive_expect_tx(&bar[st], n);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(buf + st * kCpChunk)),
                     "l"(s + off), "r"(n), "r"(smem_u32(&bar[st]))
                     : "memory");
        ++issued;
        if (c >= kCpAhead) store_chunk(c - kCpAhead);
      }
      for (long c = max(0L, n_chunks - kCpAhead); c < n_chunks; ++c) store_chunk(c);
      tma_store_wait<0>();                 // every byte of the block is in A_full ...
      asm volatile("fence.proxy.async.global;" ::: "memory");
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.ag_ready + long(src) * blocks_per_rank + rb),
                   "r"(a_epoch)
                   : "memory");            // ... before its flag says so
    }
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
