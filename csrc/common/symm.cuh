// Device-side cross-GPU synchronisation over symmetric flag pages (shared by comm.cu, ep_dispatch.cu).
//
// Flags are monotonically increasing epochs (never reset), so consecutive calls need no re-zeroing.  Every kernel
// that uses these helpers must be launched with a fully resident grid (<= SM count), because its blocks wait on one
// another / on other GPUs.
#pragma once
#include "ptx.cuh"

namespace stb {

STB_DEVICE void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
STB_DEVICE uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
STB_DEVICE uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
STB_DEVICE void st_release_gpu(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Flag layout inside each rank's flag page (uint32): [slot][world] arrival epochs written by the peers,
// followed by one local "go" word per slot at index kMaxSlots*kMaxWorld + slot.
constexpr int kMaxWorld = 16;
constexpr int kMaxSlots = 8;

// Cross-GPU barrier executed by one warp of one block: rank r writes `epoch` into slot[r] of every peer,
// then waits until every peer has written `epoch` (or later) into its own page.
STB_DEVICE void cross_gpu_barrier(uint32_t* const* flag_pages, int rank, int world, int slot, uint32_t epoch) {
  const int t = threadIdx.x;
  if (t < world) {
    __threadfence_system();
    st_release_sys(flag_pages[t] + slot * kMaxWorld + rank, epoch);
    const uint32_t* mine = flag_pages[rank] + slot * kMaxWorld + t;
    while (int32_t(ld_acquire_sys(mine) - epoch) < 0) {
    }
  }
}

// Grid-wide: block 0 performs the cross-GPU barrier and releases the other blocks through a local word.
STB_DEVICE void grid_cross_gpu_barrier(uint32_t* const* flag_pages, int rank, int world, int slot, uint32_t epoch) {
  uint32_t* go = flag_pages[rank] + kMaxSlots * kMaxWorld + slot;
  if (blockIdx.x == 0) {
    if (threadIdx.x < 32) cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
    __syncthreads();
    if (threadIdx.x == 0) st_release_gpu(go, epoch);
  } else {
    if (threadIdx.x == 0) {
      while (int32_t(ld_acquire_gpu(go) - epoch) < 0) {
      }
    }
    __syncthreads();
  }
}

// End-of-kernel variant: every block reports completion; block 0 waits for all of them, then runs the
// cross-GPU barrier.  The kernel (hence the stream) completes only after every rank finished its part.
STB_DEVICE void grid_finish_cross_gpu_barrier(uint32_t* const* flag_pages, int rank, int world, int slot,
                                              uint32_t epoch) {
  uint32_t* done = flag_pages[rank] + kMaxSlots * kMaxWorld + kMaxSlots + slot;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();
    atomicAdd(done, 1u);
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x == 0) {
      while (ld_acquire_gpu(done) < gridDim.x) {
      }
      *done = 0;   // re-arm: the next kernel using this slot starts after this one has completed
    }
    __syncthreads();
    if (threadIdx.x < 32) cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  }
}


}  // namespace stb
