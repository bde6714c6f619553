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
// local words after the arrival table: go[slot], done[slot], epoch counter[slot], check-in[slot]
constexpr int kGoBase = kMaxSlots * kMaxWorld;
constexpr int kDoneBase = kGoBase + kMaxSlots;
constexpr int kEpochBase = kDoneBase + kMaxSlots;
constexpr int kCheckinBase = kEpochBase + kMaxSlots;
constexpr int kFlagPageWords = kCheckinBase + kMaxSlots;

// Epochs can be driven by the host (`epoch` != 0: the caller passes the next unused value) or by the device
// (`epoch` == 0): the kernel reads the slot's counter word in its own flag page when it starts and the last thread to
// leave advances it.  Device epochs make a launch replayable from a CUDA graph -- the kernel parameters are then the
// same on every replay while the protocol still sees strictly increasing epochs.  All kernels of a slot are
// stream-ordered on their GPU and every rank issues the same sequence, so the counters of all ranks stay equal.
// One slot must use one mode consistently.
STB_DEVICE uint32_t symm_epoch_load(uint32_t* const* flag_pages, int rank, int slot, uint32_t epoch) {
  return epoch != 0 ? epoch : ld_acquire_gpu(flag_pages[rank] + kEpochBase + slot);
}
STB_DEVICE void symm_epoch_store(uint32_t* const* flag_pages, int rank, int slot, uint32_t host_epoch, uint32_t next) {
  if (host_epoch == 0) st_release_gpu(flag_pages[rank] + kEpochBase + slot, next);
}

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
  uint32_t* go = flag_pages[rank] + kGoBase + slot;
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
// `host_epoch` is the epoch argument the kernel was launched with: when it is 0 the slot's device counter is advanced
// past `epoch` once every block has checked out (so every block has long read the counter).
STB_DEVICE void grid_finish_cross_gpu_barrier(uint32_t* const* flag_pages, int rank, int world, int slot,
                                              uint32_t epoch, uint32_t host_epoch = 1) {
  uint32_t* done = flag_pages[rank] + kDoneBase + slot;
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
    if (threadIdx.x < 32) {
      cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
      __syncwarp();
      if (threadIdx.x == 0) symm_epoch_store(flag_pages, rank, slot, host_epoch, epoch + 1);
    }
  }
}


}  // namespace stb
