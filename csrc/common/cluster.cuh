// PTX wrappers for thread-block clusters and CTA-pair (cta_group::2) tensor-core operation.
//
// A CTA pair = the two CTAs of a (2,1,1) cluster, resident on the two SMs of one TPC.  One tcgen05.mma.cta_group::2
// issued by the leader (cluster rank 0) computes a 256-row tile: rows 0..127 accumulate in the leader's tensor memory,
// rows 128..255 in the peer's; the A operand comes from both CTAs' shared memory (same offset in each), the B operand
// is split along N between them.  Completion is signalled to both CTAs with a multicast commit.
// PTX forms cross-checked against the CUTLASS headers vendored in this image (cute/arch/copy_sm100_tma.hpp,
// cute/arch/mma_sm100_umma.hpp, cutlass/arch/barrier.h).
#pragma once
#include "ptx.cuh"

namespace stb {

STB_DEVICE uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// Full cluster barrier with release/acquire semantics (every thread of every CTA of the cluster).
STB_DEVICE void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n"
               "barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cta address of this CTA -> shared::cluster address of the same offset in CTA `rank` of the cluster.
STB_DEVICE uint32_t mapa_shared(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// Arrive on an mbarrier that lives in another (or this) CTA of the cluster.
// (default .release.cta semantics: the .release.cluster form costs MEMBAR.ALL.GPU + ERRBAR per arrive; the data these
// barriers order -- TMA bytes, tensor-memory reads -- is ordered by complete_tx / tcgen05 fences, not by this arrive)
STB_DEVICE void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// arrive + expect_tx on an mbarrier given by its shared::cluster address
STB_DEVICE void mbar_arrive_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr),
               "r"(bytes)
               : "memory");
}

// TMA tile load executed by either CTA of a pair: data lands in the *executing* CTA's shared memory, the transaction
// bytes are credited to the mbarrier at `bar_cluster_addr` (the leader's barrier).
STB_DEVICE void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint32_t bar_cluster_addr, int c0, int c1,
                                 uint64_t hint = kEvictNormal) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster_addr), "r"(c0), "r"(c1), "l"(hint)
      : "memory");
}

// Tensor-memory management for a CTA pair: executed by one warp in EACH CTA of the pair.
STB_DEVICE void tmem_alloc_pair(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
               "r"(ncols)
               : "memory");
}
STB_DEVICE void tmem_relinquish_pair() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
STB_DEVICE void tmem_dealloc_pair(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// D[tmem of both CTAs] (+)= A[smem of both CTAs] * B[smem halves]; issued by ONE thread of the leader CTA.
STB_DEVICE void umma_f16_ss_pair(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                 uint32_t accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5, %5, %5, %5, %5}, p;\n"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}
// Arrive (once all previously issued pair-MMAs have retired) on the mbarrier at this shared-memory offset in every CTA
// of `cta_mask` (0b11 = both).
STB_DEVICE void umma_commit_pair(uint64_t* bar, uint16_t cta_mask = 0b11) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"(cta_mask)
      : "memory");
}

}  // namespace stb
