// EXPERIMENTAL (not on any default path; opt in with SCALETORCH_B200_GEMM_2CTA=1): bf16 GEMM on CTA pairs.
//
//   C[M,N] (+)= op(A) * op(B)      256 x 256 output tile per CTA pair, tcgen05.mma.cta_group::2 (M = 256, N = 256, K = 16)
//
// Why: the single-CTA kernel (csrc/gemm_tcgen05.cu, 128 x 256 tile) is L2->SMEM bandwidth bound -- every CTA pulls
// 16 KB of A and 32 KB of B per 64-deep k-block = 96 B/clk/SM at full MMA rate (profiles/ncu_r1a_summary.md).  In a
// pair each CTA still loads its own 128 rows of A but only HALF of B (the other half sits in the peer's shared
// memory and the tensor core reads both), i.e. 32 KB per k-block per SM = 64 B/clk/SM.
//
// Structure per CTA (256 threads), cluster (2,1,1), persistent over 256 x 256 tiles (pair p takes tiles p, p + pairs, ...):
//   warp 0      TMA producer : own A rows [m0 + 128 r, +128), own B rows [n0 + 128 r, +128); the transaction bytes of BOTH
//                              CTAs are credited to the leader's `full` barrier (cp.async.bulk.tensor...cta_group::2)
//   warp 1      MMA issuer   : leader (rank 0) only; commits are multicast to both CTAs' `empty` / `tmem_full` barriers
//   warp 2      TMEM allocator (tcgen05.alloc.cta_group::2 in each CTA)
//   warps 4..7  epilogue     : each CTA drains ITS 128 accumulator rows (tcgen05.ld -> swizzled smem -> TMA store or
//                              reduce-add); both CTAs' epilogue threads arrive on the LEADER's `tmem_empty` barrier
// Same operand-major flexibility as the single-CTA kernel (K-major or MN-major A and B).  TMA epilogue only: C must be
// 16-byte aligned with a 16-byte-multiple row pitch.
//
// Status: compiles for sm_100a; NOT yet run on hardware (written after the round's GPU budget was spent).  The GPU
// test is gated behind SCALETORCH_B200_EXPERIMENTAL=1 (tests/test_experimental_gpu.py).
#include "../common/cluster.cuh"
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace stb {
namespace pair {

constexpr int kBMc = 128;        // rows of A / accumulator rows per CTA
constexpr int kBN = 256;         // tile columns per pair
constexpr int kBNc = kBN / 2;    // B rows loaded per CTA
constexpr int kBK = 64;
constexpr int kStages = 5;
constexpr int kThreads = 256;
constexpr int kABytes = kBMc * kBK * 2;      // 16 KB
constexpr int kBBytes = kBNc * kBK * 2;      // 16 KB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kEpiBytes = 4 * 2 * 4096;
constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
constexpr int kTmemCols = 2 * kBN;           // two accumulator stages of 256 fp32 columns

struct Params {
  int M, N, K;
  int tiles_m, tiles_n, num_tiles, k_blocks;
  int out_fp32, accumulate;
};

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmC, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kABytes;
  uint8_t* sEpi = smem + kStages * kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes + kEpiBytes);
  uint64_t* full_bar = bars;                      // used in the leader only (count 2: both producers arrive)
  uint64_t* empty_bar = bars + kStages;           // per CTA (multicast commit)
  uint64_t* tmem_full = bars + 2 * kStages;       // per CTA (multicast commit)
  uint64_t* tmem_empty = bars + 2 * kStages + 2;  // used in the leader only (count 2 x 128 epilogue threads)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair_id = int(blockIdx.x) >> 1, n_pairs = int(gridDim.x) >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 2);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 2 * 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc_pair(tmem_ptr, kTmemCols);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();                 // barrier inits and TMEM allocation of BOTH CTAs are visible
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = pair_id; tile < p.num_tiles; tile += n_pairs) {
        const int tm = tile % p.tiles_m, tn = tile / p.tiles_m;
        const int m0 = tm * 2 * kBMc + int(rank) * kBMc;      // my 128 rows of A
        const int n0 = tn * kBN + int(rank) * kBNc;           // my 128 rows of B
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);   // the leader's barrier
          if (leader) mbar_arrive_expect_tx_cluster(bar, 2 * kStageBytes);
          else mbar_arrive_cluster(bar);
          uint8_t* a_dst = sA + stage * kABytes;
          uint8_t* b_dst = sB + stage * kBBytes;
          const int k0 = kb * kBK;
          if constexpr (!A_MN) {
            tma_load_2d_pair(a_dst, &tmA, bar, k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < kBMc / 64; ++j) tma_load_2d_pair(a_dst + j * (kBK * 128), &tmA, bar, m0 + j * 64, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d_pair(b_dst, &tmB, bar, k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < kBNc / 64; ++j) tma_load_2d_pair(b_dst + j * (kBK * 128), &tmB, bar, n0 + j * 64, k0);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader only)
    if (leader && lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * kBMc, kBN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t a_lbo = A_MN ? kBK * 128 : 16, b_lbo = B_MN ? kBK * 128 : 16;
      constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32, b_kstep = B_MN ? 16 * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = pair_id; tile < p.num_tiles; tile += n_pairs, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);     // both CTAs' epilogues drained this accumulator stage
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * kBN;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);           // both CTAs' tiles of this stage have landed
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * kABytes);
          const uint32_t b_addr = smem_u32(sB + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = umma_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, 1024);
            const uint64_t db = umma_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, 1024);
            umma_f16_ss_pair(tmem_d, da, db, idesc, (kb | k) != 0);
          }
          umma_commit_pair(&empty_bar[stage]);          // frees this stage in BOTH CTAs
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(&tmem_full[acc]);              // accumulator ready in BOTH CTAs
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue (both CTAs, own 128 rows)
    const int ew = warp - 4;
    uint8_t* slots = sEpi + ew * 8192;
    int slot = 0, it = 0;
    const uint32_t sw = uint32_t(lane & 7);
    for (int tile = pair_id; tile < p.num_tiles; tile += n_pairs, ++it) {
      const int tm = tile % p.tiles_m, tn = tile / p.tiles_m;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row0 = tm * 2 * kBMc + int(rank) * kBMc + ew * 32;
      const uint32_t taddr = tmem_base + (uint32_t(ew * 32) << 16) + acc * kBN;
      const uint32_t empty_addr = mapa_shared(smem_u32(&tmem_empty[acc]), 0);
      const int step = p.out_fp32 ? 32 : 64;            // columns per 128-byte staging row
#pragma unroll 1
      for (int c = 0; c < kBN; c += step) {
        uint32_t r0[32], r1[32];
        tmem_ld_32x32b_x32(taddr + c, r0);
        if (!p.out_fp32) tmem_ld_32x32b_x32(taddr + c + 32, r1);
        tmem_ld_wait();
        if (c + step == kBN) {                          // accumulator fully in registers
          tc_fence_before();
          mbar_arrive_cluster(empty_addr);
        }
        if (lane == 0) tma_store_wait_read<1>();
        __syncwarp();
        const uint32_t dst = smem_u32(slots + slot * 4096 + lane * 128);
        if (p.out_fp32) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            st_shared_v4(dst + ((uint32_t(j) ^ sw) << 4), r0[4 * j], r0[4 * j + 1], r0[4 * j + 2], r0[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            st_shared_v4(dst + ((uint32_t(j) ^ sw) << 4),
                         pack_bf16x2(__uint_as_float(r0[8 * j + 0]), __uint_as_float(r0[8 * j + 1])),
                         pack_bf16x2(__uint_as_float(r0[8 * j + 2]), __uint_as_float(r0[8 * j + 3])),
                         pack_bf16x2(__uint_as_float(r0[8 * j + 4]), __uint_as_float(r0[8 * j + 5])),
                         pack_bf16x2(__uint_as_float(r0[8 * j + 6]), __uint_as_float(r0[8 * j + 7])));
            st_shared_v4(dst + ((uint32_t(j + 4) ^ sw) << 4),
                         pack_bf16x2(__uint_as_float(r1[8 * j + 0]), __uint_as_float(r1[8 * j + 1])),
                         pack_bf16x2(__uint_as_float(r1[8 * j + 2]), __uint_as_float(r1[8 * j + 3])),
                         pack_bf16x2(__uint_as_float(r1[8 * j + 4]), __uint_as_float(r1[8 * j + 5])),
                         pack_bf16x2(__uint_as_float(r1[8 * j + 6]), __uint_as_float(r1[8 * j + 7])));
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (p.accumulate) tma_reduce_add_2d(&tmC, slots + slot * 4096, tn * kBN + c, row0);
          else tma_store_2d(&tmC, slots + slot * 4096, tn * kBN + c, row0);
          tma_store_commit();
        }
        slot ^= 1;
      }
    }
    if (lane == 0) tma_store_wait<0>();
  }
  tc_fence_before();
  cluster_sync_all();                 // neither CTA may free tensor memory while the pair is still using it
  if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
}

template <bool A_MN, bool B_MN>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const Params& p,
                  cudaStream_t stream) {
  auto kern = gemm_pair_kernel<A_MN, B_MN>;
  STB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int grid = sm_count() & ~1;                       // whole pairs only
  const int need = 2 * p.num_tiles;
  if (need < grid) grid = need;
  kern<<<grid, kThreads, kSmemBytes, stream>>>(tmA, tmB, tmC, p);
  STB_CHECK_LAUNCH();
  return 0;
}

}  // namespace pair
}  // namespace stb

using namespace stb;

// Same contract as stb_gemm_bf16 (include/scaletorch_b200.h) restricted to TMA-storable C: 16-byte aligned base and
// row pitch.  Returns non-zero (with stb_last_error set) for shapes it does not take; callers fall back to stb_gemm_bf16.
extern "C" int stb_gemm_bf16_2cta(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                                  int a_mn, int b_mn, int out_fp32, int accumulate, cudaStream_t stream) {
  using namespace stb::pair;
  STB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_2cta: empty problem");
  const int esz = out_fp32 ? 4 : 2;
  STB_REQUIRE((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (long(ldc) * esz) % 16 == 0,
              "gemm_2cta: C must be 16-byte aligned with a 16-byte-multiple row pitch");
  STB_REQUIRE(!accumulate || out_fp32, "gemm_2cta: accumulation needs an fp32 destination");
  CUtensorMap tmA, tmB, tmC;
  int rc;
  // operand boxes: K-major -> [64 k] x [128 rows]; MN-major -> [64 rows] x [64 k] atoms (as in the single-CTA kernel)
  if (!a_mn) rc = make_tmap_2d(&tmA, A, 2, K, M, uint64_t(lda) * 2, kBK, kBMc, 1);
  else rc = make_tmap_2d(&tmA, A, 2, M, K, uint64_t(lda) * 2, 64, kBK, 1);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d(&tmB, B, 2, K, N, uint64_t(ldb) * 2, kBK, kBNc, 1);
  else rc = make_tmap_2d(&tmB, B, 2, N, K, uint64_t(ldb) * 2, 64, kBK, 1);
  if (rc) return rc;
  if (out_fp32) rc = make_tmap_2d(&tmC, C, 4, N, M, uint64_t(ldc) * 4, 32, 32, 1);
  else rc = make_tmap_2d(&tmC, C, 2, N, M, uint64_t(ldc) * 2, 64, 32, 1);
  if (rc) return rc;
  Params p;
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = (M + 2 * kBMc - 1) / (2 * kBMc);
  p.tiles_n = (N + kBN - 1) / kBN;
  p.num_tiles = p.tiles_m * p.tiles_n;
  p.k_blocks = (K + kBK - 1) / kBK;
  p.out_fp32 = out_fp32; p.accumulate = accumulate;
  if (a_mn && b_mn) return launch<true, true>(tmA, tmB, tmC, p, stream);
  if (a_mn) return launch<true, false>(tmA, tmB, tmC, p, stream);
  if (b_mn) return launch<false, true>(tmA, tmB, tmC, p, stream);
  return launch<false, false>(tmA, tmB, tmC, p, stream);
}
