// bf16 GEMM on CTA pairs (tcgen05.mma.cta_group::2): the default dense GEMM of the library for all but tiny problems.
//
//   C[M,N] (+)= op(A) * op(B)      256 x 256 output tile per CTA pair (M = 256, N = 256, K = 16 per instruction)
//
// Why pairs: the single-CTA kernel (csrc/gemm_tcgen05.cu, 128 x 256 tile) is L2->SMEM bandwidth bound -- every CTA pulls
// 16 KB of A and 32 KB of B per 64-deep k-block = 96 B/clk/SM at full MMA rate (profiles/ncu_r1a_summary.md).  In a
// pair each CTA still loads its own 128 rows of A but only HALF of B (the other half sits in the peer's shared
// memory and the tensor core reads both), i.e. 32 KB per k-block per SM = 64 B/clk/SM.  Measured (profiles/
// gemm_v3_vs_cublas.jsonl): 0.94-1.13x cuBLAS on the model shapes where the single-CTA kernel reached 0.79-1.0x.
//
// Structure per CTA (256 threads), cluster (2,1,1), persistent over work units (pair p takes units p, p + pairs, ...):
//   warp 0      TMA producer : own A rows [m0 + 128 r, +128), own B rows [n0 + 128 r, +128); the transaction bytes of BOTH
//                              CTAs are credited to the leader's `full` barrier (cp.async.bulk.tensor...cta_group::2)
//   warp 1      MMA issuer   : leader (rank 0) only; commits are multicast to both CTAs' `empty` / `tmem_full` barriers
//   warp 2      TMEM allocator (tcgen05.alloc.cta_group::2 in each CTA)
//   warps 4..7  epilogue     : each CTA drains ITS 128 accumulator rows (tcgen05.ld -> swizzled smem -> TMA store or
//                              reduce-add); both CTAs' epilogue threads arrive on the LEADER's `tmem_empty` barrier
// A work unit is (k-split, tile).  Split-K exists for weight gradients: dW[out, in] of a small layer is a handful of
// 256 x 256 tiles with a long reduction over tokens, so the reduction is cut into `splits` ranges whose epilogues
// TMA-reduce-add into the fp32 destination (zeroed first unless the caller accumulates anyway).
// Same operand-major flexibility as the single-CTA kernel (K-major or MN-major A and B).  TMA epilogue only: C must be
// 16-byte aligned with a 16-byte-multiple row pitch (stb_gemm_bf16 falls back to the single-CTA kernel otherwise).
// All TMA loads here read this GPU's memory; the fused-collective GEMMs stay on the single-CTA kernel.
#include "common/cluster.cuh"
#include "common/gemm_fused.cuh"
#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {
namespace pair {

constexpr int kBMc = 128;        // rows of A / accumulator rows per CTA
constexpr int kBK = 64;
constexpr int kThreads = 256;
constexpr int kABytes = kBMc * kBK * 2;      // 16 KB
constexpr int kEpiBytes = 4 * 2 * 4096;

// BN = tile columns per pair: 256 (default) or 128 (problems with too few 256-wide tiles to fill the pairs)
template <int BN>
struct Cfg {
  static constexpr int kBNc = BN / 2;                  // B rows loaded per CTA
  static constexpr int kBBytes = kBNc * kBK * 2;       // 16 KB / 8 KB
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = BN == 256 ? 6 : 8;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
  static constexpr int kTmemCols = 2 * BN;             // two accumulator stages of BN fp32 columns
};

struct Params {
  int M, N, K;
  int tiles_m, tiles_n, num_tiles, k_blocks;
  int splits, kb_per_split, num_units;   // unit u = (split u / num_tiles, tile u % num_tiles)
  int out_fp32, accumulate;
  // ---- fused collectives (0 / nullptr = plain GEMM); same meaning as in the single-CTA kernel (gemm_tcgen05.cu)
  int m_rotate;                 // rotation of the 256-row tile index: other ranks' rows first / local rows first
  int rows_per_dest;            // GEMM->reduce-scatter: rows [d * rows_per_dest, ...) are added into tensor map d
  const uint32_t* a_ready;      // all-gather->GEMM: one flag per a_flag_rows rows of A, ready when >= a_epoch
  uint32_t a_epoch;
  int a_flag_rows;
  int n_copy;                   // leading CTAs (whole pairs) of the grid that run the copy role
  const uint8_t* const* ag_bufs;
  uint32_t* const* ag_flag_pages;
  long ag_offset_bytes;
  uint8_t* ag_dst;
  uint32_t* ag_ready;
  int ag_rank, ag_world, ag_slot;
};

// Grouped tile order: kGroup consecutive M tiles share each sweep over N, so the tiles in flight at any time
// (one per pair) touch ~sqrt(pairs) row blocks of A and of B instead of all of A.
STB_DEVICE void unit_coords(const Params& p, int unit, int& tm, int& tn, int& kb0, int& kb1) {
  constexpr int kGroup = 8;
  const int split = unit / p.num_tiles;
  const int tile = unit - split * p.num_tiles;
  const int group_size = kGroup * p.tiles_n;
  const int g = tile / group_size;
  const int first_m = g * kGroup;
  const int gm = min(p.tiles_m - first_m, kGroup);
  const int r = tile - g * group_size;
  tm = (first_m + r % gm + p.m_rotate) % p.tiles_m;
  tn = r / gm;
  kb0 = split * p.kb_per_split;
  kb1 = min(p.k_blocks, kb0 + p.kb_per_split);
}

template <int kBN, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CMaps tmCs, const Params p) {
  // all-gather->GEMM with device-driven epochs (see gemm_tcgen05.cu): every CTA reads the slot's counter and checks in
  uint32_t a_epoch = p.a_epoch;
  if (p.a_ready != nullptr && p.a_epoch == 0) {
    uint32_t* page = p.ag_flag_pages[p.ag_rank];
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(a_epoch) : "l"(page + 8 * 16 + 16 + p.ag_slot) : "memory");
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(page + 8 * 16 + 24 + p.ag_slot, 1u);
  }
  if (int(blockIdx.x) < p.n_copy) {   // communication CTAs (whole pairs): no TMEM, no pipeline, no cluster barrier
    ag_copy_role(p, a_epoch);
    return;
  }
  constexpr int kBNc = Cfg<kBN>::kBNc, kBBytes = Cfg<kBN>::kBBytes, kStageBytes = Cfg<kBN>::kStageBytes;
  constexpr int kStages = Cfg<kBN>::kStages, kTmemCols = Cfg<kBN>::kTmemCols;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kABytes;
  uint8_t* sEpi = smem + kStages * kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes + kEpiBytes);
  uint64_t* full_bar = bars;                      // used in the leader only
  uint64_t* empty_bar = bars + kStages;           // per CTA (multicast commit)
  uint64_t* tmem_full = bars + 2 * kStages;       // per CTA (multicast commit)
  uint64_t* tmem_empty = bars + 2 * kStages + 2;  // used in the leader only (count 2 x 128 epilogue threads)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair_id = (int(blockIdx.x) - p.n_copy) >> 1, n_pairs = (int(gridDim.x) - p.n_copy) >> 1;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmCs.m[0]);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);     // the leader's producer arrives once and expects the bytes of BOTH CTAs
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
      for (int unit = pair_id; unit < p.num_units; unit += n_pairs) {
        int tm, tn, kb0, kb1;
        unit_coords(p, unit, tm, tn, kb0, kb1);
        const int m0 = tm * 2 * kBMc + int(rank) * kBMc;      // my 128 rows of A
        const int n0 = tn * kBN + int(rank) * kBNc;           // my 128 rows of B
        if (p.a_ready != nullptr) {
          // all-gather->GEMM: my 128 rows of A are being pulled from a peer by the copy CTAs of this grid
          const int last_row = min(m0 + kBMc, p.M) - 1;
          for (int fb = m0 / p.a_flag_rows; fb <= last_row / p.a_flag_rows; ++fb) {
            const uint32_t* f = p.a_ready + fb;
            uint32_t v;
            do {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            } while (int32_t(v - a_epoch) < 0);
          }
          asm volatile("fence.proxy.async.global;" ::: "memory");  // generic-proxy writes -> TMA (async proxy) reads
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);   // the leader's barrier
          // the peer's copies may complete_tx before the leader's expect_tx: the phase cannot flip until the leader
          // has arrived, and by then the expected byte count is posted
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * kStageBytes);
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
      for (int unit = pair_id; unit < p.num_units; unit += n_pairs, ++it) {
        int tm, tn, kb0, kb1;
        unit_coords(p, unit, tm, tn, kb0, kb1);
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);     // both CTAs' epilogues drained this accumulator stage
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * kBN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);           // both CTAs' tiles of this stage have landed
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * kABytes);
          const uint32_t b_addr = smem_u32(sB + stage * kBBytes);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = umma_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, 1024);
            const uint64_t db = umma_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, 1024);
            umma_f16_ss_pair(tmem_d, da, db, idesc, ((kb - kb0) | k) != 0);
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
    const bool add = p.accumulate || p.splits > 1;
    for (int unit = pair_id; unit < p.num_units; unit += n_pairs, ++it) {
      int tm, tn, kb0, kb1;
      unit_coords(p, unit, tm, tn, kb0, kb1);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      int row0 = tm * 2 * kBMc + int(rank) * kBMc + ew * 32;
      // GEMM->reduce-scatter: this CTA's 128 rows belong to rank `dest`, whose output buffer (peer memory) tensor map
      // `dest` addresses: the TMA reduce-add below travels over NVLink
      const CUtensorMap* tmC = &tmCs.m[0];
      if (p.rows_per_dest > 0) {
        const int dest = row0 / p.rows_per_dest;
        tmC = &tmCs.m[dest];
        row0 -= dest * p.rows_per_dest;
      }
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
          if (add) tma_reduce_add_2d(tmC, slots + slot * 4096, tn * kBN + c, row0);
          else tma_store_2d(tmC, slots + slot * 4096, tn * kBN + c, row0);
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

// Work model used to choose tile width and k-split, and to decide between this kernel and the single-CTA one.  Unit of
// time: one 64-deep k-block of a 128 x 256 (per SM) tile at full tensor rate (512 clk ~ 0.27 us).  Constants fitted to
// tools/gemm_sweep.py runs (profiles/gemm_sweep_r2.jsonl): the 128-wide tile moves 1.5x the operand bytes per flop and
// ran 4096^3 in 146 us where the 256-wide tile needs 99 us, so it only pays when it halves the number of waves;
// a k-split costs the zero-fill launch plus read-modify-write epilogues (~6 us) and more per extra split.
constexpr float kEpiUnits = 5.0f;        // exposed drain of a CTA's last 128 x 256 accumulator half (bf16 store)
constexpr float kSplitUnits = 40.0f;     // zero-fill launch (5.7 us measured in-step) + reduce-add epilogue instead of a store
constexpr float kPerSplitUnits = 2.0f;
constexpr int kMinKbPerSplit = 8;

float pair_cost(int M, int N, int K, int out_fp32, int bn, int splits, int n_pairs) {
  const long tiles = long((M + 2 * kBMc - 1) / (2 * kBMc)) * ((N + bn - 1) / bn);
  const int kb = (K + kBK - 1) / kBK;
  const int per = (kb + splits - 1) / splits;
  const long units = tiles * ((kb + per - 1) / per);
  const long waves = (units + n_pairs - 1) / n_pairs;
  const float width = bn == 256 ? 1.0f : 0.76f;   // the 128-wide pair tile is L2->SMEM bound like the single-CTA kernel
  float c = float(waves) * float(per) * width + kEpiUnits * width * (out_fp32 ? 1.5f : 1.0f);
  if (splits > 1) c += kSplitUnits + kPerSplitUnits * float(splits);
  return c;
}

// Best (tile width, k-split) for this problem; returns its cost.
float pick_config(int M, int N, int K, int out_fp32, int n_pairs, int* bn_out, int* splits_out) {
  const int kb = (K + kBK - 1) / kBK;
  float best = 1e30f;
  for (int bn = 256; bn >= 128; bn -= 128) {
    for (int s = 1; s <= 8; ++s) {
      if (s > 1 && (!out_fp32 || kb / s < kMinKbPerSplit)) break;
      const float c = pair_cost(M, N, K, out_fp32, bn, s, n_pairs);
      if (c < best * (s > 1 || bn == 128 ? 0.97f : 1.0f)) {      // prefer the plain wide tile on near-ties
        best = c;
        *bn_out = bn;
        *splits_out = s;
      }
    }
  }
  return best;
}

template <int BN, bool A_MN, bool B_MN>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CMaps& tmC, const Params& p,
                  cudaStream_t stream) {
  auto kern = gemm_pair_kernel<BN, A_MN, B_MN>;
  static bool attr_set[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    STB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmemBytes));
    attr_set[dev] = true;
  }
  int grid = (sm_count() - p.n_copy) & ~1;          // whole pairs only, all co-resident with the copy CTAs
  const int need = 2 * p.num_units;
  if (need < grid) grid = need;
  grid += p.n_copy;
  kern<<<grid, kThreads, Cfg<BN>::kSmemBytes, stream>>>(tmA, tmB, tmC, p);
  STB_CHECK_LAUNCH();
  return 0;
}

// bn: 0 = choose, 128 / 256 = forced tile width.  splits: 0 = choose, 1 = none, n > 1 = forced (fp32 destination only).
int run(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int a_mn, int b_mn,
        int out_fp32, int accumulate, int bn, int splits, cudaStream_t stream, const GemmFused* fxp) {
  static const GemmFused no_fusion;
  const GemmFused& fx = fxp != nullptr ? *fxp : no_fusion;
  const bool fused = fx.n_dest > 0 || fx.a_ready != nullptr;
  if (fused) {
    bn = 256;
    splits = 1;
  }
  if (fx.n_dest > 0) C = fx.c_dest[0];
  STB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_pair: empty problem");
  const int esz = out_fp32 ? 4 : 2;
  STB_REQUIRE((reinterpret_cast<uintptr_t>(C) & 15) == 0 && (long(ldc) * esz) % 16 == 0,
              "gemm_pair: C must be 16-byte aligned with a 16-byte-multiple row pitch");
  STB_REQUIRE(!accumulate || out_fp32 || fx.n_dest > 0, "gemm_pair: accumulation needs an fp32 destination");
  STB_REQUIRE(fx.n_copy % 2 == 0, "gemm_pair: the copy role takes whole CTA pairs (n_copy %d)", fx.n_copy);
  STB_REQUIRE(!(a_mn && !b_mn), "gemm_pair: (A MN-major, B K-major) is not instantiated");
  STB_REQUIRE(bn == 0 || bn == 128 || bn == 256, "gemm_pair: tile width %d", bn);
  const int n_pairs = sm_count() / 2;
  if (bn == 0 || splits <= 0) {
    int pbn = 256, ps = 1;
    pick_config(M, N, K, out_fp32, n_pairs, &pbn, &ps);
    if (bn == 0 && splits <= 0) { bn = pbn; splits = ps; }
    else if (bn == 0) bn = 256;
    else splits = 1;
  }
  if (!out_fp32) splits = 1;
  CUtensorMap tmA, tmB;
  CMaps tmC;
  int rc;
  // operand boxes: K-major -> [64 k] x [rows per CTA]; MN-major -> [64 rows] x [64 k] atoms (as in the single-CTA kernel)
  if (!a_mn) rc = make_tmap_2d(&tmA, A, 2, K, M, uint64_t(lda) * 2, kBK, kBMc, 1);
  else rc = make_tmap_2d(&tmA, A, 2, M, K, uint64_t(lda) * 2, 64, kBK, 1);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d(&tmB, B, 2, K, N, uint64_t(ldb) * 2, kBK, bn / 2, 1);
  else rc = make_tmap_2d(&tmB, B, 2, N, K, uint64_t(ldb) * 2, 64, kBK, 1);
  if (rc) return rc;
  const int nd = fx.n_dest > 0 ? fx.n_dest : 1;
  const int c_rows = fx.n_dest > 0 ? M / fx.n_dest : M;
  for (int d = 0; d < kMaxDest; ++d) {
    if (d < nd) {
      void* base = fx.n_dest > 0 ? fx.c_dest[d] : C;
      if (out_fp32) rc = make_tmap_2d(&tmC.m[d], base, 4, N, c_rows, uint64_t(ldc) * 4, 32, 32, 1);
      else rc = make_tmap_2d(&tmC.m[d], base, 2, N, c_rows, uint64_t(ldc) * 2, 64, 32, 1);
      if (rc) return rc;
    } else {
      tmC.m[d] = tmC.m[0];
    }
  }
  Params p;
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = (M + 2 * kBMc - 1) / (2 * kBMc);
  p.tiles_n = (N + bn - 1) / bn;
  p.num_tiles = p.tiles_m * p.tiles_n;
  p.k_blocks = (K + kBK - 1) / kBK;
  if (splits > p.k_blocks) splits = p.k_blocks;
  p.kb_per_split = (p.k_blocks + splits - 1) / splits;
  p.splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;     // no empty range
  p.num_units = p.num_tiles * p.splits;
  p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.m_rotate = fx.m_rotate % p.tiles_m;
  p.rows_per_dest = fx.n_dest > 0 ? M / fx.n_dest : 0;
  p.a_ready = fx.a_ready; p.a_epoch = fx.a_epoch; p.a_flag_rows = fx.a_flag_rows; p.n_copy = fx.n_copy;
  p.ag_bufs = fx.ag_bufs; p.ag_flag_pages = fx.ag_flag_pages; p.ag_offset_bytes = fx.ag_offset_bytes;
  p.ag_dst = fx.ag_dst; p.ag_ready = const_cast<uint32_t*>(fx.a_ready);
  p.ag_rank = fx.ag_rank; p.ag_world = fx.ag_world; p.ag_slot = fx.ag_slot;
  if (p.splits > 1 && !accumulate)   // every split adds: start from zero
    STB_CHECK_CUDA(cudaMemset2DAsync(C, size_t(ldc) * 4, 0, size_t(N) * 4, size_t(M), stream));
#define STB_PAIR_CASE(BN_)                                                    \
  do {                                                                        \
    if (a_mn && b_mn) return launch<BN_, true, true>(tmA, tmB, tmC, p, stream); \
    if (b_mn) return launch<BN_, false, true>(tmA, tmB, tmC, p, stream);        \
    return launch<BN_, false, false>(tmA, tmB, tmC, p, stream);                 \
  } while (0)
  if (bn == 256) STB_PAIR_CASE(256);
  STB_PAIR_CASE(128);
#undef STB_PAIR_CASE
}

}  // namespace pair
}  // namespace stb

// CTA-pair GEMM with explicit tile width and k-split; same operand contract as stb_gemm_bf16 (include/
// scaletorch_b200.h), which dispatches here by itself.  C must be TMA-storable (16-byte aligned base and row pitch).
extern "C" int stb_gemm_bf16_2cta(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                                  int a_mn, int b_mn, int out_fp32, int accumulate, int bn, int splits,
                                  cudaStream_t stream) {
  return stb::pair::run(A, B, C, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_fp32, accumulate, bn, splits, stream, nullptr);
}
