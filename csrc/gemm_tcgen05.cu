// Persistent, warp-specialised bf16 GEMM on Blackwell 5th-gen tensor cores.
//
//   C[M,N] (+)= A[M,K] * B[N,K]^T          fp32 accumulation in TMEM
//
// Each operand may be stored K-major (row-major [rows, K]) or MN-major
// (row-major [K, rows]); the three combinations a Linear layer needs are
//   forward  y  = x  W^T      : A K-major,  B K-major
//   dgrad    dx = dy W        : A K-major,  B MN-major
//   wgrad    dW = dy^T x      : A MN-major, B MN-major   (optionally dW += ...)
// so no transposed copies are ever materialised (reference: F.linear + autograd,
// scaletorch/parallel/tensor_parallel/tensor_parallel.py:240,352, tp_comms.py:265-308).
//
// Structure (one CTA per SM, 256 threads):
//   warp 0      TMA producer  : cp.async.bulk.tensor -> 128B-swizzled smem ring
//   warp 1      MMA issuer    : one thread issues tcgen05.mma (128 x BN x 16)
//   warp 2      TMEM allocator
//   warps 4..7  epilogue      : tcgen05.ld -> registers -> swizzled smem -> TMA store
//                               (TMA reduce-add when accumulating; direct stores if C is unaligned)
// The accumulator is double-buffered in TMEM so the epilogue of tile i overlaps
// the main loop of tile i+1.
#include <stdlib.h>

#include "common/gemm_fused.cuh"
#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

constexpr int kBM = 128;
constexpr int kBK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int kGemmThreads = 256;

struct GemmParams {
  void* C;
  const void* bias;  // unused for now (kept for ABI stability)
  int M, N, K;
  int ldc;
  int tiles_m, tiles_n, num_tiles, k_blocks;
  int out_fp32;
  int accumulate;
  // ---- fused collectives (0 / nullptr = plain GEMM)
  int m_rotate;                 // tile-row rotation: tiles of other ranks first, local tiles last
  int rows_per_dest;            // GEMM->reduce-scatter: rows [d*rows_per_dest, ...) go to tensor map d
  const uint32_t* a_ready;      // all-gather->GEMM: one flag per a_flag_rows rows of A, ready when >= a_epoch
  uint32_t a_epoch;
  int a_flag_rows;
  // copy CTAs of the all-gather->GEMM kernel (blockIdx < n_copy): pull the peers' A shards over NVLink
  int n_copy;
  const uint8_t* const* ag_bufs;      // device array [world]: base of every rank's symmetric buffer
  uint32_t* const* ag_flag_pages;     // device array [world]: flag pages (see comm.cu)
  long ag_offset_bytes;               // shard offset inside each symmetric buffer
  uint8_t* ag_dst;                    // local A_full
  uint32_t* ag_ready;                 // same memory as a_ready, written here
  int ag_rank, ag_world, ag_slot;
};

template <int BN>
struct GemmCfg {
  static constexpr int kABytes = kBM * kBK * 2;
  static constexpr int kBBytes = BN * kBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kTmemCols = 2 * BN;  // two accumulator stages
  // epilogue staging: per epilogue warp two 4 KB slots (32 rows x 128 B, 128B-swizzled)
  static constexpr int kEpiBytes = 4 * 2 * 4096;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 /*align*/ + 256 /*barriers*/;
};

STB_DEVICE void tile_coords(int tile, int tiles_m, int tiles_n, int& m, int& n) {
  constexpr int kGroup = 8;
  const int group_size = kGroup * tiles_n;
  const int g = tile / group_size;
  const int first_m = g * kGroup;
  const int gm = min(tiles_m - first_m, kGroup);
  const int r = tile - g * group_size;
  m = first_m + r % gm;
  n = r / gm;
}

template <int BN, bool A_MN, bool B_MN, bool EPI_TMA>
__global__ void __launch_bounds__(kGemmThreads, 1)
    gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CMaps tmCs, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  // all-gather->GEMM with device-driven epochs (a_epoch == 0: graph-replayable): every CTA reads the slot's counter
  // (csrc/common/symm.cuh) and checks in; copy CTA 0 advances the counter after all of them have
  uint32_t a_epoch = p.a_epoch;
  if (p.a_ready != nullptr && p.a_epoch == 0) {
    uint32_t* page = p.ag_flag_pages[p.ag_rank];
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(a_epoch) : "l"(page + 8 * 16 + 16 + p.ag_slot) : "memory");
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(page + 8 * 16 + 24 + p.ag_slot, 1u);
  }
  if (int(blockIdx.x) < p.n_copy) {   // all-gather->GEMM: communication CTAs (no TMEM, no smem pipeline)
    ag_copy_role(p, a_epoch);
    return;
  }
  const int cta = int(blockIdx.x) - p.n_copy;
  const int ncta = int(gridDim.x) - p.n_copy;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * Cfg::kABytes;
  uint8_t* sEpi = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes + Cfg::kEpiBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  uint64_t* tmem_empty = bars + 2 * kStages + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (EPI_TMA) tma_prefetch_desc(&tmCs.m[0]);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 128);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, Cfg::kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cta; tile < p.num_tiles; tile += ncta) {
        int tm, tn;
        tile_coords(tile, p.tiles_m, p.tiles_n, tm, tn);
        tm = (tm + p.m_rotate) % p.tiles_m;
        const int m0 = tm * kBM, n0 = tn * BN;
        if (p.a_ready != nullptr) {
          // all-gather->GEMM: these rows of A are being pulled from a peer by the copy kernel running
          // next to us; wait until every flag block covering them has landed
          const int last_row = min(m0 + kBM, p.M) - 1;
          for (int fb = m0 / p.a_flag_rows; fb <= last_row / p.a_flag_rows; ++fb) {
            const uint32_t* f = p.a_ready + fb;
            uint32_t v;
            do {
              asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
            } while (int32_t(v - a_epoch) < 0);
          }
          asm volatile("fence.proxy.async.global;" ::: "memory");  // generic-proxy writes -> TMA (async proxy) reads
        }
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          uint8_t* a_dst = sA + stage * Cfg::kABytes;
          uint8_t* b_dst = sB + stage * Cfg::kBBytes;
          const int k0 = kb * kBK;
          if constexpr (!A_MN) {
            tma_load_2d(a_dst, &tmA, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int j = 0; j < kBM / 64; ++j)
              tma_load_2d(a_dst + j * (kBK * 128), &tmA, &full_bar[stage], m0 + j * 64, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d(b_dst, &tmB, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(b_dst + j * (kBK * 128), &tmB, &full_bar[stage], n0 + j * 64, k0);
          }
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // --------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kBM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      // K-major : 8-row groups are 1024 B apart (SBO); LBO unused.
      // MN-major: 64-element MN atoms are kBK*128 B apart (LBO); 8-k groups 1024 B apart (SBO).
      constexpr uint32_t a_lbo = A_MN ? kBK * 128 : 16, b_lbo = B_MN ? kBK * 128 : 16;
      constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32, b_kstep = B_MN ? 16 * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = cta; tile < p.num_tiles; tile += ncta, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < p.k_blocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::kABytes);
          const uint32_t b_addr = smem_u32(sB + stage * Cfg::kBBytes);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {
            const uint64_t da = umma_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, 1024);
            const uint64_t db = umma_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, 1024);
            umma_f16_ss(tmem_d, da, db, idesc, (kb | k) != 0);
          }
          umma_commit(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(&tmem_full[acc]);  // accumulator ready for the epilogue
      }
    }
  } else if (warp >= 4) {
    // ----------------------------------------------------------- epilogue
    const int ew = warp - 4;  // TMEM lane quarter owned by this warp
    if constexpr (EPI_TMA) {
      // TMEM -> registers -> 128B-swizzled smem slot -> TMA store (or TMA reduce-add when
      // accumulating).  Each warp owns 32 rows and two 4 KB slots, so no cross-warp barrier.
      uint8_t* slots = sEpi + ew * 8192;
      int slot = 0;
      int it = 0;
      const uint32_t sw = uint32_t(lane & 7);
      for (int tile = cta; tile < p.num_tiles; tile += ncta, ++it) {
        int tm, tn;
        tile_coords(tile, p.tiles_m, p.tiles_n, tm, tn);
        tm = (tm + p.m_rotate) % p.tiles_m;
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&tmem_full[acc], acc_phase);
        tc_fence_after();
        int row0 = tm * kBM + ew * 32;
        // GEMM->reduce-scatter: the tile's rows belong to rank `dest`; its tensor map addresses that
        // rank's output buffer (peer memory), so the TMA reduce-add below travels over NVLink
        const CUtensorMap* tmC = &tmCs.m[0];
        if (p.rows_per_dest > 0) {
          const int dest = row0 / p.rows_per_dest;
          tmC = &tmCs.m[dest];
          row0 -= dest * p.rows_per_dest;
        }
        const uint32_t taddr = tmem_base + (uint32_t(ew * 32) << 16) + acc * BN;
        if (p.out_fp32) {
#pragma unroll 1
          for (int c = 0; c < BN / 32; ++c) {
            uint32_t r[32];
            tmem_ld_32x32b_x32(taddr + c * 32, r);
            tmem_ld_wait();
            if (c == BN / 32 - 1) {  // accumulator fully drained into registers
              tc_fence_before();
              mbar_arrive(&tmem_empty[acc]);
            }
            if (lane == 0) tma_store_wait_read<1>();
            __syncwarp();
            const uint32_t dst = smem_u32(slots + slot * 4096 + lane * 128);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              st_shared_v4(dst + ((uint32_t(j) ^ sw) << 4), r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (p.accumulate)
                tma_reduce_add_2d(tmC, slots + slot * 4096, tn * BN + c * 32, row0);
              else
                tma_store_2d(tmC, slots + slot * 4096, tn * BN + c * 32, row0);
              tma_store_commit();
            }
            slot ^= 1;
          }
        } else {
#pragma unroll 1
          for (int c = 0; c < BN / 64; ++c) {
            uint32_t r0[32], r1[32];
            tmem_ld_32x32b_x32(taddr + c * 64, r0);
            tmem_ld_32x32b_x32(taddr + c * 64 + 32, r1);
            tmem_ld_wait();
            if (c == BN / 64 - 1) {
              tc_fence_before();
              mbar_arrive(&tmem_empty[acc]);
            }
            if (lane == 0) tma_store_wait_read<1>();
            __syncwarp();
            const uint32_t dst = smem_u32(slots + slot * 4096 + lane * 128);
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
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (p.accumulate)
                tma_reduce_add_2d(tmC, slots + slot * 4096, tn * BN + c * 64, row0);
              else
                tma_store_2d(tmC, slots + slot * 4096, tn * BN + c * 64, row0);
              tma_store_commit();
            }
            slot ^= 1;
          }
        }
      }
      if (lane == 0) tma_store_wait<0>();
    } else {
    int it = 0;
    for (int tile = cta; tile < p.num_tiles; tile += ncta, ++it) {
      int tm, tn;
      tile_coords(tile, p.tiles_m, p.tiles_n, tm, tn);
      tm = (tm + p.m_rotate) % p.tiles_m;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const int row = tm * kBM + ew * 32 + lane;
      const uint32_t taddr = tmem_base + (uint32_t(ew * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(taddr + c * 32, r);
        tmem_ld_wait();
        const int col0 = tn * BN + c * 32;
        if (row < p.M && col0 < p.N) {
          const bool full = (col0 + 32 <= p.N);
          if (p.out_fp32) {
            float* dst = reinterpret_cast<float*>(p.C) + size_t(row) * p.ldc + col0;
            if (full && ((p.ldc & 3) == 0)) {
#pragma unroll
              for (int v = 0; v < 8; ++v) {
                float4 o = make_float4(__uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]),
                                       __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3]));
                if (p.accumulate) {
                  const float4 old = *reinterpret_cast<const float4*>(dst + 4 * v);
                  o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                }
                *reinterpret_cast<float4*>(dst + 4 * v) = o;
              }
            } else {
#pragma unroll
              for (int v = 0; v < 32; ++v) {
                if (col0 + v < p.N) {
                  float o = __uint_as_float(r[v]);
                  if (p.accumulate) o += dst[v];
                  dst[v] = o;
                }
              }
            }
          } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.C) + size_t(row) * p.ldc + col0;
            if (full && ((p.ldc & 7) == 0)) {
#pragma unroll
              for (int v = 0; v < 4; ++v) {
                uint4 o;
                if (p.accumulate) {
                  const uint4 old = *reinterpret_cast<const uint4*>(dst + 8 * v);
                  const float2 a0 = unpack_bf16x2(old.x), a1 = unpack_bf16x2(old.y),
                               a2 = unpack_bf16x2(old.z), a3 = unpack_bf16x2(old.w);
                  o.x = pack_bf16x2(__uint_as_float(r[8 * v + 0]) + a0.x, __uint_as_float(r[8 * v + 1]) + a0.y);
                  o.y = pack_bf16x2(__uint_as_float(r[8 * v + 2]) + a1.x, __uint_as_float(r[8 * v + 3]) + a1.y);
                  o.z = pack_bf16x2(__uint_as_float(r[8 * v + 4]) + a2.x, __uint_as_float(r[8 * v + 5]) + a2.y);
                  o.w = pack_bf16x2(__uint_as_float(r[8 * v + 6]) + a3.x, __uint_as_float(r[8 * v + 7]) + a3.y);
                } else {
                  o.x = pack_bf16x2(__uint_as_float(r[8 * v + 0]), __uint_as_float(r[8 * v + 1]));
                  o.y = pack_bf16x2(__uint_as_float(r[8 * v + 2]), __uint_as_float(r[8 * v + 3]));
                  o.z = pack_bf16x2(__uint_as_float(r[8 * v + 4]), __uint_as_float(r[8 * v + 5]));
                  o.w = pack_bf16x2(__uint_as_float(r[8 * v + 6]), __uint_as_float(r[8 * v + 7]));
                }
                *reinterpret_cast<uint4*>(dst + 8 * v) = o;
              }
            } else {
#pragma unroll
              for (int v = 0; v < 32; ++v) {
                if (col0 + v < p.N) {
                  float o = __uint_as_float(r[v]);
                  if (p.accumulate) o += __bfloat162float(dst[v]);
                  dst[v] = __float2bfloat16(o);
                }
              }
            }
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[acc]);
    }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BN, bool A_MN, bool B_MN, bool EPI_TMA>
static int launch_gemm(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                       int ldc, int out_fp32, int accumulate, cudaStream_t stream, const GemmFused& fx) {
  using Cfg = GemmCfg<BN>;
  CUtensorMap tmA, tmB;
  int rc;
  if (!A_MN)
    rc = make_tmap_2d(&tmA, A, 2, K, M, uint64_t(lda) * 2, kBK, kBM, 1);
  else
    rc = make_tmap_2d(&tmA, A, 2, M, K, uint64_t(lda) * 2, 64, kBK, 1);
  if (rc) return rc;
  if (!B_MN)
    rc = make_tmap_2d(&tmB, B, 2, K, N, uint64_t(ldb) * 2, kBK, BN, 1);
  else
    rc = make_tmap_2d(&tmB, B, 2, N, K, uint64_t(ldb) * 2, 64, kBK, 1);
  if (rc) return rc;
  CMaps tmCs;
  for (int d = 0; d < kMaxDest; ++d) tmCs.m[d] = tmA;  // placeholders
  if (EPI_TMA) {
    const int nd = fx.n_dest > 0 ? fx.n_dest : 1;
    const int rows = fx.n_dest > 0 ? M / fx.n_dest : M;
    for (int d = 0; d < nd; ++d) {
      void* base = fx.n_dest > 0 ? fx.c_dest[d] : C;
      if (out_fp32)
        rc = make_tmap_2d(&tmCs.m[d], base, 4, N, rows, uint64_t(ldc) * 4, 32, 32, 1);
      else
        rc = make_tmap_2d(&tmCs.m[d], base, 2, N, rows, uint64_t(ldc) * 2, 64, 32, 1);
      if (rc) return rc;
    }
  }

  GemmParams p;
  p.C = C;
  p.bias = nullptr;
  p.M = M; p.N = N; p.K = K; p.ldc = ldc;
  p.tiles_m = (M + kBM - 1) / kBM;
  p.tiles_n = (N + BN - 1) / BN;
  p.num_tiles = p.tiles_m * p.tiles_n;
  p.k_blocks = (K + kBK - 1) / kBK;
  p.out_fp32 = out_fp32;
  p.accumulate = accumulate;
  p.m_rotate = fx.m_rotate % p.tiles_m;
  p.rows_per_dest = fx.n_dest > 0 ? M / fx.n_dest : 0;
  p.a_ready = fx.a_ready;
  p.a_epoch = fx.a_epoch;
  p.a_flag_rows = fx.a_flag_rows;
  p.n_copy = fx.n_copy;
  p.ag_bufs = fx.ag_bufs;
  p.ag_flag_pages = fx.ag_flag_pages;
  p.ag_offset_bytes = fx.ag_offset_bytes;
  p.ag_dst = fx.ag_dst;
  p.ag_ready = const_cast<uint32_t*>(fx.a_ready);
  p.ag_rank = fx.ag_rank; p.ag_world = fx.ag_world; p.ag_slot = fx.ag_slot;

  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, EPI_TMA>;
  static bool attr_set = false;
  if (!attr_set) {
    STB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::kSmemBytes));
    attr_set = true;
  }
  // all CTAs (compute + copy) must be co-resident: one CTA per SM
  const int compute_cap = sm_count() - fx.n_copy;
  const int grid = (p.num_tiles < compute_cap ? p.num_tiles : compute_cap) + fx.n_copy;
  kern<<<grid, kGemmThreads, Cfg::kSmemBytes, stream>>>(tmA, tmB, tmCs, p);
  STB_CHECK_LAUNCH();
  return 0;
}

static int pick_bn(int M, int N) {
  // ncu (profiles/ncu_r1a_summary.md): the 128x128 tile needs 128 B/cycle/SM from L2 and starves the tensor
  // pipe; 128x256 needs 96 B/cycle.  Use the wide tile whenever N allows it and the narrow one only when it
  // avoids a mostly idle last wave on small problems.
  if (N <= 128) return 128;
  const int sms = sm_count();
  const long tm = (M + kBM - 1) / kBM;
  const long t256 = tm * ((N + 255) / 256);
  if (t256 >= sms) return 256;
  const long t128 = tm * ((N + 127) / 128);
  return (t128 <= sms) ? 128 : 256;
}

// Same time unit as pair::pair_cost (csrc/gemm_pair.cu): a 64-deep k-block of a 128 x 256 tile at full tensor rate.
// The single-CTA main loop runs below that rate because it is L2->SMEM bound (x1.12 for the 256-wide tile, x1.3 for
// the 128-wide one, from profiles/gemm_v3_vs_cublas.jsonl).
static float single_cost(int M, int N, int K, int out_fp32, int bn) {
  const long tiles = long((M + kBM - 1) / kBM) * ((N + bn - 1) / bn);
  const long waves = (tiles + sm_count() - 1) / sm_count();
  const int kb = (K + kBK - 1) / kBK;
  const float per = bn == 256 ? 1.12f : 0.65f;
  return float(waves) * float(kb) * per + (bn == 256 ? 5.0f : 2.5f) * (out_fp32 ? 1.5f : 1.0f);
}

// 0 = never, 1 = whenever the shape is legal, 2 (default) = when the work model says the pair kernel is faster
static int pair_policy() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("SCALETORCH_B200_GEMM_PAIR");
    v = e ? atoi(e) : 2;
  }
  return v;
}

}  // namespace stb

// C ABI ---------------------------------------------------------------------
// a_mn / b_mn: 0 = operand stored K-major (row-major [rows, K], ld = row stride),
//              1 = operand stored MN-major (row-major [K, rows]).
static int gemm_dispatch(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                         int a_mn, int b_mn, int out_fp32, int accumulate, int force_bn, cudaStream_t stream,
                         const stb::GemmFused& fx) {
  using namespace stb;
  STB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %d %d %d", M, N, K);
  STB_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "gemm: lda/ldb must be multiples of 8 elements");
  STB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0,
              "gemm: operands must be 16-byte aligned");
  STB_REQUIRE(!(a_mn && !b_mn), "gemm: (A MN-major, B K-major) is not instantiated");
  const int bn = force_bn ? force_bn : pick_bn(M, N);
  // TMA-store epilogue needs a 16-byte aligned C with a 16-byte multiple row pitch
  const int celt = out_fp32 ? 4 : 2;
  const void* c0 = fx.n_dest > 0 ? fx.c_dest[0] : C;
  const bool epi_tma = ((reinterpret_cast<uintptr_t>(c0) & 15) == 0) && ((size_t(ldc) * celt) % 16 == 0);
  // the CTA-pair kernel takes the GEMM when it is faster; fused-collective variants too, when their row blocks are
  // whole 256-row pair tiles and the copy CTAs come in pairs
  const bool plain = fx.n_dest == 0 && fx.a_ready == nullptr && fx.n_copy == 0;
  const bool fused_ok = !plain && (fx.n_dest == 0 || (M / fx.n_dest) % 256 == 0) &&
                        (fx.a_ready == nullptr || ((M / fx.ag_world) % 256 == 0 && fx.n_copy % 2 == 0)) && M % 256 == 0;
  if ((plain || fused_ok) && epi_tma && force_bn == 0 && pair_policy() > 0 && (!accumulate || out_fp32 || fx.n_dest > 0)) {
    const int n_pairs = (sm_count() - fx.n_copy) / 2;
    int pbn = 256, ps = 1;
    float pc = pair::pick_config(M, N, K, out_fp32, n_pairs, &pbn, &ps);
    if (!plain) pc = pair::pair_cost(M, N, K, out_fp32, 256, 1, n_pairs);
    if (pair_policy() == 1 || pc <= single_cost(M, N, K, out_fp32, bn)) {
      GemmFused f2 = fx;
      f2.m_rotate = fx.m_rotate / 2;                       // 128-row tiles -> 256-row pair tiles
      return pair::run(A, B, C, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_fp32, accumulate, pbn, ps, stream,
                       plain ? nullptr : &f2);
    }
  }
  STB_REQUIRE(epi_tma || fx.n_dest == 0, "gemm_rs: destination buffers must be 16-byte aligned");
#define STB_GEMM_CASE(BN_, AMN_, BMN_)                                                             \
  return epi_tma ? launch_gemm<BN_, AMN_, BMN_, true>(A, B, C, M, N, K, lda, ldb, ldc, out_fp32,   \
                                                      accumulate, stream, fx)                      \
                 : launch_gemm<BN_, AMN_, BMN_, false>(A, B, C, M, N, K, lda, ldb, ldc, out_fp32,  \
                                                       accumulate, stream, fx)
  if (bn == 256) {
    if (!a_mn && !b_mn) STB_GEMM_CASE(256, false, false);
    if (!a_mn && b_mn) STB_GEMM_CASE(256, false, true);
    STB_GEMM_CASE(256, true, true);
  } else {
    if (!a_mn && !b_mn) STB_GEMM_CASE(128, false, false);
    if (!a_mn && b_mn) STB_GEMM_CASE(128, false, true);
    STB_GEMM_CASE(128, true, true);
  }
#undef STB_GEMM_CASE
}

extern "C" int stb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda,
                             int ldb, int ldc, int a_mn, int b_mn, int out_fp32, int accumulate,
                             int force_bn, cudaStream_t stream) {
  stb::GemmFused fx;
  return gemm_dispatch(A, B, C, M, N, K, lda, ldb, ldc, a_mn, b_mn, out_fp32, accumulate, force_bn, stream, fx);
}

// GEMM -> reduce-scatter in one kernel.  Rank `rank` of `world` computes the partial product
// P[M,N] = A * op(B)^T and adds row block d (M/world rows) of it into c_dest[d] -- rank d's output buffer,
// mapped into this process (symmetric memory) -- with TMA reduce-add issued from the epilogue, tile by
// tile, remote ranks first.  Every destination buffer must be zeroed and every rank must have passed a
// barrier before the first add; a barrier after the kernel makes the sums visible (see runtime/fused_ops.py).
extern "C" int stb_gemm_rs_bf16(const void* A, const void* B, void* const* c_dest, int M, int N, int K, int lda,
                                int ldb, int ldc, int b_mn, int rank, int world, cudaStream_t stream) {
  using namespace stb;
  STB_REQUIRE(world >= 1 && world <= kMaxDest, "gemm_rs: world %d out of range", world);
  STB_REQUIRE(M % world == 0 && (M / world) % kBM == 0, "gemm_rs: M / world must be a multiple of %d", kBM);
  GemmFused fx;
  fx.n_dest = world;
  for (int d = 0; d < world; ++d) fx.c_dest[d] = c_dest[d];
  fx.m_rotate = ((rank + 1) % world) * (M / world / kBM);   // start with the next rank's rows, end with our own
  return gemm_dispatch(A, B, nullptr, M, N, K, lda, ldb, ldc, 0, b_mn, 0, 1, 0, stream, fx);
}

// all-gather -> GEMM in one kernel: C[M,N] = A_full[M,K] * op(B)^T where row block r of A_full is rank r's
// shard, living at `shard_offset_bytes` of that rank's symmetric buffer.  `n_copy` CTAs of the grid pull the
// shards over NVLink into the local `A_full` buffer (own shard first, then the peers in ring order) and raise
// one flag per `flag_rows` rows; the remaining CTAs are the usual persistent GEMM, whose TMA producer waits
// on the flags of the rows it is about to load.  A_full is returned to the caller (wgrad needs it).
// Consumes epochs `epoch` and `epoch + 1` of flag slot `slot`.
extern "C" int stb_ag_gemm_bf16(void* A_full, const void* B, void* C, int M, int N, int K, int ldb, int ldc,
                                int b_mn, const void* bufs_dev, const void* flag_pages_dev, long shard_offset_bytes,
                                void* ready_flags, unsigned epoch, int flag_rows, int n_copy, int rank, int world,
                                int slot, cudaStream_t stream) {
  using namespace stb;
  STB_REQUIRE(M % world == 0 && (M / world) % kBM == 0, "ag_gemm: M / world must be a multiple of %d", kBM);
  STB_REQUIRE(flag_rows % kBM == 0 || kBM % flag_rows == 0, "ag_gemm: flag_rows must divide or be a multiple of %d", kBM);
  // the copy CTAs number flags per source shard, the TMA producer by global row: both agree only if shards are whole flag blocks
  STB_REQUIRE((M / world) % flag_rows == 0, "ag_gemm: M / world (%d) must be a multiple of flag_rows (%d)", M / world, flag_rows);
  STB_REQUIRE(n_copy >= 1 && n_copy < sm_count() / 2, "ag_gemm: n_copy %d out of range", n_copy);
  STB_REQUIRE((long(K) * 2) % 16 == 0 && shard_offset_bytes % 16 == 0, "ag_gemm: rows must be 16-byte multiples");
  GemmFused fx;
  fx.a_ready = (const uint32_t*)ready_flags;
  fx.a_epoch = epoch;
  fx.a_flag_rows = flag_rows;
  fx.m_rotate = rank * (M / world / kBM);   // local rows first, then the peers in ring order
  fx.n_copy = n_copy;
  fx.ag_bufs = (const uint8_t* const*)bufs_dev;
  fx.ag_flag_pages = (uint32_t* const*)flag_pages_dev;
  fx.ag_offset_bytes = shard_offset_bytes;
  fx.ag_dst = (uint8_t*)A_full;
  fx.ag_rank = rank; fx.ag_world = world; fx.ag_slot = slot;
  return gemm_dispatch(A_full, B, C, M, N, K, K, ldb, ldc, 0, b_mn, 0, 0, 0, stream, fx);
}
