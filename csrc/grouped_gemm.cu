// Ragged grouped GEMM for MoE experts on tcgen05 tensor cores -- one launch for all experts, the per-expert row
// ranges are read on the device (no host synchronisation, no per-expert launches).
//
// Rows are grouped by expert and every expert's segment is padded to a multiple of 128 rows (padding rows are
// zero), so a 128-row tile never straddles two experts:
//   mode 0  fprop  Y[r,:]  = X[r,:] * W[e(r)]^T          X [R,K]  W [E,N,K]   (A K-major, B K-major)
//   mode 1  dgrad  dX[r,:] = dY[r,:] * W[e(r)]           dY [R,N] W [E,N,K]   (A K-major, B MN-major)
//   mode 2  wgrad  dW[e]  (+)= dY_e^T * X_e              ragged reduction over the rows of expert e (A, B MN-major)
// `tile_expert[t]` gives the expert of row tile t (or -1 past the last used tile); `offs_pad[e]` the first row
// of expert e.  Same persistent warp-specialised structure as csrc/gemm_tcgen05.cu (TMA ring -> tcgen05.mma into a
// double-buffered TMEM accumulator -> TMA-store epilogue); the expert index is the third TMA coordinate of W.
// Reference: the Python loop over experts with nonzero / gather / three nn.Linear / scatter_ in
// scaletorch/models/model_qwen3_moe.py:129-171.
#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

constexpr int kGBM = 128, kGBK = 64, kGThreads = 256;

struct GroupedParams {
  const int* tile_expert;   // modes 0/1: [tiles_m]
  const int* offs_pad;      // mode 2: [E + 1]
  int tiles_m, tiles_n, num_tiles, k_blocks;   // k_blocks: modes 0/1
  int n_out;                // mode 2: rows of one expert's weight (N)
  int out_fp32, accumulate;
};

template <int BN>
struct GGCfg {
  static constexpr int kABytes = kGBM * kGBK * 2;
  static constexpr int kBBytes = BN * kGBK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kTmemCols = 2 * BN;
  static constexpr int kEpiBytes = 4 * 2 * 4096;
  static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
};

// Decodes a tile: returns false when the tile does no work at all (modes 0/1 past the last used row tile).
template <int MODE>
STB_DEVICE bool gg_decode(const GroupedParams& p, int tile, int& e, int& tm, int& tn, int& r0, int& kblocks) {
  if (MODE == 2) {
    const int per = p.tiles_m * p.tiles_n;
    e = tile / per;
    const int r = tile - e * per;
    tm = r % p.tiles_m;
    tn = r / p.tiles_m;
    r0 = p.offs_pad[e];
    kblocks = (p.offs_pad[e + 1] - r0) / kGBK;
    return true;
  }
  tm = tile % p.tiles_m;
  tn = tile / p.tiles_m;
  e = p.tile_expert[tm];
  r0 = 0;
  kblocks = p.k_blocks;
  return e >= 0;
}

template <int BN, int MODE>
__global__ void __launch_bounds__(kGThreads, 1)
    grouped_gemm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                        const __grid_constant__ CUtensorMap tmC, const GroupedParams p) {
  using Cfg = GGCfg<BN>;
  constexpr int kStages = Cfg::kStages;
  constexpr bool A_MN = (MODE == 2), B_MN = (MODE != 0);
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * Cfg::kABytes;
  uint8_t* sEpi = smem + kStages * Cfg::kStageBytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * Cfg::kStageBytes + Cfg::kEpiBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + kStages;
  uint64_t* tmem_full = bars + 2 * kStages;
  uint64_t* tmem_empty = bars + 2 * kStages + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); tma_prefetch_desc(&tmC);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tmem_full[i], 1); mbar_init(&tmem_empty[i], 128); }
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tmem_ptr, Cfg::kTmemCols); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        int e, tm, tn, r0, kblocks;
        if (!gg_decode<MODE>(p, tile, e, tm, tn, r0, kblocks)) continue;
        const int m0 = tm * kGBM, n0 = tn * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
          uint8_t* a_dst = sA + stage * Cfg::kABytes;
          uint8_t* b_dst = sB + stage * Cfg::kBBytes;
          const int k0 = kb * kGBK;
          if (MODE == 0) {
            tma_load_2d(a_dst, &tmA, &full_bar[stage], k0, m0);
            tma_load_3d(b_dst, &tmB, &full_bar[stage], k0, n0, e);
          } else if (MODE == 1) {
            tma_load_2d(a_dst, &tmA, &full_bar[stage], k0, m0);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_3d(b_dst + j * (kGBK * 128), &tmB, &full_bar[stage], n0 + j * 64, k0, e);
          } else {
#pragma unroll
            for (int j = 0; j < kGBM / 64; ++j)
              tma_load_2d(a_dst + j * (kGBK * 128), &tmA, &full_bar[stage], m0 + j * 64, r0 + k0);
#pragma unroll
            for (int j = 0; j < BN / 64; ++j)
              tma_load_2d(b_dst + j * (kGBK * 128), &tmB, &full_bar[stage], n0 + j * 64, r0 + k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(kGBM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t a_lbo = A_MN ? kGBK * 128 : 16, b_lbo = B_MN ? kGBK * 128 : 16;
      constexpr uint32_t a_kstep = A_MN ? 16 * 128 : 32, b_kstep = B_MN ? 16 * 128 : 32;
      int stage = 0; uint32_t phase = 0; int it = 0;
      for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
        int e, tm, tn, r0, kblocks;
        if (!gg_decode<MODE>(p, tile, e, tm, tn, r0, kblocks)) continue;
        if (kblocks == 0) continue;   // wgrad of an expert without tokens: the epilogue writes zeros
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        ++it;
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * Cfg::kABytes), b_addr = smem_u32(sB + stage * Cfg::kBBytes);
#pragma unroll
          for (int k = 0; k < kGBK / 16; ++k)
            umma_f16_ss(tmem_d, umma_smem_desc_sw128(a_addr + k * a_kstep, a_lbo, 1024),
                        umma_smem_desc_sw128(b_addr + k * b_kstep, b_lbo, 1024), idesc, (kb | k) != 0);
          umma_commit(&empty_bar[stage]);
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[acc]);
      }
    }
  } else if (warp >= 4) {
    const int ew = warp - 4;
    uint8_t* slots = sEpi + ew * 8192;
    int slot = 0, it = 0;
    const uint32_t sw = uint32_t(lane & 7);
    for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
      int e, tm, tn, r0, kblocks;
      if (!gg_decode<MODE>(p, tile, e, tm, tn, r0, kblocks)) continue;
      const bool empty = (kblocks == 0);
      if (empty && p.accumulate) continue;
      int acc = 0;
      if (!empty) {
        acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        ++it;
        mbar_wait(&tmem_full[acc], acc_phase);
        tc_fence_after();
      }
      const int row0 = (MODE == 2 ? e * p.n_out : 0) + tm * kGBM + ew * 32;
      const uint32_t taddr = tmem_base + (uint32_t(ew * 32) << 16) + acc * BN;
      if (p.out_fp32) {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          if (!empty) {
            tmem_ld_32x32b_x32(taddr + c * 32, r);
            tmem_ld_wait();
            if (c == BN / 32 - 1) { tc_fence_before(); mbar_arrive(&tmem_empty[acc]); }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = 0;
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
            if (p.accumulate) tma_reduce_add_2d(&tmC, slots + slot * 4096, tn * BN + c * 32, row0);
            else tma_store_2d(&tmC, slots + slot * 4096, tn * BN + c * 32, row0);
            tma_store_commit();
          }
          slot ^= 1;
        }
      } else {
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          uint32_t r0v[32], r1v[32];
          if (!empty) {
            tmem_ld_32x32b_x32(taddr + c * 64, r0v);
            tmem_ld_32x32b_x32(taddr + c * 64 + 32, r1v);
            tmem_ld_wait();
            if (c == BN / 64 - 1) { tc_fence_before(); mbar_arrive(&tmem_empty[acc]); }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) { r0v[j] = 0; r1v[j] = 0; }
          }
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          const uint32_t dst = smem_u32(slots + slot * 4096 + lane * 128);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            st_shared_v4(dst + ((uint32_t(j) ^ sw) << 4),
                         pack_bf16x2(__uint_as_float(r0v[8 * j + 0]), __uint_as_float(r0v[8 * j + 1])),
                         pack_bf16x2(__uint_as_float(r0v[8 * j + 2]), __uint_as_float(r0v[8 * j + 3])),
                         pack_bf16x2(__uint_as_float(r0v[8 * j + 4]), __uint_as_float(r0v[8 * j + 5])),
                         pack_bf16x2(__uint_as_float(r0v[8 * j + 6]), __uint_as_float(r0v[8 * j + 7])));
            st_shared_v4(dst + ((uint32_t(j + 4) ^ sw) << 4),
                         pack_bf16x2(__uint_as_float(r1v[8 * j + 0]), __uint_as_float(r1v[8 * j + 1])),
                         pack_bf16x2(__uint_as_float(r1v[8 * j + 2]), __uint_as_float(r1v[8 * j + 3])),
                         pack_bf16x2(__uint_as_float(r1v[8 * j + 4]), __uint_as_float(r1v[8 * j + 5])),
                         pack_bf16x2(__uint_as_float(r1v[8 * j + 6]), __uint_as_float(r1v[8 * j + 7])));
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmC, slots + slot * 4096, tn * BN + c * 64, row0);
            tma_store_commit();
          }
          slot ^= 1;
        }
      }
    }
    if (lane == 0) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, Cfg::kTmemCols);
}

template <int BN, int MODE>
static int launch_grouped(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GroupedParams& p,
                          cudaStream_t stream) {
  using Cfg = GGCfg<BN>;
  auto kern = grouped_gemm_kernel<BN, MODE>;
  static bool attr_set = false;
  if (!attr_set) {
    STB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set = true;
  }
  const int grid = p.num_tiles < sm_count() ? p.num_tiles : sm_count();
  kern<<<grid, kGThreads, Cfg::kSmemBytes, stream>>>(tmA, tmB, tmC, p);
  STB_CHECK_LAUNCH();
  return 0;
}

}  // namespace stb

using namespace stb;

// mode 0 (fprop): A = X [R_max, K], W [E, N, K], C = Y [R_max, N] bf16
// mode 1 (dgrad): A = dY [R_max, N], W [E, N, K], C = dX [R_max, K] bf16
//   tile_expert: int32 [R_max / 128] (expert of each 128-row tile, -1 = unused); R_max % 128 == 0.
// mode 2 (wgrad): A = dY [R_max, N], B = X [R_max, K], C = dW [E, N, K] (fp32 or bf16; accumulate only for fp32),
//   offs_pad: int32 [E + 1] padded row offsets (multiples of 128); N % 128 == 0.
extern "C" int stb_grouped_gemm_bf16(int mode, const void* A, const void* B, void* C, const int* tile_expert,
                                     const int* offs_pad, int R_max, int N, int K, int E, int out_fp32,
                                     int accumulate, cudaStream_t stream) {
  STB_REQUIRE(mode >= 0 && mode <= 2, "grouped_gemm: bad mode %d", mode);
  STB_REQUIRE(R_max % kGBM == 0, "grouped_gemm: R_max must be a multiple of 128");
  STB_REQUIRE(N % 8 == 0 && K % 8 == 0, "grouped_gemm: N and K must be multiples of 8");
  if (R_max == 0 || E == 0) return 0;
  CUtensorMap tmA, tmB, tmC;
  GroupedParams p;
  p.tile_expert = tile_expert; p.offs_pad = offs_pad;
  p.out_fp32 = out_fp32; p.accumulate = accumulate; p.n_out = N;
  int rc = 0;
  if (mode == 0) {
    const int BN = 256;
    if ((rc = make_tmap_2d(&tmA, A, 2, K, R_max, uint64_t(K) * 2, kGBK, kGBM, 1))) return rc;
    if ((rc = make_tmap_3d(&tmB, B, 2, K, N, E, uint64_t(K) * 2, uint64_t(N) * K * 2, kGBK, BN, 1, 1))) return rc;
    if ((rc = make_tmap_2d(&tmC, C, 2, N, R_max, uint64_t(N) * 2, 64, 32, 1))) return rc;
    p.tiles_m = R_max / kGBM; p.tiles_n = (N + BN - 1) / BN; p.num_tiles = p.tiles_m * p.tiles_n;
    p.k_blocks = (K + kGBK - 1) / kGBK;
    STB_REQUIRE(!out_fp32 && !accumulate, "grouped_gemm fprop writes bf16");
    return launch_grouped<256, 0>(tmA, tmB, tmC, p, stream);
  }
  if (mode == 1) {
    const int BN = 256;
    if ((rc = make_tmap_2d(&tmA, A, 2, N, R_max, uint64_t(N) * 2, kGBK, kGBM, 1))) return rc;
    if ((rc = make_tmap_3d(&tmB, B, 2, K, N, E, uint64_t(K) * 2, uint64_t(N) * K * 2, 64, kGBK, 1, 1))) return rc;
    if ((rc = make_tmap_2d(&tmC, C, 2, K, R_max, uint64_t(K) * 2, 64, 32, 1))) return rc;
    p.tiles_m = R_max / kGBM; p.tiles_n = (K + BN - 1) / BN; p.num_tiles = p.tiles_m * p.tiles_n;
    p.k_blocks = (N + kGBK - 1) / kGBK;
    STB_REQUIRE(!out_fp32 && !accumulate, "grouped_gemm dgrad writes bf16");
    return launch_grouped<256, 1>(tmA, tmB, tmC, p, stream);
  }
  STB_REQUIRE(N % kGBM == 0, "grouped_gemm wgrad: N (rows of an expert weight) must be a multiple of 128");
  STB_REQUIRE(out_fp32 || !accumulate, "grouped_gemm wgrad: accumulation needs an fp32 destination");
  const int BN = 256;
  if ((rc = make_tmap_2d(&tmA, A, 2, N, R_max, uint64_t(N) * 2, 64, kGBK, 1))) return rc;
  if ((rc = make_tmap_2d(&tmB, B, 2, K, R_max, uint64_t(K) * 2, 64, kGBK, 1))) return rc;
  if (out_fp32) rc = make_tmap_2d(&tmC, C, 4, K, uint64_t(E) * N, uint64_t(K) * 4, 32, 32, 1);
  else rc = make_tmap_2d(&tmC, C, 2, K, uint64_t(E) * N, uint64_t(K) * 2, 64, 32, 1);
  if (rc) return rc;
  p.tiles_m = N / kGBM; p.tiles_n = (K + BN - 1) / BN; p.num_tiles = E * p.tiles_m * p.tiles_n; p.k_blocks = 0;
  return launch_grouped<256, 2>(tmA, tmB, tmC, p, stream);
}
