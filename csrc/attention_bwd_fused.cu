// Flash-attention backward in ONE tcgen05 kernel: S and dP are computed once per (key tile, query tile) pair and feed
// all three gradients.
//
//     P  = exp(scale * Q K^T - lse)      dP = dO V^T      dS = P * (dP - delta)
//     dV += P^T dO      dK += scale * dS^T Q      dQ^T(tile) = K^T dS^T  ->  reduce-added into an fp32 dQ in global memory
//
// The two-kernel backward (csrc/attention_bwd.cu: dK/dV kernel + dQ kernel) runs the S / dP GEMMs and -- more expensive on
// this machine -- the whole exp / mask / multiply / convert pass twice; both kernels are bound by those elementwise
// warps, not by the tensor pipe (profiles/ncu_r1a_summary.md).  Here one CTA per (batch, kv head, 128-key tile) keeps K, V
// resident, streams 64-row Q / dO tiles of every query head of the group, and per iteration issues five GEMMs:
//     S^T = K Q^T,  dP^T = V dO^T          [128 keys] x [64 q], accumulators in TMEM (keys on the lanes)
//     dV += P^T dO,  dK += dS^T Q          [128 keys] x [128 D], P^T / dS^T come back from the elementwise warps via smem
//     dQ^T = K^T dS^T                      [128 D] x [64 q]: A = the resident K tile read MN-major, B = dS^T read MN-major
// A fourth warpgroup drains dQ^T (double-buffered in TMEM), transposes it through shared memory and adds the [64 q][128 D]
// fp32 tile to global dQ with one TMA reduce-add; the caller zeroes dQ beforehand and consumes it as fp32 (the QK-norm /
// RoPE backward reads it directly, so no conversion pass exists).
// TMEM (512 columns): S^T 0..63 | dP^T 64..127 | dQ^T[2] 128..255 | dV 256..383 | dK 384..511.
// Warps: 0 TMA producer, 1 MMA issuer, 2 TMEM allocator, 4..11 elementwise (two threads per key row), 12..15 dQ epilogue.
// Reference: autograd of F.scaled_dot_product_attention (scaletorch/models/attention_utils.py:151-152).
#include <math_constants.h>
#include <stdlib.h>

#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {
namespace fbwd {

constexpr int kThreads = 512;
constexpr int kBT = 128;              // resident key rows
constexpr int kST = 64;               // streamed query rows
constexpr int kBigHalf = kBT * 128;   // [128 rows][64 bf16] = 16 KB
constexpr int kBigTile = 2 * kBigHalf;
constexpr int kSmallHalf = kST * 128; // [64 rows][64 bf16] = 8 KB
constexpr int kSmallTile = 2 * kSmallHalf;
constexpr int kRing = 3;
constexpr float kLog2e = 1.4426950408889634f;

struct Params {
  const float* lse;     // [B,Hq,Sq]
  const float* delta;   // [B,Hq,Sq]
  __nv_bfloat16* dk; long dk_sb, dk_ss, dk_sh;
  __nv_bfloat16* dv; long dv_sb, dv_ss, dv_sh;
  int B, Sq, Sk, Hq, Hkv;
  float scale, scale_log2;
  int causal;
};

struct Smem {
  static constexpr int kK = 0;
  static constexpr int kV = kK + kBigTile;
  static constexpr int kQ = kV + kBigTile;              // kRing x 16 KB
  static constexpr int kDO = kQ + kRing * kSmallTile;   // kRing x 16 KB
  static constexpr int kPT = kDO + kRing * kSmallTile;  // [128 keys][64 q] bf16
  static constexpr int kDST = kPT + kBigHalf;           // [128 keys][64 q] bf16
  static constexpr int kDQ = kDST + kBigHalf;           // [64 q][128 D] fp32 staging of the dQ tile (32 KB)
  static constexpr int kStat = kDQ + kST * 128 * 4;     // lse2[2][64], delta[2][64]
  static constexpr int kBar = kStat + 4 * 64 * 4;
  static constexpr int kBytes = kBar + 256 + 1024;
};

STB_DEVICE void tma_reduce_add_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];" ::"l"(
          reinterpret_cast<uint64_t>(m)),
      "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

__global__ void __launch_bounds__(kThreads, 1)
    attn_bwd_fused_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                          const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                          const __grid_constant__ CUtensorMap tmDQ, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem + Smem::kK;
  uint8_t* sV = smem + Smem::kV;
  uint8_t* sQ = smem + Smem::kQ;
  uint8_t* sDO = smem + Smem::kDO;
  uint8_t* sPT = smem + Smem::kPT;
  uint8_t* sDST = smem + Smem::kDST;
  float* sDQ = reinterpret_cast<float*>(smem + Smem::kDQ);
  float* s_lse = reinterpret_cast<float*>(smem + Smem::kStat);   // [2][64] (already * log2e)
  float* s_delta = s_lse + 128;                                  // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Smem::kBar);
  uint64_t* kv_full = bars + 0;
  uint64_t* qdo_full = bars + 1;    // [kRing]
  uint64_t* qdo_empty = bars + 4;   // [kRing]
  uint64_t* sdp_full = bars + 7;    //      S^T and dP^T of an iteration are in TMEM
  uint64_t* st_free = bars + 8;     //      the elementwise warps have drained them (256 arrivals)
  uint64_t* pds_full = bars + 9;    //      P^T / dS^T are in smem (256 arrivals)
  uint64_t* pds_free = bars + 10;   //      the dV / dK / dQ MMAs have consumed them
  uint64_t* dq_full = bars + 11;    // [2]  dQ^T tile complete in TMEM
  uint64_t* dq_free = bars + 13;    // [2]  the epilogue warps have drained it (128 arrivals)
  uint64_t* acc_full = bars + 15;   //      all MMAs retired: dV / dK final
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy key tiles first (under a causal mask key tile 0 streams every query tile)
  const int kvt = blockIdx.x / (p.B * p.Hkv);
  const int bh = blockIdx.x % (p.B * p.Hkv);
  const int b = bh / p.Hkv, hkv = bh % p.Hkv;
  const int G = p.Hq / p.Hkv;
  const int kv0 = kvt * kBT;
  const int shift = p.Sk - p.Sq;
  const int n_qt = (p.Sq + kST - 1) / kST;
  int qt_start = 0;
  if (p.causal) qt_start = max(0, kv0 - shift) / kST;
  const int per_head = max(0, n_qt - qt_start);
  const int n_iter = G * per_head;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    mbar_init(sdp_full, 1);
    mbar_init(st_free, 256);
    mbar_init(pds_full, 256);
    mbar_init(pds_free, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&dq_full[i], 1);
      mbar_init(&dq_free[i], 128);
    }
    mbar_init(acc_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  constexpr uint32_t kColS = 0, kColDP = 64, kColDQ = 128, kColDV = 256, kColDK = 384;

  // register budget per warpgroup (512 threads x 128 = the whole file): producer / issuer warps need almost nothing, the
  // elementwise warps hold 6 x 32 values per thread, the dQ epilogue 64
  // (each setmaxnreg sits at the top of its role branch so that the register allocator ties the budget to the branch)
  if (warp < 4) {
  setmaxnreg_dec<40>();
  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(kv_full, 2 * kBigTile);
      tma_load_4d(sK, &tmK, kv_full, 0, hkv, kv0, b);
      tma_load_4d(sK + kBigHalf, &tmK, kv_full, 64, hkv, kv0, b);
      tma_load_4d(sV, &tmV, kv_full, 0, hkv, kv0, b);
      tma_load_4d(sV + kBigHalf, &tmV, kv_full, 64, hkv, kv0, b);
      for (int i = 0; i < n_iter; ++i) {
        const int st = i % kRing;
        const int hq = hkv * G + i / per_head;
        const int q0 = (qt_start + i % per_head) * kST;
        if (i >= kRing) mbar_wait(&qdo_empty[st], ((i / kRing) - 1) & 1);
        mbar_arrive_expect_tx(&qdo_full[st], 2 * kSmallTile);
        tma_load_4d(sQ + st * kSmallTile, &tmQ, &qdo_full[st], 0, hq, q0, b);
        tma_load_4d(sQ + st * kSmallTile + kSmallHalf, &tmQ, &qdo_full[st], 64, hq, q0, b);
        tma_load_4d(sDO + st * kSmallTile, &tmDO, &qdo_full[st], 0, hq, q0, b);
        tma_load_4d(sDO + st * kSmallTile + kSmallHalf, &tmDO, &qdo_full[st], 64, hq, q0, b);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && n_iter > 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(kBT, kST, 0, 0);   // [128 keys] x [64 q], K = D (both K-major)
      constexpr uint32_t idesc_g = umma_idesc_bf16(kBT, 128, 0, 1);   // [128 keys] x [128 D], K = q, B MN-major
      constexpr uint32_t idesc_q = umma_idesc_bf16(128, kST, 1, 1);   // [128 D] x [64 q], K = keys, A and B MN-major
      const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV), pt_addr = smem_u32(sPT), dst_addr = smem_u32(sDST);
      auto issue_sdp = [&](int i) {
        const int st = i % kRing;
        mbar_wait(&qdo_full[st], (i / kRing) & 1);
        if (i >= 1) mbar_wait(st_free, (i - 1) & 1);      // the elementwise warps have read S^T / dP^T of tile i - 1
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + st * kSmallTile), do_addr = smem_u32(sDO + st * kSmallTile);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kBigHalf + (kk & 3) * 32, b_off = (kk >> 2) * kSmallHalf + (kk & 3) * 32;
          umma_f16_ss(tmem_base + kColS, umma_smem_desc_sw128(k_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(q_addr + b_off, 16, 1024), idesc_s, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kBigHalf + (kk & 3) * 32, b_off = (kk >> 2) * kSmallHalf + (kk & 3) * 32;
          umma_f16_ss(tmem_base + kColDP, umma_smem_desc_sw128(v_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(do_addr + b_off, 16, 1024), idesc_s, kk != 0);
        }
        umma_commit(sdp_full);
      };
      mbar_wait(kv_full, 0);
      issue_sdp(0);
      for (int i = 0; i < n_iter; ++i) {
        // S / dP of the next tile run on the tensor pipe while the elementwise warps work on this one
        if (i + 1 < n_iter) issue_sdp(i + 1);
        const int st = i % kRing, db = i & 1;
        mbar_wait(pds_full, i & 1);
        if (i >= 2) mbar_wait(&dq_free[db], ((i >> 1) - 1) & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + st * kSmallTile), do_addr = smem_u32(sDO + st * kSmallTile);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)   // dV += P^T dO     (K = 64 query rows, 16 per step)
          umma_f16_ss(tmem_base + kColDV, umma_smem_desc_sw128(pt_addr + ks * 32, 16, 1024),
                      umma_smem_desc_sw128(do_addr + ks * 2048, kSmallHalf, 1024), idesc_g, (i | ks) != 0);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)   // dK += dS^T Q
          umma_f16_ss(tmem_base + kColDK, umma_smem_desc_sw128(dst_addr + ks * 32, 16, 1024),
                      umma_smem_desc_sw128(q_addr + ks * 2048, kSmallHalf, 1024), idesc_g, (i | ks) != 0);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)   // dQ^T = K^T dS^T  (K = 128 keys, 16 per step; both operands MN-major)
          umma_f16_ss(tmem_base + kColDQ + db * kST, umma_smem_desc_sw128(k_addr + ks * 2048, kBigHalf, 1024),
                      umma_smem_desc_sw128(dst_addr + ks * 2048, kSmallHalf, 1024), idesc_q, ks != 0);
        umma_commit(&dq_full[db]);
        umma_commit(pds_free);
        umma_commit(&qdo_empty[st]);
      }
      umma_commit(acc_full);
    }
  }
  } else if (warp < 12) {
    setmaxnreg_inc<184>();
    // ------------------------------------------------------------------ elementwise warps
    // two threads per key row: `half` selects 32 of the 64 query columns of the streamed tile
    const int quarter = warp & 3;
    const int half = (warp - 4) >> 2;
    const int row = quarter * 32 + lane;
    const int tid = threadIdx.x - 128;            // 0..255
    const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
    const int key = kv0 + row;
    const bool key_ok = key < p.Sk;
    const uint32_t pt_row = smem_u32(sPT + row * 128), dst_row = smem_u32(sDST + row * 128);
    const uint32_t sw = uint32_t(row & 7);
    for (int i = 0; i < n_iter; ++i) {
      const int hq = hkv * G + i / per_head;
      const int q0 = (qt_start + i % per_head) * kST;
      float stat = 0.f;
      if (tid < 128) {
        const int c = tid & 63, q = q0 + c;
        const size_t idx = (size_t(b) * p.Hq + hq) * p.Sq + q;
        if (tid < 64) stat = (q < p.Sq) ? p.lse[idx] * kLog2e : CUDART_INF_F;
        else stat = (q < p.Sq) ? p.delta[idx] : 0.f;
      }
      const int sb = i & 1;
      mbar_wait(sdp_full, i & 1);
      tc_fence_after();
      // double-buffered statistics: a thread can only be two tiles ahead of the slowest one after passing this barrier
      if (tid < 64) s_lse[sb * 64 + tid] = stat; else if (tid < 128) s_delta[sb * 64 + (tid - 64)] = stat;
      named_bar_sync(2, 256);
      float lse2[32], dl[32];
      {
        const float4* l4 = reinterpret_cast<const float4*>(s_lse + sb * 64 + half * 32);
        const float4* d4 = reinterpret_cast<const float4*>(s_delta + sb * 64 + half * 32);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const float4 a = l4[t], c = d4[t];
          lse2[4 * t] = a.x; lse2[4 * t + 1] = a.y; lse2[4 * t + 2] = a.z; lse2[4 * t + 3] = a.w;
          dl[4 * t] = c.x; dl[4 * t + 1] = c.y; dl[4 * t + 2] = c.z; dl[4 * t + 3] = c.w;
        }
      }
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(tmem_base + lane_addr + kColS + half * 32, r0);     // S^T
      tmem_ld_32x32b_x32(tmem_base + lane_addr + kColDP + half * 32, r1);    // dP^T
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(st_free);
      float pr[32], ds[32];
      const int qpos0 = q0 + shift + half * 32;  // key visible to query column c iff key <= qpos0 + c
      const bool full_tile = (kv0 + kBT <= p.Sk) && (q0 + kST <= p.Sq) && (!p.causal || kv0 + kBT - 1 <= q0 + shift);
      if (full_tile) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          pr[j] = exp2f(fmaf(__uint_as_float(r0[j]), p.scale_log2, -lse2[j]));
          ds[j] = pr[j] * (__uint_as_float(r1[j]) - dl[j]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const float e = exp2f(fmaf(__uint_as_float(r0[j]), p.scale_log2, -lse2[j]));
          const bool ok = key_ok && (!p.causal || key <= qpos0 + j);
          pr[j] = ok ? e : 0.f;
          ds[j] = pr[j] * (__uint_as_float(r1[j]) - dl[j]);
        }
      }
      if (i > 0) mbar_wait(pds_free, (i - 1) & 1);
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const uint32_t cc = (uint32_t(half * 4 + t) ^ sw) << 4;
        st_shared_v4(pt_row + cc, pack_bf16x2(pr[8 * t + 0], pr[8 * t + 1]), pack_bf16x2(pr[8 * t + 2], pr[8 * t + 3]),
                     pack_bf16x2(pr[8 * t + 4], pr[8 * t + 5]), pack_bf16x2(pr[8 * t + 6], pr[8 * t + 7]));
        st_shared_v4(dst_row + cc, pack_bf16x2(ds[8 * t + 0], ds[8 * t + 1]), pack_bf16x2(ds[8 * t + 2], ds[8 * t + 3]),
                     pack_bf16x2(ds[8 * t + 4], ds[8 * t + 5]), pack_bf16x2(ds[8 * t + 6], ds[8 * t + 7]));
      }
      fence_proxy_async_smem();
      mbar_arrive(pds_full);
    }
    // epilogue: dV, dK rows -> bf16 -> global (row = key)
    if (n_iter > 0) {
      mbar_wait(acc_full, 0);
      tc_fence_after();
    }
#pragma unroll 1
    for (int which = 0; which < 2; ++which) {
      __nv_bfloat16* base = which == 0 ? p.dv : p.dk;
      const long sb_ = which == 0 ? p.dv_sb : p.dk_sb, ss_ = which == 0 ? p.dv_ss : p.dk_ss,
                 sh_ = which == 0 ? p.dv_sh : p.dk_sh;
      __nv_bfloat16* dst = base + b * sb_ + long(key) * ss_ + hkv * sh_;
      const float osc = which == 0 ? 1.f : p.scale;       // dK = scale * dS^T Q (dS was accumulated unscaled)
#pragma unroll 1
      for (int c = half * 2; c < half * 2 + 2; ++c) {
        uint32_t r[32];
        if (n_iter > 0) {
          tmem_ld_32x32b_x32(tmem_base + lane_addr + (which == 0 ? kColDV : kColDK) + c * 32, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0;
        }
        if (key_ok) {
#pragma unroll
          for (int t = 0; t < 4; ++t)
            *reinterpret_cast<uint4*>(dst + c * 32 + 8 * t) = make_uint4(
                pack_bf16x2(__uint_as_float(r[8 * t + 0]) * osc, __uint_as_float(r[8 * t + 1]) * osc),
                pack_bf16x2(__uint_as_float(r[8 * t + 2]) * osc, __uint_as_float(r[8 * t + 3]) * osc),
                pack_bf16x2(__uint_as_float(r[8 * t + 4]) * osc, __uint_as_float(r[8 * t + 5]) * osc),
                pack_bf16x2(__uint_as_float(r[8 * t + 6]) * osc, __uint_as_float(r[8 * t + 7]) * osc));
        }
      }
    }
  } else {
    setmaxnreg_dec<96>();
    // ------------------------------------------------------------------ dQ epilogue warps
    // TMEM lane = head-dim index d; 64 columns = the tile's query rows.  Thread d writes column d of the [64 q][128 D]
    // fp32 staging tile (consecutive lanes -> consecutive banks), then one TMA reduce-add sends the tile to global dQ.
    const int quarter = warp & 3;
    const int d = quarter * 32 + lane;
    const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
    const bool leader = (threadIdx.x == 12 * 32);
    for (int i = 0; i < n_iter; ++i) {
      const int db = i & 1;
      const int hq = hkv * G + i / per_head;
      const int q0 = (qt_start + i % per_head) * kST;
      mbar_wait(&dq_full[db], (i >> 1) & 1);
      tc_fence_after();
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(tmem_base + lane_addr + kColDQ + db * kST, r0);
      tmem_ld_32x32b_x32(tmem_base + lane_addr + kColDQ + db * kST + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&dq_free[db]);
      if (leader) tma_store_wait_read<0>();       // the previous reduce-add has finished reading the staging tile
      named_bar_sync(3, 128);
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        sDQ[j * 128 + d] = __uint_as_float(r0[j]) * p.scale;
        sDQ[(32 + j) * 128 + d] = __uint_as_float(r1[j]) * p.scale;
      }
      fence_proxy_async_smem();
      named_bar_sync(3, 128);
      if (leader) {
        tma_reduce_add_4d(&tmDQ, sDQ, 0, hq, q0, b);
        tma_store_commit();
      }
    }
    if (leader) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace fbwd
}  // namespace stb

using namespace stb;

// dout: contiguous [B,Sq,Hq,128] bf16; q,k,v: [B,S,H,128] bf16 views with element strides; lse, delta: [B,Hq,Sq] fp32
// (delta = rowsum(dO * O), e.g. from stb_attn_bwd_delta); dq: fp32 [B,Sq,Hq,128] with element strides, ACCUMULATED INTO
// (zero it first); dk, dv: bf16 with element strides, overwritten.
extern "C" int stb_flash_attn_bwd_fused(const void* dout, const void* q, const void* k, const void* v, const float* lse,
                                        const float* delta, float* dq, void* dk, void* dv, int B, int Sq, int Sk, int Hq,
                                        int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh,
                                        long v_sb, long v_ss, long v_sh, long dq_sb, long dq_ss, long dq_sh, long dk_sb,
                                        long dk_ss, long dk_sh, long dv_sb, long dv_ss, long dv_sh, float scale, int causal,
                                        cudaStream_t stream) {
  using namespace stb::fbwd;
  STB_REQUIRE(D == 128, "flash_attn_bwd_fused: head_dim must be 128 (got %d)", D);
  STB_REQUIRE(Hq % Hkv == 0, "flash_attn_bwd_fused: Hq=%d must be a multiple of Hkv=%d", Hq, Hkv);
  STB_REQUIRE(!causal || Sq <= Sk, "flash_attn_bwd_fused: causal attention needs Sq <= Sk");
  STB_REQUIRE((dk_ss % 8) == 0 && (dv_ss % 8) == 0 && (dk_sh % 8) == 0 && (dv_sh % 8) == 0 && (dq_ss % 4) == 0 &&
                  (dq_sh % 4) == 0 && (dq_sb % 4) == 0,
              "flash_attn_bwd_fused: gradient strides must be multiples of 16 bytes");
  if (B == 0 || Sq == 0 || Sk == 0) return 0;
  Params p;
  p.lse = lse; p.delta = delta;
  p.dk = (__nv_bfloat16*)dk; p.dk_sb = dk_sb; p.dk_ss = dk_ss; p.dk_sh = dk_sh;
  p.dv = (__nv_bfloat16*)dv; p.dv_sb = dv_sb; p.dv_ss = dv_ss; p.dv_sh = dv_sh;
  p.B = B; p.Sq = Sq; p.Sk = Sk; p.Hq = Hq; p.Hkv = Hkv;
  p.scale = scale; p.scale_log2 = scale * kLog2e; p.causal = causal;
  const long do_ss = long(Hq) * 128, do_sb = do_ss * Sq;
  CUtensorMap tmQ, tmK, tmV, tmDO, tmDQ;
  if (int rc = make_tmap_bshd(&tmQ, q, B, Sq, Hq, D, q_sb, q_ss, q_sh, 64, kST)) return rc;
  if (int rc = make_tmap_bshd(&tmDO, dout, B, Sq, Hq, D, do_sb, do_ss, 128, 64, kST)) return rc;
  if (int rc = make_tmap_bshd(&tmK, k, B, Sk, Hkv, D, k_sb, k_ss, k_sh, 64, kBT)) return rc;
  if (int rc = make_tmap_bshd(&tmV, v, B, Sk, Hkv, D, v_sb, v_ss, v_sh, 64, kBT)) return rc;
  if (int rc = make_tmap_bshd_f32(&tmDQ, dq, B, Sq, Hq, D, dq_sb, dq_ss, dq_sh, 128, kST)) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    STB_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem::kBytes));
    attr_set = true;
  }
  const int grid = B * Hkv * ((Sk + kBT - 1) / kBT);
  attn_bwd_fused_kernel<<<grid, kThreads, Smem::kBytes, stream>>>(tmQ, tmK, tmV, tmDO, tmDQ, p);
  STB_CHECK_LAUNCH();
  return 0;
}
