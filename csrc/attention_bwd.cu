// Flash-attention backward on tcgen05 tensor cores (native GQA, strided packed inputs/outputs).
//
// Given dO, Q, K, V, O and the row log-sum-exp of the forward (natural log, possibly the *global* one
// merged across ring-attention blocks):
//     P  = exp(scale * Q K^T - lse)          dP = dO V^T           delta = rowsum(dO * O)
//     dS = P * (dP - delta) * scale          dV = P^T dO           dK = dS^T Q          dQ = dS K
// Reference: autograd of F.scaled_dot_product_attention (scaletorch/models/attention_utils.py:151-152) and the
// eager block backward of scaletorch/parallel/context_parallel/context_parallel.py:305-364.
//
// Three launches, no atomics, deterministic:
//   1. attn_bwd_delta_kernel     delta[b,h,q] (one warp per row)
//   2. attn_bwd_dkdv_kernel      one CTA per (batch, kv-head, 128-key tile): K,V resident, streams 64-row Q/dO
//                                tiles of every q-head of the group; S^T = K Q^T and dP^T = V dO^T land in TMEM
//                                with *keys on the lanes*, so P^T / dS^T are written row-wise into swizzled smem
//                                and feed dV += P^T dO, dK += dS^T Q whose accumulators stay in TMEM.
//   3. attn_bwd_dq_kernel        one CTA per (batch, q-head, 128-query tile): Q,dO resident, streams 64-key K/V
//                                tiles; S = Q K^T, dP = dO V^T, dQ += dS K accumulated in TMEM.
// Warp roles in 2 and 3: warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 elementwise work with two threads
// per accumulator row (32 of the 64 streamed columns each).
#include <math_constants.h>

#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

constexpr int kBwdThreads = 320;  // warp 0 TMA, warp 1 MMA, warps 2..9 elementwise (2 threads per row)
constexpr int kBT = 128;        // resident tile rows
constexpr int kST = 64;         // streamed tile rows
constexpr int kBigHalf = kBT * 128;   // [128 rows][64 bf16] = 16 KB
constexpr int kBigTile = 2 * kBigHalf;
constexpr int kSmallHalf = kST * 128; // [64 rows][64 bf16] = 8 KB
constexpr int kSmallTile = 2 * kSmallHalf;
constexpr float kLog2e = 1.4426950408889634f;
constexpr int kRing = 3;        // streamed-tile ring depth (2 exposes the TMA latency every iteration)

struct AttnBwdParams {
  const __nv_bfloat16* dout;   // [B,Sq,Hq,128] contiguous
  const float* lse;            // [B,Hq,Sq]
  const float* delta;          // [B,Hq,Sq]
  __nv_bfloat16* dq; long dq_sb, dq_ss, dq_sh;
  __nv_bfloat16* dk; long dk_sb, dk_ss, dk_sh;
  __nv_bfloat16* dv; long dv_sb, dv_ss, dv_sh;
  int B, Sq, Sk, Hq, Hkv;
  float scale, scale_log2;
  int causal;
  // blockwise (ring) attention: this call covers query rows [q_lo, q_lo + Sq) of tensors with S_tot rows per batch
  // (lse / delta are [B,Hq,S_tot]); plain attention: q_lo = 0, S_tot = Sq
  int q_lo, S_tot;
  int dq_f32;        // dq is an fp32 buffer (element strides as given) ...
  int dq_acc;        // ... and this call adds to it instead of overwriting
  int dkv_f32_add;   // dk / dv are fp32 buffers, possibly in a peer GPU's memory: contributions are red.add-ed into them
};

STB_DEVICE void red_add_f32x4_sys(float* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

// --------------------------------------------------------------------------------------- delta
__global__ void __launch_bounds__(256)
    attn_bwd_delta_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ out,
                          float* __restrict__ delta, int B, int Sq, int Hq) {
  const int lane = threadIdx.x & 31;
  const long rows = long(B) * Sq * Hq;
  for (long r = long(blockIdx.x) * 8 + (threadIdx.x >> 5); r < rows; r += long(gridDim.x) * 8) {
    const uint2 a = *reinterpret_cast<const uint2*>(dout + r * 128 + lane * 4);
    const uint2 b = *reinterpret_cast<const uint2*>(out + r * 128 + lane * 4);
    const float2 a0 = unpack_bf16x2(a.x), a1 = unpack_bf16x2(a.y), b0 = unpack_bf16x2(b.x), b1 = unpack_bf16x2(b.y);
    float s = a0.x * b0.x + a0.y * b0.y + a1.x * b1.x + a1.y * b1.y;
    s = warp_sum(s);
    if (lane == 0) {
      const long bq = r / Hq;           // b * Sq + q
      const int h = int(r % Hq);
      const long bb = bq / Sq;
      const int q = int(bq % Sq);
      delta[(bb * Hq + h) * Sq + q] = s;
    }
  }
}

// ------------------------------------------------------------------------------------- dK / dV
struct DkvSmem {
  static constexpr int kK = 0;
  static constexpr int kV = kK + kBigTile;
  static constexpr int kQ = kV + kBigTile;             // kRing stages x 16 KB
  static constexpr int kDO = kQ + kRing * kSmallTile;  // kRing stages x 16 KB
  static constexpr int kPT = kDO + kRing * kSmallTile; // [128 keys][64 q] 16 KB
  static constexpr int kDST = kPT + kBigHalf;          // [128 keys][64 q] 16 KB
  static constexpr int kStat = kDST + kBigHalf;        // float lse2[2][64], delta[2][64]
  static constexpr int kBar = kStat + 4 * 64 * 4;
  static constexpr int kBytes = kBar + 256 + 1024;
};

__global__ void __launch_bounds__(kBwdThreads, 1)
    attn_bwd_dkdv_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                         const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                         const AttnBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem + DkvSmem::kK;
  uint8_t* sV = smem + DkvSmem::kV;
  uint8_t* sQ = smem + DkvSmem::kQ;
  uint8_t* sDO = smem + DkvSmem::kDO;
  uint8_t* sPT = smem + DkvSmem::kPT;
  uint8_t* sDST = smem + DkvSmem::kDST;
  float* s_lse = reinterpret_cast<float*>(smem + DkvSmem::kStat);  // [2][64] (already * log2e)
  float* s_delta = s_lse + 128;                                    // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + DkvSmem::kBar);
  uint64_t* kv_full = bars + 0;
  uint64_t* qdo_full = bars + 1;   // [kRing]
  uint64_t* qdo_empty = bars + 4;  // [kRing]
  uint64_t* sdp_full = bars + 7;   // [2]  S^T and dP^T of an iteration are in TMEM
  uint64_t* st_free = bars + 9;    // [2]  softmax threads have drained that TMEM pair
  uint64_t* pds_full = bars + 11;  //      P^T / dS^T are in smem
  uint64_t* pds_free = bars + 12;  //      dV / dK MMAs have consumed them
  uint64_t* acc_full = bars + 13;  //      all MMAs retired: dV / dK final
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_kvt = (p.Sk + kBT - 1) / kBT;
  // heavy tiles first: under a causal mask key tile 0 streams every query tile and the last one only a few, so CTAs are
  // ordered by key tile across all (batch, head) pairs and the SMs that finish early pick up the light tail
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
  }
  if (warp == 1 && lane == 0) {
    mbar_init(kv_full, 1);
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sdp_full[i], 1);
      mbar_init(&st_free[i], 256);
    }
    mbar_init(pds_full, 256);
    mbar_init(pds_free, 1);
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
  // TMEM columns: S^T[2] at 0/64, dP^T[2] at 128/192, dV at 256, dK at 384

  if (warp == 0) {
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
    if (lane == 0 && n_iter > 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(kBT, kST, 0, 0);    // [128 keys] x [64 q], K = D
      constexpr uint32_t idesc_g = umma_idesc_bf16(kBT, 128, 0, 1);    // [128 keys] x [128 D], K = q, B MN-major
      const uint32_t k_addr = smem_u32(sK), v_addr = smem_u32(sV), pt_addr = smem_u32(sPT), dst_addr = smem_u32(sDST);
      auto issue_sdp = [&](int i) {
        const int st = i % kRing, tb = i & 1;
        mbar_wait(&qdo_full[st], (i / kRing) & 1);
        if (i >= 2) mbar_wait(&st_free[tb], ((i >> 1) - 1) & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + st * kSmallTile), do_addr = smem_u32(sDO + st * kSmallTile);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kBigHalf + (kk & 3) * 32, b_off = (kk >> 2) * kSmallHalf + (kk & 3) * 32;
          umma_f16_ss(tmem_base + tb * kST, umma_smem_desc_sw128(k_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(q_addr + b_off, 16, 1024), idesc_s, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kBigHalf + (kk & 3) * 32, b_off = (kk >> 2) * kSmallHalf + (kk & 3) * 32;
          umma_f16_ss(tmem_base + 128 + tb * kST, umma_smem_desc_sw128(v_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(do_addr + b_off, 16, 1024), idesc_s, kk != 0);
        }
        umma_commit(&sdp_full[tb]);
      };
      mbar_wait(kv_full, 0);
      issue_sdp(0);
      for (int i = 0; i < n_iter; ++i) {
        if (i + 1 < n_iter) issue_sdp(i + 1);
        const int st = i % kRing;
        mbar_wait(pds_full, i & 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + st * kSmallTile), do_addr = smem_u32(sDO + st * kSmallTile);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)   // dV += P^T dO     (K = 64 query rows, 16 per step)
          umma_f16_ss(tmem_base + 256, umma_smem_desc_sw128(pt_addr + ks * 32, 16, 1024),
                      umma_smem_desc_sw128(do_addr + ks * 2048, kSmallHalf, 1024), idesc_g, (i | ks) != 0);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)   // dK += dS^T Q
          umma_f16_ss(tmem_base + 384, umma_smem_desc_sw128(dst_addr + ks * 32, 16, 1024),
                      umma_smem_desc_sw128(q_addr + ks * 2048, kSmallHalf, 1024), idesc_g, (i | ks) != 0);
        umma_commit(pds_free);
        umma_commit(&qdo_empty[st]);
      }
      umma_commit(acc_full);
    }
  } else {
    // two threads per key row: `half` selects 32 of the 64 query columns of the streamed tile
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quarter * 32 + lane;
    const int tid = threadIdx.x - 64;             // 0..255
    const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
    const int key = kv0 + row;
    const bool key_ok = key < p.Sk;
    const uint32_t pt_row = smem_u32(sPT + row * 128), dst_row = smem_u32(sDST + row * 128);
    const uint32_t sw = uint32_t(row & 7);
    for (int i = 0; i < n_iter; ++i) {
      const int sb = i & 1;
      const int hq = hkv * G + i / per_head;
      const int q0 = (qt_start + i % per_head) * kST;
      // stage this tile's lse / delta (issued before the wait so the latency hides behind the MMAs)
      float stat = 0.f;
      if (tid < 128) {
        const int c = tid & 63, q = q0 + c;
        const size_t idx = (size_t(b) * p.Hq + hq) * p.S_tot + p.q_lo + q;
        if (tid < 64) stat = (q < p.Sq) ? p.lse[idx] * kLog2e : CUDART_INF_F;
        else stat = (q < p.Sq) ? p.delta[idx] : 0.f;
      }
      mbar_wait(&sdp_full[sb], (i >> 1) & 1);
      tc_fence_after();
      if (tid < 64) s_lse[sb * 64 + tid] = stat; else if (tid < 128) s_delta[sb * 64 + (tid - 64)] = stat;
      named_bar_sync(2, 256);
      // per-column statistics of this thread's 32 query columns: 16 x LDS.128 (broadcast) instead of 64 x LDS.32
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
      tmem_ld_32x32b_x32(tmem_base + lane_addr + sb * kST + half * 32, r0);          // S^T
      tmem_ld_32x32b_x32(tmem_base + lane_addr + 128 + sb * kST + half * 32, r1);    // dP^T
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&st_free[sb]);
      float pr[32], ds[32];
      const int qpos0 = q0 + shift + half * 32;  // key visible to query column c iff key <= qpos0 + c
      // tiles strictly below the diagonal (and inside Sk / Sq) need no mask: uniform branch, 3 instructions fewer per
      // element.  dS is left unscaled here; the softmax scale is applied once per dK element in the epilogue.
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
#pragma unroll 1
      for (int c = half * 2; c < half * 2 + 2; ++c) {
        uint32_t r[32];
        if (n_iter > 0) {
          tmem_ld_32x32b_x32(tmem_base + lane_addr + 256 + which * 128 + c * 32, r);
          tmem_ld_wait();
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = 0;
        }
        const float osc = which == 0 ? 1.f : p.scale;     // dK = scale * dS^T Q (dS was accumulated unscaled)
        if (key_ok && p.dkv_f32_add) {
          // ring attention: add this rank's contribution to the block owner's fp32 accumulator (peer memory)
          float* fdst = reinterpret_cast<float*>(base) + b * sb_ + long(key) * ss_ + hkv * sh_ + c * 32;
          if (n_iter > 0) {
#pragma unroll
            for (int t = 0; t < 8; ++t)
              red_add_f32x4_sys(fdst + 4 * t, __uint_as_float(r[4 * t + 0]) * osc, __uint_as_float(r[4 * t + 1]) * osc,
                                __uint_as_float(r[4 * t + 2]) * osc, __uint_as_float(r[4 * t + 3]) * osc);
          }
        } else if (key_ok) {
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
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ------------------------------------------------------------------------------------------ dQ
struct DqSmem {
  static constexpr int kQ = 0;
  static constexpr int kDO = kQ + kBigTile;
  static constexpr int kK = kDO + kBigTile;            // kRing stages x 16 KB
  static constexpr int kV = kK + kRing * kSmallTile;   // kRing stages x 16 KB
  static constexpr int kDS = kV + kRing * kSmallTile;  // [128 q][64 keys] 16 KB
  static constexpr int kBar = kDS + kBigHalf;
  static constexpr int kBytes = kBar + 256 + 1024;
};

__global__ void __launch_bounds__(kBwdThreads, 1)
    attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmDO,
                       const AttnBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + DqSmem::kQ;
  uint8_t* sDO = smem + DqSmem::kDO;
  uint8_t* sK = smem + DqSmem::kK;
  uint8_t* sV = smem + DqSmem::kV;
  uint8_t* sDS = smem + DqSmem::kDS;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + DqSmem::kBar);
  uint64_t* qdo_full = bars + 0;
  uint64_t* kv_full = bars + 1;    // [kRing]
  uint64_t* kv_empty = bars + 4;   // [kRing]
  uint64_t* sdp_full = bars + 7;   // [2]
  uint64_t* st_free = bars + 9;    // [2]
  uint64_t* ds_full = bars + 11;
  uint64_t* ds_free = bars + 12;
  uint64_t* acc_full = bars + 13;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_qt = (p.Sq + kBT - 1) / kBT;
  const int bh = blockIdx.x % (p.B * p.Hq);
  const int qt = n_qt - 1 - blockIdx.x / (p.B * p.Hq);   // heavy tiles first
  const int b = bh / p.Hq, hq = bh % p.Hq;
  const int hkv = hq / (p.Hq / p.Hkv);
  const int q0 = qt * kBT;
  const int shift = p.Sk - p.Sq;
  int kv_end = p.Sk;
  if (p.causal) kv_end = min(p.Sk, q0 + kBT + shift);
  const int n_iter = (kv_end + kST - 1) / kST;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV); tma_prefetch_desc(&tmDO);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(qdo_full, 1);
    for (int i = 0; i < kRing; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&sdp_full[i], 1);
      mbar_init(&st_free[i], 256);
    }
    mbar_init(ds_full, 256);
    mbar_init(ds_free, 1);
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
  // TMEM columns: S[2] at 0/64, dP[2] at 128/192, dQ at 256

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(qdo_full, 2 * kBigTile);
      tma_load_4d(sQ, &tmQ, qdo_full, 0, hq, q0, b);
      tma_load_4d(sQ + kBigHalf, &tmQ, qdo_full, 64, hq, q0, b);
      tma_load_4d(sDO, &tmDO, qdo_full, 0, hq, q0, b);
      tma_load_4d(sDO + kBigHalf, &tmDO, qdo_full, 64, hq, q0, b);
      for (int i = 0; i < n_iter; ++i) {
        const int st = i % kRing;
        if (i >= kRing) mbar_wait(&kv_empty[st], ((i / kRing) - 1) & 1);
        mbar_arrive_expect_tx(&kv_full[st], 2 * kSmallTile);
        tma_load_4d(sK + st * kSmallTile, &tmK, &kv_full[st], 0, hkv, i * kST, b, kEvictLast);
        tma_load_4d(sK + st * kSmallTile + kSmallHalf, &tmK, &kv_full[st], 64, hkv, i * kST, b, kEvictLast);
        tma_load_4d(sV + st * kSmallTile, &tmV, &kv_full[st], 0, hkv, i * kST, b, kEvictLast);
        tma_load_4d(sV + st * kSmallTile + kSmallHalf, &tmV, &kv_full[st], 64, hkv, i * kST, b, kEvictLast);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(kBT, kST, 0, 0);   // [128 q] x [64 keys], K = D
      constexpr uint32_t idesc_g = umma_idesc_bf16(kBT, 128, 0, 1);   // [128 q] x [128 D], K = keys, B MN-major
      const uint32_t q_addr = smem_u32(sQ), do_addr = smem_u32(sDO), ds_addr = smem_u32(sDS);
      auto issue_sdp = [&](int i) {
        const int st = i % kRing, tb = i & 1;
        mbar_wait(&kv_full[st], (i / kRing) & 1);
        if (i >= 2) mbar_wait(&st_free[tb], ((i >> 1) - 1) & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + st * kSmallTile), v_addr = smem_u32(sV + st * kSmallTile);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kBigHalf + (kk & 3) * 32, b_off = (kk >> 2) * kSmallHalf + (kk & 3) * 32;
          umma_f16_ss(tmem_base + tb * kST, umma_smem_desc_sw128(q_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(k_addr + b_off, 16, 1024), idesc_s, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kBigHalf + (kk & 3) * 32, b_off = (kk >> 2) * kSmallHalf + (kk & 3) * 32;
          umma_f16_ss(tmem_base + 128 + tb * kST, umma_smem_desc_sw128(do_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(v_addr + b_off, 16, 1024), idesc_s, kk != 0);
        }
        umma_commit(&sdp_full[tb]);
      };
      mbar_wait(qdo_full, 0);
      issue_sdp(0);
      for (int i = 0; i < n_iter; ++i) {
        if (i + 1 < n_iter) issue_sdp(i + 1);
        const int st = i % kRing;
        mbar_wait(ds_full, i & 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + st * kSmallTile);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks)   // dQ += dS K     (K = 64 keys, 16 per step; K tile MN-major)
          umma_f16_ss(tmem_base + 256, umma_smem_desc_sw128(ds_addr + ks * 32, 16, 1024),
                      umma_smem_desc_sw128(k_addr + ks * 2048, kSmallHalf, 1024), idesc_g, (i | ks) != 0);
        umma_commit(ds_free);
        umma_commit(&kv_empty[st]);
      }
      umma_commit(acc_full);
    }
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;     // which 32 of the 64 streamed key columns
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
    const int q = q0 + row;
    const bool q_ok = q < p.Sq;
    const size_t sidx = (size_t(b) * p.Hq + hq) * p.S_tot + p.q_lo + (q_ok ? q : 0);
    const float lse2 = q_ok ? p.lse[sidx] * kLog2e : CUDART_INF_F;
    const float dl = q_ok ? p.delta[sidx] : 0.f;
    const int q_pos = q + shift;
    const uint32_t ds_row = smem_u32(sDS + row * 128);
    const uint32_t sw = uint32_t(row & 7);
    for (int i = 0; i < n_iter; ++i) {
      const int sb = i & 1;
      const int key0 = i * kST + half * 32;
      mbar_wait(&sdp_full[sb], (i >> 1) & 1);
      tc_fence_after();
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(tmem_base + lane_addr + sb * kST + half * 32, r0);          // S
      tmem_ld_32x32b_x32(tmem_base + lane_addr + 128 + sb * kST + half * 32, r1);    // dP
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&st_free[sb]);
      float ds[32];
      // tiles strictly below the diagonal need no mask (uniform branch); the softmax scale is applied once per dQ
      // element in the epilogue instead of once per dS element
      const bool full_tile = (i * kST + kST <= p.Sk) && (!p.causal || i * kST + kST - 1 <= q0 + shift);
      if (full_tile) {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          ds[j] = exp2f(fmaf(__uint_as_float(r0[j]), p.scale_log2, -lse2)) * (__uint_as_float(r1[j]) - dl);
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int key = key0 + j;
          const float e = exp2f(fmaf(__uint_as_float(r0[j]), p.scale_log2, -lse2));
          const bool ok = (key < p.Sk) && (!p.causal || key <= q_pos);
          ds[j] = (ok ? e : 0.f) * (__uint_as_float(r1[j]) - dl);
        }
      }
      if (i > 0) mbar_wait(ds_free, (i - 1) & 1);
#pragma unroll
      for (int t = 0; t < 4; ++t)
        st_shared_v4(ds_row + ((uint32_t(half * 4 + t) ^ sw) << 4), pack_bf16x2(ds[8 * t + 0], ds[8 * t + 1]),
                     pack_bf16x2(ds[8 * t + 2], ds[8 * t + 3]), pack_bf16x2(ds[8 * t + 4], ds[8 * t + 5]),
                     pack_bf16x2(ds[8 * t + 6], ds[8 * t + 7]));
      fence_proxy_async_smem();
      mbar_arrive(ds_full);
    }
    mbar_wait(acc_full, 0);
    tc_fence_after();
    __nv_bfloat16* dst = p.dq + b * p.dq_sb + long(q) * p.dq_ss + hq * p.dq_sh;
#pragma unroll 1
    for (int c = half * 2; c < half * 2 + 2; ++c) {
      uint32_t r[32];
      tmem_ld_32x32b_x32(tmem_base + lane_addr + 256 + c * 32, r);
      tmem_ld_wait();
      const float osc = p.scale;                           // dQ = scale * dS K
      if (q_ok && p.dq_f32) {
        // ring attention: dQ accumulates over the key blocks in fp32 (no per-block bf16 round trip, no eager add)
        float4* fdst = reinterpret_cast<float4*>(reinterpret_cast<float*>(p.dq) + b * p.dq_sb + long(q) * p.dq_ss +
                                                 hq * p.dq_sh + c * 32);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          float4 v = make_float4(__uint_as_float(r[4 * t + 0]) * osc, __uint_as_float(r[4 * t + 1]) * osc,
                                 __uint_as_float(r[4 * t + 2]) * osc, __uint_as_float(r[4 * t + 3]) * osc);
          if (p.dq_acc) {
            const float4 o = fdst[t];
            v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
          }
          fdst[t] = v;
        }
      } else if (q_ok) {
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
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace stb

using namespace stb;

// dout/out: contiguous [B,Sq,Hq,128]; q,k,v,dq,dk,dv: [B,S,H,128] views with element strides; lse: [B,Hq,Sq];
// delta: [B,Hq,Sq] fp32 scratch written by this call.
static int flash_attn_bwd_impl(const void* dout, const void* q, const void* k, const void* v, const void* out,
                               const float* lse, float* delta, void* dq, void* dk, void* dv, int B, int Sq,
                               int Sk, int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb,
                               long k_ss, long k_sh, long v_sb, long v_ss, long v_sh, long dq_sb, long dq_ss,
                               long dq_sh, long dk_sb, long dk_ss, long dk_sh, long dv_sb, long dv_ss, long dv_sh,
                               float scale, int causal, int q_lo, int S_tot, int compute_delta, int dq_f32, int dq_acc,
                               int dkv_f32_add, cudaStream_t stream) {
  STB_REQUIRE(D == 128, "flash_attn_bwd: head_dim must be 128 (got %d)", D);
  STB_REQUIRE(Hq % Hkv == 0, "flash_attn_bwd: Hq=%d must be a multiple of Hkv=%d", Hq, Hkv);
  STB_REQUIRE(!causal || Sq <= Sk, "flash_attn_bwd: causal attention needs Sq <= Sk");
  STB_REQUIRE((dq_ss % 8) == 0 && (dk_ss % 8) == 0 && (dv_ss % 8) == 0 && (dq_sh % 8) == 0 && (dk_sh % 8) == 0 &&
                  (dv_sh % 8) == 0,
              "flash_attn_bwd: gradient strides must be multiples of 8 elements");
  STB_REQUIRE(q_lo >= 0 && q_lo + Sq <= S_tot, "flash_attn_bwd: query window [%d, %d) outside %d rows", q_lo, q_lo + Sq, S_tot);
  if (B == 0 || Sq == 0 || Sk == 0) return 0;
  if (compute_delta) {   // over all S_tot rows of dout / out (once per backward, also for blockwise attention)
    const long rows = long(B) * S_tot * Hq;
    long blocks = (rows + 7) / 8;
    const long cap = long(sm_count()) * 16;
    attn_bwd_delta_kernel<<<int(blocks < cap ? blocks : cap), 256, 0, stream>>>(
        (const __nv_bfloat16*)dout, (const __nv_bfloat16*)out, delta, B, S_tot, Hq);
    STB_CHECK_LAUNCH();
  }
  AttnBwdParams p;
  p.dout = (const __nv_bfloat16*)dout; p.lse = lse; p.delta = delta;
  p.dq = (__nv_bfloat16*)dq; p.dq_sb = dq_sb; p.dq_ss = dq_ss; p.dq_sh = dq_sh;
  p.dk = (__nv_bfloat16*)dk; p.dk_sb = dk_sb; p.dk_ss = dk_ss; p.dk_sh = dk_sh;
  p.dv = (__nv_bfloat16*)dv; p.dv_sb = dv_sb; p.dv_ss = dv_ss; p.dv_sh = dv_sh;
  p.B = B; p.Sq = Sq; p.Sk = Sk; p.Hq = Hq; p.Hkv = Hkv;
  p.scale = scale; p.scale_log2 = scale * kLog2e; p.causal = causal;
  p.q_lo = q_lo; p.S_tot = S_tot; p.dq_f32 = dq_f32; p.dq_acc = dq_acc; p.dkv_f32_add = dkv_f32_add;
  const long do_ss = long(Hq) * 128, do_sb = do_ss * S_tot;
  dout = static_cast<const __nv_bfloat16*>(dout) + long(q_lo) * do_ss;     // first row of the query window
  static bool attr_set = false;
  if (!attr_set) {
    STB_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dkdv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DkvSmem::kBytes));
    STB_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DqSmem::kBytes));
    attr_set = true;
  }
  {  // dK / dV: resident K,V (128-row boxes), streamed Q,dO (64-row boxes)
    CUtensorMap tmQ, tmK, tmV, tmDO;
    if (int rc = make_tmap_bshd(&tmQ, q, B, Sq, Hq, D, q_sb, q_ss, q_sh, 64, kST)) return rc;
    if (int rc = make_tmap_bshd(&tmDO, dout, B, Sq, Hq, D, do_sb, do_ss, 128, 64, kST)) return rc;
    if (int rc = make_tmap_bshd(&tmK, k, B, Sk, Hkv, D, k_sb, k_ss, k_sh, 64, kBT)) return rc;
    if (int rc = make_tmap_bshd(&tmV, v, B, Sk, Hkv, D, v_sb, v_ss, v_sh, 64, kBT)) return rc;
    const int grid = B * Hkv * ((Sk + kBT - 1) / kBT);
    attn_bwd_dkdv_kernel<<<grid, kBwdThreads, DkvSmem::kBytes, stream>>>(tmQ, tmK, tmV, tmDO, p);
    STB_CHECK_LAUNCH();
  }
  {  // dQ: resident Q,dO (128-row boxes), streamed K,V (64-row boxes)
    CUtensorMap tmQ, tmK, tmV, tmDO;
    if (int rc = make_tmap_bshd(&tmQ, q, B, Sq, Hq, D, q_sb, q_ss, q_sh, 64, kBT)) return rc;
    if (int rc = make_tmap_bshd(&tmDO, dout, B, Sq, Hq, D, do_sb, do_ss, 128, 64, kBT)) return rc;
    if (int rc = make_tmap_bshd(&tmK, k, B, Sk, Hkv, D, k_sb, k_ss, k_sh, 64, kST)) return rc;
    if (int rc = make_tmap_bshd(&tmV, v, B, Sk, Hkv, D, v_sb, v_ss, v_sh, 64, kST)) return rc;
    const int grid = B * Hq * ((Sq + kBT - 1) / kBT);
    attn_bwd_dq_kernel<<<grid, kBwdThreads, DqSmem::kBytes, stream>>>(tmQ, tmK, tmV, tmDO, p);
    STB_CHECK_LAUNCH();
  }
  return 0;
}

extern "C" int stb_flash_attn_bwd(const void* dout, const void* q, const void* k, const void* v, const void* out,
                                  const float* lse, float* delta, void* dq, void* dk, void* dv, int B, int Sq,
                                  int Sk, int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb,
                                  long k_ss, long k_sh, long v_sb, long v_ss, long v_sh, long dq_sb, long dq_ss,
                                  long dq_sh, long dk_sb, long dk_ss, long dk_sh, long dv_sb, long dv_ss, long dv_sh,
                                  float scale, int causal, cudaStream_t stream) {
  return flash_attn_bwd_impl(dout, q, k, v, out, lse, delta, dq, dk, dv, B, Sq, Sk, Hq, Hkv, D, q_sb, q_ss, q_sh, k_sb,
                             k_ss, k_sh, v_sb, v_ss, v_sh, dq_sb, dq_ss, dq_sh, dk_sb, dk_ss, dk_sh, dv_sb, dv_ss, dv_sh,
                             scale, causal, 0, Sq, 1, 0, 0, 0, stream);
}

// One key block of the backward of a blockwise (ring) attention.  dout / out: contiguous [B,S_tot,Hq,128]; lse / delta:
// [B,Hq,S_tot] (the GLOBAL log-sum-exp over all key blocks); the call covers query rows [q_lo, q_lo + Sq) -- q / dq point
// at row q_lo -- against the Sk keys of this block.  delta is computed when compute_delta != 0 (first block of a
// backward).  dq: fp32, overwritten (dq_acc = 0) or accumulated (dq_acc = 1).  dk / dv: fp32 accumulators that may live
// in another GPU's memory; this rank's contributions are added with red.global.add (the owner zeroes them beforehand).
extern "C" int stb_flash_attn_bwd_block(const void* dout, const void* q, const void* k, const void* v, const void* out,
                                        const float* lse, float* delta, float* dq, float* dk, float* dv, int B, int Sq,
                                        int Sk, int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb,
                                        long k_ss, long k_sh, long v_sb, long v_ss, long v_sh, long dq_sb, long dq_ss,
                                        long dq_sh, long dk_sb, long dk_ss, long dk_sh, long dv_sb, long dv_ss,
                                        long dv_sh, float scale, int causal, int q_lo, int S_tot, int compute_delta,
                                        int dq_acc, cudaStream_t stream) {
  return flash_attn_bwd_impl(dout, q, k, v, out, lse, delta, dq, dk, dv, B, Sq, Sk, Hq, Hkv, D, q_sb, q_ss, q_sh, k_sb,
                             k_ss, k_sh, v_sb, v_ss, v_sh, dq_sb, dq_ss, dq_sh, dk_sb, dk_ss, dk_sh, dv_sb, dv_ss, dv_sh,
                             scale, causal, q_lo, S_tot, compute_delta, 1, dq_acc, 1, stream);
}

// delta[b,h,q] = rowsum(dO * O) for contiguous [B,Sq,Hq,128] dout / out (the fused backward takes it as an input).
extern "C" int stb_attn_bwd_delta(const void* dout, const void* out, float* delta, int B, int Sq, int Hq,
                                  cudaStream_t stream) {
  const long rows = long(B) * Sq * Hq;
  if (rows == 0) return 0;
  long blocks = (rows + 7) / 8;
  const long cap = long(sm_count()) * 16;
  attn_bwd_delta_kernel<<<int(blocks < cap ? blocks : cap), 256, 0, stream>>>(
      (const __nv_bfloat16*)dout, (const __nv_bfloat16*)out, delta, B, Sq, Hq);
  STB_CHECK_LAUNCH();
  return 0;
}
