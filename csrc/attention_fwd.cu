// Flash-attention forward on tcgen05 tensor cores with native GQA.
//
//   out[b,q,h,:] = softmax(scale * Q K^T + causal mask) V        lse[b,h,q] = log sum exp (natural log)
//
// Q/K/V are token-major [B, S, H, 128] views with arbitrary batch/token/head strides (they are read
// straight out of the packed QKV projection), K/V carry Hkv <= Hq heads and are never expanded.
// (out, lse) is exactly what ring attention needs to merge partial results.
// Reference: F.scaled_dot_product_attention after an explicit K/V expansion
// (scaletorch/models/attention_utils.py:130-152, scaletorch/models/model_qwen3.py:219-238) and the eager
// per-block attention of scaletorch/parallel/context_parallel/context_parallel.py:266-302.
//
// One CTA per (batch, q-head, 128-query tile), 12 warps in 3 warpgroups (setmaxnreg moves registers from
// warpgroup 0 to the softmax warpgroups):
//   warp 0      TMA producer: Q once, K/V tiles of 128 keys through a 2-stage ring (4-D tensor maps)
//   warp 1      MMA issuer  : S = Q K^T (128x128x128) and O_j = P V (128x128x128), accumulators in TMEM
//   warp 2      TMEM allocation
//   warps 4..11 softmax     : two threads per query row (64 score columns / 64 output columns each):
//               tcgen05.ld S -> online max / exp2 -> P (bf16) into 128B-swizzled smem for the PV MMA;
//               the running output is kept in registers and rescaled when the row max grows.
// TMEM: S and O_j are both double-buffered (4 x 128 columns) so QK^T of tile j+1 and PV of tile j run
// while the softmax warps work on tile j / accumulate tile j-1.
#include <math_constants.h>

#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

constexpr int kAttnThreads = 384;  // warpgroup 0: TMA / MMA / TMEM-alloc warps; warpgroups 1,2: softmax
constexpr int kTile = 128;  // query rows per CTA and keys per step
constexpr int kHeadDim = 128;
constexpr int kHalfBytes = kTile * 128;      // one [128 rows][64 bf16] swizzled half tile = 16 KB
constexpr int kTileBytes = 2 * kHalfBytes;   // 32 KB

struct AttnFwdParams {
  __nv_bfloat16* out;
  float* lse;
  float* acc;     // optional fp32 [B,Sq,Hq,128] running output of a blockwise (ring) attention; nullptr = none
  int merge;      // 1: (acc | out, lse) hold the result over earlier key blocks: fold this block in (log-sum-exp merge)
  int q_lo, S_tot;  // the call covers query rows [q_lo, q_lo + Sq) of out / acc ([B,S_tot,Hq,128]) and lse ([B,Hq,S_tot])
  int B, Sq, Sk, Hq, Hkv;
  int n_qtiles;
  float scale_log2;
  int causal;
};

struct AttnSmem {
  static constexpr int kQ = 0;
  static constexpr int kK = kQ + kTileBytes;
  static constexpr int kV = kK + 2 * kTileBytes;
  static constexpr int kP = kV + 2 * kTileBytes;
  static constexpr int kX = kP + kTileBytes;          // float xmax[2][2][128], xsum[2][128]
  static constexpr int kBar = kX + 3 * 2 * 128 * 4;
  static constexpr int kBytes = kBar + 256 + 1024;
};

__global__ void __launch_bounds__(kAttnThreads, 1)
    flash_attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                          const __grid_constant__ CUtensorMap tmV, const AttnFwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem + AttnSmem::kQ;
  uint8_t* sK = smem + AttnSmem::kK;
  uint8_t* sV = smem + AttnSmem::kV;
  uint8_t* sP = smem + AttnSmem::kP;
  float* xmax = reinterpret_cast<float*>(smem + AttnSmem::kX);   // [2 buffers][2 halves][128 rows]
  float* xsum = xmax + 2 * 2 * 128;                              // [2 halves][128 rows]
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + AttnSmem::kBar);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;    // [2]
  uint64_t* v_full = bars + 3;    // [2]
  uint64_t* k_empty = bars + 5;   // [2]  K stage reusable once its QK^T retired
  uint64_t* v_empty = bars + 7;   // [2]  V stage reusable once its PV retired
  uint64_t* s_full = bars + 9;    // [2]
  uint64_t* s_free = bars + 11;   // [2]
  uint64_t* p_full = bars + 13;
  uint64_t* p_free = bars + 14;
  uint64_t* o_full = bars + 15;   // [2]
  uint64_t* o_free = bars + 17;   // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 19);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // heavy (late, for causal) query tiles first
  const int bh = blockIdx.x % (p.B * p.Hq);
  const int qt = p.n_qtiles - 1 - blockIdx.x / (p.B * p.Hq);
  const int b = bh / p.Hq, hq = bh % p.Hq;
  const int hkv = hq / (p.Hq / p.Hkv);
  const int q0 = qt * kTile;
  const int causal_shift = p.Sk - p.Sq;  // query i may attend keys <= i + shift
  int kv_end = p.Sk;
  if (p.causal) kv_end = min(p.Sk, q0 + kTile + causal_shift);
  const int n_kv = (kv_end + kTile - 1) / kTile;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmK);
    tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], 256);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], 256);
    }
    mbar_init(p_full, 256);
    mbar_init(p_free, 1);
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

  if (warp < 4) {
  setmaxnreg_dec<40>();
  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      mbar_arrive_expect_tx(q_full, kTileBytes);
      tma_load_4d(sQ, &tmQ, q_full, 0, hq, q0, b);
      tma_load_4d(sQ + kHalfBytes, &tmQ, q_full, 64, hq, q0, b);
      auto load_k = [&](int t) {
        const int st = t & 1;
        if (t >= 2) mbar_wait(&k_empty[st], ((t >> 1) - 1) & 1);
        mbar_arrive_expect_tx(&k_full[st], kTileBytes);
        tma_load_4d(sK + st * kTileBytes, &tmK, &k_full[st], 0, hkv, t * kTile, b, kEvictLast);
        tma_load_4d(sK + st * kTileBytes + kHalfBytes, &tmK, &k_full[st], 64, hkv, t * kTile, b, kEvictLast);
      };
      load_k(0);
      // K runs one tile ahead of V: K[j+1] is needed by the QK^T issued right after PV[j-1]
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) load_k(j + 1);
        const int st = j & 1;
        if (j >= 2) mbar_wait(&v_empty[st], ((j >> 1) - 1) & 1);
        mbar_arrive_expect_tx(&v_full[st], kTileBytes);
        tma_load_4d(sV + st * kTileBytes, &tmV, &v_full[st], 0, hkv, j * kTile, b, kEvictLast);
        tma_load_4d(sV + st * kTileBytes + kHalfBytes, &tmV, &v_full[st], 64, hkv, j * kTile, b, kEvictLast);
      }
    }
  } else if (warp == 1) {
    // --------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc_bf16(kTile, kTile, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(kTile, kHeadDim, 0, 1);
      const uint32_t q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
      auto issue_qk = [&](int t) {
        const int st = t & 1;
        const uint32_t k_addr = smem_u32(sK + st * kTileBytes);
        const uint32_t d = tmem_base + st * kTile;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;
          umma_f16_ss(d, umma_smem_desc_sw128(q_addr + off, 16, 1024), umma_smem_desc_sw128(k_addr + off, 16, 1024),
                      idesc_qk, kk != 0);
        }
        umma_commit(&s_full[st]);
        umma_commit(&k_empty[st]);
      };
      mbar_wait(q_full, 0);
      mbar_wait(&k_full[0], 0);
      tc_fence_after();
      issue_qk(0);
      for (int j = 0; j < n_kv; ++j) {
        if (j + 1 < n_kv) {
          const int t = j + 1, st = t & 1;
          mbar_wait(&k_full[st], (t >> 1) & 1);
          if (t >= 2) mbar_wait(&s_free[st], ((t >> 1) - 1) & 1);
          tc_fence_after();
          issue_qk(t);
        }
        const int st = j & 1;
        mbar_wait(&v_full[st], (j >> 1) & 1);
        mbar_wait(p_full, j & 1);
        if (j >= 2) mbar_wait(&o_free[st], ((j >> 1) - 1) & 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + st * kTileBytes);
        const uint32_t d = tmem_base + 2 * kTile + st * kHeadDim;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          const uint32_t a_off = (kk >> 2) * kHalfBytes + (kk & 3) * 32;   // P: K-major, 16 keys per step
          umma_f16_ss(d, umma_smem_desc_sw128(p_addr + a_off, 16, 1024),
                      umma_smem_desc_sw128(v_addr + kk * 2048, kHalfBytes, 1024),  // V: MN-major
                      idesc_pv, kk != 0);
        }
        umma_commit(&o_full[st]);
        umma_commit(&v_empty[st]);
        umma_commit(p_free);
      }
    }
  }
  } else {
    setmaxnreg_inc<224>();
    // ----------------------------------------------------------- softmax warps
    const int quarter = warp & 3;          // TMEM lane quarter this warp may access
    const int half = (warp - 4) >> 2;      // which 64 columns of S / O this thread owns
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = uint32_t(quarter * 32) << 16;
    const int q_pos = q0 + row + causal_shift;
    float m = -CUDART_INF_F, l = 0.f, alpha_prev = 1.f;
    float o[64];
#pragma unroll
    for (int i = 0; i < 64; ++i) o[i] = 0.f;
    const uint32_t p_row = smem_u32(sP + half * kHalfBytes + row * 128);
    const uint32_t sw = uint32_t(row & 7);

    for (int j = 0; j < n_kv; ++j) {
      const int sb = j & 1;
      mbar_wait(&s_full[sb], (j >> 1) & 1);
      tc_fence_after();
      const uint32_t s_addr = tmem_base + lane_addr + sb * kTile + half * 64;
      const int key0 = j * kTile + half * 64;
      const bool need_mask = (p.causal && (key0 + 63 > q_pos)) || (key0 + 64 > p.Sk);
      // one TMEM read of this thread's 64 score columns; everything else happens in registers
      uint32_t r0[32], r1[32];
      tmem_ld_32x32b_x32(s_addr, r0);
      tmem_ld_32x32b_x32(s_addr + 32, r1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&s_free[sb]);          // S buffer may be overwritten by QK^T of tile j+2
      float sc[64];   // raw scores; the softmax scale is folded into the exp2 FFMA below
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        sc[i] = __uint_as_float(r0[i]);
        sc[32 + i] = __uint_as_float(r1[i]);
      }
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < 64; ++i) {
          const int key = key0 + i;
          const bool ok = (key < p.Sk) && (!p.causal || key <= q_pos);
          sc[i] = ok ? sc[i] : -CUDART_INF_F;
        }
      }
      float mx = sc[0];
#pragma unroll
      for (int i = 1; i < 64; ++i) mx = fmaxf(mx, sc[i]);
      mx *= p.scale_log2;   // scale > 0: max commutes with the scaling
      float* xm = xmax + sb * 256;
      xm[half * 128 + row] = mx;
      named_bar_sync(1, 256);
      const float m_new = fmaxf(m, fmaxf(mx, xm[(half ^ 1) * 128 + row]));
      const float m_safe = (m_new == -CUDART_INF_F) ? 0.f : m_new;
      const float alpha = exp2f(m - m_safe);
      float ls0 = 0.f, ls1 = 0.f, ls2 = 0.f, ls3 = 0.f;
      const float neg_m = -m_safe;
#pragma unroll
      for (int i = 0; i < 64; i += 4) {   // masked entries: exp2(-inf) = 0
        sc[i + 0] = exp2f(fmaf(sc[i + 0], p.scale_log2, neg_m)); ls0 += sc[i + 0];
        sc[i + 1] = exp2f(fmaf(sc[i + 1], p.scale_log2, neg_m)); ls1 += sc[i + 1];
        sc[i + 2] = exp2f(fmaf(sc[i + 2], p.scale_log2, neg_m)); ls2 += sc[i + 2];
        sc[i + 3] = exp2f(fmaf(sc[i + 3], p.scale_log2, neg_m)); ls3 += sc[i + 3];
      }
      const float lsum = (ls0 + ls1) + (ls2 + ls3);
      if (j > 0) mbar_wait(p_free, (j - 1) & 1);  // PV of the previous tile has consumed sP
#pragma unroll
      for (int t = 0; t < 8; ++t)
        st_shared_v4(p_row + ((uint32_t(t) ^ sw) << 4), pack_bf16x2(sc[8 * t + 0], sc[8 * t + 1]),
                     pack_bf16x2(sc[8 * t + 2], sc[8 * t + 3]), pack_bf16x2(sc[8 * t + 4], sc[8 * t + 5]),
                     pack_bf16x2(sc[8 * t + 6], sc[8 * t + 7]));
      fence_proxy_async_smem();
      mbar_arrive(p_full);
      l = l * alpha + lsum;
      // fold in the PV result of the previous tile while the tensor core works on this one
      if (j > 0) {
        const int ob = (j - 1) & 1;
        mbar_wait(&o_full[ob], ((j - 1) >> 1) & 1);
        tc_fence_after();
        const uint32_t o_addr = tmem_base + lane_addr + 2 * kTile + ob * kHeadDim + half * 64;
        tmem_ld_32x32b_x32(o_addr, r0);
        tmem_ld_32x32b_x32(o_addr + 32, r1);
        tmem_ld_wait();
        tc_fence_before();
        mbar_arrive(&o_free[ob]);
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          o[i] = o[i] * alpha_prev + __uint_as_float(r0[i]);
          o[32 + i] = o[32 + i] * alpha_prev + __uint_as_float(r1[i]);
        }
      }
      alpha_prev = alpha;
      m = m_new;
    }
    {  // last tile's PV result
      const int ob = (n_kv - 1) & 1;
      mbar_wait(&o_full[ob], ((n_kv - 1) >> 1) & 1);
      tc_fence_after();
      const uint32_t o_addr = tmem_base + lane_addr + 2 * kTile + ob * kHeadDim + half * 64;
      uint32_t r[32];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        tmem_ld_32x32b_x32(o_addr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o[c * 32 + i] = o[c * 32 + i] * alpha_prev + __uint_as_float(r[i]);
      }
    }
    xsum[half * 128 + row] = l;
    named_bar_sync(1, 256);
    const float ltot = l + xsum[(half ^ 1) * 128 + row];
    float inv = (ltot > 0.f) ? 1.f / ltot : 0.f;
    const int q = q0 + row;
    if (q < p.Sq) {
      const size_t row_off = ((size_t(b) * p.S_tot + p.q_lo + q) * p.Hq + hq) * kHeadDim + half * 64;
      const size_t lse_idx = (size_t(b) * p.Hq + hq) * p.S_tot + p.q_lo + q;
      __nv_bfloat16* dst = p.out + row_off;
      float lse_out = (ltot > 0.f) ? (m + log2f(ltot)) * 0.6931471805599453f : -CUDART_INF_F;
      if (p.merge) {
        // ring attention: (prev, lse_prev) cover the key blocks seen so far.  out = w_prev * prev + w_cur * o / ltot with
        // w = exp(lse - logaddexp(lse_prev, lse_cur)); the merge costs one extra read of the row and no extra launch
        // (reference: the eager update_out_and_lse of context_parallel.py:367-424)
        const float lse_prev = p.lse[lse_idx];
        const float mx = fmaxf(lse_prev, lse_out);
        float w_prev = 0.f, w_cur = 0.f;
        if (mx > -CUDART_INF_F) {
          const float e_prev = __expf(lse_prev - mx), e_cur = __expf(lse_out - mx);
          const float tot = e_prev + e_cur;
          w_prev = e_prev / tot;
          w_cur = e_cur / tot;
          lse_out = mx + __logf(tot);
        }
        inv *= w_cur;
        if (p.acc != nullptr) {
          const float4* prev = reinterpret_cast<const float4*>(p.acc + row_off);
#pragma unroll
          for (int t = 0; t < 16; ++t) {
            const float4 v = prev[t];
            o[4 * t + 0] = o[4 * t + 0] * inv + w_prev * v.x; o[4 * t + 1] = o[4 * t + 1] * inv + w_prev * v.y;
            o[4 * t + 2] = o[4 * t + 2] * inv + w_prev * v.z; o[4 * t + 3] = o[4 * t + 3] * inv + w_prev * v.w;
          }
        } else {
          const uint4* prev = reinterpret_cast<const uint4*>(dst);
#pragma unroll
          for (int t = 0; t < 8; ++t) {
            const uint4 v = prev[t];
            const float2 a = unpack_bf16x2(v.x), c = unpack_bf16x2(v.y), d = unpack_bf16x2(v.z), e = unpack_bf16x2(v.w);
            o[8 * t + 0] = o[8 * t + 0] * inv + w_prev * a.x; o[8 * t + 1] = o[8 * t + 1] * inv + w_prev * a.y;
            o[8 * t + 2] = o[8 * t + 2] * inv + w_prev * c.x; o[8 * t + 3] = o[8 * t + 3] * inv + w_prev * c.y;
            o[8 * t + 4] = o[8 * t + 4] * inv + w_prev * d.x; o[8 * t + 5] = o[8 * t + 5] * inv + w_prev * d.y;
            o[8 * t + 6] = o[8 * t + 6] * inv + w_prev * e.x; o[8 * t + 7] = o[8 * t + 7] * inv + w_prev * e.y;
          }
        }
        inv = 1.f;
      }
      if (p.acc != nullptr) {
        float4* a4 = reinterpret_cast<float4*>(p.acc + row_off);
#pragma unroll
        for (int t = 0; t < 16; ++t)
          a4[t] = make_float4(o[4 * t + 0] * inv, o[4 * t + 1] * inv, o[4 * t + 2] * inv, o[4 * t + 3] * inv);
      }
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        *reinterpret_cast<uint4*>(dst + 8 * t) =
            make_uint4(pack_bf16x2(o[8 * t + 0] * inv, o[8 * t + 1] * inv), pack_bf16x2(o[8 * t + 2] * inv, o[8 * t + 3] * inv),
                       pack_bf16x2(o[8 * t + 4] * inv, o[8 * t + 5] * inv), pack_bf16x2(o[8 * t + 6] * inv, o[8 * t + 7] * inv));
      }
      // both halves of a row read lse_prev above: the row's partner must have read it before it is overwritten
      if (p.merge) named_bar_sync(1, 256);
      if (half == 0) p.lse[lse_idx] = lse_out;
    } else if (p.merge) {
      named_bar_sync(1, 256);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

}  // namespace stb

using namespace stb;

// q: [B,Sq,Hq,128], k/v: [B,Sk,Hkv,128] bf16 with element strides (batch, token, head); out: contiguous
// [B,Sq,Hq,128] bf16; lse: [B,Hq,Sq] fp32.  causal: query i attends keys <= i + (Sk - Sq).
static int flash_attn_fwd_impl(const void* q, const void* k, const void* v, void* out, float* lse, float* acc, int merge,
                               int q_lo, int S_tot, int B, int Sq, int Sk, int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh,
                               long k_sb, long k_ss, long k_sh, long v_sb, long v_ss, long v_sh, float scale,
                               int causal, cudaStream_t stream) {
  STB_REQUIRE(D == kHeadDim, "flash_attn: head_dim must be 128 (got %d)", D);
  STB_REQUIRE(Hq % Hkv == 0, "flash_attn: Hq=%d must be a multiple of Hkv=%d", Hq, Hkv);
  STB_REQUIRE(!causal || Sq <= Sk, "flash_attn: causal attention needs Sq <= Sk");
  if (B == 0 || Sq == 0) return 0;
  CUtensorMap tmQ, tmK, tmV;
  if (int rc = make_tmap_bshd(&tmQ, q, B, Sq, Hq, D, q_sb, q_ss, q_sh, 64, kTile)) return rc;
  if (int rc = make_tmap_bshd(&tmK, k, B, Sk, Hkv, D, k_sb, k_ss, k_sh, 64, kTile)) return rc;
  if (int rc = make_tmap_bshd(&tmV, v, B, Sk, Hkv, D, v_sb, v_ss, v_sh, 64, kTile)) return rc;
  AttnFwdParams p;
  p.out = (__nv_bfloat16*)out;
  p.lse = lse;
  p.acc = acc;
  p.merge = merge;
  p.q_lo = q_lo;
  p.S_tot = S_tot;
  p.B = B; p.Sq = Sq; p.Sk = Sk; p.Hq = Hq; p.Hkv = Hkv;
  p.n_qtiles = (Sq + kTile - 1) / kTile;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.causal = causal;
  static bool attr_set = false;
  if (!attr_set) {
    STB_CHECK_CUDA(cudaFuncSetAttribute(flash_attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        AttnSmem::kBytes));
    attr_set = true;
  }
  const int grid = B * Hq * p.n_qtiles;
  flash_attn_fwd_kernel<<<grid, kAttnThreads, AttnSmem::kBytes, stream>>>(tmQ, tmK, tmV, p);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_flash_attn_fwd(const void* q, const void* k, const void* v, void* out, float* lse, int B,
                                  int Sq, int Sk, int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh,
                                  long k_sb, long k_ss, long k_sh, long v_sb, long v_ss, long v_sh, float scale,
                                  int causal, cudaStream_t stream) {
  return flash_attn_fwd_impl(q, k, v, out, lse, nullptr, 0, 0, Sq, B, Sq, Sk, Hq, Hkv, D, q_sb, q_ss, q_sh, k_sb, k_ss, k_sh,
                             v_sb, v_ss, v_sh, scale, causal, stream);
}

// One key block of a blockwise (ring) attention over query rows [q_lo, q_lo + Sq) of out / acc / lse (S_tot rows per
// batch; q points at row q_lo).  merge = 0: first block, (out, lse) are written; merge = 1: (out, lse)
// -- and the fp32 running output `acc` when given -- hold the result over the blocks seen so far and this block is
// folded into them in the kernel's epilogue.  `acc` (fp32 [B,Sq,Hq,128], optional) keeps the running output in full
// precision between blocks; `out` always receives the bf16 value.
extern "C" int stb_flash_attn_fwd_block(const void* q, const void* k, const void* v, void* out, float* lse, float* acc,
                                        int merge, int q_lo, int S_tot, int B, int Sq, int Sk, int Hq, int Hkv, int D,
                                        long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh, long v_sb,
                                        long v_ss, long v_sh, float scale, int causal, cudaStream_t stream) {
  STB_REQUIRE(q_lo >= 0 && q_lo + Sq <= S_tot, "flash_attn_fwd_block: query window [%d, %d) outside %d rows", q_lo,
              q_lo + Sq, S_tot);
  return flash_attn_fwd_impl(q, k, v, out, lse, acc, merge, q_lo, S_tot, B, Sq, Sk, Hq, Hkv, D, q_sb, q_ss, q_sh, k_sb, k_ss, k_sh,
                             v_sb, v_ss, v_sh, scale, causal, stream);
}
