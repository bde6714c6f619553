// Memory-bound fused kernels around the attention/MLP GEMMs (all bf16 I/O, fp32 math):
//   * residual-add + RMSNorm forward / backward           (reference: models/attention_utils.py:247-271,
//                                                            models/model_qwen3.py:260-269)
//   * per-head QK-RMSNorm + rotate-half RoPE fwd / bwd      (models/model_qwen3.py:205-217,
//                                                            models/attention_utils.py:170-192)
//   * SwiGLU forward / backward on a fused gate|up buffer   (models/model_qwen3.py:85-87)
// Each kernel touches every byte once: 16-byte vector loads/stores, one row per
// CTA (norm) or per warp (head), no shared-memory staging except block reductions.
#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

STB_DEVICE float half_warp_sum(float v) {
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
STB_DEVICE void unpack8(const uint4& r, float (&f)[8]) {
  const float2 a = unpack_bf16x2(r.x), b = unpack_bf16x2(r.y), c = unpack_bf16x2(r.z), d = unpack_bf16x2(r.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
STB_DEVICE uint4 pack8(const float (&f)[8]) {
  return make_uint4(pack_bf16x2(f[0], f[1]), pack_bf16x2(f[2], f[3]), pack_bf16x2(f[4], f[5]), pack_bf16x2(f[6], f[7]));
}
STB_DEVICE void load8f(const float* p, float (&f)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}



// ----------------------------------------------------------------- RMSNorm
template <int THREADS>
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int i = 0; i < THREADS / 32; ++i) t += red[i];
  __syncthreads();
  return t;
}

// h = x (+ res); y = w * bf16(h * rstd).   Optionally writes h back (h_out) and rstd.
template <int VPT, int kNormThreads>
__global__ void __launch_bounds__(kNormThreads)
    rmsnorm_fwd_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                       const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ h_out,
                       __nv_bfloat16* __restrict__ y, float* __restrict__ rstd_out, int rows, int H,
                       float eps) {
  __shared__ float red[kNormThreads / 32];
  const int nvec = H >> 3;
  uint4 wv[VPT];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) wv[i] = reinterpret_cast<const uint4*>(w)[v];
  }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = size_t(row) * H;
    float vals[VPT][8];
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        const uint4 a = ld_nc_v4(x + base + v * 8);
        const uint32_t au[4] = {a.x, a.y, a.z, a.w};
        uint32_t ru[4] = {0, 0, 0, 0};
        if (res != nullptr) {
          const uint4 b = ld_nc_v4(res + base + v * 8);
          ru[0] = b.x; ru[1] = b.y; ru[2] = b.z; ru[3] = b.w;
        }
        uint32_t hu[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float2 f = unpack_bf16x2(au[j]);
          if (res != nullptr) {
            const float2 g = unpack_bf16x2(ru[j]);
            f.x += g.x; f.y += g.y;
            // the residual stream is stored in bf16: round before normalising so that
            // forward and backward see the same h
            hu[j] = pack_bf16x2(f.x, f.y);
            f = unpack_bf16x2(hu[j]);
          }
          vals[i][2 * j] = f.x; vals[i][2 * j + 1] = f.y;
          ss += f.x * f.x + f.y * f.y;
        }
        if (res != nullptr && h_out != nullptr)
          *reinterpret_cast<uint4*>(h_out + base + v * 8) = make_uint4(hu[0], hu[1], hu[2], hu[3]);
      }
    }
    ss = block_sum<kNormThreads>(ss, red);
    const float rstd = rsqrtf(ss / float(H) + eps);
    if (threadIdx.x == 0 && rstd_out != nullptr) rstd_out[row] = rstd;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        const uint32_t wu[4] = {wv[i].x, wv[i].y, wv[i].z, wv[i].w};
        uint32_t ou[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 wf = unpack_bf16x2(wu[j]);
          const float2 xn = unpack_bf16x2(pack_bf16x2(vals[i][2 * j] * rstd, vals[i][2 * j + 1] * rstd));
          ou[j] = pack_bf16x2(xn.x * wf.x, xn.y * wf.y);
        }
        st_na_v4(y + base + v * 8, make_uint4(ou[0], ou[1], ou[2], ou[3]));
      }
    }
  }
}

// dx = rstd * (dy*w - xn * mean(dy*w*xn)) (+ dres);   dw_partial[cta] += dy * xn
template <int VPT, int kNormThreads>
__global__ void __launch_bounds__(kNormThreads)
    rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ h,
                       const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd_in,
                       const __nv_bfloat16* __restrict__ dres, __nv_bfloat16* __restrict__ dx,
                       float* __restrict__ dw_partial, int rows, int H) {
  __shared__ float red[kNormThreads / 32];
  const int nvec = H >> 3;
  float wf[VPT][8], dwacc[VPT][8];
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
#pragma unroll
    for (int j = 0; j < 8; ++j) { dwacc[i][j] = 0.f; wf[i][j] = 0.f; }
    if (v < nvec) {
      const uint4 a = reinterpret_cast<const uint4*>(w)[v];
      const uint32_t au[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(au[j]);
        wf[i][2 * j] = f.x; wf[i][2 * j + 1] = f.y;
      }
    }
  }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = size_t(row) * H;
    const float rstd = rstd_in[row];
    float xn[VPT][8], g[VPT][8];
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        const uint4 a = ld_nc_v4(h + base + v * 8);
        const uint4 b = ld_nc_v4(dy + base + v * 8);
        const uint32_t au[4] = {a.x, a.y, a.z, a.w}, bu[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 hf = unpack_bf16x2(au[j]);
          const float2 df = unpack_bf16x2(bu[j]);
          xn[i][2 * j] = hf.x * rstd; xn[i][2 * j + 1] = hf.y * rstd;
          dwacc[i][2 * j] += df.x * xn[i][2 * j];
          dwacc[i][2 * j + 1] += df.y * xn[i][2 * j + 1];
          g[i][2 * j] = df.x * wf[i][2 * j]; g[i][2 * j + 1] = df.y * wf[i][2 * j + 1];
          dot += g[i][2 * j] * xn[i][2 * j] + g[i][2 * j + 1] * xn[i][2 * j + 1];
        }
      }
    }
    dot = block_sum<kNormThreads>(dot, red) / float(H);
#pragma unroll
    for (int i = 0; i < VPT; ++i) {
      const int v = threadIdx.x + i * kNormThreads;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (g[i][j] - xn[i][j] * dot);
        if (dres != nullptr) {
          const uint4 c = ld_nc_v4(dres + base + v * 8);
          const uint32_t cu[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 f = unpack_bf16x2(cu[j]);
            o[2 * j] += f.x; o[2 * j + 1] += f.y;
          }
        }
        st_na_v4(dx + base + v * 8,
                 make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]),
                            pack_bf16x2(o[6], o[7])));
      }
    }
  }
  // one fp32 partial row per CTA (scratch), reduced by rmsnorm_dw_reduce_kernel
  float* out = dw_partial + size_t(blockIdx.x) * H;
#pragma unroll
  for (int i = 0; i < VPT; ++i) {
    const int v = threadIdx.x + i * kNormThreads;
    if (v < nvec) {
      *reinterpret_cast<float4*>(out + v * 8) = make_float4(dwacc[i][0], dwacc[i][1], dwacc[i][2], dwacc[i][3]);
      *reinterpret_cast<float4*>(out + v * 8 + 4) = make_float4(dwacc[i][4], dwacc[i][5], dwacc[i][6], dwacc[i][7]);
    }
  }
}

// dw[c] (+)= sum_p partial[p][c].  Block = 32 columns x 32 row lanes: every warp reads 128 contiguous bytes of one
// partial row, each thread sums nparts/32 rows with independent loads, then a shared-memory tree over the row lanes
// (the first version walked all partial rows serially per column: 22 us per call, 4 % of a Qwen3-0.6B step).
__global__ void __launch_bounds__(1024)
rmsnorm_dw_reduce_kernel(const float* __restrict__ partial, float* __restrict__ dw, int nparts, int H, int accumulate) {
  __shared__ float red[32][33];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + cx;
  float s = 0.f;
  if (c < H) {
#pragma unroll 4
    for (int p = ry; p < nparts; p += 32) s += partial[size_t(p) * H + c];
  }
  red[ry][cx] = s;
  __syncthreads();
  if (ry == 0 && c < H) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 32; ++r) t += red[r][cx];
    dw[c] = accumulate ? dw[c] + t : t;
  }
}

// ------------------------------------------------------- QK-norm + RoPE
// head_dim == 128.  A head is handled by HALF a warp: 16 lanes x 8 consecutive bf16 (one 16-byte load / store per lane),
// so a warp works on two heads of a token at a time and four head pairs are in flight per warp.  rotate-half pairs element
// j with j + 64, i.e. lane l with lane l ^ 8 of the same half-warp; the norm reductions stay inside the 16 lanes.
// in : qkv [T, (Hq + 2*Hkv) * 128]   (q heads, then k heads, then v heads)
// out: qk  [T, (Hq + Hkv) * 128]     rotated q and k (v is consumed in place by attention)
template <bool NORM>
__global__ void __launch_bounds__(128)
    qknorm_rope_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, const __nv_bfloat16* __restrict__ wq,
                           const __nv_bfloat16* __restrict__ wk, const float* __restrict__ cos_t,
                           const float* __restrict__ sin_t, const int* __restrict__ pos,
                           __nv_bfloat16* __restrict__ out, float* __restrict__ rstd_out, int T, int S,
                           int Hq, int Hkv, int pos_offset, float eps) {
  const int lane = threadIdx.x & 31, l16 = lane & 15, grp = lane >> 4;
  const int nheads = Hq + Hkv;
  const int in_ld = (Hq + 2 * Hkv) * 128, out_ld = nheads * 128;
  float wqf[8], wkf[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) wqf[j] = wkf[j] = 1.f;
  if (NORM) {
    unpack8(*reinterpret_cast<const uint4*>(wq + l16 * 8), wqf);
    unpack8(*reinterpret_cast<const uint4*>(wk + l16 * 8), wkf);
  }
  const float sign = (l16 < 8) ? -1.f : 1.f;      // rotate-half: first half takes -x[j + 64], second half +x[j - 64]
  for (int t = blockIdx.x * 4 + (threadIdx.x >> 5); t < T; t += gridDim.x * 4) {
    const int p = pos ? pos[t] : (t % S) + pos_offset;
    float cf[8], sf[8];
    load8f(cos_t + size_t(p) * 128 + l16 * 8, cf);
    load8f(sin_t + size_t(p) * 128 + l16 * 8, sf);
    const __nv_bfloat16* src = qkv + size_t(t) * in_ld + l16 * 8;
    __nv_bfloat16* dst = out + size_t(t) * out_ld + l16 * 8;
    constexpr int kB = 6;                          // head pairs whose loads are issued before any of them is consumed
    for (int hb = 0; hb < nheads; hb += 2 * kB) {
      uint4 raw[kB];
#pragma unroll
      for (int i = 0; i < kB; ++i) {
        const int hd = hb + 2 * i + grp;
        raw[i] = hd < nheads ? *reinterpret_cast<const uint4*>(src + hd * 128) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int i = 0; i < kB; ++i) {
        const int hd = hb + 2 * i + grp;
        const bool live = hd < nheads;
        if (hb + 2 * i >= nheads) break;           // warp-uniform
        float v[8];
        unpack8(raw[i], v);
        if (NORM) {
          float ss = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) ss += v[j] * v[j];
          ss = half_warp_sum(ss);
          const float rstd = rsqrtf(ss * (1.f / 128.f) + eps);
          if (l16 == 0 && live && rstd_out) rstd_out[long(t) * nheads + hd] = rstd;
          const bool is_q = hd < Hq;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float xn = __bfloat162float(__float2bfloat16(v[j] * rstd));
            v[j] = __bfloat162float(__float2bfloat16(xn * (is_q ? wqf[j] : wkf[j])));
          }
        }
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float other = __shfl_xor_sync(0xffffffffu, v[j], 8);
          o[j] = v[j] * cf[j] + sign * other * sf[j];
        }
        if (live) *reinterpret_cast<uint4*>(dst + hd * 128) = pack8(o);
      }
    }
  }
}

// dqk: [T, (Hq+Hkv)*128] gradients w.r.t. rotated q,k ; writes d(qkv)[:, :(Hq+Hkv)*128] into dqkv
// (row stride (Hq+2Hkv)*128) and accumulates dwq/dwk[128] (fp32) with atomics, one per CTA.
template <bool NORM>
__global__ void __launch_bounds__(128)
    qknorm_rope_bwd_kernel(const __nv_bfloat16* __restrict__ dqk, const float* __restrict__ dq32,
                           const __nv_bfloat16* __restrict__ qkv,
                           const __nv_bfloat16* __restrict__ wq, const __nv_bfloat16* __restrict__ wk,
                           const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                           const int* __restrict__ pos, const float* __restrict__ rstd_in,
                           __nv_bfloat16* __restrict__ dqkv, float* __restrict__ dwq,
                           float* __restrict__ dwk, int T, int S, int Hq, int Hkv, int pos_offset) {
  __shared__ float sdw[2][128];
  const int lane = threadIdx.x & 31, l16 = lane & 15, grp = lane >> 4;
  if (NORM) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) (&sdw[0][0])[i] = 0.f;
    __syncthreads();
  }
  const int nheads = Hq + Hkv;
  const int in_ld = (Hq + 2 * Hkv) * 128, g_ld = nheads * 128;
  float dwl[2][8];
  float wqf[8], wkf[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    dwl[0][j] = dwl[1][j] = 0.f;
    wqf[j] = wkf[j] = 1.f;
  }
  if (NORM) {
    unpack8(*reinterpret_cast<const uint4*>(wq + l16 * 8), wqf);
    unpack8(*reinterpret_cast<const uint4*>(wk + l16 * 8), wkf);
  }
  // transpose of rotate-half: dv[j] = g[j] cos[j] + (j < 64 ? +1 : -1) * (g sin)[j ^ 64]
  const float sign = (l16 < 8) ? 1.f : -1.f;
  for (int t = blockIdx.x * 4 + (threadIdx.x >> 5); t < T; t += gridDim.x * 4) {
    const int p = pos ? pos[t] : (t % S) + pos_offset;
    float cf[8], sf[8];
    load8f(cos_t + size_t(p) * 128 + l16 * 8, cf);
    load8f(sin_t + size_t(p) * 128 + l16 * 8, sf);
    constexpr int kB = 4;                          // head pairs whose loads are issued before any of them is consumed
    for (int hb = 0; hb < nheads; hb += 2 * kB) {
      uint4 graw[kB], xraw[kB];
      float4 gq0[kB], gq1[kB];                     // q-head gradients when they arrive as fp32 (fused attention backward)
      float rs[kB];
#pragma unroll
      for (int i = 0; i < kB; ++i) {
        const int hd = hb + 2 * i + grp;
        const bool live = hd < nheads;
        if (dq32 != nullptr && live && hd < Hq) {
          const float* gp = dq32 + (size_t(t) * Hq + hd) * 128 + l16 * 8;
          gq0[i] = *reinterpret_cast<const float4*>(gp);
          gq1[i] = *reinterpret_cast<const float4*>(gp + 4);
          graw[i] = make_uint4(0, 0, 0, 0);
        } else {
          graw[i] = live ? *reinterpret_cast<const uint4*>(dqk + size_t(t) * g_ld + hd * 128 + l16 * 8) : make_uint4(0, 0, 0, 0);
        }
        if (NORM) {
          xraw[i] = live ? *reinterpret_cast<const uint4*>(qkv + size_t(t) * in_ld + hd * 128 + l16 * 8) : make_uint4(0, 0, 0, 0);
          rs[i] = live ? rstd_in[long(t) * nheads + hd] : 0.f;
        }
      }
#pragma unroll
      for (int i = 0; i < kB; ++i) {
        const int hd = hb + 2 * i + grp;
        const bool live = hd < nheads;
        if (hb + 2 * i >= nheads) break;           // warp-uniform
        float go[8], dv[8];
        if (dq32 != nullptr && live && hd < Hq) {
          go[0] = gq0[i].x; go[1] = gq0[i].y; go[2] = gq0[i].z; go[3] = gq0[i].w;
          go[4] = gq1[i].x; go[5] = gq1[i].y; go[6] = gq1[i].z; go[7] = gq1[i].w;
        } else {
          unpack8(graw[i], go);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float other = __shfl_xor_sync(0xffffffffu, go[j] * sf[j], 8);
          dv[j] = go[j] * cf[j] + sign * other;
        }
        float o[8];
        if (NORM) {
          float x[8];
          unpack8(xraw[i], x);
          const float rstd = rs[i];
          const bool is_q = hd < Hq;
          float xn[8], g[8], dot = 0.f;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            xn[j] = x[j] * rstd;
            g[j] = dv[j] * (is_q ? wqf[j] : wkf[j]);
            dot += g[j] * xn[j];
            const float contrib = dv[j] * xn[j];
            dwl[0][j] += is_q ? contrib : 0.f;
            dwl[1][j] += is_q ? 0.f : contrib;
          }
          dot = half_warp_sum(dot) * (1.f / 128.f);
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = rstd * (g[j] - xn[j] * dot);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) o[j] = dv[j];
        }
        if (live) *reinterpret_cast<uint4*>(dqkv + size_t(t) * in_ld + hd * 128 + l16 * 8) = pack8(o);
      }
    }
  }
  if (NORM) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {   // the two half-warps hold the same element positions: fold them first
      dwl[0][j] += __shfl_xor_sync(0xffffffffu, dwl[0][j], 16);
      dwl[1][j] += __shfl_xor_sync(0xffffffffu, dwl[1][j], 16);
    }
    if (grp == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        atomicAdd(&sdw[0][l16 * 8 + j], dwl[0][j]);
        atomicAdd(&sdw[1][l16 * 8 + j], dwl[1][j]);
      }
    }
    __syncthreads();
    atomicAdd(dwq + threadIdx.x, sdw[0][threadIdx.x]);
    atomicAdd(dwk + threadIdx.x, sdw[1][threadIdx.x]);
  }
}

// ------------------------------------------------------------------ SwiGLU
// gu: [T, 2*I] with gate in columns [0,I) and up in [I,2I);  act[T, I] = silu(gate) * up
__global__ void __launch_bounds__(256)
    swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act, long T,
                      int I) {
  const int vec_per_row = I >> 3;
  const long total = T * vec_per_row;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const long t = i / vec_per_row;
    const int c = int(i % vec_per_row) * 8;
    const uint4 g = ld_nc_v4(gu + t * 2 * I + c);
    const uint4 u = ld_nc_v4(gu + t * 2 * I + I + c);
    const uint32_t gg[4] = {g.x, g.y, g.z, g.w}, uu[4] = {u.x, u.y, u.z, u.w};
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 a = unpack_bf16x2(gg[j]), b = unpack_bf16x2(uu[j]);
      const float s0 = a.x / (1.f + __expf(-a.x)), s1 = a.y / (1.f + __expf(-a.y));
      o[j] = pack_bf16x2(s0 * b.x, s1 * b.y);
    }
    st_na_v4(act + t * I + c, make_uint4(o[0], o[1], o[2], o[3]));
  }
}

// dgu[:, :I] = dact * up * dsilu(gate);  dgu[:, I:] = dact * silu(gate)
__global__ void __launch_bounds__(256)
    swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* __restrict__ gu,
                      __nv_bfloat16* __restrict__ dgu, long T, int I) {
  const int vec_per_row = I >> 3;
  const long total = T * vec_per_row;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const long t = i / vec_per_row;
    const int c = int(i % vec_per_row) * 8;
    const uint4 g = ld_nc_v4(gu + t * 2 * I + c);
    const uint4 u = ld_nc_v4(gu + t * 2 * I + I + c);
    const uint4 d = ld_nc_v4(dact + t * I + c);
    const uint32_t gg[4] = {g.x, g.y, g.z, g.w}, uu[4] = {u.x, u.y, u.z, u.w}, dd[4] = {d.x, d.y, d.z, d.w};
    uint32_t og[4], ou[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 a = unpack_bf16x2(gg[j]), b = unpack_bf16x2(uu[j]), e = unpack_bf16x2(dd[j]);
      const float sg0 = 1.f / (1.f + __expf(-a.x)), sg1 = 1.f / (1.f + __expf(-a.y));
      const float si0 = a.x * sg0, si1 = a.y * sg1;
      const float ds0 = sg0 * (1.f + a.x * (1.f - sg0)), ds1 = sg1 * (1.f + a.y * (1.f - sg1));
      og[j] = pack_bf16x2(e.x * b.x * ds0, e.y * b.y * ds1);
      ou[j] = pack_bf16x2(e.x * si0, e.y * si1);
    }
    st_na_v4(dgu + t * 2 * I + c, make_uint4(og[0], og[1], og[2], og[3]));
    st_na_v4(dgu + t * 2 * I + I + c, make_uint4(ou[0], ou[1], ou[2], ou[3]));
  }
}

static int norm_grid(int rows) {
  const int cap = sm_count() * 8;
  return rows < cap ? rows : cap;
}

}  // namespace stb

using namespace stb;

#define STB_DISPATCH_VPT(H, CALL)                                                        \
  do {                                                                                   \
    const int _nvec = (H) / 8;                                                           \
    if (_nvec <= 128) { constexpr int VPT = 1; constexpr int NT = 128; CALL; }           \
    else if (_nvec <= 256) { constexpr int VPT = 2; constexpr int NT = 128; CALL; }      \
    else if (_nvec <= 512) { constexpr int VPT = 2; constexpr int NT = 256; CALL; }      \
    else { constexpr int VPT = 4; constexpr int NT = 256; CALL; }                        \
  } while (0)

extern "C" int stb_rmsnorm_fwd(const void* x, const void* res, const void* w, void* h_out, void* y,
                               float* rstd, int rows, int H, float eps, cudaStream_t stream) {
  STB_REQUIRE(H % 8 == 0 && H <= 8192, "rmsnorm: H=%d must be a multiple of 8 and <= 8192", H);
  if (rows == 0) return 0;
  STB_DISPATCH_VPT(H, (rmsnorm_fwd_kernel<VPT, NT><<<norm_grid(rows), NT, 0, stream>>>(
                          (const __nv_bfloat16*)x, (const __nv_bfloat16*)res, (const __nv_bfloat16*)w,
                          (__nv_bfloat16*)h_out, (__nv_bfloat16*)y, rstd, rows, H, eps)));
  STB_CHECK_LAUNCH();
  return 0;
}

// Number of fp32 scratch elements stb_rmsnorm_bwd needs for (rows, H).
extern "C" long stb_rmsnorm_bwd_scratch_elems(int rows, int H) {
  const int cap = sm_count() * 4;
  const int grid = rows < cap ? rows : cap;
  return long(grid) * H;
}

extern "C" int stb_rmsnorm_bwd(const void* dy, const void* h, const void* w, const float* rstd,
                               const void* dres, void* dx, float* dw, int dw_accumulate,
                               float* scratch, int rows, int H, cudaStream_t stream) {
  STB_REQUIRE(H % 8 == 0 && H <= 8192, "rmsnorm: H=%d must be a multiple of 8 and <= 8192", H);
  if (rows == 0) return 0;
  const int cap = sm_count() * 4;
  const int grid = rows < cap ? rows : cap;
  STB_DISPATCH_VPT(H, (rmsnorm_bwd_kernel<VPT, NT><<<grid, NT, 0, stream>>>(
                          (const __nv_bfloat16*)dy, (const __nv_bfloat16*)h, (const __nv_bfloat16*)w,
                          rstd, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, rows, H)));
  STB_CHECK_LAUNCH();
  rmsnorm_dw_reduce_kernel<<<(H + 31) / 32, 1024, 0, stream>>>(scratch, dw, grid, H, dw_accumulate);
  STB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------
// RMSNorm backward for callers that must not allocate device memory.
//
// stb_rmsnorm_bwd() needs a caller-provided fp32 scratch buffer for the per-CTA partial sums of the
// weight gradient.  Some users of this C ABI have nowhere to get one: a host program that hands the
// whole of HBM to its own arena at start-up (so a later cudaMalloc fails or fragments the arena), or
// glue code running where an allocation -- and the device-wide synchronisation cudaMalloc/cudaFree
// imply -- is not acceptable, such as the inner loop of a latency-sensitive service.  For them this
// variant keeps the scratch inside the library image: a file-scope __device__ array of
// kNoAllocScratchElems floats (16 MiB).  Every GPU that loads the library carries that array until the
// process exits, whether or not this function is ever called.
//
// What to know before calling it:
//  * Limit: the partial sums take min(rows, 4 * SM count) * H floats.  If that exceeds
//    kNoAllocScratchElems (4 Mi floats) the call is refused with an error naming both numbers; use
//    stb_rmsnorm_bwd() with your own buffer for such shapes.  On a 148-SM B200 every H <= 7080 fits
//    for any row count.
//  * The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
//    that device current) and caches it.  Make that first call during start-up, before any stream
//    capture begins on the device.
//  * It calls no cudaMalloc / cudaFree and no library routine that allocates.
//  * All calls on one device share the one array, which still holds whatever the previous call left
//    there.  Nothing in it has to start at zero for this kernel -- every partial row that the reduction
//    reads is fully rewritten by the same call first -- but two calls whose kernels overlap on the GPU
//    (different streams, no ordering between them) overwrite each other's partial sums and both
//    produce a wrong dw.  Serialise them: same stream, or an event between the streams.
//  * Results are the same as stb_rmsnorm_bwd() with the same arguments (same grid, same summation
//    order).
constexpr long kNoAllocScratchElems = 4l << 20;
static __device__ __align__(256) float g_rmsnorm_noalloc_scratch[kNoAllocScratchElems];

// In host code the symbol above is only a handle, not a device address: the address each device gave
// the array is fetched once per device and kept here.
static float* g_noalloc_scratch_ptr[64] = {nullptr};

static int noalloc_scratch_for_current_device(float** out) {
  int dev = 0;
  STB_CHECK_CUDA(cudaGetDevice(&dev));
  STB_REQUIRE(dev >= 0 && dev < 64, "rmsnorm_bwd_noalloc: device index %d out of range", dev);
  if (g_noalloc_scratch_ptr[dev] == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_rmsnorm_noalloc_scratch);
    STB_REQUIRE(e == cudaSuccess && p != nullptr,
                "rmsnorm_bwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                cudaGetErrorString(e));
    g_noalloc_scratch_ptr[dev] = p;
  }
  *out = g_noalloc_scratch_ptr[dev];
  return 0;
}

extern "C" long stb_rmsnorm_bwd_noalloc_max_elems() { return kNoAllocScratchElems; }

extern "C" int stb_rmsnorm_bwd_noalloc(const void* dy, const void* h, const void* w, const float* rstd,
                                       const void* dres, void* dx, float* dw, int dw_accumulate,
                                       int rows, int H, cudaStream_t stream) {
  STB_REQUIRE(H % 8 == 0 && H <= 8192, "rmsnorm: H=%d must be a multiple of 8 and <= 8192", H);
  if (rows == 0) return 0;
  const int cap = sm_count() * 4;
  const int grid = rows < cap ? rows : cap;
  const long need = long(grid) * H;
  STB_REQUIRE(need <= kNoAllocScratchElems,
              "rmsnorm_bwd_noalloc: rows=%d, H=%d needs %ld scratch floats but the built-in scratch holds "
              "%ld; call stb_rmsnorm_bwd with a caller-provided buffer instead",
              rows, H, need, kNoAllocScratchElems);
  float* scratch = nullptr;
  if (int rc = noalloc_scratch_for_current_device(&scratch)) return rc;
  STB_DISPATCH_VPT(H, (rmsnorm_bwd_kernel<VPT, NT><<<grid, NT, 0, stream>>>(
                          (const __nv_bfloat16*)dy, (const __nv_bfloat16*)h, (const __nv_bfloat16*)w,
                          rstd, (const __nv_bfloat16*)dres, (__nv_bfloat16*)dx, scratch, rows, H)));
  STB_CHECK_LAUNCH();
  rmsnorm_dw_reduce_kernel<<<(H + 31) / 32, 1024, 0, stream>>>(scratch, dw, grid, H, dw_accumulate);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_qknorm_rope_fwd(const void* qkv, const void* wq, const void* wk, const float* cos_t,
                                   const float* sin_t, const int* pos, void* out, float* rstd, int T,
                                   int S, int Hq, int Hkv, int head_dim, int pos_offset, int use_norm,
                                   float eps, cudaStream_t stream) {
  STB_REQUIRE(head_dim == 128, "qknorm_rope: head_dim must be 128 (got %d)", head_dim);
  if (T == 0) return 0;
  long blocks = (long(T) + 3) / 4;
  const long cap = long(sm_count()) * 16;
  const int grid = int(blocks < cap ? blocks : cap);
  if (use_norm)
    qknorm_rope_fwd_kernel<true><<<grid, 128, 0, stream>>>(
        (const __nv_bfloat16*)qkv, (const __nv_bfloat16*)wq, (const __nv_bfloat16*)wk, cos_t, sin_t, pos,
        (__nv_bfloat16*)out, rstd, T, S, Hq, Hkv, pos_offset, eps);
  else
    qknorm_rope_fwd_kernel<false><<<grid, 128, 0, stream>>>(
        (const __nv_bfloat16*)qkv, nullptr, nullptr, cos_t, sin_t, pos, (__nv_bfloat16*)out, nullptr, T,
        S, Hq, Hkv, pos_offset, eps);
  STB_CHECK_LAUNCH();
  return 0;
}

static int qknorm_rope_bwd_impl(const void* dqk, const float* dq32, const void* qkv, const void* wq, const void* wk,
                                const float* cos_t, const float* sin_t, const int* pos,
                                const float* rstd, void* dqkv, float* dwq, float* dwk, int T, int S,
                                int Hq, int Hkv, int head_dim, int pos_offset, int use_norm,
                                cudaStream_t stream) {
  STB_REQUIRE(head_dim == 128, "qknorm_rope: head_dim must be 128 (got %d)", head_dim);
  if (T == 0) return 0;
  long blocks = (long(T) + 3) / 4;
  // every CTA ends with 256 global atomics on the same 256 weight-gradient words: few persistent CTAs (the atomics of
  // 1024 short-lived CTAs serialised in L2 and cost more than the kernel's memory traffic)
  const long cap = long(sm_count()) * 3;
  const int grid = int(blocks < cap ? blocks : cap);
  if (use_norm)
    qknorm_rope_bwd_kernel<true><<<grid, 128, 0, stream>>>(
        (const __nv_bfloat16*)dqk, dq32, (const __nv_bfloat16*)qkv, (const __nv_bfloat16*)wq,
        (const __nv_bfloat16*)wk, cos_t, sin_t, pos, rstd, (__nv_bfloat16*)dqkv, dwq, dwk, T, S, Hq, Hkv,
        pos_offset);
  else
    qknorm_rope_bwd_kernel<false><<<grid, 128, 0, stream>>>(
        (const __nv_bfloat16*)dqk, dq32, nullptr, nullptr, nullptr, cos_t, sin_t, pos, nullptr,
        (__nv_bfloat16*)dqkv, nullptr, nullptr, T, S, Hq, Hkv, pos_offset);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_qknorm_rope_bwd(const void* dqk, const void* qkv, const void* wq, const void* wk,
                                   const float* cos_t, const float* sin_t, const int* pos,
                                   const float* rstd, void* dqkv, float* dwq, float* dwk, int T, int S,
                                   int Hq, int Hkv, int head_dim, int pos_offset, int use_norm,
                                   cudaStream_t stream) {
  return qknorm_rope_bwd_impl(dqk, nullptr, qkv, wq, wk, cos_t, sin_t, pos, rstd, dqkv, dwq, dwk, T, S, Hq, Hkv, head_dim,
                              pos_offset, use_norm, stream);
}

// Same, with the gradient of the rotated QUERY heads given as a contiguous fp32 [T, Hq, 128] tensor (what the fused
// attention backward accumulates); `dqk` then only supplies the key heads (columns [Hq*128, (Hq+Hkv)*128) of its rows).
extern "C" int stb_qknorm_rope_bwd_f32q(const void* dqk, const float* dq32, const void* qkv, const void* wq,
                                        const void* wk, const float* cos_t, const float* sin_t, const int* pos,
                                        const float* rstd, void* dqkv, float* dwq, float* dwk, int T, int S, int Hq,
                                        int Hkv, int head_dim, int pos_offset, int use_norm, cudaStream_t stream) {
  return qknorm_rope_bwd_impl(dqk, dq32, qkv, wq, wk, cos_t, sin_t, pos, rstd, dqkv, dwq, dwk, T, S, Hq, Hkv, head_dim,
                              pos_offset, use_norm, stream);
}

extern "C" int stb_swiglu_fwd(const void* gu, void* act, long T, int I, cudaStream_t stream) {
  STB_REQUIRE(I % 8 == 0, "swiglu: I=%d must be a multiple of 8", I);
  if (T == 0) return 0;
  const long total = T * (I / 8);
  long blocks = (total + 255) / 256;
  const long cap = long(sm_count()) * 16;
  swiglu_fwd_kernel<<<int(blocks < cap ? blocks : cap), 256, 0, stream>>>(
      (const __nv_bfloat16*)gu, (__nv_bfloat16*)act, T, I);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_swiglu_bwd(const void* dact, const void* gu, void* dgu, long T, int I,
                              cudaStream_t stream) {
  STB_REQUIRE(I % 8 == 0, "swiglu: I=%d must be a multiple of 8", I);
  if (T == 0) return 0;
  const long total = T * (I / 8);
  long blocks = (total + 255) / 256;
  const long cap = long(sm_count()) * 16;
  swiglu_bwd_kernel<<<int(blocks < cap ? blocks : cap), 256, 0, stream>>>(
      (const __nv_bfloat16*)dact, (const __nv_bfloat16*)gu, (__nv_bfloat16*)dgu, T, I);
  STB_CHECK_LAUNCH();
  return 0;
}
