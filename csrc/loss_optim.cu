// Loss / optimiser / embedding kernels (bf16 activations, fp32 statistics):
//   * cross-entropy over the (optionally vocab-sharded) logits: a statistics pass
//     (row max, sum-exp, target logit) and an in-place gradient pass that overwrites
//     the logits with (softmax - onehot) * scale, so full-vocab fp32 logits and a
//     separate softmax-backward never exist (reference: trainer/train_step.py:89-103,
//     parallel/pipeline_parallel/pipeline_parallel.py:326-333).
//   * fused AdamW over the flat parameter / gradient / state buffers with the clip
//     coefficient read from device memory (reference: trainer/model_builder.py:119-134,
//     trainer/train_step.py:122-136 -- torch fused AdamW + clip_grad_norm_).
//   * squared-L2-norm reduction of the flat gradient buffer.
//   * embedding gather and scatter-add backward (reference: models/model_qwen3.py:109-110).
#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

constexpr int kCeThreads = 512;

struct MaxSum {
  float m, s;
};
__device__ __forceinline__ MaxSum ms_merge(MaxSum a, MaxSum b) {
  const float m = fmaxf(a.m, b.m);
  MaxSum r;
  r.m = m;
  r.s = (m == -INFINITY) ? 0.f : a.s * __expf(a.m - m) + b.s * __expf(b.m - m);
  return r;
}
__device__ __forceinline__ MaxSum block_ms_reduce(MaxSum v, MaxSum* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    MaxSum other;
    other.m = __shfl_xor_sync(0xffffffffu, v.m, o);
    other.s = __shfl_xor_sync(0xffffffffu, v.s, o);
    v = ms_merge(v, other);
  }
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  MaxSum t = red[0];
  for (int i = 1; i < nw; ++i) t = ms_merge(t, red[i]);
  __syncthreads();
  return t;
}

// stats[row] = {max, sumexp (relative to max), target logit or 0 when the target is not local}
__global__ void __launch_bounds__(kCeThreads)
    ce_stats_kernel(const __nv_bfloat16* __restrict__ logits, const long* __restrict__ targets,
                    float* __restrict__ stats, int rows, int V, long ld, int vocab_start, int merge) {
  __shared__ MaxSum red[kCeThreads / 32];
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const __nv_bfloat16* lr = logits + size_t(row) * ld;
    MaxSum acc;
    acc.m = -INFINITY;
    acc.s = 0.f;
    const int nvec = V >> 3;
    for (int v = threadIdx.x; v < nvec; v += kCeThreads) {
      const uint4 a = ld_nc_v4(lr + v * 8);
      const uint32_t au[4] = {a.x, a.y, a.z, a.w};
      float f[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 t = unpack_bf16x2(au[j]);
        f[2 * j] = t.x; f[2 * j + 1] = t.y;
      }
      float m = f[0];
#pragma unroll
      for (int j = 1; j < 8; ++j) m = fmaxf(m, f[j]);
      const float nm = fmaxf(acc.m, m);
      float s = acc.s * __expf(acc.m - nm);
#pragma unroll
      for (int j = 0; j < 8; ++j) s += __expf(f[j] - nm);
      acc.m = nm;
      acc.s = s;
    }
    for (int c = (nvec << 3) + threadIdx.x; c < V; c += kCeThreads) {
      MaxSum e;
      e.m = __bfloat162float(lr[c]);
      e.s = 1.f;
      acc = ms_merge(acc, e);
    }
    acc = block_ms_reduce(acc, red);
    if (threadIdx.x == 0) {
      const long t = targets[row] - vocab_start;
      float tl = (t >= 0 && t < V) ? __bfloat162float(lr[t]) : 0.f;
      if (merge) {   // vocabulary processed in column chunks: fold this chunk into the running (max, sum-exp, target)
        MaxSum old;
        old.m = stats[3 * row + 0];
        old.s = stats[3 * row + 1];
        acc = ms_merge(old, acc);
        tl += stats[3 * row + 2];
      }
      stats[3 * row + 0] = acc.m;
      stats[3 * row + 1] = acc.s;
      stats[3 * row + 2] = tl;
    }
  }
}

// logits <- (exp(logit - lse) - onehot) * scale ; loss_rows[row] = lse - target_logit
// lse_in: per-row log-sum-exp (already combined across vocab shards if any).
__global__ void __launch_bounds__(kCeThreads)
    ce_grad_kernel(__nv_bfloat16* __restrict__ logits, const long* __restrict__ targets,
                   const float* __restrict__ lse_in, int rows, int V, long ld, int vocab_start,
                   float scale, const float* __restrict__ scale_ptr, long ignore_index) {
  if (scale_ptr != nullptr) scale *= *scale_ptr;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    __nv_bfloat16* lr = logits + size_t(row) * ld;
    const float lse = lse_in[row];
    const long tgt = targets[row];
    const float sc = (tgt == ignore_index) ? 0.f : scale;
    const long t = tgt - vocab_start;
    const int nvec = V >> 3;
    for (int v = threadIdx.x; v < nvec; v += kCeThreads) {
      const uint4 a = *reinterpret_cast<const uint4*>(lr + v * 8);
      const uint32_t au[4] = {a.x, a.y, a.z, a.w};
      uint32_t o[4];
      const int c0 = v * 8;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(au[j]);
        float p0 = __expf(f.x - lse), p1 = __expf(f.y - lse);
        if (c0 + 2 * j == t) p0 -= 1.f;
        if (c0 + 2 * j + 1 == t) p1 -= 1.f;
        o[j] = pack_bf16x2(p0 * sc, p1 * sc);
      }
      *reinterpret_cast<uint4*>(lr + v * 8) = make_uint4(o[0], o[1], o[2], o[3]);
    }
    for (int c = (nvec << 3) + threadIdx.x; c < V; c += kCeThreads) {
      float pr = __expf(__bfloat162float(lr[c]) - lse);
      if (c == t) pr -= 1.f;
      lr[c] = __float2bfloat16(pr * sc);
    }
  }
}

// lse[row] = max + log(sumexp); loss[row] = lse - target (0 if ignored)
__global__ void ce_finalize_kernel(const float* __restrict__ stats, const long* __restrict__ targets,
                                   float* __restrict__ lse, float* __restrict__ loss_rows, int rows,
                                   long ignore_index) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const float l = stats[3 * r] + __logf(stats[3 * r + 1]);
  lse[r] = l;
  loss_rows[r] = (targets[r] == ignore_index) ? 0.f : (l - stats[3 * r + 2]);
}

// ------------------------------------------------------------------ AdamW
// p (bf16) / g (fp32 main grad) / m, v (fp32).  clip_coef_ptr (optional, device) multiplies g.
template <typename PT>
__global__ void __launch_bounds__(256)
    adamw_kernel(PT* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                 float* __restrict__ v, long n, float lr, float beta1, float beta2, float eps,
                 float weight_decay, float bc1, float bc2, const float* __restrict__ clip_coef_ptr,
                 float grad_scale, const float* __restrict__ hyper) {
  if (hyper != nullptr) {   // step-dependent scalars from device memory: the launch is replayable from a CUDA graph
    lr = hyper[0];
    bc1 = hyper[1];
    bc2 = hyper[2];
  }
  const float clip = (clip_coef_ptr ? *clip_coef_ptr : 1.f) * grad_scale;
  const long nvec = n >> 2;
  const float inv_bc1 = 1.f / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += long(gridDim.x) * blockDim.x) {
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 mm = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float pf[4];
    if constexpr (sizeof(PT) == 2) {
      const uint2 pr = reinterpret_cast<const uint2*>(p)[i];
      const float2 a = unpack_bf16x2(pr.x), b = unpack_bf16x2(pr.y);
      pf[0] = a.x; pf[1] = a.y; pf[2] = b.x; pf[3] = b.y;
    } else {
      const float4 pr = reinterpret_cast<const float4*>(p)[i];
      pf[0] = pr.x; pf[1] = pr.y; pf[2] = pr.z; pf[3] = pr.w;
    }
    const float ga[4] = {gg.x * clip, gg.y * clip, gg.z * clip, gg.w * clip};
    float ma[4] = {mm.x, mm.y, mm.z, mm.w}, va[4] = {vv.x, vv.y, vv.z, vv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      ma[j] = beta1 * ma[j] + (1.f - beta1) * ga[j];
      va[j] = beta2 * va[j] + (1.f - beta2) * ga[j] * ga[j];
      const float denom = sqrtf(va[j]) * inv_sqrt_bc2 + eps;
      pf[j] = pf[j] * (1.f - lr * weight_decay) - lr * (ma[j] * inv_bc1) / denom;
    }
    reinterpret_cast<float4*>(m)[i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(v)[i] = make_float4(va[0], va[1], va[2], va[3]);
    if constexpr (sizeof(PT) == 2) {
      reinterpret_cast<uint2*>(p)[i] = make_uint2(pack_bf16x2(pf[0], pf[1]), pack_bf16x2(pf[2], pf[3]));
    } else {
      reinterpret_cast<float4*>(p)[i] = make_float4(pf[0], pf[1], pf[2], pf[3]);
    }
  }
  // scalar tail
  const long tail0 = nvec << 2;
  for (long i = tail0 + long(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
    const float gi = g[i] * clip;
    const float mi = beta1 * m[i] + (1.f - beta1) * gi;
    const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
    m[i] = mi; v[i] = vi;
    float pi;
    if constexpr (sizeof(PT) == 2) pi = __bfloat162float(reinterpret_cast<__nv_bfloat16*>(p)[i]);
    else pi = reinterpret_cast<float*>(p)[i];
    pi = pi * (1.f - lr * weight_decay) - lr * (mi * inv_bc1) / (sqrtf(vi) * inv_sqrt_bc2 + eps);
    if constexpr (sizeof(PT) == 2) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(pi);
    else reinterpret_cast<float*>(p)[i] = pi;
  }
}

// partial[cta] = sum g^2 over this CTA's slice; last CTA to finish reduces all partials into
// out[0] (sum of squares) -- deterministic order, single launch.
__global__ void __launch_bounds__(256)
    sumsq_kernel(const float* __restrict__ g, long n, float* __restrict__ partial,
                 unsigned int* __restrict__ counter, float* __restrict__ out) {
  __shared__ float red[8];
  __shared__ bool is_last;
  float s = 0.f;
  const long nvec = n >> 2;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += long(gridDim.x) * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(g)[i];
    s += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
  }
  for (long i = (nvec << 2) + long(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x)
    s += g[i] * g[i];
  s = warp_sum(s);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    partial[blockIdx.x] = t;
    __threadfence();
    const unsigned int done = atomicAdd(counter, 1u);
    is_last = (done == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float t = 0.f;
    for (int i = threadIdx.x; i < gridDim.x; i += blockDim.x) t += reinterpret_cast<volatile float*>(partial)[i];
    t = warp_sum(t);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
      float tot = 0.f;
      for (int i = 0; i < 8; ++i) tot += red[i];
      out[0] = tot;
      *counter = 0;  // re-arm for the next call
    }
  }
}

// coef = min(1, max_norm / (sqrt(sumsq) + 1e-6)); norm_out = sqrt(sumsq)
__global__ void clip_coef_kernel(const float* __restrict__ sumsq, float max_norm, float* __restrict__ coef,
                                 float* __restrict__ norm_out) {
  const float nrm = sqrtf(sumsq[0]);
  if (norm_out) norm_out[0] = nrm;
  coef[0] = (max_norm > 0.f) ? fminf(1.f, max_norm / (nrm + 1e-6f)) : 1.f;
}

// -------------------------------------------------------------- embedding
// out[t, :] = W[ids[t] - vocab_start, :]   (zero row when the id is not local)
__global__ void __launch_bounds__(256)
    embedding_fwd_kernel(const long* __restrict__ ids, const __nv_bfloat16* __restrict__ W,
                         __nv_bfloat16* __restrict__ out, long T, int H, int vocab_start, int vocab_local) {
  const int vec_per_row = H >> 3;
  const long total = T * vec_per_row;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const long t = i / vec_per_row;
    const int c = int(i % vec_per_row) * 8;
    const long id = ids[t] - vocab_start;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (id >= 0 && id < vocab_local) v = *reinterpret_cast<const uint4*>(W + id * H + c);
    st_na_v4(out + t * H + c, v);
  }
}

// dW[ids[t], :] += dout[t, :]    (fp32 main-grad accumulate)
__global__ void __launch_bounds__(256)
    embedding_bwd_kernel(const long* __restrict__ ids, const __nv_bfloat16* __restrict__ dout,
                         float* __restrict__ dW, long T, int H, int vocab_start, int vocab_local) {
  const int vec_per_row = H >> 3;
  const long total = T * vec_per_row;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const long t = i / vec_per_row;
    const int c = int(i % vec_per_row) * 8;
    const long id = ids[t] - vocab_start;
    if (id < 0 || id >= vocab_local) continue;
    const uint4 d = ld_nc_v4(dout + t * H + c);
    const uint32_t du[4] = {d.x, d.y, d.z, d.w};
    float* dst = dW + id * H + c;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(du[j]);
      atomicAdd(dst + 2 * j, f.x);
      atomicAdd(dst + 2 * j + 1, f.y);
    }
  }
}

// dst (bf16) = src (fp32) * scale  -- used for bf16-on-the-wire gradient buckets
__global__ void __launch_bounds__(256)
    cast_scale_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long n,
                               float scale) {
  const long nvec = n >> 3;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += long(gridDim.x) * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
    const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
    reinterpret_cast<uint4*>(dst)[i] =
        make_uint4(pack_bf16x2(a.x * scale, a.y * scale), pack_bf16x2(a.z * scale, a.w * scale),
                   pack_bf16x2(b.x * scale, b.y * scale), pack_bf16x2(b.z * scale, b.w * scale));
  }
  for (long i = (nvec << 3) + long(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x)
    dst[i] = __float2bfloat16(src[i] * scale);
}

static int ew_grid(long work_items, int threads, int per_sm) {
  long blocks = (work_items + threads - 1) / threads;
  const long cap = long(sm_count()) * per_sm;
  if (blocks < 1) blocks = 1;
  return int(blocks < cap ? blocks : cap);
}

}  // namespace stb

using namespace stb;

extern "C" int stb_ce_stats(const void* logits, const long* targets, float* stats, int rows, int V,
                            long ld, int vocab_start, cudaStream_t stream) {
  STB_REQUIRE(ld % 8 == 0, "ce: logits row stride must be a multiple of 8");
  if (rows == 0) return 0;
  const int cap = sm_count() * 4;
  ce_stats_kernel<<<rows < cap ? rows : cap, kCeThreads, 0, stream>>>(
      (const __nv_bfloat16*)logits, targets, stats, rows, V, ld, vocab_start, 0);
  STB_CHECK_LAUNCH();
  return 0;
}

// Statistics of one column chunk [rows, Vc] of the logits (first column = global vocabulary id `vocab_start`); with
// merge != 0 the chunk is folded into the running statistics already in `stats` (online softmax over chunks).
extern "C" int stb_ce_stats_chunk(const void* logits, const long* targets, float* stats, int rows, int Vc, long ld,
                                  int vocab_start, int merge, cudaStream_t stream) {
  STB_REQUIRE(ld % 8 == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0, "ce: chunk must be 16-byte aligned, ld %% 8");
  if (rows == 0) return 0;
  const int cap = sm_count() * 4;
  ce_stats_kernel<<<rows < cap ? rows : cap, kCeThreads, 0, stream>>>(
      (const __nv_bfloat16*)logits, targets, stats, rows, Vc, ld, vocab_start, merge);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_ce_finalize(const float* stats, const long* targets, float* lse, float* loss_rows,
                               int rows, long ignore_index, cudaStream_t stream) {
  if (rows == 0) return 0;
  ce_finalize_kernel<<<(rows + 255) / 256, 256, 0, stream>>>(stats, targets, lse, loss_rows, rows,
                                                              ignore_index);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_ce_grad(void* logits, const long* targets, const float* lse, int rows, int V,
                           long ld, int vocab_start, float scale, const float* scale_ptr,
                           long ignore_index, cudaStream_t stream) {
  STB_REQUIRE(ld % 8 == 0, "ce: logits row stride must be a multiple of 8");
  if (rows == 0) return 0;
  const int cap = sm_count() * 4;
  ce_grad_kernel<<<rows < cap ? rows : cap, kCeThreads, 0, stream>>>(
      (__nv_bfloat16*)logits, targets, lse, rows, V, ld, vocab_start, scale, scale_ptr, ignore_index);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_adamw(void* p, int p_is_fp32, const float* g, float* m, float* v, long n, float lr,
                         float beta1, float beta2, float eps, float weight_decay, int step,
                         const float* clip_coef, float grad_scale, cudaStream_t stream) {
  if (n == 0) return 0;
  STB_REQUIRE(step >= 1, "adamw: step must be >= 1");
  const float bc1 = 1.f - powf(beta1, float(step)), bc2 = 1.f - powf(beta2, float(step));
  const int grid = ew_grid(n / 4 + 1, 256, 8);
  if (p_is_fp32)
    adamw_kernel<float><<<grid, 256, 0, stream>>>((float*)p, g, m, v, n, lr, beta1, beta2, eps,
                                                  weight_decay, bc1, bc2, clip_coef, grad_scale, nullptr);
  else
    adamw_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((__nv_bfloat16*)p, g, m, v, n, lr, beta1,
                                                          beta2, eps, weight_decay, bc1, bc2,
                                                          clip_coef, grad_scale, nullptr);
  STB_CHECK_LAUNCH();
  return 0;
}

// Same update with the step-dependent scalars read from device memory: hyper = {lr, 1 - beta1^step, 1 - beta2^step}.
// The host refreshes `hyper` (a 12-byte copy) before each step, so the launch itself never changes and can live in a
// CUDA graph.
extern "C" int stb_adamw_dyn(void* p, int p_is_fp32, const float* g, float* m, float* v, long n, float beta1,
                             float beta2, float eps, float weight_decay, const float* clip_coef,
                             float grad_scale, const float* hyper, cudaStream_t stream) {
  if (n == 0) return 0;
  STB_REQUIRE(hyper != nullptr, "adamw_dyn: hyper must point at {lr, bc1, bc2} in device memory");
  const int grid = ew_grid(n / 4 + 1, 256, 8);
  if (p_is_fp32)
    adamw_kernel<float><<<grid, 256, 0, stream>>>((float*)p, g, m, v, n, 0.f, beta1, beta2, eps, weight_decay,
                                                  1.f, 1.f, clip_coef, grad_scale, hyper);
  else
    adamw_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((__nv_bfloat16*)p, g, m, v, n, 0.f, beta1, beta2, eps,
                                                          weight_decay, 1.f, 1.f, clip_coef, grad_scale, hyper);
  STB_CHECK_LAUNCH();
  return 0;
}

// scratch: at least stb_sumsq_scratch_bytes() bytes; the first 4 bytes (the arrival counter) must be
// zero before the first call (the kernel re-arms it afterwards).
extern "C" long stb_sumsq_scratch_bytes() { return 16 + long(sm_count()) * 8 * sizeof(float); }

extern "C" int stb_sumsq(const float* g, long n, void* scratch, float* out, cudaStream_t stream) {
  const int grid = ew_grid(n / 4 + 1, 256, 8);
  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + 16);
  sumsq_kernel<<<grid, 256, 0, stream>>>(g, n, partial, counter, out);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_clip_coef(const float* sumsq, float max_norm, float* coef, float* norm_out,
                             cudaStream_t stream) {
  clip_coef_kernel<<<1, 1, 0, stream>>>(sumsq, max_norm, coef, norm_out);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_embedding_fwd(const long* ids, const void* W, void* out, long T, int H,
                                 int vocab_start, int vocab_local, cudaStream_t stream) {
  STB_REQUIRE(H % 8 == 0, "embedding: H must be a multiple of 8");
  if (T == 0) return 0;
  embedding_fwd_kernel<<<ew_grid(T * (H / 8), 256, 16), 256, 0, stream>>>(
      ids, (const __nv_bfloat16*)W, (__nv_bfloat16*)out, T, H, vocab_start, vocab_local);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_embedding_bwd(const long* ids, const void* dout, float* dW, long T, int H,
                                 int vocab_start, int vocab_local, cudaStream_t stream) {
  STB_REQUIRE(H % 8 == 0, "embedding: H must be a multiple of 8");
  if (T == 0) return 0;
  embedding_bwd_kernel<<<ew_grid(T * (H / 8), 256, 16), 256, 0, stream>>>(
      ids, (const __nv_bfloat16*)dout, dW, T, H, vocab_start, vocab_local);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_cast_scale_f32_bf16(const float* src, void* dst, long n, float scale,
                                       cudaStream_t stream) {
  if (n == 0) return 0;
  cast_scale_f32_bf16_kernel<<<ew_grid(n / 8 + 1, 256, 8), 256, 0, stream>>>(src, (__nv_bfloat16*)dst,
                                                                             n, scale);
  STB_CHECK_LAUNCH();
  return 0;
}
