// MoE routing on the device, one kernel per stage, no library sort / scan and no host synchronisation:
//
//   stb_moe_router_fwd   gate logits [T, E] -> top-k expert ids, softmax(top-k logits) gate weights, per-expert pair
//                        counts, the position of every pair inside its expert's segment, and the per-expert sum of
//                        the full-softmax probabilities (the Switch load-balancing loss needs its mean).
//   stb_moe_router_bwd   d(gate weights) [T, k] and d(aux loss) -> d(logits) [T, E]
//   stb_ep_plan          after the count exchange ([ep, E] table): destination rank and row of every (token, k) pair
//                        in the receivers' 128-row-padded expert-major regions, the padded offsets / tile->expert table
//                        of this rank's own experts for the grouped GEMM, the number of dropped pairs.
//   stb_ep_zero_padding  zeroes only the padding rows (< 128 per expert) of a receive region.
//   stb_copy_rows_dyn    copies the first *n_rows (device scalar) rows of a region into a private buffer.
//
// Reference: scaletorch/models/model_qwen3_moe.py:30-171 (torch.topk / softmax / bincount, two .tolist() syncs in
// ep_comms.py:96,102) -- and this repository's round-1 path, which built the plan from argsort + four cumsums + scatter.
// One warp per token; E <= 256 (8 logits per lane), k <= 16.
#include <math_constants.h>

#include "common/host.h"
#include "common/ptx.cuh"

namespace stb {

constexpr int kRouterMaxPerLane = 8;   // E <= 256
constexpr int kRouterMaxK = 16;

// warp-wide argmax with smallest-index tie break: returns (value, index) on every lane
STB_DEVICE void warp_argmax(float& v, int& i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, i, o);
    if (ov > v || (ov == v && oi < i)) {
      v = ov;
      i = oi;
    }
  }
}

__global__ void __launch_bounds__(256)
    moe_router_fwd_kernel(const __nv_bfloat16* __restrict__ logits, long T, int E, int k, float* __restrict__ weights,
                          int* __restrict__ idx, int* __restrict__ pos, int* __restrict__ counts,
                          float* __restrict__ psum) {
  extern __shared__ float s_psum[];                 // [E] block-local sum of softmax probabilities
  int* s_cnt = reinterpret_cast<int*>(s_psum + E);  // [E] block-local pair counts (only to pre-aggregate the atomics)
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    s_psum[e] = 0.f;
    s_cnt[e] = 0;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warps = blockDim.x >> 5;
  const int per = (E + 31) / 32;
  for (long t = long(blockIdx.x) * warps + (threadIdx.x >> 5); t < T; t += long(gridDim.x) * warps) {
    float v[kRouterMaxPerLane];
    float mx = -CUDART_INF_F;
#pragma unroll
    for (int j = 0; j < kRouterMaxPerLane; ++j) {
      const int e = j * 32 + lane;
      v[j] = (j < per && e < E) ? __bfloat162float(logits[t * E + e]) : -CUDART_INF_F;
      mx = fmaxf(mx, v[j]);
    }
    mx = warp_max(mx);
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < kRouterMaxPerLane; ++j) sum += (v[j] > -CUDART_INF_F) ? __expf(v[j] - mx) : 0.f;
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
#pragma unroll
    for (int j = 0; j < kRouterMaxPerLane; ++j) {
      const int e = j * 32 + lane;
      if (j < per && e < E) atomicAdd(&s_psum[e], __expf(v[j] - mx) * inv);
    }
    // top-k by k rounds of warp argmax
    float sel_v[kRouterMaxK];
    int sel_i[kRouterMaxK];
#pragma unroll
    for (int r = 0; r < kRouterMaxK; ++r) {
      sel_v[r] = -CUDART_INF_F;
      sel_i[r] = 0;
      if (r < k) {
        float bv = -CUDART_INF_F;
        int bi = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < kRouterMaxPerLane; ++j) {
          const int e = j * 32 + lane;
          if (v[j] > bv) {       // ascending e within a lane: first maximum wins
            bv = v[j];
            bi = e;
          }
        }
        warp_argmax(bv, bi);
        sel_v[r] = bv;
        sel_i[r] = bi;
#pragma unroll
        for (int j = 0; j < kRouterMaxPerLane; ++j)
          if (j * 32 + lane == bi) v[j] = -CUDART_INF_F;
      }
    }
    // gate weights: softmax over the k selected logits (sel_v[0] is the maximum); lane r keeps selection r
    float wsum = 0.f, mv = 0.f;
    int mi = 0;
#pragma unroll
    for (int r = 0; r < kRouterMaxK; ++r) {
      if (r < k) wsum += __expf(sel_v[r] - sel_v[0]);
      if (r == lane) {
        mv = sel_v[r];
        mi = sel_i[r];
      }
    }
    if (lane < k) {
      weights[t * k + lane] = __expf(mv - sel_v[0]) / wsum;
      idx[t * k + lane] = mi;
      atomicAdd(&s_cnt[mi], 1);
    }
  }
  __syncthreads();
  // one global atomic per (block, expert): the block claims a contiguous range of positions inside every expert ...
  int* s_base = s_cnt;   // reuse: after the claim s_cnt[e] holds the block's first position
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    const int c = s_cnt[e];
    if (s_psum[e] != 0.f) atomicAdd(&psum[e], s_psum[e]);
    s_base[e] = c > 0 ? atomicAdd(&counts[e], c) : 0;
  }
  __syncthreads();
  // ... and hands them out to its pairs in a second pass over the ids it wrote
  for (long t = long(blockIdx.x) * warps + (threadIdx.x >> 5); t < T; t += long(gridDim.x) * warps) {
    if (lane < k) pos[t * k + lane] = atomicAdd(&s_base[idx[t * k + lane]], 1);
  }
}

// dlogits[t, e] = [e selected as j] w_j (dw_j - sum_i w_i dw_i)  +  g_aux * coef * E / T * p_te (f_e - sum_e' f_e' p_te')
__global__ void __launch_bounds__(256)
    moe_router_bwd_kernel(const __nv_bfloat16* __restrict__ logits, const float* __restrict__ weights,
                          const int* __restrict__ idx, const float* __restrict__ dweights,
                          const int* __restrict__ counts, const float* __restrict__ g_aux, float aux_scale, long T,
                          int E, int k, __nv_bfloat16* __restrict__ dlogits) {
  const int lane = threadIdx.x & 31, warps = blockDim.x >> 5;
  const int per = (E + 31) / 32;
  const float ga = (g_aux != nullptr ? *g_aux : 0.f) * aux_scale;      // aux_scale = coef * E / T
  const float inv_T = 1.f / float(T);
  for (long t = long(blockIdx.x) * warps + (threadIdx.x >> 5); t < T; t += long(gridDim.x) * warps) {
    float out[kRouterMaxPerLane];
#pragma unroll
    for (int j = 0; j < kRouterMaxPerLane; ++j) out[j] = 0.f;
    if (ga != 0.f) {
      float v[kRouterMaxPerLane], f[kRouterMaxPerLane];
      float mx = -CUDART_INF_F;
#pragma unroll
      for (int j = 0; j < kRouterMaxPerLane; ++j) {
        const int e = j * 32 + lane;
        const bool ok = j < per && e < E;
        v[j] = ok ? __bfloat162float(logits[t * E + e]) : -CUDART_INF_F;
        f[j] = ok ? float(counts[e]) * inv_T : 0.f;
        mx = fmaxf(mx, v[j]);
      }
      mx = warp_max(mx);
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < kRouterMaxPerLane; ++j) {
        v[j] = (v[j] > -CUDART_INF_F) ? __expf(v[j] - mx) : 0.f;
        sum += v[j];
      }
      sum = warp_sum(sum);
      const float inv = 1.f / sum;
      float dot = 0.f;
#pragma unroll
      for (int j = 0; j < kRouterMaxPerLane; ++j) {
        v[j] *= inv;
        dot += f[j] * v[j];
      }
      dot = warp_sum(dot);
#pragma unroll
      for (int j = 0; j < kRouterMaxPerLane; ++j) out[j] = ga * v[j] * (f[j] - dot);
    }
    float wd = 0.f;
    for (int r = 0; r < k; ++r) wd += weights[t * k + r] * dweights[t * k + r];
    for (int r = 0; r < k; ++r) {
      const int e = idx[t * k + r];
      const float w = weights[t * k + r];
      const float dz = w * (dweights[t * k + r] - wd);
#pragma unroll
      for (int j = 0; j < kRouterMaxPerLane; ++j)
        if (j * 32 + lane == e) out[j] += dz;
    }
#pragma unroll
    for (int j = 0; j < kRouterMaxPerLane; ++j) {
      const int e = j * 32 + lane;
      if (j < per && e < E) dlogits[t * E + e] = __float2bfloat16(out[j]);
    }
  }
}

// Every block rebuilds the (tiny) per-expert base table in shared memory, then the grid walks the pairs.
// table [ep, E]: pairs source rank r sends to global expert e.  Receive regions are expert-major with each expert's
// segment padded to 128 rows; inside a segment rows are ordered by source rank, then by the sender's `pos`.
__global__ void __launch_bounds__(256)
    ep_plan_kernel(const int* __restrict__ table, const int* __restrict__ flat, const int* __restrict__ pos, long P,
                   int E, int ep, int rank, int rows_cap, int* __restrict__ dst_row, int* __restrict__ dst_rank,
                   int* __restrict__ dropped, int* __restrict__ offs_pad, int* __restrict__ tokens_per_expert,
                   int* __restrict__ tile_expert, int n_tiles) {
  extern __shared__ int s_plan[];
  int* s_base = s_plan;          // [E] first row of MY pairs inside expert e's segment on its owner
  int* s_tot = s_plan + E;       // [E] rows of expert e over all sources
  int* s_start = s_plan + 2 * E; // [E] padded start of expert e on its owner
  const int e_local = E / ep;
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    int tot = 0, before = 0;
    for (int r = 0; r < ep; ++r) {
      const int c = table[r * E + e];
      tot += c;
      if (r < rank) before += c;
    }
    s_tot[e] = tot;
    s_base[e] = before;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < E; e += blockDim.x) {
    const int first = (e / e_local) * e_local;
    int start = 0;
    for (int j = first; j < e; ++j) start += (s_tot[j] + 127) / 128 * 128;
    s_start[e] = start;
  }
  __syncthreads();
  if (blockIdx.x == 0) {         // layout of my own experts for the grouped GEMM
    const int mine = rank * e_local;
    for (int j = threadIdx.x; j <= e_local; j += blockDim.x) {
      int v;
      if (j < e_local) {
        v = s_start[mine + j];
        tokens_per_expert[j] = s_tot[mine + j];
      } else {
        v = s_start[mine + e_local - 1] + (s_tot[mine + e_local - 1] + 127) / 128 * 128;
      }
      offs_pad[j] = min(v, rows_cap);
    }
    const int end = min(s_start[mine + e_local - 1] + (s_tot[mine + e_local - 1] + 127) / 128 * 128, rows_cap);
    for (int t = threadIdx.x; t < n_tiles; t += blockDim.x) {
      const int row = t * 128;
      int te = -1;
      if (row < end) {
        te = 0;
        for (int j = 1; j < e_local; ++j)
          if (s_start[mine + j] <= row) te = j;
      }
      tile_expert[t] = te;
    }
  }
  int over = 0;
  for (long p = long(blockIdx.x) * blockDim.x + threadIdx.x; p < P; p += long(gridDim.x) * blockDim.x) {
    const int e = flat[p];
    const int row = s_start[e] + s_base[e] + pos[p];
    const bool ok = row < rows_cap;
    dst_row[p] = ok ? row : -1;
    dst_rank[p] = e / e_local;
    over += ok ? 0 : 1;
  }
  over = int(warp_sum(float(over)));
  if ((threadIdx.x & 31) == 0 && over > 0) atomicAdd(dropped, over);
}

// region [rows_cap, H] bf16; rows [offs_pad[e] + tokens_per_expert[e], offs_pad[e + 1]) of every local expert become 0
__global__ void __launch_bounds__(256)
    ep_zero_padding_kernel(__nv_bfloat16* __restrict__ region, const int* __restrict__ offs_pad,
                           const int* __restrict__ tokens_per_expert, int e_local, int H, int rows_cap) {
  const int vec = H >> 3;
  const int warps = blockDim.x >> 5, lane = threadIdx.x & 31;
  for (int w = blockIdx.x * warps + (threadIdx.x >> 5); w < e_local * 128; w += gridDim.x * warps) {
    const int e = w >> 7, j = w & 127;
    const int row = offs_pad[e] + tokens_per_expert[e] + j;
    if (row >= offs_pad[e + 1] || row >= rows_cap) continue;
    uint4* d = reinterpret_cast<uint4*>(region + long(row) * H);
    for (int c = lane; c < vec; c += 32) d[c] = make_uint4(0, 0, 0, 0);
  }
}

__global__ void __launch_bounds__(256)
    copy_rows_dyn_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, const int* __restrict__ n_rows,
                         long vec_per_row) {
  const long n = long(*n_rows) * vec_per_row;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) dst[i] = src[i];
}

}  // namespace stb

using namespace stb;

// logits: bf16 [T, E] contiguous.  weights fp32 [T, k]; idx / pos int32 [T, k]; counts int32 [E] and psum fp32 [E] must
// be zero on entry (they are accumulated with atomics).
extern "C" int stb_moe_router_fwd(const void* logits, long T, int E, int k, float* weights, int* idx, int* pos,
                                  int* counts, float* psum, cudaStream_t stream) {
  STB_REQUIRE(E >= 1 && E <= 32 * kRouterMaxPerLane, "moe_router: E = %d out of range (1..256)", E);
  STB_REQUIRE(k >= 1 && k <= kRouterMaxK && k <= E, "moe_router: k = %d out of range", k);
  if (T == 0) return 0;
  long blocks = (T + 7) / 8;
  const long cap = long(sm_count()) * 4;
  moe_router_fwd_kernel<<<int(blocks < cap ? blocks : cap), 256, E * 8, stream>>>(
      (const __nv_bfloat16*)logits, T, E, k, weights, idx, pos, counts, psum);
  STB_CHECK_LAUNCH();
  return 0;
}

// aux_scale = coef * E / T; g_aux: device scalar (gradient of the aux-loss output) or NULL.
extern "C" int stb_moe_router_bwd(const void* logits, const float* weights, const int* idx, const float* dweights,
                                  const int* counts, const float* g_aux, float aux_scale, long T, int E, int k,
                                  void* dlogits, cudaStream_t stream) {
  STB_REQUIRE(E >= 1 && E <= 32 * kRouterMaxPerLane && k >= 1 && k <= kRouterMaxK, "moe_router_bwd: E %d / k %d", E, k);
  if (T == 0) return 0;
  long blocks = (T + 7) / 8;
  const long cap = long(sm_count()) * 4;
  moe_router_bwd_kernel<<<int(blocks < cap ? blocks : cap), 256, 0, stream>>>(
      (const __nv_bfloat16*)logits, weights, idx, dweights, counts, g_aux, aux_scale, T, E, k, (__nv_bfloat16*)dlogits);
  STB_CHECK_LAUNCH();
  return 0;
}

// table int32 [ep, E]; flat / pos int32 [P]; outputs: dst_row / dst_rank [P], dropped [1] (zero on entry), offs_pad
// [E / ep + 1], tokens_per_expert [E / ep], tile_expert [n_tiles] (n_tiles = rows_cap / 128).
extern "C" int stb_ep_plan(const int* table, const int* flat, const int* pos, long P, int E, int ep, int rank,
                           int rows_cap, int* dst_row, int* dst_rank, int* dropped, int* offs_pad,
                           int* tokens_per_expert, int* tile_expert, int n_tiles, cudaStream_t stream) {
  STB_REQUIRE(ep >= 1 && E % ep == 0 && E <= 4096, "ep_plan: E %d / ep %d", E, ep);
  long blocks = (P + 255) / 256;
  if (blocks < 1) blocks = 1;
  const long cap = long(sm_count()) * 2;
  ep_plan_kernel<<<int(blocks < cap ? blocks : cap), 256, 3 * E * 4, stream>>>(
      table, flat, pos, P, E, ep, rank, rows_cap, dst_row, dst_rank, dropped, offs_pad, tokens_per_expert, tile_expert,
      n_tiles);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_ep_zero_padding(void* region, const int* offs_pad, const int* tokens_per_expert, int e_local, int H,
                                   int rows_cap, cudaStream_t stream) {
  STB_REQUIRE(H % 8 == 0, "ep_zero_padding: H %% 8");
  const int blocks = (e_local * 128 + 7) / 8;
  const int cap = sm_count() * 4;
  ep_zero_padding_kernel<<<blocks < cap ? blocks : cap, 256, 0, stream>>>((__nv_bfloat16*)region, offs_pad,
                                                                          tokens_per_expert, e_local, H, rows_cap);
  STB_CHECK_LAUNCH();
  return 0;
}

// dst[0 : *n_rows] = src[0 : *n_rows]  (rows of row_bytes bytes, a multiple of 16); *n_rows is read on the device
extern "C" int stb_copy_rows_dyn(const void* src, void* dst, const int* n_rows, long row_bytes, int max_rows,
                                 cudaStream_t stream) {
  STB_REQUIRE(row_bytes % 16 == 0, "copy_rows_dyn: row_bytes %% 16");
  if (max_rows <= 0) return 0;
  copy_rows_dyn_kernel<<<sm_count() * 4, 256, 0, stream>>>((const uint4*)src, (uint4*)dst, n_rows, row_bytes / 16);
  STB_CHECK_LAUNCH();
  return 0;
}
