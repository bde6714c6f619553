// Communication kernels over NVLink peer memory (symmetric buffers: the same allocation size on every rank,
// every peer's copy mapped into this process, optionally one NVLS multicast address for all of them).
//
//   stb_symm_barrier      all ranks of the group rendezvous on the device (release/acquire flags in peer memory)
//   stb_allreduce_f32     in-place sum (x scale) of an fp32 range of the symmetric buffer: each rank reduces its
//                         1/world slice by loading it from every peer and stores the result back to every peer
//                         (two-shot, one kernel); with an NVLS multicast mapping the loads/stores become
//                         multimem.ld_reduce / multimem.st and the reduction happens inside the switch.
//                         Used for the data-parallel gradient buckets (the 1/world scale is fused) and the
//                         non-SP tensor-parallel all-reduce.
//   stb_allreduce_bf16    same for bf16 ranges (tensor-parallel activations), fp32 accumulation.
//   stb_allgather_pull    each rank pulls the other ranks' shards out of their symmetric buffers into a local
//                         contiguous buffer and raises one ready flag per row block (consumed by the AG->GEMM
//                         producer warp of stb_gemm_bf16_ex).
//
// Reference counterparts: dist.all_reduce / all_gather / reduce_scatter through c10d-NCCL in
// scaletorch/parallel/data_parallel/bucket.py:73-78, tensor_parallel/tp_comms.py:144-147,
// sequence_parallel/sp_comms.py:46-47,79-82.
//
// Flags are monotonically increasing epochs (never reset), so consecutive calls need no re-zeroing.
// Every kernel here is launched with a grid that is fully resident (<= SM count), because its blocks wait
// on one another / on other GPUs.
#include "common/host.h"
#include "common/ptx.cuh"
#include "common/symm.cuh"

namespace stb {

__global__ void symm_barrier_kernel(uint32_t* const* flag_pages, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  __syncwarp();
  if (threadIdx.x == 0) symm_epoch_store(flag_pages, rank, slot, epoch_arg, epoch + 1);
}

STB_DEVICE float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
STB_DEVICE void st_peer_f4(float* p, const float4& v) {
  asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
STB_DEVICE float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
STB_DEVICE void multimem_st_f4(float* mc, const float4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
STB_DEVICE uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
STB_DEVICE void multimem_st_u4(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}

// In-place all-reduce of `n` fp32 elements starting `offset` elements into every rank's symmetric buffer.
template <bool MULTIMEM>
__global__ void __launch_bounds__(512)
    allreduce_f32_kernel(float* const* bufs, float* mc, uint32_t* const* flag_pages, long offset, long n, float scale,
                         int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);   // every rank's data is in place
  const long nvec = n >> 2;
  const long per = (nvec + world - 1) / world;
  const long lo = per * rank, hi = min(nvec, lo + per);
  if (MULTIMEM) {
    // NVLS: the switch reduces; per thread 8 independent 16-byte ld_reduce in flight, then 8 broadcast stores, so a
    // few CTAs saturate the link and the rest of the SMs stay with the backward pass
    constexpr int U = 8;
    const long stride = long(gridDim.x) * blockDim.x;
    for (long i0 = lo + long(blockIdx.x) * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * U) {
      float4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < hi) acc[u] = multimem_ld_reduce_f4(mc + offset + (i << 2));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < hi) {
          acc[u].x *= scale; acc[u].y *= scale; acc[u].z *= scale; acc[u].w *= scale;
          multimem_st_f4(mc + offset + (i << 2), acc[u]);
        }
      }
    }
  } else {
    // peer loads/stores: 8 independent 16-byte loads in flight per thread and peer (64 KB per CTA), so ~24 CTAs keep
    // the NVLink ports busy and the remaining SMs stay with the backward pass (one load in flight per thread needed
    // all 148 SMs to reach the same bandwidth)
    constexpr int U = 8;
    const long stride = long(gridDim.x) * blockDim.x;
    for (long i0 = lo + long(blockIdx.x) * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * U) {
      float4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        acc[u] = (i < hi) ? ld_peer_f4(bufs[rank] + offset + (i << 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
      for (int p = 1; p < world; ++p) {
        const float* src = bufs[(rank + p) % world] + offset;
        float4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long i = i0 + u * stride;
          v[u] = (i < hi) ? ld_peer_f4(src + (i << 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          acc[u].x += v[u].x; acc[u].y += v[u].y; acc[u].z += v[u].z; acc[u].w += v[u].w;
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        acc[u].x *= scale; acc[u].y *= scale; acc[u].z *= scale; acc[u].w *= scale;
      }
      for (int p = 0; p < world; ++p) {
        float* dst = bufs[(rank + p) % world] + offset;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long i = i0 + u * stride;
          if (i < hi) st_peer_f4(dst + (i << 2), acc[u]);
        }
      }
    }
  }
  if (rank == 0) {  // scalar tail (n % 4): rank 0 alone sums and stores it with plain accesses while the others run ahead -- correct only because every rank waits in the closing barrier below before anyone reads the result
    for (long i = (nvec << 2) + long(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += long(gridDim.x) * blockDim.x) {
      float acc = 0.f;
      for (int p = 0; p < world; ++p) acc += bufs[p][offset + i];
      acc *= scale;
      for (int p = 0; p < world; ++p) bufs[p][offset + i] = acc;
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);  // results have landed everywhere
}

template <bool MULTIMEM>
__global__ void __launch_bounds__(512)
    allreduce_bf16_kernel(__nv_bfloat16* const* bufs, __nv_bfloat16* mc, uint32_t* const* flag_pages, long offset,
                          long n, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  const long nvec = n >> 3;
  const long per = (nvec + world - 1) / world;
  const long lo = per * rank, hi = min(nvec, lo + per);
  if (MULTIMEM) {
    // NVLS: 8 independent 16-byte ld_reduce in flight per thread, then 8 broadcast stores (as in the fp32 kernel)
    constexpr int U = 8;
    const long stride = long(gridDim.x) * blockDim.x;
    for (long i0 = lo + long(blockIdx.x) * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * U) {
      uint4 acc[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < hi) acc[u] = multimem_ld_reduce_bf16x8(mc + offset + (i << 3));
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < hi) multimem_st_u4(mc + offset + (i << 3), acc[u]);
      }
    }
  } else {
    for (long i = lo + long(blockIdx.x) * blockDim.x + threadIdx.x; i < hi; i += long(gridDim.x) * blockDim.x) {
      const long e = offset + (i << 3);
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int p = 0; p < world; ++p) {
        const float4 vf = ld_peer_f4(reinterpret_cast<const float*>(bufs[(rank + p) % world] + e));
        const uint32_t u[4] = {__float_as_uint(vf.x), __float_as_uint(vf.y), __float_as_uint(vf.z),
                               __float_as_uint(vf.w)};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = unpack_bf16x2(u[j]);
          acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
        }
      }
      const uint4 out = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]), pack_bf16x2(acc[4], acc[5]),
                                   pack_bf16x2(acc[6], acc[7]));
      for (int p = 0; p < world; ++p) *reinterpret_cast<uint4*>(bufs[(rank + p) % world] + e) = out;
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);
}

// out (local, fresh) <- sum over ranks of bufs[p][offset .. offset+n): every rank reads all peers (one-shot).
__global__ void __launch_bounds__(512)
    allreduce_oneshot_bf16_kernel(const __nv_bfloat16* const* bufs, uint32_t* const* flag_pages, long offset,
                                  __nv_bfloat16* out, long n, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  const long nvec = n >> 3;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += long(gridDim.x) * blockDim.x) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int p = 0; p < world; ++p) {
      const uint4 v = ld_peer_v4(bufs[(rank + p) % world] + offset + (i << 3));
      const uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = unpack_bf16x2(u[j]);
        acc[2 * j] += f.x; acc[2 * j + 1] += f.y;
      }
    }
    *reinterpret_cast<uint4*>(out + (i << 3)) = make_uint4(pack_bf16x2(acc[0], acc[1]), pack_bf16x2(acc[2], acc[3]),
                                                           pack_bf16x2(acc[4], acc[5]), pack_bf16x2(acc[6], acc[7]));
  }
  // the partial buffers may be overwritten only after every rank has read them
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);
}

// dst[world * rows_per_rank, row_bytes] <- shard of every rank (rank r's shard sits at offset_bytes of its buffer).
// One ready flag (epoch) per block of `flag_rows` rows is released in `ready` (local memory) when its bytes landed.
__global__ void __launch_bounds__(512)
    allgather_pull_kernel(const uint8_t* const* bufs, uint32_t* const* flag_pages, long offset_bytes, uint8_t* dst,
                          long rows_per_rank, long row_bytes, int flag_rows, uint32_t* ready, int rank, int world,
                          int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);   // every rank's shard is written
  const long blocks_per_rank = (rows_per_rank + flag_rows - 1) / flag_rows;
  const long total_blocks = blocks_per_rank * world;
  const long vec_per_row = row_bytes >> 4;
  for (long blk = blockIdx.x; blk < total_blocks; blk += gridDim.x) {
    // local shard first, then the peers in ring order
    const int src = int((rank + blk / blocks_per_rank) % world);
    const long rb = blk % blocks_per_rank;
    const long r0 = rb * flag_rows, r1 = min(rows_per_rank, r0 + flag_rows);
    const uint4* s = reinterpret_cast<const uint4*>(bufs[src] + offset_bytes + r0 * row_bytes);
    uint4* d = reinterpret_cast<uint4*>(dst + (long(src) * rows_per_rank + r0) * row_bytes);
    const long nv = (r1 - r0) * vec_per_row;
    for (long i = threadIdx.x; i < nv; i += blockDim.x) d[i] = ld_peer_v4(s + i);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) st_release_gpu(ready + long(src) * blocks_per_rank + rb, epoch);
  }
  // peers may reuse their shard buffers only after everybody has pulled them
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);
}

// Embedding gradient as token rows: every rank has staged (ids[T], dout[T, H]) in its symmetric window; each rank walks
// the rows of *all* ranks and scatter-adds scale * dout into its local fp32 gradient of the (vocab-shard of the)
// embedding matrix.  The gradient of an embedding touches at most T rows per rank, so exchanging the rows costs
// world * T * H * 2 bytes where all-reducing the [V, H] fp32 gradient costs V * H * 4 bytes (8 MB vs 622 MB per rank
// for Qwen3-0.6B) -- and that all-reduce sat at the very end of backward where nothing can hide it.
// Reference: the embedding weight goes through the same bucket all-reduce as every other parameter
// (scaletorch/parallel/data_parallel/bucket.py:58-78).
__global__ void __launch_bounds__(256)
    embedding_bwd_peers_kernel(const uint8_t* const* bufs, uint32_t* const* flag_pages, long ids_off, long dout_off,
                               float* __restrict__ dW, long T, int H, int vocab_start, int vocab_local, float scale,
                               int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);       // every rank has staged its rows
  const int vec_per_row = H >> 3;
  const long per_rank = T * vec_per_row;
  const long total = per_rank * world;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += long(gridDim.x) * blockDim.x) {
    const int src = int((rank + i / per_rank) % world);               // own rows first, then the peers in ring order
    const long r = i % per_rank;
    const long t = r / vec_per_row;
    const int c = int(r % vec_per_row) * 8;
    const long id = reinterpret_cast<const long*>(bufs[src] + ids_off)[t] - vocab_start;
    if (id < 0 || id >= vocab_local) continue;
    const uint4 d = ld_peer_v4(reinterpret_cast<const __nv_bfloat16*>(bufs[src] + dout_off) + t * H + c);
    const uint32_t du[4] = {d.x, d.y, d.z, d.w};
    float* dst = dW + id * H + c;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(du[j]);
      atomicAdd(dst + 2 * j, f.x * scale);
      atomicAdd(dst + 2 * j + 1, f.y * scale);
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);   // staging areas may be reused
}

// ---- ZeRO-1 over peer memory: reduce-scatter of gradient buckets + sharded AdamW that broadcasts the new parameters ----
// Every rank owns one contiguous shard [own_lo, own_hi) of the flat fp32 gradient / bf16 parameter buffers.  A bucket
// [offset, offset + n) is reduced only on the rank(s) whose shard it intersects: that rank sums the range over all
// ranks (multimem.ld_reduce through the switch, or peer loads), scales it and stores it into ITS buffer alone -- the
// other ranks never need the reduced gradient, so the broadcast half of an all-reduce (n x 4 bytes of NVLink ingress
// and HBM writes per rank) disappears, and AdamW touches 1/world of the elements.  What travels instead is the bf16
// parameter shard after the update (2 bytes per parameter).
// Reference counterpart: the bucket all-reduce of scaletorch/parallel/data_parallel/bucket.py:58-78 followed by a
// replicated torch.optim.AdamW step (trainer/model_builder.py:119-134).
template <bool MULTIMEM>
__global__ void __launch_bounds__(512)
    reduce_scatter_f32_kernel(float* const* bufs, const float* mc, uint32_t* const* flag_pages, long offset, long n,
                              long own_lo, long own_hi, float scale, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);   // every rank's bucket is complete
  const long lo = max(offset, own_lo), hi = min(offset + n, own_hi);       // multiples of 4 except possibly `hi`
  if (lo < hi) {
    float* mine = bufs[rank];
    const long v0 = lo >> 2, v1 = hi >> 2;
    constexpr int U = 8;
    const long stride = long(gridDim.x) * blockDim.x;
    for (long i0 = v0 + long(blockIdx.x) * blockDim.x + threadIdx.x; i0 < v1; i0 += stride * U) {
      float4 acc[U];
      if (MULTIMEM) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long i = i0 + u * stride;
          if (i < v1) acc[u] = multimem_ld_reduce_f4(mc + (i << 2));
        }
      } else {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const long i = i0 + u * stride;
          acc[u] = (i < v1) ? ld_peer_f4(mine + (i << 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int p = 1; p < world; ++p) {
          const float* src = bufs[(rank + p) % world];
          float4 v[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const long i = i0 + u * stride;
            v[u] = (i < v1) ? ld_peer_f4(src + (i << 2)) : make_float4(0.f, 0.f, 0.f, 0.f);
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            acc[u].x += v[u].x; acc[u].y += v[u].y; acc[u].z += v[u].z; acc[u].w += v[u].w;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < v1) {
          acc[u].x *= scale; acc[u].y *= scale; acc[u].z *= scale; acc[u].w *= scale;
          *reinterpret_cast<float4*>(mine + (i << 2)) = acc[u];
        }
      }
    }
    // scalar tail (hi not a multiple of 4): plain peer loads
    for (long i = (v1 << 2) + long(blockIdx.x) * blockDim.x + threadIdx.x; i < hi; i += stride) {
      float acc = 0.f;
      for (int p = 0; p < world; ++p) acc += bufs[p][i];
      mine[i] = acc * scale;
    }
  }
  // peers may overwrite their gradient buffers (next backward) only after every owner has read them
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);
}

STB_DEVICE void multimem_st_b128(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w)
               : "memory");
}
STB_DEVICE void st_peer_b128(void* p, const uint4& v) {
  asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// AdamW on this rank's shard: g / m / v are local fp32 arrays of `n` elements (n % 8 == 0), the bf16 parameters are
// read from the local copy at element `p_off` of the symmetric parameter buffer and the updated values are written to
// EVERY rank's copy (one multimem.st per 16 bytes through the switch, or one store per peer).  Opening barrier: no
// rank may still be reading parameters (its backward has finished); closing barrier: when the kernel completes on a
// rank, every rank's writes into that rank's copy have landed.  hyper = {lr, 1 - beta1^t, 1 - beta2^t} on the device.
template <bool MULTIMEM>
__global__ void __launch_bounds__(512)
    adamw_shard_bcast_kernel(__nv_bfloat16* const* pbufs, __nv_bfloat16* pmc, uint32_t* const* flag_pages, long p_off,
                             const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long n,
                             float beta1, float beta2, float eps, float weight_decay,
                             const float* __restrict__ clip_coef_ptr, float grad_scale, const float* __restrict__ hyper,
                             int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  const float lr = hyper[0], inv_bc1 = 1.f / hyper[1], inv_sqrt_bc2 = rsqrtf(hyper[2]);
  const float clip = (clip_coef_ptr ? *clip_coef_ptr : 1.f) * grad_scale;
  const float decay = 1.f - lr * weight_decay;
  const __nv_bfloat16* p_local = pbufs[rank] + p_off;
  const long nvec = n >> 3;
  for (long i = long(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += long(gridDim.x) * blockDim.x) {
    const float4 g0 = reinterpret_cast<const float4*>(g)[2 * i], g1 = reinterpret_cast<const float4*>(g)[2 * i + 1];
    const float4 m0 = reinterpret_cast<const float4*>(m)[2 * i], m1 = reinterpret_cast<const float4*>(m)[2 * i + 1];
    const float4 v0 = reinterpret_cast<const float4*>(v)[2 * i], v1 = reinterpret_cast<const float4*>(v)[2 * i + 1];
    const uint4 pr = reinterpret_cast<const uint4*>(p_local)[i];
    float ga[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    float ma[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    float va[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    const uint32_t pu[4] = {pr.x, pr.y, pr.z, pr.w};
    float pf[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = unpack_bf16x2(pu[j]);
      pf[2 * j] = f.x; pf[2 * j + 1] = f.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = ga[j] * clip;
      ma[j] = beta1 * ma[j] + (1.f - beta1) * gj;
      va[j] = beta2 * va[j] + (1.f - beta2) * gj * gj;
      pf[j] = pf[j] * decay - lr * (ma[j] * inv_bc1) / (sqrtf(va[j]) * inv_sqrt_bc2 + eps);
    }
    reinterpret_cast<float4*>(m)[2 * i] = make_float4(ma[0], ma[1], ma[2], ma[3]);
    reinterpret_cast<float4*>(m)[2 * i + 1] = make_float4(ma[4], ma[5], ma[6], ma[7]);
    reinterpret_cast<float4*>(v)[2 * i] = make_float4(va[0], va[1], va[2], va[3]);
    reinterpret_cast<float4*>(v)[2 * i + 1] = make_float4(va[4], va[5], va[6], va[7]);
    const uint4 out = make_uint4(pack_bf16x2(pf[0], pf[1]), pack_bf16x2(pf[2], pf[3]), pack_bf16x2(pf[4], pf[5]),
                                 pack_bf16x2(pf[6], pf[7]));
    if (MULTIMEM) {
      multimem_st_b128(pmc + p_off + (i << 3), out);
    } else {
      for (int p = 0; p < world; ++p) st_peer_b128(pbufs[(rank + p) % world] + p_off + (i << 3), out);
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);
}

// Broadcast of one owner's parameter shard: elements [p_off, p_off + n) of the symmetric bf16 parameter buffer are
// copied from rank `owner`'s copy into every rank's copy (multimem.st through the switch: the owner sends each byte
// once, or one store per peer).  Every rank launches the kernel (opening barrier: nobody still reads the old values;
// closing barrier: the values have landed everywhere); only the owner moves data.  Used by the sharded optimiser to
// all-gather the updated parameters one owner after the other at the START of the next step, on the communication
// stream, so the forward pass begins as soon as the first shards are in and the rest of the gather hides behind it.
template <bool MULTIMEM>
__global__ void __launch_bounds__(512)
    param_bcast_kernel(__nv_bfloat16* const* pbufs, __nv_bfloat16* pmc, uint32_t* const* flag_pages, long p_off, long n,
                       int owner, int rank, int world, int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  grid_cross_gpu_barrier(flag_pages, rank, world, slot, epoch);
  if (rank == owner) {
    const uint4* src = reinterpret_cast<const uint4*>(pbufs[rank] + p_off);
    const long nvec = n >> 3;
    constexpr int U = 4;
    const long stride = long(gridDim.x) * blockDim.x;
    for (long i0 = long(blockIdx.x) * blockDim.x + threadIdx.x; i0 < nvec; i0 += stride * U) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < nvec) v[u] = src[i];
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long i = i0 + u * stride;
        if (i < nvec) {
          if (MULTIMEM) {
            multimem_st_b128(pmc + p_off + (i << 3), v[u]);
          } else {
            for (int p = 1; p < world; ++p) st_peer_b128(pbufs[(rank + p) % world] + p_off + (i << 3), v[u]);
          }
        }
      }
    }
  }
  grid_finish_cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1, epoch_arg);
}

// In-place sum of `n` (<= 32) fp32 values living at element `offset` of every rank's symmetric buffer: every rank
// reads all copies and writes the total into its OWN copy.  Used for the squared-norm partials of the sharded
// optimiser; one CTA, two device barriers, no library call in the step.
__global__ void __launch_bounds__(32)
    allreduce_tiny_f32_kernel(float* const* bufs, uint32_t* const* flag_pages, long offset, int n, int rank, int world,
                              int slot, uint32_t epoch_arg) {
  const uint32_t epoch = symm_epoch_load(flag_pages, rank, slot, epoch_arg);
  cross_gpu_barrier(flag_pages, rank, world, slot, epoch);          // every rank's partials are written
  __syncwarp();
  float acc = 0.f;
  if (int(threadIdx.x) < n)
    for (int p = 0; p < world; ++p) {
      float x;
      asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(x) : "l"(bufs[p] + offset + threadIdx.x) : "memory");
      acc += x;
    }
  __syncwarp();
  cross_gpu_barrier(flag_pages, rank, world, slot, epoch + 1);      // everybody has read everybody's partials
  __syncwarp();
  if (int(threadIdx.x) < n) bufs[rank][offset + threadIdx.x] = acc;
  if (threadIdx.x == 0) symm_epoch_store(flag_pages, rank, slot, epoch_arg, epoch + 2);
}

static int comm_grid(int sm_budget) {
  const int sms = sm_count();
  if (sm_budget <= 0 || sm_budget > sms) sm_budget = sms;
  return sm_budget;
}

}  // namespace stb

using namespace stb;

extern "C" int stb_symm_flag_page_bytes() { return kFlagPageWords * 4; }

extern "C" int stb_symm_barrier(const void* flag_pages_dev, int rank, int world, int slot, unsigned epoch,
                                cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "symm_barrier: world %d / slot %d out of range", world, slot);
  symm_barrier_kernel<<<1, 32, 0, stream>>>((uint32_t* const*)flag_pages_dev, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes epochs `epoch` and `epoch + 1` of `slot`.
extern "C" int stb_allreduce_f32(const void* bufs_dev, void* multicast_ptr, const void* flag_pages_dev,
                                 long offset_elems, long n, float scale, int rank, int world, int slot,
                                 unsigned epoch, int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "allreduce: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(offset_elems % 4 == 0, "allreduce_f32: offset must be a multiple of 4 elements");
  if (n == 0) return 0;
  const int grid = comm_grid(sm_budget);
  if (multicast_ptr != nullptr)
    allreduce_f32_kernel<true><<<grid, 512, 0, stream>>>((float* const*)bufs_dev, (float*)multicast_ptr,
                                                         (uint32_t* const*)flag_pages_dev, offset_elems, n, scale, rank,
                                                         world, slot, epoch);
  else
    allreduce_f32_kernel<false><<<grid, 512, 0, stream>>>((float* const*)bufs_dev, nullptr,
                                                          (uint32_t* const*)flag_pages_dev, offset_elems, n, scale, rank,
                                                          world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_allreduce_bf16(const void* bufs_dev, void* multicast_ptr, const void* flag_pages_dev,
                                  long offset_elems, long n, int rank, int world, int slot, unsigned epoch,
                                  int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "allreduce: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(offset_elems % 8 == 0 && n % 8 == 0, "allreduce_bf16: offset and n must be multiples of 8 elements");
  if (n == 0) return 0;
  const int grid = comm_grid(sm_budget);
  if (multicast_ptr != nullptr)
    allreduce_bf16_kernel<true><<<grid, 512, 0, stream>>>((__nv_bfloat16* const*)bufs_dev, (__nv_bfloat16*)multicast_ptr,
                                                          (uint32_t* const*)flag_pages_dev, offset_elems, n, rank, world,
                                                          slot, epoch);
  else
    allreduce_bf16_kernel<false><<<grid, 512, 0, stream>>>((__nv_bfloat16* const*)bufs_dev, nullptr,
                                                           (uint32_t* const*)flag_pages_dev, offset_elems, n, rank,
                                                           world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_allreduce_oneshot_bf16(const void* bufs_dev, const void* flag_pages_dev, long offset_elems, void* out,
                                          long n, int rank, int world, int slot, unsigned epoch, int sm_budget,
                                          cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "allreduce: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(offset_elems % 8 == 0 && n % 8 == 0, "allreduce_oneshot_bf16: offset and n must be multiples of 8");
  if (n == 0) return 0;
  allreduce_oneshot_bf16_kernel<<<comm_grid(sm_budget), 512, 0, stream>>>(
      (const __nv_bfloat16* const*)bufs_dev, (uint32_t* const*)flag_pages_dev, offset_elems, (__nv_bfloat16*)out, n,
      rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

extern "C" int stb_allgather_pull(const void* bufs_dev, const void* flag_pages_dev, long offset_bytes, void* dst,
                                  long rows_per_rank, long row_bytes, int flag_rows, void* ready, int rank, int world,
                                  int slot, unsigned epoch, int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "allgather: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(row_bytes % 16 == 0 && offset_bytes % 16 == 0, "allgather: rows must be 16-byte multiples");
  const int grid = comm_grid(sm_budget);
  allgather_pull_kernel<<<grid, 512, 0, stream>>>((const uint8_t* const*)bufs_dev, (uint32_t* const*)flag_pages_dev,
                                                  offset_bytes, (uint8_t*)dst, rows_per_rank, row_bytes, flag_rows,
                                                  (uint32_t*)ready, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes 2 epochs.  ids (int64 [T]) at ids_off and dout (bf16 [T, H]) at dout_off of every rank's symmetric buffer;
// dW: local fp32 [vocab_local, H], accumulated into (the caller zeroes it when this is the first contribution).
extern "C" int stb_embedding_bwd_peers(const void* bufs_dev, const void* flag_pages_dev, long ids_off, long dout_off,
                                       float* dW, long T, int H, int vocab_start, int vocab_local, float scale,
                                       int rank, int world, int slot, unsigned epoch, int sm_budget,
                                       cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "embedding_bwd_peers: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(H % 8 == 0 && ids_off % 8 == 0 && dout_off % 16 == 0, "embedding_bwd_peers: H %% 8 and aligned offsets required");
  embedding_bwd_peers_kernel<<<comm_grid(sm_budget), 256, 0, stream>>>(
      (const uint8_t* const*)bufs_dev, (uint32_t* const*)flag_pages_dev, ids_off, dout_off, dW, T, H, vocab_start,
      vocab_local, scale, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes 2 epochs.  The reduced, scaled range [max(offset, own_lo), min(offset + n, own_hi)) lands in this rank's
// buffer only; `multicast_ptr` is the NVLS address of the SAME symmetric buffer (element 0), or NULL for peer loads.
extern "C" int stb_reduce_scatter_f32(const void* bufs_dev, const void* multicast_ptr, const void* flag_pages_dev,
                                      long offset_elems, long n, long own_lo, long own_hi, float scale, int rank,
                                      int world, int slot, unsigned epoch, int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "reduce_scatter: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(offset_elems % 4 == 0 && own_lo % 4 == 0, "reduce_scatter_f32: offsets must be multiples of 4 elements");
  if (n == 0) return 0;
  const int grid = comm_grid(sm_budget);
  if (multicast_ptr != nullptr)
    reduce_scatter_f32_kernel<true><<<grid, 512, 0, stream>>>((float* const*)bufs_dev, (const float*)multicast_ptr,
                                                              (uint32_t* const*)flag_pages_dev, offset_elems, n, own_lo,
                                                              own_hi, scale, rank, world, slot, epoch);
  else
    reduce_scatter_f32_kernel<false><<<grid, 512, 0, stream>>>((float* const*)bufs_dev, nullptr,
                                                               (uint32_t* const*)flag_pages_dev, offset_elems, n, own_lo,
                                                               own_hi, scale, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes 2 epochs.  See adamw_shard_bcast_kernel.
extern "C" int stb_adamw_shard_bcast(const void* pbufs_dev, void* p_multicast_ptr, const void* flag_pages_dev,
                                     long p_offset_elems, const float* g, float* m, float* v, long n, float beta1,
                                     float beta2, float eps, float weight_decay, const float* clip_coef,
                                     float grad_scale, const float* hyper, int rank, int world, int slot,
                                     unsigned epoch, int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "adamw_shard_bcast: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(p_offset_elems % 8 == 0 && n % 8 == 0, "adamw_shard_bcast: offset and n must be multiples of 8 elements");
  STB_REQUIRE(hyper != nullptr, "adamw_shard_bcast: hyper must point at {lr, bc1, bc2} in device memory");
  const int grid = comm_grid(sm_budget);
  if (p_multicast_ptr != nullptr)
    adamw_shard_bcast_kernel<true><<<grid, 512, 0, stream>>>(
        (__nv_bfloat16* const*)pbufs_dev, (__nv_bfloat16*)p_multicast_ptr, (uint32_t* const*)flag_pages_dev,
        p_offset_elems, g, m, v, n, beta1, beta2, eps, weight_decay, clip_coef, grad_scale, hyper, rank, world, slot,
        epoch);
  else
    adamw_shard_bcast_kernel<false><<<grid, 512, 0, stream>>>(
        (__nv_bfloat16* const*)pbufs_dev, nullptr, (uint32_t* const*)flag_pages_dev, p_offset_elems, g, m, v, n, beta1,
        beta2, eps, weight_decay, clip_coef, grad_scale, hyper, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes 2 epochs.  n <= 32 fp32 values at element `offset` of every rank's symmetric buffer, summed in place.
extern "C" int stb_allreduce_tiny_f32(const void* bufs_dev, const void* flag_pages_dev, long offset_elems, int n,
                                      int rank, int world, int slot, unsigned epoch, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "allreduce_tiny: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(n >= 1 && n <= 32, "allreduce_tiny: n %d out of range (1..32)", n);
  allreduce_tiny_f32_kernel<<<1, 32, 0, stream>>>((float* const*)bufs_dev, (uint32_t* const*)flag_pages_dev,
                                                  offset_elems, n, rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}

// Consumes 2 epochs.  See param_bcast_kernel.
extern "C" int stb_param_bcast(const void* pbufs_dev, void* p_multicast_ptr, const void* flag_pages_dev,
                               long p_offset_elems, long n, int owner, int rank, int world, int slot, unsigned epoch,
                               int sm_budget, cudaStream_t stream) {
  STB_REQUIRE(world <= kMaxWorld && slot < kMaxSlots, "param_bcast: world %d / slot %d out of range", world, slot);
  STB_REQUIRE(p_offset_elems % 8 == 0 && n % 8 == 0, "param_bcast: offset and n must be multiples of 8 elements");
  STB_REQUIRE(owner >= 0 && owner < world, "param_bcast: owner %d", owner);
  const int grid = rank == owner ? comm_grid(sm_budget) : 1;
  if (p_multicast_ptr != nullptr)
    param_bcast_kernel<true><<<grid, 512, 0, stream>>>((__nv_bfloat16* const*)pbufs_dev, (__nv_bfloat16*)p_multicast_ptr,
                                                       (uint32_t* const*)flag_pages_dev, p_offset_elems, n, owner, rank,
                                                       world, slot, epoch);
  else
    param_bcast_kernel<false><<<grid, 512, 0, stream>>>((__nv_bfloat16* const*)pbufs_dev, nullptr,
                                                        (uint32_t* const*)flag_pages_dev, p_offset_elems, n, owner,
                                                        rank, world, slot, epoch);
  STB_CHECK_LAUNCH();
  return 0;
}
