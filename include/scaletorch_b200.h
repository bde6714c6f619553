/* C ABI of libscaletorch_b200.so -- hand-written sm_100a kernels for the
 * scaletorch_b200 training framework.
 *
 * Conventions
 *   - every function returns 0 on success; on failure a non-zero code is
 *     returned and stb_last_error() holds a message (thread-local);
 *   - all pointers are device pointers unless stated otherwise; bf16 tensors are
 *     passed as void*; `stream` is the CUDA stream the work is enqueued on;
 *   - nothing here allocates device memory: scratch buffers are passed in by the
 *     caller (the *_scratch_* helpers say how large they must be);
 *   - the library targets B200 (sm_100a) only.
 *
 * Each group cites the reference (jianzhnie/ScaleTorch) code whose behaviour it
 * provides.
 */
#ifndef SCALETORCH_B200_H_
#define SCALETORCH_B200_H_

#include <cuda_runtime_api.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* stb_last_error(void);
int stb_abi_version(void);

/* ---- GEMM (tcgen05 / TMEM / TMA), csrc/gemm_tcgen05.cu -------------------
 * C[M,N] (+)= A[M,K] * B[N,K]^T, bf16 inputs, fp32 accumulation.
 * a_mn / b_mn = 0: operand stored K-major (row-major [rows,K]); 1: MN-major
 * (row-major [K,rows]).  out_fp32: C is fp32 (else bf16).  accumulate: C += .
 * force_bn: 0 = auto, else 128 or 256 (N tile).
 * Reference: F.linear + autograd in scaletorch/parallel/tensor_parallel/
 * tensor_parallel.py:240,352 and tp_comms.py:265-308. */
int stb_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                  int ldc, int a_mn, int b_mn, int out_fp32, int accumulate, int force_bn,
                  cudaStream_t stream);

/* ---- RMSNorm (+ residual add), csrc/norm_rope.cu --------------------------
 * fwd: h = x (+ res); y = w * bf16(h * rsqrt(mean(h^2)+eps)); h_out/rstd optional.
 * bwd: dx = rmsnorm_bwd(dy) (+ dres); dw (+)= sum_rows dy * xn.  `scratch` needs
 *      stb_rmsnorm_bwd_scratch_elems(rows,H) floats.
 * Reference: scaletorch/models/attention_utils.py:247-271. */
int stb_rmsnorm_fwd(const void* x, const void* res, const void* w, void* h_out, void* y, float* rstd,
                    int rows, int H, float eps, cudaStream_t stream);
long stb_rmsnorm_bwd_scratch_elems(int rows, int H);
int stb_rmsnorm_bwd(const void* dy, const void* h, const void* w, const float* rstd, const void* dres,
                    void* dx, float* dw, int dw_accumulate, float* scratch, int rows, int H,
                    cudaStream_t stream);

/* RMSNorm backward without a caller-provided scratch buffer -- this entry point has no counterpart in
 * the reference; it exists for users of this C ABI who cannot allocate device memory (their process
 * gave all of HBM to its own arena at start-up, or they run where cudaMalloc's device-wide
 * synchronisation is not acceptable).  Same arguments and same result as stb_rmsnorm_bwd minus
 * `scratch`: the partial sums go to a 16 MiB __device__ array inside the library, which every GPU that
 * loads the library holds until exit.  Shapes with min(rows, 4*SMs) * H > stb_rmsnorm_bwd_noalloc_max_elems()
 * (4 Mi floats) are refused with an error.  The first call on a device resolves the array's address:
 * make it before any stream capture.  Never allocates.  Calls that overlap on the GPU share the array
 * and race; order them on one stream or with events.  csrc/norm_rope.cu. */
long stb_rmsnorm_bwd_noalloc_max_elems(void);
int stb_rmsnorm_bwd_noalloc(const void* dy, const void* h, const void* w, const float* rstd,
                            const void* dres, void* dx, float* dw, int dw_accumulate, int rows, int H,
                            cudaStream_t stream);

/* ---- per-head QK-RMSNorm + rotate-half RoPE, csrc/norm_rope.cu -----------
 * qkv: [T, (Hq+2Hkv)*128]; out: [T, (Hq+Hkv)*128]; cos/sin: fp32 [max_pos,128];
 * pos: optional int32 [T] (else position = t % S + pos_offset).
 * Reference: scaletorch/models/model_qwen3.py:205-217, attention_utils.py:170-239. */
int stb_qknorm_rope_fwd(const void* qkv, const void* wq, const void* wk, const float* cos_t,
                        const float* sin_t, const int* pos, void* out, float* rstd, int T, int S,
                        int Hq, int Hkv, int head_dim, int pos_offset, int use_norm, float eps,
                        cudaStream_t stream);
int stb_qknorm_rope_bwd(const void* dqk, const void* qkv, const void* wq, const void* wk,
                        const float* cos_t, const float* sin_t, const int* pos, const float* rstd,
                        void* dqkv, float* dwq, float* dwk, int T, int S, int Hq, int Hkv,
                        int head_dim, int pos_offset, int use_norm, cudaStream_t stream);

/* ---- SwiGLU on a fused gate|up buffer [T, 2I], csrc/norm_rope.cu ---------
 * Reference: scaletorch/models/model_qwen3.py:85-87. */
int stb_swiglu_fwd(const void* gu, void* act, long T, int I, cudaStream_t stream);
int stb_swiglu_bwd(const void* dact, const void* gu, void* dgu, long T, int I, cudaStream_t stream);

/* ---- cross-entropy on bf16 logits, csrc/loss_optim.cu --------------------
 * stats[row] = {max, sumexp, target_logit} over this vocab shard;
 * finalize: lse = max + log(sumexp), loss = lse - target;
 * grad: logits <- (softmax - onehot) * scale (* *scale_ptr if given, a device scalar), in place.
 * Reference: scaletorch/trainer/train_step.py:89-103. */
int stb_ce_stats(const void* logits, const long* targets, float* stats, int rows, int V, long ld,
                 int vocab_start, cudaStream_t stream);
int stb_ce_finalize(const float* stats, const long* targets, float* lse, float* loss_rows, int rows,
                    long ignore_index, cudaStream_t stream);
int stb_ce_grad(void* logits, const long* targets, const float* lse, int rows, int V, long ld,
                int vocab_start, float scale, const float* scale_ptr, long ignore_index,
                cudaStream_t stream);

/* ---- optimiser, csrc/loss_optim.cu ----------------------------------------
 * Fused AdamW over flat buffers; clip_coef (device scalar, optional) and
 * grad_scale multiply the gradient.  stb_sumsq: out[0] = sum g^2 (scratch of
 * stb_sumsq_scratch_bytes() bytes whose first 4 bytes start at zero).
 * Reference: scaletorch/trainer/model_builder.py:119-134, train_step.py:122-136. */
int stb_adamw(void* p, int p_is_fp32, const float* g, float* m, float* v, long n, float lr,
              float beta1, float beta2, float eps, float weight_decay, int step,
              const float* clip_coef, float grad_scale, cudaStream_t stream);
/* graph-replayable variant: hyper = {lr, 1 - beta1^step, 1 - beta2^step} in device memory */
int stb_adamw_dyn(void* p, int p_is_fp32, const float* g, float* m, float* v, long n, float beta1, float beta2,
                  float eps, float weight_decay, const float* clip_coef, float grad_scale, const float* hyper,
                  cudaStream_t stream);
long stb_sumsq_scratch_bytes(void);
int stb_sumsq(const float* g, long n, void* scratch, float* out, cudaStream_t stream);
int stb_clip_coef(const float* sumsq, float max_norm, float* coef, float* norm_out,
                  cudaStream_t stream);
int stb_cast_scale_f32_bf16(const float* src, void* dst, long n, float scale, cudaStream_t stream);

/* ---- embedding, csrc/loss_optim.cu -----------------------------------------
 * Reference: scaletorch/models/model_qwen3.py:109-110,
 * scaletorch/parallel/tensor_parallel/tensor_parallel.py:483-502. */
int stb_embedding_fwd(const long* ids, const void* W, void* out, long T, int H, int vocab_start,
                      int vocab_local, cudaStream_t stream);
int stb_embedding_bwd(const long* ids, const void* dout, float* dW, long T, int H, int vocab_start,
                      int vocab_local, cudaStream_t stream);

/* ---- flash attention (tcgen05, head_dim 128, GQA, causal), csrc/attention_fwd.cu / attention_bwd.cu ------------
 * Tensors are [B, S, H, D] views described by element strides (sb, ss, sh); D is contiguous. lse / delta are
 * [B, Hq, Sq] fp32. The backward runs three kernels (delta, dK/dV, dQ) and writes dq/dk/dv through their own strides
 * (so a packed dqkv buffer can be filled in place).
 * Reference: F.scaled_dot_product_attention in scaletorch/models/attention_utils.py:130-152 and the eager per-block
 * ring attention in scaletorch/parallel/context_parallel/context_parallel.py:266-364. */
int stb_flash_attn_fwd(const void* q, const void* k, const void* v, void* out, float* lse, int B, int Sq, int Sk,
                       int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh,
                       long v_sb, long v_ss, long v_sh, float scale, int causal, cudaStream_t stream);
int stb_flash_attn_bwd(const void* dout, const void* q, const void* k, const void* v, const void* out,
                       const float* lse, float* delta, void* dq, void* dk, void* dv, int B, int Sq, int Sk, int Hq,
                       int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh, long v_sb,
                       long v_ss, long v_sh, long dq_sb, long dq_ss, long dq_sh, long dk_sb, long dk_ss, long dk_sh,
                       long dv_sb, long dv_ss, long dv_sh, float scale, int causal, cudaStream_t stream);

/* ---- collectives over NVLink peer memory, csrc/comm.cu ---------------------------------------------------------
 * `bufs_dev` / `flag_pages_dev`: device arrays of `world` peer base pointers of a symmetric data buffer / of the 4 KiB
 * flag page (stb_symm_flag_page_bytes) of every rank. Flags are monotonically increasing epochs per `slot` (< 8):
 * the caller passes the next unused epoch; the comment of each function says how many it consumes. Grids are sized
 * <= SM count (`sm_budget`, 0 = all) because blocks wait on other GPUs.
 * Reference: dist.all_reduce / all_gather / reduce_scatter through c10d-NCCL (scaletorch/parallel/data_parallel/
 * bucket.py:73-78, tensor_parallel/tp_comms.py:144-147, sequence_parallel/sp_comms.py:46-47,79-82). */
int stb_symm_flag_page_bytes(void);
int stb_symm_barrier(const void* flag_pages_dev, int rank, int world, int slot, unsigned epoch,
                     cudaStream_t stream);                                                   /* 1 epoch  */
int stb_allreduce_f32(const void* bufs_dev, void* multicast_ptr, const void* flag_pages_dev, long offset_elems,
                      long n, float scale, int rank, int world, int slot, unsigned epoch, int sm_budget,
                      cudaStream_t stream);                                                  /* 2 epochs; in place, x scale */
int stb_allreduce_bf16(const void* bufs_dev, void* multicast_ptr, const void* flag_pages_dev, long offset_elems,
                       long n, int rank, int world, int slot, unsigned epoch, int sm_budget,
                       cudaStream_t stream);                                                 /* 2 epochs */
int stb_allreduce_oneshot_bf16(const void* bufs_dev, const void* flag_pages_dev, long offset_elems, void* out,
                               long n, int rank, int world, int slot, unsigned epoch, int sm_budget,
                               cudaStream_t stream);                                         /* 2 epochs; out is local */
int stb_allgather_pull(const void* bufs_dev, const void* flag_pages_dev, long offset_bytes, void* dst,
                       long rows_per_rank, long row_bytes, int flag_rows, void* ready, int rank, int world, int slot,
                       unsigned epoch, int sm_budget, cudaStream_t stream);                  /* 2 epochs */

/* embedding gradient exchanged as token rows instead of all-reducing the [V, H] matrix; 2 epochs */
int stb_embedding_bwd_peers(const void* bufs_dev, const void* flag_pages_dev, long ids_off, long dout_off, float* dW,
                            long T, int H, int vocab_start, int vocab_local, float scale, int rank, int world,
                            int slot, unsigned epoch, int sm_budget, cudaStream_t stream);

/* ---- GEMM fused with its collective, csrc/gemm_tcgen05.cu -------------------------------------------------------
 * stb_gemm_rs_bf16: C = A . op(B) whose epilogue TMA-reduce-adds row block d of the product into `c_dest[d]` (rank d's
 *   zeroed [M / world, N] bf16 buffer, peer-mapped) -> GEMM + reduce-scatter in one kernel; tiles are visited
 *   starting at the caller's own block + 1 so the NVLink traffic is spread over the peers.
 * stb_ag_gemm_bf16: `n_copy` CTAs of the grid pull the other ranks' [M / world, K] shards out of their symmetric
 *   windows into A_full and raise one flag per `flag_rows` rows; the GEMM CTAs' TMA producer waits on the flags of the
 *   tile it is about to load (own shard first) -> all-gather + GEMM in one kernel. Consumes 2 epochs.
 * Reference: ColumnParallelLinear / RowParallelLinear + separate NCCL calls (tensor_parallel.py:196-362). */
int stb_gemm_rs_bf16(const void* A, const void* B, void* const* c_dest, int M, int N, int K, int lda, int ldb,
                     int ldc, int b_mn, int rank, int world, cudaStream_t stream);
int stb_ag_gemm_bf16(void* A_full, const void* B, void* C, int M, int N, int K, int ldb, int ldc, int b_mn,
                     const void* bufs_dev, const void* flag_pages_dev, long shard_offset_bytes, void* ready_flags,
                     unsigned epoch, int flag_rows, int n_copy, int rank, int world, int slot, cudaStream_t stream);

/* ---- MoE: ragged grouped GEMM, csrc/grouped_gemm.cu -------------------------------------------------------------
 * Rows are expert-major with every expert segment padded to 128 rows (padding rows zero). `offs_pad` [E + 1] and
 * `tile_expert` [R_max / 128] (-1 = unused tile) live on the device: no host synchronisation.
 *   mode 0 (fprop)  C[R, N]   = A[R, K]  . B[e][N, K]^T        mode 1 (dgrad)  C[R, K] = A[R, N] . B[e][N, K]
 *   mode 2 (wgrad)  C[e][N,K] (+)= A[rows of e, N]^T . B[rows of e, K]   (fp32 or bf16 C, optional accumulate)
 * Reference: per-expert Python loop, scaletorch/models/model_qwen3_moe.py:129-171. */
int stb_grouped_gemm_bf16(int mode, const void* A, const void* B, void* C, const int* tile_expert,
                          const int* offs_pad, int R_max, int N, int K, int E, int out_fp32, int accumulate,
                          cudaStream_t stream);

/* ---- MoE: expert-parallel dispatch / combine over peer memory, csrc/ep_dispatch.cu ------------------------------
 * Reference: scaletorch/parallel/expert_parallel/ep_comms.py:30-171 (all_to_all_single x5, two host syncs). */
int stb_ep_exchange_counts(const void* bufs_dev, const void* flag_pages_dev, long table_off, const int* counts,
                           int E, int* out, int rank, int world, int slot, unsigned epoch,
                           cudaStream_t stream);                                             /* 1 epoch  */
int stb_ep_push(const void* bufs_dev, const void* flag_pages_dev, long region_off, const void* src,
                const int* dst_rank, const int* dst_row, const float* scale, long P, int k, int H, int rank,
                int world, int slot, unsigned epoch, int sm_budget, cudaStream_t stream);    /* 2 epochs */
int stb_ep_pull(const void* bufs_dev, const void* flag_pages_dev, long region_off, void* out, void* rows_out,
                const int* src_rank, const int* src_row, const float* weights, long T, int k, int H, int rank,
                int world, int slot, unsigned epoch, int sm_budget, cudaStream_t stream);    /* 2 epochs */

/* ---- round-2 entry points ---------------------------------------------------------------------------------------- */

/* CTA-pair GEMM (csrc/gemm_pair.cu): cluster (2,1,1), tcgen05.mma.cta_group::2 with a 256-row UMMA tile, each CTA
 * TMA-loads half of B, accumulators in both CTAs' TMEM, split-K by TMA reduce-add.  bn = 128 / 256 / 0 (cost model),
 * splits = 0 (cost model) or the number of K partitions.  stb_gemm_bf16 dispatches here when its cost model says so
 * (SCALETORCH_B200_GEMM_PAIR=0/1/2: never / cost model / always). */
int stb_gemm_bf16_2cta(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int a_mn,
                       int b_mn, int out_fp32, int accumulate, int bn, int splits, cudaStream_t stream);

/* Attention on a (query block, key block) pair of a longer sequence -- the building block of ring attention
 * (csrc/attention_fwd.cu, csrc/attention_bwd.cu).  q_lo / S_tot place the query rows inside the full sequence for the
 * causal mask.  Forward: merge = 0 writes (out, lse, acc); merge = 1 folds this block into the running fp32 `acc` and
 * `lse` (log-sum-exp merge in the epilogue) and rewrites `out`.  Backward: dq / dk / dv are fp32 accumulators; dq_acc = 1
 * adds into dq; dk / dv are always added with red.global.add.v4.f32 at system scope, so they may live in a peer's window.
 * Reference: scaletorch/parallel/context_parallel/context_parallel.py:40-244 (Python LSE merge, P2P per step). */
int stb_flash_attn_fwd_block(const void* q, const void* k, const void* v, void* out, float* lse, float* acc, int merge,
                             int q_lo, int S_tot, int B, int Sq, int Sk, int Hq, int Hkv, int D, long q_sb, long q_ss,
                             long q_sh, long k_sb, long k_ss, long k_sh, long v_sb, long v_ss, long v_sh, float scale,
                             int causal, cudaStream_t stream);
int stb_flash_attn_bwd_block(const void* dout, const void* q, const void* k, const void* v, const void* out,
                             const float* lse, float* delta, float* dq, float* dk, float* dv, int B, int Sq, int Sk,
                             int Hq, int Hkv, int D, long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh,
                             long v_sb, long v_ss, long v_sh, long dq_sb, long dq_ss, long dq_sh, long dk_sb, long dk_ss,
                             long dk_sh, long dv_sb, long dv_ss, long dv_sh, float scale, int causal, int q_lo,
                             int S_tot, int compute_delta, int dq_acc, cudaStream_t stream);
/* delta[b, h, q] = rowsum(dO * O), contiguous [B, Sq, Hq, 128] operands. */
int stb_attn_bwd_delta(const void* dout, const void* out, float* delta, int B, int Sq, int Hq, cudaStream_t stream);
/* One-kernel backward (csrc/attention_bwd_fused.cu): S / dP and the exp pass are computed once per tile pair and feed
 * dV, dK and dQ; dQ tiles are TMA reduce-added into the fp32 `dq`, which the caller zeroes beforehand. */
int stb_flash_attn_bwd_fused(const void* dout, const void* q, const void* k, const void* v, const float* lse,
                             const float* delta, float* dq, void* dk, void* dv, int B, int Sq, int Sk, int Hq, int Hkv,
                             int D, long q_sb, long q_ss, long q_sh, long k_sb, long k_ss, long k_sh, long v_sb,
                             long v_ss, long v_sh, long dq_sb, long dq_ss, long dq_sh, long dk_sb, long dk_ss,
                             long dk_sh, long dv_sb, long dv_ss, long dv_sh, float scale, int causal,
                             cudaStream_t stream);
/* QK-norm + RoPE backward that takes dQ as the fp32 tensor written by stb_flash_attn_bwd_fused (no conversion pass). */
int stb_qknorm_rope_bwd_f32q(const void* dqk, const float* dq32, const void* qkv, const void* wq, const void* wk,
                             const float* cos_t, const float* sin_t, const int* pos, const float* rstd, void* dqkv,
                             float* dwq, float* dwk, int T, int S, int Hq, int Hkv, int head_dim, int pos_offset,
                             int use_norm, cudaStream_t stream);

/* Vocabulary-chunked cross-entropy statistics: merge = 0 starts, merge = 1 folds chunk [vocab_start, vocab_start + Vc)
 * into the running (max, sum-exp, target logit) rows (csrc/loss_optim.cu). */
int stb_ce_stats_chunk(const void* logits, const long* targets, float* stats, int rows, int Vc, long ld,
                       int vocab_start, int merge, cudaStream_t stream);

/* ZeRO-1 over symmetric memory (csrc/comm.cu).  Reference: torch DDP all-reduce + replicated AdamW,
 * the modules under scaletorch/parallel/data_parallel and scaletorch/trainer.
 *   stb_reduce_scatter_f32   each rank reduces the [own_lo, own_hi) slice of bucket [offset, offset + n) from all
 *                            peers' fp32 gradient windows (multimem.ld_reduce when multicast_ptr != NULL, peer loads
 *                            otherwise) into its own window, scaled.                                   2 epochs
 *   stb_allreduce_tiny_f32   <= 32 floats (sum of squares partials), one warp.                       1 epoch
 *   stb_adamw_shard_bcast    AdamW on the owned fp32 shard (g, m, v local) and broadcast of the new bf16 parameters
 *                            into every rank's parameter window (multimem.st or peer stores); clip_coef / hyper =
 *                            {lr, bias-correction 1, bias-correction 2} are read from device memory.   2 epochs
 *   stb_param_bcast          owner -> all broadcast of a parameter range (deferred gather).            2 epochs */
int stb_reduce_scatter_f32(const void* bufs_dev, const void* multicast_ptr, const void* flag_pages_dev,
                           long offset_elems, long n, long own_lo, long own_hi, float scale, int rank, int world,
                           int slot, unsigned epoch, int sm_budget, cudaStream_t stream);
int stb_allreduce_tiny_f32(const void* bufs_dev, const void* flag_pages_dev, long offset_elems, int n, int rank,
                           int world, int slot, unsigned epoch, cudaStream_t stream);
int stb_adamw_shard_bcast(const void* pbufs_dev, void* p_multicast_ptr, const void* flag_pages_dev, long p_offset_elems,
                          const float* g, float* m, float* v, long n, float beta1, float beta2, float eps,
                          float weight_decay, const float* clip_coef, float grad_scale, const float* hyper, int rank,
                          int world, int slot, unsigned epoch, int sm_budget, cudaStream_t stream);
int stb_param_bcast(const void* pbufs_dev, void* p_multicast_ptr, const void* flag_pages_dev, long p_offset_elems,
                    long n, int owner, int rank, int world, int slot, unsigned epoch, int sm_budget,
                    cudaStream_t stream);

/* MoE routing and the expert-parallel plan on the device (csrc/moe_router.cu): no library sort / scan, no host sync.
 * Reference: scaletorch/models/model_qwen3_moe.py:30-171, scaletorch/parallel/expert_parallel/ep_comms.py:96-102.
 *   router_fwd   logits [T, E] bf16 -> top-k ids, softmax(top-k) weights, per-expert counts, position of each pair in
 *                its expert's segment, per-expert sum of full-softmax probabilities (aux loss).  E <= 256, k <= 16.
 *   router_bwd   d(weights), d(aux) -> d(logits).
 *   ep_plan      [ep, E] count table -> destination (rank, row) of every pair in the receivers' 128-row-padded regions,
 *                this rank's padded offsets / tile->expert table, number of pairs beyond rows_cap (dropped).
 *   ep_zero_padding   zeroes only the padding rows of a receive region;  copy_rows_dyn  copies *n_rows rows. */
int stb_moe_router_fwd(const void* logits, long T, int E, int k, float* weights, int* idx, int* pos, int* counts,
                       float* psum, cudaStream_t stream);
int stb_moe_router_bwd(const void* logits, const float* weights, const int* idx, const float* dweights,
                       const int* counts, const float* g_aux, float aux_scale, long T, int E, int k, void* dlogits,
                       cudaStream_t stream);
int stb_ep_plan(const int* table, const int* flat, const int* pos, long P, int E, int ep, int rank, int rows_cap,
                int* dst_row, int* dst_rank, int* dropped, int* offs_pad, int* tokens_per_expert, int* tile_expert,
                int n_tiles, cudaStream_t stream);
int stb_ep_zero_padding(void* region, const int* offs_pad, const int* tokens_per_expert, int e_local, int H,
                        int rows_cap, cudaStream_t stream);
int stb_copy_rows_dyn(const void* src, void* dst, const int* n_rows, long row_bytes, int max_rows, cudaStream_t stream);

/* ---- host runtime: native micro-batch producer, csrc/runtime/token_loader.cpp -----------------------------------
 * Worker threads cut (input, target) windows out of a flat int32 token array, apply the context-parallel slice
 * (contiguous or zig-zag) and fill a caller-owned ring of n_slots int64 [2][mbs][S / cp] host buffers (pin them).
 * Sampling: sample w = tokens[w (S+1) : (w+1)(S+1)], per-epoch Fisher-Yates permutation from splitmix64(seed + epoch),
 * data rank r takes perm[r], perm[r + world], ...  One consumer thread: next() blocks for the next batch in order and
 * returns its slot, release() hands the slot back once its H2D copy has completed.
 * Reference: torch DataLoader + DistributedSampler + Python collate, scaletorch/data/dataloader.py:107-292. */
void* stb_loader_create(const int* tokens, long n_tokens, int seq_len, int mbs, int cp_rank, int cp_size, int zigzag,
                        int data_rank, int data_world, int shuffle, unsigned long long seed, int n_slots,
                        long long* ring, int n_threads);
long stb_loader_batches_per_epoch(void* handle);
long stb_loader_num_samples(void* handle);
int stb_loader_next(void* handle, long* batch_out);
int stb_loader_release(void* handle, int slot);
void stb_loader_destroy(void* handle);

#ifdef __cplusplus
}
#endif
#endif /* SCALETORCH_B200_H_ */
