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

#ifdef __cplusplus
}
#endif
#endif /* SCALETORCH_B200_H_ */
