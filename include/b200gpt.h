/* C ABI of libb200gpt.so - hand-written sm_100a kernels for GPT training on B200.
 *
 * Conventions: every function returns 0 on success, a negative code on failure
 * (b200_last_error() has the message); pointers are device pointers unless stated; `stream`
 * is a cudaStream_t; bf16 tensors are passed as void*.  No entry point allocates device
 * memory, so all of them may be captured in a CUDA graph once warmed up.
 *
 * Each declaration cites the reference call site it serves (/root/reference).
 */
#ifndef B200GPT_H_
#define B200GPT_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* b200_last_error(void);
/* sticky device-side error word (bounded waits that timed out); host-synchronous */
int b200_device_error(int* out, int clear);
int b200_smoke(void* stream);

/* token + position embedding, fp32 out [n_tokens, dim]  (models/gpt.py:180-185) */
int b200_embedding_fwd(const int64_t* ids, const int64_t* pos, const float* tok_w, const float* pos_w,
                       float* out, int n_tokens, int dim, int vocab, int max_pos, void* stream);
/* scatter-add of dout into both tables (accumulates)  (autograd of models/gpt.py:185) */
int b200_embedding_bwd(const int64_t* ids, const int64_t* pos, const float* dout, float* dtok,
                       float* dpos, int n_tokens, int dim, int vocab, int max_pos, void* stream);

/* LayerNorm fp32 -> bf16, saves mean/rstd  (models/gpt.py:119,122,217) */
int b200_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* y_bf16, float* mean,
                       float* rstd, int rows, int dim, float eps, void* stream);
/* dx = LN'(dy) [+ dres]; optional bf16 copy of dx; dgamma/dbeta accumulate */
/* scratch: [16][2*dim] floats + one int counter, both zero on entry and left zero on exit; launches
 * sharing a scratch buffer must not overlap on the GPU */
int b200_layernorm_bwd(const void* dy_bf16, const float* x, const float* gamma, const float* mean,
                       const float* rstd, const float* dres, float* dx, void* dx_bf16, float* dgamma,
                       float* dbeta, int rows, int dim, float* scratch, int* counter, void* stream);
/* ... and, with fuse != 0, the first step of the backward of the Linear / MLP whose residual epilogue produced x
 * (models/gpt.py:33-41,105-123): the bf16 output becomes dz = dx * act'(mask_src) (mask_src NULL: plain copy; act 1:
 * mask_src is the ReLU output, 2 / 3: the GELU pre-activation) and dbias_z += column sums of dz (NULL: no bias).
 * scratch: [16][3*dim] floats. */
int b200_layernorm_bwd_fused(const void* dy_bf16, const float* x, const float* gamma, const float* mean,
                             const float* rstd, const float* dres, float* dx, void* dz_bf16, float* dgamma,
                             float* dbeta, int rows, int dim, float* scratch, int* counter, const void* mask_src,
                             int act, float* dbias_z, int fuse, void* stream);

/* fused AdamW over a flat arena; hyper = {lr, step, skipped, -} in device memory  (main-single.py:42,99-101).
 * Non-finite gradient elements are skipped one by one (parameter and moments keep their values) and counted in
 * hyper[2]; the reference's GradScaler instead drops the whole step after a host sync - a deliberate deviation. */
int b200_adamw(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n,
               const float* hyper, float beta1, float beta2, float eps, float wd, void* stream);
int b200_adamw_tick(float* hyper, void* stream);
int b200_cast_bf16(const float* in, void* out_bf16, int64_t n, void* stream);

/* dz = dy * (y_act > 0) as bf16 (y_act NULL: plain cast); dbias += colsum(dz)  (models/gpt.py:35,38) */
/* scratch: [16][cols] floats + one int counter (same contract as b200_layernorm_bwd) */
int b200_relu_bwd(const void* dy, const void* y_act_bf16, void* dz_bf16, float* dbias, int rows, int cols,
                  int dy_is_fp32, float* scratch, int* counter, void* stream);
/* same with the activation selectable: act 1 ReLU (y_or_z = activated output), 2 / 3 GELU erf / tanh form
 * (y_or_z = pre-activation)  (models/gpt.py:14-15: activation is a constructor argument) */
int b200_act_bwd(const void* dy, const void* y_or_z_bf16, void* dz_bf16, float* dbias, int rows, int cols,
                 int dy_is_fp32, int act, float* scratch, int* counter, void* stream);
int b200_colsum_bf16(const void* x_bf16, float* out_accum, int rows, int cols, int ld, float* scratch,
                     int* counter, void* stream);
/* No counterpart in the reference (an addition of this library).  Same result as b200_relu_bwd, for
 * callers that cannot provide scratch memory: a host program whose device memory is entirely owned by
 * its own pool or by another framework, or a call site (a callback, a latency-critical loop) where the
 * CUDA allocator must not be entered.  The scratch is a 256 KiB array built into the library - every
 * GPU that loads the library carries it until the process exits - so cols is limited to 4096; wider
 * inputs are refused with an error.  Nothing is allocated or freed by this call.  The first call on a
 * device also looks up the array's address on that device; make that first call outside any stream
 * capture.  All calls on a device share the one array: two calls that overlap on the GPU (different
 * streams, no ordering between them) corrupt each other's sums - order them. */
int b200_relu_bwd_static(const void* dy, const void* y_act_bf16, void* dz_bf16, float* dbias, int rows, int cols,
                         int dy_is_fp32, void* stream);

/* cross entropy over a bf16 logits chunk, in place  (main-single.py:95-96,127-131)
 * stats[0] = n_valid (input, from b200_count_valid), stats[1] += loss sum, stats[2] += #correct */
int b200_count_valid(const int64_t* targets, float* stats, int ignore_index, int64_t n, void* stream);
int b200_ce_fwd_bwd(void* logits_bf16, const int64_t* targets, float* stats, int rows, int vocab, int ld,
                    int ignore_index, int write_grad, float grad_scale, void* stream);

/* Fused LM head + cross entropy (models/gpt.py:219,231 + main-single.py:95-96,127-131).  The logits are never
 * written: the logits GEMM's epilogue stores E = exp(logit - shift[m]) (bf16) with a shift fixed before the GEMM
 * (upper bound |x||w| of the row's logits minus an observed gap), and leaves per-row statistics; `finish` derives
 * loss / accuracy and prepares the two gradient GEMMs, recomputing exactly any row the shift estimate missed.
 *   b200_lmhead_begin   stats[0] = #targets != ignore_index, stats[1..3] = 0, rotate lmstate
 *   b200_lmhead_norms   xnorm[m] = |X[m,:]|, lmstate[0] = max_v |W[v,:]|
 *   b200_lmhead_gemm_bf16  E (NULL: statistics only), part[span][M] = (sum E, max logit), tgt_val[m]
 *   b200_lmhead_finish  stats[1] += loss sum, stats[2] += #correct, rscale, E[m,target] -= sum_m, Xs = rscale * X
 *   then  dX = b200_gemm_rowscale_bf16(E, W, rscale)   and   dW += E^T . Xs  (b200_gemm_bf16)
 * lmstate: 4 floats that live with the weight (zero once, then leave alone): wmax, gap in use, gap observed,
 * rows recomputed by the last call.  part: max_parts x M float2 with max_parts >= ceil(V / 128) + 2. */
int b200_lmhead_begin(const int64_t* targets, float* stats, float* lmstate, int ignore_index, int64_t n,
                      void* stream);
int b200_lmhead_norms(const void* X_bf16, int ldx, const void* W_bf16, int ldw, int M, int V, int D, float* xnorm,
                      float* lmstate, void* stream);
int b200_lmhead_gemm_bf16(const void* X, const void* W, void* E_out, int M, int V, int K, int ldx, int ldw, int lde,
                          const float* xnorm, const float* lmstate, void* part, int max_parts, int* num_parts,
                          const int64_t* targets, float* tgt_val, int cfg, void* stream);
int b200_lmhead_finish(const void* part, int num_parts, int M, const float* xnorm, const float* tgt_val,
                       const int64_t* targets, int ignore_index, float* stats, float* lmstate, void* E, int lde,
                       int V, const void* X, int ldx, int D, const void* W, int ldw, void* Xs, float* rscale,
                       float grad_scale, int write_grad, void* stream);
/* C (bf16) = diag(row_scale) . (A . op(B)) */
int b200_gemm_rowscale_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                            int b_mn_major, const float* row_scale, int cfg, void* stream);

/* Masked dgrad GEMM with the bias gradient fused into its epilogue (MLP backward of reference
 * models/gpt.py:33-41): C (bf16) = (A . B) * (mask_src > 0), colsum[n] += sum over rows of C.  colsum is
 * accumulated with fp32 reductions, like every other gradient of this library. */
int b200_gemm_mask_colsum_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                               int b_mn_major, const void* mask_src, float* colsum, int cfg, void* stream);

/* same, activation selectable: act 0 / 1 ReLU mask (mask_src = activated output), 2 / 3 GELU (mask_src = pre-activation,
 * C = (A . B) * gelu'(mask_src)); colsum may be NULL */
int b200_gemm_dact_colsum_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                               int b_mn_major, const void* mask_src, float* colsum, int act, int cfg, void* stream);

/* C[M,N] (+)= A[M,K] . B[N,K]^T on tcgen05 with a fused epilogue  (models/gpt.py:29-38,60-64,102,219)
 * a/b_mn_major: 0 = operand stored [rows][K]; 1 = stored [K][rows] (dgrad / wgrad, no transposes).
 * epilogue order: +bias -> act (1 ReLU, 2 GELU erf form, 3 GELU tanh form) -> *(mask_src>0) -> +residual(fp32) -> store;
 * aux_out (bf16) receives the activated value for ReLU and the PRE-activation for GELU (what each backward needs). */
int b200_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                   int ldc, int a_mn_major, int b_mn_major, const float* bias, const float* residual,
                   int ldr, int act, int out_fp32, int accumulate, void* aux_out_bf16,
                   const void* mask_src_bf16, int cfg, void* stream);
/* Weight gradients of up to 4 Linear layers in ONE launch: C_i[M_i,N_i] (fp32) += A_i^T . B_i with A_i stored [K][M_i]
 * and B_i stored [K][N_i] (dZ and X as they sit in memory, K = tokens; main-single.py:97 loss.backward()).  The
 * problems share one persistent grid of CTA pairs and one split of K.  -100: CTA pairs are off, launch one by one. */
int b200_gemm_wgrad_group(int n, const void* const* A, const void* const* B, void* const* C, const int* M,
                          const int* N, int K, const int* lda, const int* ldb, const int* ldc, void* stream);
/* GEMM with the FSDP all-gather of its weight fused in (pre-forward unshard of main-fsdp.py:131, folded
 * into the first consuming GEMM): the kernel pulls bytes [ag_off, ag_off+ag_bytes) of a flat, evenly
 * sharded parameter group from the owning peers' shard buffers over NVLink into full_base, then
 * computes C = A . W^T (W = B = full_base + ag_off, K-major).  peers_dev: device array [world] of shard
 * base pointers; ag_sync: 2 + ag_sync_tiles ints of device memory (>= N/256 tiles), zero on entry and exit. */
int b200_ag_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                      const float* bias, const float* residual, int ldr, int act, int out_fp32, void* aux_out_bf16,
                      const void* peers_dev, void* full_base, int64_t ag_off, int64_t ag_bytes,
                      int64_t shard_bytes, int* ag_sync, int ag_sync_tiles, void* stream);
int b200_gemm_debug(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb,
                    int ldc, int a_mn_major, int b_mn_major, int a_lbo, int a_sbo, int a_kstep,
                    int b_lbo, int b_sbo, int b_kstep, int force_bn, int force_splits, void* stream);

/* causal flash attention over packed qkv[B,T,3,H,D] (D in 32/64/128)  (models/gpt.py:75-100)
 * key_mask: optional uint8 [B,T], 1 = padding key.  lse = log-sum-exp of the scaled scores. */
int b200_attn_fwd(const void* qkv_bf16, void* out_bf16, float* lse, int B, int T, int H, int D, float scale,
                  const void* key_mask, void* stream);
/* same, forward kernel chosen by the caller: ws = 0 per-tile kernel (short sequences), 1 warp-specialised TMA-fed
 * pipelined kernel (long sequences), -1 automatic (what b200_attn_fwd does) */
int b200_attn_fwd_select(const void* qkv_bf16, void* out_bf16, float* lse, int B, int T, int H, int D, float scale,
                         const void* key_mask, int ws, void* stream);
/* scratch (caller-provided, overwritten): delta fp32 [B,H,T], dq_acc fp32 [B,T,H*D] */
int b200_attn_bwd(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                  void* dqkv_bf16, float* delta, float* dq_acc, int B, int T, int H, int D, float scale,
                  const void* key_mask, void* stream);

/* same, kernels chosen by the caller: mode 0 per-tile kernel, 1 warp-specialised (dQ by fp32 vector reductions into
 * dq_acc), 2 warp-specialised dK / dV kernel + query-owned dQ kernel (no atomics, dq_acc unused), -1 automatic */
int b200_attn_bwd_select(const void* qkv_bf16, const void* out_bf16, const void* dout_bf16, const float* lse,
                         void* dqkv_bf16, float* delta, float* dq_acc, int B, int T, int H, int D, float scale,
                         const void* key_mask, int mode, void* stream);

/* ---- NVLink / NVSwitch collectives over symmetric memory (no NCCL) --------------------------
 * bufs_dev : device array [world] of every rank's base pointer of the symmetric data buffer
 * mc_ptr   : NVLS multicast mapping of that buffer (NULL -> peer load/store fallback)
 * flags_dev: device array [world] of every rank's flag buffer (256 channels x 16 uint32, zeroed)
 * kernels that may run concurrently must use disjoint channel ranges [channel_base, +nblocks). */

/* in-place bucket all-reduce (sum * scale) of elements [offset, offset+n) of the fp32 gradient
 * arena  (the DDP reducer behind main-ddp.py:55,124) */
int b200_allreduce(const void* bufs_dev, void* mc_ptr, const void* flags_dev, int rank, int world,
                   int64_t offset_elems, int64_t n_elems, float scale, int nblocks, int channel_base,
                   void* stream);
/* same reduction with bf16 on the wire and fp32 accumulation in the switch: grad is the LOCAL fp32 arena, stage a
 * symmetric bf16 buffer of the same element geometry (needs its multicast mapping)  (main-ddp.py:55,124) */
int b200_allreduce_bf16(float* grad, const void* stage_bufs_dev, void* stage_mc_ptr, const void* flags_dev, int rank,
                        int world, int64_t offset_elems, int64_t n_elems, float scale, int nblocks, int channel_base,
                        void* stream);
/* reduce-scatter of this rank's shard fused with the sharded AdamW update and the bf16 shard
 * write  (FSDP post-backward reduce-scatter + optimizer step, main-fsdp.py:138-140) */
int b200_rs_adamw(const void* grad_bufs_dev, void* grad_mc_ptr, const void* flags_dev, int rank, int world,
                  int64_t shard_off, int64_t shard_n, float scale, float* p, float* m, float* v,
                  void* p_bf16, const float* hyper, float beta1, float beta2, float eps, float wd,
                  int end_barrier, int nblocks, int channel_base, void* stream);
/* reduce-scatter alone: out[i] = scale * sum_ranks grad[shard_off + i], for a small grid (<= 16 blocks; the
 * switch adds), followed by b200_adamw on `out` as an ordinary kernel  (main-fsdp.py:138-140) */
int b200_reduce_scatter(const void* grad_bufs_dev, void* grad_mc_ptr, const void* flags_dev, int rank, int world,
                        int64_t shard_off, int64_t shard_n, float scale, float* out, int end_barrier, int nblocks,
                        int channel_base, void* stream);
/* pull all-gather: dst[peer * shard_bytes ...] = peer's bytes [src_off, src_off + shard_bytes)
 * (FSDP pre-forward unshard, main-fsdp.py:131) */
int b200_allgather(const void* src_bufs_dev, const void* flags_dev, int rank, int world,
                   int64_t src_off_bytes, int64_t shard_bytes, void* dst, int start_barrier, int nblocks,
                   int channel_base, void* stream);
int b200_barrier(const void* flags_dev, int rank, int world, int channel, void* stream);
/* test aids: busy-wait on the stream; bound of every cross-GPU wait on the current device (default ~20 s) */
int b200_spin(int64_t ns, void* stream);
int b200_set_comm_timeout_ms(int ms);

#ifdef __cplusplus
}
#endif
#endif /* B200GPT_H_ */
