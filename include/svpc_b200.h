/* svpc_b200 - C ABI of the B200-native SVPC hot path (libsvpc_b200.so, hand-written sm_100a CUDA).
 *
 * The reference (awkrail/svpc) has no FFI: its hot path is the Python object protocol of
 * StateAwareRecursiveTransformer (src/rtransformer/model.py:826-1189) and everything below it is a stock
 * ATen/cuBLAS/cuDNN call.  Each entry point here replaces the library calls of one piece of that path; the
 * reference lines it stands for are cited per function.  The reference-side binding is svpc_b200/cuda_backend.py
 * (ctypes) - see INTEGRATION.md.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller (torch); nothing is allocated or freed here;
 *   - no entry point synchronises; work is enqueued on `stream` (a cudaStream_t);
 *   - return 0 (SVPC_OK) or a negative SVPC_ERR_*; svpc_last_error() gives the message (thread local);
 *   - activations are fp32 (SVPC_F32) or bf16 (SVPC_BF16), selected by a dtype code; parameters, statistics,
 *     probabilities and losses are always fp32; ids / offsets are int32;
 *   - "ragged" arrays are concatenations over videos / clips described by offset arrays (svpc_b200/model.py StepPlan).
 */
#ifndef SVPC_B200_H
#define SVPC_B200_H
#ifdef __cplusplus
extern "C" {
#endif

#define SVPC_OK 0
#define SVPC_ERR_ARG (-1)
#define SVPC_ERR_CUDA (-2)
#define SVPC_ERR_UNSUPPORTED (-3)

#define SVPC_F32 0
#define SVPC_BF16 1

#define SVPC_ACT_NONE 0
#define SVPC_ACT_RELU 1
#define SVPC_ACT_GELU 2     /* erf GELU, model.py:58-64 */
#define SVPC_ACT_SIGMOID 3

#define SVPC_GEMM_AUTO 0
#define SVPC_GEMM_SIMT 1    /* fp32 FFMA tiles (parity path, odd shapes) */
#define SVPC_GEMM_TC 2      /* tcgen05 / TMEM / TMA bf16 (fails if the shape / layout is not eligible) */

typedef void* svpc_stream_t; /* cudaStream_t */

const char* svpc_last_error(void);
int svpc_abi_version(void);
/* which kernel svpc_gemm would pick for this problem: SVPC_GEMM_SIMT or SVPC_GEMM_TC */
int svpc_gemm_route(int M, int N, int K, int dtA, long long sa_m, long long sa_k, int dtB, long long sb_n, long long sb_k,
                    int dtD, long long ldd);
/* counters: number of kernels launched by this library since the last reset (bench.py "gpu_launches") */
long long svpc_launch_count(void);
void svpc_reset_launch_count(void);

/* Train-mode dropout (nn.Dropout, model.py:214, 231, 287, 495, 522, 550, 757) is a counter-based mask that every kernel
 * recomputes instead of storing: keep(seed, site, element index), kept elements scaled by 1 / (1 - p).  `drop_seed` is a
 * DEVICE pointer to the step seed (bumped by the caller once per forward, so a captured CUDA graph draws fresh masks on
 * every replay); `drop_site` distinguishes the call sites of one step; drop_p == 0 or a NULL seed disables it.
 * svpc_dropout_mask writes the scale (0 or 1/(1-p)) of elements [0, n) - used by the tests to state the contract. */
int svpc_dropout_mask(long long n, const unsigned int* drop_seed, unsigned int drop_site, float drop_p, float* out,
                      svpc_stream_t stream);
/* Gumbel(0,1) noise of F.gumbel_softmax (model.py:1018; torch draws -log(Exp(1)) on the device inside the step): element i is
 * -log(-log(u_i)), u_i from the counter hash of (*seed, site, i), so that a replayed CUDA graph draws new noise when the
 * device seed is bumped.  The model uses it when no noise is injected (tests inject the oracle's noise instead). */
int svpc_gumbel_noise(long long n, const unsigned int* seed, unsigned int site, float* out, svpc_stream_t stream);

/* D[M,N] = drop(act(A[M,K] . B[N,K]^T + bias[N])) + residual[M,N]  (+ D when accumulate); optional copy of the
 * pre-activation.  Dropout element index = m * N + n.  A and B are addressed by element strides (one of each pair must be 1).  Replaces every
 * nn.Linear / matmul on the path: model.py:195-197,230,263,286,496,551,706-708,738,797-811,899,939-940 and their
 * autograd dgrad / wgrad.  gref != NULL (dtype of D, row pitch ldg): the result is finally multiplied by gact'(gref[m,n])
 * - an input-gradient GEMM then writes the gradient of the PREVIOUS Linear's pre-activation directly (gref = that
 * pre-activation for GELU, its output for RELU / SIGMOID), residual = the gradient arriving over a skip connection. */
int svpc_gemm(int M, int N, int K,
              const void* A, int dtA, long long sa_m, long long sa_k,
              const void* B, int dtB, long long sb_n, long long sb_k,
              void* D, int dtD, long long ldd,
              const float* bias, int act,
              const void* residual, long long ldr,
              void* pre, long long ldp,
              int accumulate, int route,
              const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
              const void* gref, long long ldg, int gact, svpc_stream_t stream);

/* The same contraction for fp32 operands on the tensor cores (the 1e-4 parity path of model.py:195-197, 230, 263, 286, 551,
 * 706-708, 738): every fp32 operand is split into two bf16 planes, x = hi + lo with hi = bf16(x), lo = bf16(x - hi)
 * (svpc_split_bf16, row pitch `ldd` elements, a multiple of 8), and
 *     D = A_hi.B_hi^T + A_lo.B_hi^T + A_hi.B_lo^T        (fp32 accumulation in one TMEM accumulator; A_lo.B_lo^T ~ 2^-18 dropped)
 * Epilogue arguments as svpc_gemm.  svpc_gemm_split_eligible: 1 when the tcgen05 kernel can take the problem (M >= 64, N >= 16,
 * K >= 32, one unit stride per operand and the other a multiple of 8 elements); otherwise run svpc_gemm on the fp32 data. */
int svpc_split_bf16(int rows, int cols, const float* src, long long lds, void* hi, void* lo, long long ldd, svpc_stream_t stream);
int svpc_gemm_split_eligible(int M, int N, int K, long long sa_m, long long sa_k, long long sb_n, long long sb_k);
int svpc_gemm_split(int M, int N, int K,
                    const void* A_hi, const void* A_lo, long long sa_m, long long sa_k,
                    const void* B_hi, const void* B_lo, long long sb_n, long long sb_k,
                    void* D, int dtD, long long ldd,
                    const float* bias, int act,
                    const void* residual, long long ldr,
                    void* pre, long long ldp,
                    int accumulate,
                    const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
                    const void* gref, long long ldg, int gact, svpc_stream_t stream);

/* dz = dy * dropscale(r * cols + c) * act'(ref), dbias[c] = sum_r dz[r,c].  ref = output for RELU/SIGMOID,
 * pre-activation for GELU. */
int svpc_act_bwd(int rows, int cols, const void* dy, const void* ref, int dt, int act, void* dz, float* dbias,
                 const unsigned int* drop_seed, unsigned int drop_site, float drop_p, svpc_stream_t stream);
int svpc_colsum(int rows, int cols, const void* x, int dt, float* out, svpc_stream_t stream);
int svpc_cast(long long n, const void* src, int dts, void* dst, int dtd, svpc_stream_t stream);
int svpc_cast2d(int rows, int cols, const void* src, int dts, long long lds, void* dst, int dtd, long long ldd,
                svpc_stream_t stream);
int svpc_sum(long long n, const float* x, float* out, svpc_stream_t stream);

/* BertLayerNorm (model.py:152-156) on x (+ res), optional + pe[row % pe_period] + tt[tt_ids[row]]
 * (model.py:561, 890-891).  pre (= x + res) is written only when res != NULL. */
int svpc_ln_fwd(int rows, int D, const void* x, const void* res, int dt_in, const float* gamma, const float* beta,
                float eps, const float* pe, int pe_period, const float* tt, const int* tt_ids, void* y, int dt_out,
                void* pre, float* mean, float* rstd, svpc_stream_t stream);
/* dz != NULL fuses the activation / dropout backward of the Linear whose output (after act and dropout, plus an optional
 * residual) is the LayerNorm's input (model.py:229-233, 285-289, 496-499, 551-553, 704-709):
 *   dz[r, c] = dpre[r, c] * dropscale(r * D + c) * act'(ref[r, c]),  dbias[c] = sum_r dz[r, c]   (dbias optional)
 * dz / ref have the dtype and layout of dpre / pre; ref NULL with SVPC_ACT_RELU uses the LayerNorm input itself (valid when
 * no residual was added); the drop_* arguments are that Linear's dropout site. */
int svpc_ln_bwd(int rows, int D, const void* dy, int dt_dy, const void* pre, int dt_pre, const float* mean,
                const float* rstd, const float* gamma, void* dpre, float* dgamma, float* dbeta,
                void* dz, int act, const void* ref, float* dbias,
                const unsigned int* drop_seed, unsigned int drop_site, float drop_p, svpc_stream_t stream);
/* out[bin, :] = sum over rows with ids[row] == bin of dy[row, :]  (token-type embedding gradient, model.py:890) */
int svpc_index_colsum(int rows, int D, const void* dy, int dt, const int* ids, int nbins, float* out,
                      svpc_stream_t stream);

/* Embedding gather + LN(W) (model.py:492-494 / 519-521) and its backward (scatter-add into the table).  ldy: row
 * pitch of y in elements (>= W; padded to 16 bytes so that y is a TMA-addressable GEMM operand). */
int svpc_embed_ln_fwd(int rows, int W, const float* table, const int* ids, const float* gamma, const float* beta,
                      float eps, void* y, long long ldy, int dt, float* mean, float* rstd,
                      const unsigned int* drop_seed, unsigned int drop_site, float drop_p, svpc_stream_t stream);
int svpc_embed_ln_bwd(int rows, int W, int V, const void* dy, int dt, const float* table, const int* ids,
                      const float* mean, const float* rstd, const float* gamma, int padding_idx, float* dtable,
                      float* dgamma, float* dbeta, const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
                      svpc_stream_t stream);

/* Gathers the first Lv rows of clip (t,b) = (clip_t[n], clip_b[n]) from the S per-step feature tensors
 * (B, L, F) (dt_feat: SVPC_F32 as the reference's loader delivers them, or SVPC_BF16 - the half-size upload of the bf16
 * path) and writes (x - mean) * rstd (LN(3072) without affine, model.py:549). */
int svpc_video_ln_pack(int n_clips, int B, int L, int Lv, int F, const long long* feat_ptrs, int dt_feat, const int* clip_t,
                       const int* clip_b, float eps, void* xhat, int dt, float* mean, float* rstd, const float* gamma,
                       const float* beta, const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
                       svpc_stream_t stream);
/* train mode (gamma / beta given): xhat <- drop(xhat * gamma + beta), element index row * F + c (model.py:549-550); its
 * backward towards gamma / beta (the features are inputs), xhat recomputed from the features and the statistics */
int svpc_video_ln_affine_bwd(int n_clips, int B, int L, int Lv, int F, const long long* feat_ptrs, int dt_feat, const int* clip_t,
                             const int* clip_b, const float* mean, const float* rstd, const void* dy, int dt,
                             float* dgamma, float* dbeta, const unsigned int* drop_seed, unsigned int drop_site,
                             float drop_p, svpc_stream_t stream);

/* BertSelfAttention core (model.py:204-216) over ragged short sequences; one CTA per (sequence, head).  Dropout on the
 * probabilities (model.py:214): element index ((seq * H + head) * max_lq + i) * lks + j, lks = max_lk rounded up to even. */
int svpc_attn_fwd(int n_seq, int H, int dh, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                  long long ldv, int dt, const int* q_off, const int* q_len, const int* k_off, const int* k_len,
                  const float* key_mask, int causal, int max_lq, int max_lk, void* ctx, long long ldo,
                  const unsigned int* drop_seed, unsigned int drop_site, float drop_p, svpc_stream_t stream);
int svpc_attn_bwd(int n_seq, int H, int dh, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                  long long ldv, int dt, const int* q_off, const int* q_len, const int* k_off, const int* k_len,
                  const float* key_mask, int causal, int max_lq, int max_lk, const void* dctx, long long lddo, void* dq,
                  long long lddq, void* dk, long long lddk, void* dv, long long lddv, const unsigned int* drop_seed,
                  unsigned int drop_site, float drop_p, svpc_stream_t stream);

/* IngredientPositionEncoding (model.py:116-140): segment means between [SEP]s + pe[j]. */
int svpc_segmean_fwd(int B, int Li, int D, int NI, const void* x, int dt, const int* sep_mask, const int* cu_ingr,
                     const float* pe, float* E0, int* scratch_row_seg, int* scratch_seg_lohi, svpc_stream_t stream);
int svpc_segmean_bwd(int B, int Li, int D, int NI, const float* dE0, const int* sep_mask, const int* cu_ingr, void* dx,
                     int dt, int* scratch_row_seg, int* scratch_seg_lohi, svpc_stream_t stream);
/* svpc_segmean_fwd for callers that cannot allocate (CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).
 * No counterpart in the reference.  Same inputs, outputs and results as
 * svpc_segmean_fwd; the two scratch tables live in a 320 KB file-scope device array inside the library (held by every GPU
 * that loads it until exit) and the call performs no cudaMalloc / cudaFree.  Limits: B * Li <= SVPC_SEGMEAN_STATIC_MAX_ROWS
 * and NI <= SVPC_SEGMEAN_STATIC_MAX_SEGS, beyond which it returns SVPC_ERR_ARG (svpc_last_error names the limit).  All
 * calls share the array: calls that overlap on the GPU (two streams, two host threads) race; serialise them on one stream. */
#define SVPC_SEGMEAN_STATIC_MAX_ROWS 65536
#define SVPC_SEGMEAN_STATIC_MAX_SEGS 8192
int svpc_segmean_fwd_static(int B, int Li, int D, int NI, const void* x, int dt, const int* sep_mask, const int* cu_ingr,
                            const float* pe, float* E0, svpc_stream_t stream);
int svpc_ragged_mean_fwd(int B, int D, const float* x, const int* cu, float* y, svpc_stream_t stream);
int svpc_ragged_mean_bwd(int B, int D, const float* dy, const int* cu, float* dx, svpc_stream_t stream);
int svpc_row_normalize_fwd(int rows, int cols, const float* x, float* y, svpc_stream_t stream);
int svpc_row_normalize_bwd(int rows, int cols, const float* dy, const float* x, float* dx, svpc_stream_t stream);

/* Entity-state recurrence of EntitiyReasoningNetwork.forward (model.py:792-820); one CTA per video.  max_steps: the largest
 * number of clips of any video (sizes the shared-memory staging of the per-step rows; 0 = unknown, un-pipelined kernel). */
int svpc_reasoner_fwd(int B, int D, int max_ni, int max_steps, const float* E0, const int* cu_ingr, const int* cu_steps,
                      const int* clip_ent_off, const float* w2v, const float* clog, const float* w4s, float* ent,
                      float* ebar, float* Es, float* a_saved, svpc_stream_t stream);
int svpc_reasoner_bwd(int B, int D, int max_ni, int max_steps, const float* E0, const int* cu_ingr, const int* cu_steps,
                      const int* clip_ent_off, const float* w2v, const float* clog, const float* w4s, const float* ent,
                      const float* ebar, const float* Es, const float* a_saved, const float* d_ent, const float* d_ebar,
                      const float* d_Es, float* dE0, float* dw2v, float* dclog, float* dw4s, svpc_stream_t stream);

/* pointer_generator_network + LabelSmoothingLoss (model.py:896-923, 37-55); one CTA per clip. */
int svpc_pointer_fwd(int n_clips, int Lt, int D, int V, int max_ni, int max_vx, const void* dec, int dt, const float* G,
                     const float* Esrc, const int* clip_src_off, const int* clip_ni, const int* clip_gi,
                     const int* clip_vx, const int* clip_P_off, const int* tok_ptr, const int* tok_ids,
                     const float* wpgen, const float* bpgen, const float* logits, long long ldl, const int* labels,
                     float smoothing, float* P, float* loss_rows, float* alpha, float* pgen, float* ctx,
                     svpc_stream_t stream);
int svpc_pointer_bwd(int n_clips, int Lt, int D, int V, int max_ni, int max_vx, const void* dec, int dt, const float* G,
                     const float* Esrc, const int* clip_src_off, const int* clip_ni, const int* clip_gi,
                     const int* clip_vx, const int* clip_P_off, const int* tok_ptr, const int* tok_ids,
                     const float* wpgen, const float* logits, long long ldl, const int* labels, float smoothing,
                     const float* P, const float* alpha, const float* pgen, const float* ctx, const float* dP,
                     const float* dloss, void* ddec, float* dG, float* dE, float* du, float* dwpgen, float* dbpgen,
                     float* dlogits, float* dscore_ws /* scratch: n_clips * Lt * max_ni floats */, svpc_stream_t stream);
/* 'video' mode: softmax over the vocabulary + LabelSmoothingLoss (model.py:1011-1014). */
int svpc_softmax_ls_fwd(int rows, int V, const float* logits, long long ldl, const int* labels, float smoothing, float* P,
                        float* loss_rows, svpc_stream_t stream);
int svpc_softmax_ls_bwd(int rows, int V, const float* dP, const float* dloss, const float* P, const int* labels,
                        float smoothing, float* dlogits, long long ldl, svpc_stream_t stream);

/* Textual re-simulator front end (model.py:1018-1021): hard Gumbel-softmax with INJECTED noise g, word-vector
 * lookup, masked mean over the text positions. */
int svpc_gumbel_fwd(int n_clips, int Lt, int V, int W, const float* P, const float* g, const int* clip_P_off,
                    const int* clip_vx, float tau, const float* emb, const float* text_mask, float* xmean, int* idx,
                    float* yval, svpc_stream_t stream);
int svpc_gumbel_bwd(int n_clips, int Lt, int V, int W, const float* P, const float* g, const int* clip_P_off,
                    const int* clip_vx, float tau, const float* emb, const float* text_mask, const int* idx,
                    const float* yval, const float* dxmean, float* dP, float* demb,
                    float* tv_ws /* scratch: n_clips * V floats */, svpc_stream_t stream);

/* Recurrence of nn.LSTM(bidirectional) (model.py:865, 1022-1024), gate order i,f,g,o: ONE launch per time step j for
 * both directions and all videos, recurrent product fused with the cell.  rows (2, Smax, B): clip row of video b at
 * step j of direction dir, or -1.  xg (2, Nc, 4D) = x W_ih^T + b_ih + b_hh; whh (2, 4D, D); hs / cs / hprev (2, Nc, D),
 * gates (2, Nc, 4D) are saved for the backward; out (Nc, D) must be zero-initialised (both directions are summed into it).
 * Backward: call with j = Smax-1 .. 0; whh_t (2, D, 4D) = W_hh^T (svpc_transpose_f32); dc (2, B, D) zero-initialised. */
int svpc_lstm_step_fwd(int B, int D, int Nc, int Smax, int j, const float* xg, const float* whh, const int* rows,
                       float* hs, float* cs, float* gates, float* out, float* hprev, svpc_stream_t stream);
int svpc_lstm_step_bwd(int B, int D, int Nc, int Smax, int j, const float* whh_t, const int* rows, const float* cs,
                       const float* gates, const float* dout, float* dgates, float* dc, svpc_stream_t stream);
int svpc_transpose_f32(int R, int C, const float* in, float* out, svpc_stream_t stream);
/* The same recurrence as ONE persistent launch per sweep (all Smax steps): one CTA per SM, half of them per direction,
 * each keeping the W_hh rows of its 10-11 hidden units resident in shared memory; the steps are separated by a barrier
 * among the CTAs of a direction on `sync` (>= 3 uint32: the two arrival counters sync[0..1] are reset by the call itself
 * (a memset node ahead of the kernel); sync[2] is set if a barrier timed out and is STICKY - only the host clears it, so
 * one buffer can be polled once per epoch).  Arguments as the per-step functions.  Requires ceil(D / (SMs / 2)) <= 12 and
 * (48 + 16) D floats of shared memory; svpc_lstm_seq_supported(D) says whether the current device can hold the whole
 * grid co-resident (occupancy calculator) - when it returns 0 use the per-step functions. */
int svpc_lstm_seq_supported(int D);
int svpc_lstm_seq_fwd(int B, int D, int Nc, int Smax, const float* xg, const float* whh, const int* rows, float* hs,
                      float* cs, float* gates, float* out, float* hprev, unsigned int* sync, svpc_stream_t stream);
int svpc_lstm_seq_bwd(int B, int D, int Nc, int Smax, const float* whh_t, const int* rows, const float* cs,
                      const float* gates, const float* dout, float* dgates, float* dc, unsigned int* sync,
                      svpc_stream_t stream);

/* Greedy decoding step (src/translator.py:106-112): per clip argmax of the current position's scores with column `unk`
 * forced to -1e10; ext_out[r, pos] = argmax (extended id), next_in[r] = argmax, or unk when it is a copied OOV id
 * (>= width - n_oov).  Row r starts at scores + row_off[r] (or r * ld when row_off is NULL) and is row_width[r] (or
 * `width`) wide. */
int svpc_greedy_pick(int rows, const float* scores, const int* row_off, long long ld, const int* row_width, int width,
                     const int* row_noov, int unk, int* next_in, long long* ext_out, long long ld_ext, int pos,
                     svpc_stream_t stream);

/* nn.BCELoss(sum) on probabilities (model.py:871) and AsymmetricLoss (libs/ASL/src/loss_functions/losses.py:15-50). */
int svpc_bce_fwd(long long n, const float* p, const float* y, float* loss, svpc_stream_t stream);
int svpc_bce_bwd(long long n, const float* p, const float* y, const float* g, float* dp, svpc_stream_t stream);
int svpc_asl_fwd(int rows, int A, const float* p, const float* y, float* loss, svpc_stream_t stream);
int svpc_asl_bwd(int rows, int A, const float* p, const float* y, const float* g, float* dp, svpc_stream_t stream);

/* Training-iteration tail, multi-tensor over separately allocated fp32 tensors.  `grads`, `params`, `m`,
 * `v`, `shadow` are DEVICE arrays of n_tensors DEVICE pointers; numel (n_tensors); a chunk table gives every CTA up to
 * chunk_elems consecutive elements of one tensor: chunk_tensor / chunk_start (n_chunks), the chunks of tensor t being
 * tensor_first[t] .. tensor_first[t+1]-1.
 *   svpc_optim_sqnorm  chunk_sq[c] = sum of squares of chunk c of the gradients (fixed summation order)
 *   svpc_optim_coef    coef[t] = c_global * c_param(t): c_global = min(1, global_max_norm / (||g|| + 1e-6)) is
 *                      nn.utils.clip_grad_norm_ over all parameters (src/train.py:141-142), c_param = min(1, param_max_norm /
 *                      (c_global ||g_t|| + 1e-6)) is BertAdam's per-parameter clip (optimization.py:307-308); a max <= 0
 *                      disables that clip; norms[0] = ||g|| before clipping, norms[1] = c_global
 *   svpc_optim_apply   flags bit0: BertAdam.step (optimization.py:310-328, no bias correction): g' = coef g,
 *                      m = b1 m + (1-b1) g', v = b2 v + (1-b2) g'^2, p -= lr_t (m / (sqrt v + eps) + wd_t p);
 *                      bit1: EMA (optimization.py:196-203) shadow = (1-d) p + d shadow with d = ema_decay[0],
 *                      1-d = ema_decay[1] (device scalars, so a captured graph sees the per-step value);
 *                      bit2: also write g' back (the reference clips p.grad in place).
 *                      tensor_lr / tensor_wd (n_tensors): scheduled learning rate and weight decay of each tensor. */
int svpc_optim_sqnorm(int n_chunks, const int* chunk_tensor, const long long* chunk_start, int chunk_elems,
                      const void* const* grads, const long long* numel, float* chunk_sq, svpc_stream_t stream);
int svpc_optim_coef(int n_tensors, const int* tensor_first, const float* chunk_sq, float global_max_norm,
                    float param_max_norm, float* coef, float* norms, svpc_stream_t stream);
int svpc_optim_apply(int n_chunks, const int* chunk_tensor, const long long* chunk_start, int chunk_elems,
                     void* const* grads, void* const* params, void* const* m, void* const* v, void* const* shadow,
                     const long long* numel, const float* coef, const float* tensor_lr, const float* tensor_wd,
                     double b1, double b2, double eps, const float* ema_decay, int flags, svpc_stream_t stream);

/* Pitched asynchronous copy of `height` blocks of `width` bytes: the per-step input upload without the 25 text rows of
 * every (125, 3072) clip tensor that the path never reads (model.py:889).  kind 1: `src` is PINNED HOST
 * memory (the only host pointer in this ABI); kind 3: device to device.  Replaces the per-tensor `.cuda()` calls of
 * src/train.py:91-112 (prepare_batch_inputs).  Not counted as a kernel launch. */
int svpc_memcpy2d_async(void* dst, long long dpitch, const void* src, long long spitch, long long width, long long height,
                        int kind, svpc_stream_t stream);

/* Refresh of the compute-dtype copies of many fp32 weight tensors in one launch (functional.WeightCache; replaces the
 * per-use `.to(bf16)` / torch.cat of fused QKV weights).  src / dst: DEVICE arrays of DEVICE pointers; tensor t has
 * numel[t] elements in rows of cols[t], element i is written to dst[t][(i / cols[t]) * dst_ld[t] + i % cols[t]];
 * chunk table as svpc_optim_*.  dt_dst: SVPC_F32 or SVPC_BF16. */
int svpc_cast_multi(int n_chunks, const int* chunk_tensor, const long long* chunk_start, int chunk_elems,
                    const void* const* src, void* const* dst, const long long* numel, const int* cols,
                    const long long* dst_ld, int dt_dst, svpc_stream_t stream);

/* Training-loop bookkeeping without host synchronisation.  svpc_caption_accuracy = cal_performance +
 * n_word of src/train.py:32-38, 153-157 over all videos of a step: scores / gold are DEVICE arrays of DEVICE pointers
 * (video b: rows[b] rows of width[b] fp32 scores, rows[b] int64 labels), row_start = exclusive prefix sum of rows;
 * acc[0] += correct, acc[1] += labelled rows (label != ignore).  svpc_f1_counts = calculate_f1 (train.py:40-49):
 * acc[0] += sum gold[p > 0.5], acc[1] += sum gold, acc[2] += count(p > 0.5).  acc is fp64 (exact for counts). */
int svpc_caption_accuracy(int n_videos, const void* const* scores, const void* const* gold, const int* rows, const int* width,
                          const int* row_start, int total_rows, long long ignore, double* acc, svpc_stream_t stream);
int svpc_f1_counts(long long n, const float* p, const float* y, double* acc, svpc_stream_t stream);

/* Clip features from the per-video .npy files straight into the pinned input arena.  HOST entry points
 * (every pointer below is a host pointer; no CUDA call is made).  Together they replace, per training step,
 * `np.concatenate([np.load(<v>_resnet.npy), np.load(<v>_bn.npy)], axis=1)` (src/rtransformer/recursive_caption_dataset.py
 * :187-189), the float64 `(max_v_len + max_t_len, D)` scratch of `_load_indexed_video_feature` (:389-416), `.astype(
 * np.float32)` (:332) and the collate stack: each needed row is read once and written once, as fp32, at its final place.
 * svpc_npy_info: shape / dtype code (0 '<f4', 1 '<f8', 2 '<f2') / byte offset of the data of a rank-2 C-ordered .npy.
 * svpc_features_gather: job j copies rows[row_start[j] .. row_start[j+1]) of file paths[j] to rows first_row.. of the clip
 * block dst[j] (block_rows rows of dst_pitch elements) at columns [col_off[j], col_off[j] + file columns) and zero-fills
 * the block's other rows in that column range; jobs run on n_threads host threads.  The row lists come from
 * svpc_b200/features.py `clip_rows` (the host mirror of `_convert_to_feat_index_st_ed` / `_load_indexed_video_feature`). */
int svpc_npy_info(const char* path, long long* rows, long long* cols, int* dtype, long long* data_offset);
int svpc_features_gather(int n_jobs, const char* const* paths, const long long* row_start, const long long* rows,
                         void* const* dst, int dst_dtype /* SVPC_F32, or SVPC_BF16 (round to nearest even) for the half-size upload */,
                         long long dst_pitch /* elements */, const long long* col_off, int first_row, int block_rows,
                         int n_threads);

#ifdef __cplusplus
}
#endif
#endif
