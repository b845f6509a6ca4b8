/* C ABI of libddl25_kernels.so — the hand-written sm_100a (NVIDIA B200) kernel library of
 * ddl25spring_b200.  Python binds these through ctypes (ddl25spring_b200/ops/_lib.py); any
 * other host program can link the library and call them directly.
 *
 * Conventions
 *  - every entry point returns 0 on success, non-zero on failure; ddl_last_error() then
 *    returns a human-readable reason (thread-local string, valid until the next call).
 *  - `stream` is a cudaStream_t passed as void* (NULL = default stream). Launches are
 *    asynchronous; nothing here synchronises unless stated.
 *  - all tensors are fp32, contiguous, 16-byte aligned unless stated; "flat" vectors follow
 *    the FlatArena layout (ddl25spring_b200/core/arena.py).
 *  - "(ref: path:lines)" names the reference code the entry point replaces, relative to the
 *    DDL25Spring tree; "[north star]" marks ops the reference does not contain.
 */
#ifndef DDL25_KERNELS_H_
#define DDL25_KERNELS_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDL_MAX_ROWS 128 /* update rows per aggregation launch */
#define DDL_MAX_OUTS 80  /* broadcast destinations per aggregation launch */
#define DDL_MAX_WORLD 16 /* ranks (GPUs) in a symmetric-memory group */

/* ---- library -------------------------------------------------------------------------- */
const char* ddl_last_error(void);
/* number of kernels this library has launched in this process (bench.py: gpu_launches) */
int64_t ddl_launch_count(void);
void ddl_reset_launch_count(void);
/* compiled-for architecture string, e.g. "sm_100a" */
const char* ddl_arch(void);

/* ---- K1: FedAvg / FedSGD aggregation (ref: lab/tutorial_1a/hfl_complete.py:291-306,370-383) --
 * out[d] = alpha*base[d] + beta * sum_k weights[k]*updates[k*ld + d],  d in [0, D).
 * `weights` is a HOST array (n_k / sum n) or NULL; `dev_weights` is a DEVICE array that, when
 * non-NULL, overrides it (Krum's on-GPU selection weights); base may be NULL (alpha ignored). */
int ddl_weighted_sum(const float* updates, int64_t ld, int K, const float* weights, const float* dev_weights,
                     const float* base, float alpha, float beta, float* out, int64_t D, void* stream);
/* Row-pointer form used by the coordinate-sharded multi-GPU exchange: rows[k] / outs[o] are HOST
 * arrays of DEVICE pointers (local or peer-mapped); only coordinates [d0, d1) are touched and
 * the result is stored to every outs[o].  peer_rows != 0 uses coherent loads (rows written by
 * another GPU); max_ctas > 0 caps the grid (keep SMs free for overlapped compute). */
int ddl_weighted_sum_rows(const float* const* rows, const float* weights, const float* dev_weights, int K,
                          const float* base, float alpha, float beta, float* const* outs, int n_out, int64_t d0,
                          int64_t d1, int peer_rows, int max_ctas, void* stream);

/* ---- K4 / K5: coordinate-wise median, trimmed mean [north star] --------------------------- */
/* lower != 0: lower median for even K (torch.median); else mean of the middle pair. K <= 128. */
int ddl_coordinate_median(const float* updates, int64_t ld, int K, int lower, float* out, int64_t D, void* stream);
/* per coordinate drop the `trim` smallest and largest values, mean the rest. K <= 128. */
int ddl_trimmed_mean(const float* updates, int64_t ld, int K, int trim, float* out, int64_t D, void* stream);
/* general form: mean of sorted ranks [lo, hi) over rows[0..K), coordinates [d0, d1), broadcast to outs[] */
int ddl_rank_select_mean_rows(const float* const* rows, int K, int lo, int hi, float* const* outs, int n_out,
                              int64_t d0, int64_t d1, void* stream);

/* ---- K6: Byzantine perturbation [north star] ---------------------------------------------- */
/* out = g - scale*(u - g) (+ uniform noise in [-noise, noise]); g == NULL treats g as 0. */
int ddl_attack_perturb(const float* update, const float* global_or_null, float* out, float scale, float noise,
                       uint32_t seed, int64_t D, void* stream);

/* ---- K3 epilogue: Krum / Multi-Krum selection from a Gram matrix [north star] -------------- */
/* gram: K x K (row stride ldg) device matrix G = U U^T.  Writes scores[K], sel[multi] (indices of
 * the best rows, best first) and selw[K] (1/multi for selected rows else 0). */
int ddl_krum_select(const float* gram, int ldg, int K, int n_byzantine, int multi, float* scores, int* sel,
                    float* selw, void* stream);

/* ---- K15: optimizer steps over flat arenas ----------------------------------------------------
 * (ref: SGD lab/tutorial_1a/hfl_complete.py:196,306,319; Adam lab/tutorial_1b/DP/gradient_aggr/
 *  intro_DP_GA.py:33,67; AdamW lab/tutorial_2b/vfl.py:50,82).  grads are multiplied by grad_scale
 * first (folds the 1/world_size of a DP all-reduce into the step). n % 4 == 0. */
int ddl_sgd_step(float* params, const float* grads, float* momentum_buf, int64_t n, float lr, float weight_decay,
                 float momentum, float grad_scale, void* stream);
int ddl_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, int step,
                  float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled,
                  float grad_scale, void* stream);

/* same update, bias-correction step read from a DEVICE int (>= 1): for optimizer steps captured into a CUDA graph */
int ddl_adam_step_dev(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, const int* step_dev,
                      float lr, float beta1, float beta2, float eps, float weight_decay, int decoupled, float grad_scale,
                      void* stream);

/* ---- K13: softmax cross-entropy, value + gradient in one pass ---------------------------------
 * (ref: log_softmax+nll_loss hfl_complete.py:62,78; soft-target CE vfl.py:51,79; causalLLMLoss
 *  lab/tutorial_1b/primer/intro.py:29).  Exactly one of target_idx (int64, element stride
 * target_stride) / target_prob (float [N,C]) is non-NULL.  *loss_sum += loss_scale * sum_rows(loss_row)
 * (caller zeroes it); grad (may be NULL) = (softmax*sum(t) - t) * grad_scale; rows whose index
 * equals ignore_index are skipped; *n_valid (may be NULL) += number of contributing rows. */
int ddl_softmax_xent(const float* logits, int64_t ld, const int64_t* target_idx, int64_t target_stride,
                     const float* target_prob, float* grad, float* loss_sum, float* n_valid, int64_t N, int C,
                     float loss_scale, float grad_scale, int64_t ignore_index, void* stream);
/* binary cross-entropy on n logits against one constant target (DCGAN: real = 1, fake = 0) [north star]:
 * *loss_accum += scale * sum_i bce(x_i, target) (caller zeroes it); grad (may be NULL)_i = scale * (sigmoid(x_i) - target). */
int ddl_bce_logits_const(const float* logits, int64_t n, float target, float scale, float* loss_accum, float* grad, void* stream);
/* ---- K17: *correct += #rows whose argmax equals target (ref: hfl_complete.py:172-183) ---------- */
int ddl_eval_top1(const float* logits, int64_t ld, const int64_t* target, int64_t N, int C, int* correct,
                  void* stream);
/* ---- K14: VAE loss MSE(sum)+KL and its gradients (ref: lab/tutorial_2a/generative-modeling.py:
 *  121-130; lab/tutorial_2b/exercise_3.py:140-147).  *loss += value (caller zeroes it). */
int ddl_vae_loss(const float* x_recon, const float* x, int64_t nx, const float* mu, const float* logvar, int64_t nz,
                 float* loss, float* grad_x_recon, float* grad_mu, float* grad_logvar, float grad_scale,
                 void* stream);

/* ---- K10 / K3: grouped GEMM on tcgen05 tensor cores (TMA-fed, TMEM accumulators) ---------------
 * (ref: every nn.Linear, SURVEY 2.5 - e.g. lab/tutorial_1a/hfl_complete.py:58-61, lab/tutorial_2b/
 *  vfl.py:21-22,37-39; Krum's Gram matrix U U^T is [north star]).
 * C[g] = act(alpha * A[g] * B[g]^T + beta * C[g] + bias[g]),  A: M x K, B: N x K, C: M x N fp32, g < G
 * (beta != 0 needs n_split == 1; alpha = -lr, beta = 1 fuses the SGD step into a weight-gradient GEMM).
 *  - a_mn_major / b_mn_major == 0: operand stored [rows, K] (K contiguous, ld = row stride);
 *    != 0: stored [K, rows] (rows contiguous, ld = k-row stride; rows % (128/dtype_bytes) == 0).
 *  - dtype_bytes 4: fp32 operands, TF32 math; precise != 0 selects the 3xTF32 error-compensated
 *    mode (fp32-grade products).  dtype_bytes 2: bf16 operands.
 *  - act: 0 none, 1 relu, 2 leaky relu(slope).  bias may be NULL.
 *  - n_split > 1: split-K; `scratch` must hold ddl_gemm_scratch_floats(M, N, G, n_split) floats
 *    (partials are reduced in a fixed order, so results are deterministic). */
int64_t ddl_gemm_scratch_floats(int M, int N, int G, int n_split);
int ddl_gemm_tcgen05(const void* A, int64_t lda, int64_t a_group_stride, int a_mn_major, const void* B, int64_t ldb,
                     int64_t b_group_stride, int b_mn_major, float* C, int64_t ldc, int64_t c_group_stride,
                     const float* bias, int64_t bias_group_stride, int act, float slope, float alpha, float beta,
                     int M, int N, int K, int G, int dtype_bytes, int precise, int n_split, float* scratch,
                     void* stream);

/* Same GEMM, split-K reduced INSIDE the kernel when the grid is one wave (tiles x G x n_split <= 148 CTAs): the CTAs of a tile
 * store their partial tiles, meet at the tile's counter, and each sums its share of the rows over all partials in split order
 * (deterministic), applies bias + activation and writes C - no second launch.  Larger grids use the separate reduce kernel.
 * tile_counters: ddl_gemm_tile_count(M, N, G) ints, zero before the call (the kernel leaves them zero again); launches that
 * can overlap on the GPU must not share them. */
int ddl_gemm_tile_count(int M, int N, int G);
int ddl_gemm_tcgen05_fixup(const void* A, int64_t lda, int64_t a_group_stride, int a_mn_major, const void* B, int64_t ldb,
                           int64_t b_group_stride, int b_mn_major, float* C, int64_t ldc, int64_t c_group_stride,
                           const float* bias, int64_t bias_group_stride, int act, float slope, float alpha, float beta,
                           int M, int N, int K, int G, int dtype_bytes, int precise, int n_split, float* scratch,
                           int* tile_counters, void* stream);

/* K3: G = (U - 1 c^T)(U - 1 c^T)^T for U = [K <= 128, D] (row stride ld): the Gram matrix of the update DELTAS u_k - c
 * [north star: Krum].  The centre c (the global model, [D], 16-byte aligned; NULL = plain U U^T) is subtracted from the
 * staged tiles in shared memory (3xTF32 mode is implied), so no K x D delta matrix is written to HBM.  Split-K scratch:
 * n_split * K * K floats when n_split > 1. */
int ddl_gram_tcgen05(const float* updates, int64_t ld, int K, int64_t D, const float* center, float* gram, int64_t ldg,
                     int precise, int n_split, float* scratch, void* stream);

/* ---- operand preparation for ragged layer widths (csrc/elementwise.cu) ------------------------------------------------
 * (ref: the nn.Linear layers of the tabular models, lab/tutorial_2a/centralized.py:24-28, lab/tutorial_2b/vfl.py:21-22,37-39,
 *  lab/tutorial_2a/generative-modeling.py:50-75: widths 2..256 that are not multiples of the GEMM's 32-column k-block.)
 * ddl_pad2d: dst[rows_dst, cols_dst] (row stride ld_dst) = src[rows, cols] zero-padded; column fill_col (>= cols, or -1 for
 *            none) of the first `rows` rows is set to fill_vec[r] (if non-NULL) else fill_val - a ones column in X / the bias
 *            in W folds the bias into the GEMM and makes the weight-gradient GEMM produce the bias gradient as well.
 * ddl_col_sum: out[c] = alpha * sum_r x[r, c] + beta * out[c]  (deterministic). */
int ddl_pad2d(const float* src, int64_t ld_src, int rows, int cols, float* dst, int64_t ld_dst, int rows_dst, int cols_dst,
              int fill_col, const float* fill_vec, float fill_val, void* stream);
int ddl_col_sum(const float* x, int64_t ld, int rows, int cols, float* out, float alpha, float beta, void* stream);

/* ---- Packed MnistCnn engine: G clients trained in lock-step, grouped kernels -------------------
 * (ref: MnistCnn forward/backward + train_epoch, lab/tutorial_1a/hfl_complete.py:39-80, executed
 *  for all clients packed on the GPU at once instead of one client at a time, :365-375).
 * W: [G][1 200 000] flat weight arena (FlatArena layout of MnistCnn); rows = G*B activation rows;
 * bsz[g] = valid rows of client g in this step; perm (may be NULL) = per-client sample order. */
int ddl_mnist_conv1_fwd(const float* W, const float* X, int64_t x_client_stride, const int* perm, int64_t perm_stride,
                        int sample_off, float* a1, int G, int B, const int* bsz, void* stream);
int ddl_mnist_prep_conv2(const float* W, float* wt_fwd, float* wt_dgrad, int G, void* stream);
/* K9 fprop on tcgen05: conv2 (implicit GEMM, TF32) + bias + ReLU + 2x2 max-pool (+ dropout) */
int ddl_mnist_conv2_pool_fwd(const float* W, const float* wt_fwd, const float* a1, float* pooled, uint8_t* pmask,
                             int G, int B, const int* bsz, int train, float p_drop, uint32_t seed, void* stream);
int ddl_mnist_head_scratch_floats(int G);
int ddl_mnist_head_train(float* W, const float* h1, const int64_t* Y, int64_t y_client_stride, const int* perm,
                         int64_t perm_stride, int sample_off, float* dz1, float* loss_out, float* scratch, int G, int B,
                         const int* bsz, float p_drop, uint32_t seed, float lr, void* stream);
int ddl_mnist_head_eval(const float* W, const float* h1, const int64_t* Y, int sample_off, int nrows, float* logits_out,
                        int* correct, void* stream);

/* backward of the packed engine (ref: autograd backward + optimizer.step, hfl_complete.py:79-80) */
int ddl_mnist_pool_bwd(const float* dpool, const uint8_t* pmask, float* dy2, float* c2b_part, int G, int B,
                       const int* bsz, float p_drop, void* stream);
int ddl_mnist_conv2_dgrad(const float* wt_dgrad, const float* dy2, const float* a1, float* da1, int G, int B,
                          const int* bsz, void* stream);
/* wg_part: [G][ddl_mnist_wgrad_ctas(G)][3][128][64] floats of scratch */
int ddl_mnist_wgrad_ctas(int G);
int ddl_mnist_conv2_wgrad(const float* a1, const float* dy2, float* wg_part, int G, int B, const int* bsz,
                          void* stream);
int ddl_mnist_conv1_wgrad(const float* X, int64_t x_client_stride, const int* perm, int64_t perm_stride, int sample_off,
                          const float* da1, float* c1_part, int G, int B, const int* bsz, void* stream);
int ddl_mnist_small_update(float* W, const float* wg_part, int n_cta, const float* c2b_part, const float* c1_part,
                           float* wt_fwd, float* wt_dgrad, int G, int B, const int* bsz, float lr, void* stream);

/* ---- Gram matrix without any allocation (no counterpart in the reference) ----------------------
 * G = U U^T like ddl_gemm_tcgen05(A = B = U, n_split > 1), for callers that cannot provide a scratch
 * buffer at call time: plain C/C++ hosts without PyTorch's caching allocator, or services whose
 * memory is fixed at start-up and for which a cudaMalloc in the round loop (it synchronises the
 * device and may fail late) is not acceptable.  The split-K partial tiles go to a static device
 * array that ships inside the library: DDL_GRAM_STATIC_SCRATCH_FLOATS floats (4 MiB) held by every
 * GPU that loads the library until the process exits.  This call never allocates or frees.
 *  - limit: n_split * K * K <= DDL_GRAM_STATIC_SCRATCH_FLOATS (K <= 128 clients), else an error;
 *  - all calls on a device share the one array: calls that may overlap on the GPU race - keep them
 *    on one stream (or order them with events); leftover contents never need clearing because every
 *    partial tile is fully rewritten before it is read;
 *  - the first call on each device looks up the array's device address (cudaGetSymbolAddress) and
 *    does the one-time kernel set-up: make that first call before starting any stream capture;
 *    the address is cached per device ordinal for the life of the process, so it does not survive a
 *    cudaDeviceReset() - do not reset a device between calls;
 *  - D must be a multiple of 4 (the tensor map's inner extent is a whole number of 16-byte units).
 * updates: [K, D] fp32, row stride ld (ld*4 and the pointer 16-byte aligned); gram: [K, K], row stride ldg. */
#define DDL_GRAM_STATIC_SCRATCH_FLOATS (1 << 20)
int ddl_gram_static_scratch(const float* updates, int64_t ld, int K, int64_t D, float* gram, int64_t ldg,
                            int n_split, int precise, void* stream);

/* ---- Symmetric peer memory + fused cross-GPU reduce (K1 / K2 / K2b / K4 / K5 over NVLink) ------
 * (ref: gloo all_reduce of flattened grads/weights, lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:53-67,
 *  weight_aggr/intro_DP_WA.py:54-67; client-message reduction lab/tutorial_1a/hfl_complete.py:370-383).
 * ddl_symm_alloc/ipc_*: cudaMalloc'd buffers shared between the ranks of one node with CUDA IPC.
 * ddl_symm_reduce: ONE kernel per GPU = entry barrier + reduce of coordinates [d0, d1) over all rows
 * (local or peer pointers; HOST arrays of device pointers) + store to every outs[] replica + exit barrier.
 *  mode 0: out = alpha*base + beta*sum_k weights[k]*rows[k]  (dev_weights: optional DEVICE weights that override
 *          `weights`, e.g. Krum's selection weights; zero-weight rows are not read);  with opt->kind != 0 the sum
 *          is a gradient: out = the new parameters after an SGD / Adam / AdamW step on base (the local parameter
 *          shard) - reduce-scatter, optimizer step and all-gather of a DP bucket in one launch;  with mc_in /
 *          mc_out (NVLS multicast addresses of the summed / destination buffers) the peer loads / stores go
 *          through the switch (multimem.ld_reduce / multimem.st) and rows / outs may be NULL / 0.
 *  mode 1: mean of sorted ranks [rank_lo, rank_hi) (K <= 64; a window that is the median uses a pruned network);
 *  mode 2: gather of this GPU's coordinate shard of every row (minus base) into outs[0].
 * A CHANNEL is pads[p] (int[2*world] signal pad on GPU p) + state (int[4] on this GPU, zeroed once): the sequence
 * number of the collective lives in `state` and is advanced by the kernel, so launches can be captured into CUDA
 * graphs; collectives of one channel must be stream-ordered and issued in the same order on every rank.
 * *err (device int, zeroed once) becomes non-zero if a peer did not arrive within 5 s; from then on every
 * collective given that flag skips its stores. */
typedef struct ddl_opt_args {
  int kind;            /* 0 none, 1 SGD, 2 Adam (L2 decay), 3 AdamW (decoupled decay) */
  float lr, weight_decay, momentum, beta1, beta2, eps;
  float* m;            /* SGD momentum buffer / Adam exp_avg of the owned shard (index: coordinate - 4*state_off4) */
  float* v;            /* Adam exp_avg_sq of the owned shard */
  int64_t state_off4;  /* first owned coordinate / 4 */
  int* step_dev;       /* DEVICE int: completed optimizer steps (bias correction uses *step_dev + 1) */
  int bump_step;       /* != 0 on the LAST bucket of a step: the kernel increments *step_dev when it finishes */
} ddl_opt_args;
int ddl_symm_alloc(void** ptr, int64_t bytes);
int ddl_symm_free(void* ptr);
int ddl_ipc_export(void* ptr, unsigned char* handle64);
int ddl_ipc_import(const unsigned char* handle64, void** ptr);
int ddl_ipc_close(void* ptr);
int ddl_symm_barrier(int* const* pads, int rank, int world, int* state, int* err, void* stream);
int ddl_symm_reduce(int mode, const float* const* rows, const float* weights, const float* dev_weights, int K, const float* base,
                    float alpha, float beta, int rank_lo, int rank_hi, float* const* outs, int n_out, int64_t d0, int64_t d1,
                    int* const* pads, int rank, int world, int* state, int* err, int max_ctas, const ddl_opt_args* opt,
                    const float* mc_in, float* mc_out, void* stream);

/* ---- K7: vertical-FL embedding gather fused into the top model's first GEMM --------------------
 * (ref: torch.cat(party_outs, 1) + fc1, lab/tutorial_2b/vfl.py:36-37; exercise_3.py:128,85).
 * C[M,N] = act(cat_p(parts[p])[M, sum dims] * W^T + bias); W is [N, sum_p round_up(dims[p], 4)]: the slab of part p
 * starts at column sum_{q<p} round_up(dims[q], 4) (TMA box origins must be 16-byte aligned), padding columns are ignored; parts[p] are device pointers that may be
 * peer-mapped (each party's embedding stays in its owner GPU's HBM; k-blocks are TMA-loaded over NVLink).
 * parts / part_dims / part_lds are HOST arrays (n_parts <= 8); part_lds[p] >= round_up(part_dims[p], 4) and the
 * padding columns [dims, round_up(dims, 4)) of every part must be zero. */
int ddl_gather_gemm_tcgen05(const float* const* parts, const int* part_dims, const int64_t* part_lds, int n_parts,
                            const float* W, int64_t ldw, float* C, int64_t ldc, const float* bias, int act, float slope,
                            int M, int N, int precise, int n_split, float* scratch, void* stream);

/* ---- K16 (part): RMSNorm forward / backward, one warp per row (ref: the LLaMA blocks of `simplellm`,
 * lab/tutorial_1b/primer/intro.py:17,28).  rstd [rows] is written by fwd and read by bwd; dw must be zeroed. */
int ddl_rmsnorm_fwd(const float* x, const float* w, float* y, float* rstd, int64_t rows, int d, float eps, void* stream);
int ddl_rmsnorm_bwd(const float* x, const float* w, const float* dy, const float* rstd, float* dx, float* dw, int64_t rows, int d,
                    void* stream);

/* ---- K11: BatchNorm 1d / 2d (ref: lab/tutorial_2a/generative-modeling.py:21-45, lab/tutorial_2b/exercise_3.py; ResNet-18 /
 * DCGAN [north star]).  x, y, dy, dx: [N, C, S] (S = H*W or 1).  Training forward writes save_mean / save_rstd [C] and
 * updates running_mean / running_var (may be NULL) like torch (momentum, unbiased running variance). */
int ddl_batchnorm_fwd(const float* x, const float* gamma, const float* beta, float* y, float* save_mean, float* save_rstd,
                      float* running_mean, float* running_var, int N, int C, int S, float eps, float momentum, int training,
                      void* stream);
int ddl_batchnorm_bwd(const float* x, const float* dy, const float* gamma, const float* save_mean, const float* save_rstd, float* dx,
                      float* dgamma, float* dbeta, int N, int C, int S, void* stream);

/* ---- K16: causal self-attention, flash-style on CUDA cores for the labs' shape (T = 256, head dim 48) (ref: the LLaMA blocks of
 * `simplellm`, lab/tutorial_1b/primer/intro.py:17,28).  q, k, v, o, dO, dq, dk, dv: [BH, T, D] fp32, D in {16, 32, 48, 64};
 * lse, delta_scratch: [BH, T].  Nothing of size T x T is materialised; backward recomputes probabilities from lse. */
int ddl_causal_attention_fwd(const float* q, const float* k, const float* v, float* o, float* lse, int BH, int T, int D, float scale,
                             void* stream);
int ddl_causal_attention_bwd(const float* q, const float* k, const float* v, const float* o, const float* dO, const float* lse,
                             float* dq, float* dk, float* dv, float* delta_scratch, int BH, int T, int D, float scale, void* stream);

/* ---- K16 on the tensor cores (csrc/attention_tc.cu): the row-wise kernels between the grouped tcgen05 GEMMs of causal attention
 * (ref: attention inside simplellm's LLaMA blocks, lab/tutorial_1b/primer/intro.py:17,28).  S, P, dP, dS: [rows = G*T, T];
 * O, dO_padded: [G*T, HP] (head dim zero-padded to HP).
 * ddl_causal_softmax: row r of S (already scaled) becomes softmax over columns 0 .. r % T, zeros after (in place).
 * ddl_attn_bwd_prep : dO_padded = zero-padded copy of dO[b, h, t, 0:hd] (element strides given); delta[row] = sum_d dO * O.
 * ddl_attn_ds       : dP <- scale * P * (dP - delta[row])  (in place). */
int ddl_causal_softmax(float* S, int64_t rows, int T, void* stream);
int ddl_attn_bwd_prep(const float* dO, int64_t stride_b, int64_t stride_h, int64_t stride_t, const float* O_padded, float* dO_padded,
                      float* delta, int B, int H, int T, int hd, int HP, void* stream);
int ddl_attn_ds(const float* P, float* dP_inout, const float* delta, int64_t rows, int T, float scale, void* stream);

/* ---- K3 fused: all-gather of client updates + Gram GEMM in one kernel [north star] --------------------------------------
 * gram[K x K] = U[:, d0:d1] U[:, d0:d1]^T with K = n_parts * rows_per_part and the rows of U spread over GPUs: part_ptrs[p]
 * (HOST array of device pointers, may be peer-mapped) is the base of GPU p's rows_per_part rows (row stride ld floats).
 * Every k-block is TMA-loaded straight from the owning GPUs into one shared-memory tile that serves as both MMA operands.
 * Row order of the result: p * rows_per_part + i.  rows_per_part % 8 == 0, K <= 128; split-K scratch: n_split*K*K floats.
 * center (may be NULL; [>= d1] floats, precise != 0 required): subtracted from every row while the tiles sit in shared
 * memory, so the Gram matrix is that of the deltas u_k - center although raw weights cross NVLink. */
int ddl_gram_rowgather_tcgen05(const float* const* part_ptrs, int n_parts, int rows_per_part, int64_t ld, int64_t d0, int64_t d1,
                               const float* center, float* gram, int64_t ldg, int precise, int n_split, float* scratch, void* stream);

/* ---- NVLS all-reduce: in-switch reduction through a multicast address (multimem.ld_reduce / multimem.st) -----------------
 * (ref: gloo all_reduce, lab/tutorial_1b/DP/gradient_aggr/intro_DP_GA.py:63).  multicast_ptr addresses the same symmetric
 * buffer on every GPU; this GPU reduces + re-broadcasts coordinates [d0, d1) (its shard), scaled by `scale`.  The caller
 * orders the GPUs with barriers before (inputs written) and after (results visible). */
int ddl_nvls_allreduce(float* multicast_ptr, int64_t d0, int64_t d1, float scale, int max_ctas, void* stream);

/* ---- input pipeline: out = (in / 255 - mean) / std, bit-identical to the host float transform (csrc/data.cu) ---------------
 * (ref: ToTensor + Normalize in the DataLoader, lab/tutorial_1a/hfl_complete.py:19-24.)  n % 4 == 0. */
int ddl_u8_normalize(const unsigned char* in, float* out, int64_t n, float mean, float std, void* stream);

/* ---- K16 element-wise: fused QKV split + head transpose + rotary embedding, SwiGLU (csrc/llama_ops.cu) ---------------------
 * (ref: simplellm LLaMA block internals used by lab/tutorial_1b/primer/intro.py:17,28.)
 * qkv: [B, T, 3, H, hd]; q, k, v: [B, H, T, hd]; cos/sin: [>= T, hd/2].  backward == 0: qkv -> q, k, v (q, k rotated);
 * backward != 0: dq, dk, dv (passed as q, k, v) -> d(qkv) written into `qkv`. */
int ddl_rope_qkv(float* qkv, const float* cos_tab, const float* sin_tab, float* q, float* k, float* v, int B, int T, int H, int hd,
                 int hd_pitch /* row pitch of q, k, v (>= hd; pad columns are written as zeros); <= 0: hd */, int backward, void* stream);
/* embedding gather out[n, :] = W[tokens[n], :] and its scatter-add backward dW[tokens[n], :] += grad_out[n, :] (dW_accum is
 * accumulated into: zero it first; rows equal to padding_idx, or -1 for none, receive nothing).  d % 4 == 0. */
int ddl_embedding_fwd(const float* W, const int64_t* tokens, float* out, int64_t n_tokens, int d, int64_t vocab, void* stream);
int ddl_embedding_bwd(const float* grad_out, const int64_t* tokens, float* dW_accum, int64_t n_tokens, int d, int64_t vocab,
                      int64_t padding_idx, void* stream);
int ddl_swiglu_fwd(const float* g, const float* u, float* h, int64_t n, void* stream);               /* h = silu(g) * u */
int ddl_swiglu_bwd(const float* g, const float* u, const float* dh, float* dg, float* du, int64_t n, void* stream);
/* stacked layout: gu = [M, 2F] (gate | up per row, the output of ONE GEMM against [Wg; Wu]); h, dh = [M, F]; dgu = [M, 2F] */
int ddl_swiglu_gu_fwd(const float* gu, float* h, int64_t M, int F, void* stream);
int ddl_swiglu_gu_bwd(const float* gu, const float* dh, float* dgu, int64_t M, int F, void* stream);

/* ---- K9 general: NHWC / OHWI convolutions as tap GEMMs on tcgen05 (csrc/conv_tcgen05.cu) -------------------------------
 * (ref: nn.Conv2d via ATen/cuDNN, lab/tutorial_1a/hfl_complete.py:51,53; ResNet-18 / DCGAN convs of the north star.)
 * A convolution is a list of taps; tap t reads activation map tap_map[t] (an NHWC tensor [N, Ha, Wa, Ca]) at spatial
 * offset (tap_dh[t], tap_dw[t]) and multiplies by weight slice Wt[:, tap_w[t], :] of the OHWI weight [Cout, T, Cin].
 * w_mn_major = 0: forward (reduce over Cin, produce Cout channels); 1: input gradient (reduce over Cout, produce Cin).
 * Output pixel (n, i, j), n < N, i < Hh, j < Ww, is stored at out + o_off + n*o_sn + i*o_sh + j*o_sw (elements):
 * out = act(alpha * conv + beta * out + bias).  act: 0 none, 1 relu, 2 leaky(slope), 3 tanh.
 * n_phases > 1 (stride-2 input gradient / transposed conv): phase f uses taps [phase_begin[f], phase_begin[f+1]) and adds
 * phase_off[f] to o_off; all phases run in one launch.  n_phases <= 1: pass NULL for both arrays.
 * Channel counts: multiples of 4, and of 32 when above 32.  scratch (optional) enables split-K on small grids. */
int ddl_conv_tap_gemm(const float* const* acts, int n_maps, int N, int Ha, int Wa, int Ca, const float* Wt, int Cout, int T, int Cin,
                      int w_mn_major, const int* tap_map, const int* tap_dh, const int* tap_dw, const int* tap_w, int n_taps,
                      int n_phases, const int* phase_begin, const int64_t* phase_off, float* out, int Hh, int Ww, int64_t o_off, int64_t o_sn, int64_t o_sh, int64_t o_sw, const float* bias,
                      float alpha, float beta, int act, float slope, float* scratch, int64_t scratch_floats, void* stream);
/* dW[co, tap_w[t], ci] = alpha * sum_{n,i,j} dY[n,i,j,co] * X_{tap_map[t]}[n, i+dh, j+dw, ci] + beta * dW   (dY: [N,Ho,Wo,Cout]) */
int ddl_conv_wgrad(const float* dY, int N, int Ho, int Wo, int Cout, const float* const* xs, int n_maps, int Hx, int Wx, int Cin,
                   const int* tap_map, const int* tap_dh, const int* tap_dw, const int* tap_w, int n_taps, float* dW, int T,
                   float alpha, float beta, float* scratch, int64_t scratch_floats, void* stream);

/* ---- NHWC companions (csrc/nhwc.cu): BatchNorm(+ReLU+residual), stride-2 phase split, pools, stem im2col ------------------
 * (ref: ATen batch_norm / relu / add / max_pool2d / adaptive_avg_pool2d of the north-star ResNet-18.)
 * x, y, dy, ...: [R = N*H*W, C] fp32.  partial: ddl_bn_nhwc_ws_floats(R, C) floats; counters: (C+31)/32 zeroed unsigned. */
int64_t ddl_bn_nhwc_ws_floats(int64_t R, int C);
/* n_tracked (may be NULL): the module's int64 num_batches_tracked, incremented by one on the device */
int ddl_bn_nhwc_stats(const float* x, int64_t R, int C, float* partial, unsigned* counters, float* mean, float* rstd, float* run_mean,
                      float* run_var, long long* n_tracked, float eps, float momentum, void* stream);
int ddl_bn_nhwc_apply(const float* x, const float* res, float* y, const float* mean, const float* rstd, const float* gamma,
                      const float* beta, int64_t R, int C, int relu /* 0 none, 1 relu, 2 leaky(slope) */, float slope, void* stream);
int ddl_bn_nhwc_bwd(const float* x, const float* y, const float* dy, float* dx, float* dres, const float* mean, const float* rstd,
                    const float* gamma, float* dgamma, float* dbeta, float* partial, unsigned* counters, int64_t R, int C, int relu,
                    float slope, void* stream);
int ddl_phase_split(const float* x, float* ph, int N, int H, int W, int C, void* stream);      /* ph: [4][N,H/2,W/2,C] */
int ddl_maxpool3x3s2_fwd(const float* x, float* y, unsigned char* code, int N, int H, int W, int C, void* stream);
int ddl_maxpool3x3s2_bwd(const float* dy, const unsigned char* code, float* dx, int N, int H, int W, int C, void* stream);
int ddl_avgpool_nhwc_fwd(const float* x, float* y, int N, int S, int C, void* stream);
int ddl_avgpool_nhwc_bwd(const float* dy, float* dx, int N, int S, int C, void* stream);
int ddl_im2col_stem7(const float* x, float* col, int N, int Cin, int H, int W, void* stream);  /* x NCHW -> col [N*Ho*Wo, 224] (196 + 28 zeros) */

#ifdef __cplusplus
}
#endif
#endif /* DDL25_KERNELS_H_ */
