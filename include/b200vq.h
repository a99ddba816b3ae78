/*
 * b200vq.h -- C ABI of the B200-native quantization bottleneck (libb200vq.so).
 *
 * Drop-in boundary for the hot path of zhxie0117/video-tokenizer's
 * models/bottleneck.py.  The reference has no FFI of its own (it is pure PyTorch), so each
 * entry point cites the reference code it replaces; the Python binding a maintainer adds
 * (ctypes, no torch types cross this line) is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer on the current CUDA device unless it says "host";
 *   - the library never allocates: scratch comes from a caller workspace
 *     (b200vq_workspace_bytes), outputs are caller-owned;
 *   - every call only ENQUEUES work on `stream` (a cudaStream_t); no host sync, no host
 *     read of device data, CUDA-graph capturable;
 *   - return value: 0 on success, a negative B200VQ_E* code otherwise
 *     (b200vq_error_string gives the text);
 *   - z / grad_z may be fp32, bf16 or fp16 (B200VQ_DT_*); all quantizer arithmetic and all
 *     other tensors are fp32 (reference bottleneck.py:252,262-264 force fp32), indices are
 *     int64 (torch.argmin / multinomial dtype).
 */
#ifndef B200VQ_H_
#define B200VQ_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200VQ_VERSION 122

/* element types of z / grad_z */
#define B200VQ_DT_F32 0
#define B200VQ_DT_BF16 1
#define B200VQ_DT_F16 2

/* code selection rule */
#define B200VQ_MODE_ARGMIN_DIST 0 /* argmin_k |z|^2+|e_k|^2-2 z.e_k   (bottleneck.py:282-290) */
#define B200VQ_MODE_ARGMAX_COS 1  /* argmax_k softmax(z.e_k/T) == argmax z.e_k (:277-278)    */
#define B200VQ_MODE_SAMPLE 2      /* k ~ softmax(z.e_k/T), Gumbel-max       (:275-276,:280)  */

/* error codes */
#define B200VQ_OK 0
#define B200VQ_EINVAL (-1)    /* bad shape / size / enum                       */
#define B200VQ_EDIM (-2)      /* unsupported code dimension D                  */
#define B200VQ_EALIGN (-3)    /* pointer or stride not aligned as required     */
#define B200VQ_ECUDA (-4)     /* a CUDA call failed; see b200vq_last_cuda_error */
#define B200VQ_EWORKSPACE (-5) /* workspace too small / null                    */
#define B200VQ_ENOALLOC_SIZE (-6) /* (C, D) needs more than B200VQ_NOALLOC_SCRATCH_BYTES of built-in scratch */

typedef void* b200vq_stream_t; /* cudaStream_t */

int b200vq_version(void);
const char* b200vq_error_string(int code);
/* cudaError_t of the last failing CUDA call made by this library on this thread */
int b200vq_last_cuda_error(void);
/* largest supported code dimension */
int b200vq_max_dim(void);
/* number of kernels this library has launched in this process (bench.py's gpu_launches) */
uint64_t b200vq_launch_count(void);
/* Measurement aid: a register-resident fp32 FMA loop over every SM, `iters` rounds of 8 independent
 * chains per thread.  *out_flops_per_launch (host, nullable) = 2 * 8 * iters * threads launched; out (device, 1 float)
 * keeps the result live.  Time it with CUDA events to get the box's fp32 FMA peak (the roofline of the exact-fp32
 * kernels) instead of quoting the nominal 128 lanes x 2 x clock. */
int b200vq_fma_probe(int64_t iters, float* out, double* out_flops_per_launch, b200vq_stream_t stream);

/* Scratch bytes the forward/backward calls need for N tokens, D dims, K codes. */
size_t b200vq_workspace_bytes(int64_t N, int D, int64_t K);

/*
 * get_emb (bottleneck.py:252-259) -- the codebook as the quantizer sees it, once per step.
 *   weight (K,D) fp32 raw embedding.weight
 *   emb    (K,D) fp32 out: weight / max(|weight|,1e-12) if l2norm else a copy
 *   e2     (K)   fp32 out: sum_d emb^2   (the |e_k|^2 term of :283-285)
 *   wnorm  (K)   fp32 out: |weight_k| (unclamped; the backward's normalise Jacobian needs it)
 */
int b200vq_prepare_codebook(const float* weight, int64_t K, int D, int l2norm,
                            float* emb, float* e2, float* wnorm, b200vq_stream_t stream);

/*
 * Selection keys.  key[i] = (orderable(score_i,best) << 32) | ~k_best ; unsigned max over
 * keys == best score, lowest index on ties.  keys must be reset before the first partial
 * call of a step.
 */
int b200vq_keys_reset(uint64_t* keys, int64_t N, b200vq_stream_t stream);

/*
 * Scoring pass over a slice [k_offset, k_offset+K_local) of the codebook
 * (bottleneck.py:264-290 without materialising the N x K matrix).  Normalises z rows on
 * the fly, streams the slice through shared memory, keeps a per-row running best and
 * merges it into keys[] with a 64-bit atomic max.  Token-sharded / single GPU: one call
 * with the whole codebook.  Codebook-sharded: one call per rank on its slice, then an
 * all-reduce(MAX, uint64) of keys[].
 *   z         (N,D) rows, element type z_dtype, row stride z_row_stride ELEMENTS
 *   emb, e2   slice base pointers (already offset by the caller: emb + k_offset*D)
 *   t_inv     1/T (MODE_SAMPLE only)
 *   seed, offset  Philox key / stream offset (MODE_SAMPLE only), taken from the torch CUDA
 *             generator by the host wrapper; noise is keyed by (row_offset+i, k_offset+k)
 *             so draws do not depend on tiling or sharding
 */
int b200vq_scores_partial(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D,
                          const float* emb, const float* e2, int64_t K_local, int64_t k_offset,
                          int mode, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                          int64_t row_offset, uint64_t* keys, b200vq_stream_t stream);

/*
 * Tensor-core (tcgen05) scoring path.  Same contract as b200vq_scores_partial -- keys[] get
 * the per-row winner -- but the N x K contraction runs on the 5th-gen tensor cores with
 * split-bf16 operands (error <= 1.2e-5 on unit vectors) and the accumulator tile in TMEM.
 * Deterministic modes: rows whose best and second-best approximate scores are closer than
 * 4e-5 are re-scored in exact fp32 inside the same call, so the indices are bit-identical
 * to b200vq_scores_partial.  MODE_SAMPLE: lazy Gumbel-max over the approximate scores with
 * the same per-(row, code) noise as the exact path.
 *   b200vq_tc_supported   1 if (D, l2norm, mode) can take this path (D in {8,16,24,32}, l2norm)
 *   b200vq_tc_image_bytes size of the codebook image for K codes
 *   b200vq_prepare_codebook_tc  emb (K,D) fp32 normalised codebook -> image (16-byte aligned)
 *   scratch     b200vq_tc_scratch_bytes(N) bytes, zeroed by the caller with keys (deterministic
 *               modes): runner-up scores, then the list of undecided rows
 *   n_resolved  nullable device counter, += number of rows re-scored exactly
 *   k_offset    multiple of 256
 */
int b200vq_tc_supported(int D, int l2norm, int mode);
size_t b200vq_tc_image_bytes(int64_t K, int D);
size_t b200vq_tc_scratch_bytes(int64_t N);
int b200vq_prepare_codebook_tc(const float* emb, int64_t K, int D, void* image, b200vq_stream_t stream);
int b200vq_scores_tc(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D,
                     const void* image, const float* emb, const float* e2, int64_t K, int64_t k_offset,
                     int mode, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                     int64_t row_offset, uint64_t* keys, float* scratch, unsigned int* n_resolved,
                     b200vq_stream_t stream);

/*
 * Gather + straight-through + losses + usage histograms from keys
 * (bottleneck.py:292-298, :304-310; trainer bincounts larp_tokenizer_trainer.py:337-343).
 *   emb        full (K,D) normalised codebook (may be null when q_rows is given)
 *   q_rows     nullable (N,D): the selected code rows already gathered (codebook-sharded
 *              mode, where no rank holds the whole codebook); else rows come from emb
 *   n_first    number of leading rows that belong to sample 0 of the batch (token_nums);
 *              those also count into hist_first (the trainer's "first instance" stats)
 *   znorm_out  (N,D) fp32  'unregularized_z'  (normalised z)
 *   ste_out    (N,D) fp32  'regularized_z'    z + (q - z)
 *   idx_out    (N)   int64 'bottleneck_rep'
 *   zlen_out   (N)   fp32  |z_i| (unclamped) for the backward
 *   hist, hist_first (K) int32, ACCUMULATED into (caller zeroes them)
 *   sqerr      (1) double, ACCUMULATED: sum_i |q_i - z_i|^2 (caller zeroes)
 *   losses     (3) fp32 out when finalize!=0: loss_q, loss_commit, loss_codebook computed
 *              from sqerr / (N_total*D) with beta, w_cb  (:295-298,:304)
 */
int b200vq_finish(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D,
                  const float* emb, int64_t K, const uint64_t* keys, const float* q_rows, int l2norm,
                  int64_t n_first,
                  float* znorm_out, float* ste_out, int64_t* idx_out, float* zlen_out,
                  int32_t* hist, int32_t* hist_first, double* sqerr,
                  int finalize, int64_t N_total, float beta, float w_cb, float* losses,
                  b200vq_stream_t stream);

/*
 * One-call forward for the unsharded case: keys_reset + scores_partial(whole codebook) +
 * finish(finalize=1).  workspace >= b200vq_workspace_bytes(N,D,K); hist/hist_first/sqerr
 * are zeroed inside.  image: nullable tensor-core codebook image (b200vq_prepare_codebook_tc);
 * when given and b200vq_tc_supported(D,l2norm,mode), scores come from b200vq_scores_tc,
 * otherwise from the exact-fp32 b200vq_scores_partial.  Deterministic indices are identical
 * either way.
 */
int b200vq_forward(const void* z, int z_dtype, int64_t z_row_stride, int64_t N, int D,
                   const float* emb, const float* e2, const void* image, int64_t K,
                   int mode, int l2norm, float t_inv, uint64_t seed, uint64_t offset,
                   int64_t n_first, float beta, float w_cb,
                   float* znorm_out, float* ste_out, int64_t* idx_out, float* zlen_out,
                   int32_t* hist, int32_t* hist_first, float* losses,
                   void* workspace, size_t workspace_bytes, b200vq_stream_t stream);

/*
 * Fused backward of SimpleVectorQuantizer.forward (autograd of bottleneck.py:292-307;
 * closed form in SURVEY.md 8a-5): straight-through + commitment gradient pushed through
 * the normalise Jacobian into grad_z, and the codebook-loss gradient scatter-added into
 * grad_emb (K,D) (zeroed inside).
 *   g_ste      (N,D) fp32 grad of 'regularized_z'      (nullable = 0)
 *   g_znorm    (N,D) fp32 grad of 'unregularized_z'    (nullable = 0)
 *   g_loss     (3)   fp32 DEVICE grads of loss_q, loss_commit, loss_codebook (nullable = 0)
 *   znorm      (N,D) fp32 saved normalised z ; zlen (N) saved |z|
 *   idx        (N)   int64 saved (global) indices
 *   emb        (K,D) the codebook rows [k_offset, k_offset+K) this rank owns (K = full
 *              codebook and k_offset = 0 when unsharded); grad_emb covers the same rows and
 *              only owned codes are scatter-added
 *   q_rows     nullable (N,D) selected code rows (required when the codebook is sharded)
 *   grad_z     (N,D) out, element type gz_dtype, contiguous
 *   grad_emb   (K,D) fp32 out: d loss / d emb  (before the codebook normalise Jacobian)
 */
int b200vq_backward(const float* g_ste, const float* g_znorm, const float* g_loss,
                    const float* znorm, const float* zlen, const int64_t* idx, const float* emb,
                    const float* q_rows, int64_t k_offset,
                    int64_t N, int D, int64_t K, int l2norm, int64_t N_total, float beta, float w_cb,
                    void* grad_z, int gz_dtype, float* grad_emb, b200vq_stream_t stream);

/*
 * Codebook side of the backward: grad_weight = J_normalise(weight)^T (grad_emb + g_emb)
 * (autograd of F.normalize at bottleneck.py:255).  g_emb = upstream grad on 'emb', nullable.
 */
int b200vq_codebook_backward(const float* grad_emb, const float* g_emb, const float* emb,
                             const float* wnorm, int64_t K, int D, int l2norm, float* grad_weight,
                             b200vq_stream_t stream);

/*
 * get_codebook_entry / decode (bottleneck.py:327-344): out[i] = normalise(weight[idx[i]]).
 * Out-of-range indices make the call write NaN rows and set *bad_flag (device int, nullable).
 */
int b200vq_decode(const int64_t* idx, int64_t n, const float* weight, int64_t K, int D, int l2norm,
                  float* out, int* bad_flag, b200vq_stream_t stream);

/*
 * Index -> hidden in one pass (Bottleneck.decode, bottleneck.py:166-168 = get_codebook_entry :327-341 +
 * out_linear :162-164; LARPTokenizer.decode_from_bottleneck + the decoder's additive latent positional
 * table, models/larp_tokenizer.py:463-464, 484-487):
 *   out[i, :] = normalise(weight[idx[i]]) w_out^T + b_out (+ add[i % add_rows, :])
 * weight (K,D), w_out (C,D), b_out (C, nullable), add (add_rows, C) nullable, out (n, C; row stride ldo,
 * 16-byte aligned).  The (n, D) code rows never touch HBM.  (C, D) as b200vq_linear_supported.
 * Out-of-range indices write NaN rows and set *bad_flag (device int, nullable), as b200vq_decode.
 */
int b200vq_decode_project(const int64_t* idx, int64_t n, const float* weight, int64_t K, int D, int l2norm,
                          const float* w_out, const float* b_out, int C, const float* add, int64_t add_rows,
                          float* out, int64_t ldo, int* bad_flag, b200vq_stream_t stream);

/*
 * Trainer code-usage statistics from the fused histograms
 * (larp_tokenizer_trainer.py:335-352, utils/statistics.py:5-32), on device:
 *   out[0]=index_usage  out[1]=index_usage_batch  out[2]=kl_uni  out[3]=perplexity
 */
int b200vq_usage_stats(const int32_t* hist_first, const int32_t* hist_batch, int64_t K,
                       float* out4, b200vq_stream_t stream);

/*
 * Codebook-sharded helper: out[i] = emb_local[idx_i - k_offset] if this rank owns idx_i else 0
 * (rows are then summed across ranks).  keys are the all-reduced keys.
 */
int b200vq_gather_owned(const uint64_t* keys, int64_t N, const float* emb_local, int64_t K_local,
                        int64_t k_offset, int D, float* out, b200vq_stream_t stream);

/*
 * The Bottleneck's two skinny linears and their backward (bottleneck.py:109-110 in_linear C->D,
 * out_linear D->C; :140-142 project_in, :162-164 project_out), fp32 in / fp32 out, D in
 * {8,16,24,32}, C a multiple of 16.  HBM-streaming kernels: every activation byte is touched
 * once, the weight matrix sits in shared memory.
 *   b200vq_linear_supported  1 if (C, D) can take these kernels (else use the framework's GEMM)
 *   b200vq_linear_in    y (N,D) = x (N,C; row stride ldx) w^T + bias, w (D,C); 3xTF32 (fp32-class
 *                       accuracy: y decides the code indices).  Also grad of out_linear's input
 *                       (x = grad_output, w = out_linear.weight^T).
 *   b200vq_linear_out   y (N,C; row stride ldy) = a (N,D) w'^T + bias (nullable); w' is (C,D)
 *                       when w_is_dc == 0 (out_linear.weight) or (D,C) when w_is_dc == 1
 *                       (in_linear.weight: grad_x = grad_z W_in).  TF32, fp32 accumulate.
 *   b200vq_weight_grad  dw = x^T y  (x (N,C; ldx), y (N,D)), written as (C,D) or, dw_is_dc == 1,
 *                       (D,C); colsum (C, nullable) = column sums of x (bias grad).  Outputs are
 *                       zeroed inside; accumulation across CTAs uses fp32 atomics.
 *   b200vq_weight_grad_ws  the same result without atomics: every CTA writes its partial (C*D + C floats)
 *                       to its own slab of `workspace` (b200vq_weight_grad_workspace_bytes(C, D), 16-byte
 *                       aligned) and a second kernel adds the slabs in a fixed order (run-to-run
 *                       deterministic, and faster: no contended atomics).  B200VQ_EINVAL if the
 *                       workspace is too small.
 *   b200vq_colsum       out (D) = column sums of y (N,D), D <= 32 (in_linear.bias.grad)
 */
int b200vq_linear_supported(int C, int D);
int b200vq_linear_in(const float* x, int64_t ldx, int64_t N, int C, const float* w, const float* bias, int D,
                     float* y, b200vq_stream_t stream);
int b200vq_linear_out(const float* a, int64_t N, int D, const float* w, int w_is_dc, const float* bias, int C,
                      float* y, int64_t ldy, b200vq_stream_t stream);
int b200vq_weight_grad(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw,
                       int dw_is_dc, float* colsum, b200vq_stream_t stream);
size_t b200vq_weight_grad_workspace_bytes(int C, int D);
int b200vq_weight_grad_ws(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw,
                          int dw_is_dc, float* colsum, void* workspace, size_t workspace_bytes,
                          b200vq_stream_t stream);
/*
 * b200vq_weight_grad_noalloc -- b200vq_weight_grad_ws without a workspace argument and without any allocation.
 * No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *   - same arguments, kernels, path selection and results as b200vq_weight_grad_ws; the per-CTA slabs live in a
 *     static device array of B200VQ_NOALLOC_SCRATCH_BYTES (24 MiB) inside the library, held by every GPU that
 *     loads it until the process exits.  The call makes no cudaMalloc / cudaFree / allocating library call.
 *   - (C, D) with b200vq_weight_grad_workspace_bytes(C, D) above that limit (e.g. C = 1024 with D >= 24 on a
 *     148-SM part) are refused with B200VQ_ENOALLOC_SIZE; N is not limited.
 *   - every call on a device shares the one array: calls that overlap on the GPU (different streams, or graphs
 *     replayed concurrently) race on it -- keep them on one stream or order them with events.  The array keeps
 *     the previous call's slabs; nothing has to start at zero, because each slab the reduction reads is written
 *     in full by its CTA in the same call.
 *   - the array's device address is resolved with cudaGetSymbolAddress on the first call per device (with the
 *     caller's device current) and cached; make that first call outside stream capture.
 */
#define B200VQ_NOALLOC_SCRATCH_BYTES (24u << 20)
int b200vq_weight_grad_noalloc(const float* x, int64_t ldx, const float* y, int64_t N, int C, int D, float* dw,
                               int dw_is_dc, float* colsum, b200vq_stream_t stream);
/*
 * Backward of out_linear in ONE pass over grad_output g (N,C; row stride ldg), instead of
 * b200vq_linear_in (grad of the input) + b200vq_weight_grad_ws (weight / bias grads) reading it twice:
 *   ga (N,D) = g w_out (+ ga_add, nullable (N,D));  dw (C,D) = g^T a;  colsum (C, nullable) = column sums of g
 * a (N,D) = out_linear's input ('regularized_z'), w_out (C,D).  Needs C in {256,512,768,1024} and the
 * weight matrix + a 96 KiB ring in shared memory (b200vq_grad_out_supported); workspace as
 * b200vq_weight_grad_workspace_bytes(C, D).  TF32 operands, fp32 accumulate, no atomics.
 */
int b200vq_grad_out_supported(int C, int D);
int b200vq_grad_out(const float* g, int64_t ldg, const float* a, const float* w_out, int64_t N, int C, int D,
                    const float* ga_add, float* ga, float* dw, float* colsum, void* workspace,
                    size_t workspace_bytes, b200vq_stream_t stream);
int b200vq_colsum(const float* y, int64_t N, int D, float* out, b200vq_stream_t stream);

/*
 * Whole-Bottleneck step in ONE host call per direction (Bottleneck.forward, bottleneck.py:170-188:
 * in_linear -> quantizer -> out_linear, plus the trainer's usage statistics; and its autograd).
 * Same kernels as the piecewise entry points above, enqueued back to back on `stream`; exists so
 * that a training step costs two host calls instead of ~40.  fp32 x, b200vq_linear_supported(C, D).
 *   forward  outputs: z (N,D) 'projected_z'; emb (K,D) + wnorm (K) (saved for backward); znorm, ste (N,D);
 *            idx (N) int64; zlen (N); hist, hist_first (K) int32; losses (3); usage4 (4, nullable:
 *            index_usage, index_usage_batch, kl_uni, perplexity); norms2 (2, nullable: batch means of
 *            |x[b,0,:]| and |x[b,n_first-1,:]|, the 'input_norm_first/last' diagnostics; rows of one sample
 *            are the n_first consecutive rows); out (N,C; row stride ldo) 'output'
 *   backward inputs: g_out (N,C; ldg) grad of 'output' (nullable); g_losses (3, device, nullable);
 *            g_ste / g_znorm / g_z (N,D) and g_emb (K,D): extra grads on 'regularized_z',
 *            'unregularized_z', 'projected_z', 'emb' (all nullable); the tensors saved by forward
 *   backward outputs (each nullable = not needed): grad_x (N,C; ldgx), grad_w_in (D,C), grad_b_in (D),
 *            grad_weight (K,D), grad_w_out (C,D), grad_b_out (C)
 *   workspace >= b200vq_bottleneck_workspace_bytes(N,C,D,K), 16-byte aligned, same buffer both ways
 *   act_dtype (B200VQ_DT_*): element type of z ('projected_z'), out ('output') and g_out -- fp32, or the caller's
 *            autocast dtype (bf16 / fp16): under torch.autocast the reference's nn.Linears hand back that dtype while
 *            x, the quantizer's tensors and every gradient this call returns stay fp32 (SURVEY.md 8b "dtype flow under
 *            AMP"; trainers/larp_tokenizer_trainer.py:218-227, 244).  The contractions keep fp32 accumulation; the 2-byte
 *            outputs are rounded once, from the fp32 result.
 *   add (forward, nullable; add_rows rows of C fp32): a table added to 'output' in the out_linear store, row
 *            (token index % add_rows) -- the tokenizer's `z + decoder_latent_pe` (models/larp_tokenizer.py:463-464) for
 *            callers that want it folded into this pass.
 *   philox_dev (forward, nullable): DEVICE {seed, offset} (2 x uint64).  When given (stochastic mode on the
 *            tensor-core path only) it overrides seed / offset and the library adds 4 to the offset on the device
 *            after the draw (a one-thread kernel) -- the host-side generator advance of functional.philox_state, moved to the GPU so that a
 *            captured CUDA graph draws fresh noise on every replay (SURVEY.md 8b "Threading").
 */
size_t b200vq_bottleneck_workspace_bytes(int64_t N, int C, int D, int64_t K);
int b200vq_bottleneck_forward(const float* x, int64_t ldx, int64_t N, int C, const float* w_in, const float* b_in,
                              const float* weight, int64_t K, int D, const float* w_out, const float* b_out, int mode,
                              int l2norm, float t_inv, uint64_t seed, uint64_t offset, int64_t n_first, float beta,
                              float w_cb, int use_tc, void* z, float* emb, float* wnorm, float* znorm, float* ste,
                              int64_t* idx, float* zlen, int32_t* hist, int32_t* hist_first, float* losses, float* usage4,
                              float* norms2, void* out, int64_t ldo, void* workspace, size_t workspace_bytes,
                              uint64_t* philox_dev, int act_dtype, const float* add, int64_t add_rows,
                              b200vq_stream_t stream);
int b200vq_bottleneck_backward(const void* g_out, int act_dtype, int64_t ldg, const float* g_losses, const float* g_ste,
                               const float* g_znorm, const float* g_emb, const float* g_z, const float* x, int64_t ldx,
                               int64_t N, int C, const float* w_in, const float* w_out, const float* emb,
                               const float* wnorm, const float* znorm, const float* zlen, const int64_t* idx,
                               const float* ste, int64_t K, int D, int l2norm, float beta, float w_cb, float* grad_x,
                               int64_t ldgx, float* grad_w_in, float* grad_b_in, float* grad_weight, float* grad_w_out,
                               float* grad_b_out, void* workspace, size_t workspace_bytes, void* grads_ready_event,
                               b200vq_stream_t stream);

/* grads_ready_event (cudaEvent_t, nullable): recorded once every PARAMETER gradient (grad_w_in, grad_b_in,
 * grad_weight, grad_w_out, grad_b_out) has been enqueued and BEFORE the grad_x pass, so that a data-parallel caller can
 * start its gradient all-reduce on another stream while grad_x (one pass of stores over (N, C)) is still being written
 * (the reference's DDP overlaps bucket all-reduces with backward the same way, trainers/base_trainer.py:388). */

/* out3 = {*a, *b, *c} with 0 for null pointers: packs the upstream gradients of loss_q / loss_commit /
 * loss_codebook (three separate 0-dim tensors under autograd) into the g_losses vector of the backward calls. */
int b200vq_pack3(const float* a, const float* b, const float* c, float* out3, b200vq_stream_t stream);

/* Streaming entropy regulariser: reference `entropy_loss(-d, temperature=T_e)` (models/bottleneck.py:12-33 as
 * called at :300-303; the "softmax" loss type, the only one the call site uses) without the N x K affinity,
 * softmax and log-softmax matrices.  z (N,D): the latents the reference ranks with (l2-normalised when
 * l2_normalized, i.e. 'unregularized_z'); emb (K,D): 'emb'; both fp32, contiguous, 16-byte aligned; t_inv = 1/T_e.
 *   forward : out3 = {loss_entropy, per_sample_entropy, codebook_entropy};  saves lse (N), ent (N) =
 *             sum_k p log p per row, w (K) = d codebook_entropy / d mean-prob; psum (K) and ent_sum (1 double)
 *             are scratch; `scratch` (>= b200vq_entropy_scratch_bytes(N,K), 16-byte aligned) holds the per-split
 *             row statistics when few rows meet many codes and the code stream is split across CTAs.
 *   backward: g3 = device {g_loss_entropy, g_per_sample_entropy, g_codebook_entropy}; c (N) scratch;
 *             grad_z (N,D) / grad_emb (K,D) written (each nullable = not needed).
 * b200vq_entropy_supported(D): D in {8,16,24,32,64}. */
int b200vq_entropy_supported(int D);
size_t b200vq_entropy_scratch_bytes(int64_t N, int64_t K);
int b200vq_entropy_forward(const float* z, int64_t N, int D, const float* emb, int64_t K, float t_inv, float* lse,
                           float* ent, float* psum, float* w, double* ent_sum, float* out3, void* scratch,
                           size_t scratch_bytes, b200vq_stream_t stream);
int b200vq_entropy_backward(const float* z, int64_t N, int D, const float* emb, int64_t K, float t_inv,
                            const float* lse, const float* ent, const float* w, const float* g3, float* c,
                            float* grad_z, float* grad_emb, b200vq_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200VQ_H_ */
