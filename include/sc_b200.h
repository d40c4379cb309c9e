/* sc_b200.h - C ABI of the B200 (sm_100a) engine for ensemble sparse-autoencoder training.
 *
 * The reference (johnathan217/sparse_coding_) is pure Python/PyTorch and has no FFI of its own;
 * the "operator API" of this path is the Python duck type consumed by
 * autoencoders/ensemble.py:68-193.  Each entry point below replaces the torch.vmap / torchopt /
 * cuBLAS call chain of the cited reference lines.  The Python mirror in sparse_coding__b200/
 * binds these with ctypes (see INTEGRATION.md for the stub a reference maintainer would add).
 *
 * Conventions
 *   - every pointer is a DEVICE pointer to contiguous row-major fp32 unless stated otherwise,
 *     16-byte aligned; d and N must be multiples of 4 (TMA strides are multiples of 16 bytes)
 *   - the caller owns every buffer, including the workspace; the library allocates nothing on
 *     the device and only caches immutable per-process state (function attributes)
 *   - calls enqueue work on `stream` (a cudaStream_t) and return without synchronising
 *   - return 0 = ok, < 0 = argument error caught before any launch, > 0 = cudaError_t;
 *     sc_last_error() gives the message for the calling thread
 *   - M = ensemble members, B = batch rows, d = activation width, N = dictionary size
 */
#ifndef SC_B200_H
#define SC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

int sc_version(void);
const char* sc_last_error(void);

/* Adam hyper-parameters as torchopt 0.7.1 `adam(lr, betas, eps, eps_root)` takes them
 * (reference call sites ensemble.py:94-95,123,182).  bc1 = 1 - b1^t and bc2 = 1 - b2^t are the
 * bias-correction denominators: the reference never advances its stored step count
 * (ensemble.py:185-189 copies into a clone), so reference-faithful callers pass t = 1 forever. */
typedef struct {
    float lr, b1, b2, eps, eps_root;
    float bc1, bc2;
} ScAdam;

/* One FunctionalEnsemble.step_batch (ensemble.py:175-193) for sig in
 * {FunctionalSAE (sae_ensemble.py:13-77), FunctionalFista (fista.py:18-84, same loss),
 *  FunctionalTiedSAE with identity or pre-applied centering (sae_ensemble.py:80-160)}:
 * forward, loss terms, backward and the in-place Adam update of every member. */
typedef struct {
    int32_t M, B, d, N;
    int32_t tied;           /* 0: encoder + bias + decoder, 1: encoder + bias, D = rownorm(encoder) both ways */
    int32_t x_per_member;   /* 0: x is [B,d] shared by all members (expand_dims=True), 1: x is [M,B,d] */
    const float* x;
    float *enc, *enc_mu, *enc_nu;    /* [M,N,d] params["encoder"] and its Adam moments */
    float *bias, *bias_mu, *bias_nu; /* [M,N]   params["encoder_bias"] */
    float *dec, *dec_mu, *dec_nu;    /* [M,N,d] params["decoder"] (ignored when tied) */
    const float* l1_alpha;           /* [M] buffers["l1_alpha"] */
    const float* bias_decay;         /* [M] buffers["bias_decay"] */
    ScAdam adam;
    float* out_losses;      /* [4,M]: loss, l_reconstruction, l_l1, l_bias_decay (sae_ensemble.py:66-71) */
    float* out_c;           /* [M,B,N] aux "c" (sae_ensemble.py:73-75); values are TF32-rounded, an entry
                               whose pre-activation was negative is -0.0f */
    void* workspace;        /* >= sc_sae_step_workspace_bytes(...) bytes, 1024-byte aligned */
    size_t workspace_bytes;
    void* stream;
    /* Gradient-only mode (data-parallel training of one large dictionary, reference analogue
     * experiments/huge_batch_size.py:259-335 DDP): when grad_enc is non-NULL the gradients of the total
     * loss are WRITTEN to grad_enc / grad_bias / grad_dec (grad_dec ignored when tied) and no Adam
     * update happens; the caller all-reduces them and calls sc_adam_apply.  B_total > 0 replaces B in
     * the 1/B and 1/(B d) loss normalisations so that per-rank results add up to the global-batch ones. */
    float *grad_enc, *grad_bias, *grad_dec;
    int32_t B_total;
    /* operand_format 0: GEMM operands and c / R / dz in fp32 storage rounded to TF32 (kind::tf32).
     * operand_format 1: the same tensors in fp16 (kind::f16, same 11-bit significand, twice the MMA rate, half
     * the bytes; residual and dz kept unscaled so nothing underflows); needs sc_sae_step_f16_supported(B,d,N).
     * In format 1 the engine keeps the codes TRANSPOSED: out_c_half (required) is c^T = [M][N][ldb] fp16 with
     * ldb = B rounded up to a multiple of 8 (columns >= B are never written or read); out_c may be NULL (when
     * given it is filled from the fp16 codes, as [M,B,N] fp32; sc_codes_to_float does the same later).
     * Inputs with |x| > 65504 overflow fp16: use format 0. */
    int32_t operand_format;
    void* out_c_half;
    /* Masked ensembles (FunctionalMaskedSAE / FunctionalMaskedTiedSAE, sae_ensemble.py:307-442): member m only
     * uses its first dict_size[m] features; the rest produce zero codes and zero gradients.  [M] int32 on the
     * device, or NULL for unmasked. */
    const int32_t* dict_size;
    /* Fused reduce-scatter for data-parallel training of ONE dictionary (M == 1; replaces the NCCL all-reduce of
     * the gradient-only mode; reference analogue: DDP in experiments/huge_batch_size.py:259-335).  rs_world > 0:
     * rank r owns dictionary rows [r * rs_rows, (r + 1) * rs_rows), rs_rows = N / rs_world a multiple of 32.  The
     * weight-gradient kernel adds every 32 x 32 tile of this rank's gradient straight into the OWNER's
     * accumulation buffer over NVLink (TMA reduce-add to peer memory): rs_grad_enc[r] / rs_grad_dec[r] point at rank
     * r's [rs_rows, d] buffers, rs_grad_bias[r] at its [rs_rows] buffer, rs_losses[r] at its [3] buffer (sum over
     * ranks of l_reconstruction, l_l1, l_bias_decay; every rank adds to every rank's).  All must be zero before
     * the step, peer-mapped in this process, and every rank must pass the same lists.  No Adam update happens here:
     * after a cross-rank barrier the owner calls sc_dp_adam_broadcast on its slice.  grad_* are ignored. */
    int32_t rs_world;
    int32_t rs_rows;
    float* rs_grad_enc[8];
    float* rs_grad_dec[8];
    float* rs_grad_bias[8];
    float* rs_losses[8];
    /* operand_format 1, untied, full update (no grad_* / rs_*): the Adam epilogue leaves fp16 copies of the UPDATED
     * encoder / decoder and the decoder's row norms in the workspace, so the next step does not have to re-read the
     * fp32 weights.  operands_valid = 1 tells the step that this hand-over is intact: same workspace, same M, d, N, and
     * nothing but sc_sae_step has written the parameters since the previous full step; 0 (always safe) rebuilds the
     * copies.  The library itself remembers per workspace whether its last fp16 step ran in this mode (it is used up
     * to 3072 rows per step; larger batches run the preparation pass, which is cheaper there); an operand_format 0 step
     * in the same workspace overwrites the copies, pass 0 after one.
     * In this mode the stored codes are c / ||W_d|| per feature; out_code_scale ([M,N] floats; operand_format 1 only;
     * required with out_c, otherwise optional) receives the factor that sc_codes_to_float applies to the stored codes:
     * ||W_d|| here, 1 in every other mode. */
    int32_t operands_valid;
    float* out_code_scale;
    /* operand_format 1 range guard: fp16 holds |x| <= 65504.  Every conversion to fp16 in the engine saturates (no
     * infinity is ever stored, so an outlier costs accuracy on that element instead of turning the step into NaNs),
     * and when the batch holds a value beyond the range (or a non-finite one) the step does
     * atomicMax(*range_flag, step_tag).  range_flag: device int32 or NULL; the caller reads it when it likes (no sync
     * in the step) and switches to operand_format 0 - the host mirror does that in its "auto" mode. */
    int32_t* range_flag;
    int32_t step_tag;
} ScStepArgs;

size_t sc_sae_step_workspace_bytes(int32_t M, int32_t B, int32_t d, int32_t N, int32_t tied, int32_t x_per_member);
int sc_sae_step(const ScStepArgs* args);
int sc_sae_step_f16_supported(int32_t B, int32_t d, int32_t N);
/* aux "c" of step_batch (sae_ensemble.py:73-75) from the fp16-mode codes: out [M,B,N] fp32 from codes_t =
 * ScStepArgs.out_c_half ([M][N][ldb] fp16, ldb = B rounded up to 8) times code_scale (ScStepArgs.out_code_scale, [M,N],
 * or NULL for 1); entries with a negative pre-activation are -0. */
int sc_codes_to_float(const void* codes_t, const float* code_scale, float* out, int32_t M, int32_t B, int32_t N,
                      void* stream);

/* Elementwise Adam exactly as the fused epilogues apply it (torchopt 0.7.1 semantics, see ScAdam):
 * p, mu, nu updated in place from gradient g, n elements (multiple of 4). */
int sc_adam_apply(float* p, float* mu, float* nu, const float* g, int64_t n, const ScAdam* adam, void* stream);

/* LearnedDict.encode (learned_dict.py:127-131 untied, :186-196 tied) for G dictionaries at once:
 * c[g] = relu(x W[g]^T + b[g]).  W is used as given (the tied caller passes the row-normalised
 * dictionary).  Optionally counts, per feature, the rows with c != 0
 * (standard_metrics.py:303-306 mean_nonzero_activations, :439-452 n-ever-active). */
typedef struct {
    int32_t G, B, d, N;
    int32_t x_per_group;
    const float* x;         /* [B,d] or [G,B,d] */
    const float* W;         /* [G,N,d] */
    const float* bias;      /* [G,N] or NULL */
    float* out_c;           /* [G,B,N] or NULL (statistics only) */
    float* out_nonzero;     /* [G,N] count of rows with c != 0, accumulated (+=) into the buffer, or NULL */
    float* out_moments;     /* [G,4,N] sums of c, c^2, c^3, c^4 over rows, accumulated, or NULL
                               (standard_metrics.py:480-509 calc_moments_streaming) */
    void* workspace;
    size_t workspace_bytes;
    void* stream;
    int32_t operand_format; /* 0: TF32 operands; 1: fp16 operands (same 11-bit significand, twice the MMA rate);
                               needs sc_infer_f16_supported(B, d, N, transposed = statistics requested) */
    void* out_c_half;       /* [G,B,N] fp16 codes or NULL (operand_format 1, no statistics): the form sc_decode's c_half
                               takes, so that predict / FVU never materialise fp32 codes (learned_dict.py:42-47) */
} ScEncodeArgs;
size_t sc_encode_workspace_bytes(int32_t G, int32_t B, int32_t d, int32_t N, int32_t x_per_group);
int sc_encode(const ScEncodeArgs* args);

/* LearnedDict.decode (learned_dict.py:29-32): x_hat[g] = c[g] @ rownorm(Wd[g]) (normalise = 1) or
 * c[g] @ Wd[g] (normalise = 0).  With `target` also returns sum((x_hat - target)^2) per group,
 * the numerator of fraction_variance_unexplained (standard_metrics.py:308-312). */
typedef struct {
    int32_t G, B, d, N;
    int32_t normalise;
    int32_t target_per_group;
    const float* c;         /* [G,B,N] */
    const float* Wd;        /* [G,N,d] */
    const float* target;    /* [B,d] or [G,B,d] or NULL */
    float* out_x_hat;       /* [G,B,d] or NULL */
    float* out_sq_err;      /* [G] or NULL */
    void* workspace;
    size_t workspace_bytes;
    void* stream;
    int32_t operand_format; /* as in ScEncodeArgs; the codes are converted to fp16 (|c| <= 65504) */
    const void* c_half;     /* [G,B,N] fp16 codes from sc_encode's out_c_half, used instead of c (operand_format 1) */
} ScDecodeArgs;
size_t sc_decode_workspace_bytes(int32_t G, int32_t B, int32_t d, int32_t N);
/* 1 when sc_encode / sc_decode can run with fp16 operands for this shape (d, N multiples of 64 and >= 128;
 * the transposed statistics variant of sc_encode also needs B > 64). */
int sc_infer_f16_supported(int32_t B, int32_t d, int32_t N, int32_t transposed);
int sc_decode(const ScDecodeArgs* args);

/* FunctionalFista.fista (fista.py:99-128; method twin :274-301): num_iter FISTA iterations of
 * min_a 0.5||x - a D||^2 + l1 ||a||_1, a >= 0, from the warm start a0.  Products run as split
 * 3xTF32 (hi*hi + hi*lo + lo*hi) so the codes agree with the fp32 reference to ~1e-6 relative.
 * eta <= 0 computes eta = 1/lambda_max(D^T D) by power iteration (fista.py:105-106 takes it from
 * eigvalsh(D D^T): same value).  momentum[k] = (t_k - 1)/t_{k+1}, computed by the caller in
 * float64 exactly as fista.py:117-118 does. */
typedef struct {
    int32_t B, d, N, num_iter;
    const float* x;         /* [B,d] */
    const float* D;         /* [N,d] dictionary, used as given */
    const float* a0;        /* [B,N] warm start */
    float l1_coef;
    float eta;              /* <= 0: compute */
    const float* momentum;  /* HOST pointer, [num_iter] */
    float* out_a;           /* [B,N] */
    float* out_res;         /* [B,d] x - a D */
    float* out_eta;         /* device scalar, the step size used, or NULL */
    void* workspace;
    size_t workspace_bytes;
    void* stream;
    const float* l1_coef_dev; /* device scalar or NULL: when given it replaces l1_coef, so a caller that holds the
                                 coefficient as a device tensor (buffers["l1_alpha"], big_sweep.py:186) need not sync */
    int32_t pair_format;      /* 0: fp16 (hi, lo) operand pairs when d and N are multiples of 64 (values must stay below
                                 65504 in magnitude: codes are bounded by the row norms of x for a unit-row dictionary),
                                 else TF32 pairs; 1: always TF32 pairs (fp32 range) */
} ScFistaArgs;
size_t sc_fista_workspace_bytes(int32_t B, int32_t d, int32_t N);
int sc_fista(const ScFistaArgs* args);

/* LearnedDict.get_learned_dict (learned_dict.py:118-120, :172-174): out[g][n][:] = W[g][n][:] /
 * max(||W[g][n]||_2, 1e-8), full fp32. */
int sc_row_normalise(const float* W, float* out, int32_t G, int32_t N, int32_t d, void* stream);

/* Denominator of fraction_variance_unexplained (standard_metrics.py:311): *out_sum =
 * sum_{b,j} (x[b][j] - mean_b x[b][j])^2 for x [B,d].  workspace: at least (d + 1) floats. */
int sc_column_variance(const float* x, float* out_sum, int32_t B, int32_t d, void* workspace, void* stream);

/* Batch assembly from an activation chunk resident in HBM - replaces `dataset[batch_idxs].to(device)` of
 * the reference train loops (big_sweep.py:170, basic_l1_sweep.py:103, cluster_runs.py:37) and the chunk
 * centring of big_sweep.py:402-407:  out[b][:] = float(chunk[idx[b]][:]) - mean[:].
 * chunk is [n_rows][d] in its storage format: fp16 (elem_bytes 2, as activation_dataset.py:393-397 writes
 * it; d % 8 == 0) or fp32 (elem_bytes 4; d % 4 == 0).  idx: B int64 row numbers on the device, negative
 * values count from the end as in torch indexing.  mean: [d] fp32 or NULL.  A row number outside
 * [-n_rows, n_rows) cannot raise from a kernel: that output row is zero-filled and *err_flag (device int,
 * may be NULL) is OR-ed with 1 - the host mirror checks the flag and raises IndexError like the reference. */
typedef struct {
    const void* chunk;
    int32_t elem_bytes;
    int64_t n_rows;
    int32_t d;
    const int64_t* idx;
    int32_t B;
    const float* mean;
    float* out; /* [B][d] fp32 */
    int32_t* err_flag;
    void* stream;
} ScGatherArgs;
int sc_gather_rows(const ScGatherArgs* args);

/* Column means of a resident chunk (`chunk.mean(dim=0)`, big_sweep.py:402-405): out_mean[j] =
 * (1/n_rows) sum_r float(chunk[r][j]), fp32 accumulation in a fixed order (deterministic). */
size_t sc_column_mean_workspace_bytes(int32_t d);
int sc_column_mean(const void* chunk, int32_t elem_bytes, int64_t n_rows, int32_t d, float* out_mean, void* workspace,
                   void* stream);

/* Owner-side half of the fused data-parallel step: Adam on n contiguous elements whose summed gradient sits in
 * `grad` (zeroed afterwards for the next step), moments updated in place, and the NEW parameter values stored to
 * `n_peers` destinations - this rank's own copy and every peer's copy of the same slice (peer-mapped pointers), so
 * no all-gather follows.  n must be a multiple of 4, pointers 16-byte aligned. */
int sc_dp_adam_broadcast(float* grad, float* mu, float* nu, const float* p_local, float* const* p_dst, int32_t n_peers,
                         int64_t n, const ScAdam* adam, void* stream);

/* Encode for callers that cannot allocate device memory.  (No counterpart in the reference: it is
 * an addition of this library.)  Same result as sc_encode with G = 1 and only out_c requested -
 * c = relu(x W^T + bias), x [B,d], W [N,d], bias [N] or NULL, out_c [B,N] - but the TF32-rounded
 * operand copies live in a fixed array inside the library image instead of a caller workspace, so the
 * call makes no cudaMalloc / cudaFree and needs no allocator: meant for a monitoring or serving loop
 * that has given all device memory to its own pool, or cannot take cudaMalloc's device-wide
 * synchronisation.  Limit: pad256(B*d) + pad256(N*d) <= SC_NOALLOC_SCRATCH_FLOATS, else the call
 * returns -4 with a message and launches nothing.  Every GPU that loads the library holds the array
 * (16 MiB) until exit.  The first call on each device looks up the array's address
 * (cudaGetSymbolAddress) - make that call before any stream capture.  All calls share the array and
 * simply overwrite it; calls that overlap on the GPU race on it, so serialise them. */
#define SC_NOALLOC_SCRATCH_FLOATS (4 * 1024 * 1024)
int sc_encode_noalloc(const float* x, const float* W, const float* bias, float* out_c, int32_t B, int32_t d, int32_t N,
                      void* stream);

/* Per-launch timing for bench.py: between sc_profile_begin() and sc_profile_end() every kernel the
 * library launches is bracketed by CUDA events on the caller's stream.  sc_profile_end synchronises and
 * returns up to `cap` (kernel name, milliseconds) pairs in launch order. */
int sc_profile_begin(void);
int sc_profile_end(const char** names, float* ms, int cap);

/* Debug / validation hook: C[g] = A[g] B[g]^T through the tcgen05 core with plain store epilogue.
 * a_mn / b_mn select MN-major storage ([K,M] / [K,N]) for the operand.  Inputs are used as given
 * (the tensor core truncates to TF32). */
int sc_gemm_tf32(const float* A, const float* Bm, float* C, int32_t G, int32_t M, int32_t N, int32_t K,
                 int32_t a_mn, int32_t b_mn, void* stream);

/* C[g] = A[g] B[g]^T at fp32 accuracy for the small dense products AROUND the hot path that the reference does with
 * torch.einsum / torch.mm in fp32: the centring transform of FunctionalTiedSAE / TiedSAE (sae_ensemble.py:126-131,
 * learned_dict.py:176-184: (x - t) Rot^T and its inverse) and the dictionary update of fista.py:131-138 (Res^T ahat).
 * Operands are split into TF32 hi + lo parts and multiplied as hi*lo + lo*hi + hi*hi into one accumulator (relative
 * 2^-21).  Layout flags and alignment rules as sc_gemm_tf32; workspace holds the four split copies. */
size_t sc_gemm_f32_workspace_bytes(int32_t G, int32_t M, int32_t N, int32_t K);
int sc_gemm_f32(const float* A, const float* Bm, float* C, int32_t G, int32_t M, int32_t N, int32_t K, int32_t a_mn,
                int32_t b_mn, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SC_B200_H */
