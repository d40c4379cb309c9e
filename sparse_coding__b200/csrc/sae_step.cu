// sc_sae_step: one FunctionalEnsemble.step_batch (reference autoencoders/ensemble.py:175-193) for
// the SAE / Fista / TiedSAE signatures, TF32-operand mode: seven launches on the caller's stream:
//   round x -> prep params -> ||b|| -> G1 encode -> G2 decode+MSE -> G3 dL/dz -> G4/G5 dW with fused Adam
//   (the finalize work - loss scalars + bias Adam - runs on that launch's idle warps).
// The default fp16-operand mode is sae_step_h.cu (five launches).
// Five tcgen05 TF32 GEMMs per member carry 10 B N d flops; everything else is fused epilogue work.
#include "../../include/sc_b200.h"
#include "host_common.cuh"
#include "sae_epilogues.cuh"

namespace sc {

thread_local char g_last_error[512] = "";
thread_local int g_dbg_sm_limit = 0;
thread_local int g_dbg_phase = 3;
Profiler g_prof;

// ------------------------------------------------------------------------------------------------
// out = rn_tf32(in), elementwise (the activation batch; n multiple of 4, 16-byte aligned)
// ------------------------------------------------------------------------------------------------
__global__ void round_copy_kernel(const float4* __restrict__ in, float4* __restrict__ out, long long n4) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 v = in[i];
        v.x = rn_tf32(v.x); v.y = rn_tf32(v.y); v.z = rn_tf32(v.z); v.w = rn_tf32(v.w);
        out[i] = v;
    }
}

// ------------------------------------------------------------------------------------------------
// Per-step operand preparation, one warp per dictionary row (grid: ceil(N/8) x G, 256 threads).
//   enc_r = rn_tf32(enc)                                   (when enc_r != nullptr)
//   D_r   = rn_tf32(W / max(||W||_2, 1e-8)) per row        (sae_ensemble.py:58-59 / :135-136)
//   inv_norm = 1 / max(||W||, 1e-8), stored NEGATIVE when the clamp is active
// ------------------------------------------------------------------------------------------------
__global__ void prep_params_kernel(const float* __restrict__ enc, float* __restrict__ enc_r,
                                   const float* __restrict__ W, float* __restrict__ D_r,
                                   float* __restrict__ inv_norm, int N, int d, int normalise) {
    pdl_launch_dependents();
    pdl_wait();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * 8 + warp;
    if (n >= N) return;
    const long long row = ((long long)blockIdx.y * N + n) * d;
    const int d4 = d >> 2;
    if (enc_r) {
        const float4* src = reinterpret_cast<const float4*>(enc + row);
        float4* dst = reinterpret_cast<float4*>(enc_r + row);
        for (int i = lane; i < d4; i += 32) {
            float4 v = src[i];
            v.x = rn_tf32(v.x); v.y = rn_tf32(v.y); v.z = rn_tf32(v.z); v.w = rn_tf32(v.w);
            dst[i] = v;
        }
    }
    if (!D_r) return;
    const float4* src = reinterpret_cast<const float4*>(W + row);
    float inv = 1.f;
    if (normalise) {
        float ss = 0.f;
        for (int i = lane; i < d4; i += 32) {
            const float4 v = src[i];
            ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
        ss = warp_sum(ss);
        const float nrm = sqrtf(ss);
        const float cl = fmaxf(nrm, 1e-8f);
        inv = 1.f / cl;
        if (lane == 0 && inv_norm) inv_norm[(long long)blockIdx.y * N + n] = nrm < 1e-8f ? -inv : inv;
    }
    float4* dst = reinterpret_cast<float4*>(D_r + row);
    for (int i = lane; i < d4; i += 32) {
        float4 v = src[i];
        v.x = rn_tf32(v.x * inv); v.y = rn_tf32(v.y * inv); v.z = rn_tf32(v.z * inv); v.w = rn_tf32(v.w * inv);
        dst[i] = v;
    }
}

// deterministic block-wide sum (256 threads)
SC_DEVICE float block_sum(float v, float* sh) {
    v = warp_sum(v);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sh[i];
    return t;
}

// ------------------------------------------------------------------------------------------------
// Finalize, one block per member: loss scalars (sae_ensemble.py:62-71) from the per-warp partials
// of G1/G2, then the bias gradient db + beta b/||b|| (zero sub-gradient at b == 0) and its Adam step.
// ------------------------------------------------------------------------------------------------
// elementwise Adam for the data-parallel path (gradients all-reduced by the caller)
__global__ void adam_apply_kernel(float4* __restrict__ p, float4* __restrict__ mu, float4* __restrict__ nu,
                                  const float4* __restrict__ g, long long n4, AdamConst a) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 pv = p[i], mv = mu[i], nv = nu[i];
        const float4 gv = g[i];
        adam_update(pv.x, mv.x, nv.x, gv.x, a);
        adam_update(pv.y, mv.y, nv.y, gv.y, a);
        adam_update(pv.z, mv.z, nv.z, gv.z, a);
        adam_update(pv.w, mv.w, nv.w, gv.w, a);
        p[i] = pv; mu[i] = mv; nu[i] = nv;
    }
}

// Owner's Adam on its slice, new parameters stored to every rank's copy (peer-mapped), gradient buffer re-zeroed
struct PeerDst {
    float4* p[8];
    int n;
};
__global__ void dp_adam_broadcast_kernel(float4* __restrict__ g, float4* __restrict__ mu, float4* __restrict__ nu,
                                         const float4* __restrict__ p_local, PeerDst dst, long long n4, AdamConst a) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 pv = p_local[i], mv = mu[i], nv = nu[i];
        const float4 gv = g[i];
        adam_update(pv.x, mv.x, nv.x, gv.x, a);
        adam_update(pv.y, mv.y, nv.y, gv.y, a);
        adam_update(pv.z, mv.z, nv.z, gv.z, a);
        adam_update(pv.w, mv.w, nv.w, gv.w, a);
        mu[i] = mv; nu[i] = nv;
        g[i] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < 8; ++r)
            if (r < dst.n) dst.p[r][i] = pv;
    }
}

// stand-alone form of the finalize work (sae_epilogues.cuh); the fused steps run it inside the weight-gradient launch
__global__ void finalize_kernel(FinalizeArgs f) {
    pdl_launch_dependents();
    pdl_wait();
    finalize_work(f, blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), gridDim.x * (blockDim.x >> 5));
}

// ||b||_2 per member (sae_ensemble.py:64), one block per member
__global__ void bias_norm_kernel(const float* __restrict__ bias, int N, float* __restrict__ out) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    float acc = 0.f;
    for (int n = threadIdx.x; n < N; n += blockDim.x) {
        const float b = bias[(long long)blockIdx.x * N + n];
        acc += b * b;
    }
    const float t = block_sum(acc, sh);
    if (threadIdx.x == 0) out[blockIdx.x] = sqrtf(t);
}

}  // namespace sc

// ================================================================================================
// host side
// ================================================================================================
namespace sc {

size_t step_h_workspace_bytes(int M, int B, int d, int N, int tied, int xpm);  // sae_step_h.cu

struct StepWs {
    float *x_r, *enc_r, *D_r, *inv_norm, *R_r, *dz, *l1_part, *mse_part, *db_part, *s_part, *bnorm;
    int n_l1, n_mse, tiles_b;
    size_t bytes;
};

static StepWs carve_step(void* base, int M, int B, int d, int N, int tied, int xpm) {
    StepWs w;
    Carver c(base);
    const size_t Mz = M, Bz = B, dz_ = d, Nz = N;
    w.x_r = c.floats((xpm ? Mz : 1) * Bz * dz_);
    w.enc_r = tied ? nullptr : c.floats(Mz * Nz * dz_);
    w.D_r = c.floats(Mz * Nz * dz_);
    w.inv_norm = c.floats(Mz * Nz);
    w.R_r = c.floats(Mz * Bz * dz_);
    w.dz = c.floats(Mz * Bz * Nz);
    w.n_l1 = row_tiles128(B) * ceil_div(N, pick_bn(N)) * 4 * kEpiHalves;
    w.n_mse = row_tiles128(B) * ceil_div(d, pick_bn(d)) * 4 * kEpiHalves;
    w.tiles_b = ceil_div(B, pick_bn(B)) * kEpiHalves;  // partial slots per feature
    w.l1_part = c.floats(Mz * w.n_l1);
    w.mse_part = c.floats(Mz * w.n_mse);
    w.db_part = c.floats(Mz * w.tiles_b * Nz);
    w.s_part = c.floats(Mz * w.tiles_b * Nz);
    w.bnorm = c.floats(Mz);
    w.bytes = c.off;
    return w;
}

static AdamConst make_adam(const ScAdam& a) {
    AdamConst k;
    k.lr = a.lr; k.b1 = a.b1; k.b2 = a.b2;
    k.omb1 = static_cast<float>(1.0 - static_cast<double>(a.b1));
    k.omb2 = static_cast<float>(1.0 - static_cast<double>(a.b2));
    k.eps = a.eps; k.eps_root = a.eps_root; k.rbc1 = 1.0f / a.bc1; k.rbc2 = 1.0f / a.bc2;
    k.nlr_rbc1 = -a.lr * k.rbc1;
    return k;
}

static int io_map(CUtensorMap* out, const float* base, long long rows, long long cols, long long g, long long gstride) {
    const char* err = nullptr;
    const int rc = make_io_map(out, base, rows, cols, g, gstride, &err);
    if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
    return 0;
}

static int check_ptr(const void* p, const char* name) {
    if (!p) return fail(-1, "%s is NULL", name);
    if (reinterpret_cast<uintptr_t>(p) & 15) return fail(-1, "%s is not 16-byte aligned", name);
    return 0;
}

int check_rs(const ScStepArgs& a) {
    if (a.rs_world == 0) return 0;
    if (a.rs_world < 0 || a.rs_world > 8) return fail(-1, "rs_world must be in [0, 8] (got %d)", a.rs_world);
    if (a.M != 1) return fail(-1, "the fused reduce-scatter mode trains ONE dictionary (M == 1, got %d)", a.M);
    if (a.rs_rows <= 0 || (a.rs_rows & 31) || (long long)a.rs_rows * a.rs_world != a.N)
        return fail(-1, "rs_rows must be N / rs_world and a multiple of 32 (N=%d rs_world=%d rs_rows=%d)", a.N, a.rs_world, a.rs_rows);
    for (int r = 0; r < a.rs_world; ++r) {
        SC_TRY(check_ptr(a.rs_grad_enc[r], "rs_grad_enc[r]"));
        SC_TRY(check_ptr(a.rs_grad_bias[r], "rs_grad_bias[r]"));
        SC_TRY(check_ptr(a.rs_losses[r], "rs_losses[r]"));
        if (!a.tied) SC_TRY(check_ptr(a.rs_grad_dec[r], "rs_grad_dec[r]"));
    }
    return 0;
}

// reduce-scatter destinations of the weight-gradient kernel and of finalize
int fill_rs(AdamParams& e, const ScStepArgs& a) {
    e.rs_world = a.rs_world;
    e.rs_rows = a.rs_rows;
    for (int r = 0; r < a.rs_world; ++r) {
        SC_TRY(io_map(&e.rs_map[0][r], a.rs_grad_enc[r], a.rs_rows, a.d, 1, 0));
        if (!a.tied) SC_TRY(io_map(&e.rs_map[1][r], a.rs_grad_dec[r], a.rs_rows, a.d, 1, 0));
    }
    return 0;
}
RsFinal make_rs_final(const ScStepArgs& a) {
    RsFinal f{};
    f.world = a.rs_world;
    f.rows = a.rs_rows;
    for (int r = 0; r < a.rs_world; ++r) { f.bias[r] = a.rs_grad_bias[r]; f.losses[r] = a.rs_losses[r]; }
    return f;
}
FinalizeArgs make_finalize(const ScStepArgs& a, const float* l1_part, int n_l1, const float* mse_part, int n_mse,
                           const float* db_part, int tiles_b, const float* bnorm, int Bn, const AdamConst& ad, float db_scale) {
    FinalizeArgs f{};
    f.l1_part = l1_part; f.n_l1 = n_l1; f.mse_part = mse_part; f.n_mse = n_mse; f.db_part = db_part; f.tiles_b = tiles_b;
    f.bias = a.bias; f.bias_mu = a.bias_mu; f.bias_nu = a.bias_nu; f.l1_alpha = a.l1_alpha; f.bias_decay = a.bias_decay;
    f.out_losses = a.out_losses; f.M = a.M; f.B = Bn; f.d = a.d; f.N = a.N; f.a = ad;
    f.grad_bias = (a.grad_enc != nullptr && a.rs_world == 0) ? a.grad_bias : nullptr;
    f.bnorm = bnorm; f.db_scale = db_scale; f.rs = make_rs_final(a);
    return f;
}

int sae_step_f16_impl(const ScStepArgs& a);

static int sae_step_impl(const ScStepArgs& a) {
    const int M = a.M, B = a.B, d = a.d, N = a.N;
    if (M <= 0 || B <= 0 || d <= 0 || N <= 0) return fail(-1, "M, B, d, N must be positive (got %d %d %d %d)", M, B, d, N);
    if ((d & 3) || (N & 3)) return fail(-1, "d and N must be multiples of 4 (got d=%d N=%d)", d, N);
    SC_TRY(check_ptr(a.x, "x"));
    SC_TRY(check_ptr(a.enc, "enc")); SC_TRY(check_ptr(a.enc_mu, "enc_mu")); SC_TRY(check_ptr(a.enc_nu, "enc_nu"));
    SC_TRY(check_ptr(a.bias, "bias")); SC_TRY(check_ptr(a.bias_mu, "bias_mu")); SC_TRY(check_ptr(a.bias_nu, "bias_nu"));
    if (!a.tied) { SC_TRY(check_ptr(a.dec, "dec")); SC_TRY(check_ptr(a.dec_mu, "dec_mu")); SC_TRY(check_ptr(a.dec_nu, "dec_nu")); }
    if (a.operand_format == 0) SC_TRY(check_ptr(a.out_c, "out_c"));
    SC_TRY(check_ptr(a.workspace, "workspace"));
    if (!a.l1_alpha || !a.bias_decay || !a.out_losses) return fail(-1, "l1_alpha / bias_decay / out_losses is NULL");
    if (reinterpret_cast<uintptr_t>(a.workspace) & 1023) return fail(-1, "workspace must be 1024-byte aligned");
    SC_TRY(check_rs(a));
    if (a.operand_format == 1) return sae_step_f16_impl(a);
    if (a.operand_format != 0) return fail(-1, "unknown operand_format %d", a.operand_format);
    const StepWs w = carve_step(a.workspace, M, B, d, N, a.tied, a.x_per_member);
    if (a.workspace_bytes < w.bytes)
        return fail(-1, "workspace too small: %zu bytes given, %zu needed", a.workspace_bytes, w.bytes);
    cudaStream_t st = static_cast<cudaStream_t>(a.stream);
    const AdamConst ad = make_adam(a.adam);
    const int xg = a.x_per_member ? M : 1;
    const int Bn = a.B_total > 0 ? a.B_total : B;  // normalising batch size
    const bool grad_mode = a.grad_enc != nullptr || a.rs_world > 0;
    if (grad_mode && a.rs_world == 0) {
        SC_TRY(check_ptr(a.grad_enc, "grad_enc")); SC_TRY(check_ptr(a.grad_bias, "grad_bias"));
        if (!a.tied) SC_TRY(check_ptr(a.grad_dec, "grad_dec"));
    }
    const long long Bd = (long long)B * d, BN_ = (long long)B * N, Nd = (long long)N * d;

    prof_mark(st, "begin");
    // 1. x_r = rn(x)
    {
        const long long n4 = (long long)xg * Bd / 4;
        const int grid = static_cast<int>(n4 / 256 + 1 < 2048 ? n4 / 256 + 1 : 2048);
        launch_k(round_copy_kernel, dim3(grid), dim3(256), 0, st, reinterpret_cast<const float4*>(a.x), reinterpret_cast<float4*>(w.x_r), n4);
    }
    prof_mark(st, "round_x");
    // 2. enc_r, D_r, 1/||W||
    {
        dim3 grid(ceil_div(N, 8), M);
        if (a.tied) launch_k(prep_params_kernel, dim3(grid), dim3(256), 0, st, nullptr, nullptr, a.enc, w.D_r, w.inv_norm, N, d, 1);
        else launch_k(prep_params_kernel, dim3(grid), dim3(256), 0, st, a.enc, w.enc_r, a.dec, w.D_r, w.inv_norm, N, d, 1);
        SC_CUDA(cudaGetLastError(), "prep launch");
    }
    launch_k(bias_norm_kernel, dim3(M), dim3(256), 0, st, a.bias, N, w.bnorm);  // ||b|| of the pre-update bias
    SC_CUDA(cudaGetLastError(), "bias_norm launch");
    prof_mark(st, "prep_params");
    // 3. G1: c = relu(x W_e^T + b)                                  [B x N], K = d
    {
        Seg s{kmajor(w.x_r, B, d, xg, Bd), kmajor(a.tied ? w.D_r : w.enc_r, N, d, M, Nd)};
        EpiEncode::Params ep{};
        SC_TRY(io_map(&ep.c_map, a.out_c, B, N, M, BN_));
        ep.bias = a.bias; ep.l1_part = w.l1_part; ep.signed_zero = 1; ep.dict_size = a.dict_size;
        SC_TRY((run_gemm<false, false, EpiEncode>(B, N, M, &s, 1, ep, st)));
    }
    prof_mark(st, "G1_encode");
    // 4. G2: r = c D - x ; R = 2 r/(B d)                             [B x d], K = N
    {
        Seg s{kmajor(a.out_c, B, N, M, BN_), mnmajor(w.D_r, d, N, M, Nd)};
        EpiDecode::Params ep{a.x, w.R_r, nullptr, w.mse_part, a.x_per_member ? Bd : 0, Bd,
                             2.0f / ((float)Bn * (float)d)};
        SC_TRY((run_gemm<false, true, EpiDecode>(B, d, M, &s, 1, ep, st)));
    }
    prof_mark(st, "G2_decode_mse");
    // 5. G3 (transposed): dz^T = f(D R^T, c)                         [N x B], K = d
    {
        Seg s{kmajor(w.D_r, N, d, M, Nd), kmajor(w.R_r, B, d, M, Bd)};
        EpiBackCode::Params ep{a.out_c, w.dz, a.l1_alpha, a.bias, w.db_part, w.s_part, BN_, 1.0f / (float)Bn, a.tied, N};
        SC_TRY((run_gemm<false, false, EpiBackCode>(N, B, M, &s, 1, ep, st)));
    }
    prof_mark(st, "G3_back_code");
    // 6/7. weight gradients with fused Adam                          [N x d], K = B
    {
        EpiAdam::Params e{};
        e.tiles_b = w.tiles_b; e.a = ad; e.g_scale = 1.0f;
        auto fill = [&](EpiAdam::Sub& u, float* p, float* mu, float* nu, float* grad, bool projected) -> int {
            SC_TRY(io_map(&u.p_map, p, N, d, M, Nd));
            SC_TRY(io_map(&u.mu_map, mu, N, d, M, Nd));
            SC_TRY(io_map(&u.nu_map, nu, N, d, M, Nd));
            SC_TRY(io_map(&u.grad_map, grad ? grad : p, N, d, M, Nd));
            u.s_part = projected ? w.s_part : nullptr;
            u.inv_norm = projected ? w.inv_norm : nullptr;
            u.projected = projected ? 1 : 0;
            u.grad_only = (grad || a.rs_world > 0) ? 1 : 0;
            return 0;
        };
        SC_TRY(fill_rs(e, a));
        e.fin = make_finalize(a, w.l1_part, w.n_l1, w.mse_part, w.n_mse, w.db_part, w.tiles_b, w.bnorm, Bn, ad, 1.0f);
        e.fin_on = 1;
        if (!a.tied) {
            // one launch, two independent problems: dW_e = dz^T x (plain) and dW_d = c^T R (projected)
            Seg s45[2] = {{mnmajor(w.dz, N, B, M, BN_), mnmajor(w.x_r, d, B, xg, Bd)},
                          {mnmajor(a.out_c, N, B, M, BN_), mnmajor(w.R_r, d, B, M, Bd)}};
            SC_TRY(fill(e.sub[0], a.enc, a.enc_mu, a.enc_nu, a.grad_enc, false));
            SC_TRY(fill(e.sub[1], a.dec, a.dec_mu, a.dec_nu, grad_mode ? a.grad_dec : nullptr, true));
            SC_TRY((run_gemm<true, true, EpiAdam>(N, d, M, s45, 2, e, st, /*segs_are_problems=*/true)));
        } else {
            Seg s[2] = {{mnmajor(a.out_c, N, B, M, BN_), mnmajor(w.R_r, d, B, M, Bd)},
                        {mnmajor(w.dz, N, B, M, BN_), mnmajor(w.x_r, d, B, xg, Bd)}};
            SC_TRY(fill(e.sub[0], a.enc, a.enc_mu, a.enc_nu, a.grad_enc, true));
            e.sub[1] = e.sub[0];
            SC_TRY((run_gemm<true, true, EpiAdam>(N, d, M, s, 2, e, st)));
        }
    }
    prof_mark(st, "G45_dW_adam");  // includes the finalize work (loss scalars + bias Adam) on the idle warps
    return 0;
}

}  // namespace sc

extern "C" {

int sc_version(void) { return 100; }
const char* sc_last_error(void) { return sc::g_last_error; }

size_t sc_sae_step_workspace_bytes(int32_t M, int32_t B, int32_t d, int32_t N, int32_t tied, int32_t x_per_member) {
    if (M <= 0 || B <= 0 || d <= 0 || N <= 0) return 0;
    const size_t a = sc::carve_step(nullptr, M, B, d, N, tied, x_per_member).bytes;
    const size_t b = sc::step_h_workspace_bytes(M, B, d, N, tied, x_per_member);
    return a > b ? a : b;
}

/* profiling hooks (used by bench.py to time individual kernels with CUDA events on the caller's stream) */
int sc_profile_begin(void) {
    sc::g_prof.on = true;
    sc::g_prof.n = 0;
    return 0;
}

/* Synchronises, writes up to `cap` (name, milliseconds since the previous mark) pairs, returns their count. */
int sc_profile_end(const char** names, float* ms, int cap) {
    using namespace sc;
    g_prof.on = false;
    int out = 0;
    if (g_prof.n > 0) cudaEventSynchronize(g_prof.ev[g_prof.n - 1]);
    for (int i = 1; i < g_prof.n; ++i) {
        if (strcmp(g_prof.name[i], "begin") == 0) continue;
        float t = 0.f;
        cudaEventElapsedTime(&t, g_prof.ev[i - 1], g_prof.ev[i]);
        if (out < cap) { names[out] = g_prof.name[i]; ms[out] = t; ++out; }
    }
    for (int i = 0; i < g_prof.n; ++i) cudaEventDestroy(g_prof.ev[i]);
    g_prof.n = 0;
    return out;
}

int sc_dp_adam_broadcast(float* grad, float* mu, float* nu, const float* p_local, float* const* p_dst, int32_t n_peers,
                         int64_t n, const ScAdam* adam, void* stream) {
    using namespace sc;
    if (!grad || !mu || !nu || !p_local || !p_dst || !adam || n <= 0 || (n & 3) || n_peers < 1 || n_peers > 8)
        return fail(-1, "bad sc_dp_adam_broadcast argument");
    PeerDst dst{};
    dst.n = n_peers;
    for (int r = 0; r < n_peers; ++r) {
        if (!p_dst[r] || (reinterpret_cast<uintptr_t>(p_dst[r]) & 15)) return fail(-1, "p_dst[%d] is NULL or not 16-byte aligned", r);
        dst.p[r] = reinterpret_cast<float4*>(p_dst[r]);
    }
    const long long n4 = n / 4;
    const int grid = static_cast<int>(n4 / 256 + 1 < 148 * 16 ? n4 / 256 + 1 : 148 * 16);
    launch_k(dp_adam_broadcast_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), reinterpret_cast<float4*>(grad),
             reinterpret_cast<float4*>(mu), reinterpret_cast<float4*>(nu), reinterpret_cast<const float4*>(p_local), dst, n4,
             make_adam(*adam));
    SC_CUDA(cudaGetLastError(), "dp_adam_broadcast launch");
    return 0;
}

int sc_adam_apply(float* p, float* mu, float* nu, const float* g, int64_t n, const ScAdam* adam, void* stream) {
    using namespace sc;
    if (!p || !mu || !nu || !g || !adam || n <= 0 || (n & 3)) return fail(-1, "bad sc_adam_apply argument");
    const long long n4 = n / 4;
    const int grid = static_cast<int>(n4 / 256 + 1 < 148 * 16 ? n4 / 256 + 1 : 148 * 16);
    launch_k(adam_apply_kernel, dim3(grid), dim3(256), 0, static_cast<cudaStream_t>(stream), 
        reinterpret_cast<float4*>(p), reinterpret_cast<float4*>(mu), reinterpret_cast<float4*>(nu),
        reinterpret_cast<const float4*>(g), n4, make_adam(*adam));
    SC_CUDA(cudaGetLastError(), "adam_apply launch");
    return 0;
}

int sc_sae_step(const ScStepArgs* args) {
    if (!args) return sc::fail(-1, "args is NULL");
    return sc::sae_step_impl(*args);
}

}  // extern "C"

extern "C" void sc_debug_set_sm_limit(int n) { sc::g_dbg_sm_limit = n; }
extern "C" void sc_debug_set_phase(int mask) { sc::g_dbg_phase = mask; }
