// Fused epilogues of the ensemble-SAE train step (reference: autoencoders/sae_ensemble.py:53-77 and
// :134-160 for the forward, functorch's backward of it, torchopt adam for the update).
// Notation: x [B,d] batch, W_e/W_d [N,d], b [N], c = relu(x W_e^T + b) [B,N], D = rownorm(W_d),
// r = c D - x, R = 2 r/(B d), q = R D^T, dz = (q + lam/B [c>0]) [z>=0].
//
// Every tensor an epilogue writes for a later GEMM is rounded to TF32 (round to nearest) here, so
// the tensor core's truncation of the low 13 mantissa bits never sees a non-zero bit.
#pragma once
#include "epi_io.cuh"
#include "epi_store.cuh"
#include <cuda_fp16.h>

namespace sc {

SC_DEVICE float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// "Fold" mode of the untied fp16 step: the decoder operand copy is half(W_d), NOT row-normalised (the Adam epilogue
// of the previous step wrote it together with per-column-tile partial sums of w^2, so no kernel re-reads the fp32
// weights), and the normalisation D = W_d / max(||W_d||, 1e-8) lives in per-feature scalars of the epilogues:
//   codes stored as   c' = c  * inv        (x_hat = c D = c' W_d;   s = sum_b c q = sum_b c' (r W_d^T))
//   dz    stored as   dz'' = dz * nrm      (dz = inv (r W_d^T) + l1 ... ;  dW_e = inv dz''^T x,  db = inv sum_b dz'')
//   dD = nrm c'^T r, and the projected gradient (dD - s D) inv = c'^T r - s inv^2 W_d.
// Rows under the clamp (||W_d|| < 1e-8) use scale 1 everywhere; their gradient is dD / 1e-8 as in the reference.
struct RowNorm {
    float nrm, inv;   // max(||W_n||, 1e-8) and its reciprocal
    float ns, is;     // scales applied to stored tensors: nrm / inv, or 1 / 1 for a clamped row (and outside fold mode)
    bool clamped;
};
// inv_norm[g][row] = 1 / max(||W_d row||, 1e-8), negative when the clamp is active (written by the step's prologue);
// nullptr outside fold mode: every scale is 1
SC_DEVICE RowNorm row_norm_of(const float* inv_norm, int g, int feats, int row) {
    RowNorm r{1.f, 1.f, 1.f, 1.f, false};
    if (!inv_norm || row >= feats) return r;
    const float iv = inv_norm[(long long)g * feats + row];
    r.clamped = iv < 0.f;
    r.inv = fabsf(iv);
    r.nrm = 1.f / r.inv;
    if (!r.clamped) { r.ns = r.nrm; r.is = r.inv; }
    return r;
}

// ---------------------------------------------------------------------------------------------
// G1  z = x W_e^T + b ; c = relu(z)            rows = batch b, cols = feature n
//   c is stored rounded to TF32; an element with z < 0 is stored as -0.0f so the backward can tell
//   z < 0 from z == 0 (the reference's clamp passes gradient at z == 0) without keeping z.
//   Also emits the per-warp partial sums of sum(c) for the L1 term (sae_ensemble.py:63).
// ---------------------------------------------------------------------------------------------
struct EpiEncode {
    static constexpr int kIoBufs = 1;  // one staging tile per warp: c leaves through TMA, fully coalesced
    struct alignas(64) Params {
        CUtensorMap c_map;  // io map of c [G][B][N]
        const float* bias;  // [G][N] or nullptr
        float* l1_part;     // [G][tiles*8] or nullptr
        int signed_zero;    // 1 (training): TF32-rounded, z<0 -> -0.0f; 0 (inference API): plain fp32 relu
        const int* dict_size;  // [G] or nullptr: features >= dict_size[g] are masked out (sae_ensemble.py:356, :425)
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float sum;
    SC_DEVICE EpiEncode(const Params& p_, const RowCtx& c_, const EpiIo& io_) : p(p_), c(c_), io(io_), sum(0.f) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        const int ncols = c.N - col;  // may be <= 0 on a ragged tile: the TMA store clips
        const int lane = threadIdx.x & 31;
        float bv[32];
        if (p.bias && ncols > 0) {
            load_row32(p.bias + (long long)c.g * c.N + col, bv, ncols);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) bv[j] = 0.f;
        }
        const bool full = c.row < c.M && ncols >= 32;
        if (p.signed_zero) {
            // masked features behave like z < 0: code -0.0, no L1 term, and the backward then gives dz = 0
            const int live_cols = p.dict_size ? p.dict_size[c.g] - col : 32;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const float z = j < live_cols ? v[j] + bv[j] : -1.f;
                if (full || (c.row < c.M && j < ncols)) sum += fmaxf(z, 0.f);
                v[j] = (z >= 0.f) ? rn_tf32(z) : -0.0f;
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j] + bv[j], 0.f);
        }
        io_wait_stores_read();  // the previous chunk's store has read the staging tile
        io_write_row(io.buf, lane, v);
        io_store_tile(&p.c_map, io.buf, col, c.row0, c.g);
        io_commit();
    }
    SC_DEVICE void finish() {
        if (!p.l1_part) return;
        const float s = warp_sum(sum);
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.l1_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * kEpiHalves) + c.part] = s;
        }
    }
};

// ---------------------------------------------------------------------------------------------
// G2  x_hat = c D ; r = x_hat - x              rows = batch b, cols = activation dim j
//   emits R = rn(2 r / (B d)) for the backward GEMMs, per-warp partials of sum(r^2)
//   (sae_ensemble.py:62) and, when asked, x_hat itself (LearnedDict.predict, learned_dict.py:42-47).
// ---------------------------------------------------------------------------------------------
struct EpiDecode {
    struct Params {
        const float* x;     // [Gx][B][d] target (nullptr: no residual, plain decode)
        float* R;           // [G][B][d] or nullptr
        float* x_hat;       // [G][B][d] or nullptr
        float* mse_part;    // [G][tiles*4] or nullptr
        long long x_gstride, out_gstride;
        float r_scale;      // 2/(B d)
    };
    const Params& p;
    const RowCtx c;
    float sum;
    static constexpr int kIoBufs = 0;
    SC_DEVICE EpiDecode(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_), sum(0.f) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        if (c.row >= c.M || col >= c.N) return;
        const int ncols = c.N - col;
        const long long off = (long long)c.row * c.N + col;
        if (p.x_hat) store_row32(p.x_hat + c.g * p.out_gstride + off, v, ncols);
        if (!p.x) return;
        float xv[32];
        load_row32(p.x + c.g * p.x_gstride + off, xv, ncols);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const float r = v[j] - xv[j];
            if (j < ncols) sum += r * r;
            v[j] = rn_tf32(r * p.r_scale);
        }
        if (p.R) store_row32(p.R + c.g * p.out_gstride + off, v, ncols);
    }
    SC_DEVICE void finish() {
        if (!p.mse_part) return;
        const float s = warp_sum(sum);
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.mse_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * kEpiHalves) + c.part] = s;
        }
    }
};

// ---------------------------------------------------------------------------------------------
// G3  q^T = D R^T, computed TRANSPOSED         rows = feature n, cols = batch b
//   dz = (q + lam/B [c>0]) [z>=0]  written to dz[b][n] (coalesced: the 32 lanes are 32 features);
//   each thread owns a feature, so the two batch reductions the backward needs are plain
//   per-thread sums, stored per column tile (deterministic, no atomics):
//     db_part[g][tile_b*2+half][n] = sum_b dz                       (bias gradient)
//     s_part [g][tile_b][n] = sum_b c q   (+ dz (c - b_n) when tied: encode also goes through D)
//   s is the radial component <dD_n, D_n> that the row normalisation projects out.
// ---------------------------------------------------------------------------------------------
struct EpiBackCode {
    struct Params {
        const float* c;       // [G][B][N]
        float* dz;            // [G][B][N]
        const float* l1;      // [G] l1_alpha
        const float* bias;    // [G][N] (tied only)
        float* db_part;       // [G][tiles_b][N]
        float* s_part;        // [G][tiles_b][N]
        long long cb_gstride; // B*N
        float inv_B;
        int tied;
        int feats;            // N (= ctx.M of this transposed GEMM)
    };
    const Params& p;
    const RowCtx c;
    float db, s, l1_over_B, bn;
    float nxt[32];  // c values of the chunk after the current one, loads in flight
    static constexpr int kIoBufs = 0;
    SC_DEVICE EpiBackCode(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_), db(0.f), s(0.f) {
        l1_over_B = p.l1[c.g] * p.inv_B;
        bn = (p.tied && c.row < c.M) ? p.bias[(long long)c.g * p.feats + c.row] : 0.f;
    }
    // 32 coalesced loads (lanes = 32 consecutive features of one batch row each) left in flight
    SC_DEVICE void prefetch(int co) {
        if (c.row >= c.M) return;
        const int col = c.col0 + co;
        const float* cp = p.c + c.g * p.cb_gstride + (long long)col * p.feats + c.row;
        const int nb = c.N - col;
        if (nb >= 32) {
#pragma unroll
            for (int j = 0; j < 32; ++j) nxt[j] = __ldg(cp + (long long)j * p.feats);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) nxt[j] = (j < nb) ? __ldg(cp + (long long)j * p.feats) : 0.f;
        }
    }
    // branch-free body, ~10 instructions per element; FULL = all 32 batch columns of the chunk exist
    template <bool FULL, bool TIED>
    SC_DEVICE void body(const float (&cvs)[32], const float (&v)[32], float* dp, int nb) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (FULL || j < nb) {
                const float cv = cvs[j];
                const float q = v[j];
                const float t = q + (cv > 0.f ? l1_over_B : 0.f);
                const float dzv = (__float_as_int(cv) < 0) ? 0.f : t;  // sign bit set <=> z < 0
                dp[(long long)j * p.feats] = rn_tf32(dzv);
                db += dzv;
                s = fmaf(cv, q, s);  // cv is +-0 when z <= 0
                if (TIED) s = fmaf(dzv, cv - bn, s);
            }
        }
    }
    SC_DEVICE void chunk(int co, float (&v)[32], int next_co) {
        if (c.row >= c.M) return;
        const int col = c.col0 + co;  // batch index of v[0]
        float* dp = p.dz + c.g * p.cb_gstride + (long long)col * p.feats + c.row;
        const int nb = c.N - col;  // c.N = B here
        float cvs[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) cvs[j] = nxt[j];
        if (next_co >= 0) prefetch(next_co);
        if (nb >= 32) {
            if (p.tied) body<true, true>(cvs, v, dp, nb);
            else body<true, false>(cvs, v, dp, nb);
        } else {
            if (p.tied) body<false, true>(cvs, v, dp, nb);
            else body<false, false>(cvs, v, dp, nb);
        }
    }
    SC_DEVICE void finish() {
        if (c.row >= c.M) return;
        const long long o = ((long long)c.g * c.tiles_n * kEpiHalves + c.tile_n * kEpiHalves + c.half) * p.feats + c.row;
        p.db_part[o] = db;
        p.s_part[o] = s;
    }
};

// ---------------------------------------------------------------------------------------------
// G4 / G5  weight gradient with Adam fused      rows = feature n, cols = activation dim j
//   plain    : g = acc                                   (untied encoder: dW_e = dz^T x)
//   projected: g = (acc - s_n D_nj) / max(||W_n||,1e-8)  (decoder / tied dictionary: gradient
//              through D = W/||W||; rows below the clamp have D = W/1e-8 and no projection)
//   then torchopt-0.7.1 Adam exactly as the reference executes it (ensemble.py:182-191):
//   mu,nu updated in place, bias correction with the caller-supplied denominators (t = 1 in
//   reference-faithful mode), p += -lr * mu_hat / (sqrt(nu_hat + eps_root) + eps).
//   The gradient never touches HBM.
// ---------------------------------------------------------------------------------------------
struct AdamConst {
    float lr, b1, b2, omb1, omb2, eps, eps_root, rbc1, rbc2;  // omb = 1 - beta, rbc = 1 / (1 - beta^t)
    float nlr_rbc1;  // -lr * rbc1
};

SC_DEVICE float sqrt_approx(float x) {
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
SC_DEVICE float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// 11 instructions per element (MUFU sqrt / rcp, 1-2 ulp): the update is lr * O(1), so the
// approximation is far below fp32 resolution of the parameter and keeps the epilogue small enough
// for the instruction cache.  The two bias corrections ride on an FFMA each: nu_hat + eps_root = fma(nu, rbc2, eps_root)
// and p -= lr * mu_hat / den = fma(mu / den, -lr * rbc1, p).
SC_DEVICE void adam_update(float& p, float& mu, float& nu, float g, const AdamConst& a) {
    mu = fmaf(mu, a.b1, g * a.omb1);
    nu = fmaf(nu, a.b2, (g * g) * a.omb2);
    const float den = sqrt_approx(fmaf(nu, a.rbc2, a.eps_root)) + a.eps;
    p = fmaf(mu * rcp_approx(den), a.nlr_rbc1, p);
}

// ---------------------------------------------------------------------------------------------
// Finalize: loss scalars (sae_ensemble.py:62-71) from the per-warp partials of G1 / G2, bias gradient
// db + beta b/||b|| (zero sub-gradient at b == 0) and its Adam step.  Small, latency-bound work that depends
// on G1-G3 only: it runs on the otherwise idle fourth warp of every CTA of the weight-gradient kernel
// (EpiAdam::side_job), concurrently with that kernel's first tile, instead of in a launch of its own.
// ---------------------------------------------------------------------------------------------
struct RsFinal {  // destinations of the fused reduce-scatter mode (ScStepArgs.rs_*): per owner, bias gradient and losses
    int world, rows;
    float* bias[8];
    float* losses[8];
};
struct FinalizeArgs {
    const float* l1_part; int n_l1;
    const float* mse_part; int n_mse;
    const float* db_part; int tiles_b;
    float *bias, *bias_mu, *bias_nu;
    const float *l1_alpha, *bias_decay;
    float* out_losses;
    int M, B, d, N;       // B: the normalising (global) batch size
    AdamConst a;
    float* grad_bias;     // gradient-only mode: write the bias gradient here instead of updating
    const float* bnorm;   // [M] ||b||_2 of the pre-update bias (prologue)
    float db_scale;
    RsFinal rs;
};
// `warp` of `nwarps` cooperating warps (any number >= 1); deterministic for a given nwarps
SC_DEVICE void finalize_work(const FinalizeArgs& f, int warp, int nwarps) {
    const int lane = threadIdx.x & 31;
    for (int m = warp; m < f.M; m += nwarps) {
        float acc = 0.f;
        for (int i = lane; i < f.n_l1; i += 32) acc += f.l1_part[(long long)m * f.n_l1 + i];
        const float l1_sum = warp_sum(acc);
        acc = 0.f;
        for (int i = lane; i < f.n_mse; i += 32) acc += f.mse_part[(long long)m * f.n_mse + i];
        const float sq_sum = warp_sum(acc);
        if (lane == 0) {
            const float l_rec = sq_sum / ((float)f.B * (float)f.d);
            const float l_l1 = f.l1_alpha[m] * (l1_sum / (float)f.B);
            const float l_bd = f.bias_decay[m] * f.bnorm[m];
            f.out_losses[0 * f.M + m] = l_rec + l_l1 + l_bd;
            f.out_losses[1 * f.M + m] = l_rec;
            f.out_losses[2 * f.M + m] = l_l1;
            f.out_losses[3 * f.M + m] = l_bd;
            // fused data-parallel mode (M == 1): every rank adds its share of the loss terms to every rank's buffer
            for (int r = 0; r < f.rs.world; ++r) {
                atomicAdd(f.rs.losses[r] + 0, l_rec);
                atomicAdd(f.rs.losses[r] + 1, l_l1);
                atomicAdd(f.rs.losses[r] + 2, l_bd);
            }
        }
    }
    const long long total = (long long)f.M * f.N;
    for (long long o = (long long)warp * 32 + lane; o < total; o += (long long)nwarps * 32) {
        const int m = static_cast<int>(o / f.N), n = static_cast<int>(o - (long long)m * f.N);
        const float bnorm = f.bnorm[m];
        const float bscale = bnorm > 0.f ? f.bias_decay[m] / bnorm : 0.f;
        float g = 0.f;
        for (int t = 0; t < f.tiles_b; ++t) g += f.db_part[((long long)m * f.tiles_b + t) * f.N + n];
        g *= f.db_scale;
        float p = f.bias[o];
        g += bscale * p;
        if (f.rs.world > 0) {
            const int owner = n / f.rs.rows;  // M == 1: o == n
            atomicAdd(f.rs.bias[owner] + (n - owner * f.rs.rows), g);
        } else if (f.grad_bias) {
            f.grad_bias[o] = g;
        } else {
            float mu = f.bias_mu[o], nu = f.bias_nu[o];
            adam_update(p, mu, nu, g, f.a);
            f.bias[o] = p; f.bias_mu[o] = mu; f.bias_nu[o] = nu;
        }
    }
}

// Parameters of the weight-gradient + Adam epilogue (EpiAdamT below)
struct alignas(64) AdamSub {
    CUtensorMap p_map, mu_map, nu_map;  // io maps of [G][N][d]
    CUtensorMap grad_map;               // io map of the gradient output (gradient-only mode)
    const float* s_part;                // [G][parts][N] (projected only)
    const float* inv_norm;              // [G][N]  1/max(||W_n||,1e-8), negative when the clamp is active
    int projected;
    int grad_only;
    // fold mode (untied fp16 step, see RowNorm): the accumulators carry the decoder normalisation as per-row scales,
    // and the epilogue leaves behind what the next step's GEMMs need, so that no kernel re-reads the fp32 weights
    const float* fold_inv;              // [G][N] 1/max(||W_d row||,1e-8) of the DECODER before this step (sign: clamp)
    float* wsq_out;                     // [G][wsq_parts][N] partial sums of w^2 of the updated decoder (decoder problem only)
    CUtensorMap half_map;               // fp16 io map (box 64 x 32) of the [G][N][d] copy of the updated parameter
    int half_on;
    CUtensorMap half_t_map;             // fp16 io map (box 32 x 64) of the TRANSPOSED copy [G][d][N] (decoder: the decode
    int half_t_on;                      // GEMM reads the dictionary K-major from it)
};
struct alignas(64) AdamParams {
AdamSub sub[2];   // one per batched problem (untied: [0] encoder, [1] decoder)
    int tiles_b;  // partial slots per feature in s_part
    float g_scale;  // applied to the accumulator and to s (fp16 mode keeps them scaled by B d / 2; else 1)
    AdamConst a;
    // fused reduce-scatter (ScStepArgs.rs_world > 0): gradient tiles are ADDED into their owner's buffer
    int rs_world;                  // 0: off
    int rs_rows;                   // dictionary rows per owner (multiple of 32)
    CUtensorMap rs_map[2][8];      // [problem][owner]: io map of the owner's [1][rs_rows][d] accumulation buffer
    int fin_on;                    // 1: the idle warps of this launch also do the step's finalize work
    int fold;                      // 1: fold mode
    int acc_split;                 // 1: split-K launch (GemmShape::ksplit > 1): every partial tile is reduce-ADDED into the
                                   // (zeroed) gradient buffer behind grad_map; Adam follows in a pass of its own
    int wsq_parts;                 // partial slots per feature in wsq_out: column tiles x SPLIT
    FinalizeArgs fin;
};

// SPLIT = 1: four epilogue warps, each with its own staging tiles.  SPLIT = 2 (gemm_core.cuh "column split"): eight
// warps, the two of a row quarter share the quarter's staging tiles and each handles 16 of a chunk's 32 columns - the
// Adam arithmetic (11 ops + 6 shared-memory vector accesses per element, two MUFU among them) is more than one warp
// per scheduler issues in the time the tile's HBM traffic takes, two warps per scheduler overlap each other's
// latencies.  Warp half 0 of the quarter owns the TMA traffic; one 64-thread named barrier per chunk orders the
// other warp's row writes before the tile stores.
// FIO = 1 adds the two fp16 staging tiles of fold mode (8 KB per quarter: one operand-pipeline stage less); launches
// that never hand operands over (TF32 mode, tied, gradient-only, split-K) use FIO = 0.
template <int SPLIT, int FIO = 0>
struct EpiAdamT {
    // HBM-bound epilogue: what matters is how many tile loads and stores are in flight.  Per quarter TWO sets of
    // p / mu / nu staging tiles: while one set is computed on and stored, the other is already loading the next chunk -
    // across tile boundaries too (static tile schedule).
    static constexpr int kHalves = SPLIT;
    static constexpr int kColSplit = SPLIT;
    static constexpr int W = 32 / SPLIT;  // columns per thread and chunk
    static constexpr int kIoBufs = 6 + 2 * FIO;  // 2 sets x (p, mu, nu) x 4 KB (+ fp16 tiles [32][64] and, transposed, [64][32])
    using Sub = AdamSub;
    using Params = AdamParams;
    // warp 3 of CTA `cta` of `nctas` (no role in the GEMM)
    static SC_DEVICE void side_job(const Params& p, int cta, int nctas) {
        if (p.fin_on) finalize_work(p.fin, cta, nctas);
    }
    const Params& p;
    const Sub& u;
    const RowCtx c;
    const EpiIo io;
    float s, inv, gsc, wsq;
    bool clamped, half_store, owner;  // owner: this warp issues the quarter's TMA loads and stores
    int half_store_col;
    SC_DEVICE EpiAdamT(const Params& p_, const RowCtx& c_, const EpiIo& io_)
        : p(p_), u(p_.sub[c_.prob]), c(c_), io(io_), s(0.f), inv(1.f), gsc(p_.g_scale), wsq(0.f), clamped(false),
          half_store(false), owner(SPLIT == 1 || c_.half == 0), half_store_col(0) {
        if (u.projected && c.row < c.M) {
            for (int t = 0; t < p.tiles_b; ++t) s += u.s_part[((long long)c.g * p.tiles_b + t) * c.M + c.row];
            s *= p.g_scale;
        }
        if (p.fold) {
            const RowNorm rn = row_norm_of(u.fold_inv, c.g, c.M, c.row);
            clamped = rn.clamped;
            inv = rn.inv;
            if (u.projected) s *= inv * inv;   // g = acc - (s inv^2) W   (clamped rows: g = acc inv, s unused)
            else gsc *= rn.is;                 // dW_e = inv dz''^T x
        } else if (u.projected && c.row < c.M) {
            const float iv = u.inv_norm[(long long)c.g * c.M + c.row];
            clamped = iv < 0.f;
            inv = fabsf(iv);
        }
    }
    // lane 0 of the owner warp: tile loads of one chunk (problem `sub`, group g, rows row0.., columns col..) into set `set`
    SC_DEVICE void request(int set, const Sub& sub, int g, int row0, int col) const {
        uint8_t* b = io.buf + set * (3 * kIoTileBytes);
        if (sub.grad_only) {
            io_expect(io, kIoTileBytes, set);
            tma_load_3d(&sub.p_map, io.bar + set, b, col, row0, g);
        } else {
            io_expect(io, 3 * kIoTileBytes, set);
            tma_load_3d(&sub.p_map, io.bar + set, b, col, row0, g);
            tma_load_3d(&sub.mu_map, io.bar + set, b + kIoTileBytes, col, row0, g);
            tma_load_3d(&sub.nu_map, io.bar + set, b + 2 * kIoTileBytes, col, row0, g);
        }
    }
    // Called by the core before it waits for the accumulator.  Only the warp's very first tile needs it: every
    // later first chunk was requested by the last chunk of the tile before.
    SC_DEVICE void prefetch(int co) {
        if (c.first && owner && (threadIdx.x & 31) == 0) request((co >> 5) & 1, u, c.g, c.row0, c.col0 + co);
    }
    // every thread of the quarter has written its part of the staging rows: make them visible to the async proxy
    // and let the owner's lane 0 go on to the tile stores
    SC_DEVICE void rows_written() const {
        fence_proxy_async_smem();
        if (SPLIT == 1) __syncwarp();
        else bar_sync_named(1 + c.warp_q, 64);
    }
    SC_DEVICE void chunk(int co, float (&v)[W], int next_co) {
        const int lane = threadIdx.x & 31;
        const int h = SPLIT == 1 ? 0 : c.half;
        const int set = (co >> 5) & 1;  // consecutive chunks alternate between the two sets
        const int col = c.col0 + co, row0 = c.row0;
        uint8_t* b = io.buf + set * (3 * kIoTileBytes);
        io_wait_loads(io, set);
        float pv[W], mv[W], nv[W];
        io_read_cols<W>(b, lane, h, pv);
        if (!u.grad_only) {
            io_read_cols<W>(b + kIoTileBytes, lane, h, mv);
            io_read_cols<W>(b + 2 * kIoTileBytes, lane, h, nv);
        }
        // the other set was stored one chunk ago (both warps of the quarter had written their rows before that
        // store was issued); once the store has read it, the next chunk loads into it
        if (owner) {
            io_wait_stores_read();
            if (lane == 0) {
                if (next_co >= 0) request(set ^ 1, u, c.g, row0, c.col0 + next_co);
                else if (c.next_g >= 0) request(set ^ 1, p.sub[c.next_prob], c.next_g, c.next_row0, c.next_col0);
            }
        }
        // the owner has seen the previous chunk's stores read their tiles: only now may the other warp of the quarter
        // overwrite the fp16 tile those stores included
        if (SPLIT == 2 && FIO && p.fold) bar_sync_named(5 + c.warp_q, 64);
        if (p.g_scale != 1.f || p.fold) {
#pragma unroll
            for (int j = 0; j < W; ++j) v[j] *= gsc;
        }
        if (u.projected) {
            if (p.fold) {
#pragma unroll
                for (int j = 0; j < W; ++j) v[j] = clamped ? v[j] * inv : fmaf(-s, pv[j], v[j]);
            } else if (c.split == 0) {
#pragma unroll
                for (int j = 0; j < W; ++j) v[j] = clamped ? v[j] * inv : (v[j] - s * (pv[j] * inv)) * inv;
            } else {  // partial tile of a split K loop: the projection is linear, its s D term goes in once (split 0)
#pragma unroll
                for (int j = 0; j < W; ++j) v[j] *= inv;
            }
        }
        if (u.grad_only) {
            io_write_cols<W>(b, lane, h, v);
            rows_written();
            if (owner && lane == 0) {
                if (p.rs_world > 0) {
                    // the quarter's 32 rows belong to one owner rank (rs_rows is a multiple of 32); rows past M are
                    // clipped by the map of the last owner
                    const int rk = min(row0 / p.rs_rows, p.rs_world - 1);
                    tma_reduce_add_3d(&p.rs_map[c.prob][rk], b, col, row0 - rk * p.rs_rows, 0);
                } else if (p.acc_split) {
                    tma_reduce_add_3d(&u.grad_map, b, col, row0, c.g);
                } else {
                    tma_store_3d(&u.grad_map, b, col, row0, c.g);
                }
            }
        } else {
#pragma unroll
            for (int j = 0; j < W; ++j) adam_update(pv[j], mv[j], nv[j], v[j], p.a);
            if (FIO && p.fold) leave_operands(pv, co, h, next_co < 0);
            io_write_cols<W>(b, lane, h, pv);
            io_write_cols<W>(b + kIoTileBytes, lane, h, mv);
            io_write_cols<W>(b + 2 * kIoTileBytes, lane, h, nv);
            rows_written();
            if (owner && lane == 0) {
                tma_store_3d(&u.p_map, b, col, row0, c.g);
                tma_store_3d(&u.mu_map, b + kIoTileBytes, col, row0, c.g);
                tma_store_3d(&u.nu_map, b + 2 * kIoTileBytes, col, row0, c.g);
                if (half_store) {
                    tma_store_3d(&u.half_map, io.buf + 6 * kIoTileBytes, half_store_col, row0, c.g);
                    if (u.half_t_on) tma_store_3d(&u.half_t_map, io.buf + 7 * kIoTileBytes, row0, half_store_col, c.g);
                }
            }
            half_store = false;
        }
        if (owner) io_commit();
    }
    // fold mode: fp16 copy of the updated parameters - two 32-column chunks share one [32][64] staging tile (128-byte
    // rows: one L2 request per row and chunk pair) - and, for the decoder, this thread's share of the row's sum of squares
    SC_DEVICE void leave_operands(const float (&pv)[W], int co, int h, bool last) {
        const int lane = threadIdx.x & 31;
        if (u.wsq_out) {
            const int ncols = c.N - (c.col0 + co) - h * W;
#pragma unroll
            for (int j = 0; j < W; ++j) wsq = (j < ncols) ? fmaf(pv[j], pv[j], wsq) : wsq;
        }
        if (!u.half_on) return;
        uint8_t* tile = io.buf + 6 * kIoTileBytes;
        const int hsel = (co >> 5) & 1;
        const uint32_t row = smem_u32(tile) + lane * 128;
#pragma unroll
        for (int j = 0; j < W / 8; ++j) {
            uint32_t hh[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const __half2 x = half2_sat(pv[8 * j + 2 * t], pv[8 * j + 2 * t + 1]);
                hh[t] = *reinterpret_cast<const uint32_t*>(&x);
            }
            sts_128u(row + (((4 * hsel + h * (W / 8) + j) ^ (lane & 7)) << 4), hh[0], hh[1], hh[2], hh[3]);
        }
        if (u.half_t_on) {
            // transposed copy: element (row = lane, column j) goes to row (32 hsel + j) of a [64][32] fp16 tile, 64-byte
            // rows with the 64-byte swizzle; the 32 lanes write one whole row per instruction
            const uint32_t tt = smem_u32(io.buf + 7 * kIoTileBytes) + ((lane & 7) << 1);
#pragma unroll
            for (int j = 0; j < W; ++j) {
                const int r = 32 * hsel + h * W + j;
                sts_u16(tt + r * 64 + ((((lane >> 3) ^ (r >> 1)) & 3) << 4), __half_as_ushort(half_sat(pv[j])));
            }
        }
        if (hsel == 1 || last) {  // the caller's fence + barrier precede the store it issues for us
            half_store_col = c.col0 + (co & ~63);
            half_store = true;
        }
    }
    SC_DEVICE void finish() {
        // two slots per column tile (wsq_parts = 2 x column tiles): one per warp of the quarter, or value and zero
        if (p.fold && u.wsq_out && c.row < c.M) {
            float* o = u.wsq_out + ((long long)c.g * p.wsq_parts + c.tile_n * 2) * c.M + c.row;
            if (SPLIT == 1) { o[0] = wsq; o[c.M] = 0.f; }
            else o[(long long)c.half * c.M] = wsq;
        }
    }
};
using EpiAdam = EpiAdamT<1, 0>;
using EpiAdamF = EpiAdamT<1, 1>;  // fold mode
using EpiAdamS = EpiAdamT<2, 1>;  // fold mode, column-split experiment

}  // namespace sc
