// fp16-operand mode of sc_sae_step (ScStepArgs.operand_format == 1): five launches (prologue, G1, G2, G3,
// G4+G5 with the finalize work on its idle warps), GEMM operands
// and the intermediate tensors c / r / dz in fp16 (see sae_epilogues_h.cuh for the scaling), master
// weights, Adam state, accumulators and all loss arithmetic in fp32.  Needs the CTA-pair kernel, d and N
// multiples of 64 and >= 128, B > 64 (sc_sae_step_f16_supported); other shapes use the TF32 mode.
// The codes and dz are kept TRANSPOSED, c^T / dz^T = [M][N][ldb] with ldb = B rounded up to 8 (TMA strides are
// multiples of 16 bytes): G1 and G3 are computed with the feature as the accumulator row, so their epilogues
// touch whole rows, G2 reads c^T MN-major and the weight-gradient GEMMs read c^T / dz^T K-major.
#include <mutex>

#include "../../include/sc_b200.h"
#include "host_common.cuh"
#include "sae_epilogues_h.cuh"

namespace sc {

__global__ void to_half_kernel(const float4* __restrict__ in, uint2* __restrict__ out, long long n4) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 v = in[i];
        __half2 h[2] = {half2_sat(v.x, v.y), half2_sat(v.z, v.w)};
        out[i] = *reinterpret_cast<uint2*>(h);
    }
}

// aux output: c [M][B][N] fp32 from the engine's c^T [M][N][ldb] fp16 (64 x 64 tiles through shared memory)
// row_scale [M][N] (or nullptr): what the stored codes of feature n have to be multiplied by (fold mode keeps c / ||W_d||)
__global__ void __launch_bounds__(256) codes_to_float_kernel(const __half* __restrict__ cT, float* __restrict__ out, int B,
                                                             int N, int ldb, const float* __restrict__ row_scale) {
    pdl_launch_dependents();
    pdl_wait();
    // 64 x 64 tile: 128-byte rows in (half2 per lane), two 128-byte row halves out per warp and batch row
    __shared__ float t[64][65];
    const int g = blockIdx.z, n0 = blockIdx.y * 64, b0 = blockIdx.x * 64;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = warp; i < 64; i += 8) {
        const int n = n0 + i, b = b0 + 2 * lane;
        float v0 = 0.f, v1 = 0.f;
        if (n < N && b < B) {  // ldb is B rounded up to 8, so the pair at an even b < B lies inside the row
            const float sc = row_scale ? row_scale[(long long)g * N + n] : 1.f;
            const float2 v = __half22float2(*reinterpret_cast<const __half2*>(cT + ((long long)g * N + n) * ldb + b));
            v0 = v.x * sc;
            v1 = v.y * sc;
        }
        t[i][2 * lane] = v0;
        t[i][2 * lane + 1] = v1;
    }
    __syncthreads();
    for (int i = warp; i < 64; i += 8) {
        const int b = b0 + i;
        if (b >= B) continue;
        float* row = out + ((long long)g * B + b) * N + n0;
        if (n0 + lane < N) row[lane] = t[lane][i];
        if (n0 + lane + 32 < N) row[lane + 32] = t[lane + 32][i];
    }
}

// one warp per dictionary row: enc_h = half(enc), D_h = half(W / max(||W||,1e-8)), inv_norm as in prep_params_kernel
__global__ void prep_params_h_kernel(const float* __restrict__ enc, __half* __restrict__ enc_h,
                                     const float* __restrict__ W, __half* __restrict__ D_h,
                                     float* __restrict__ inv_norm, int N, int d) {
    pdl_launch_dependents();
    pdl_wait();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * 8 + warp;
    if (n >= N) return;
    const long long row = ((long long)blockIdx.y * N + n) * d;
    const int d4 = d >> 2;
    if (enc_h) {
        const float4* src = reinterpret_cast<const float4*>(enc + row);
        uint2* dst = reinterpret_cast<uint2*>(enc_h + row);
        for (int i = lane; i < d4; i += 32) {
            const float4 v = src[i];
            __half2 h[2] = {half2_sat(v.x, v.y), half2_sat(v.z, v.w)};
            dst[i] = *reinterpret_cast<uint2*>(h);
        }
    }
    const float4* src = reinterpret_cast<const float4*>(W + row);
    float ss = 0.f;
    for (int i = lane; i < d4; i += 32) {
        const float4 v = src[i];
        ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    const float nrm = sqrtf(ss);
    const float inv = 1.f / fmaxf(nrm, 1e-8f);
    if (lane == 0 && inv_norm) inv_norm[(long long)blockIdx.y * N + n] = nrm < 1e-8f ? -inv : inv;
    uint2* dst = reinterpret_cast<uint2*>(D_h + row);
    for (int i = lane; i < d4; i += 32) {
        const float4 v = src[i];
        __half2 h[2] = {half2_sat(v.x * inv, v.y * inv), half2_sat(v.z * inv, v.w * inv)};
        dst[i] = *reinterpret_cast<uint2*>(h);
    }
}

// One launch for everything a step needs before its first GEMM (three launches and two kernel boundaries less):
//   blocks [0, M)                  ||b||_2 of member m (pre-update bias, for the bias-decay term in finalize)
//   blocks [M, M + x_blocks)       x_h = half(x)
//   the rest, ceil(N/8) per member operand preparation, one warp per dictionary row (prep_params_h_kernel)
// The long-latency single-block jobs come first so that they are scheduled first.
__global__ void __launch_bounds__(256) step_prologue_h_kernel(const float* __restrict__ bias, float* __restrict__ bnorm, int M,
                                                              const float4* __restrict__ x, uint2* __restrict__ x_h,
                                                              long long x_n4, int x_blocks, const float* __restrict__ enc,
                                                              __half* __restrict__ enc_h, const float* __restrict__ W,
                                                              __half* __restrict__ D_h, float* __restrict__ inv_norm, int N,
                                                              int d, float4* __restrict__ zero, long long zero_n4,
                                                              int fold, const float* __restrict__ wsq, int wsq_parts,
                                                              int norm_blocks, int* __restrict__ range_flag, int step_tag,
                                                              __half* __restrict__ D_ht) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    int blk = blockIdx.x;
    if (blk < M) {
        float acc = 0.f;
        for (int n = threadIdx.x; n < N; n += blockDim.x) {
            const float b = bias[(long long)blk * N + n];
            acc = fmaf(b, b, acc);
        }
        const float t = block_sum(acc, sh);
        if (threadIdx.x == 0) bnorm[blk] = sqrtf(t);
        return;
    }
    blk -= M;
    if (blk < x_blocks) {
        bool out_of_range = false;
        for (long long i = blk * (long long)blockDim.x + threadIdx.x; i < x_n4; i += (long long)x_blocks * blockDim.x) {
            const float4 v = x[i];
            // fp16 holds |x| <= 65504: larger (or non-finite) activations are saturated by the conversion and reported
            out_of_range |= !(fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))) <= 65504.f);
            __half2 h[2] = {half2_sat(v.x, v.y), half2_sat(v.z, v.w)};
            x_h[i] = *reinterpret_cast<uint2*>(h);
        }
        if (range_flag && __any_sync(0xffffffffu, out_of_range) && (threadIdx.x & 31) == 0) atomicMax(range_flag, step_tag);
        // split-K decode accumulates into this buffer (EpiReduceTile): clear it here instead of in a memset node
        for (long long i = blk * (long long)blockDim.x + threadIdx.x; i < zero_n4; i += (long long)x_blocks * blockDim.x)
            zero[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        return;
    }
    blk -= x_blocks;
    if (norm_blocks > 0) {
        // fold mode, operands left by the previous step: only 1/||W_d row|| is missing, from the partial sums of w^2 the
        // Adam epilogue wrote per column tile
        const long long o = (long long)blk * blockDim.x + threadIdx.x;
        if (o < (long long)M * N) {
            const long long m = o / N, n = o - m * N;
            float ss = 0.f;
            for (int t = 0; t < wsq_parts; ++t) ss += wsq[(m * wsq_parts + t) * N + n];
            const float nrm = sqrtf(ss);
            const float inv = 1.f / fmaxf(nrm, 1e-8f);
            inv_norm[o] = nrm < 1e-8f ? -inv : inv;
        }
        return;
    }
    // eight dictionary rows per block, one warp each; the transposed copy leaves through shared memory so that every
    // thread stores eight consecutive features (16 bytes) of one activation dimension
    extern __shared__ __align__(16) uint8_t prologue_smem[];
    __half* tile = reinterpret_cast<__half*>(prologue_smem);  // [8 rows][d] when D_ht is wanted
    const int row_blocks = (N + 7) >> 3;
    const int m = blk / row_blocks;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = (blk - m * row_blocks) * 8;
    const int n = n0 + warp;
    const int d4 = d >> 2;
    if (n < N) {
        const long long row = ((long long)m * N + n) * d;
        if (enc_h) {
            const float4* src = reinterpret_cast<const float4*>(enc + row);
            uint2* dst = reinterpret_cast<uint2*>(enc_h + row);
            for (int i = lane; i < d4; i += 32) {
                const float4 v = src[i];
                __half2 h[2] = {half2_sat(v.x, v.y), half2_sat(v.z, v.w)};
                dst[i] = *reinterpret_cast<uint2*>(h);
            }
        }
        const float4* src = reinterpret_cast<const float4*>(W + row);
        float ss = 0.f;
        for (int i = lane; i < d4; i += 32) {
            const float4 v = src[i];
            ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
        ss = warp_sum(ss);
        const float nrm = sqrtf(ss);
        float inv = 1.f / fmaxf(nrm, 1e-8f);
        if (lane == 0) inv_norm[(long long)m * N + n] = nrm < 1e-8f ? -inv : inv;
        if (fold) inv = 1.f;  // fold mode: the copies stay un-normalised (RowNorm in sae_epilogues.cuh)
        uint2* dst = reinterpret_cast<uint2*>(D_h + row);
        for (int i = lane; i < d4; i += 32) {
            const float4 v = src[i];
            __half2 h[2] = {half2_sat(v.x * inv, v.y * inv), half2_sat(v.z * inv, v.w * inv)};
            dst[i] = *reinterpret_cast<uint2*>(h);
            if (D_ht) reinterpret_cast<uint2*>(tile + warp * d)[i] = *reinterpret_cast<uint2*>(h);
        }
    } else if (D_ht) {
        for (int i = lane; i < d; i += 32) tile[warp * d + i] = __float2half(0.f);
    }
    if (D_ht) {  // (block-uniform)
        __syncthreads();
        // D_ht[m][j][n0 .. n0 + 7]: N is a multiple of 8 in fp16 mode, so the eight features exist or none does
        if (n0 + 8 <= N)
            for (int j = threadIdx.x; j < d; j += blockDim.x) {
                __half col[8];
#pragma unroll
                for (int r = 0; r < 8; ++r) col[r] = tile[r * d + j];
                *reinterpret_cast<uint4*>(D_ht + ((long long)m * d + j) * N + n0) = *reinterpret_cast<const uint4*>(col);
            }
    }
}

// r = half(x_hat - x), mse_part[m][j] = sum of r^2 over slab j (n_mse slabs per member; block (j, m))
__global__ void __launch_bounds__(256) residual_h_kernel(const float4* __restrict__ x_hat, const float4* __restrict__ x,
                                                         uint2* __restrict__ r_h, float* __restrict__ mse_part, long long n4,
                                                         int n_mse, long long x_mstride4) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    const int m = blockIdx.y;
    const long long per = (n4 + n_mse - 1) / n_mse;
    const long long lo = blockIdx.x * per, hi = lo + per < n4 ? lo + per : n4;
    float acc = 0.f;
    for (long long i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const float4 a = x_hat[m * n4 + i], b = x[m * x_mstride4 + i];
        const float r0 = a.x - b.x, r1 = a.y - b.y, r2 = a.z - b.z, r3 = a.w - b.w;
        acc += (r0 * r0 + r1 * r1) + (r2 * r2 + r3 * r3);
        __half2 h[2] = {half2_sat(r0, r1), half2_sat(r2, r3)};
        r_h[m * n4 + i] = *reinterpret_cast<uint2*>(h);
    }
    const float t = block_sum(acc, sh);
    if (threadIdx.x == 0) mse_part[(long long)m * n_mse + blockIdx.x] = t;
}

struct StepWsH {
    __half *x_h, *enc_h, *D_h, *r_h, *dz_h;
    __half* dec_ht;  // fold mode: transposed fp16 copy [M][d][N] of the decoder (K-major operand of the decode GEMM)
    float *inv_norm, *l1_part, *mse_part, *db_part, *s_part, *bnorm, *x_hat;
    float *gacc_enc, *gacc_dec;  // split-K weight gradients of small ensembles: fp32 accumulation buffers [M][N][d]
    int g45_split;               // > 1: the weight-gradient GEMMs run split-K into them, Adam follows as its own pass
    float* zero_from;            // start and length (bytes) of the workspace span the prologue has to clear
    size_t zero_bytes;
    float* wsq;  // fold mode: [M][wsq_parts][N] partial sums of w^2 of the updated decoder rows (Adam epilogue -> next prologue)
    int wsq_parts;
    int g2_split;  // > 1: G2 runs split-K into x_hat
    int n_l1, n_mse, tiles_b;
    int ldb;  // row stride of c^T / dz^T in elements
    size_t bytes;
};

static StepWsH carve_step_h(void* base, int M, int B, int d, int N, int tied, int xpm) {
    StepWsH w;
    Carver c(base);
    const size_t Mz = M, Bz = B, dz_ = d, Nz = N;
    auto halves = [&](size_t n) { return reinterpret_cast<__half*>(c.floats((n + 1) / 2)); };
    // what a step may leave for the next one (ScStepArgs.operands_valid) comes first: its offsets depend on M, N, d only
    w.enc_h = tied ? nullptr : halves(Mz * Nz * dz_);
    w.D_h = halves(Mz * Nz * dz_);
    w.dec_ht = halves(Mz * Nz * dz_);
    w.wsq_parts = ceil_div(d, pick_bn(d)) * 2;  // one slot per column tile and epilogue warp of a row quarter
    w.wsq = c.floats(Mz * w.wsq_parts * Nz);
    w.inv_norm = c.floats(Mz * Nz);
    w.x_h = halves((xpm ? Mz : 1) * Bz * dz_);
    w.r_h = halves(Mz * Bz * dz_);
    w.ldb = (B + 7) & ~7;
    w.dz_h = halves(Mz * Nz * (size_t)w.ldb);
    w.n_l1 = row_tiles128(N) * ceil_div(B, pick_bn(B)) * 4 * EpiHalvesOf<EpiEncodeTH>::v;  // G1 runs transposed: rows = features
    w.n_mse = row_tiles128(B) * ceil_div(d, pick_bn(d)) * 4 * kEpiHalves;
    w.tiles_b = ceil_div(B, pick_bn(B)) * EpiHalvesOf<EpiBackCodeH>::v;  // partial slots per feature
    w.l1_part = c.floats(Mz * w.n_l1);
    w.mse_part = c.floats(Mz * w.n_mse);
    w.db_part = c.floats(Mz * w.tiles_b * Nz);
    w.s_part = c.floats(Mz * w.tiles_b * Nz);
    w.bnorm = c.floats(Mz);
    // G2's output has M ceil(B/256) ceil(d/bn) tiles; with fewer than half as many as CTA pairs most of the machine
    // would idle through its long K = N loop: split K so that the partial tiles fill it (>= 8 k-blocks each)
    w.g2_split = 1;
    w.x_hat = nullptr;
    {
        const long long tiles = (long long)M * ceil_div(B, 256) * ceil_div(d, pick_bn(d));
        int sms = 148;
        (void)device_sm_count(&sms);  // falls back to 148 when no device is current (workspace sizing on the host)
        const int pairs = sms / 2, kblocks = ceil_div(N, 64);
        int sp = static_cast<int>(pairs / (tiles > 0 ? tiles : 1));
        if (sp > kblocks / 8) sp = kblocks / 8;
        if (sp > 8) sp = 8;
        if (sp >= 2 && (d & 3) == 0) {
            while (sp > 1 && (long long)(sp - 1) * ceil_div(kblocks, sp) >= kblocks) --sp;  // no empty K range
            if (sp >= 2) {
                w.g2_split = sp;
                w.x_hat = c.floats(Mz * Bz * dz_);
            }
        }
    }
    // Weight gradients: M ceil(N/256) ceil(d/bn) tiles per problem with the LONGEST K loop of the step (K = B).  Small
    // ensembles leave most CTA pairs idle there (toy config: 8 tiles on 74 pairs); split K so that the partial tiles
    // fill the machine, reduce-add them into fp32 buffers and run Adam as an elementwise pass afterwards.
    w.g45_split = 1;
    w.gacc_enc = w.gacc_dec = nullptr;
    {
        int sms = 148;
        (void)device_sm_count(&sms);
        const long long tiles = (long long)M * ceil_div(N, 256) * ceil_div(d, pick_bn(d)) * (tied ? 1 : 2);
        const int kblocks = ceil_div(B, 64);
        int sp = static_cast<int>((sms / 2) / (tiles > 0 ? tiles : 1));
        if (sp > kblocks / 8) sp = kblocks / 8;
        if (sp > 8) sp = 8;
        while (sp > 1 && (long long)(sp - 1) * ceil_div(kblocks, sp) >= kblocks) --sp;  // no empty K range
        if (sp >= 2) {
            w.g45_split = sp;
            w.gacc_enc = c.floats(Mz * Nz * dz_);
            if (!tied) w.gacc_dec = c.floats(Mz * Nz * dz_);
        }
    }
    w.zero_from = w.x_hat ? w.x_hat : w.gacc_enc;
    w.zero_bytes = w.zero_from ? static_cast<size_t>(c.base + c.off - reinterpret_cast<uint8_t*>(w.zero_from)) : 0;
    w.bytes = c.off;
    return w;
}

size_t step_h_workspace_bytes(int M, int B, int d, int N, int tied, int xpm) { return carve_step_h(nullptr, M, B, d, N, tied, xpm).bytes; }

// SC_B200_ADAM_SPLIT=1: the Adam epilogue with eight warps (two per row quarter sharing every chunk) instead of four.
// Measured at the headline shape: no difference (0.472 vs 0.471 ms per step) - the kernel waits for HBM, not for the
// arithmetic of its epilogue warps - so the simpler four-warp form stays the default.
static bool adam_split() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SC_B200_ADAM_SPLIT");
        v = (e && e[0] == '1') ? 1 : 0;
    }
    return v == 1;
}

// Operand hand-over (fold mode) or the operand-preparation pass in every step?  The hand-over saves a constant (the
// pass over the fp32 weights); what it costs - a pipeline stage of the weight-gradient kernel for its fp16 staging
// tiles - grows with the batch.  Same-box step times, hand-over / preparation pass: 0.4715 / 0.483 ms at 2048 rows,
// 0.830 / 0.826 at 4096, 1.507 / 1.481 at 8192.  SC_B200_FOLD: 0 never, 1 (default) up to 3072 rows, 2 always.
static bool fold_enabled(int B) {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SC_B200_FOLD");
        v = e ? atoi(e) : 1;
    }
    return v == 2 || (v == 1 && B <= 3072);
}

// What the last fp16 step left in a workspace: the caller's operands_valid only says that nobody else touched the
// parameters - whether the copies in THIS workspace are hand-over copies (un-normalised, fold mode) or the preparation
// pass's (normalised) depends on which mode that step ran in, and the mode depends on the batch size.
struct HandoverRecord {
    const void* ws;
    int M, d, N;
    bool fold;
};
static HandoverRecord g_handover[16];
static int g_handover_next = 0;
static std::mutex g_handover_mutex;
static bool handover_intact(const void* ws, int M, int d, int N) {
    std::lock_guard<std::mutex> lock(g_handover_mutex);
    for (const HandoverRecord& r : g_handover)
        if (r.ws == ws) return r.fold && r.M == M && r.d == d && r.N == N;
    return false;
}
static void handover_record(const void* ws, int M, int d, int N, bool fold) {
    std::lock_guard<std::mutex> lock(g_handover_mutex);
    for (HandoverRecord& r : g_handover)
        if (r.ws == ws) { r = HandoverRecord{ws, M, d, N, fold}; return; }
    g_handover[g_handover_next] = HandoverRecord{ws, M, d, N, fold};
    g_handover_next = (g_handover_next + 1) % 16;
}

static bool f16_supported(int B, int d, int N) {
    return use_cta_pairs() && d >= 128 && N >= 128 && (d % 64) == 0 && (N % 64) == 0 && B > 64;
}

static int io_map_h(CUtensorMap* out, const void* base, long long rows, long long cols, long long g, long long gstride,
                    long long ld = 0) {
    const char* err = nullptr;
    const int rc = make_io_map_h(out, base, rows, cols, g, gstride, &err, ld);
    if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
    return 0;
}

int sae_step_f16_impl(const ScStepArgs& a) {
    const int M = a.M, B = a.B, d = a.d, N = a.N;
    if (!f16_supported(B, d, N))
        return fail(-1, "fp16 operand mode needs the CTA-pair kernels, d and N multiples of 64 and >= 128, B > 64 "
                        "(got B=%d d=%d N=%d); use operand_format 0", B, d, N);
    if (!a.out_c_half) return fail(-1, "out_c_half is NULL in fp16 operand mode");
    if (a.out_c && !a.out_code_scale) return fail(-1, "out_c in fp16 operand mode needs out_code_scale ([M,N] floats) as well");
    if (reinterpret_cast<uintptr_t>(a.out_c_half) & 15) return fail(-1, "out_c_half is not 16-byte aligned");
    const StepWsH w = carve_step_h(a.workspace, M, B, d, N, a.tied, a.x_per_member);
    if (a.workspace_bytes < w.bytes) return fail(-1, "workspace too small: %zu bytes given, %zu needed", a.workspace_bytes, w.bytes);
    cudaStream_t st = static_cast<cudaStream_t>(a.stream);
    const AdamConst ad = make_adam(a.adam);
    const int xg = a.x_per_member ? M : 1;
    const int Bn = a.B_total > 0 ? a.B_total : B;
    const bool grad_mode = a.grad_enc != nullptr || a.rs_world > 0;
    const long long Bd = (long long)B * d, Nd = (long long)N * d;
    const long long ldb = w.ldb, NB = (long long)N * ldb;  // c^T / dz^T: [M][N][ldb]
    // c^T and dz^T as GEMM operands: MN-major with m = batch row (G2's A), K-major with k = batch row (G4 / G5's A)
    auto codesT_mn = [&](const void* base) { return Operand{base, B, N, M, ldb, NB, true, 2}; };
    auto codesT_k = [&](const void* base) { return Operand{base, N, B, M, ldb, NB, false, 2}; };
    const float sc_ = 2.0f / ((float)Bn * (float)d);
    __half* c_h = static_cast<__half*>(a.out_c_half);
    // Fold mode (untied full step): the Adam epilogue of this step leaves the fp16 operand copies and the decoder row
    // norms of the UPDATED parameters in the workspace, and the next step (operands_valid) skips the pass over the fp32
    // weights: its prologue only turns the epilogue's partial sums of w^2 into 1/||W_d row||.
    const bool g45_split = w.g45_split > 1 && !grad_mode;
    const bool fold = !a.tied && !grad_mode && !g45_split && fold_enabled(B);
    const float* fold_inv = fold ? w.inv_norm : nullptr;
    // SC_B200_DICT_T=1: the decode GEMM reads the dictionary K-major from a transposed fp16 copy (made by the prologue
    // when it rebuilds the operand copies, kept current by the Adam epilogue in fold mode) instead of MN-major from the
    // plain copy.  That was worth 12 % of the decode mainloop while the MMA issue loop was the bottleneck (its MN-major
    // descriptors cost more to build); with the converged issue loops the two read the same (0.790 / 0.787 ms at 4096
    // rows, 1.402 / 1.403 at 8192), and in fold mode the extra 33 MB of epilogue stores cost 1.6 % of the step
    // (0.4497 / 0.4423 ms at 2048 rows), so it is off by default.
    static const bool dict_t_env = getenv("SC_B200_DICT_T") && getenv("SC_B200_DICT_T")[0] == '1';
    const bool dict_t = dict_t_env && (size_t)d * 16 <= 48 * 1024;  // the prologue's transposing tile: 16 d bytes of shared memory
    const bool params_ready = fold && a.operands_valid != 0 && handover_intact(a.workspace, M, d, N);
    handover_record(a.workspace, M, d, N, fold);

    // SC_B200_DUP=k (debug): launch kernel k (0 prologue, 1 G1, 2 G2, 3 G3; all idempotent) twice, so that the
    // difference in step time is that kernel's marginal cost inside the real launch chain
    static const int dup_k = getenv("SC_B200_DUP") ? atoi(getenv("SC_B200_DUP")) : -1;
    prof_mark(st, "begin");
    if (g_dbg_phase & 1) {
    for (int rep = 0; rep < (dup_k == 0 ? 2 : 1); ++rep) {
        const long long x_n4 = xg * Bd / 4;
        const int x_blocks = static_cast<int>(x_n4 / 1024 + 1 < 1024 ? x_n4 / 1024 + 1 : 1024);
        // operands left by the previous step: only the batch conversion, the bias norms and the zeroing remain
        const int norm_blocks = params_ready ? ceil_div((long long)M * N, 256) : 0;
        const int blocks = M + x_blocks + (params_ready ? norm_blocks : ceil_div(N, 8) * M);
        // split-K GEMMs accumulate into workspace buffers (x_hat of G2, the weight-gradient buffers): cleared here
        float4* zero = reinterpret_cast<float4*>(w.zero_from);
        const long long zero_n4 = static_cast<long long>(w.zero_bytes / 16);
        if (a.tied)
            launch_k(step_prologue_h_kernel, dim3(blocks), dim3(256), dict_t ? (size_t)d * 16 : 0, st, a.bias, w.bnorm, M,
                     reinterpret_cast<const float4*>(a.x), reinterpret_cast<uint2*>(w.x_h), x_n4, x_blocks, nullptr, nullptr,
                     a.enc, w.D_h, w.inv_norm, N, d, zero, zero_n4, 0, nullptr, 0, 0, a.range_flag, a.step_tag,
                     dict_t ? w.dec_ht : nullptr);
        else
            launch_k(step_prologue_h_kernel, dim3(blocks), dim3(256), (dict_t && !params_ready) ? (size_t)d * 16 : 0, st, a.bias,
                     w.bnorm, M, reinterpret_cast<const float4*>(a.x), reinterpret_cast<uint2*>(w.x_h), x_n4, x_blocks, a.enc,
                     w.enc_h, a.dec, w.D_h, w.inv_norm, N, d, zero, zero_n4, fold ? 1 : 0, w.wsq, w.wsq_parts, norm_blocks,
                     a.range_flag, a.step_tag, dict_t ? w.dec_ht : nullptr);
        SC_CUDA(cudaGetLastError(), "prologue launch");
    }
    prof_mark(st, "prologue");
    for (int rep = 0; rep < (dup_k == 1 ? 2 : 1); ++rep) {   // G1 (transposed: rows = features)
        Seg s{kmajor_h(a.tied ? w.D_h : w.enc_h, N, d, M, Nd), kmajor_h(w.x_h, B, d, xg, Bd)};
        EpiEncodeTH::Params ep{};
        SC_TRY(io_map_h(&ep.c_map, c_h, N, B, M, NB, ldb));
        ep.bias = a.bias; ep.l1_part = w.l1_part; ep.dict_size = a.dict_size;
        ep.fold_inv = fold_inv; ep.code_scale = a.out_code_scale;
        SC_TRY((run_gemm_short_k<EpiEncodeTH>(N, B, M, &s, ep, st, 1)));
    }
    prof_mark(st, "G1_encode");
    // fold mode: the dictionary comes K-major from its transposed copy (both operands MN-major cost 12 % of the mainloop)
    const Operand g2_dict = dict_t ? kmajor_h(w.dec_ht, d, N, M, Nd) : mnmajor_h(w.D_h, d, N, M, Nd);
    for (int rep = 0; rep < (dup_k == 2 ? 2 : 1); ++rep)
    if (w.g2_split > 1) {  // G2, split K: partial tiles reduce-added into x_hat, then the residual pass
        Seg s{codesT_mn(c_h), g2_dict};
        EpiReduceTile::Params ep{};
        {
            const char* err = nullptr;
            const int rc = make_io_map(&ep.out_map, w.x_hat, B, d, M, Bd, &err);
            if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
        }
        if (dict_t) SC_TRY((run_gemm<true, false, EpiReduceTile, true>(B, d, M, &s, 1, ep, st, false, 0, w.g2_split)));
        else SC_TRY((run_gemm<true, true, EpiReduceTile, true>(B, d, M, &s, 1, ep, st, false, 0, w.g2_split)));
        launch_k(residual_h_kernel, dim3(w.n_mse, M), dim3(256), 0, st, reinterpret_cast<const float4*>(w.x_hat),
                 reinterpret_cast<const float4*>(a.x), reinterpret_cast<uint2*>(w.r_h), w.mse_part, Bd / 4, w.n_mse,
                 a.x_per_member ? Bd / 4 : 0LL);
        SC_CUDA(cudaGetLastError(), "residual launch");
    } else {   // G2
        Seg s{codesT_mn(c_h), g2_dict};
        EpiDecodeH::Params ep{a.x, w.r_h, w.mse_part, a.x_per_member ? Bd : 0, Bd};
        if (dict_t) SC_TRY((run_gemm<true, false, EpiDecodeH, true>(B, d, M, &s, 1, ep, st)));
        else SC_TRY((run_gemm<true, true, EpiDecodeH, true>(B, d, M, &s, 1, ep, st)));
    }
    prof_mark(st, "G2_decode_mse");
    // G3 chunk width: 64 batch columns per epilogue chunk (streaming kernel, 12 KB of staging per warp) or 32 (6 KB per
    // warp, which fits beside a resident A operand).  SC_B200_G3W=32|64.
    static const int g3w_env = getenv("SC_B200_G3W") ? atoi(getenv("SC_B200_G3W")) : 64;
    const bool g3_wide = g3w_env == 64 && pick_bn(B) == 256;
    for (int rep = 0; rep < (dup_k == 3 ? 2 : 1); ++rep) {   // G3 (transposed)
        Seg s{kmajor_h(w.D_h, N, d, M, Nd), kmajor_h(w.r_h, B, d, M, Bd)};
        auto fill = [&](auto& ep) {
            ep.l1 = a.l1_alpha; ep.bias = a.bias; ep.db_part = w.db_part; ep.s_part = w.s_part;
            ep.l1_scale = 0.5f * (float)d; ep.tied = a.tied; ep.feats = N;
            ep.fold_inv = fold_inv;
        };
        if (g3_wide) {
            EpiBackCodeH64::Params ep{};
            const char* err = nullptr;
            int rc = make_io_map_h64(&ep.c_map, c_h, N, B, M, NB, &err, ldb);
            if (rc == 0) rc = make_io_map_h64(&ep.dz_map, w.dz_h, N, B, M, NB, &err, ldb);
            if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
            fill(ep);
            SC_TRY((run_gemm<false, false, EpiBackCodeH64, true>(N, B, M, &s, 1, ep, st)));
        } else {
            EpiBackCodeH::Params ep{};
            SC_TRY(io_map_h(&ep.c_map, c_h, N, B, M, NB, ldb));
            SC_TRY(io_map_h(&ep.dz_map, w.dz_h, N, B, M, NB, ldb));
            fill(ep);
            SC_TRY((run_gemm_short_k<EpiBackCodeH>(N, B, M, &s, ep, st, 2)));
        }
    }
    }  // phase 1
    prof_mark(st, "G3_back_code");
    if (!(g_dbg_phase & 2)) return 0;
    {
        AdamParams e{};
        e.tiles_b = w.tiles_b; e.a = ad; e.g_scale = sc_;
        e.fold = fold ? 1 : 0; e.wsq_parts = w.wsq_parts;
        e.acc_split = g45_split ? 1 : 0;
        const int ks = g45_split ? w.g45_split : 0;
        auto fill = [&](AdamSub& u, float* p, float* mu, float* nu, float* grad, bool projected) -> int {
            u.fold_inv = fold_inv;
            u.wsq_out = (fold && projected) ? w.wsq : nullptr;
            u.half_on = fold ? 1 : 0;
            if (fold) {
                const char* err = nullptr;
                int rc = make_io_map_h64(&u.half_map, projected ? w.D_h : w.enc_h, N, d, M, Nd, &err);
                u.half_t_on = (projected && dict_t) ? 1 : 0;
                if (rc == 0 && projected && dict_t) rc = make_io_map_ht(&u.half_t_map, w.dec_ht, d, N, M, Nd, &err);
                if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
            }
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
        e.fin = make_finalize(a, w.l1_part, w.n_l1, w.mse_part, w.n_mse, w.db_part, w.tiles_b, w.bnorm, Bn, ad, sc_);
        e.fin_on = 1;
        if (!a.tied) {
            Seg s45[2] = {{codesT_k(w.dz_h), mnmajor_h(w.x_h, d, B, xg, Bd)},
                          {codesT_k(c_h), mnmajor_h(w.r_h, d, B, M, Bd)}};
            SC_TRY(fill(e.sub[0], a.enc, a.enc_mu, a.enc_nu, g45_split ? w.gacc_enc : a.grad_enc, false));
            SC_TRY(fill(e.sub[1], a.dec, a.dec_mu, a.dec_nu, g45_split ? w.gacc_dec : (grad_mode ? a.grad_dec : nullptr), true));
            if (fold && adam_split()) SC_TRY((run_gemm<false, true, EpiAdamS, true>(N, d, M, s45, 2, e, st, true)));
            else if (fold) SC_TRY((run_gemm<false, true, EpiAdamF, true>(N, d, M, s45, 2, e, st, true)));
            else SC_TRY((run_gemm<false, true, EpiAdam, true>(N, d, M, s45, 2, e, st, true, 0, ks)));
        } else {
            Seg s[2] = {{codesT_k(c_h), mnmajor_h(w.r_h, d, B, M, Bd)},
                        {codesT_k(w.dz_h), mnmajor_h(w.x_h, d, B, xg, Bd)}};
            SC_TRY(fill(e.sub[0], a.enc, a.enc_mu, a.enc_nu, g45_split ? w.gacc_enc : a.grad_enc, true));
            e.sub[1] = e.sub[0];
            SC_TRY((run_gemm<false, true, EpiAdam, true>(N, d, M, s, 2, e, st, false, 0, ks)));  // tied: never fold mode
        }
        if (g45_split) {  // Adam from the accumulated gradients (elementwise, the tensors are small in this regime)
            const long long n4 = (long long)M * Nd / 4;
            const int grid = static_cast<int>(n4 / 256 + 1 < 1184 ? n4 / 256 + 1 : 1184);
            launch_k(adam_apply_kernel, dim3(grid), dim3(256), 0, st, reinterpret_cast<float4*>(a.enc), reinterpret_cast<float4*>(a.enc_mu),
                     reinterpret_cast<float4*>(a.enc_nu), reinterpret_cast<const float4*>(w.gacc_enc), n4, ad);
            if (!a.tied)
                launch_k(adam_apply_kernel, dim3(grid), dim3(256), 0, st, reinterpret_cast<float4*>(a.dec), reinterpret_cast<float4*>(a.dec_mu),
                         reinterpret_cast<float4*>(a.dec_nu), reinterpret_cast<const float4*>(w.gacc_dec), n4, ad);
            SC_CUDA(cudaGetLastError(), "adam apply launch");
        }
    }
    prof_mark(st, "G45_dW_adam");  // includes the finalize work (loss scalars + bias Adam) on the idle warps
    if (a.out_c) {  // caller wants the fp32 aux tensor as well
        launch_k(codes_to_float_kernel, dim3(ceil_div(B, 64), ceil_div(N, 64), M), dim3(256), 0, st, c_h, a.out_c, B, N, w.ldb,
                 static_cast<const float*>(a.out_code_scale));
        SC_CUDA(cudaGetLastError(), "c conversion launch");
    }
    return 0;
}

}  // namespace sc

extern "C" int sc_sae_step_f16_supported(int32_t B, int32_t d, int32_t N) { return sc::f16_supported(B, d, N) ? 1 : 0; }

extern "C" int sc_codes_to_float(const void* codes_t, const float* code_scale, float* out, int32_t M, int32_t B, int32_t N,
                                 void* stream) {
    if (!codes_t || !out || M <= 0 || B <= 0 || N <= 0) return sc::fail(-1, "sc_codes_to_float: bad arguments");
    if (M > 65535 || (N + 63) / 64 > 65535) return sc::fail(-1, "sc_codes_to_float: M or N too large for one launch");
    sc::launch_k(sc::codes_to_float_kernel, dim3(sc::ceil_div(B, 64), sc::ceil_div(N, 64), M), dim3(256), 0,
                 static_cast<cudaStream_t>(stream), static_cast<const __half*>(codes_t), out, B, N, (B + 7) & ~7, code_scale);
    const cudaError_t e = cudaGetLastError();
    return e == cudaSuccess ? 0 : sc::cuda_fail(e, "codes conversion launch");
}
