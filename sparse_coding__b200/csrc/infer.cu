// Inference-side entry points: LearnedDict.encode / decode (reference autoencoders/learned_dict.py:
// 29-32, 118-131, 172-196) and the reductions of standard_metrics.py:303-312, 439-452, 480-509 fused
// into their epilogues, plus the raw GEMM hook used by the tests.
#include "../../include/sc_b200.h"
#include "host_common.cuh"
#include "sae_epilogues_h.cuh"

namespace sc {

__global__ void round_copy_kernel(const float4* __restrict__ in, float4* __restrict__ out, long long n4);
__global__ void prep_params_kernel(const float* __restrict__ enc, float* __restrict__ enc_r,
                                   const float* __restrict__ W, float* __restrict__ D_r,
                                   float* __restrict__ inv_norm, int N, int d, int normalise);

static int round_copy(const float* in, float* out, long long n, cudaStream_t st) {
    const long long n4 = n / 4;
    const int grid = static_cast<int>(n4 / 256 + 1 < 4096 ? n4 / 256 + 1 : 4096);
    round_copy_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(out), n4);
    SC_CUDA(cudaGetLastError(), "round_copy launch");
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Transposed encode with per-feature statistics: rows = feature n, cols = batch b.
// A thread owns one feature, so "how many rows fire this feature" and the raw moments are
// in-thread sums; they are added to the caller's accumulators with one atomic per tile.
// Counts are integers below 2^24, so the float atomics are exact and order independent.
// ---------------------------------------------------------------------------------------------
struct EpiEncodeT {
    struct Params {
        const float* bias;  // [G][N] or nullptr
        float* c;           // [G][B][N] or nullptr
        float* nonzero;     // [G][N] or nullptr
        float* moments;     // [G][4][N] or nullptr
        long long c_gstride;
        int feats;
    };
    const Params& p;
    const RowCtx c;
    float bn, cnt, m1, m2, m3, m4;
    static constexpr int kIoBufs = 0;
    SC_DEVICE EpiEncodeT(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_), cnt(0.f), m1(0.f), m2(0.f), m3(0.f), m4(0.f) {
        bn = (p.bias && c.row < c.M) ? p.bias[(long long)c.g * p.feats + c.row] : 0.f;
    }
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        if (c.row >= c.M) return;
        const int col = c.col0 + co;
        const int nb = c.N - col;
        float* cp = p.c ? p.c + c.g * p.c_gstride + (long long)col * p.feats + c.row : nullptr;
        if (!cp && !p.moments && nb >= 32) {  // counts only (mean L0, dead features): the common metric call
#pragma unroll
            for (int j = 0; j < 32; ++j) cnt += (v[j] + bn > 0.f) ? 1.f : 0.f;
            return;
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < nb) {
                const float cv = fmaxf(v[j] + bn, 0.f);
                if (cp) cp[(long long)j * p.feats] = cv;
                cnt += cv != 0.f ? 1.f : 0.f;
                const float c2 = cv * cv;
                m1 += cv; m2 += c2; m3 += c2 * cv; m4 += c2 * c2;
            }
        }
    }
    SC_DEVICE void finish() {
        if (c.row >= c.M) return;
        const long long o = (long long)c.g * p.feats + c.row;
        if (p.nonzero) atomicAdd(p.nonzero + o, cnt);
        if (p.moments) {
            float* mp = p.moments + (long long)c.g * 4 * p.feats + c.row;
            atomicAdd(mp, m1);
            atomicAdd(mp + p.feats, m2);
            atomicAdd(mp + 2 * p.feats, m3);
            atomicAdd(mp + 3 * p.feats, m4);
        }
    }
};

// out = W / max(||W||_2, 1e-8) per row, fp32 exact division (get_learned_dict); warp per row.
__global__ void row_normalise_kernel(const float* __restrict__ W, float* __restrict__ out, int N, int d) {
    pdl_launch_dependents();
    pdl_wait();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.x * 8 + warp;
    if (n >= N) return;
    const long long row = ((long long)blockIdx.y * N + n) * d;
    float ss = 0.f;
    for (int i = lane; i < d; i += 32) { const float v = W[row + i]; ss += v * v; }
    ss = warp_sum(ss);
    const float cl = fmaxf(sqrtf(ss), 1e-8f);
    for (int i = lane; i < d; i += 32) out[row + i] = __fdiv_rn(W[row + i], cl);
}

// column sums of x [B,d]: blockDim (32,8), grid (ceil(d/32), row_blocks)
__global__ void column_sum_kernel(const float* __restrict__ x, float* __restrict__ colsum, int B, int d) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8][33];
    const int j = blockIdx.x * 32 + threadIdx.x;
    float acc = 0.f;
    if (j < d)
        for (long long b = blockIdx.y * 8 + threadIdx.y; b < B; b += (long long)gridDim.y * 8) acc += x[b * d + j];
    sh[threadIdx.y][threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.y == 0 && j < d) {
        float t = 0.f;
        for (int r = 0; r < 8; ++r) t += sh[r][threadIdx.x];
        atomicAdd(colsum + j, t);
    }
}

// sum over all elements of (x - colmean)^2
__global__ void column_sqdev_kernel(const float* __restrict__ x, const float* __restrict__ colsum, float* __restrict__ out,
                                    int B, int d) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    const int j = blockIdx.x * 32 + threadIdx.x;
    float acc = 0.f;
    if (j < d) {
        const float mean = colsum[j] / (float)B;
        for (long long b = blockIdx.y * 8 + threadIdx.y; b < B; b += (long long)gridDim.y * 8) {
            const float t = x[b * d + j] - mean;
            acc += t * t;
        }
    }
    acc = warp_sum(acc);
    if (threadIdx.x == 0) sh[threadIdx.y] = acc;
    __syncthreads();
    if (threadIdx.x == 0 && threadIdx.y == 0) {
        float t = 0.f;
        for (int r = 0; r < 8; ++r) t += sh[r];
        atomicAdd(out, t);
    }
}

__global__ void sum_parts_kernel(const float* __restrict__ part, int n, float* __restrict__ out) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    float acc = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) acc += part[(long long)blockIdx.x * n + i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < 8; ++i) t += sh[i];
        out[blockIdx.x] = t;
    }
}

// fp16 operand copies for the inference entry points (kernels of sae_step_h.cu)
__global__ void to_half_kernel(const float4* __restrict__ in, uint2* __restrict__ out, long long n4);
__global__ void prep_params_h_kernel(const float* __restrict__ enc, __half* __restrict__ enc_h,
                                     const float* __restrict__ W, __half* __restrict__ D_h,
                                     float* __restrict__ inv_norm, int N, int d);

static int to_half(const float* in, __half* out, long long n, cudaStream_t st) {
    const long long n4 = n / 4;
    const int grid = static_cast<int>(n4 / 256 + 1 < 4096 ? n4 / 256 + 1 : 4096);
    launch_k(to_half_kernel, dim3(grid), dim3(256), 0, st, reinterpret_cast<const float4*>(in), reinterpret_cast<uint2*>(out), n4);
    SC_CUDA(cudaGetLastError(), "to_half launch");
    return 0;
}

static bool infer_f16_supported(int B, int d, int N, int transposed) {
    return use_cta_pairs() && d >= 128 && N >= 128 && (d % 64) == 0 && (N % 64) == 0 && (!transposed || B > 64);
}

static int encode_impl(const ScEncodeArgs& a) {
    const int G = a.G, B = a.B, d = a.d, N = a.N;
    if (G <= 0 || B <= 0 || d <= 0 || N <= 0) return fail(-1, "G, B, d, N must be positive");
    if ((d & 3) || (N & 3)) return fail(-1, "d and N must be multiples of 4 (got d=%d N=%d)", d, N);
    if (!a.x || !a.W || !a.workspace) return fail(-1, "x / W / workspace is NULL");
    if (!a.out_c && !a.out_nonzero && !a.out_moments && !a.out_c_half) return fail(-1, "no output requested");
    if (a.out_c_half && (a.operand_format != 1 || a.out_nonzero || a.out_moments))
        return fail(-1, "out_c_half needs operand_format 1 and no statistics outputs");
    const size_t need = sc_encode_workspace_bytes(G, B, d, N, a.x_per_group);
    if (a.workspace_bytes < need) return fail(-1, "workspace too small: %zu given, %zu needed", a.workspace_bytes, need);
    cudaStream_t st = static_cast<cudaStream_t>(a.stream);
    Carver cv(a.workspace);
    const int xg = a.x_per_group ? G : 1;
    const long long Bd = (long long)B * d, Nd = (long long)N * d, BN_ = (long long)B * N;
    const bool stats = a.out_nonzero || a.out_moments;
    if (a.operand_format == 1) {
        if (!infer_f16_supported(B, d, N, stats))
            return fail(-1, "fp16 operand mode needs d, N multiples of 64 and >= 128%s (got B=%d d=%d N=%d)",
                        stats ? " and B > 64" : "", B, d, N);
        __half* x_h = reinterpret_cast<__half*>(cv.floats(((size_t)xg * Bd + 1) / 2));
        __half* W_h = reinterpret_cast<__half*>(cv.floats(((size_t)G * Nd + 1) / 2));
        SC_TRY(to_half(a.x, x_h, xg * Bd, st));
        SC_TRY(to_half(a.W, W_h, G * Nd, st));
        if (stats) {
            Seg s{kmajor_h(W_h, N, d, G, Nd), kmajor_h(x_h, B, d, xg, Bd)};
            EpiEncodeT::Params ep{a.bias, a.out_c, a.out_nonzero, a.out_moments, BN_, N};
            return run_gemm<false, false, EpiEncodeT, true>(N, B, G, &s, 1, ep, st);
        }
        Seg s{kmajor_h(x_h, B, d, xg, Bd), kmajor_h(W_h, N, d, G, Nd)};
        if (a.out_c_half) {  // fp16 codes for sc_decode (predict / FVU): the train step's encode epilogue
            if (reinterpret_cast<uintptr_t>(a.out_c_half) & 15) return fail(-1, "out_c_half is not 16-byte aligned");
            EpiEncodeH::Params eh{};
            const char* err = nullptr;
            const int rc = make_io_map_h(&eh.c_map, a.out_c_half, B, N, G, BN_, &err);
            if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
            eh.bias = a.bias; eh.l1_part = nullptr; eh.dict_size = nullptr;
            SC_TRY((run_gemm<false, false, EpiEncodeH, true>(B, N, G, &s, 1, eh, st)));
            if (!a.out_c) return 0;
        }
        EpiEncode::Params ep{};
        const char* err = nullptr;
        const int rc = make_io_map(&ep.c_map, a.out_c, B, N, G, BN_, &err);
        if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
        ep.bias = a.bias; ep.l1_part = nullptr; ep.signed_zero = 0; ep.dict_size = nullptr;
        return run_gemm<false, false, EpiEncode, true>(B, N, G, &s, 1, ep, st);
    }
    float* x_r = cv.floats((size_t)xg * Bd);
    float* W_r = cv.floats((size_t)G * Nd);
    SC_TRY(round_copy(a.x, x_r, xg * Bd, st));
    SC_TRY(round_copy(a.W, W_r, G * Nd, st));
    if (stats) {
        Seg s{kmajor(W_r, N, d, G, Nd), kmajor(x_r, B, d, xg, Bd)};
        EpiEncodeT::Params ep{a.bias, a.out_c, a.out_nonzero, a.out_moments, BN_, N};
        SC_TRY((run_gemm<false, false, EpiEncodeT>(N, B, G, &s, 1, ep, st)));
    } else {
        Seg s{kmajor(x_r, B, d, xg, Bd), kmajor(W_r, N, d, G, Nd)};
        EpiEncode::Params ep{};
        {
            const char* err = nullptr;
            const int rc = make_io_map(&ep.c_map, a.out_c, B, N, G, BN_, &err);
            if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
        }
        ep.bias = a.bias; ep.l1_part = nullptr; ep.signed_zero = 0; ep.dict_size = nullptr;
        SC_TRY((run_gemm<false, false, EpiEncode>(B, N, G, &s, 1, ep, st)));
    }
    return 0;
}

static int decode_impl(const ScDecodeArgs& a) {
    const int G = a.G, B = a.B, d = a.d, N = a.N;
    if (G <= 0 || B <= 0 || d <= 0 || N <= 0) return fail(-1, "G, B, d, N must be positive");
    if ((d & 3) || (N & 3)) return fail(-1, "d and N must be multiples of 4 (got d=%d N=%d)", d, N);
    if ((!a.c && !a.c_half) || !a.Wd || !a.workspace) return fail(-1, "c / Wd / workspace is NULL");
    if (a.c_half && a.operand_format != 1) return fail(-1, "c_half needs operand_format 1");
    if (a.out_sq_err && !a.target) return fail(-1, "out_sq_err needs a target");
    const size_t need = sc_decode_workspace_bytes(G, B, d, N);
    if (a.workspace_bytes < need) return fail(-1, "workspace too small: %zu given, %zu needed", a.workspace_bytes, need);
    cudaStream_t st = static_cast<cudaStream_t>(a.stream);
    Carver cv(a.workspace);
    const long long Bd = (long long)B * d, Nd = (long long)N * d, BN_ = (long long)B * N;
    const int n_part = row_tiles128(B) * ceil_div(d, pick_bn(d)) * 4 * kEpiHalves;
    if (a.operand_format == 1) {
        if (!infer_f16_supported(B, d, N, 0))
            return fail(-1, "fp16 operand mode needs d, N multiples of 64 and >= 128 (got d=%d N=%d)", d, N);
        __half* c_h = reinterpret_cast<__half*>(cv.floats(((size_t)G * BN_ + 1) / 2));
        __half* D_h = reinterpret_cast<__half*>(cv.floats(((size_t)G * Nd + 1) / 2));
        float* part = cv.floats((size_t)G * n_part);
        if (a.c_half) c_h = const_cast<__half*>(static_cast<const __half*>(a.c_half));
        else SC_TRY(to_half(a.c, c_h, G * BN_, st));
        if (a.normalise) {
            launch_k(prep_params_h_kernel, dim3(ceil_div(N, 8), G), dim3(256), 0, st, nullptr, nullptr, a.Wd, D_h, nullptr, N, d);
            SC_CUDA(cudaGetLastError(), "prep launch");
        } else {
            SC_TRY(to_half(a.Wd, D_h, G * Nd, st));
        }
        Seg s{kmajor_h(c_h, B, N, G, BN_), mnmajor_h(D_h, d, N, G, Nd)};
        EpiDecode::Params ep{a.target, nullptr, a.out_x_hat, a.out_sq_err ? part : nullptr, a.target_per_group ? Bd : 0, Bd, 1.f};
        SC_TRY((run_gemm<false, true, EpiDecode, true>(B, d, G, &s, 1, ep, st)));
        if (a.out_sq_err) {
            launch_k(sum_parts_kernel, dim3(G), dim3(256), 0, st, part, n_part, a.out_sq_err);
            SC_CUDA(cudaGetLastError(), "sum_parts launch");
        }
        return 0;
    }
    float* c_r = cv.floats((size_t)G * BN_);
    float* D_r = cv.floats((size_t)G * Nd);
    float* part = cv.floats((size_t)G * n_part);
    SC_TRY(round_copy(a.c, c_r, G * BN_, st));
    dim3 grid(ceil_div(N, 8), G);
    prep_params_kernel<<<grid, 256, 0, st>>>(nullptr, nullptr, a.Wd, D_r, nullptr, N, d, a.normalise);
    SC_CUDA(cudaGetLastError(), "prep launch");
    Seg s{kmajor(c_r, B, N, G, BN_), mnmajor(D_r, d, N, G, Nd)};
    EpiDecode::Params ep{a.target, nullptr, a.out_x_hat, a.out_sq_err ? part : nullptr, a.target_per_group ? Bd : 0, Bd, 1.f};
    SC_TRY((run_gemm<false, true, EpiDecode>(B, d, G, &s, 1, ep, st)));
    if (a.out_sq_err) {
        sum_parts_kernel<<<G, 256, 0, st>>>(part, n_part, a.out_sq_err);
        SC_CUDA(cudaGetLastError(), "sum_parts launch");
    }
    return 0;
}

}  // namespace sc

extern "C" {

size_t sc_encode_workspace_bytes(int32_t G, int32_t B, int32_t d, int32_t N, int32_t x_per_group) {
    sc::Carver cv(nullptr);
    cv.floats((size_t)(x_per_group ? G : 1) * B * d);
    cv.floats((size_t)G * N * d);
    return cv.off;
}

int sc_encode(const ScEncodeArgs* args) {
    if (!args) return sc::fail(-1, "args is NULL");
    return sc::encode_impl(*args);
}

size_t sc_decode_workspace_bytes(int32_t G, int32_t B, int32_t d, int32_t N) {
    sc::Carver cv(nullptr);
    cv.floats((size_t)G * B * N);
    cv.floats((size_t)G * N * d);
    cv.floats((size_t)G * sc::row_tiles128(B) * sc::ceil_div(d, sc::pick_bn(d)) * 4 * sc::kEpiHalves);
    return cv.off;
}

int sc_infer_f16_supported(int32_t B, int32_t d, int32_t N, int32_t transposed) {
    return sc::infer_f16_supported(B, d, N, transposed) ? 1 : 0;
}

int sc_decode(const ScDecodeArgs* args) {
    if (!args) return sc::fail(-1, "args is NULL");
    return sc::decode_impl(*args);
}

int sc_row_normalise(const float* W, float* out, int32_t G, int32_t N, int32_t d, void* stream) {
    using namespace sc;
    if (!W || !out || G <= 0 || N <= 0 || d <= 0) return fail(-1, "bad argument");
    dim3 grid(ceil_div(N, 8), G);
    row_normalise_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(W, out, N, d);
    SC_CUDA(cudaGetLastError(), "row_normalise launch");
    return 0;
}

int sc_column_variance(const float* x, float* out_sum, int32_t B, int32_t d, void* workspace, void* stream) {
    using namespace sc;
    if (!x || !out_sum || !workspace || B <= 0 || d <= 0) return fail(-1, "bad argument");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    float* colsum = static_cast<float*>(workspace);
    SC_CUDA(cudaMemsetAsync(colsum, 0, sizeof(float) * d, st), "memset");
    SC_CUDA(cudaMemsetAsync(out_sum, 0, sizeof(float), st), "memset");
    const int rb = B / 64 + 1 < 512 ? B / 64 + 1 : 512;
    dim3 grid(ceil_div(d, 32), rb), block(32, 8);
    column_sum_kernel<<<grid, block, 0, st>>>(x, colsum, B, d);
    column_sqdev_kernel<<<grid, block, 0, st>>>(x, colsum, out_sum, B, d);
    SC_CUDA(cudaGetLastError(), "column_variance launch");
    return 0;
}

int sc_gemm_tf32(const float* A, const float* Bm, float* C, int32_t G, int32_t M, int32_t N, int32_t K, int32_t a_mn,
                 int32_t b_mn, void* stream) {
    using namespace sc;
    if (!A || !Bm || !C) return fail(-1, "NULL operand");
    if (M <= 0 || N <= 0 || K <= 0 || G <= 0) return fail(-1, "bad shape");
    if ((a_mn ? M : K) & 3 || (b_mn ? N : K) & 3 || (N & 3)) return fail(-1, "leading dimensions must be multiples of 4");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Seg s{a_mn ? mnmajor(A, M, K, G, (long long)M * K) : kmajor(A, M, K, G, (long long)M * K),
          b_mn ? mnmajor(Bm, N, K, G, (long long)N * K) : kmajor(Bm, N, K, G, (long long)N * K)};
    EpiStore::Params ep{C, N, (long long)M * N};
    if (!a_mn && !b_mn) return run_gemm<false, false, EpiStore>(M, N, G, &s, 1, ep, st);
    if (a_mn && !b_mn) return run_gemm<true, false, EpiStore>(M, N, G, &s, 1, ep, st);
    if (!a_mn && b_mn) return run_gemm<false, true, EpiStore>(M, N, G, &s, 1, ep, st);
    return run_gemm<true, true, EpiStore>(M, N, G, &s, 1, ep, st);
}

}  // extern "C"
