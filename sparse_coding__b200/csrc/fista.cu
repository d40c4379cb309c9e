// sc_fista: batched FISTA sparse inference (reference autoencoders/fista.py:99-128, twin :274-301).
//
// Per iteration k (fista.py:116-125):
//   Res = x - y D                      F1: [B x d] GEMM, K = N, epilogue forms the residual
//   y   = y + eta Res D^T              F2: [B x N] GEMM, K = d, epilogue does the rest:
//   a   = max(y - eta l1, 0)
//   y   = a + mom_k (a - a_prev)
// The reference runs these in fp32 and the parity gate is 1e-4 absolute on the codes after 300
// iterations, so plain TF32 products are not enough.  Every operand is kept as an exact hi + lo
// pair of TF32 numbers (hi = rn_tf32(v), lo = rn_tf32(v - hi)) and each product runs as three
// tensor-core passes hi*hi + hi*lo + lo*hi accumulated in the same TMEM tile ("3xTF32", relative
// error ~2^-21); the GEMM core's operand segments do this without extra kernels.
// The tensor core adds each instruction's products into the TMEM accumulator with truncation, a bias of
// about 3e-8 x |accumulator| per instruction for same-sign sums (measured: -2.7e-5 on a K = 2000 Gram
// diagonal over 750 instructions).  The two small cross terms therefore run FIRST, while the
// accumulator is still tiny, and hi*hi last: only K/8 instructions see the full magnitude.
#include "../../include/sc_b200.h"
#include "host_common.cuh"
#include "sae_epilogues.cuh"

namespace sc {

// v -> (hi, lo) TF32 pair, optionally also a plain copy
__global__ void split_kernel(const float4* __restrict__ in, float4* __restrict__ hi, float4* __restrict__ lo,
                             float4* __restrict__ copy, long long n4) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 v = in[i];
        float4 h, l;
        h.x = rn_tf32(v.x); h.y = rn_tf32(v.y); h.z = rn_tf32(v.z); h.w = rn_tf32(v.w);
        l.x = rn_tf32(v.x - h.x); l.y = rn_tf32(v.y - h.y); l.z = rn_tf32(v.z - h.z); l.w = rn_tf32(v.w - h.w);
        hi[i] = h;
        lo[i] = l;
        if (copy) copy[i] = v;
    }
}

// F1 epilogue: res = x - acc -> (res_hi, res_lo) for F2, or the full-precision residual at the end
struct EpiFistaRes {
    struct Params {
        const float* x;
        float *res_hi, *res_lo, *res_full;
    };
    const Params& p;
    const RowCtx c;
    static constexpr int kIoBufs = 0;
    SC_DEVICE EpiFistaRes(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        if (c.row >= c.M || col >= c.N) return;
        const int ncols = c.N - col;
        const long long off = (long long)c.row * c.N + col;
        float xv[32];
        load_row32(p.x + off, xv, ncols);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = xv[j] - v[j];
        if (p.res_full) store_row32(p.res_full + off, v, ncols);
        if (p.res_hi) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                xv[j] = rn_tf32(v[j]);
                v[j] = rn_tf32(v[j] - xv[j]);
            }
            store_row32(p.res_hi + off, xv, ncols);
            store_row32(p.res_lo + off, v, ncols);
        }
    }
    SC_DEVICE void finish() {}
};

// F2 epilogue: gradient step, shrink + ReLU, momentum (fista.py:121-123).
// y lives only as its (hi, lo) TF32 pair (hi + lo carries 22+ bits, a perturbation below fp32 rounding of
// the update itself); the three [32 x 32] tiles y_hi, y_lo, a come in and go out through the warp's
// staging tiles by TMA, in place, so the global traffic is fully coalesced.
struct EpiFistaStep {
    // HBM-heavy epilogue (six 4 KB tiles per chunk): four epilogue warps with two staging sets each, the next
    // chunk's y_hi / y_lo / a_prev tiles load while the current chunk is processed and stored (as in EpiAdam).
    static constexpr int kHalves = 1;
    static constexpr int kIoBufs = 6;
    struct alignas(64) Params {
        CUtensorMap yhi_map, ylo_map, a_map;  // io maps of [B][N]
        const float* eta;                      // device scalar
        const float* l1_dev;                   // device scalar, used instead of l1 when not null
        float l1, mom;
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float eta, thr;
    SC_DEVICE EpiFistaStep(const Params& p_, const RowCtx& c_, const EpiIo& io_) : p(p_), c(c_), io(io_) {
        eta = *p.eta;
        thr = eta * (p.l1_dev ? *p.l1_dev : p.l1);
    }
    SC_DEVICE void request(int set, int row0, int col) const {  // lane 0
        uint8_t* b = io.buf + set * (3 * kIoTileBytes);
        io_expect(io, 3 * kIoTileBytes, set);
        tma_load_3d(&p.yhi_map, io.bar + set, b, col, row0, 0);
        tma_load_3d(&p.ylo_map, io.bar + set, b + kIoTileBytes, col, row0, 0);
        tma_load_3d(&p.a_map, io.bar + set, b + 2 * kIoTileBytes, col, row0, 0);
    }
    SC_DEVICE void prefetch(int co) {  // only the warp's very first chunk; later ones are requested one chunk ahead
        if (c.first && (threadIdx.x & 31) == 0) request((co >> 5) & 1, c.row0, c.col0 + co);
    }
    SC_DEVICE void chunk(int co, float (&v)[32], int next_co) {
        const int lane = threadIdx.x & 31;
        const int set = (co >> 5) & 1;
        const int col = c.col0 + co, row0 = c.row0;
        uint8_t* b = io.buf + set * (3 * kIoTileBytes);
        io_wait_loads(io, set);
        float yh[32], yl[32], av[32];
        io_read_row(b, lane, yh);
        io_read_row(b + kIoTileBytes, lane, yl);
        io_read_row(b + 2 * kIoTileBytes, lane, av);
        io_wait_stores_read();  // the other set's store (one chunk ago) has read it
        if (lane == 0) {
            if (next_co >= 0) request(set ^ 1, row0, c.col0 + next_co);
            else if (c.next_g >= 0) request(set ^ 1, c.next_row0, c.next_col0);
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const float yn = (yh[j] + yl[j]) + eta * v[j];   // ahat_y.add(eta * Res D^T)
            const float an = fmaxf(yn - thr, 0.f);            // .sub(eta*l1).clamp(min=0)
            const float y2 = an + (an - av[j]) * p.mom;       // ahat.add(ahat.sub(ahat_pre).mul(mom))
            av[j] = an;
            yh[j] = rn_tf32(y2);
            yl[j] = rn_tf32(y2 - yh[j]);
        }
        io_write_row(b, lane, yh);
        io_write_row(b + kIoTileBytes, lane, yl);
        io_write_row(b + 2 * kIoTileBytes, lane, av);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_3d(&p.yhi_map, b, col, row0, 0);
            tma_store_3d(&p.ylo_map, b + kIoTileBytes, col, row0, 0);
            tma_store_3d(&p.a_map, b + 2 * kIoTileBytes, col, row0, 0);
        }
        io_commit();
    }
    SC_DEVICE void finish() {}
};

// ---------------------------------------------------------------------------------------------------------------
// fp16 hi / lo variant (default when the shape allows it): the same split-product scheme with the pair kept as two
// fp16 numbers, hi = half(v), lo = half(v - hi), multiplied by kind::f16 MMAs - twice the tensor rate and half the
// operand bytes of the TF32 pair.  hi + lo carries 22 significand bits wherever lo is a normal fp16 number
// (|v| >= 2^-2); below that lo is subnormal and the pair is exact to 2^-25 ABSOLUTE (3e-8), far inside the 1e-4
// absolute gate on the codes (the dictionary has unit rows, the codes are O(1)).  |v| must stay below 65504
// (conversions saturate).
__global__ void split_h_kernel(const float4* __restrict__ in, uint2* __restrict__ hi, uint2* __restrict__ lo, long long n4) {
    pdl_launch_dependents();
    pdl_wait();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        const float4 v = in[i];
        __half2 h[2] = {half2_sat(v.x, v.y), half2_sat(v.z, v.w)};
        const float2 h0 = __half22float2(h[0]), h1 = __half22float2(h[1]);
        __half2 l[2] = {half2_sat(v.x - h0.x, v.y - h0.y), half2_sat(v.z - h1.x, v.w - h1.y)};
        hi[i] = *reinterpret_cast<uint2*>(h);
        lo[i] = *reinterpret_cast<uint2*>(l);
    }
}

// 32 floats -> fp16 (hi, lo) pairs, packed two per word
SC_DEVICE void split_row_h(const float (&v)[32], uint32_t (&hi)[16], uint32_t (&lo)[16]) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const __half2 h = half2_sat(v[2 * j], v[2 * j + 1]);
        const float2 hf = __half22float2(h);
        const __half2 l = half2_sat(v[2 * j] - hf.x, v[2 * j + 1] - hf.y);
        hi[j] = *reinterpret_cast<const uint32_t*>(&h);
        lo[j] = *reinterpret_cast<const uint32_t*>(&l);
    }
}

// F1 epilogue, fp16 pair out
struct EpiFistaResH {
    struct Params {
        const float* x;
        __half *res_hi, *res_lo;
        float* res_full;
    };
    const Params& p;
    const RowCtx c;
    static constexpr int kIoBufs = 0;
    SC_DEVICE EpiFistaResH(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        if (c.row >= c.M || col >= c.N) return;
        const int ncols = c.N - col;
        const long long off = (long long)c.row * c.N + col;
        float xv[32];
        load_row32(p.x + off, xv, ncols);
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = xv[j] - v[j];
        if (p.res_full) store_row32(p.res_full + off, v, ncols);
        if (p.res_hi) {
            uint32_t hi[16], lo[16];
            split_row_h(v, hi, lo);
            if (ncols >= 32) {  // N (= d here) is a multiple of 8 and col of 32: 16-byte aligned
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    reinterpret_cast<uint4*>(p.res_hi + off)[j] = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
                    reinterpret_cast<uint4*>(p.res_lo + off)[j] = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
                }
            } else {
                const __half* hh = reinterpret_cast<const __half*>(hi);
                const __half* ll = reinterpret_cast<const __half*>(lo);
                for (int j = 0; j < ncols; ++j) { p.res_hi[off + j] = hh[j]; p.res_lo[off + j] = ll[j]; }
            }
        }
    }
    SC_DEVICE void finish() {}
};

// F2 epilogue with y kept as an fp16 (hi, lo) pair: per chunk two 2 KB fp16 tiles and one 4 KB fp32 tile (a) come in
// and go out through the warp's staging memory, two sets of 8 KB
struct EpiFistaStepH {
    static constexpr int kHalves = 1;
    static constexpr int kIoBufs = 4;  // 2 sets x (y_hi 2 KB + y_lo 2 KB + a 4 KB)
    struct alignas(64) Params {
        CUtensorMap yhi_map, ylo_map;  // fp16 io maps of [B][N]
        CUtensorMap a_map;             // fp32 io map of [B][N]
        const float* eta;
        const float* l1_dev;
        float l1, mom;
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float eta, thr;
    SC_DEVICE EpiFistaStepH(const Params& p_, const RowCtx& c_, const EpiIo& io_) : p(p_), c(c_), io(io_) {
        eta = *p.eta;
        thr = eta * (p.l1_dev ? *p.l1_dev : p.l1);
    }
    SC_DEVICE void request(int set, int row0, int col) const {  // lane 0
        uint8_t* b = io.buf + set * 8192;
        io_expect(io, 8192, set);
        tma_load_3d(&p.yhi_map, io.bar + set, b, col, row0, 0);
        tma_load_3d(&p.ylo_map, io.bar + set, b + 2048, col, row0, 0);
        tma_load_3d(&p.a_map, io.bar + set, b + 4096, col, row0, 0);
    }
    SC_DEVICE void prefetch(int co) {
        if (c.first && (threadIdx.x & 31) == 0) request((co >> 5) & 1, c.row0, c.col0 + co);
    }
    SC_DEVICE void chunk(int co, float (&v)[32], int next_co) {
        const int lane = threadIdx.x & 31;
        const int set = (co >> 5) & 1;
        const int col = c.col0 + co, row0 = c.row0;
        uint8_t* b = io.buf + set * 8192;
        io_wait_loads(io, set);
        const uint32_t rh = smem_u32(b) + lane * 64, rl = rh + 2048;
        const int sw = (lane >> 1) & 3;  // SWIZZLE_64B
        uint4 yh[4], yl[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            yh[j] = lds_128u(rh + ((j ^ sw) << 4));
            yl[j] = lds_128u(rl + ((j ^ sw) << 4));
        }
        float av[32];
        io_read_row(b + 4096, lane, av);
        io_wait_stores_read();  // the other set's store (one chunk ago) has read it
        if (lane == 0) {
            if (next_co >= 0) request(set ^ 1, row0, c.col0 + next_co);
            else if (c.next_g >= 0) request(set ^ 1, c.next_row0, c.next_col0);
        }
        float y2[32];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t wh[4] = {yh[j].x, yh[j].y, yh[j].z, yh[j].w}, wl[4] = {yl[j].x, yl[j].y, yl[j].z, yl[j].w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const float2 fh = __half22float2(*reinterpret_cast<const __half2*>(&wh[t]));
                const float2 fl = __half22float2(*reinterpret_cast<const __half2*>(&wl[t]));
                const int e = 8 * j + 2 * t;
                const float yn0 = (fh.x + fl.x) + eta * v[e], yn1 = (fh.y + fl.y) + eta * v[e + 1];
                const float an0 = fmaxf(yn0 - thr, 0.f), an1 = fmaxf(yn1 - thr, 0.f);
                y2[e] = an0 + (an0 - av[e]) * p.mom;
                y2[e + 1] = an1 + (an1 - av[e + 1]) * p.mom;
                av[e] = an0; av[e + 1] = an1;
            }
        }
        uint32_t hi[16], lo[16];
        split_row_h(y2, hi, lo);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sts_128u(rh + ((j ^ sw) << 4), hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
            sts_128u(rl + ((j ^ sw) << 4), lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
        }
        io_write_row(b + 4096, lane, av);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_3d(&p.yhi_map, b, col, row0, 0);
            tma_store_3d(&p.ylo_map, b + 2048, col, row0, 0);
            tma_store_3d(&p.a_map, b + 4096, col, row0, 0);
        }
        io_commit();
    }
    SC_DEVICE void finish() {}
};

// eta = 1 / lambda_max(G), G = D^T D [d x d] symmetric PSD (fista.py:105-106 takes it from
// eigvalsh(D D^T): same non-zero spectrum).  Plain power iteration stalls when the top eigenvalues
// are close (random dictionaries: Marchenko-Pastur edge), so the matrix is first squared kSquarings
// times on the tensor cores (T <- (T/tr T)^2, 3xTF32): T becomes the 2^k-th power of G up to scale and
// a handful of power steps on it isolate the top eigenvector; lambda is then the Rayleigh quotient of
// that vector with the ORIGINAL fp32 G, whose error is second order in the vector's.
constexpr int kSquarings = 14;

__global__ void normalise_split_kernel(const float* __restrict__ T, int d, float* __restrict__ hi, float* __restrict__ lo) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    float acc = 0.f;
    for (int i = threadIdx.x; i < d; i += blockDim.x) acc += T[(long long)i * d + i];
    const float inv = 1.0f / block_sum(acc, sh);
    const long long n = (long long)d * d;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = T[i] * inv;
        const float h = rn_tf32(v);
        hi[i] = h;
        lo[i] = rn_tf32(v - h);
    }
}

// One CTA: v <- T v (iters times, normalised), then eta = (v.v) / (v.G v).
__global__ void top_eigen_kernel(const float* __restrict__ T, const float* __restrict__ G, int d, float* __restrict__ v0,
                                 float* __restrict__ v1, float* __restrict__ out_eta, int iters) {
    pdl_launch_dependents();
    pdl_wait();
    __shared__ float sh[8];
    const int tid = threadIdx.x, nthreads = blockDim.x, warp = tid >> 5, lane = tid & 31, nwarps = nthreads >> 5;
    for (int i = tid; i < d; i += nthreads) v0[i] = 1.0f + 0.25f * (float)((i * 37) % 101) / 101.f;
    __syncthreads();
    float* vin = v0;
    float* vout = v1;
    for (int it = 0; it <= iters; ++it) {
        const float* Mx = it < iters ? T : G;
        for (int r = warp; r < d; r += nwarps) {
            float acc = 0.f;
            for (int j = lane; j < d; j += 32) acc += Mx[(long long)r * d + j] * vin[j];
            acc = warp_sum(acc);
            if (lane == 0) vout[r] = acc;
        }
        __syncthreads();
        if (it == iters) {
            float num = 0.f, den = 0.f;
            for (int i = tid; i < d; i += nthreads) { num += vin[i] * vout[i]; den += vin[i] * vin[i]; }
            num = block_sum(num, sh);
            den = block_sum(den, sh);
            if (tid == 0) *out_eta = den / num;
        } else {
            float nn = 0.f;
            for (int i = tid; i < d; i += nthreads) nn += vout[i] * vout[i];
            const float inv = rsqrtf(block_sum(nn, sh));
            for (int i = tid; i < d; i += nthreads) vout[i] *= inv;
            __syncthreads();
            float* t = vin; vin = vout; vout = t;
        }
    }
}

__global__ void set_scalar_kernel(float* p, float v) {
    pdl_launch_dependents();
    pdl_wait(); *p = v; }

struct FistaWs {
    float *D_hi, *D_lo, *y, *y_hi, *y_lo, *res_hi, *res_lo, *gram, *pw, *pw_hi, *pw_lo, *v0, *v1, *eta;
    size_t bytes;
};

// SC_B200_FISTA_F16=0 keeps the TF32 pair everywhere
static bool fista_f16(int B, int d, int N) {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SC_B200_FISTA_F16");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1 && use_cta_pairs() && d >= 128 && N >= 128 && (d % 64) == 0 && (N % 64) == 0 && B >= 1;
}

static FistaWs carve_fista(void* base, int B, int d, int N) {
    FistaWs w;
    Carver c(base);
    const size_t Bz = B, dz = d, Nz = N;
    w.D_hi = c.floats(Nz * dz); w.D_lo = c.floats(Nz * dz);
    w.y = c.floats(Bz * Nz); w.y_hi = c.floats(Bz * Nz); w.y_lo = c.floats(Bz * Nz);
    w.res_hi = c.floats(Bz * dz); w.res_lo = c.floats(Bz * dz);
    w.gram = c.floats(dz * dz); w.pw = c.floats(dz * dz); w.pw_hi = c.floats(dz * dz); w.pw_lo = c.floats(dz * dz);
    w.v0 = c.floats(dz); w.v1 = c.floats(dz); w.eta = c.floats(4);
    w.bytes = c.off;
    return w;
}

static int split(const float* in, float* hi, float* lo, float* copy, long long n, cudaStream_t st) {
    const long long n4 = n / 4;
    const int grid = static_cast<int>(n4 / 256 + 1 < 4096 ? n4 / 256 + 1 : 4096);
    launch_k(split_kernel, dim3(grid), dim3(256), 0, st, reinterpret_cast<const float4*>(in), reinterpret_cast<float4*>(hi),
                                       reinterpret_cast<float4*>(lo), reinterpret_cast<float4*>(copy), n4);
    SC_CUDA(cudaGetLastError(), "split launch");
    return 0;
}

static int split_h(const float* in, __half* hi, __half* lo, long long n, cudaStream_t st) {
    const long long n4 = n / 4;
    const int grid = static_cast<int>(n4 / 256 + 1 < 4096 ? n4 / 256 + 1 : 4096);
    launch_k(split_h_kernel, dim3(grid), dim3(256), 0, st, reinterpret_cast<const float4*>(in), reinterpret_cast<uint2*>(hi),
             reinterpret_cast<uint2*>(lo), n4);
    SC_CUDA(cudaGetLastError(), "split launch");
    return 0;
}

// The iterations with fp16 (hi, lo) operand pairs.  The fp16 arrays live in the front halves of the workspace blocks
// the TF32 pairs would occupy (D_hi / D_lo are re-split after the step-size computation, which used the TF32 pairs).
static int fista_iterations_h(const ScFistaArgs& a, const FistaWs& w, cudaStream_t st) {
    const int B = a.B, d = a.d, N = a.N;
    const long long BN_ = (long long)B * N, Nd = (long long)N * d;
    __half *D_hi = reinterpret_cast<__half*>(w.D_hi), *D_lo = reinterpret_cast<__half*>(w.D_lo);
    __half *y_hi = reinterpret_cast<__half*>(w.y_hi), *y_lo = reinterpret_cast<__half*>(w.y_lo);
    __half *r_hi = reinterpret_cast<__half*>(w.res_hi), *r_lo = reinterpret_cast<__half*>(w.res_lo);
    SC_TRY(split_h(a.D, D_hi, D_lo, Nd, st));
    SC_TRY(split_h(a.a0, y_hi, y_lo, BN_, st));
    GemmPlan p1, p2;
    {   // F1: acc = y D     A = y [B,N] K-major, B = D as an MN-major view (mn = d, k = N); cross terms first
        Seg s[3] = {{kmajor_h(y_hi, B, N, 1, 0), mnmajor_h(D_lo, d, N, 1, 0)},
                    {kmajor_h(y_lo, B, N, 1, 0), mnmajor_h(D_hi, d, N, 1, 0)},
                    {kmajor_h(y_hi, B, N, 1, 0), mnmajor_h(D_hi, d, N, 1, 0)}};
        SC_TRY(plan_gemm(&p1, B, d, 1, s, 3, false, true));
    }
    {   // F2: acc = Res D^T  A = Res [B,d] K-major, B = D [N,d] K-major
        Seg s[3] = {{kmajor_h(r_hi, B, d, 1, 0), kmajor_h(D_lo, N, d, 1, 0)},
                    {kmajor_h(r_lo, B, d, 1, 0), kmajor_h(D_hi, N, d, 1, 0)},
                    {kmajor_h(r_hi, B, d, 1, 0), kmajor_h(D_hi, N, d, 1, 0)}};
        SC_TRY(plan_gemm(&p2, B, N, 1, s, 3, false, false));
    }
    EpiFistaStepH::Params e2{};
    {
        const char* err = nullptr;
        int rc = make_io_map_h(&e2.yhi_map, y_hi, B, N, 1, 0, &err);
        if (rc == 0) rc = make_io_map_h(&e2.ylo_map, y_lo, B, N, 1, 0, &err);
        if (rc == 0) rc = make_io_map(&e2.a_map, a.out_a, B, N, 1, 0, &err);
        if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
    }
    e2.eta = w.eta; e2.l1 = a.l1_coef; e2.l1_dev = a.l1_coef_dev;
    for (int k = 0; k < a.num_iter; ++k) {
        EpiFistaResH::Params e1{a.x, r_hi, r_lo, nullptr};
        SC_TRY((launch_plan<false, true, EpiFistaResH, true>(p1, e1, st)));
        e2.mom = a.momentum[k];
        SC_TRY((launch_plan<false, false, EpiFistaStepH, true>(p2, e2, st)));
    }
    // final residual from ahat (fista.py:126)
    SC_TRY(split_h(a.out_a, y_hi, y_lo, BN_, st));
    EpiFistaResH::Params e3{a.x, nullptr, nullptr, a.out_res};
    SC_TRY((launch_plan<false, true, EpiFistaResH, true>(p1, e3, st)));
    if (a.out_eta) SC_CUDA(cudaMemcpyAsync(a.out_eta, w.eta, sizeof(float), cudaMemcpyDeviceToDevice, st), "copy eta");
    return 0;
}

static int fista_impl(const ScFistaArgs& a) {
    const int B = a.B, d = a.d, N = a.N;
    if (B <= 0 || d <= 0 || N <= 0 || a.num_iter < 0) return fail(-1, "bad FISTA shape");
    if ((d & 3) || (N & 3)) return fail(-1, "d and N must be multiples of 4 (got d=%d N=%d)", d, N);
    if (!a.x || !a.D || !a.a0 || !a.out_a || !a.out_res || !a.workspace) return fail(-1, "NULL pointer argument");
    if (a.num_iter > 0 && !a.momentum) return fail(-1, "momentum is NULL");
    const FistaWs w = carve_fista(a.workspace, B, d, N);
    if (a.workspace_bytes < w.bytes) return fail(-1, "workspace too small: %zu given, %zu needed", a.workspace_bytes, w.bytes);
    cudaStream_t st = static_cast<cudaStream_t>(a.stream);
    const long long BN_ = (long long)B * N, Nd = (long long)N * d;

    SC_TRY(split(a.D, w.D_hi, w.D_lo, nullptr, Nd, st));
    const bool f16 = a.pair_format == 0 && fista_f16(B, d, N);
    if (!f16) SC_TRY(split(a.a0, w.y_hi, w.y_lo, nullptr, BN_, st));
    SC_CUDA(cudaMemcpyAsync(a.out_a, a.a0, BN_ * sizeof(float), cudaMemcpyDeviceToDevice, st), "copy a0");

    if (a.eta > 0.f) {
        launch_k(set_scalar_kernel, dim3(1), dim3(1), 0, st, w.eta, a.eta);
    } else {
        // G = D^T D in 3xTF32: both operands MN-major views of D (mn = d, k = N)
        Seg s[3] = {{mnmajor(w.D_hi, d, N, 1, 0), mnmajor(w.D_lo, d, N, 1, 0)},
                    {mnmajor(w.D_lo, d, N, 1, 0), mnmajor(w.D_hi, d, N, 1, 0)},
                    {mnmajor(w.D_hi, d, N, 1, 0), mnmajor(w.D_hi, d, N, 1, 0)}};
        EpiStore::Params ep{w.gram, d, 0};
        SC_TRY((run_gemm<true, true, EpiStore>(d, d, 1, s, 3, ep, st)));
        // repeated squaring of the trace-normalised matrix (symmetric, so B = A works K-major)
        Seg q[3] = {{kmajor(w.pw_hi, d, d, 1, 0), kmajor(w.pw_lo, d, d, 1, 0)},
                    {kmajor(w.pw_lo, d, d, 1, 0), kmajor(w.pw_hi, d, d, 1, 0)},
                    {kmajor(w.pw_hi, d, d, 1, 0), kmajor(w.pw_hi, d, d, 1, 0)}};
        GemmPlan sq;
        SC_TRY(plan_gemm(&sq, d, d, 1, q, 3, false, false));
        EpiStore::Params eq{w.pw, d, 0};
        const int ngrid = ceil_div((long long)d * d, 256 * 4) < 148 ? ceil_div((long long)d * d, 256 * 4) : 148;
        for (int k = 0; k < kSquarings; ++k) {
            launch_k(normalise_split_kernel, dim3(ngrid), dim3(256), 0, st, k == 0 ? w.gram : w.pw, d, w.pw_hi, w.pw_lo);
            SC_TRY((launch_plan<false, false, EpiStore>(sq, eq, st)));
        }
        launch_k(top_eigen_kernel, dim3(1), dim3(256), 0, st, w.pw, w.gram, d, w.v0, w.v1, w.eta, 6);
    }
    SC_CUDA(cudaGetLastError(), "eta launch");
    if (f16) return fista_iterations_h(a, w, st);

    // F1: acc = y D     A = y [B,N] K-major (k = N), B = D MN-major view (mn = d, k = N)
    GemmPlan p1, p2;
    {
        Seg s[3] = {{kmajor(w.y_hi, B, N, 1, 0), mnmajor(w.D_lo, d, N, 1, 0)},
                    {kmajor(w.y_lo, B, N, 1, 0), mnmajor(w.D_hi, d, N, 1, 0)},
                    {kmajor(w.y_hi, B, N, 1, 0), mnmajor(w.D_hi, d, N, 1, 0)}};
        SC_TRY(plan_gemm(&p1, B, d, 1, s, 3, false, true));
    }
    // F2: acc = Res D^T  A = Res [B,d] K-major, B = D [N,d] K-major
    {
        Seg s[3] = {{kmajor(w.res_hi, B, d, 1, 0), kmajor(w.D_lo, N, d, 1, 0)},
                    {kmajor(w.res_lo, B, d, 1, 0), kmajor(w.D_hi, N, d, 1, 0)},
                    {kmajor(w.res_hi, B, d, 1, 0), kmajor(w.D_hi, N, d, 1, 0)}};
        SC_TRY(plan_gemm(&p2, B, N, 1, s, 3, false, false));
    }
    EpiFistaStep::Params e2{};
    {
        const char* err = nullptr;
        int rc = make_io_map(&e2.yhi_map, w.y_hi, B, N, 1, 0, &err);
        if (rc == 0) rc = make_io_map(&e2.ylo_map, w.y_lo, B, N, 1, 0, &err);
        if (rc == 0) rc = make_io_map(&e2.a_map, a.out_a, B, N, 1, 0, &err);
        if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
    }
    e2.eta = w.eta; e2.l1 = a.l1_coef; e2.l1_dev = a.l1_coef_dev;
    for (int k = 0; k < a.num_iter; ++k) {
        EpiFistaRes::Params e1{a.x, w.res_hi, w.res_lo, nullptr};
        SC_TRY((launch_plan<false, true, EpiFistaRes>(p1, e1, st)));
        e2.mom = a.momentum[k];
        SC_TRY((launch_plan<false, false, EpiFistaStep>(p2, e2, st)));
    }
    // final residual from ahat (fista.py:126): reuse the y_hi/y_lo operand buffers for the split of a
    SC_TRY(split(a.out_a, w.y_hi, w.y_lo, nullptr, BN_, st));
    EpiFistaRes::Params e3{a.x, nullptr, nullptr, a.out_res};
    SC_TRY((launch_plan<false, true, EpiFistaRes>(p1, e3, st)));
    if (a.out_eta) SC_CUDA(cudaMemcpyAsync(a.out_eta, w.eta, sizeof(float), cudaMemcpyDeviceToDevice, st), "copy eta");
    return 0;
}

}  // namespace sc

extern "C" {

size_t sc_fista_workspace_bytes(int32_t B, int32_t d, int32_t N) {
    if (B <= 0 || d <= 0 || N <= 0) return 0;
    return sc::carve_fista(nullptr, B, d, N).bytes;
}

int sc_fista(const ScFistaArgs* args) {
    if (!args) return sc::fail(-1, "args is NULL");
    return sc::fista_impl(*args);
}

size_t sc_gemm_f32_workspace_bytes(int32_t G, int32_t M, int32_t N, int32_t K) {
    if (G <= 0 || M <= 0 || N <= 0 || K <= 0) return 0;
    sc::Carver c(nullptr);
    c.floats((size_t)G * M * K); c.floats((size_t)G * M * K);
    c.floats((size_t)G * N * K); c.floats((size_t)G * N * K);
    return c.off;
}

int sc_gemm_f32(const float* A, const float* Bm, float* C, int32_t G, int32_t M, int32_t N, int32_t K, int32_t a_mn,
                int32_t b_mn, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace sc;
    if (!A || !Bm || !C || !workspace) return fail(-1, "NULL operand");
    if (M <= 0 || N <= 0 || K <= 0 || G <= 0) return fail(-1, "bad shape");
    if ((a_mn ? M : K) & 3 || (b_mn ? N : K) & 3 || (N & 3)) return fail(-1, "leading dimensions must be multiples of 4");
    if (workspace_bytes < sc_gemm_f32_workspace_bytes(G, M, N, K)) return fail(-1, "workspace too small");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    Carver c(workspace);
    const long long na = (long long)G * M * K, nb = (long long)G * N * K;
    float *a_hi = c.floats(na), *a_lo = c.floats(na), *b_hi = c.floats(nb), *b_lo = c.floats(nb);
    SC_TRY(split(A, a_hi, a_lo, nullptr, na, st));
    SC_TRY(split(Bm, b_hi, b_lo, nullptr, nb, st));
    auto opa = [&](const float* p) { return a_mn ? mnmajor(p, M, K, G, (long long)M * K) : kmajor(p, M, K, G, (long long)M * K); };
    auto opb = [&](const float* p) { return b_mn ? mnmajor(p, N, K, G, (long long)N * K) : kmajor(p, N, K, G, (long long)N * K); };
    // cross terms first: the tensor core adds each instruction's products into the accumulator with truncation
    Seg s[3] = {{opa(a_hi), opb(b_lo)}, {opa(a_lo), opb(b_hi)}, {opa(a_hi), opb(b_hi)}};
    EpiStore::Params ep{C, N, (long long)M * N};
    if (!a_mn && !b_mn) return run_gemm<false, false, EpiStore>(M, N, G, s, 3, ep, st);
    if (a_mn && !b_mn) return run_gemm<true, false, EpiStore>(M, N, G, s, 3, ep, st);
    if (!a_mn && b_mn) return run_gemm<false, true, EpiStore>(M, N, G, s, 3, ep, st);
    return run_gemm<true, true, EpiStore>(M, N, G, s, 3, ep, st);
}

}  // extern "C"
