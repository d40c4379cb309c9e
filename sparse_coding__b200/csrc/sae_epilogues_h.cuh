// Epilogues of the train step in fp16-operand mode.  Same mathematics as sae_epilogues.cuh; the tensors
// that only feed later GEMMs (codes c, residual r, dL/dz) are stored as fp16 - the same 11-bit
// significand as the TF32 rounding of the default mode, at half the bytes and twice the MMA rate.
// fp16 has a narrow exponent, so nothing tiny is stored: the residual is kept UNSCALED (r, not
// R = 2r/(B d)) and the 2/(B d) factor is applied in fp32 where the gradients leave the accumulators:
//   q' = r D^T = q (B d / 2);  dz' = (q' + lam d/2 [c>0]) [z>=0] = dz (B d / 2);
//   dW_e = (2/(B d)) dz'^T x;  dD = (2/(B d)) c^T r;  s = (2/(B d)) sum_b c q';  db = (2/(B d)) sum_b dz'.
#pragma once
#include "sae_epilogues.cuh"

namespace sc {

// G1: c = relu(x W^T + b) -> fp16 codes (z < 0 stored as -0), L1 partial sums
struct EpiEncodeH {
    static constexpr int kIoBufs = 1;
    struct alignas(64) Params {
        CUtensorMap c_map;  // fp16 io map of c [G][B][N]
        const float* bias;  // [G][N]
        float* l1_part;
        const int* dict_size;  // [G] or nullptr: features >= dict_size[g] are masked out
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float sum;
    SC_DEVICE EpiEncodeH(const Params& p_, const RowCtx& c_, const EpiIo& io_) : p(p_), c(c_), io(io_), sum(0.f) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        const int ncols = c.N - col;
        const int lane = threadIdx.x & 31;
        float bv[32];
        if (ncols > 0) {
            load_row32(p.bias + (long long)c.g * c.N + col, bv, ncols);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) bv[j] = 0.f;
        }
        const bool full = c.row < c.M && ncols >= 32;
        const int live_cols = p.dict_size ? p.dict_size[c.g] - col : 32;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const float z = j < live_cols ? v[j] + bv[j] : -1.f;
            if (full || (c.row < c.M && j < ncols)) sum += fmaxf(z, 0.f);
            v[j] = (z >= 0.f) ? z : -0.0f;
        }
        io_wait_stores_read();
        io_write_row_h(io.buf, lane, v);
        io_store_tile(&p.c_map, io.buf, col, c.row - lane, c.g);
        io_commit();
    }
    SC_DEVICE void finish() {
        const float s = warp_sum(sum);
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.l1_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * kEpiHalves) + c.part] = s;
        }
    }
};

// G2: r = c D - x -> fp16 residual (unscaled), MSE partial sums
struct EpiDecodeH {
    static constexpr int kIoBufs = 0;
    struct Params {
        const float* x;  // [Gx][B][d] fp32 target
        __half* r;       // [G][B][d]
        float* mse_part;
        long long x_gstride, out_gstride;
    };
    const Params& p;
    const RowCtx c;
    float sum;
    SC_DEVICE EpiDecodeH(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_), sum(0.f) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        if (c.row >= c.M || col >= c.N) return;
        const int ncols = c.N - col;
        const long long off = (long long)c.row * c.N + col;
        float xv[32];
        load_row32(p.x + c.g * p.x_gstride + off, xv, ncols);
        __half* dst = p.r + c.g * p.out_gstride + off;
        if (ncols >= 32) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                __half2 h[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float r0 = v[8 * j + 2 * t] - xv[8 * j + 2 * t], r1 = v[8 * j + 2 * t + 1] - xv[8 * j + 2 * t + 1];
                    sum += r0 * r0 + r1 * r1;
                    h[t] = __floats2half2_rn(r0, r1);
                }
                reinterpret_cast<uint4*>(dst)[j] = *reinterpret_cast<uint4*>(h);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < ncols) {
                    const float r = v[j] - xv[j];
                    sum += r * r;
                    dst[j] = __float2half_rn(r);
                }
            }
        }
    }
    SC_DEVICE void finish() {
        const float s = warp_sum(sum);
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.mse_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * kEpiHalves) + c.part] = s;
        }
    }
};

// G3 (transposed: rows = feature n, cols = batch b): dz' from q' = r D^T and the fp16 codes
struct EpiBackCodeH {
    static constexpr int kIoBufs = 0;
    struct Params {
        const __half* c;      // [G][B][N]
        __half* dz;           // [G][B][N]  (scaled by B d / 2)
        const float* l1;      // [G]
        const float* bias;    // [G][N] (tied only)
        float* db_part;       // [G][parts][N]  (scaled)
        float* s_part;        // [G][parts][N]  (scaled)
        long long cb_gstride;
        float l1_scale;       // d / 2:  lam/B * (B d / 2)
        int tied;
        int feats;
    };
    const Params& p;
    const RowCtx c;
    float db, s, l1_term, bn;
    __half nxt[32];
    SC_DEVICE EpiBackCodeH(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_), db(0.f), s(0.f) {
        l1_term = p.l1[c.g] * p.l1_scale;
        bn = (p.tied && c.row < c.M) ? p.bias[(long long)c.g * p.feats + c.row] : 0.f;
    }
    SC_DEVICE void prefetch(int co) {
        if (c.row >= c.M) return;
        const int col = c.col0 + co;
        const __half* cp = p.c + c.g * p.cb_gstride + (long long)col * p.feats + c.row;
        const int nb = c.N - col;
        if (nb >= 32) {
#pragma unroll
            for (int j = 0; j < 32; ++j) nxt[j] = cp[(long long)j * p.feats];
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) nxt[j] = (j < nb) ? cp[(long long)j * p.feats] : __float2half(0.f);
        }
    }
    template <bool FULL, bool TIED>
    SC_DEVICE void body(const __half (&cvs)[32], const float (&v)[32], __half* dp, int nb) {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (FULL || j < nb) {
                const bool neg = (__half_as_ushort(cvs[j]) & 0x8000u) != 0;  // sign bit set <=> z < 0
                const float cv = __half2float(cvs[j]);
                const float q = v[j];
                const float t = q + (cv > 0.f ? l1_term : 0.f);
                const float dzv = neg ? 0.f : t;
                dp[(long long)j * p.feats] = __float2half_rn(dzv);
                db += dzv;
                s = fmaf(cv, q, s);
                if (TIED) s = fmaf(dzv, cv - bn, s);
            }
        }
    }
    SC_DEVICE void chunk(int co, float (&v)[32], int next_co) {
        if (c.row >= c.M) return;
        const int col = c.col0 + co;
        __half* dp = p.dz + c.g * p.cb_gstride + (long long)col * p.feats + c.row;
        const int nb = c.N - col;
        __half cvs[32];
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

}  // namespace sc
