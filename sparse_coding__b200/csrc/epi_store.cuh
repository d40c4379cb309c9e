// Plain "write the accumulator" epilogue: C[g][row][col] = acc.  Used by the GEMM self-test
// and as the pattern every fused epilogue follows (constructor, chunk() per 32 columns, finish()).
#pragma once
#include "gemm_core.cuh"

namespace sc {

// store 32 consecutive floats of one row; vectorised when the row pointer allows it
SC_DEVICE void store_row32(float* dst, const float (&v)[32], int ncols) {
    if (ncols >= 32 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            reinterpret_cast<float4*>(dst)[j] = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
            if (j < ncols) dst[j] = v[j];
    }
}

SC_DEVICE void load_row32(const float* src, float (&v)[32], int ncols) {
    if (ncols >= 32 && (reinterpret_cast<uintptr_t>(src) & 15) == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 t = reinterpret_cast<const float4*>(src)[j];
            v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = (j < ncols) ? src[j] : 0.f;
    }
}

struct EpiStore {
    struct Params {
        float* C;
        long long ld, gstride;
    };
    const Params& p;
    const RowCtx c;
    static constexpr int kIoBufs = 0;
    SC_DEVICE EpiStore(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        if (c.row >= c.M) return;
        const int col = c.col0 + co;
        if (col >= c.N) return;
        store_row32(p.C + c.g * p.gstride + (long long)c.row * p.ld + col, v, c.N - col);
    }
    SC_DEVICE void finish() {}
};

}  // namespace sc
