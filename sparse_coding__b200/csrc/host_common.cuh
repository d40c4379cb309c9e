// Host-side plumbing shared by the C-ABI entry points: error reporting, workspace carving and the
// "describe operands -> build TMA maps -> launch the tcgen05 core" helper.
#pragma once
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <stdlib.h>

#include "gemm_core.cuh"
#include "gemm_core2.cuh"
#include "gemm_core2_res.cuh"
#include "tmap.h"

namespace sc {

extern thread_local char g_last_error[512];
extern thread_local int g_dbg_sm_limit;   // probe: > 0 caps the number of CTAs of the GEMM launches of this thread
extern thread_local int g_dbg_phase;      // probe: bit 0 = prologue..G3, bit 1 = weight gradients + Adam (default 3)

inline int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_last_error, sizeof g_last_error, fmt, ap);
    va_end(ap);
    return code;
}

inline int cuda_fail(cudaError_t e, const char* what) {
    snprintf(g_last_error, sizeof g_last_error, "%s: %s", what, cudaGetErrorString(e));
    return static_cast<int>(e);
}

#define SC_TRY(expr)            \
    do {                        \
        int rc_ = (expr);       \
        if (rc_ != 0) return rc_; \
    } while (0)

#define SC_CUDA(expr, what)                              \
    do {                                                 \
        cudaError_t e_ = (expr);                         \
        if (e_ != cudaSuccess) return cuda_fail(e_, what); \
    } while (0)

inline int device_sm_count(int* out) {
    static int cached[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice");
    if (!cached[dev & 63]) {
        int cc_major = 0;
        e = cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev);
        if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
        if (cc_major != 10)
            return fail(-2, "this library is built for sm_100a only; device %d has compute capability %d.x", dev, cc_major);
        int n = 0;
        e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceGetAttribute");
        cached[dev & 63] = n;
    }
    *out = cached[dev & 63];
    return 0;
}

// Optional per-launch timing: when enabled (sc_profile_begin), every mark records an event on the
// caller's stream; sc_profile_end returns the elapsed time between consecutive marks.
struct Profiler {
    bool on = false;
    int n = 0;
    static constexpr int kMax = 4096;
    cudaEvent_t ev[kMax];
    const char* name[kMax];
};
extern Profiler g_prof;
inline void prof_mark(cudaStream_t st, const char* name) {
    if (!g_prof.on || g_prof.n >= Profiler::kMax) return;
    cudaEvent_t e;
    if (cudaEventCreate(&e) != cudaSuccess) return;
    cudaEventRecord(e, st);
    g_prof.ev[g_prof.n] = e;
    g_prof.name[g_prof.n] = name;
    ++g_prof.n;
}

// Plain (non-GEMM) kernels of a step are launched with the programmatic-stream-serialization attribute so that
// their launch latency overlaps the tail of the previous kernel; each of them starts with pdl_wait().
template <class... KArgs, class... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = use_pdl_plain() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// Bump allocator over the caller's workspace; every block is 1024-byte aligned.
struct Carver {
    uint8_t* base;
    size_t off;
    explicit Carver(void* p) : base(static_cast<uint8_t*>(p)), off(0) {}
    float* floats(size_t n) {
        float* r = reinterpret_cast<float*>(base + off);
        off += (n * sizeof(float) + 1023) & ~size_t(1023);
        return r;
    }
};

// CTA-pair (cta_group::2) kernels are the default; SC_B200_CTA_PAIRS=0 selects the single-CTA core.
inline bool use_cta_pairs() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SC_B200_CTA_PAIRS");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}
// number of 128-row tile slots the epilogues index for an output with `rows` rows
inline int row_tiles128(long long rows) {
    return use_cta_pairs() ? 2 * static_cast<int>((rows + 255) / 256) : static_cast<int>((rows + 127) / 128);
}

inline int pick_bn(long long n) { return n > 128 ? 256 : (n > 64 ? 128 : 64); }
inline int ceil_div(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

struct Seg {
    Operand a, b;
};

inline Operand kmajor(const float* base, long long mn, long long k, long long g, long long gstride) {
    return Operand{base, mn, k, g, k, gstride, false};
}
inline Operand mnmajor(const float* base, long long mn, long long k, long long g, long long gstride) {
    return Operand{base, mn, k, g, mn, gstride, true};
}
// fp16 operands (base points at __half data)
inline Operand kmajor_h(const void* base, long long mn, long long k, long long g, long long gstride) {
    return Operand{base, mn, k, g, k, gstride, false, 2};
}
inline Operand mnmajor_h(const void* base, long long mn, long long k, long long g, long long gstride) {
    return Operand{base, mn, k, g, mn, gstride, true, 2};
}

// A prepared launch: TMA maps + shape; reusable while the operand addresses and shapes stay the same.
struct GemmPlan {
    GemmMaps maps;
    GemmShape sh;
    int bn;
    int sms;
    bool cta2;
    int eb;  // operand element bytes (4: TF32 path, 2: fp16 path - CTA-pair kernel only)
};

inline int plan_gemm(GemmPlan* pl, int M, int N, int G, const Seg* segs, int nseg, bool a_mn, bool b_mn,
                     bool segs_are_problems = false, int bn = 0, int ksplit = 0) {
    memset(pl, 0, sizeof *pl);
    SC_TRY(device_sm_count(&pl->sms));
    if (g_dbg_sm_limit > 0 && g_dbg_sm_limit < pl->sms) pl->sms = g_dbg_sm_limit & ~1;
    pl->bn = bn > 0 ? bn : pick_bn(N);
    if (ksplit > 1 && segs[0].a.elem_bytes != 2) return fail(-3, "split-K needs fp16 operands (CTA-pair kernel)");
    pl->sh.ksplit = ksplit > 1 ? ksplit : 1;
    pl->eb = segs[0].a.elem_bytes == 2 ? 2 : 4;
    pl->cta2 = use_cta_pairs() || pl->eb == 2;
    GemmShape& sh = pl->sh;
    sh.M = M; sh.N = N; sh.G = G; sh.nseg = nseg;
    sh.nprob = segs_are_problems ? nseg : 1;
    for (int s = 0; s < nseg; ++s) {
        if (segs[s].a.mn_major != a_mn || segs[s].b.mn_major != b_mn) return fail(-3, "operand major mismatch");
        if (segs[s].a.k != segs[s].b.k || segs[s].a.k <= 0) return fail(-3, "segment K mismatch");
        if ((segs[s].a.elem_bytes == 2) != (pl->eb == 2) || (segs[s].b.elem_bytes == 2) != (pl->eb == 2))
            return fail(-3, "operand element size mismatch");
        const char* err = nullptr;
        bool a1 = false, b1 = false;
        int rc = make_operand_map(&pl->maps.a[s], segs[s].a, kBlockM, &a1, &err);
        if (rc == 0) rc = make_operand_map(&pl->maps.b[s], segs[s].b, pl->cta2 ? pl->bn / 2 : pl->bn, &b1, &err);
        sh.a_one_box[s] = a1;
        sh.b_one_box[s] = b1;
        if (rc != 0) return fail(rc, "%s", err ? err : "tensor map creation failed");
        sh.K[s] = static_cast<int>(segs[s].a.k);
        sh.a_gmul[s] = segs[s].a.g > 1 ? 1 : 0;
        sh.b_gmul[s] = segs[s].b.g > 1 ? 1 : 0;
    }
    return 0;
}

template <bool A_MN, bool B_MN, class Epi, bool F16 = false>
int launch_plan(const GemmPlan& pl, const typename Epi::Params& ep, cudaStream_t stream) {
    cudaError_t e;
    if constexpr (F16) {  // fp16 epilogues exist for the CTA-pair kernel only
        if (pl.eb != 2 || pl.bn < 128) return fail(-3, "fp16 operand path needs bn >= 128");
        if (pl.bn == 256) e = launch_gemm_2cta<256, A_MN, B_MN, Epi, 2>(pl.maps, pl.sh, ep, pl.sms, stream);
        else e = launch_gemm_2cta<128, A_MN, B_MN, Epi, 2>(pl.maps, pl.sh, ep, pl.sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "gemm (fp16, CTA pair) launch");
        return 0;
    } else {
        if (pl.eb != 4) return fail(-3, "fp16 operands given to a TF32 launch");
        if (pl.cta2) {
            if (pl.bn == 256) e = launch_gemm_2cta<256, A_MN, B_MN, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
            else if (pl.bn == 128) e = launch_gemm_2cta<128, A_MN, B_MN, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
            else e = launch_gemm_2cta<64, A_MN, B_MN, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
            if (e != cudaSuccess) return cuda_fail(e, "gemm (CTA pair) launch");
            return 0;
        }
        if (pl.bn == 256) e = launch_gemm<256, A_MN, B_MN, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
        else if (pl.bn == 128) e = launch_gemm<128, A_MN, B_MN, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
        else e = launch_gemm<64, A_MN, B_MN, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
        if (e != cudaSuccess) return cuda_fail(e, "gemm launch");
        return 0;
    }
}

// ACC[g] = sum_seg A_seg[g] B_seg[g]^T over G groups, epilogue Epi.  Operand majors are compile-time.
template <bool A_MN, bool B_MN, class Epi, bool F16 = false>
int run_gemm(int M, int N, int G, const Seg* segs, int nseg, const typename Epi::Params& ep, cudaStream_t stream,
             bool segs_are_problems = false, int bn = 0, int ksplit = 0) {
    if (M <= 0 || N <= 0 || G <= 0) return 0;
    GemmPlan pl;
    SC_TRY(plan_gemm(&pl, M, N, G, segs, nseg, A_MN, B_MN, segs_are_problems, bn, ksplit));
    return launch_plan<A_MN, B_MN, Epi, F16>(pl, ep, stream);
}

// SC_B200_RESA: bit mask of the GEMMs that may use the resident-A kernel (1: encode, 2: dL/dz; 0: never).  Default 2:
// same-box runs at the headline shape give 0.482 ms per step with neither, 0.485 with the encode GEMM only, 0.478 with
// the dL/dz GEMM only, 0.479 with both.
inline bool use_resident_a(int which) {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SC_B200_RESA");
        v = e ? atoi(e) : 2;
    }
    return (v & which) != 0;
}

// fp16 K-major GEMM with a short K loop: the resident-A kernel (gemm_core2_res.cuh) when it applies, else run_gemm
template <class Epi>
int run_gemm_short_k(int M, int N, int G, const Seg* seg, const typename Epi::Params& ep, cudaStream_t stream, int which) {
    if (M <= 0 || N <= 0 || G <= 0) return 0;
    // measured with the dL/dz GEMM (ms per step, resident / streaming): 0.478 / 0.482 with 8 column tiles per row tile,
    // 0.827 / 0.824 with 16, 1.527 / 1.500 with 32 - beyond ~8 the contiguous tile ranges cost more than the saved
    // operand traffic gives back
    const bool ok = use_resident_a(which) && seg->a.elem_bytes == 2 && !seg->a.mn_major && !seg->b.mn_major &&
                    seg->a.k <= 64 * kResMaxKBlocks && pick_bn(N) >= 128 && ceil_div(N, pick_bn(N)) <= 8;
    const int kblocks = ceil_div(seg->a.k, 64);
    const bool fits = pick_bn(N) == 256 ? GemmCfgRes<256, Epi>::stages(kblocks) >= 2 : GemmCfgRes<128, Epi>::stages(kblocks) >= 2;
    if (!ok || !fits) return run_gemm<false, false, Epi, true>(M, N, G, seg, 1, ep, stream);
    GemmPlan pl;
    SC_TRY(plan_gemm(&pl, M, N, G, seg, 1, false, false));
    cudaError_t e;
    if (pl.bn == 256) e = launch_gemm_resa<256, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
    else e = launch_gemm_resa<128, Epi>(pl.maps, pl.sh, ep, pl.sms, stream);
    if (e != cudaSuccess) return cuda_fail(e, "gemm (resident A) launch");
    return 0;
}

}  // namespace sc
