// sc_encode_noalloc: dictionary encode for callers that must not allocate device memory.
//
// Who that is: a host program that scores activations against a trained dictionary from inside a
// latency-critical loop (an online feature monitor next to a serving model, a plain C/C++ harness
// without torch's caching allocator) and has either handed all device memory to its own pool or
// cannot afford cudaMalloc's implicit device-wide synchronisation in that loop.  sc_encode needs a
// caller-provided workspace for the TF32-rounded copies of x and W; this variant keeps that scratch
// in a fixed array inside the library image instead, so the call touches no allocator at all.
//
// Cost and limits: every GPU that loads libsc_b200.so carries the array (SC_NOALLOC_SCRATCH_FLOATS
// floats = 16 MiB) until the process exits.  One dictionary per call; B*d + N*d (each rounded up to
// 256 floats) must fit, larger problems are refused with an error - use sc_encode for those.
// All calls share the one array: its previous contents are simply overwritten (nothing in it has
// to start at zero - both regions are fully written by the rounding pass before the GEMM reads
// them), and two calls that overlap on the GPU (different streams, or different host threads)
// race on it.  Serialise them.
#include "../../include/sc_b200.h"
#include "host_common.cuh"
#include "sae_epilogues.cuh"

#include <mutex>

namespace sc {

__global__ void round_copy_kernel(const float4* __restrict__ in, float4* __restrict__ out, long long n4);

constexpr long long kNoallocFloats = SC_NOALLOC_SCRATCH_FLOATS;

// The scratch itself.  In host code this name is only a handle for the runtime - NOT a device
// address; kernels get the real address obtained once per device by cudaGetSymbolAddress below.
static __device__ __align__(256) float g_noalloc_scratch[kNoallocFloats];

static std::mutex g_noalloc_mu;
static float* g_noalloc_ptr[64] = {};

// First call on a device resolves the array's device address (this may load the module, so make
// that first call outside any stream capture); later calls reuse the cached pointer.
static int noalloc_scratch(float** out) {
    int dev = 0;
    SC_CUDA(cudaGetDevice(&dev), "cudaGetDevice");
    if (dev < 0 || dev >= 64) return fail(-1, "device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        float* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        if (!p) return fail(-1, "cudaGetSymbolAddress returned a null scratch address");
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return 0;
}

static long long pad256(long long n) { return (n + 255) & ~255LL; }

static int encode_noalloc_impl(const float* x, const float* W, const float* bias, float* out_c, int B, int d, int N,
                               cudaStream_t st) {
    if (B <= 0 || d <= 0 || N <= 0) return fail(-1, "B, d, N must be positive (got %d %d %d)", B, d, N);
    if ((d & 3) || (N & 3)) return fail(-1, "d and N must be multiples of 4 (got d=%d N=%d)", d, N);
    if (!x || !W || !out_c) return fail(-1, "x / W / out_c is NULL");
    if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(W) | reinterpret_cast<uintptr_t>(out_c)) & 15)
        return fail(-1, "x, W and out_c must be 16-byte aligned");
    const long long nx = pad256((long long)B * d), nw = pad256((long long)N * d);
    if (nx + nw > kNoallocFloats)
        return fail(-4, "sc_encode_noalloc: B*d + N*d = %lld floats exceeds the built-in scratch of %lld floats "
                        "(SC_NOALLOC_SCRATCH_FLOATS); use sc_encode with a caller workspace",
                    nx + nw, kNoallocFloats);
    float* scratch = nullptr;
    SC_TRY(noalloc_scratch(&scratch));
    float* x_r = scratch;
    float* W_r = scratch + nx;
    const long long x4 = (long long)B * d / 4, w4 = (long long)N * d / 4;
    round_copy_kernel<<<static_cast<int>(x4 / 256 + 1 < 4096 ? x4 / 256 + 1 : 4096), 256, 0, st>>>(
        reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(x_r), x4);
    round_copy_kernel<<<static_cast<int>(w4 / 256 + 1 < 4096 ? w4 / 256 + 1 : 4096), 256, 0, st>>>(
        reinterpret_cast<const float4*>(W), reinterpret_cast<float4*>(W_r), w4);
    SC_CUDA(cudaGetLastError(), "round_copy launch");
    Seg s{kmajor(x_r, B, d, 1, 0), kmajor(W_r, N, d, 1, 0)};
    EpiEncode::Params ep{};
    {
        const char* err = nullptr;
        const int rc = make_io_map(&ep.c_map, out_c, B, N, 1, (long long)B * N, &err);
        if (rc != 0) return fail(rc, "%s", err ? err : "io tensor map creation failed");
    }
    ep.bias = bias; ep.l1_part = nullptr; ep.signed_zero = 0; ep.dict_size = nullptr;
    return run_gemm<false, false, EpiEncode>(B, N, 1, &s, 1, ep, st);
}

}  // namespace sc

extern "C" int sc_encode_noalloc(const float* x, const float* W, const float* bias, float* out_c, int32_t B, int32_t d,
                                 int32_t N, void* stream) {
    return sc::encode_noalloc_impl(x, W, bias, out_c, B, d, N, static_cast<cudaStream_t>(stream));
}
