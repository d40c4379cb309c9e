// sc_gather_rows: batch assembly from an activation chunk that stays resident in HBM.
//
// Replaces the caller-side `dataset[batch_idxs].to(device)` of the reference train loops
// (big_sweep.py:170, basic_l1_sweep.py:103, cluster_runs.py:37): instead of a host-side random gather of
// fp32 rows followed by a pageable H2D copy every step, the chunk is uploaded once in its storage
// format (fp16 as written by activation_dataset.py:393-397, or fp32) and each step gathers its rows on
// the device, converting to fp32 and subtracting the chunk means (big_sweep.py:402-407) on the way.
//
// HBM-bound byte work: elem_bytes*B*d read + 4*B*d written, one warp per row, 16-byte accesses.
#include "../../include/sc_b200.h"
#include "host_common.cuh"

#include <cuda_fp16.h>

namespace sc {

template <bool HALF>
__global__ void __launch_bounds__(256) gather_rows_kernel(const void* __restrict__ chunk, const int64_t* __restrict__ idx,
                                                          const float* __restrict__ mean, float* __restrict__ out,
                                                          long long n_rows, int d, int B, int* __restrict__ err_flag) {
    pdl_launch_dependents();
    pdl_wait();
    const int lane = threadIdx.x & 31;
    const int warps = (blockDim.x >> 5) * gridDim.x;
    for (int b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < B; b += warps) {
        long long r = idx[b];
        if (r < 0) r += n_rows;  // python-style negative indices, as torch indexing accepts
        const bool ok = r >= 0 && r < n_rows;
        if (!ok && lane == 0 && err_flag) atomicOr(err_flag, 1);
        float* dst = out + (long long)b * d;
        if (HALF) {
            const __half* src = static_cast<const __half*>(chunk) + r * d;
            for (int j = lane * 8; j < d; j += 256) {  // d % 8 == 0 checked by the host
                float v[8];
                if (ok) {
                    const uint4 q = *reinterpret_cast<const uint4*>(src + j);
                    const __half2* h = reinterpret_cast<const __half2*>(&q);
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float2 f = __half22float2(h[t]);
                        v[2 * t] = f.x;
                        v[2 * t + 1] = f.y;
                    }
                } else {
#pragma unroll
                    for (int t = 0; t < 8; ++t) v[t] = 0.f;
                }
                if (mean && ok) {
                    const float4 m0 = *reinterpret_cast<const float4*>(mean + j);
                    const float4 m1 = *reinterpret_cast<const float4*>(mean + j + 4);
                    v[0] -= m0.x; v[1] -= m0.y; v[2] -= m0.z; v[3] -= m0.w;
                    v[4] -= m1.x; v[5] -= m1.y; v[6] -= m1.z; v[7] -= m1.w;
                }
                *reinterpret_cast<float4*>(dst + j) = make_float4(v[0], v[1], v[2], v[3]);
                *reinterpret_cast<float4*>(dst + j + 4) = make_float4(v[4], v[5], v[6], v[7]);
            }
        } else {
            const float* src = static_cast<const float*>(chunk) + r * d;
            for (int j = lane * 4; j < d; j += 128) {  // d % 4 == 0 checked by the host
                float4 v = ok ? *reinterpret_cast<const float4*>(src + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                if (mean && ok) {
                    const float4 m = *reinterpret_cast<const float4*>(mean + j);
                    v.x -= m.x; v.y -= m.y; v.z -= m.z; v.w -= m.w;
                }
                *reinterpret_cast<float4*>(dst + j) = v;
            }
        }
    }
}

// Column means of a resident chunk (big_sweep.py:402-405, `chunk.mean(dim=0)`): two deterministic passes,
// kMeanSlabs row slabs -> partial sums [slab][d], then one thread per column adds the slabs in order.
constexpr int kMeanSlabs = 1024;

template <bool HALF>
__global__ void __launch_bounds__(256) column_partial_kernel(const void* __restrict__ chunk, float* __restrict__ part,
                                                             long long n_rows, int d) {
    pdl_launch_dependents();
    pdl_wait();
    const long long per = (n_rows + kMeanSlabs - 1) / kMeanSlabs;
    const long long r0 = blockIdx.x * per, r1 = (r0 + per < n_rows) ? r0 + per : n_rows;
    for (int j = threadIdx.x; j < d; j += blockDim.x) {
        float acc = 0.f;
        for (long long r = r0; r < r1; ++r)
            acc += HALF ? __half2float(static_cast<const __half*>(chunk)[r * d + j]) : static_cast<const float*>(chunk)[r * d + j];
        part[(long long)blockIdx.x * d + j] = acc;
    }
}

__global__ void column_mean_finish_kernel(const float* __restrict__ part, float* __restrict__ out, long long n_rows, int d) {
    pdl_launch_dependents();
    pdl_wait();
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d) return;
    float acc = 0.f;
    for (int s = 0; s < kMeanSlabs; ++s) acc += part[(long long)s * d + j];
    out[j] = acc / static_cast<float>(n_rows);
}

static int column_mean_impl(const void* chunk, int elem_bytes, long long n_rows, int d, float* out, void* workspace,
                            cudaStream_t st) {
    if (n_rows <= 0 || d <= 0) return fail(-1, "sc_column_mean: bad sizes (n_rows=%lld d=%d)", n_rows, d);
    if (elem_bytes != 2 && elem_bytes != 4) return fail(-1, "sc_column_mean: elem_bytes must be 2 (fp16) or 4 (fp32)");
    if (!chunk || !out || !workspace) return fail(-1, "sc_column_mean: chunk / out / workspace is NULL");
    float* part = static_cast<float*>(workspace);
    if (elem_bytes == 2) column_partial_kernel<true><<<kMeanSlabs, 256, 0, st>>>(chunk, part, n_rows, d);
    else column_partial_kernel<false><<<kMeanSlabs, 256, 0, st>>>(chunk, part, n_rows, d);
    column_mean_finish_kernel<<<(d + 127) / 128, 128, 0, st>>>(part, out, n_rows, d);
    SC_CUDA(cudaGetLastError(), "column_mean launch");
    return 0;
}

static int gather_impl(const ScGatherArgs& a) {
    if (a.B == 0) return 0;
    if (a.B < 0 || a.d <= 0 || a.n_rows <= 0) return fail(-1, "sc_gather_rows: bad sizes (B=%d d=%d n_rows=%lld)", a.B, a.d, a.n_rows);
    if (a.elem_bytes != 2 && a.elem_bytes != 4) return fail(-1, "sc_gather_rows: elem_bytes must be 2 (fp16) or 4 (fp32)");
    if (!a.chunk || !a.idx || !a.out) return fail(-1, "sc_gather_rows: chunk / idx / out is NULL");
    const int vec = a.elem_bytes == 2 ? 8 : 4;
    if (a.d % vec) return fail(-1, "sc_gather_rows: d must be a multiple of %d for %d-byte elements (got %d)", vec, a.elem_bytes, a.d);
    if ((reinterpret_cast<uintptr_t>(a.chunk) | reinterpret_cast<uintptr_t>(a.out) | reinterpret_cast<uintptr_t>(a.mean)) & 15)
        return fail(-1, "sc_gather_rows: chunk, out and mean must be 16-byte aligned");
    int sms = 0;
    SC_TRY(device_sm_count(&sms));
    cudaStream_t st = static_cast<cudaStream_t>(a.stream);
    const int rows_per_block = 8;
    int blocks = (a.B + rows_per_block - 1) / rows_per_block;
    if (blocks > 8 * sms) blocks = 8 * sms;
    if (a.elem_bytes == 2)
        gather_rows_kernel<true><<<blocks, 256, 0, st>>>(a.chunk, a.idx, a.mean, a.out, a.n_rows, a.d, a.B, a.err_flag);
    else
        gather_rows_kernel<false><<<blocks, 256, 0, st>>>(a.chunk, a.idx, a.mean, a.out, a.n_rows, a.d, a.B, a.err_flag);
    SC_CUDA(cudaGetLastError(), "gather_rows launch");
    return 0;
}

}  // namespace sc

extern "C" int sc_gather_rows(const ScGatherArgs* args) {
    if (!args) return sc::fail(-1, "sc_gather_rows: NULL args");
    return sc::gather_impl(*args);
}

extern "C" size_t sc_column_mean_workspace_bytes(int32_t d) {
    return d > 0 ? sizeof(float) * static_cast<size_t>(sc::kMeanSlabs) * static_cast<size_t>(d) : 0;
}

extern "C" int sc_column_mean(const void* chunk, int32_t elem_bytes, int64_t n_rows, int32_t d, float* out_mean,
                              void* workspace, void* stream) {
    return sc::column_mean_impl(chunk, elem_bytes, n_rows, d, out_mean, workspace, static_cast<cudaStream_t>(stream));
}
