#include "tmap.h"

#include <cudaTypedefs.h>
#include <stdio.h>

#include <mutex>

namespace sc {

static PFN_cuTensorMapEncodeTiled get_encode(const char** err) {
    static PFN_cuTensorMapEncodeTiled fn = nullptr;
    static std::once_flag once;
    static const char* init_err = nullptr;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr) {
            init_err = "cuTensorMapEncodeTiled not available from the driver";
            (void)cudaGetLastError();
        } else {
            fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled>(p);
        }
    });
    if (!fn && err) *err = init_err;
    return fn;
}

static thread_local char g_msg[256];

static int encode(CUtensorMap* out, const void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides_bytes,
                  const cuuint32_t* box, CUtensorMapSwizzle swz, const char** err,
                  CUtensorMapDataType dtype = CU_TENSOR_MAP_DATA_TYPE_FLOAT32) {
    PFN_cuTensorMapEncodeTiled fn = get_encode(err);
    if (!fn) return -10;
    if (reinterpret_cast<uintptr_t>(base) & 15) {
        if (err) *err = "tensor base address is not 16-byte aligned";
        return -11;
    }
    for (int i = 0; i < (rank - 1 < 2 ? rank - 1 : 2); ++i) {
        if (strides_bytes[i] % 16 != 0 || strides_bytes[i] == 0) {
            snprintf(g_msg, sizeof g_msg, "tensor stride %d = %llu bytes is not a positive multiple of 16", i,
                     (unsigned long long)strides_bytes[i]);
            if (err) *err = g_msg;
            return -12;
        }
    }
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    if (rank == 4 && (strides_bytes[2] % 16 != 0 || strides_bytes[2] == 0)) {
        if (err) *err = "tensor group stride is not a positive multiple of 16 bytes";
        return -12;
    }
    CUresult r = fn(out, dtype, rank, const_cast<void*>(base), dims, strides_bytes, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        snprintf(g_msg, sizeof g_msg,
                 "cuTensorMapEncodeTiled failed (%d): dims {%llu,%llu,%llu} strides {%llu,%llu} box {%u,%u,%u} swizzle %d",
                 (int)r, (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
                 (unsigned long long)strides_bytes[0], (unsigned long long)strides_bytes[1], box[0], box[1], box[2], (int)swz);
        if (err) *err = g_msg;
        return -13;
    }
    return 0;
}

int make_operand_map(CUtensorMap* out, const Operand& op, int box_mn, bool* one_box, const char** err) {
    if (one_box) *one_box = false;
    const int eb = op.elem_bytes == 2 ? 2 : 4;
    const CUtensorMapDataType dt = eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    const cuuint32_t kelems = 128 / eb;  // one 128-byte swizzle span of K (or of M/N when MN-major)
    const int64_t g = op.g < 1 ? 1 : op.g;
    // a never-used outer stride still has to be a legal one
    const int64_t gs = (g > 1 && op.gstride > 0) ? op.gstride : op.ld * (op.mn_major ? op.k : op.mn);
    cuuint64_t strides[2] = {(cuuint64_t)op.ld * eb, (cuuint64_t)gs * eb};
    if (!op.mn_major) {
        cuuint64_t dims[3] = {(cuuint64_t)op.k, (cuuint64_t)op.mn, (cuuint64_t)g};
        cuuint32_t box[3] = {kelems, (cuuint32_t)box_mn, 1};
        return encode(out, op.base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, err, dt);
    }
    // MN-major: fp32 needs the 32-byte-atom flavour of the 128-byte swizzle, fp16 the plain one
    const CUtensorMapSwizzle swz = eb == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B;
    const cuuint32_t chunk = kelems;  // m/n elements per 128-byte chunk: 32 fp32 or 64 fp16
    if (one_box && op.mn % chunk == 0 && box_mn % (int)chunk == 0) {
        cuuint64_t dims4[4] = {chunk, (cuuint64_t)op.k, (cuuint64_t)(op.mn / chunk), (cuuint64_t)g};
        cuuint64_t strides4[3] = {(cuuint64_t)op.ld * eb, 128, (cuuint64_t)gs * eb};
        cuuint32_t box4[4] = {chunk, kelems, (cuuint32_t)(box_mn / chunk), 1};
        *one_box = true;
        return encode(out, op.base, 4, dims4, strides4, box4, swz, err, dt);
    }
    cuuint64_t dims[3] = {(cuuint64_t)op.mn, (cuuint64_t)op.k, (cuuint64_t)g};
    cuuint32_t box[3] = {chunk, kelems, 1};
    return encode(out, op.base, 3, dims, strides, box, swz, err, dt);
}

int make_rowmajor_map(CUtensorMap* out, const float* base, int64_t rows, int64_t cols, int64_t g, int64_t ld,
                      int64_t gstride, int box_rows, int box_cols, const char** err) {
    if (g < 1) g = 1;
    const int64_t gs = (g > 1 && gstride > 0) ? gstride : ld * rows;
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)g};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 4, (cuuint64_t)gs * 4};
    cuuint32_t box[3] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1};
    return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE, err);
}

int make_io_map(CUtensorMap* out, const float* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                const char** err) {
    if (g < 1) g = 1;
    const int64_t gs = (g > 1 && gstride > 0) ? gstride : cols * rows;
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)g};
    cuuint64_t strides[2] = {(cuuint64_t)cols * 4, (cuuint64_t)gs * 4};
    cuuint32_t box[3] = {32, 32, 1};
    return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, err);
}

int make_io_map_h(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                  const char** err, int64_t ld) {
    if (g < 1) g = 1;
    if (ld <= 0) ld = cols;
    const int64_t gs = (g > 1 && gstride > 0) ? gstride : ld * rows;
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)g};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)gs * 2};
    cuuint32_t box[3] = {32, 32, 1};
    return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B, err, CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
}

int make_io_map_h64(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                    const char** err, int64_t ld) {
    if (g < 1) g = 1;
    if (ld <= 0) ld = cols;
    const int64_t gs = (g > 1 && gstride > 0) ? gstride : ld * rows;
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)g};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 2, (cuuint64_t)gs * 2};
    cuuint32_t box[3] = {64, 32, 1};
    return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B, err, CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
}

int make_io_map_ht(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                   const char** err) {
    if (g < 1) g = 1;
    const int64_t gs = (g > 1 && gstride > 0) ? gstride : cols * rows;
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)g};
    cuuint64_t strides[2] = {(cuuint64_t)cols * 2, (cuuint64_t)gs * 2};
    cuuint32_t box[3] = {32, 64, 1};
    return encode(out, base, 3, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_64B, err, CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
}

}  // namespace sc
