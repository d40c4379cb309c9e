// Host-side TMA descriptor ("tensor map") construction.  The driver entry point is
// resolved through the runtime (cudaGetDriverEntryPoint) so the library does not link libcuda
// and still loads on a machine without a driver.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace sc {

// Operand for the GEMM core, fp32 storage, logical shape [G][MN][K].
//   K-major : element (g, mn, k) at base[g*gstride + mn*ld + k]    (k contiguous)
//   MN-major: element (g, mn, k) at base[g*gstride + k*ld + mn]    (mn contiguous)
// ld and gstride in elements, multiples of 4 (TMA strides are multiples of 16 bytes);
// base 16-byte aligned.  G == 1 and gstride == 0 describe an operand shared by every group.
struct Operand {
    const void* base;
    int64_t mn, k, g;
    int64_t ld, gstride;
    bool mn_major;
    int elem_bytes = 4;  // 4: fp32 storage (TF32 MMA), 2: fp16
};

// box_mn: rows of the operand tile (128 for the A side, BLOCK_N for the B side); BLOCK_K = 32.
// K-major : one box {32 k, box_mn rows}, 128-byte swizzle.
// MN-major: 128-byte swizzle with 32-byte atoms (the only layout tcgen05 accepts for MN-major 32-bit
//           operands).  When mn is a multiple of 32 the tensor is described in 4-D as
//           {32, k, mn/32, g} and ONE box {32, 32 k, box_mn/32} fills the whole tile (*one_box = true);
//           otherwise boxes {32 mn, 32 k} and the kernel issues box_mn/32 loads per stage.
// Returns 0 on success, else a negative sc error code with the message in *err.
int make_operand_map(CUtensorMap* out, const Operand& op, int box_mn, bool* one_box, const char** err);

// Plain row-major fp32 [G][rows][cols] map for TMA stores / epilogue tile loads, no swizzle.
int make_rowmajor_map(CUtensorMap* out, const float* base, int64_t rows, int64_t cols, int64_t g, int64_t ld,
                      int64_t gstride, int box_rows, int box_cols, const char** err);

// Row-major fp32 [G][rows][cols] map for the epilogue staging tiles: box {32 cols, 32 rows}, 128-byte swizzle.
int make_io_map(CUtensorMap* out, const float* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                const char** err);
// Same for fp16 tensors: box {32 cols, 32 rows} = 64-byte rows, 64-byte swizzle.  ld: row stride in elements
// (multiple of 8; 0 = cols).
int make_io_map_h(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                  const char** err, int64_t ld = 0);

// fp16 [G][rows][cols] map with box {64 cols, 32 rows} = 128-byte rows, 128-byte swizzle (two 32-column chunks per store)
int make_io_map_h64(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                    const char** err, int64_t ld = 0);

// fp16 [G][rows][cols] map with box {32 cols, 64 rows} = 64-byte rows, 64-byte swizzle: the TRANSPOSED copy of a
// [32 x 64] register tile (rows of the map = columns of the accumulator)
int make_io_map_ht(CUtensorMap* out, const void* base, int64_t rows, int64_t cols, int64_t g, int64_t gstride,
                   const char** err);

}  // namespace sc
