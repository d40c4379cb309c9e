// Coalesced global I/O for epilogue warps.  A thread of the epilogue owns one accumulator ROW, so
// touching global memory directly costs one 128-byte line per lane per instruction (32 L1 wavefronts
// for 512 bytes).  Instead every epilogue warp owns a few 4 KB staging tiles in shared memory -
// [32 rows][32 fp32], 128-byte swizzled exactly as TMA's SWIZZLE_128B lays them out - and moves whole
// tiles with TMA: tile loads complete on the warp's own mbarrier, tile stores are bulk-async groups.
// Row reads / writes of the swizzled tile are bank-conflict free (a quarter warp covers all 32 banks).
#pragma once
#include <cuda_fp16.h>

#include "sm100_ptx.cuh"

namespace sc {

constexpr int kIoTileBytes = 32 * 32 * 4;
constexpr int kIoBars = 4;  // tile-load mbarriers per epilogue warp

struct EpiIo {
    uint8_t* buf;     // this warp's staging tiles (Epi::kIoBufs x 4 KB, 1024-byte aligned)
    uint64_t* bar;    // this warp's kIoBars mbarriers for tile loads (one per load slot)
    uint32_t* phase;  // bit s (< kIoBars) = parity of the next completion on bar[s]; bits >= 8 are free for the functor
                      // (one word per warp, lives across tiles)
};

// element (r, c) of a staging tile sits at r*128 + ((c/4) ^ (r & 7))*16 + (c % 4)*4
SC_DEVICE void io_read_row(const uint8_t* tile, int r, float (&v)[32]) {
    const uint32_t row = smem_u32(tile) + r * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 t = lds_128(row + ((j ^ (r & 7)) << 4));
        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
    }
}
SC_DEVICE void io_write_row(uint8_t* tile, int r, const float (&v)[32]) {
    const uint32_t row = smem_u32(tile) + r * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j) sts_128(row + ((j ^ (r & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}

// W consecutive columns (W = 32: the whole row, 16: columns [16 h, 16 h + 16)) of row r
template <int W>
SC_DEVICE void io_read_cols(const uint8_t* tile, int r, int h, float (&v)[W]) {
    const uint32_t row = smem_u32(tile) + r * 128;
#pragma unroll
    for (int j = 0; j < W / 4; ++j) {
        const float4 t = lds_128(row + (((j + h * (W / 4)) ^ (r & 7)) << 4));
        v[4 * j] = t.x; v[4 * j + 1] = t.y; v[4 * j + 2] = t.z; v[4 * j + 3] = t.w;
    }
}
template <int W>
SC_DEVICE void io_write_cols(uint8_t* tile, int r, int h, const float (&v)[W]) {
    const uint32_t row = smem_u32(tile) + r * 128;
#pragma unroll
    for (int j = 0; j < W / 4; ++j)
        sts_128(row + (((j + h * (W / 4)) ^ (r & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
}

// Whole warp.  Make the rows just written visible to the async proxy, then lane 0 stores the tile.
// The caller commits the group (io_commit) once all tiles of the step are issued.
SC_DEVICE void io_store_tile(const CUtensorMap* map, const uint8_t* tile, int col, int row, int g) {
    fence_proxy_async_smem();
    __syncwarp();
    if ((threadIdx.x & 31) == 0) tma_store_3d(map, tile, col, row, g);
}
SC_DEVICE void io_commit() {
    if ((threadIdx.x & 31) == 0) tma_store_commit();
}
// Whole warp.  The staging tiles may be overwritten once earlier stores have READ them.
SC_DEVICE void io_wait_stores_read() {
    if ((threadIdx.x & 31) == 0) tma_store_wait_read<0>();
    __syncwarp();
}
// Same, but the most recent N store groups may still be reading (staging tiles used in rotation).
template <int N>
SC_DEVICE void io_wait_stores_read_n() {
    if ((threadIdx.x & 31) == 0) tma_store_wait_read<N>();
    __syncwarp();
}
// Whole warp, before kernel exit: the stores themselves must be complete.
SC_DEVICE void io_wait_stores_done() {
    if ((threadIdx.x & 31) == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    __syncwarp();
}
// Lane 0 arms the warp's barrier for `bytes` and the caller (lane 0) issues the tile loads.
SC_DEVICE void io_expect(const EpiIo& io, uint32_t bytes, int slot = 0) { mbar_arrive_expect_tx(io.bar + slot, bytes); }
SC_DEVICE void io_wait_loads(const EpiIo& io, int slot = 0) {
    mbar_wait(io.bar + slot, (*io.phase >> slot) & 1u);
    *io.phase ^= 1u << slot;
}
// element (r, c) of a staging tile (see io_read_row)
SC_DEVICE uint32_t io_elem_offset(int r, int c) { return r * 128 + (((c >> 2) ^ (r & 7)) << 4) + ((c & 3) << 2); }

// fp16 staging tile [32 rows][32 halves] = 64-byte rows, laid out as TMA's SWIZZLE_64B expects:
// 16-byte chunk j of row r sits at r*64 + ((j ^ ((r >> 1) & 3)) << 4).  Conflict free for row writes.
SC_DEVICE void io_write_row_h(uint8_t* tile, int r, const float (&v)[32]) {
    const uint32_t row = smem_u32(tile) + r * 64;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        uint32_t h[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const __half2 p = half2_sat(v[8 * j + 2 * t], v[8 * j + 2 * t + 1]);
            h[t] = *reinterpret_cast<const uint32_t*>(&p);
        }
        sts_128u(row + ((j ^ ((r >> 1) & 3)) << 4), h[0], h[1], h[2], h[3]);
    }
}

}  // namespace sc
