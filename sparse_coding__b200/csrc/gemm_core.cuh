// Persistent, warp-specialised tcgen05 GEMM core for sm_100a with pluggable fused epilogues.
//
//   ACC[g][m][n] = sum_seg sum_k A_seg[g][m][k] * B_seg[g][n][k]        (TF32 in, fp32 accumulate)
//
// * one CTA per SM, 256 threads: warp 0 = TMA producer, warp 1 = MMA issuer, warp 2 = TMEM
//   allocator, warps 4..11 = epilogue: two warps per 32-row quarter of the tile (TMEM lanes are
//   only reachable from warps with the matching warp%4), the pair splitting the 32-column chunks
// * operand tiles [128 x 32] / [BN x 32] fp32 arrive by TMA into 128-byte-swizzled shared memory,
//   a ring of kStages stages guarded by full/empty mbarriers
// * one elected thread issues tcgen05.mma.kind::tf32 128 x BN x 8; accumulators live in TMEM,
//   double buffered (2 x BN columns) so the epilogue of tile i overlaps the MMAs of tile i+1
// * up to three (A,B) operand pairs ("segments") accumulate into the same tile: used for the tied
//   SAE gradient (c^T R + dz^T x) and for the 3xTF32 split products of FISTA
// * either operand may be K-major (k contiguous) or MN-major (m/n contiguous); MN-major fp32 tiles
//   use the 128-byte swizzle with 32-byte atoms, the only MN-major layout tcgen05 takes for 32-bit
// * the epilogue is a functor that sees its accumulator row 32 columns at a time:
//     Epi(params, ctx); prefetch(col); chunk(col, v[32], next_col or -1); finish()
//
// Group (member) index g lets one launch cover every SAE of the ensemble; an operand that all
// members share (the activation batch) has gmul = 0 and is fetched from the same address by all.
#pragma once
#include <stdlib.h>
#include "epi_io.cuh"
#include "sm100_ptx.cuh"

namespace sc {

constexpr int kBlockM = 128;
constexpr int kBlockK = 32;  // fp32 elements: 128 bytes, one swizzle span
constexpr int kUmmaK = 8;    // tf32: 32 bytes of K per instruction
constexpr int kMaxSeg = 3;
constexpr int kGemmThreads = 384;
constexpr int kEpiHalves = 2;  // epilogue warps per row quarter
constexpr int kATileBytes = kBlockM * kBlockK * 4;

struct alignas(64) GemmMaps {
    CUtensorMap a[kMaxSeg];
    CUtensorMap b[kMaxSeg];
};

struct GemmShape {
    int M, N, G;
    int nseg;
    int K[kMaxSeg];
    int a_gmul[kMaxSeg], b_gmul[kMaxSeg];
    int a_one_box[kMaxSeg], b_one_box[kMaxSeg];  // MN-major operand described 4-D: one TMA box per tile
    int tiles_m, tiles_n;
    int nprob;  // > 1: segment p is an independent problem p (own operands, own epilogue parameters) sharing
                // the launch; groups are then laid out as [problem][member]
    int res_stages;  // resident-A kernel (gemm_core2_res.cuh): stages of its B ring, set by the launcher
    int ksplit; // > 1 (CTA-pair kernel, one segment): every output tile is computed as ksplit partial tiles over
                // disjoint ranges of the K loop, each by its own CTA pair; the epilogue sees RowCtx::split and must
                // combine the partials itself (reduce-add).  For outputs with fewer tiles than CTA pairs.
};

constexpr int kEpiWarps = 4 * 2;  // 4 row quarters x kEpiHalves

template <int BN, int IO = 0, int H = kEpiHalves, int OWNERS = 4 * H>
struct GemmCfg {
    static constexpr int kBTileBytes = BN * kBlockK * 4;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
    static constexpr int kWarpsEpi = 4 * H;
    static constexpr int kThreads = 128 + 32 * kWarpsEpi;
    static constexpr int kIoBytes = OWNERS * IO * kIoTileBytes;  // epilogue staging tiles (epi_io.cuh)
    static constexpr int kStagesRaw = (227 * 1024 - 2048 - kIoBytes) / kStageBytes;
    static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
    static constexpr int kTmemCols = (2 * BN <= 32) ? 32 : (2 * BN <= 64) ? 64 : (2 * BN <= 128) ? 128 : (2 * BN <= 256) ? 256 : 512;
    static constexpr int kBarrierBytes = (2 * kStages + 4 + kIoBars * kWarpsEpi) * 8 + 16;
    static constexpr int kSmemBytes = kStages * kStageBytes + kIoBytes + kBarrierBytes + 1024;  // + alignment slack
    static_assert(BN % 32 == 0 && BN >= 32 && BN <= 256, "BN must be a multiple of 32 in [32,256]");
    static_assert(kStages >= 2, "not enough shared memory left for the operand pipeline");
};

// Epilogue warps per 32-row quarter of the tile: kEpiHalves unless the functor sets Epi::kHalves (1: four epilogue
// warps, each with twice the staging memory - room to double-buffer; 4: sixteen).
template <class E, class = void>
struct EpiHalvesOf { static constexpr int v = kEpiHalves; };
template <class E>
struct EpiHalvesOf<E, decltype((void)E::kHalves)> { static constexpr int v = E::kHalves; };

// Column split (Epi::kColSplit == 2, with kHalves == 2): the two warps of a row quarter do not alternate chunks but
// SHARE every chunk - warp `half` takes columns [16 half, 16 half + 16) of each 32-column chunk - and share one set of
// staging tiles and load barriers per quarter.  For epilogues whose arithmetic per element, not their memory
// traffic, is what a single warp per scheduler cannot get through (the Adam update).
template <class E, class = void>
struct EpiColSplitOf { static constexpr int v = 1; };
template <class E>
struct EpiColSplitOf<E, decltype((void)E::kColSplit)> { static constexpr int v = E::kColSplit; };
// Columns per chunk() call (Epi::kChunkW: 32 or 64).  64 halves the per-chunk bookkeeping (fences, TMA issue, barrier
// arming) of epilogues that move a staging tile per chunk.
template <class E, class = void>
struct EpiChunkWOf { static constexpr int v = 32; };
template <class E>
struct EpiChunkWOf<E, decltype((void)E::kChunkW)> { static constexpr int v = E::kChunkW; };
// number of owners of staging memory: one per epilogue warp, or one per row quarter when the columns are split
template <class E>
struct EpiIoOwnersOf { static constexpr int v = EpiColSplitOf<E>::v == 2 ? 4 : 4 * EpiHalvesOf<E>::v; };

// Optional work for warp 3, which has no role in the GEMM: Epi::side_job(params, cta, nctas) when the functor has one.
template <class E>
SC_DEVICE auto run_side_job(const typename E::Params& p, int cta, int nctas, int) -> decltype(E::side_job(p, cta, nctas), void()) {
    E::side_job(p, cta, nctas);
}
template <class E>
SC_DEVICE void run_side_job(const typename E::Params&, int, int, long) {}

// Context handed to an epilogue functor for one accumulator row of one tile.
struct RowCtx {
    int g;       // group / ensemble member
    int row;     // global accumulator row (may be >= M on the ragged last tile)
    int row0;    // row of lane 0 = first row of this warp's 32-row slice (a warp-uniform value: TMA coordinates come from it)
    int col0;    // first global column of the tile
    int M, N;    // logical extents
    int tile_m, tile_n;
    int tiles_m, tiles_n;
    int warp_q;  // which 32-row quarter of the tile this warp owns (0..3)
    int prob;    // problem index when the launch batches independent problems (else 0)
    int half;    // which of the `halves` warps of that quarter (takes chunks half, half+halves, ...)
    int halves;  // epilogue warps per quarter in this launch (EpiHalvesOf<Epi>)
    int part;    // warp_q + 4*half: slot for per-warp partial results (0..4*halves-1)
    // the tile this warp processes next (static schedule), for functors that prefetch across tiles:
    int split;      // which K range this partial tile covers (0 when GemmShape::ksplit <= 1)
    int bn;         // tile width in columns
    int first;      // 1 on the warp's first tile (nothing was prefetched for it)
    int next_g;     // group of the next tile, -1 when this is the last one
    int next_prob;  // its problem index
    int next_row0;  // first global row of this warp's 32-row slice in the next tile
    int next_col0;  // first global column of the next tile
};

// One accumulator tile through an epilogue functor (shared by the single-CTA and the CTA-pair kernels).
//   Epi(params, ctx, io); prefetch(col); chunk(col, v[32], next_col or -1); finish()
// `release` signals the MMA warp that the TMEM stage may be overwritten (differs between the variants).
template <int BN, class Epi, class Release>
SC_DEVICE void epilogue_tile(const typename Epi::Params& ep, const RowCtx& ctx, const EpiIo& io, uint64_t* full_bar,
                             uint32_t parity, uint32_t taddr, Release release) {
    Epi epi(ep, ctx, io);
    constexpr int kH = EpiHalvesOf<Epi>::v;
    // The accumulator stage goes back to the MMA warp as soon as this warp has READ its last chunk of the tile (the values
    // are in registers then), not after it has processed it: the stage is held for one chunk's arithmetic and stores
    // less per tile, which matters where a warp has few chunks per tile (two in the dL/dz epilogue).
    bool released = false;
    auto give_back = [&] {
        tc_fence_before();
        release();
        released = true;
    };
    if constexpr (EpiColSplitOf<Epi>::v == 2) {
        static_assert(kH == 2, "column split pairs the two warps of a row quarter");
        epi.prefetch(0);
        mbar_wait(full_bar, parity);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            float v[16];
            tmem_ld16(taddr + c * 32 + ctx.half * 16, v);
            tmem_ld_wait();
            if (c + 1 == BN / 32) give_back();
            epi.chunk(c * 32, v, (c + 1 < BN / 32) ? (c + 1) * 32 : -1);
        }
    } else if constexpr (EpiChunkWOf<Epi>::v == 64) {
        constexpr int kUnits = BN / 64, kChunks = (kUnits + kH - 1) / kH;
        static_assert(BN % 64 == 0, "64-column chunks need a tile width that is a multiple of 64");
        mbar_wait(full_bar, parity);
        tc_fence_after();
#pragma unroll 1
        for (int i = 0; i < kChunks; ++i) {
            const int c = ctx.half + i * kH;
            if (c < kUnits) {
                float v[64];
                tmem_ld32(taddr + c * 64, *reinterpret_cast<float(*)[32]>(&v[0]));
                tmem_ld32(taddr + c * 64 + 32, *reinterpret_cast<float(*)[32]>(&v[32]));
                tmem_ld_wait();
                if (c + kH >= kUnits) give_back();
                epi.chunk(c * 64, v, (c + kH < kUnits) ? (c + kH) * 64 : -1);
            }
        }
    } else {
        constexpr int kChunks = (BN / 32 + kH - 1) / kH;
        // global loads the functor needs for its first chunk do not depend on the accumulator: issue them
        // before waiting for the MMAs, and the next chunk's while the current one is processed
        if (ctx.half < BN / 32) epi.prefetch(ctx.half * 32);
        mbar_wait(full_bar, parity);
        tc_fence_after();
#pragma unroll 1  // keep the (large) functor body in the instruction cache
        for (int i = 0; i < kChunks; ++i) {
            const int c = ctx.half + i * kH;
            if (c < BN / 32) {
                float v[32];
                tmem_ld32(taddr + c * 32, v);
                tmem_ld_wait();
                if (c + kH >= BN / 32) give_back();
                epi.chunk(c * 32, v, (c + kH < BN / 32) ? (c + kH) * 32 : -1);
            }
        }
    }
    if (!released) give_back();  // a warp without a chunk in this tile
    epi.finish();
}

template <int BN, bool A_MN, bool B_MN, class Epi>
__global__ void __launch_bounds__(128 + 128 * EpiHalvesOf<Epi>::v, 1)
gemm_tf32_kernel(const __grid_constant__ GemmMaps maps, const GemmShape shape,
                 const __grid_constant__ typename Epi::Params ep) {
    constexpr int kH = EpiHalvesOf<Epi>::v;
    using Cfg = GemmCfg<BN, Epi::kIoBufs, kH, EpiIoOwnersOf<Epi>::v>;
    constexpr int kStages = Cfg::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* io_smem = smem + kStages * Cfg::kStageBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(io_smem + Cfg::kIoBytes);
    uint64_t* empty_bar = full_bar + kStages;
    uint64_t* tmem_full = empty_bar + kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint64_t* io_bar = tmem_empty + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(io_bar + kIoBars * Cfg::kWarpsEpi);

    const int warp = warp_id_uniform();
    const int lane = threadIdx.x & 31;

    pdl_launch_dependents();
    if (warp == 0 && lane == 0) {
        for (int s = 0; s < shape.nseg; ++s) {
            tma_prefetch_desc(&maps.a[s]);
            tma_prefetch_desc(&maps.b[s]);
        }
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full_bar[i], 1);
            mbar_init(&empty_bar[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 128 * kH);
        }
        for (int i = 0; i < kIoBars * Cfg::kWarpsEpi; ++i) mbar_init(&io_bar[i], 1);
        fence_mbar_init();
    }
    if (warp == 2) {
        tmem_alloc(tmem_ptr, Cfg::kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = uniform_u32(*tmem_ptr);
    pdl_wait();  // everything above touched only shared memory / TMEM; global memory from here on

    const int tiles_per_group = shape.tiles_m * shape.tiles_n;
    const int nprob = shape.nprob > 1 ? shape.nprob : 1;
    const int total_tiles = tiles_per_group * shape.G * nprob;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                const int gg = tile / tiles_per_group;
                const int prob = gg / shape.G;
                const int g = gg - prob * shape.G;
                const int rem = tile - gg * tiles_per_group;
                const int tn = rem / shape.tiles_m;
                const int tm = rem - tn * shape.tiles_m;
                const int m0 = tm * kBlockM, n0 = tn * BN;
                const int seg_lo = nprob > 1 ? prob : 0, seg_hi = nprob > 1 ? prob + 1 : shape.nseg;
                for (int seg = seg_lo; seg < seg_hi; ++seg) {
                    const int kblocks = (shape.K[seg] + kBlockK - 1) / kBlockK;
                    const int ga = g * shape.a_gmul[seg], gb = g * shape.b_gmul[seg];
                    for (int kb = 0; kb < kblocks; ++kb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        mbar_arrive_expect_tx(&full_bar[stage], Cfg::kStageBytes);
                        uint8_t* sa = smem + stage * Cfg::kStageBytes;
                        uint8_t* sb = sa + kATileBytes;
                        const int k0 = kb * kBlockK;
                        if (!A_MN) {
                            tma_load_3d(&maps.a[seg], &full_bar[stage], sa, k0, m0, ga);
                        } else if (shape.a_one_box[seg]) {
                            tma_load_4d(&maps.a[seg], &full_bar[stage], sa, 0, k0, m0 >> 5, ga);
                        } else {
#pragma unroll
                            for (int i = 0; i < kBlockM / 32; ++i)
                                tma_load_3d(&maps.a[seg], &full_bar[stage], sa + i * 4096, m0 + 32 * i, k0, ga);
                        }
                        if (!B_MN) {
                            tma_load_3d(&maps.b[seg], &full_bar[stage], sb, k0, n0, gb);
                        } else if (shape.b_one_box[seg]) {
                            tma_load_4d(&maps.b[seg], &full_bar[stage], sb, 0, k0, n0 >> 5, gb);
                        } else {
#pragma unroll
                            for (int i = 0; i < BN / 32; ++i)
                                tma_load_3d(&maps.b[seg], &full_bar[stage], sb + i * 4096, n0 + 32 * i, k0, gb);
                        }
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc_tf32(kBlockM, BN, A_MN, B_MN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
                const int as = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                mbar_wait(&tmem_empty[as], aph ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                uint32_t accumulate = 0;
                const int prob = (tile / tiles_per_group) / shape.G;
                const int seg_lo = nprob > 1 ? prob : 0, seg_hi = nprob > 1 ? prob + 1 : shape.nseg;
                for (int seg = seg_lo; seg < seg_hi; ++seg) {
                    const int kblocks = (shape.K[seg] + kBlockK - 1) / kBlockK;
                    for (int kb = 0; kb < kblocks; ++kb) {
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        const uint32_t a_base = smem_u32(smem + stage * Cfg::kStageBytes);
                        const uint32_t b_base = a_base + kATileBytes;
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                            // K-major : 8-row groups 1024 B apart (SBO); K advances 32 B inside the swizzle span
                            // MN-major: 32-wide MN chunks 4096 B apart (LBO); 4-k-row groups 512 B apart (SBO);
                            //           K advances 8 rows = 1024 B
                            const uint64_t da = A_MN ? make_smem_desc(a_base + k * 1024, 4096, 512, kLayoutSw128Base32)
                                                     : make_smem_desc(a_base + k * 32, 16, 1024, kLayoutSw128);
                            const uint64_t db = B_MN ? make_smem_desc(b_base + k * 1024, 4096, 512, kLayoutSw128Base32)
                                                     : make_smem_desc(b_base + k * 32, 16, 1024, kLayoutSw128);
                            umma_tf32(d_tmem, da, db, idesc, accumulate);
                            accumulate = 1;
                        }
                        umma_commit(&empty_bar[stage]);  // frees the smem stage once these MMAs retire
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
                umma_commit(&tmem_full[as]);  // accumulator ready for the epilogue
            }
        }
    } else if (warp == 3) {
        run_side_job<Epi>(ep, blockIdx.x, gridDim.x, 0);
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue
        const int q = warp & 3;
        const int half = (warp - 4) >> 2;
        uint32_t io_phase = 0;
        const int io_owner = EpiColSplitOf<Epi>::v == 2 ? q : warp - 4;
        EpiIo io{io_smem + io_owner * (Epi::kIoBufs * kIoTileBytes), &io_bar[kIoBars * io_owner], &io_phase};
        int it = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
            const int gg = tile / tiles_per_group;
            const int prob = gg / shape.G;
            const int g = gg - prob * shape.G;
            const int rem = tile - gg * tiles_per_group;
            const int tn = rem / shape.tiles_m;
            const int tm = rem - tn * shape.tiles_m;
            const int as = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            RowCtx ctx;
            ctx.g = g; ctx.row0 = tm * kBlockM + q * 32; ctx.row = ctx.row0 + lane; ctx.col0 = tn * BN;
            ctx.M = shape.M; ctx.N = shape.N; ctx.tile_m = tm; ctx.tile_n = tn; ctx.tiles_m = shape.tiles_m; ctx.tiles_n = shape.tiles_n; ctx.warp_q = q; ctx.prob = prob; ctx.half = half; ctx.halves = kH; ctx.part = q + 4 * half;
            ctx.split = 0; ctx.bn = BN; ctx.first = it == 0;
            ctx.next_g = -1; ctx.next_prob = 0; ctx.next_row0 = 0; ctx.next_col0 = 0;
            if (tile + static_cast<int>(gridDim.x) < total_tiles) {
                const int nt = tile + gridDim.x;
                const int ngg = nt / tiles_per_group;
                const int nrem = nt - ngg * tiles_per_group;
                const int ntn = nrem / shape.tiles_m;
                ctx.next_g = ngg % shape.G; ctx.next_prob = ngg / shape.G; ctx.next_row0 = (nrem - ntn * shape.tiles_m) * kBlockM + q * 32; ctx.next_col0 = ntn * BN;
            }
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
            epilogue_tile<BN, Epi>(ep, ctx, io, &tmem_full[as], aph, taddr, [&] { mbar_arrive(&tmem_empty[as]); });
        }
        if (Epi::kIoBufs > 0) io_wait_stores_done();
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc(tmem_base, Cfg::kTmemCols);
    }
}

// SC_B200_PDL=0 switches programmatic dependent launch off (every kernel then waits for its predecessor to drain);
// 1 / 2 enable it for the plain / the GEMM kernels only, 3 (default) for both.
inline int pdl_mask() {  // bit 0: plain kernels, bit 1: GEMM kernels
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("SC_B200_PDL");
        v = e ? atoi(e) : 3;
    }
    return v;
}
inline bool use_pdl() { return (pdl_mask() & 2) != 0; }
inline bool use_pdl_plain() { return (pdl_mask() & 1) != 0; }

// Host-side launcher: grid = min(#tiles, #SMs) persistent CTAs.
template <int BN, bool A_MN, bool B_MN, class Epi>
cudaError_t launch_gemm(const GemmMaps& maps, GemmShape shape, const typename Epi::Params& ep, int num_sms,
                        cudaStream_t stream) {
    using Cfg = GemmCfg<BN, Epi::kIoBufs, EpiHalvesOf<Epi>::v, EpiIoOwnersOf<Epi>::v>;
    auto kern = gemm_tf32_kernel<BN, A_MN, B_MN, Epi>;
    static bool attr_set[64] = {};  // per template instantiation, per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_set[dev & 63] = true;
    }
    shape.tiles_m = (shape.M + kBlockM - 1) / kBlockM;
    shape.tiles_n = (shape.N + BN - 1) / BN;
    const long long total = 1LL * shape.tiles_m * shape.tiles_n * shape.G * (shape.nprob > 1 ? shape.nprob : 1);
    if (total <= 0) return cudaSuccess;
    const int grid = static_cast<int>(total < num_sms ? total : num_sms);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(Cfg::kThreads);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = use_pdl() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, maps, shape, ep);
}

}  // namespace sc
