// CTA-pair variant of the GEMM core: tcgen05.mma.cta_group::2, one 256 x BN accumulator tile per
// cluster of two CTAs (two SMs of a TPC).  Each CTA stages its own 128 rows of A and HALF of the B
// tile (BN/2 rows), so a pipeline stage is 16 KB + BN/2*128 B instead of 16 KB + BN*128 B: for BN = 256
// that is 32 KB instead of 48 KB per 128x256x32 MACs of work per SM - a third less L2 -> SM traffic
// and 7 stages instead of 4 in the same shared memory, i.e. more latency tolerance where the
// single-CTA kernel stalls on operand delivery (ncu: tensor pipe ~37 %, L2 ~40 %, nothing saturated).
//
// Differences to gemm_core.cuh, following the 2-SM pattern of the public sm_100 GEMMs:
//   * TMA loads of both CTAs complete on the LEADER's full barrier (2 arrivals: leader's
//     arrive.expect_tx for both CTAs' bytes + the peer's remote arrive)
//   * only the leader issues MMAs; tcgen05.commit is multicast to both CTAs' empty / tmem_full barriers
//   * epilogue warps of both CTAs arrive remotely on the leader's tmem_empty barrier
//   * TMEM alloc / dealloc with cta_group::2 in both CTAs, cluster barriers around set-up and tear-down
// Epilogue functors, operand descriptions and the RowCtx are shared with the single-CTA core.
#pragma once
#include <cuda_runtime.h>

#include "gemm_core.cuh"

namespace sc {

template <int BN, int IO = 0, int H = kEpiHalves, int OWNERS = 4 * H>
struct GemmCfg2 {
    static constexpr int kWarpsEpi = 4 * H;
    static constexpr int kThreads = 128 + 32 * kWarpsEpi;
    static constexpr int kBHalf = BN / 2;
    static constexpr int kBTileBytes = kBHalf * kBlockK * 4;
    static constexpr int kStageBytes = kATileBytes + kBTileBytes;
    static constexpr int kIoBytes = OWNERS * IO * kIoTileBytes;
    static constexpr int kStagesRaw = (227 * 1024 - 2048 - kIoBytes) / kStageBytes;
    static constexpr int kStages = kStagesRaw > 8 ? 8 : kStagesRaw;
    static constexpr int kTmemCols = GemmCfg<BN>::kTmemCols;
    static constexpr int kBarrierBytes = (2 * kStages + 4 + kIoBars * kWarpsEpi) * 8 + 16;
    static constexpr int kSmemBytes = kStages * kStageBytes + kIoBytes + kBarrierBytes + 1024;
    static_assert(BN % 64 == 0 && BN >= 64 && BN <= 256, "paired kernel: BN multiple of 64 in [64,256]");
    static_assert(kStages >= 2, "not enough shared memory left for the operand pipeline");
};

// EB = bytes per operand element: 4 (fp32 storage, .kind::tf32, 32 elements per 128-byte k-block row) or
// 2 (fp16, .kind::f16, 64 elements).  Tile bytes, stage bytes and the four MMAs per k-block are the same.
// MN-major fp16 tiles use the plain 128-byte swizzle: chunks of 64 m/n x 64 k (8 KB), 8-row groups 1 KB apart.
template <int BN, bool A_MN, bool B_MN, class Epi, int EB = 4>
__global__ void __launch_bounds__(128 + 128 * EpiHalvesOf<Epi>::v, 1)
gemm_tf32_2cta_kernel(const __grid_constant__ GemmMaps maps, const GemmShape shape,
                      const __grid_constant__ typename Epi::Params ep) {
    constexpr int kH = EpiHalvesOf<Epi>::v;
    using Cfg = GemmCfg2<BN, Epi::kIoBufs, kH, EpiIoOwnersOf<Epi>::v>;
    constexpr int kStages = Cfg::kStages;
    constexpr int kBHalf = Cfg::kBHalf;
    constexpr int kKElems = 128 / EB;            // operand elements per k-block
    constexpr int kMnChunk = EB == 4 ? 32 : 64;  // m/n elements per 128-byte MN-major chunk
    constexpr int kMnChunkBytes = kMnChunk == 32 ? 4096 : 8192;
    constexpr int kMnShift = EB == 4 ? 5 : 6;
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
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;

    pdl_launch_dependents();
    cluster_sync_all();  // both CTAs resident before any cross-CTA state is touched
    if (warp == 0 && lane == 0) {
        for (int s = 0; s < shape.nseg; ++s) {
            tma_prefetch_desc(&maps.a[s]);
            tma_prefetch_desc(&maps.b[s]);
        }
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(&full_bar[i], 2);   // leader's expect_tx arrive + peer's remote arrive (used on the leader only)
            mbar_init(&empty_bar[i], 1);  // multicast commit from the leader's MMA thread
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 2 * 128 * kH);  // epilogue threads of BOTH CTAs (leader's copy is used)
        }
        for (int i = 0; i < kIoBars * Cfg::kWarpsEpi; ++i) mbar_init(&io_bar[i], 1);
        fence_mbar_init();
    }
    if (warp == 2) {
        tmem_alloc_2cta(tmem_ptr, Cfg::kTmemCols);
        tmem_relinquish_2cta();
    }
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = uniform_u32(*tmem_ptr);
    pdl_wait();  // everything above touched only shared memory / TMEM; global memory from here on

    const int nprob = shape.nprob > 1 ? shape.nprob : 1;
    const int tiles_per_group = shape.tiles_m * shape.tiles_n;  // tiles_m counts 256-row tiles here
    const int ksplit = shape.ksplit > 1 ? shape.ksplit : 1;  // partial tiles per output tile (split fastest)
    const int total_tiles = tiles_per_group * shape.G * nprob * ksplit;
    const int cluster_id = blockIdx.x >> 1;
    const int num_clusters = gridDim.x >> 1;

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        {   // the whole (converged) warp walks the loops; one elected lane issues the copies
            int stage = 0;
            uint32_t phase = 0;
            for (int ptile = cluster_id; ptile < total_tiles; ptile += num_clusters) {
                const int tile = ptile / ksplit, split = ptile - tile * ksplit;
                const int gg = tile / tiles_per_group;
                const int prob = gg / shape.G;
                const int g = gg - prob * shape.G;
                const int rem = tile - gg * tiles_per_group;
                const int tn = rem / shape.tiles_m;
                const int tm = rem - tn * shape.tiles_m;
                const int m0 = tm * (2 * kBlockM) + rank * kBlockM;
                const int n0 = tn * BN + rank * kBHalf;
                const int seg_lo = nprob > 1 ? prob : 0, seg_hi = nprob > 1 ? prob + 1 : shape.nseg;
                for (int seg = seg_lo; seg < seg_hi; ++seg) {
                    const int kblocks = (shape.K[seg] + kKElems - 1) / kKElems;
                    const int kper = (kblocks + ksplit - 1) / ksplit;
                    const int kb_lo = split * kper, kb_hi = kb_lo + kper < kblocks ? kb_lo + kper : kblocks;
                    const int ga = g * shape.a_gmul[seg], gb = g * shape.b_gmul[seg];
                    for (int kb = kb_lo; kb < kb_hi; ++kb) {
                        mbar_wait(&empty_bar[stage], phase ^ 1);
                        uint8_t* sa = smem + stage * Cfg::kStageBytes;
                        uint8_t* sb = sa + kATileBytes;
                        const int k0 = kb * kKElems;
                        if (elect_one()) {
                        if (!A_MN) {
                            tma_load_3d_2cta(&maps.a[seg], &full_bar[stage], sa, k0, m0, ga);
                        } else if (shape.a_one_box[seg]) {
                            tma_load_4d_2cta(&maps.a[seg], &full_bar[stage], sa, 0, k0, m0 >> kMnShift, ga);
                        } else {
#pragma unroll
                            for (int i = 0; i < kBlockM / kMnChunk; ++i)
                                tma_load_3d_2cta(&maps.a[seg], &full_bar[stage], sa + i * kMnChunkBytes, m0 + kMnChunk * i, k0, ga);
                        }
                        if (!B_MN) {
                            tma_load_3d_2cta(&maps.b[seg], &full_bar[stage], sb, k0, n0, gb);
                        } else if (shape.b_one_box[seg]) {
                            tma_load_4d_2cta(&maps.b[seg], &full_bar[stage], sb, 0, k0, n0 >> kMnShift, gb);
                        } else {
#pragma unroll
                            for (int i = 0; i < (kBHalf + kMnChunk - 1) / kMnChunk; ++i)
                                tma_load_3d_2cta(&maps.b[seg], &full_bar[stage], sb + i * kMnChunkBytes, n0 + kMnChunk * i, k0, gb);
                        }
                        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::kStageBytes);
                        else mbar_arrive_remote(&full_bar[stage], 0);
                        }
                        __syncwarp();
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA only)
        if (leader) {  // the whole (converged) warp walks the loops; one elected lane issues
            constexpr uint32_t idesc = EB == 4 ? make_idesc_tf32(2 * kBlockM, BN, A_MN, B_MN) : make_idesc_f16(2 * kBlockM, BN, A_MN, B_MN);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int ptile = cluster_id; ptile < total_tiles; ptile += num_clusters, ++it) {
                const int tile = ptile / ksplit, split = ptile - tile * ksplit;
                const int as = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                mbar_wait(&tmem_empty[as], aph ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                uint32_t accumulate = 0;
                const int prob = (tile / tiles_per_group) / shape.G;
                const int seg_lo = nprob > 1 ? prob : 0, seg_hi = nprob > 1 ? prob + 1 : shape.nseg;
                for (int seg = seg_lo; seg < seg_hi; ++seg) {
                    const int kblocks = (shape.K[seg] + kKElems - 1) / kKElems;
                    const int kper = (kblocks + ksplit - 1) / ksplit;
                    const int kb_lo = split * kper, kb_hi = kb_lo + kper < kblocks ? kb_lo + kper : kblocks;
                    for (int kb = kb_lo; kb < kb_hi; ++kb) {
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        const uint32_t a_base = smem_u32(smem + stage * Cfg::kStageBytes);
                        const uint32_t b_base = a_base + kATileBytes;
                        if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                            // one MMA consumes 32 bytes of K: 8 tf32 or 16 fp16 elements
                            uint64_t da, db;
                            const uint32_t acc = k > 0 ? 1u : accumulate;  // only a tile's very first MMA overwrites
                            if (EB == 4) {
                                da = A_MN ? make_smem_desc(a_base + k * 1024, 4096, 512, kLayoutSw128Base32)
                                          : make_smem_desc(a_base + k * 32, 16, 1024, kLayoutSw128);
                                db = B_MN ? make_smem_desc(b_base + k * 1024, 4096, 512, kLayoutSw128Base32)
                                          : make_smem_desc(b_base + k * 32, 16, 1024, kLayoutSw128);
                                umma_tf32_2cta(d_tmem, da, db, idesc, acc);
                            } else {
                                // MN-major fp16: 16 k-rows per MMA = two 8-row groups (SBO 1024); 64-wide chunks 8 KB apart (LBO)
                                da = A_MN ? make_smem_desc(a_base + k * 2048, 8192, 1024, kLayoutSw128)
                                          : make_smem_desc(a_base + k * 32, 16, 1024, kLayoutSw128);
                                db = B_MN ? make_smem_desc(b_base + k * 2048, 8192, 1024, kLayoutSw128)
                                          : make_smem_desc(b_base + k * 32, 16, 1024, kLayoutSw128);
                                umma_f16_2cta(d_tmem, da, db, idesc, acc);
                            }
                        }
                        umma_commit_2cta(&empty_bar[stage]);
                        }
                        __syncwarp();
                        accumulate = 1;
                        if (++stage == kStages) { stage = 0; phase ^= 1; }
                    }
                }
                if (elect_one()) umma_commit_2cta(&tmem_full[as]);
                __syncwarp();
            }
        }
    } else if (warp == 3) {
        run_side_job<Epi>(ep, blockIdx.x, gridDim.x, 0);
    } else if (warp >= 4) {
        // ------------------------------------------------------------------ epilogue (both CTAs, own 128 rows)
        const int q = warp & 3;
        const int half = (warp - 4) >> 2;
        uint32_t io_phase = 0;
        const int io_owner = EpiColSplitOf<Epi>::v == 2 ? q : warp - 4;
        EpiIo io{io_smem + io_owner * (Epi::kIoBufs * kIoTileBytes), &io_bar[kIoBars * io_owner], &io_phase};
        int it = 0;
        for (int ptile = cluster_id; ptile < total_tiles; ptile += num_clusters, ++it) {
            const int tile = ptile / ksplit, split = ptile - tile * ksplit;
            const int gg = tile / tiles_per_group;
            const int prob = gg / shape.G;
            const int g = gg - prob * shape.G;
            const int rem = tile - gg * tiles_per_group;
            const int tn = rem / shape.tiles_m;
            const int tm = rem - tn * shape.tiles_m;
            const int as = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            RowCtx ctx;
            ctx.g = g; ctx.row0 = tm * (2 * kBlockM) + rank * kBlockM + q * 32; ctx.row = ctx.row0 + lane; ctx.col0 = tn * BN;
            ctx.M = shape.M; ctx.N = shape.N;
            // partial-sum slots are indexed by 128-row tile: report this CTA's half as its own row tile
            ctx.tile_m = tm * 2 + rank; ctx.tile_n = tn; ctx.tiles_m = shape.tiles_m * 2; ctx.tiles_n = shape.tiles_n;
            ctx.warp_q = q; ctx.prob = prob; ctx.half = half; ctx.halves = kH; ctx.part = q + 4 * half;
            ctx.split = split; ctx.bn = BN; ctx.first = it == 0;
            ctx.next_g = -1; ctx.next_prob = 0; ctx.next_row0 = 0; ctx.next_col0 = 0;
            if (ptile + num_clusters < total_tiles) {
                const int nt = (ptile + num_clusters) / ksplit;
                const int ngg = nt / tiles_per_group;
                const int nrem = nt - ngg * tiles_per_group;
                const int ntn = nrem / shape.tiles_m;
                ctx.next_g = ngg % shape.G;
                ctx.next_prob = ngg / shape.G;
                ctx.next_row0 = (nrem - ntn * shape.tiles_m) * (2 * kBlockM) + rank * kBlockM + q * 32;
                ctx.next_col0 = ntn * BN;
            }
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
            epilogue_tile<BN, Epi>(ep, ctx, io, &tmem_full[as], aph, taddr, [&] {
                if (leader) mbar_arrive(&tmem_empty[as]);
                else mbar_arrive_remote(&tmem_empty[as], 0);
            });
        }
        if (Epi::kIoBufs > 0) io_wait_stores_done();
    }

    tc_fence_before();
    cluster_sync_all();  // every remote arrive has landed and every TMEM read is done in both CTAs
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, Cfg::kTmemCols);
    }
}

// grid = 2 * min(#tiles, #SMs / 2) persistent CTAs in clusters of two.
template <int BN, bool A_MN, bool B_MN, class Epi, int EB = 4>
cudaError_t launch_gemm_2cta(const GemmMaps& maps, GemmShape shape, const typename Epi::Params& ep, int num_sms,
                             cudaStream_t stream) {
    using Cfg = GemmCfg2<BN, Epi::kIoBufs, EpiHalvesOf<Epi>::v, EpiIoOwnersOf<Epi>::v>;
    auto kern = gemm_tf32_2cta_kernel<BN, A_MN, B_MN, Epi, EB>;
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_set[dev & 63] = true;
    }
    shape.tiles_m = (shape.M + 2 * kBlockM - 1) / (2 * kBlockM);
    shape.tiles_n = (shape.N + BN - 1) / BN;
    const long long total = 1LL * shape.tiles_m * shape.tiles_n * shape.G * (shape.nprob > 1 ? shape.nprob : 1) *
                            (shape.ksplit > 1 ? shape.ksplit : 1);
    if (total <= 0) return cudaSuccess;
    const int pairs = num_sms / 2;
    const int clusters = static_cast<int>(total < pairs ? total : pairs);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(Cfg::kThreads);
    cfg.dynamicSmemBytes = Cfg::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = use_pdl() ? 2 : 1;
    return cudaLaunchKernelEx(&cfg, kern, maps, shape, ep);
}

}  // namespace sc
