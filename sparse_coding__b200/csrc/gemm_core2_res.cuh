// CTA-pair GEMM with a RESIDENT A operand, for fp16 GEMMs whose K loop is short (K <= 512: the encode and the
// dL/dz GEMMs of the train step, K = activation width).
//
// Why: with 256 x 256 tiles both operands stream from L2 for every tile, 512 KB per tile and CTA pair, and at the
// tensor-core rate that is ~11 TB/s chip-wide - the L2's own throughput limit (measured: the mainloop alone runs
// at the tensor peak, and every byte the epilogue moves through L2 on top of it costs time 1:1).  Here every CTA
// pair walks a CONTIGUOUS range of tiles in the order (group, row tile, column tile), so consecutive tiles share
// their A rows: the pair's 2 x 128 rows x K of A stay in shared memory (K/64 k-blocks of 16 KB per CTA) and are
// loaded once per row tile, only B streams through the ring.  Operand traffic per tile drops from 512 KB to
// 256 KB + 256 KB / (column tiles per row tile).
//
// Protocol on top of gemm_core2.cuh's (same warp roles, same TMEM double buffering, same epilogue functors):
//   a_full[kb]  A k-block kb of the current row tile has landed (both CTAs' halves; leader's barrier)
//   a_free[kb]  every MMA that reads k-block kb of the CURRENT row tile has completed (multicast commit, issued
//               in the last tile of the row tile) - the producer may overwrite it with the next row tile's
// so the next row tile's A streams in k-block by k-block while the last tile of the old one is still computing.
#pragma once
#include "gemm_core2.cuh"

namespace sc {

// bytes of staging memory per owner: Epi::kIoBytes when the functor states it, else kIoBufs x 4 KB
template <class E, class = void>
struct EpiIoBytesOf { static constexpr int v = E::kIoBufs * kIoTileBytes; };
template <class E>
struct EpiIoBytesOf<E, decltype((void)E::kIoBytes)> { static constexpr int v = E::kIoBytes; };

constexpr int kResMaxKBlocks = 8;

template <int BN, class Epi>
struct GemmCfgRes {
    static constexpr int kH = EpiHalvesOf<Epi>::v;
    static constexpr int kWarpsEpi = 4 * kH;
    static constexpr int kThreads = 128 + 32 * kWarpsEpi;
    static constexpr int kBHalf = BN / 2;
    static constexpr int kBTileBytes = kBHalf * 128;  // 64 fp16 of K per row
    static constexpr int kIoBytes = EpiIoOwnersOf<Epi>::v * EpiIoBytesOf<Epi>::v;
    static constexpr int kBarrierBytes = (4 * kResMaxKBlocks + 4 + kIoBars * kWarpsEpi) * 8 + 16;
    static constexpr int kTmemCols = GemmCfg<BN>::kTmemCols;
    static constexpr int kSmemMax = 227 * 1024;
    // stages of the B ring that fit beside `kblocks` resident k-blocks of A
    static int stages(int kblocks) {
        const int left = kSmemMax - 1024 - kBarrierBytes - kIoBytes - kblocks * kATileBytes;
        int s = left / kBTileBytes;
        return s > kResMaxKBlocks ? kResMaxKBlocks : s;
    }
    static int smem_bytes(int kblocks, int stages_) {
        return kblocks * kATileBytes + stages_ * kBTileBytes + kIoBytes + kBarrierBytes + 1024;
    }
};

template <int BN, class Epi>
__global__ void __launch_bounds__(128 + 128 * EpiHalvesOf<Epi>::v, 1)
gemm_f16_resa_kernel(const __grid_constant__ GemmMaps maps, const GemmShape shape,
                     const __grid_constant__ typename Epi::Params ep) {
    using Cfg = GemmCfgRes<BN, Epi>;
    constexpr int kH = Cfg::kH;
    constexpr int kBHalf = Cfg::kBHalf;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int kblocks = (shape.K[0] + 63) / 64;
    const int stages = shape.res_stages;
    uint8_t* a_res = smem;
    uint8_t* b_ring = a_res + kblocks * kATileBytes;
    uint8_t* io_smem = b_ring + stages * Cfg::kBTileBytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(io_smem + Cfg::kIoBytes);
    uint64_t* empty_bar = full_bar + kResMaxKBlocks;
    uint64_t* a_full = empty_bar + kResMaxKBlocks;
    uint64_t* a_free = a_full + kResMaxKBlocks;
    uint64_t* tmem_full = a_free + kResMaxKBlocks;
    uint64_t* tmem_empty = tmem_full + 2;
    uint64_t* io_bar = tmem_empty + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(io_bar + kIoBars * Cfg::kWarpsEpi);

    const int warp = warp_id_uniform();
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = rank == 0;

    pdl_launch_dependents();
    cluster_sync_all();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&maps.a[0]);
        tma_prefetch_desc(&maps.b[0]);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kResMaxKBlocks; ++i) {
            mbar_init(&full_bar[i], 2);
            mbar_init(&empty_bar[i], 1);
            mbar_init(&a_full[i], 2);
            mbar_init(&a_free[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(&tmem_full[i], 1);
            mbar_init(&tmem_empty[i], 2 * 128 * kH);
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
    pdl_wait();

    // contiguous range of tiles for this CTA pair; tile t = (g * tiles_m + tm) * tiles_n + tn
    const int tiles_n = shape.tiles_n, tiles_m = shape.tiles_m;
    const long long total = (long long)tiles_m * tiles_n * shape.G;
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    const int t_lo = static_cast<int>(total * cluster_id / num_clusters);
    const int t_hi = static_cast<int>(total * (cluster_id + 1) / num_clusters);

    if (warp == 0) {
        // ------------------------------------------------------------------ TMA producer (both CTAs)
        {   // the whole (converged) warp walks the loops; one elected lane issues the copies
            int stage = 0;
            uint32_t phase = 0;
            int a_gen = 0, cur_group = -1;
            for (int t = t_lo; t < t_hi; ++t) {
                const int group = t / tiles_n, tn = t - group * tiles_n;
                const int g = group / tiles_m, tm = group - g * tiles_m;
                const int m0 = tm * (2 * kBlockM) + rank * kBlockM;
                const int n0 = tn * BN + rank * kBHalf;
                const bool new_a = group != cur_group;
                const int ga = g * shape.a_gmul[0], gb = g * shape.b_gmul[0];
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (new_a) {
                        if (a_gen > 0) mbar_wait(&a_free[kb], (a_gen - 1) & 1);  // the old row tile's MMAs on this k-block are done
                        if (elect_one()) {
                            tma_load_3d_2cta(&maps.a[0], &a_full[kb], a_res + kb * kATileBytes, kb * 64, m0, ga);
                            if (leader) mbar_arrive_expect_tx(&a_full[kb], 2 * kATileBytes);
                            else mbar_arrive_remote(&a_full[kb], 0);
                        }
                        __syncwarp();
                    }
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (elect_one()) {
                        tma_load_3d_2cta(&maps.b[0], &full_bar[stage], b_ring + stage * Cfg::kBTileBytes, kb * 64, n0, gb);
                        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::kBTileBytes);
                        else mbar_arrive_remote(&full_bar[stage], 0);
                    }
                    __syncwarp();
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
                if (new_a) { ++a_gen; cur_group = group; }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------------ MMA issuer (leader CTA only)
        if (leader) {  // whole warp, one elected lane issues
            constexpr uint32_t idesc = make_idesc_f16(2 * kBlockM, BN, false, false);
            int stage = 0;
            uint32_t phase = 0;
            int a_gen = 0, cur_group = -1;
            int it = 0;
            for (int t = t_lo; t < t_hi; ++t, ++it) {
                const int group = t / tiles_n;
                const bool new_a = group != cur_group;
                const bool last_of_group = (t + 1 == t_hi) || ((t + 1) / tiles_n != group);
                const int as = it & 1;
                const uint32_t aph = (it >> 1) & 1;
                mbar_wait(&tmem_empty[as], aph ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                uint32_t accumulate = 0;
                for (int kb = 0; kb < kblocks; ++kb) {
                    if (new_a) mbar_wait(&a_full[kb], a_gen & 1);
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(a_res + kb * kATileBytes);
                    const uint32_t b_base = smem_u32(b_ring + stage * Cfg::kBTileBytes);
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < 4; ++k) {  // 16 fp16 of K per MMA, 32 bytes inside the 128-byte swizzle span
                            const uint64_t da = make_smem_desc(a_base + k * 32, 16, 1024, kLayoutSw128);
                            const uint64_t db = make_smem_desc(b_base + k * 32, 16, 1024, kLayoutSw128);
                            umma_f16_2cta(d_tmem, da, db, idesc, k > 0 ? 1u : accumulate);
                        }
                        umma_commit_2cta(&empty_bar[stage]);
                        if (last_of_group) umma_commit_2cta(&a_free[kb]);
                    }
                    __syncwarp();
                    accumulate = 1;
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
                if (elect_one()) umma_commit_2cta(&tmem_full[as]);
                __syncwarp();
                if (new_a) { ++a_gen; cur_group = group; }
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
        EpiIo io{io_smem + io_owner * EpiIoBytesOf<Epi>::v, &io_bar[kIoBars * io_owner], &io_phase};
        int it = 0;
        for (int t = t_lo; t < t_hi; ++t, ++it) {
            const int group = t / tiles_n, tn = t - group * tiles_n;
            const int g = group / tiles_m, tm = group - g * tiles_m;
            const int as = it & 1;
            const uint32_t aph = (it >> 1) & 1;
            RowCtx ctx;
            ctx.g = g; ctx.row0 = tm * (2 * kBlockM) + rank * kBlockM + q * 32; ctx.row = ctx.row0 + lane; ctx.col0 = tn * BN;
            ctx.M = shape.M; ctx.N = shape.N;
            ctx.tile_m = tm * 2 + rank; ctx.tile_n = tn; ctx.tiles_m = tiles_m * 2; ctx.tiles_n = tiles_n;
            ctx.warp_q = q; ctx.prob = 0; ctx.half = half; ctx.halves = kH; ctx.part = q + 4 * half;
            ctx.split = 0; ctx.bn = BN; ctx.first = t == t_lo;
            ctx.next_g = -1; ctx.next_prob = 0; ctx.next_row0 = 0; ctx.next_col0 = 0;
            if (t + 1 < t_hi) {
                const int ngroup = (t + 1) / tiles_n, ntn = (t + 1) - ngroup * tiles_n;
                const int ng = ngroup / tiles_m, ntm = ngroup - ng * tiles_m;
                ctx.next_g = ng;
                ctx.next_row0 = ntm * (2 * kBlockM) + rank * kBlockM + q * 32;
                ctx.next_col0 = ntn * BN;
            }
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
            epilogue_tile<BN, Epi>(ep, ctx, io, &tmem_full[as], aph, taddr, [&] {
                if (leader) mbar_arrive(&tmem_empty[as]);
                else mbar_arrive_remote(&tmem_empty[as], 0);
            });
        }
        if (EpiIoBytesOf<Epi>::v > 0) io_wait_stores_done();
    }

    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc_2cta(tmem_base, Cfg::kTmemCols);
    }
}

// grid = 2 * min(#tiles, #SMs / 2) persistent CTAs in clusters of two; shape.K[0] <= 512, K-major fp16 operands.
template <int BN, class Epi>
cudaError_t launch_gemm_resa(const GemmMaps& maps, GemmShape shape, const typename Epi::Params& ep, int num_sms,
                             cudaStream_t stream) {
    using Cfg = GemmCfgRes<BN, Epi>;
    auto kern = gemm_f16_resa_kernel<BN, Epi>;
    static bool attr_set[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!attr_set[dev & 63]) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemMax);
        if (e != cudaSuccess) return e;
        attr_set[dev & 63] = true;
    }
    const int kblocks = (shape.K[0] + 63) / 64;
    shape.res_stages = Cfg::stages(kblocks);
    if (kblocks > kResMaxKBlocks || shape.res_stages < 2) return cudaErrorInvalidConfiguration;
    shape.tiles_m = (shape.M + 2 * kBlockM - 1) / (2 * kBlockM);
    shape.tiles_n = (shape.N + BN - 1) / BN;
    const long long total = 1LL * shape.tiles_m * shape.tiles_n * shape.G;
    if (total <= 0) return cudaSuccess;
    const int pairs = num_sms / 2;
    const int clusters = static_cast<int>(total < pairs ? total : pairs);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters);
    cfg.blockDim = dim3(Cfg::kThreads);
    cfg.dynamicSmemBytes = Cfg::smem_bytes(kblocks, shape.res_stages);
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
