// Epilogues of the train step in fp16-operand mode.  Same mathematics as sae_epilogues.cuh; the tensors
// that only feed later GEMMs (codes c, residual r, dL/dz) are stored as fp16 - the same 11-bit
// significand as the TF32 rounding of the default mode, at half the bytes and twice the MMA rate.
// fp16 has a narrow exponent, so nothing tiny is stored: the residual is kept UNSCALED (r, not
// R = 2r/(B d)) and the 2/(B d) factor is applied in fp32 where the gradients leave the accumulators:
//   q' = r D^T = q (B d / 2);  dz' = (q' + lam d/2 [c>0]) [z>=0] = dz (B d / 2);
//   dW_e = (2/(B d)) dz'^T x;  dD = (2/(B d)) c^T r;  s = (2/(B d)) sum_b c q';  db = (2/(B d)) sum_b dz'.
#pragma once
#include "sae_epilogues.cuh"

namespace sc {

// G1: c = relu(x W^T + b) -> fp16 codes (z < 0 stored as -0), L1 partial sums
struct EpiEncodeH {
    static constexpr int kIoBufs = 1;  // one 4 KB buffer = two 2 KB fp16 staging tiles, used alternately
    struct alignas(64) Params {
        CUtensorMap c_map;  // fp16 io map of c [G][B][N]
        const float* bias;  // [G][N]
        float* l1_part;
        const int* dict_size;  // [G] or nullptr: features >= dict_size[g] are masked out
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float sum, sum1, sum2, sum3;  // four chains: the L1 sum is not a 32-long dependent add per chunk
    int live_end;  // first column that does not count (N, or dict_size[g] when masked)
    SC_DEVICE EpiEncodeH(const Params& p_, const RowCtx& c_, const EpiIo& io_)
        : p(p_), c(c_), io(io_), sum(0.f), sum1(0.f), sum2(0.f), sum3(0.f) {
        live_end = p.dict_size ? min(p.dict_size[c.g], c.N) : c.N;
    }
    float nxt[32];  // bias of the next chunk, loaded while the current one is processed
    SC_DEVICE void prefetch(int co) {
        const int col = c.col0 + co;
        const int ncols = c.N - col;
        if (ncols > 0 && p.bias) {
            load_row32(p.bias + (long long)c.g * c.N + col, nxt, ncols);
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) nxt[j] = 0.f;
        }
    }
    SC_DEVICE void chunk(int co, float (&v)[32], int next_co) {
        const int col = c.col0 + co;
        const int lane = threadIdx.x & 31;
        float bv[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) bv[j] = nxt[j];
        if (next_co >= 0) prefetch(next_co);
        const int live = live_end - col;
        // max(z, -0.0) is z for z >= 0 (max(+0, -0) = +0) and -0.0 for z < 0; adding -0.0 to the L1 sum is a no-op
        if (c.row < c.M && live >= 32) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                v[j] = fmaxf(v[j] + bv[j], -0.0f);
                v[j + 1] = fmaxf(v[j + 1] + bv[j + 1], -0.0f);
                v[j + 2] = fmaxf(v[j + 2] + bv[j + 2], -0.0f);
                v[j + 3] = fmaxf(v[j + 3] + bv[j + 3], -0.0f);
                sum += v[j]; sum1 += v[j + 1]; sum2 += v[j + 2]; sum3 += v[j + 3];
            }
        } else {
            const bool row_ok = c.row < c.M;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                v[j] = (row_ok && j < live) ? fmaxf(v[j] + bv[j], -0.0f) : -0.0f;  // masked features behave like z < 0
                sum += v[j];
            }
        }
        uint8_t* tile = io.buf + (((co >> 5) / kEpiHalves) & 1) * 2048;  // this warp's chunks alternate between the two tiles
        io_wait_stores_read_n<1>();  // the store issued two chunks ago has read this staging tile
        io_write_row_h(tile, lane, v);
        io_store_tile(&p.c_map, tile, col, c.row0, c.g);
        io_commit();
    }
    SC_DEVICE void finish() {
        if (!p.l1_part) return;  // inference use: codes only
        const float s = warp_sum((sum + sum1) + (sum2 + sum3));
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.l1_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * kEpiHalves) + c.part] = s;
        }
    }
};

// G2: r = c D - x -> fp16 residual (unscaled), MSE partial sums
struct EpiDecodeH {
    static constexpr int kIoBufs = 0;
    struct Params {
        const float* x;  // [Gx][B][d] fp32 target
        __half* r;       // [G][B][d]
        float* mse_part;
        long long x_gstride, out_gstride;
    };
    const Params& p;
    const RowCtx c;
    float sum;
    SC_DEVICE EpiDecodeH(const Params& p_, const RowCtx& c_, const EpiIo&) : p(p_), c(c_), sum(0.f) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;
        if (c.row >= c.M || col >= c.N) return;
        const int ncols = c.N - col;
        const long long off = (long long)c.row * c.N + col;
        float xv[32];
        load_row32(p.x + c.g * p.x_gstride + off, xv, ncols);
        __half* dst = p.r + c.g * p.out_gstride + off;
        if (ncols >= 32) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                __half2 h[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float r0 = v[8 * j + 2 * t] - xv[8 * j + 2 * t], r1 = v[8 * j + 2 * t + 1] - xv[8 * j + 2 * t + 1];
                    sum += r0 * r0 + r1 * r1;
                    h[t] = half2_sat(r0, r1);
                }
                reinterpret_cast<uint4*>(dst)[j] = *reinterpret_cast<uint4*>(h);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                if (j < ncols) {
                    const float r = v[j] - xv[j];
                    sum += r * r;
                    dst[j] = half_sat(r);
                }
            }
        }
    }
    SC_DEVICE void finish() {
        const float s = warp_sum(sum);
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.mse_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * kEpiHalves) + c.part] = s;
        }
    }
};

// G2 with split K (few output tiles: small ensembles): every partial accumulator tile is ADDED into a zeroed fp32
// x_hat buffer by TMA reduce-add; residual_h_kernel then forms r = x_hat - x and the MSE partial sums.
struct EpiReduceTile {
    static constexpr int kIoBufs = 1;
    struct alignas(64) Params {
        CUtensorMap out_map;  // fp32 io map of the accumulation buffer [G][rows][cols]
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    SC_DEVICE EpiReduceTile(const Params& p_, const RowCtx& c_, const EpiIo& io_) : p(p_), c(c_), io(io_) {}
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int lane = threadIdx.x & 31;
        io_wait_stores_read();
        io_write_row(io.buf, lane, v);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) tma_reduce_add_3d(&p.out_map, io.buf, c.col0 + co, c.row0, c.g);  // clipped at the tensor's edge
        io_commit();
    }
    SC_DEVICE void finish() {}
};

// G1 of the TRAIN step, computed transposed (rows = feature n, cols = batch b): the codes are kept as
// c^T [G][N][ldb] fp16 (z < 0 stored as -0), which makes them row accesses for every later consumer - the dL/dz
// epilogue (thread = feature) reads its own row, the weight-gradient GEMMs read c^T / dz^T K-major.  The bias is one
// scalar per thread.  Batch columns >= B (zero-filled x rows on a ragged tile) are written as -0 and clipped by the store.
template <int H>
struct EpiEncodeTHT {
    static constexpr int kHalves = H;  // epilogue warps per row quarter
    static constexpr int kIoBufs = 1;  // one 4 KB buffer = two 2 KB fp16 staging tiles, used alternately
    struct alignas(64) Params {
        CUtensorMap c_map;     // fp16 io map of c^T [G][N][B] (row stride ldb)
        const float* bias;     // [G][N]
        float* l1_part;
        const int* dict_size;  // [G] or nullptr: features >= dict_size[g] are masked out
        const float* fold_inv; // fold mode: [G][N] 1/max(||W_d row||,1e-8) (sign: clamp), else nullptr
        float* code_scale;     // fold mode, optional: [G][N] factor that turns the stored codes back into c (ScStepArgs.out_code_scale)
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float sum, sum1, sum2, sum3;
    float bcs, cs, ns;  // bias * cs; scale of the stored code (fold mode: 1/||W_d||, else 1) and its reciprocal
    bool live_row;
    SC_DEVICE EpiEncodeTHT(const Params& p_, const RowCtx& c_, const EpiIo& io_)
        : p(p_), c(c_), io(io_), sum(0.f), sum1(0.f), sum2(0.f), sum3(0.f) {
        const int live_end = p.dict_size ? min(p.dict_size[c.g], c.M) : c.M;
        live_row = c.row < live_end;
        const RowNorm rn = row_norm_of(p.fold_inv, c.g, c.M, c.row);
        cs = rn.is; ns = rn.ns;
        bcs = live_row ? p.bias[(long long)c.g * c.M + c.row] * cs : 0.f;
    }
    SC_DEVICE void prefetch(int) {}
    SC_DEVICE void chunk(int co, float (&v)[32], int /*next_co*/) {
        const int col = c.col0 + co;  // batch index of v[0]
        const int lane = threadIdx.x & 31;
        const int nb = c.N - col;     // c.N = B here
        // stored code c' = max(z, -0) * cs = max(z * cs, -0): -0.0 marks z < 0 (adding it to the L1 sum is a no-op);
        // outside fold mode cs = 1 and fma(acc, 1, b) is acc + b exactly
        if (live_row && nb >= 32) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
                v[j] = fmaxf(fmaf(v[j], cs, bcs), -0.0f);
                v[j + 1] = fmaxf(fmaf(v[j + 1], cs, bcs), -0.0f);
                v[j + 2] = fmaxf(fmaf(v[j + 2], cs, bcs), -0.0f);
                v[j + 3] = fmaxf(fmaf(v[j + 3], cs, bcs), -0.0f);
                sum += v[j]; sum1 += v[j + 1]; sum2 += v[j + 2]; sum3 += v[j + 3];
            }
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                v[j] = (live_row && j < nb) ? fmaxf(fmaf(v[j], cs, bcs), -0.0f) : -0.0f;  // masked features behave like z < 0
                sum += v[j];
            }
        }
        // the two staging tiles alternate per stored chunk (bit 8 of the warp's persistent phase word: a tile can have an
        // odd number of chunks per warp), so the one used now is the one whose store is two bulk groups back
        uint8_t* tile = io.buf + ((uniform_u32(*io.phase) >> 8) & 1u) * 2048;  // a warp-uniform address for the TMA store
        *io.phase ^= 1u << 8;
        io_wait_stores_read_n<1>();
        io_write_row_h(tile, lane, v);
        io_store_tile(&p.c_map, tile, col, c.row0, c.g);
        io_commit();
    }
    SC_DEVICE void finish() {
        if (p.code_scale && c.tile_n == 0 && c.half == 0 && c.row < c.M) p.code_scale[(long long)c.g * c.M + c.row] = ns;
        const float s = warp_sum(((sum + sum1) + (sum2 + sum3)) * ns);  // L1 term counts the unscaled codes
        if ((threadIdx.x & 31) == 0) {
            const int tiles = c.tiles_m * c.tiles_n;
            p.l1_part[((long long)c.g * tiles + c.tile_n * c.tiles_m + c.tile_m) * (4 * H) + c.part] = s;
        }
    }
};
using EpiEncodeTH = EpiEncodeTHT<2>;

// G3 (transposed: rows = feature n, cols = batch b): dz' from q' = r D^T and the fp16 codes.
// c^T and dz^T are [N][ldb] in memory, so the thread of an accumulator row (one feature, walking along the batch)
// touches them ROW-wise: 32 x 32 fp16 staging tiles moved by TMA (box = 32 batch columns x 32 features), each lane
// reading / writing its own 64-byte row with four 16-byte shared accesses.  Three 2 KB slots per warp rotate: a
// slot receives a c tile, dz overwrites it in place (a thread rewrites exactly the bytes it read), the tile is
// stored from there, and once that store has read it the slot takes the c tile two chunks ahead - across tile
// boundaries too (static schedule) - so loads fly during two chunks of arithmetic and 6 KB per warp is all the
// staging this epilogue needs (the resident-A kernel has little shared memory to spare).
template <int H, int SLOTS = 3, int W = 32>
struct EpiBackCodeHT {
    // H warps per row quarter (4 H epilogue warps): 256 / W / H chunks of W columns per warp and 256-wide tile.
    // SLOTS staging slots per warp: SLOTS - 1 chunks of look-ahead (3: what fits beside a resident A operand).
    // W = 64: chunks of 64 batch columns (128-byte rows, SWIZZLE_128B, 4 KB slots) - half the per-chunk bookkeeping.
    static_assert(W == 32 || W == 64, "chunk width");
    static constexpr int kHalves = H;
    static constexpr int kChunkW = W;
    static constexpr int kPerTile = 256 / W / H;
    static constexpr int kSlots = SLOTS;
    static constexpr int kAhead = SLOTS - 1;
    static constexpr int kSlotBytes = 32 * W * 2;
    static constexpr int kIoBytes = kSlots * kSlotBytes;  // per warp (kernels that count staging in 4 KB units round up)
    static constexpr int kIoBufs = (kIoBytes + 4095) / 4096;
    static_assert(kSlots <= kIoBars, "one load barrier per slot");
    struct alignas(64) Params {
        CUtensorMap c_map;    // fp16 io map of c^T  [G][N][B]  (box W batch columns x 32 features)
        CUtensorMap dz_map;   // fp16 io map of dz^T [G][N][B]  (scaled by B d / 2)
        const float* l1;      // [G]
        const float* bias;    // [G][N] (tied only)
        float* db_part;       // [G][parts][N]  (scaled)
        float* s_part;        // [G][parts][N]  (scaled)
        float l1_scale;       // d / 2:  lam/B * (B d / 2)
        int tied;
        int feats;
        const float* fold_inv; // fold mode (untied): [G][N] 1/max(||W_d row||,1e-8) (sign: clamp), else nullptr
    };
    const Params& p;
    const RowCtx c;
    const EpiIo io;
    float db, db1, s, s1, l1_term, bn, db_scale;
    // warp state that outlives the tile, in the free bits of the warp's phase word: bits 8-10 the slot of the next
    // chunk to consume, bits 12-14 how many chunks after the current one have been requested (0..kAhead)
    SC_DEVICE void request(int slot, int g, int row0, int bcol) const {
        io_expect(io, kSlotBytes, slot);
        tma_load_3d(&p.c_map, io.bar + slot, io.buf + slot * kSlotBytes, bcol, row0, g);
    }
    SC_DEVICE EpiBackCodeHT(const Params& p_, const RowCtx& c_, const EpiIo& io_)
        : p(p_), c(c_), io(io_), db(0.f), db1(0.f), s(0.f), s1(0.f) {
        // fold mode: the accumulator is r W_d^T (not normalised); dz is stored multiplied by ||W_d|| (see RowNorm)
        const RowNorm rn = row_norm_of(p.fold_inv, c.g, p.feats, c.row);
        l1_term = p.l1[c.g] * p.l1_scale * rn.ns;
        db_scale = rn.is;
        bn = (p.tied && c.row < c.M) ? p.bias[(long long)c.g * p.feats + c.row] : 0.f;
    }
    SC_DEVICE void prefetch(int) {}
    // chunk j of this warp in a tile that starts at batch column col0 exists (live chunks are a prefix in j)
    SC_DEVICE bool live(int col0, int j) const {
        const int cw = (c.half + j * H) * W;
        return j < kPerTile && cw < c.bn && col0 + cw < c.N;
    }
    // the k-th (1..kAhead) live chunk after chunk i of the current tile, in processing order
    SC_DEVICE bool nth_next(int i, int k, int& g, int& row0, int& bcol) const {
        for (int j = i + 1; live(c.col0, j); ++j)
            if (--k == 0) { g = c.g; row0 = c.row0; bcol = c.col0 + (c.half + j * H) * W; return true; }
        if (c.next_g < 0) return false;
        for (int j = 0; live(c.next_col0, j); ++j)
            if (--k == 0) { g = c.next_g; row0 = c.next_row0; bcol = c.next_col0 + (c.half + j * H) * W; return true; }
        return false;
    }
    template <bool TIED>
    SC_DEVICE void body(uint8_t* slot, const float (&v)[W]) {
        const int lane = threadIdx.x & 31;
        const uint32_t rs = smem_u32(slot) + lane * (2 * W);
        // SWIZZLE_64B (W = 32): 16-byte piece j of row r sits at piece j ^ ((r >> 1) & 3); SWIZZLE_128B (W = 64): j ^ (r & 7)
        const int sw = W == 32 ? (lane >> 1) & 3 : lane & 7;
#pragma unroll
        for (int jb = 0; jb < W / 8; jb += 4) {
            uint4 hc[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) hc[j] = lds_128u(rs + (((jb + j) ^ sw) << 4));  // all loads in flight first
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t w[4] = {hc[j].x, hc[j].y, hc[j].z, hc[j].w};
                uint32_t o[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float2 cf = __half22float2(*reinterpret_cast<const __half2*>(&w[t]));
                    const float q0 = v[8 * (jb + j) + 2 * t], q1 = v[8 * (jb + j) + 2 * t + 1];
                    // sign bit set <=> z < 0 (code -0);  > 0 <=> c > 0
                    const float t0 = cf.x > 0.f ? q0 + l1_term : q0;
                    const float t1 = cf.y > 0.f ? q1 + l1_term : q1;
                    const float d0 = __float_as_int(cf.x) < 0 ? 0.f : t0;
                    const float d1 = __float_as_int(cf.y) < 0 ? 0.f : t1;
                    const __half2 dh = half2_sat(d0, d1);
                    o[t] = *reinterpret_cast<const uint32_t*>(&dh);
                    db += d0; db1 += d1;
                    s = fmaf(cf.x, q0, s); s1 = fmaf(cf.y, q1, s1);
                    if (TIED) { s = fmaf(d0, cf.x - bn, s); s1 = fmaf(d1, cf.y - bn, s1); }
                }
                sts_128u(rs + (((jb + j) ^ sw) << 4), o[0], o[1], o[2], o[3]);  // dz over the codes just read
            }
        }
    }
    SC_DEVICE void chunk(int co, float (&v)[W], int /*next_co*/) {
        const int lane = threadIdx.x & 31;
        const int i = (co / W) / H;  // index of this chunk among the warp's chunks of the tile
        const int bcol = c.col0 + co;
        if (bcol >= c.N) return;      // the whole chunk lies beyond the batch (warp-uniform): nothing was requested
        const int row0 = c.row0;
        const uint32_t ph = uniform_u32(*io.phase);  // the warp's state word, as a value known to be warp-uniform
        int cons = (ph >> 8) & 7, ahead = (ph >> 12) & 7;
        if (ahead == 0) {  // nothing in flight for this chunk (first chunk of the kernel, or the look-ahead could not see it)
            io_wait_stores_read_n<1>();
            if (lane == 0) request(cons, c.g, row0, bcol);
            ahead = 1;
        }
        uint8_t* slot = io.buf + cons * kSlotBytes;
        io_wait_loads(io, cons);
        if (p.tied) body<true>(slot, v);
        else body<false>(slot, v);
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
            tma_store_3d(&p.dz_map, slot, bcol, row0, c.g);
            tma_store_commit();
        }
        --ahead;
        cons = cons == kSlots - 1 ? 0 : cons + 1;
        // keep kAhead chunks in flight: the slots they go to were stored at least one bulk group before the store above
        constexpr int kFullW = kPerTile * H * W;
        if (ahead == kAhead - 1 && c.bn == kFullW && c.col0 + kFullW <= c.N &&
            (i + kAhead < kPerTile || c.next_g < 0 || c.next_col0 + kFullW <= c.N)) {
            // steady state on full tiles (ncu: the general search below was a third of this epilogue's instructions):
            // exactly one request, the chunk kAhead places on - in this tile or at the start of the next one
            static_assert(kAhead <= kPerTile, "look-ahead reaches at most into the next tile");
            const int j = i + kAhead;
            const bool here = j < kPerTile;
            if (here || c.next_g >= 0) {
                io_wait_stores_read_n<1>();
                const int jj = here ? j : j - kPerTile;
                const int sl = cons + kAhead - 1 >= kSlots ? cons + kAhead - 1 - kSlots : cons + kAhead - 1;
                if (lane == 0)
                    request(sl, here ? c.g : c.next_g, here ? row0 : c.next_row0,
                            (here ? c.col0 : c.next_col0) + (c.half + jj * H) * W);
                ahead = kAhead;
            }
        } else {
            bool waited = false;
            for (int k = ahead + 1; k <= kAhead; ++k) {
                int g, r0, bc;
                if (!nth_next(i, k, g, r0, bc)) break;
                if (!waited) { io_wait_stores_read_n<1>(); waited = true; }
                if (lane == 0) request((cons + k - 1) % kSlots, g, r0, bc);
                ahead = k;
            }
        }
        *io.phase = (*io.phase & ~0x7700u) | (static_cast<uint32_t>(cons) << 8) | (static_cast<uint32_t>(ahead) << 12);
    }
    SC_DEVICE void finish() {
        if (c.row >= c.M) return;
        const long long o = ((long long)c.g * c.tiles_n * c.halves + c.tile_n * c.halves + c.half) * p.feats + c.row;
        p.db_part[o] = (db + db1) * db_scale;
        p.s_part[o] = s + s1;
    }
};
// Measured at the headline shape: 16 epilogue warps (H = 4, 96-register cap) are 1 - 1.5 % slower per step than 8, with
// the 12 KB-per-warp staging of round 1 as well as with the 6 KB rotating slots (0.488 / 0.491 vs 0.483 ms).
// Four slots (three chunks of look-ahead, streaming kernel) measured 0.483 - 0.486 ms per step against 0.478 - 0.480 with
// three slots beside a resident A operand.
using EpiBackCodeH = EpiBackCodeHT<2>;
using EpiBackCodeH64 = EpiBackCodeHT<2, 3, 64>;  // streaming kernel only (12 KB of staging per warp)

}  // namespace sc
