// tcgen05 GEMM / implicit-GEMM 3x3 conv for sm_100a.
//
// One kernel serves every GEMM-shaped layer of the diffusion UNets, text encoders and VAEs:
//   * linear / 1x1 conv:  A [M, K] row-major, loaded by a 2-D TMA tensor map
//   * 3x3 conv (pad 1):   A is never materialised; each K step is one (tap, 64-channel) slab of
//                         the NHWC image fetched by a 4-D TMA box whose start coordinate is shifted
//                         by the tap offset — TMA's out-of-bounds zero fill IS the padding.
//   * virtual concat:     a second A source appended along K (skip connections never get copied)
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + single-thread MMA issuer,
// warps 2..9 = epilogue (tcgen05.ld -> bias/time-emb/activation/GEGLU/residual -> global), two warps per
// TMEM lane quarter on alternate 32-column chunks with the next chunk's tcgen05.ld in flight.
// Accumulators live in TMEM (128 lanes x BN fp32 columns). Two CTAs fit per SM (<=113 KB smem,
// <=256 TMEM columns each) so one CTA's epilogue overlaps the other's main loop.
// Split-K is a fixed, caller-chosen factor with an in-order second pass: results never depend on
// scheduling, grid size or GPU count.
#include <stdio.h>
#include <stdlib.h>

#include <mutex>
#include <unordered_map>

#include "../include/arbius_b200.h"
#include "common.cuh"
#include "gemm_common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

// MT = 2: one CTA owns two M-adjacent 128-row tiles (two accumulators in TMEM) that share every weight tile it
// loads. Shared-memory ingest per MMA cycle drops from (A + B) / t to (2A + B) / 2t: the 128 x 256 tile needs
// ~94 B/clk at tensor peak against the ~64 B/clk an SM can take from L2 (680 cycles per K step measured, 512 ideal),
// the 256 x 256 super-tile needs 62. Used when the problem still fills the GPU for at least two waves.
template <int BN, int STAGES, typename T, int EPI, int MT>
__global__ void __launch_bounds__(GEMM_THREADS, ((BN > 160 || MT > 1) ? 1 : 2))
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ GemmParams p) {
    constexpr int A1_BYTES = BM * BK * 2;
    constexpr int A_BYTES = MT * A1_BYTES;
    constexpr int B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr int ACC_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
    constexpr int TMEM_COLS = MT * ACC_COLS;
    static_assert(BN % 32 == 0 && BN >= 32 && BN <= 256, "UMMA N for M=128; epilogue reads 32-column chunks");
    static_assert(B_BYTES % 1024 == 0, "B stage must keep 1024-byte alignment");
    static_assert(MT * (BN / 32) * BM * 64 <= STAGES * STAGE_BYTES, "output staging boxes reuse the pipeline stages");
    // STAGES is the smallest depth this instantiation runs with (what fits twice on an SM); p.stages >= STAGES is the
    // depth of THIS launch: a grid of at most one CTA per SM takes all the shared memory, because its K loop is paced
    // by (TMA round trip) / depth, ~500 cycles per 64-deep step at depth 3-4 whatever the tile width.
    const int stages = p.stages;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + stages * STAGE_BYTES);
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* acc_bar = empty_bar + stages;
    uint64_t* acc_free = acc_bar + 1;             // epilogue -> MMA: the accumulator has been read (persistent launches)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_free + 1);
    float* sbias = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(tmem_slot) + 4 + 15) & ~uintptr_t(15));  // [MT][BN]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const long long t_start = p.dbg ? clock64() : 0;
    long long* dbg = p.dbg ? p.dbg + 8LL * (blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) : nullptr;

    // ---- tile coordinates -------------------------------------------------------------------
    const int n0 = blockIdx.y * BN;
    // sub-tile s of tile group tb is M tile tb * MT + s; a CTA starts with group blockIdx.x (and, in a persistent launch,
    // goes on with blockIdx.x + gridDim.x, ...: every role calls set_tile at the top of each of its iterations)
    int tile_ms[MT], i0s[MT], h0s[MT], w0s[MT], m0s[MT], pys[MT], pxs[MT];
    auto set_tile = [&](int tb) {
#pragma unroll
    for (int s = 0; s < MT; ++s) {
        int tm = tb * MT + s;
        tile_ms[s] = tm;
        i0s[s] = h0s[s] = w0s[s] = m0s[s] = pys[s] = pxs[s] = 0;
        if (p.resample == 2) {            // four parity classes per low-resolution tile
            pys[s] = (tm >> 1) & 1;
            pxs[s] = tm & 1;
            tm >>= 2;
        }
        if (p.conv) {
            const int tw = tm % p.tiles_w;
            const int th = (tm / p.tiles_w) % p.tiles_h;
            const int ti = tm / (p.tiles_w * p.tiles_h);
            w0s[s] = tw << p.bw_log2;
            h0s[s] = th << p.bh_log2;
            i0s[s] = ti << (7 - p.bw_log2 - p.bh_log2);
        } else {
            m0s[s] = tm * BM;
        }
    }
    };
    set_tile(blockIdx.x);
    const int chunks_per_tap = p.k1_chunks + p.k2_chunks;
    const int total_chunks = p.taps * chunks_per_tap;
    const int kc_begin = blockIdx.z * p.chunks_per_split;
    const int kc_end = min(total_chunks, kc_begin + p.chunks_per_split);

    // ---- one-time setup ---------------------------------------------------------------------
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if (p.k2_chunks) tma_prefetch_desc(&tmA2);
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        mbar_init(acc_bar, 1);
        mbar_init(acc_free, GEMM_THREADS - 64);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // everything above overlapped the previous kernel's tail; operands/outputs are only touched past this point
    pdl_wait();
    pdl_launch_dependents();
    if (dbg && threadIdx.x == 32) dbg[0] = clock64() - t_start;   // setup done

    if (warp == 0) {
        // ================= TMA producer =================
        // ONE lane chosen by elect.sync runs the whole loop. The compiler then knows the region is single-lane and keeps
        // addresses, coordinates and descriptors in uniform registers. Under `if (lane == 0)` it cannot, and moves every
        // UTMALDG / UTCHMMA operand out of a "divergent" context with an ELECT / R2UR / BRA.U.ANY loop (~100 cycles per
        // instruction); a warp-uniform loop with `if (leader)` inside pays ~230 cycles per iteration for the branch and
        // the reconvergence (scripts/micro/gemm_loop.cu: 470-500 -> 369 cycles per K step).
        if (elect_one()) {
            if (p.sync_wait) {
                // only a tile that touches a halo row has to wait for the neighbour that fills it: interior tiles start
                // at once and hide the hand-off's NVLink latency
                unsigned int need = p.sync_wait_mask;
                if (p.conv && p.halo) {
                    need = 0;
#pragma unroll
                    for (int s = 0; s < MT; ++s) {
                        if (h0s[s] == 0) need |= p.sync_wait_mask & p.sync_top_mask;
                        if (h0s[s] + (1 << p.bh_log2) >= p.H) need |= p.sync_wait_mask & ~p.sync_top_mask;
                    }
                }
                if (need) {
                    sync_wait_flags(p.sync_wait, need, *p.sync_tick);
                    fence_proxy_async_all();  // the peers' plain stores (acquired above) before this thread's TMA reads
                }
            }
            int stage = 0;
            uint32_t phase = 0;
            bool vacant = mbar_test_wait(&empty_bar[0], 1);
            for (int tb = blockIdx.x; tb < p.tiles_x; tb += gridDim.x) {
            set_tile(tb);
            // (tap, channel chunk) advance by counting: no division in the loop
            int tap = kc_begin / chunks_per_tap;
            int cc = kc_begin - tap * chunks_per_tap;
            for (int kc = kc_begin; kc < kc_end; ++kc) {
                if (!vacant) mbar_wait(&empty_bar[stage], phase ^ 1);
                // look at the next stage's barrier now: its round trip overlaps this stage's loads
                int nstage = stage + 1;
                uint32_t nphase = phase;
                if (nstage == stages) { nstage = 0; nphase ^= 1; }
                vacant = (kc + 1 < kc_end || tb + gridDim.x < p.tiles_x) && mbar_test_wait(&empty_bar[nstage], nphase ^ 1);
                uint8_t* sa = smem + stage * STAGE_BYTES;
                uint8_t* sb = sa + A_BYTES;
                mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
                const bool second = cc >= p.k1_chunks;
                const CUtensorMap* ma = second ? &tmA2 : &tmA;
                const int c0 = (second ? cc - p.k1_chunks : cc) * BK;
#pragma unroll
                for (int s = 0; s < MT; ++s) {
                    uint8_t* sas = sa + s * A1_BYTES;
                    if (p.conv) {
                        // conv == 1: 3x3 taps; conv == 2: 3x1 taps along H only (temporal Conv3d (3,1,1) on [B, F, HW, C])
                        const int dh = p.conv == 2 ? tap - 1 : tap / 3 - 1;
                        const int dw = p.conv == 2 ? 0 : tap % 3 - 1;
                        if (p.resample == 0) {
                            tma_load_4d(sas, ma, &full_bar[stage], c0, w0s[s] + dw, h0s[s] + dh + p.halo, i0s[s]);
                        } else if (p.resample == 1) {
                            // stride 2: the map steps two pixels per element along W and H
                            tma_load_4d(sas, ma, &full_bar[stage], c0, 2 * w0s[s] + dw + 1 - p.pad_lo,
                                        2 * h0s[s] + dh + 1 - p.pad_lo, i0s[s]);
                        } else {
                            // nearest 2x upsample: upsampled pixel 2y + py + dh is low-resolution pixel y + ((py + dh) >> 1)
                            tma_load_4d(sas, ma, &full_bar[stage], c0, w0s[s] + ((pxs[s] + dw) >> 1),
                                        h0s[s] + ((pys[s] + dh) >> 1), i0s[s]);
                        }
                    } else {
                        if (p.nsrc > 1) {
                            const int src = tile_ms[s] / p.tiles_per_src;
                            tma_load_2d(sas, &p.amaps[src], &full_bar[stage], c0, (tile_ms[s] - src * p.tiles_per_src) * BM);
                        } else {
                            tma_load_2d(sas, ma, &full_bar[stage], c0, m0s[s]);
                        }
                    }
                }
                tma_load_2d(sb, &tmB, &full_bar[stage], kc * BK, n0);
                stage = nstage;
                phase = nphase;
                if (++cc == chunks_per_tap) { cc = 0; ++tap; }
            }
            }   // tiles of this CTA
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================= MMA issuer (one elected lane runs the loop, see the producer) =================
        if (elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(BM, BN, DT<T>::kBf16, false, false);
            const uint64_t desc0 = make_smem_desc(0, 16, 1024, kSwizzle128);
            const uint32_t smem_base = smem_u32(smem);
            int stage = 0;
            uint32_t phase = 0;
            bool ready = mbar_test_wait(&full_bar[0], 0);
            int it = 0;
            for (int tb = blockIdx.x; tb < p.tiles_x; tb += gridDim.x, ++it) {
            if (it > 0) {                         // the epilogue warps have read the previous tile out of tensor memory
                mbar_wait(acc_free, (it - 1) & 1);
                tc_fence_after();
            }
            for (int kc = kc_begin; kc < kc_end; ++kc) {
                if (!ready) mbar_wait(&full_bar[stage], phase);
                tc_fence_after();
                if (dbg && kc == kc_begin && it == 0) dbg[1] = clock64() - t_start;   // first operands landed
                const uint32_t sa = smem_base + stage * STAGE_BYTES;
                const uint32_t sb = sa + A_BYTES;
                // look at the next stage's barrier now: its round trip overlaps the issue of this stage's MMAs
                int nstage = stage + 1;
                uint32_t nphase = phase;
                if (nstage == stages) { nstage = 0; nphase ^= 1; }
                ready = (kc + 1 < kc_end || tb + gridDim.x < p.tiles_x) && mbar_test_wait(&full_bar[nstage], nphase);
#pragma unroll
                for (int k = 0; k < BK / UK; ++k) {
                    const uint64_t bdesc = smem_desc_at(desc0, sb + k * UK * 2);
#pragma unroll
                    for (int s = 0; s < MT; ++s) {
                        const uint64_t adesc = smem_desc_at(desc0, sa + s * A1_BYTES + k * UK * 2);
                        umma_ss(tmem_base + s * ACC_COLS, adesc, bdesc, idesc, (kc > kc_begin || k > 0) ? 1u : 0u);
                    }
                }
                umma_commit(&empty_bar[stage]);   // frees this smem stage once the MMAs have read it
                stage = nstage;
                phase = nphase;
            }
            umma_commit(acc_bar);                 // accumulator complete
            if (dbg && it == 0) dbg[2] = clock64() - t_start;          // all MMAs issued
            }   // tiles of this CTA
        }
        __syncwarp();
    } else {
        // ================= epilogue warps =================
        const int q = warp & 3;                   // TMEM lane quarter this warp may read
        const int hf = (warp - 2) >> 2;           // which half of the 32-column chunks this warp handles
        const int r = q * 32 + lane;              // row inside the tile
        int it = 0;
        for (int tb = blockIdx.x; tb < p.tiles_x; tb += gridDim.x, ++it) {
        set_tile(tb);
        // While the main loop runs, stage bias (+ row bias if a whole sub-tile shares one group) in smem.
        bool single_group[MT];
        const bool use_sb = p.splitk <= 1;
#pragma unroll
        for (int s = 0; s < MT; ++s) {
            int tile_grp;
            if (p.conv) { single_group[s] = (p.bw_log2 + p.bh_log2) == 7; tile_grp = i0s[s]; }
            else if (p.rows_per_group > 0) {
                const long long last = min(static_cast<long long>(m0s[s]) + BM - 1, static_cast<long long>(p.M) - 1);
                tile_grp = m0s[s] / p.rows_per_group;
                single_group[s] = (last / p.rows_per_group) == tile_grp;
            } else { single_group[s] = true; tile_grp = 0; }
            // (a persistent CTA keeps its N column: without a row bias the staged values do not change between tiles)
            if (use_sb && (it == 0 || p.rowbias)) {
                const T* bias = reinterpret_cast<const T*>(p.bias);
                const T* rowbias = reinterpret_cast<const T*>(p.rowbias);
                for (int n = threadIdx.x - 64; n < BN; n += GEMM_THREADS - 64) {
                    float b = 0.f;
                    if (n0 + n < p.N) {
                        if (bias) b = DT<T>::to_f(bias[n0 + n]);
                        if (rowbias && single_group[s])
                            b += DT<T>::to_f(rowbias[static_cast<long long>(tile_grp) * p.ldrb + n0 + n]);
                    }
                    sbias[s * BN + n] = b;
                }
            }
        }
        if (use_sb && (it == 0 || p.rowbias)) asm volatile("bar.sync 1, 256;\n" ::: "memory");
        mbar_wait(acc_bar, it & 1);
        tc_fence_after();
        if (dbg && threadIdx.x == 64 && it == 0) dbg[3] = clock64() - t_start;   // accumulator ready
#pragma unroll
        for (int s = 0; s < MT; ++s) {
        const int w0 = w0s[s], h0 = h0s[s], i0 = i0s[s], m0 = m0s[s];
        const bool rowbias_global = !single_group[s];
        float* const sbias_s = sbias + s * BN;
        uint8_t* const stage_base = smem + s * ((BN / 32) * BM * 64);
        long long out_row;
        int grp;
        bool row_ok;
        if (p.conv) {
            const int wl = r & ((1 << p.bw_log2) - 1);
            const int hl = (r >> p.bw_log2) & ((1 << p.bh_log2) - 1);
            const int il = r >> (p.bw_log2 + p.bh_log2);
            const int ww = w0 + wl, hh = h0 + hl, ii = i0 + il;
            row_ok = (ww < p.W) && (hh < p.H) && (ii < p.nimg);
            out_row = (static_cast<long long>(ii) * p.H + hh) * p.W + ww;
            if (p.resample == 2)   // this tile's parity class of the [nimg, 2H, 2W] output
                out_row = (static_cast<long long>(ii) * (2 * p.H) + 2 * hh + pys[s]) * (2 * p.W) + 2 * ww + pxs[s];
            grp = ii;
        } else {
            out_row = m0 + r;
            row_ok = out_row < p.M;
            grp = p.rows_per_group > 0 ? static_cast<int>(out_row / p.rows_per_group) : 0;
        }
        const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + s * ACC_COLS;
        // Two warps share each TMEM lane quarter and take alternate 32-column chunks; the tcgen05.ld of a warp's
        // next chunk is in flight while it converts and stores the current one.
        constexpr int NCHUNK = BN / 32;
        constexpr int MY_MAX = (NCHUNK + 1) / 2;
        // (the prefetch needs a second 32-register buffer: only where one CTA per SM leaves the registers free)
        constexpr bool PF = BN > 160;
        uint32_t raw[PF ? 2 : 1][32];
        if (PF && hf < NCHUNK) tmem_ld32(taddr + hf * 32, raw[0]);
#pragma unroll
        for (int j = 0; j < MY_MAX; ++j) {
            const int c = 2 * j + hf;
            if (c >= NCHUNK) break;
            if (!PF) tmem_ld32(taddr + c * 32, raw[0]);
            tmem_wait_ld();
            if (dbg && threadIdx.x == 64 && it == 0 && j == 0 && s == 0) dbg[6] = clock64() - t_start;   // first chunk in registers
            if (PF && j + 1 < MY_MAX && c + 2 < NCHUNK) tmem_ld32(taddr + (c + 2) * 32, raw[PF ? ((j + 1) & 1) : 0]);
            const int n_base = n0 + c * 32;
            if (p.tma_store) {
                if (n_base >= p.N) continue;                           // warp-uniform: the whole box is out of range
                uint8_t* box = stage_base + c * (BM * 64);
                if (row_ok) {
                    float v[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[PF ? (j & 1) : 0][i]);
                    epilogue_chunk32<T, EPI>(p, out_row, grp, n_base, v, sbias_s + c * 32, rowbias_global, box + r * 64,
                                             (r >> 1) & 3);
                }
                if (dbg && threadIdx.x == 64 && it == 0 && j == 0 && s == 0) dbg[7] = clock64() - t_start;   // first chunk staged
                continue;
            }
            if (row_ok && n_base < p.N) {
                float v[32];
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[PF ? (j & 1) : 0][i]);
                if (p.splitk > 1) {
                    const int owner = static_cast<int>(out_row / p.rows_per_owner);
                    float* ws = p.ws_peer[owner] + (static_cast<long long>(p.z_base + blockIdx.z) * p.rows_per_owner +
                                                    (out_row - static_cast<long long>(owner) * p.rows_per_owner)) * p.N + n_base;
#pragma unroll
                    for (int g = 0; g < 8; ++g)
                        if (n_base + 4 * g < p.N)
                            *reinterpret_cast<float4*>(ws + 4 * g) = make_float4(v[4 * g], v[4 * g + 1], v[4 * g + 2], v[4 * g + 3]);
                } else {
                    epilogue_chunk32<T, EPI>(p, out_row, grp, n_base, v, sbias_s + c * 32, rowbias_global);
                }
            }
            if (dbg && threadIdx.x == 64 && it == 0 && j == 0 && s == 0) dbg[7] = clock64() - t_start;   // first chunk stored
        }
        if (p.tma_store) {
            // every chunk has its own box, so the proxy fence and the barrier are paid once per tile, not once per chunk
            // (they were ~1 us of each chunk's 1.14 us), and one lane issues the tile's stores back to back
            fence_proxy_async();
            asm volatile("bar.sync %0, 128;\n" ::"r"(2 + hf) : "memory");   // the four warps of this chunk parity
            if (q == 0) {
                if (elect_one()) {
                    for (int c = hf; c < NCHUNK; c += 2) {
                        const int n_base = n0 + c * 32;
                        if (n_base >= p.N) break;
                        uint8_t* box = stage_base + c * (BM * 64);
                        if (p.conv) tma_store_4d(&p.omap, box, n_base, w0, h0, i0);
                        else tma_store_2d(&p.omap, box, n_base, m0);
                    }
                    tma_store_commit();
                }
                __syncwarp();
            }
        }
        }   // sub-tiles
        if (tb + gridDim.x < p.tiles_x) {
            // persistent launch: hand the accumulator back to the MMA lane (every epilogue thread's tcgen05.ld has
            // completed: each chunk ends with tcgen05.wait::ld), and keep sbias until everyone has used it
            tc_fence_before();
            mbar_arrive(acc_free);
            if (p.rowbias) asm volatile("bar.sync 1, 256;\n" ::: "memory");   // sbias is rewritten for the next tile
        }
        }   // tiles of this CTA
        if (p.tma_store && q == 0) {
            if (elect_one()) tma_store_wait_read();   // boxes must outlive the bulk reads (same lane: elect.sync with the
            __syncwarp();                             // full mask always names the same leader)
        }
    }

    if (dbg && threadIdx.x == 64) dbg[4] = clock64() - t_start;       // epilogue done
    // ---- teardown -----------------------------------------------------------------------------
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
    if (dbg && threadIdx.x == 32) dbg[5] = clock64() - t_start;       // TMEM released
}

// Second pass of split-K: sum the fp32 partial tiles in split order, then the normal epilogue.
template <typename T, int EPI>
__global__ void splitk_reduce_kernel(const GemmParams p) {
    pdl_wait();
    pdl_launch_dependents();
    const int chunks_n = (p.N + 31) / 32;
    const long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= static_cast<long long>(p.M) * chunks_n) return;
    const long long row = idx / chunks_n;
    const int n_base = static_cast<int>(idx % chunks_n) * 32;
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0.f;
    for (int z = 0; z < p.splitk; ++z) {
        const float* ws = p.workspace + (static_cast<long long>(z) * p.M + row) * p.N + n_base;
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            if (n_base + 4 * g < p.N) {
                float4 t = *reinterpret_cast<const float4*>(ws + 4 * g);
                v[4 * g] += t.x; v[4 * g + 1] += t.y; v[4 * g + 2] += t.z; v[4 * g + 3] += t.w;
            }
        }
    }
    const int grp = p.conv ? static_cast<int>(row / (static_cast<long long>(p.H) * p.W))
                           : (p.rows_per_group > 0 ? static_cast<int>(row / p.rows_per_group) : 0);
    epilogue_chunk32<T, EPI>(p, row, grp, n_base, v, nullptr, true);
}

// ---------------------------------------------------------------------------------------------
// host side
// This file is compiled twice (build.sh compiles the .cu files in parallel): as is, it holds the host entry points
// and the fp16 kernels; gemm_tc_bf16.cu includes it with AB_GEMM_BF16_TU defined and contributes only the bf16
// main-kernel instantiations behind ab_gemm_main_bf16() — ~110 large kernels are too slow for one nvcc process.
// ---------------------------------------------------------------------------------------------
#ifndef AB_GEMM_BF16_TU
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    });
    return fn;
}

}  // namespace ab

// Exposed for the other translation units (attention uses the same encoder).
extern "C" int ab_make_tensor_map_strided(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                                          const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes,
                                          const uint32_t* elem_strides);
extern "C" int ab_make_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                                  const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
    return ab_make_tensor_map_strided(out, dtype, rank, base, dims, strides_bytes, box, swizzle_bytes, nullptr);
}
// elem_strides (or NULL = all 1): traversal stride per dimension; a box of `box[i]` tensor elements then delivers
// ceil(box[i] / elem_strides[i]) of them (the strided conv reads every second pixel this way).
extern "C" int ab_make_tensor_map_strided(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                                          const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes,
                                          const uint32_t* elem_strides) {
    auto fn = ab::get_encode_fn();
    if (!fn) { ab_set_error("cuTensorMapEncodeTiled entry point not available"); return -1; }
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    if (elem_strides) for (int i = 0; i < rank && i < 5; ++i) estr[i] = elem_strides[i];
    CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
    CUresult r = fn(out, dtype == AB_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16,
                    static_cast<cuuint32_t>(rank), const_cast<void*>(base), reinterpret_cast<const cuuint64_t*>(dims),
                    reinterpret_cast<const cuuint64_t*>(strides_bytes), reinterpret_cast<const cuuint32_t*>(box), estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[256];
        snprintf(buf, sizeof(buf), "cuTensorMapEncodeTiled failed: CUresult %d (rank %d dims %llu,%llu box %u,%u)",
                 static_cast<int>(r), rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                 box[0], rank > 1 ? box[1] : 0);
        ab_set_error(buf);
        return -1;
    }
    return 0;
}
namespace ab {
#endif  // !AB_GEMM_BF16_TU
}  // namespace ab

namespace ab {

static int auto_block_n(int M_tiles, int N, bool geglu) {
    // Prefer the tile width that fills 148 SMs x 2 resident CTAs with the least idle tail.
    static const int cands[] = {256, 160, 128, 96, 64, 32};
    int best = 64; double best_cost = 1e30;
    for (int bn : cands) {
        if (geglu && bn % 32) continue;
        int tiles = M_tiles * ((N + bn - 1) / bn);
        int slots = 148 * (bn > 160 ? 1 : 2);
        int waves = (tiles + slots - 1) / slots;
        // cost ~ waves x per-tile time (proportional to bn, plus a fixed overhead) incl. padded columns
        double cost = waves * (bn + 48.0) * (bn > 160 ? 0.55 : 1.0);
        if (cost < best_cost) { best_cost = cost; best = bn; }
    }
    return best;
}

constexpr int gemm_smem_bytes(int stages, int stage_bytes, int bias_floats) {
    return stages * stage_bytes + (2 * stages + 1) * 8 + 32 + bias_floats * 4 + 1024;
}
constexpr int gemm_deepest(int at_least, int stage_bytes, int bias_floats) {
    int st = at_least;
    while (st < 12 && gemm_smem_bytes(st + 1, stage_bytes, bias_floats) <= 227 * 1024) ++st;
    return st;
}

template <int BN, int STAGES, typename T, int EPI, int MT>
static int launch_gemm_mt(const CUtensorMap& tmA, const CUtensorMap& tmA2, const CUtensorMap& tmB, const GemmParams& p,
                          dim3 grid, cudaStream_t stream) {
    constexpr int kStage = MT * BM * BK * 2 + BN * BK * 2;
    static_assert(gemm_smem_bytes(STAGES, kStage, MT * BN) <= 227 * 1024, "stage count does not fit shared memory");
    constexpr int kDeep = gemm_deepest(STAGES, kStage, MT * BN);   // deepest pipeline one CTA per SM can hold
    auto smem_for = [](int st) { return gemm_smem_bytes(st, kStage, MT * BN); };
    static const int deep_mode = [] { const char* e = getenv("AB_GEMM_DEEP"); return e ? atoi(e) : 1; }();
    const long long ctas = static_cast<long long>(grid.x) * grid.y * grid.z;
    GemmParams q = p;
    q.stages = (deep_mode && ctas <= ab_sm_count()) ? kDeep : STAGES;
    const int smem = smem_for(q.stages);
    if (ab_ensure_max_smem(reinterpret_cast<const void*>(gemm_tc_kernel<BN, STAGES, T, EPI, MT>), smem_for(kDeep))) return -1;
    AB_CHECK_CUDA(launch_pdl(gemm_tc_kernel<BN, STAGES, T, EPI, MT>, grid, dim3(GEMM_THREADS), smem, stream, tmA, tmA2, tmB, q));
    ab_count_launch(1);
    return 0;
}

// STAGES is the depth for one 128-row tile per CTA; the two-tile variant exists for BN = 256 only (three 64 KB stages).
// At BN <= 160 two co-resident CTAs already overlap each other's epilogue and loads: the ZeroScope convs measured
// 1137-1236 TFLOP/s with one tile per CTA and 1121-1238 with two, so those widths keep one.
template <int BN, int STAGES, typename T, int EPI>
static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmA2, const CUtensorMap& tmB, const GemmParams& p,
                       dim3 grid, cudaStream_t stream) {
    if constexpr (BN == 256) {
        if (p.mt == 2) return launch_gemm_mt<BN, 3, T, EPI, 2>(tmA, tmA2, tmB, p, grid, stream);
    }
    return launch_gemm_mt<BN, STAGES, T, EPI, 1>(tmA, tmA2, tmB, p, grid, stream);
}

template <typename T, int EPI>
static int dispatch_bn(int bn, const CUtensorMap& tmA, const CUtensorMap& tmA2, const CUtensorMap& tmB,
                       const GemmParams& p, dim3 grid, cudaStream_t stream) {
    switch (bn) {
        case 32:  return launch_gemm<32, 4, T, EPI>(tmA, tmA2, tmB, p, grid, stream);
        case 64:  return launch_gemm<64, 4, T, EPI>(tmA, tmA2, tmB, p, grid, stream);
        case 96:  return launch_gemm<96, 3, T, EPI>(tmA, tmA2, tmB, p, grid, stream);
        case 128: return launch_gemm<128, 3, T, EPI>(tmA, tmA2, tmB, p, grid, stream);
        case 160: return launch_gemm<160, 3, T, EPI>(tmA, tmA2, tmB, p, grid, stream);
        case 256: return launch_gemm<256, 4, T, EPI>(tmA, tmA2, tmB, p, grid, stream);
        default: ab_set_error("unsupported block_n"); return -2;
    }
}

template <typename T>
static int gemm_main_t(int main_epi, int bn, const CUtensorMap& tmA, const CUtensorMap& tmA2, const CUtensorMap& tmB,
                       const GemmParams& p, dim3 grid, cudaStream_t stream) {
    switch (main_epi) {
        case EPI_NONE: return dispatch_bn<T, EPI_NONE>(bn, tmA, tmA2, tmB, p, grid, stream);
        case EPI_SILU: return dispatch_bn<T, EPI_SILU>(bn, tmA, tmA2, tmB, p, grid, stream);
        case EPI_GELU: return dispatch_bn<T, EPI_GELU>(bn, tmA, tmA2, tmB, p, grid, stream);
        case EPI_QGELU: return dispatch_bn<T, EPI_QGELU>(bn, tmA, tmA2, tmB, p, grid, stream);
        case EPI_GEGLU: return dispatch_bn<T, EPI_GEGLU>(bn, tmA, tmA2, tmB, p, grid, stream);
        default: return dispatch_bn<T, EPI_ANY>(bn, tmA, tmA2, tmB, p, grid, stream);
    }
}

}  // namespace ab

#ifdef AB_GEMM_BF16_TU
extern "C" int ab_gemm_main_bf16(int main_epi, int bn, const CUtensorMap* tmA, const CUtensorMap* tmA2, const CUtensorMap* tmB,
                                 const void* params, unsigned gx, unsigned gy, unsigned gz, void* stream) {
    return ab::gemm_main_t<__nv_bfloat16>(main_epi, bn, *tmA, *tmA2, *tmB, *reinterpret_cast<const ab::GemmParams*>(params),
                                          dim3(gx, gy, gz), reinterpret_cast<cudaStream_t>(stream));
}
#else
extern "C" int ab_gemm_main_bf16(int main_epi, int bn, const CUtensorMap* tmA, const CUtensorMap* tmA2, const CUtensorMap* tmB,
                                 const void* params, unsigned gx, unsigned gy, unsigned gz, void* stream);

static long long* g_gemm_dbg = nullptr;
// Profiling aid: point at a device buffer of >= 8 * grid_size int64 to receive per-CTA phase cycle counts.
extern "C" void ab_gemm_set_debug(long long* buf) { g_gemm_dbg = buf; }

extern "C" size_t ab_gemm_workspace_bytes(const ab_gemm_args* a) {
    if (!a || a->splitk <= 1) return 0;
    return static_cast<size_t>(a->splitk) * a->M * a->N * sizeof(float);
}

extern "C" int ab_gemm_2sm_try(const ab_gemm_args* a, void* stream_);   // gemm_2sm.cu: 0 = done, 1 = not its shape, -1 = error
extern "C" int ab_gemm(const ab_gemm_args* a, void* stream_) {
    if (a) {
        const int rc = ab_gemm_2sm_try(a, stream_);
        if (rc <= 0) return rc;
    }
    return ab_gemm_dist(a, nullptr, 0, stream_);
}

extern "C" int ab_gemm_dist(const ab_gemm_args* a, const ab_gemm_dist_args* d, int phase, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const bool ag = d && d->nsrc > 1;
    const bool rs = d && d->rs_world >= 1;
    AB_REQUIRE(a && (a->a || ag) && a->w && a->out, "null pointer");
    AB_REQUIRE(a->M > 0 && a->N > 0 && a->K1 > 0, "bad shape");
    AB_REQUIRE(a->N % 8 == 0, "N must be a multiple of 8");
    const int Ktot = a->K1 + a->K2;
    AB_REQUIRE(Ktot % 8 == 0, "K must be a multiple of 8 (16-byte rows for TMA)");
    if (a->K2 > 0) AB_REQUIRE(a->K1 % BK == 0 && a->a2, "virtual concat needs K1 % 64 == 0 and a2");
    if (a->conv) AB_REQUIRE(a->K1 % BK == 0 && a->K2 % BK == 0, "conv needs channel counts that are multiples of 64");
    if (a->conv) AB_REQUIRE(a->M == (a->resample == 2 ? 4 : 1) * a->nimg * a->H * a->W, "conv: M != nimg*H*W (x4 with the 2x upsample)");
    if (a->geglu) AB_REQUIRE(a->N % 32 == 0, "geglu needs N % 32 == 0");

    GemmParams p{};
    p.out = a->out; p.bias = a->bias; p.rowbias = a->rowbias; p.residual = a->residual;
    p.workspace = reinterpret_cast<float*>(a->workspace);
    p.dbg = g_gemm_dbg;
    p.sync_tick = a->sync_tick; p.sync_wait = a->sync_wait; p.sync_wait_mask = a->sync_wait_mask;
    p.sync_top_mask = a->sync_top_mask;
    AB_REQUIRE(!a->sync_wait || a->sync_tick, "sync_wait needs sync_tick");
    p.M = a->M; p.N = a->N; p.Nout = a->geglu ? a->N / 2 : a->N;
    p.taps = a->conv == 2 ? 3 : (a->conv ? 9 : 1);
    p.k1_chunks = (a->K2 > 0 || a->conv) ? a->K1 / BK : (a->K1 + BK - 1) / BK;
    p.k2_chunks = a->K2 > 0 ? (a->conv ? a->K2 / BK : (a->K2 + BK - 1) / BK) : 0;
    p.conv = a->conv; p.nimg = a->nimg; p.H = a->H; p.W = a->W; p.halo = a->halo;
    p.resample = a->resample; p.pad_lo = a->pad_lo;
    AB_REQUIRE(a->resample >= 0 && a->resample <= 2, "resample: 0, 1 (stride 2) or 2 (nearest 2x upsample)");
    if (a->resample) {
        AB_REQUIRE(a->conv == 1 && a->halo == 0 && a->K2 == 0 && !a->sync_wait && !d, "resample: plain 3x3 conv only");
        AB_REQUIRE(a->pad_lo == 0 || a->pad_lo == 1, "pad_lo is 0 or 1");
        if (a->resample == 1)
            AB_REQUIRE(a->Hin > 0 && a->Win > 0 && 2 * (a->H - 1) + 2 - a->pad_lo < a->Hin + 2 &&
                       2 * (a->W - 1) + 2 - a->pad_lo < a->Win + 2, "stride 2: input size does not match the output size");
    }
    AB_REQUIRE(a->halo == 0 || (a->halo == 1 && a->conv != 0), "halo rows (0 or 1) exist in the conv modes only");
    p.rows_per_group = a->conv ? a->H * a->W : a->rows_per_group;
    p.act = a->act; p.geglu = a->geglu;
    p.out_scale = a->out_scale == 0.f ? 1.f : a->out_scale;
    p.ldo = a->ldo > 0 ? a->ldo : p.Nout;
    p.ldr = a->ldr > 0 ? a->ldr : p.Nout;
    p.ldrb = a->ldrb > 0 ? a->ldrb : a->N;
    AB_REQUIRE(p.ldo % 8 == 0 && p.ldr % 8 == 0 && p.ldrb % 8 == 0, "row strides must be multiples of 8 elements");

    int tiles_m;
    ConvTile ct{128, 1, 1};
    if (a->conv) {
        ct = pick_conv_tile(a->H, a->W);
        p.bw_log2 = ilog2(ct.bw); p.bh_log2 = ilog2(ct.bh);
        p.tiles_w = (a->W + ct.bw - 1) / ct.bw;
        p.tiles_h = (a->H + ct.bh - 1) / ct.bh;
        tiles_m = p.tiles_w * p.tiles_h * ((a->nimg + ct.bi - 1) / ct.bi);
        if (a->resample == 2) tiles_m *= 4;   // one tile per parity class of each low-resolution box
    } else {
        tiles_m = (a->M + BM - 1) / BM;
    }
    const int total_chunks = p.taps * (p.k1_chunks + p.k2_chunks);
    int splitk = a->splitk > 1 ? a->splitk : 1;
    if (splitk > total_chunks) splitk = total_chunks;
    p.chunks_per_split = (total_chunks + splitk - 1) / splitk;
    splitk = (total_chunks + p.chunks_per_split - 1) / p.chunks_per_split;   // no empty splits
    p.splitk = splitk;
    int grid_z = splitk;
    p.rows_per_owner = a->M; p.z_base = 0; p.ws_peer[0] = p.workspace;
    if (rs) {
        // canonical 8 K shards; this rank computes 8 / rs_world of them over its local K slice
        AB_REQUIRE(!a->conv && (d->rs_world == 1 || d->rs_world == 2 || d->rs_world == 4 || d->rs_world == 8),
                   "reduce-scatter mode: linear GEMM, 1/2/4/8 ranks");
        AB_REQUIRE(a->M % d->rs_world == 0, "reduce-scatter: M must divide evenly over the ranks");
        grid_z = 8 / d->rs_world;
        // shard boundaries must be the same K offsets for every rank count: global K % 512 == 0
        AB_REQUIRE(total_chunks % grid_z == 0 && Ktot % BK == 0,
                   "reduce-scatter: global K must be a multiple of 512 (8 canonical shards of whole 64-wide chunks)");
        p.chunks_per_split = total_chunks / grid_z;
        p.splitk = splitk = 8;
        p.rows_per_owner = a->M / d->rs_world;
        p.z_base = d->rs_rank * grid_z;
        for (int r = 0; r < d->rs_world; ++r) {
            AB_REQUIRE(d->ws_peers[r], "reduce-scatter: missing workspace pointer");
            p.ws_peer[r] = reinterpret_cast<float*>(d->ws_peers[r]);
        }
        p.workspace = p.ws_peer[d->rs_rank];
    }
    if (splitk > 1) AB_REQUIRE(p.workspace, "split-K needs a workspace");
    if (ag) {
        AB_REQUIRE(!a->conv && a->K2 == 0, "all-gather mode: linear GEMM without virtual concat");
        AB_REQUIRE(d->nsrc == 2 || d->nsrc == 4 || d->nsrc == 8, "all-gather mode: 2/4/8 sources");
        AB_REQUIRE(d->rows_per_src % BM == 0 && a->M == d->nsrc * d->rows_per_src, "all-gather mode: rows per source must be a multiple of 128");
        p.nsrc = d->nsrc;
        p.tiles_per_src = d->rows_per_src / BM;
        for (int s = 0; s < d->nsrc; ++s) {
            uint64_t dims[2] = {(uint64_t)a->K1, (uint64_t)d->rows_per_src};
            uint64_t str[1] = {(uint64_t)a->K1 * 2};
            uint32_t box[2] = {BK, BM};
            if (ab_make_tensor_map(&p.amaps[s], a->dtype, 2, d->a_srcs[s], dims, str, box, 128)) return -1;
        }
    }

    int bn = a->block_n > 0 ? a->block_n : auto_block_n(tiles_m * splitk, a->N, a->geglu != 0);

    // ---- tensor maps ----
    CUtensorMap tmA, tmA2, tmB;
    const int esz = 2;
    if (a->conv) {
        const uint64_t Hin = (uint64_t)a->H + 2 * (uint64_t)a->halo;   // halo rows are real data, not the zero fill
        uint64_t dims[4] = {(uint64_t)a->K1, (uint64_t)a->W, Hin, (uint64_t)a->nimg};
        uint64_t str[3] = {(uint64_t)a->K1 * esz, (uint64_t)a->W * a->K1 * esz, Hin * a->W * a->K1 * esz};
        uint32_t box[4] = {BK, (uint32_t)ct.bw, (uint32_t)ct.bh, (uint32_t)ct.bi};
        if (a->resample == 1) {
            // stride 2: the map is over the INPUT image and steps two pixels per delivered element along W and H
            uint64_t dims_s[4] = {(uint64_t)a->K1, (uint64_t)a->Win, (uint64_t)a->Hin, (uint64_t)a->nimg};
            uint64_t str_s[3] = {(uint64_t)a->K1 * esz, (uint64_t)a->Win * a->K1 * esz, (uint64_t)a->Hin * a->Win * a->K1 * esz};
            uint32_t box_s[4] = {BK, 2u * ct.bw, 2u * ct.bh, (uint32_t)ct.bi};
            uint32_t es[4] = {1, 2, 2, 1};
            if (ab_make_tensor_map_strided(&tmA, a->dtype, 4, a->a, dims_s, str_s, box_s, 128, es)) return -1;
        } else if (ab_make_tensor_map(&tmA, a->dtype, 4, a->a, dims, str, box, 128)) return -1;
        tmA2 = tmA;
        if (a->K2 > 0) {
            uint64_t dims2[4] = {(uint64_t)a->K2, (uint64_t)a->W, Hin, (uint64_t)a->nimg};
            uint64_t str2[3] = {(uint64_t)a->K2 * esz, (uint64_t)a->W * a->K2 * esz, Hin * a->W * a->K2 * esz};
            if (ab_make_tensor_map(&tmA2, a->dtype, 4, a->a2, dims2, str2, box, 128)) return -1;
        }
    } else {
        uint64_t dims[2] = {(uint64_t)a->K1, (uint64_t)a->M};
        uint64_t str[1] = {(uint64_t)a->K1 * esz};
        uint32_t box[2] = {BK, BM};
        if (ag) tmA = p.amaps[0];
        else if (ab_make_tensor_map(&tmA, a->dtype, 2, a->a, dims, str, box, 128)) return -1;
        tmA2 = tmA;
        if (a->K2 > 0) {
            uint64_t dims2[2] = {(uint64_t)a->K2, (uint64_t)a->M};
            uint64_t str2[1] = {(uint64_t)a->K2 * esz};
            if (ab_make_tensor_map(&tmA2, a->dtype, 2, a->a2, dims2, str2, box, 128)) return -1;
        }
    }
    {
        const uint64_t Kw = (uint64_t)p.taps * Ktot;
        uint64_t dims[2] = {Kw, (uint64_t)a->N};
        uint64_t str[1] = {Kw * esz};
        uint32_t box[2] = {BK, (uint32_t)bn};
        if (ab_make_tensor_map(&tmB, a->dtype, 2, a->w, dims, str, box, 128)) return -1;
    }

    static const bool tma_store_on = [] { const char* e = getenv("AB_TMA_STORE"); return !(e && e[0] == '0'); }();
    p.tma_store = 0;
    if (tma_store_on && !a->geglu && splitk <= 1 && !rs && a->resample != 2 && (reinterpret_cast<uintptr_t>(a->out) & 15) == 0) {
        const uint64_t ld = (uint64_t)p.ldo * esz;
        if (a->conv) {
            uint64_t dims[4] = {(uint64_t)a->N, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->nimg};
            uint64_t str[3] = {ld, (uint64_t)a->W * ld, (uint64_t)a->H * a->W * ld};
            uint32_t box[4] = {32, (uint32_t)ct.bw, (uint32_t)ct.bh, (uint32_t)ct.bi};
            if (ab_make_tensor_map(&p.omap, a->dtype, 4, a->out, dims, str, box, 64)) return -1;
        } else {
            uint64_t dims[2] = {(uint64_t)a->N, (uint64_t)a->M};
            uint64_t str[1] = {ld};
            uint32_t box[2] = {32, BM};
            if (ab_make_tensor_map(&p.omap, a->dtype, 2, a->out, dims, str, box, 64)) return -1;
        }
        p.tma_store = 1;
    }

    static const bool mt_on = [] { const char* e = getenv("AB_GEMM_MT2"); return !(e && e[0] == '0'); }();
    const int n_tiles = (a->N + bn - 1) / bn;
    // two tiles per CTA: only for long main loops (the doubled epilogue is serial behind them: 512x512x128->128 with
    // 18 K steps got 20 % slower) on problems that still give every SM two waves
    p.mt = (mt_on && !ag && bn == 256 && tiles_m % 2 == 0 && p.chunks_per_split >= 24 &&
            static_cast<long long>(tiles_m / 2) * n_tiles * grid_z >= 2LL * ab_sm_count()) ? 2 : 1;
    p.tiles_x = tiles_m / p.mt;
    int grid_x = p.tiles_x;
    {
        // M-persistent launch for short K loops on grids of several waves (GemmParams.tiles_x): as many CTAs as are
        // resident at once, each walking the M tiles of its N column. The TMA-store staging boxes live in the pipeline
        // stages, which a persistent CTA is already refilling, so these launches store from registers.
        static const int persist_mode = [] { const char* e = getenv("AB_GEMM_PERSIST"); return e ? atoi(e) : 1; }();
        const int per_sm = (bn <= 160 && p.mt == 1) ? 2 : 1;
        const long long slots = static_cast<long long>(ab_sm_count()) * per_sm;
        const int gx = static_cast<int>(slots / n_tiles);
        if (persist_mode && splitk <= 1 && !rs && !ag && !a->sync_wait && p.chunks_per_split <= 16 && gx >= 1 &&
            p.tiles_x >= 3 * gx) {
            grid_x = gx;
            p.tma_store = 0;
        }
    }
    dim3 grid(grid_x, n_tiles, grid_z);
    int epi = a->geglu ? EPI_GEGLU
            : a->act == ACT_NONE ? EPI_NONE
            : a->act == ACT_SILU ? EPI_SILU
            : a->act == ACT_GELU ? EPI_GELU
            : a->act == ACT_QUICK_GELU ? EPI_QGELU : EPI_ANY;
    const int main_epi = splitk > 1 ? EPI_NONE : epi;     // with split-K the main kernel only stores fp32 partials
    int rc = 0;
    if (phase != 2) {
        if (a->dtype == AB_BF16) rc = ab_gemm_main_bf16(main_epi, bn, &tmA, &tmA2, &tmB, &p, grid.x, grid.y, grid.z, stream_);
        else rc = gemm_main_t<__half>(main_epi, bn, tmA, tmA2, tmB, p, grid, stream);
    }
    if (rc) return rc;
    if (splitk > 1 && phase != 1) {
        if (rs) { p.M = p.rows_per_owner; p.splitk = 8; }      // owners reduce their own row band
        // one thread per (row, 32-column chunk); 64-thread blocks keep many CTAs in flight for small M*N
        const long long work = static_cast<long long>(p.M) * ((p.N + 31) / 32);
        const int threads = 64;
        const int blocks = static_cast<int>((work + threads - 1) / threads);
#define AB_RED_EPI(T)                                                                                      \
    switch (epi) {                                                                                         \
        case EPI_NONE: launch_pdl(splitk_reduce_kernel<T, EPI_NONE>, dim3(blocks), dim3(threads), 0, stream, p); break;        \
        case EPI_SILU: launch_pdl(splitk_reduce_kernel<T, EPI_SILU>, dim3(blocks), dim3(threads), 0, stream, p); break;        \
        case EPI_GELU: launch_pdl(splitk_reduce_kernel<T, EPI_GELU>, dim3(blocks), dim3(threads), 0, stream, p); break;        \
        case EPI_QGELU: launch_pdl(splitk_reduce_kernel<T, EPI_QGELU>, dim3(blocks), dim3(threads), 0, stream, p); break;      \
        case EPI_GEGLU: launch_pdl(splitk_reduce_kernel<T, EPI_GEGLU>, dim3(blocks), dim3(threads), 0, stream, p); break;      \
        default: launch_pdl(splitk_reduce_kernel<T, EPI_ANY>, dim3(blocks), dim3(threads), 0, stream, p); break;               \
    }
        if (a->dtype == AB_BF16) { AB_RED_EPI(__nv_bfloat16) } else { AB_RED_EPI(__half) }
#undef AB_RED_EPI
        AB_CHECK_CUDA(cudaGetLastError());
        ab_count_launch(1);
    }
    return 0;
}

#endif  // !AB_GEMM_BF16_TU
