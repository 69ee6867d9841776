// Two-SM tcgen05 GEMM for large linear layers and convs: out[M, N] = act(A[M, K] . W[N, K]^T + bias)
// (+ res), the GEGLU form, or the implicit-GEMM 3x3 / temporal 3x1 conv (4-D TMA boxes, row bias, virtual concat) of
// gemm_tc.cu. A pair of CTAs on neighbouring SMs (cluster 2 x 1 x 1) owns a 256-row x BN tile:
//   * each CTA loads ITS 128 rows of A and ITS HALF of the weight tile (BN / 2 rows) — `tcgen05.mma.cta_group::2` reads
//     both halves, so per SM the shared-memory traffic per FLOP is what a 128 x 2BN one-SM tile would need;
//   * the leader CTA's elected lane issues the MMAs for both SMs; each CTA's accumulator (its 128 rows x BN columns)
//     lives in its own tensor memory, double-buffered, so the epilogue of tile i runs under the main loop of tile i + 1;
//   * the launch is M-persistent: one pair per two SMs walks the 256-row tiles of its N column.
// Why: one-SM SS-mode MMAs on 128 x 160 tiles are bound by shared-memory bandwidth (TMA writes + operand reads, and the
// epilogue behind them): 1.5-1.8x behind cuBLAS on ZeroScope's projections.
// The K order of every output element is the one-SM kernel's (64-deep chunks, four K=16 MMAs each).
// TMA loads here read this GPU's memory only.
#include <stdio.h>
#include <stdlib.h>

#include "../include/arbius_b200.h"
#include "common.cuh"
#include "gemm_common.cuh"

extern "C" void ab_count_launch(int n);
extern "C" int ab_make_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                                  const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

namespace ab {

constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;   // shared::cluster address of the same offset in the pair's even CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// TMA load issued by either CTA of the pair; the bytes are counted on the LEADER's mbarrier
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n" ::
            "r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tma_load_4d_2sm(void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], "
        "[%2];\n" ::"r"(smem_u32(dst)),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void umma_ss_2sm(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs of the pair once the MMAs issued so far have completed
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(
                     smem_u32(bar)),
                 "h"(static_cast<uint16_t>(3))
                 : "memory");
}
// arrive on the LEADER's copy of a barrier from either CTA
__device__ __forceinline__ void mbar_arrive_leader(uint64_t* bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(smem_u32(bar) & kPeerBitMask) : "memory");
}

constexpr int G2_THREADS = 320;

template <int BN, typename T, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(G2_THREADS, 1)
gemm2sm_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ CUtensorMap tmB, const __grid_constant__ GemmParams p) {
    constexpr int A_BYTES = BM * BK * 2;                 // this CTA's 128 rows
    constexpr int B_BYTES = (BN / 2) * BK * 2;           // this CTA's half of the weight tile
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr int ACC_COLS = 256;
    static_assert(BN % 32 == 0 && BN <= 256 && (BN / 2) % 8 == 0, "UMMA N for the pair; the epilogue reads 32-column chunks");
    static_assert(B_BYTES % 1024 == 0, "B stage must keep 1024-byte alignment");
    const int stages = p.stages;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    // output staging for the bulk tensor stores: two buffers of BN / 32 boxes (128 rows x 64 bytes, SWIZZLE_64B)
    constexpr int STG_BYTES = (BN / 32) * BM * 64;
    uint8_t* staging = smem + stages * STAGE_BYTES;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + (p.tma_store ? 2 * STG_BYTES : 0));
    uint64_t* empty_bar = full_bar + stages;
    uint64_t* acc_full = empty_bar + stages;     // [2]: MMA -> epilogue (both CTAs)
    uint64_t* acc_free = acc_full + 2;           // [2]: epilogues of both CTAs -> the leader's MMA lane
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_free + 2);
    float* sbias = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(tmem_slot) + 4 + 15) & ~uintptr_t(15));   // [BN]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const int pair0 = blockIdx.x >> 1;
    const int pairs = gridDim.x >> 1;
    const int n0 = blockIdx.y * BN;
    const int chunks_per_tap = p.k1_chunks + p.k2_chunks;
    const int kc_end = p.taps * chunks_per_tap;
    // this CTA's 128-row tile of pair tile tp is M tile 2 * tp + rank; conv modes: a bw x bh x bi pixel box of the images
    int w0 = 0, h0 = 0, i0 = 0;
    auto set_tile = [&](int tp) {
        if (p.conv) {
            const int tm = 2 * tp + static_cast<int>(rank);
            w0 = (tm % p.tiles_w) << p.bw_log2;
            h0 = ((tm / p.tiles_w) % p.tiles_h) << p.bh_log2;
            i0 = (tm / (p.tiles_w * p.tiles_h)) << (7 - p.bw_log2 - p.bh_log2);
        }
    };

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        if (p.k2_chunks) tma_prefetch_desc(&tmA2);
        for (int s = 0; s < stages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_free[b], 2);
        }
        fence_barrier_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();              // the peer's barriers exist before anything is signalled across the pair
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        // ================= TMA producer (both CTAs, one elected lane) =================
        if (elect_one()) {
            int stage = 0;
            uint32_t phase = 0;
            for (int tp = pair0; tp < p.tiles_x; tp += pairs) {
                const int m_row = tp * 2 * BM + static_cast<int>(rank) * BM;
                set_tile(tp);
                int tap = 0, cc = 0;
                for (int kc = 0; kc < kc_end; ++kc) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem + stage * STAGE_BYTES;
                    if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * STAGE_BYTES);   // both CTAs' bytes land on this barrier
                    const bool second = cc >= p.k1_chunks;
                    const CUtensorMap* ma = second ? &tmA2 : &tmA;
                    const int c0 = (second ? cc - p.k1_chunks : cc) * BK;
                    if (p.conv) {
                        // conv == 1: 3x3 taps; conv == 2: 3x1 taps along H only (temporal conv on [B, F, HW, C])
                        const int dh = p.conv == 2 ? tap - 1 : tap / 3 - 1;
                        const int dw = p.conv == 2 ? 0 : tap % 3 - 1;
                        tma_load_4d_2sm(sa, ma, &full_bar[stage], c0, w0 + dw, h0 + dh, i0);
                    } else {
                        tma_load_2d_2sm(sa, ma, &full_bar[stage], c0, m_row);
                    }
                    tma_load_2d_2sm(sa + A_BYTES, &tmB, &full_bar[stage], kc * BK, n0 + static_cast<int>(rank) * (BN / 2));
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                    if (++cc == chunks_per_tap) { cc = 0; ++tap; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ================= MMA issuer: the leader CTA's elected lane drives both SMs =================
        if (rank == 0 && elect_one()) {
            constexpr uint32_t idesc = make_idesc_f16(2 * BM, BN, DT<T>::kBf16, false, false);
            const uint64_t desc0 = make_smem_desc(0, 16, 1024, kSwizzle128);
            const uint32_t smem_base = smem_u32(smem);
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int tp = pair0; tp < p.tiles_x; tp += pairs, ++it) {
                const int b = it & 1;
                if (it >= 2) {                        // both epilogues have read this buffer's previous tile
                    mbar_wait(&acc_free[b], ((it >> 1) - 1) & 1);
                    tc_fence_after();
                }
                const uint32_t acc = tmem_base + b * ACC_COLS;
                for (int kc = 0; kc < kc_end; ++kc) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_base + stage * STAGE_BYTES;
                    const uint32_t sb = sa + A_BYTES;
#pragma unroll
                    for (int k = 0; k < BK / UK; ++k)
                        umma_ss_2sm(acc, smem_desc_at(desc0, sa + k * UK * 2), smem_desc_at(desc0, sb + k * UK * 2), idesc,
                                    (kc > 0 || k > 0) ? 1u : 0u);
                    umma_commit_2sm(&empty_bar[stage]);    // both producers may refill this stage
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
                umma_commit_2sm(&acc_full[b]);             // both epilogues may read this buffer
            }
        }
        __syncwarp();
    } else {
        // ================= epilogue warps (both CTAs): this CTA's 128 rows x BN columns =================
        const int q = warp & 3;
        const int hf = (warp - 2) >> 2;
        const int r = q * 32 + lane;
        {
            const T* bias = reinterpret_cast<const T*>(p.bias);
            for (int n = threadIdx.x - 64; n < BN; n += G2_THREADS - 64)
                sbias[n] = (bias && n0 + n < p.N) ? DT<T>::to_f(bias[n0 + n]) : 0.f;
            asm volatile("bar.sync 1, 256;\n" ::: "memory");
        }
        int it = 0;
        for (int tp = pair0; tp < p.tiles_x; tp += pairs, ++it) {
            const int b = it & 1;
            mbar_wait(&acc_full[b], (it >> 1) & 1);
            tc_fence_after();
            const long long row0 = static_cast<long long>(tp) * 2 * BM + static_cast<long long>(rank) * BM;
            long long out_row = row0 + r;
            bool row_ok = out_row < p.M;
            int grp = 0;
            if (p.conv) {
                set_tile(tp);
                const int ww = w0 + (r & ((1 << p.bw_log2) - 1));
                const int hh = h0 + ((r >> p.bw_log2) & ((1 << p.bh_log2) - 1));
                const int ii = i0 + (r >> (p.bw_log2 + p.bh_log2));
                row_ok = (ww < p.W) && (hh < p.H) && (ii < p.nimg);
                out_row = (static_cast<long long>(ii) * p.H + hh) * p.W + ww;
                grp = ii;
            }
            // the row bias (time embedding of the tile's image): staged with the bias when the tile lies in one image — the
            // same fp32 order as the one-SM kernel (bias + row bias first, then added to the accumulator) — else per row
            const bool restage = p.conv && p.rowbias && (p.bw_log2 + p.bh_log2) == 7;
            const bool rb_global = p.rowbias != nullptr && !restage;
            if (restage) {
                const T* bias = reinterpret_cast<const T*>(p.bias);
                const T* rowbias = reinterpret_cast<const T*>(p.rowbias);
                for (int n = threadIdx.x - 64; n < BN; n += G2_THREADS - 64) {
                    float bv = 0.f;
                    if (n0 + n < p.N) {
                        if (bias) bv = DT<T>::to_f(bias[n0 + n]);
                        if (i0 < p.nimg) bv += DT<T>::to_f(rowbias[static_cast<long long>(i0) * p.ldrb + n0 + n]);
                    }
                    sbias[n] = bv;
                }
            }
            uint8_t* const stg = staging + (it & 1) * STG_BYTES;
            if (p.tma_store && it >= 2 && warp == 2) {
                // this staging buffer was last used two tiles ago: its bulk stores must have finished READING it
                if (elect_one()) asm volatile("cp.async.bulk.wait_group.read 1;\n" ::: "memory");
                __syncwarp();
            }
            if (restage || (p.tma_store && it >= 2)) asm volatile("bar.sync 1, 256;\n" ::: "memory");
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + b * ACC_COLS;
            constexpr int NCHUNK = BN / 32;
            uint32_t raw[2][32];
            if (hf < NCHUNK) tmem_ld32(taddr + hf * 32, raw[0]);
#pragma unroll
            for (int j = 0; j < (NCHUNK + 1) / 2; ++j) {
                const int c = 2 * j + hf;
                if (c >= NCHUNK) break;
                tmem_wait_ld();
                if (c + 2 < NCHUNK) tmem_ld32(taddr + (c + 2) * 32, raw[(j + 1) & 1]);
                const int n_base = n0 + c * 32;
                if (row_ok && n_base < p.N) {
                    float v[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[j & 1][i]);
                    if (p.tma_store && EPI == EPI_GEGLU)      // chunks 2k and 2k+1 fill the halves of output box k
                        epilogue_chunk32<T, EPI>(p, out_row, grp, n_base, v, sbias + c * 32, rb_global,
                                                 stg + (c >> 1) * (BM * 64) + r * 64, (r >> 1) & 3, (c & 1) * 2);
                    else if (p.tma_store)
                        epilogue_chunk32<T, EPI>(p, out_row, grp, n_base, v, sbias + c * 32, rb_global, stg + c * (BM * 64) + r * 64,
                                                 (r >> 1) & 3);
                    else
                        epilogue_chunk32<T, EPI>(p, out_row, grp, n_base, v, sbias + c * 32, rb_global);
                }
            }
            // the accumulator buffer goes back to the leader's MMA lane: one arrival per CTA
            if (p.tma_store) fence_proxy_async();
            tc_fence_before();
            asm volatile("bar.sync 1, 256;\n" ::: "memory");
            if (threadIdx.x == 64) mbar_arrive_leader(&acc_free[b]);
            if (p.tma_store && warp == 2) {
                if (elect_one()) {
                    if (EPI == EPI_GEGLU) {
                        for (int c = 0; c < NCHUNK / 2; ++c) {
                            if (n0 + c * 64 >= p.N) break;
                            tma_store_2d(&p.omap, stg + c * (BM * 64), (n0 >> 1) + c * 32, static_cast<int>(row0));
                        }
                    } else {
                        for (int c = 0; c < NCHUNK; ++c) {
                            if (n0 + c * 32 >= p.N) break;
                            if (p.conv) tma_store_4d(&p.omap, stg + c * (BM * 64), n0 + c * 32, w0, h0, i0);
                            else tma_store_2d(&p.omap, stg + c * (BM * 64), n0 + c * 32, static_cast<int>(row0));
                        }
                    }
                    tma_store_commit();
                }
                __syncwarp();
            }
        }
    }

    if (p.tma_store && warp == 2) {
        if (elect_one()) tma_store_wait_read();      // the staging buffers outlive the bulk reads
        __syncwarp();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();              // neither CTA frees tensor memory the pair's MMAs may still write
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, 512;\n" ::"r"(tmem_base) : "memory");
}

template <int BN, typename T, int EPI>
static int launch_gemm2sm(const ab_gemm_args* a, GemmParams& p, cudaStream_t stream) {
    constexpr int kStage = BM * BK * 2 + (BN / 2) * BK * 2;
    static const bool tma_store_on = [] { const char* e = getenv("AB_TMA_STORE"); return !(e && e[0] == '0'); }();
    // (GEGLU writes 16 output columns per accumulator chunk: whole 32-column boxes only when BN / 64 is an integer)
    p.tma_store = (tma_store_on && (EPI != EPI_GEGLU || BN % 64 == 0) && (reinterpret_cast<uintptr_t>(a->out) & 15) == 0) ? 1 : 0;
    const int stg = p.tma_store ? 2 * (BN / 32) * BM * 64 : 0;
    int stages = 2;
    while (stages < 12 && (stages + 1) * kStage + stg + (2 * (stages + 1) + 4) * 8 + 32 + BN * 4 + 1024 <= 227 * 1024) ++stages;
    const int smem = stages * kStage + stg + (2 * stages + 4) * 8 + 32 + BN * 4 + 1024;
    const int esz = 2;
    CUtensorMap tmA, tmA2, tmB;
    int tiles_m;
    if (a->conv) {
        const ConvTile ct = pick_conv_tile(a->H, a->W);
        p.bw_log2 = ilog2(ct.bw); p.bh_log2 = ilog2(ct.bh);
        p.tiles_w = (a->W + ct.bw - 1) / ct.bw;
        p.tiles_h = (a->H + ct.bh - 1) / ct.bh;
        tiles_m = p.tiles_w * p.tiles_h * ((a->nimg + ct.bi - 1) / ct.bi);
        uint64_t dims[4] = {(uint64_t)a->K1, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->nimg};
        uint64_t str[3] = {(uint64_t)a->K1 * esz, (uint64_t)a->W * a->K1 * esz, (uint64_t)a->H * a->W * a->K1 * esz};
        uint32_t box[4] = {BK, (uint32_t)ct.bw, (uint32_t)ct.bh, (uint32_t)ct.bi};
        if (ab_make_tensor_map(&tmA, a->dtype, 4, a->a, dims, str, box, 128)) return -1;
        tmA2 = tmA;
        if (a->K2 > 0) {
            uint64_t dims2[4] = {(uint64_t)a->K2, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->nimg};
            uint64_t str2[3] = {(uint64_t)a->K2 * esz, (uint64_t)a->W * a->K2 * esz, (uint64_t)a->H * a->W * a->K2 * esz};
            if (ab_make_tensor_map(&tmA2, a->dtype, 4, a->a2, dims2, str2, box, 128)) return -1;
        }
        if (p.tma_store) {
            const uint64_t ld = (uint64_t)p.ldo * esz;
            uint64_t odims[4] = {(uint64_t)a->N, (uint64_t)a->W, (uint64_t)a->H, (uint64_t)a->nimg};
            uint64_t ostr[3] = {ld, (uint64_t)a->W * ld, (uint64_t)a->H * a->W * ld};
            uint32_t obox[4] = {32, (uint32_t)ct.bw, (uint32_t)ct.bh, (uint32_t)ct.bi};
            if (ab_make_tensor_map(&p.omap, a->dtype, 4, a->out, odims, ostr, obox, 64)) return -1;
        }
    } else {
        tiles_m = (a->M + BM - 1) / BM;
        uint64_t dims[2] = {(uint64_t)a->K1, (uint64_t)a->M};
        uint64_t str[1] = {(uint64_t)a->K1 * esz};
        uint32_t box[2] = {BK, BM};
        if (ab_make_tensor_map(&tmA, a->dtype, 2, a->a, dims, str, box, 128)) return -1;
        tmA2 = tmA;
        if (p.tma_store) {
            uint64_t odims[2] = {(uint64_t)p.Nout, (uint64_t)a->M};
            uint64_t ostr[1] = {(uint64_t)p.ldo * esz};
            uint32_t obox[2] = {32, BM};
            if (ab_make_tensor_map(&p.omap, a->dtype, 2, a->out, odims, ostr, obox, 64)) return -1;
        }
    }
    {
        const uint64_t Kw = (uint64_t)p.taps * (a->K1 + a->K2);
        uint64_t dimsw[2] = {Kw, (uint64_t)a->N};
        uint64_t strw[1] = {Kw * esz};
        uint32_t boxw[2] = {BK, BN / 2};
        if (ab_make_tensor_map(&tmB, a->dtype, 2, a->w, dimsw, strw, boxw, 128)) return -1;
    }
    p.stages = stages;
    p.tiles_x = (tiles_m + 1) / 2;
    const int n_tiles = (a->N + BN - 1) / BN;
    int pairs = (ab_sm_count() / 2) / n_tiles;
    if (pairs < 1) pairs = 1;
    if (pairs > p.tiles_x) pairs = p.tiles_x;
    if (ab_ensure_max_smem(reinterpret_cast<const void*>(gemm2sm_kernel<BN, T, EPI>), smem)) return -1;
    dim3 grid(2 * pairs, n_tiles, 1);
    AB_CHECK_CUDA(launch_pdl(gemm2sm_kernel<BN, T, EPI>, grid, dim3(G2_THREADS), smem, stream, tmA, tmA2, tmB, p));
    ab_count_launch(1);
    return 0;
}

template <int BN, typename T>
static int dispatch_epi2sm(const ab_gemm_args* a, GemmParams& p, cudaStream_t stream) {
    // the activation is a template parameter where it is common: the runtime switch of EPI_ANY costs more than the GEMM
    if (a->geglu) return launch_gemm2sm<BN, T, EPI_GEGLU>(a, p, stream);
    switch (a->act) {
        case ACT_NONE: return launch_gemm2sm<BN, T, EPI_NONE>(a, p, stream);
        case ACT_SILU: return launch_gemm2sm<BN, T, EPI_SILU>(a, p, stream);
        case ACT_GELU: return launch_gemm2sm<BN, T, EPI_GELU>(a, p, stream);
        default: return launch_gemm2sm<BN, T, EPI_ANY>(a, p, stream);
    }
}

template <typename T>
static int dispatch_gemm2sm(const ab_gemm_args* a, int bn, GemmParams& p, cudaStream_t stream) {
    if (bn == 256) return dispatch_epi2sm<256, T>(a, p, stream);
    return dispatch_epi2sm<160, T>(a, p, stream);
}

}  // namespace ab

// Called by ab_gemm for the shapes this kernel serves; returns 1 if it did not take the call (the one-SM kernel runs).
// mode: 1 = on (default; AB_GEMM_2SM=0 in the environment turns it off), 0 = off; min_rows: smallest M it takes
// (default 12288: at 8192 rows the one-SM kernel's 128-row tiles fill the GPU better, 8.5 vs 9.6 us; at 12288 the pair tiles win, 18.4 vs 23.3). Both kernels return the same bytes,
// so this only moves time; the setter exists for tests and A/B timing.
static int g2_mode = -1, g2_min_rows_v = -1;
extern "C" void ab_gemm_set_2sm(int mode, int min_rows) { g2_mode = mode; g2_min_rows_v = min_rows; }
static int g2_enabled() {
    static const int env = [] { const char* e = getenv("AB_GEMM_2SM"); return e ? atoi(e) : 1; }();
    return g2_mode >= 0 ? g2_mode : env;
}
static int g2_min_rows() {
    static const int env = [] { const char* e = getenv("AB_GEMM_2SM_MIN_ROWS"); return e ? atoi(e) : 12288; }();
    return g2_min_rows_v >= 0 ? g2_min_rows_v : env;
}
extern "C" int ab_gemm_2sm_try(const ab_gemm_args* a, void* stream_) {
    using namespace ab;
    if (!g2_enabled()) return 1;
    // anything malformed goes to the one-SM entry point, which owns the argument checks and their error texts
    if (!a->a || !a->w || !a->out || (a->K2 > 0 && !a->a2) || a->conv < 0 || a->conv > 2 || a->M <= 0 || a->N <= 0 || a->K1 <= 0 ||
        (a->dtype != AB_F16 && a->dtype != AB_BF16))
        return 1;
    if (a->splitk > 1 || a->sync_wait || a->halo || a->resample) return 1;
    if (!a->conv && (a->K2 > 0 || a->rowbias)) return 1;          // (linear concat / grouped row bias: one-SM kernel)
    if (a->conv && (a->geglu || a->K2 % BK != 0 || a->M != a->nimg * a->H * a->W)) return 1;
    // convs need more rows to fill the 74 pairs evenly: at 32768 rows (four SD tasks at 64 x 64) the one-SM kernel's 128-row
    // tiles still win, at ZeroScope's 110k-442k rows the pair tiles do (12.91 -> 12.69 s per task)
    const int min_rows = (a->conv && g2_min_rows_v < 0) ? (g2_min_rows() > 65536 ? g2_min_rows() : 65536) : g2_min_rows();
    if (a->K1 % BK != 0 || a->N % 32 != 0 || a->M < min_rows) return 1;
    if (a->geglu && a->N % 64 != 0) return 1;
    // the tile width is free (it does not change bytes): 256 for wide outputs and for GEGLU (whose bulk stores need whole
    // 32-column output boxes), 160 where it divides N
    int bn;
    if (a->N % 256 == 0 && (a->geglu || a->N >= 2048)) bn = 256;
    else if (a->N % 160 == 0) bn = 160;
    else if (a->N % 256 == 0) bn = 256;
    else bn = 160;
    GemmParams p{};
    p.out = a->out; p.bias = a->bias; p.residual = a->residual; p.rowbias = a->rowbias;
    p.M = a->M; p.N = a->N; p.Nout = a->geglu ? a->N / 2 : a->N;
    p.k1_chunks = a->K1 / BK; p.k2_chunks = a->K2 / BK;
    p.taps = a->conv == 2 ? 3 : (a->conv ? 9 : 1);
    p.conv = a->conv; p.nimg = a->nimg; p.H = a->H; p.W = a->W;
    p.act = a->act; p.geglu = a->geglu;
    p.out_scale = a->out_scale == 0.f ? 1.f : a->out_scale;
    p.ldo = a->ldo > 0 ? a->ldo : p.Nout;
    p.ldr = a->ldr > 0 ? a->ldr : p.Nout;
    p.ldrb = a->ldrb > 0 ? a->ldrb : a->N;
    p.splitk = 1;
    if (p.ldo % 8 != 0 || p.ldr % 8 != 0 || p.ldrb % 8 != 0) return 1;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    const int rc = a->dtype == AB_BF16 ? dispatch_gemm2sm<__nv_bfloat16>(a, bn, p, stream) : dispatch_gemm2sm<__half>(a, bn, p, stream);
    return rc == 0 ? 0 : -1;
}
