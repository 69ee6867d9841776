// Temporal self-attention of the video UNets, tcgen05/TMEM.
//
// Shape: B*H*W independent sequences of only F (16..32) frames. One CTA packs PX = 128 / F pixels into a single
// 128-row tile whose rows are ordered (frame, pixel); S = Q K^T is a 128x128 MMA and the softmax applies a
// block-diagonal mask (row and column must belong to the same pixel). One KV tile => no online-softmax loop.
//
// The tile is assembled by TMA boxes of shape (D, 1 head, PX pixels, F_s frames): with the frame axis sharded
// over `nsrc` GPUs (sequence parallelism: spatial layers own frames, temporal layers need all frames of a
// pixel), source s simply contributes rows [s*F_s*PX, (s+1)*F_s*PX) of the tile through its own (peer-mapped)
// tensor map, and the output rows are scattered back to the buffer of the GPU that owns each frame. The
// frame<->pixel all-to-all therefore happens inside the tile loads/stores of the attention kernel itself.
#include <math.h>
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);
extern "C" int ab_make_tensor_map(CUtensorMap* out, int dtype, int rank, const void* base, const uint64_t* dims,
                                  const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

namespace ab {

constexpr int TA_THREADS = 160;
constexpr int TA_TILE_BYTES = 128 * 128;

struct TaMaps { CUtensorMap q[8], k[8], v[8]; };

__device__ __forceinline__ float ex2_approx_ta(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

struct TaParams {
    void* out[8];            // per-source output buffers [B, F_s, HW, heads*D] (row stride o_ls elements)
    int batch, heads, frames, frames_per_src, nsrc, hw, px, head_dim, pix_begin, pix_end;
    long long o_ls;
    float scale_log2;
};

// PM (pixel-major rows): row = pixel * F + frame, so a row's valid keys are the contiguous columns of its own pixel
// (two compares per score instead of an integer modulo) — needs F_s % 8 == 0 so every per-pixel TMA box starts on a
// 1024-byte swizzle atom. Otherwise rows stay frame-major (one box per source) with the modulo mask.
// Shared memory: P aliases Q|K (both are dead once S is complete) and O aliases S's first columns (S is dead once P
// is written): 48 KB + 128 TMEM columns per CTA, four CTAs per SM hide each other's load -> MMA -> softmax chain.
template <typename T, bool PM>
__global__ void __launch_bounds__(TA_THREADS, 4)
temporal_attn_kernel(const __grid_constant__ TaMaps maps, const TaParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + TA_TILE_BYTES;
    uint8_t* sV = sK + TA_TILE_BYTES;
    uint8_t* sP = sQ;                                      // 2 x 16 KB (128 keys = two 64-wide K chunks), over Q|K
    uint64_t* bars = reinterpret_cast<uint64_t*>(sV + TA_TILE_BYTES);
    uint64_t* ld_full = bars;
    uint64_t* s_full = bars + 1;
    uint64_t* p_full = bars + 2;
    uint64_t* o_full = bars + 3;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int pix0 = p.pix_begin + blockIdx.x * p.px;
    const int h = blockIdx.y;
    const int b = blockIdx.z;
    const int D = p.head_dim;
    const int DP = (D + 15) & ~15;
    const int R = p.frames * p.px;                         // valid rows of the tile (<= 128)
    const int rows_per_src = p.frames_per_src * p.px;

    // rows [R, 128) are never written by TMA: clear them once so masked lanes multiply zeros, not stale smem
    for (int i = threadIdx.x; i < (128 - R) * 8 * 3; i += TA_THREADS) {
        const int t = i / ((128 - R) * 8), rem = i % ((128 - R) * 8);
        uint8_t* base = (t == 0 ? sQ : t == 1 ? sK : sV) + (R + rem / 8) * 128 + (rem % 8) * 16;
        *reinterpret_cast<uint4*>(base) = make_uint4(0, 0, 0, 0);
    }
    if (warp == 0 && lane == 0) {
        for (int s = 0; s < p.nsrc; ++s) { tma_prefetch_desc(&maps.q[s]); tma_prefetch_desc(&maps.k[s]); tma_prefetch_desc(&maps.v[s]); }
        mbar_init(ld_full, 1); mbar_init(s_full, 1); mbar_init(p_full, 128); mbar_init(o_full, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc<128>(tmem_slot);
    fence_proxy_async();                                   // the zero fill above must be visible to the MMA
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();
    pdl_launch_dependents();

    if (warp == 0) {
        if (elect_one()) {   // single-lane region the compiler can see: operands stay in uniform registers
            mbar_expect_tx(ld_full, 3u * static_cast<uint32_t>(R) * 128u);
            for (int s = 0; s < p.nsrc; ++s) {
                if constexpr (PM) {
                    for (int px = 0; px < p.px; ++px) {       // box = (D, 1 head, 1 pixel, F_s frames)
                        const int off = (px * p.frames + s * p.frames_per_src) * 128;
                        tma_load_5d(sQ + off, &maps.q[s], ld_full, 0, h, pix0 + px, 0, b);
                        tma_load_5d(sK + off, &maps.k[s], ld_full, 0, h, pix0 + px, 0, b);
                        tma_load_5d(sV + off, &maps.v[s], ld_full, 0, h, pix0 + px, 0, b);
                    }
                } else {
                    const int off = s * rows_per_src * 128;
                    tma_load_5d(sQ + off, &maps.q[s], ld_full, 0, h, pix0, 0, b);
                    tma_load_5d(sK + off, &maps.k[s], ld_full, 0, h, pix0, 0, b);
                    tma_load_5d(sV + off, &maps.v[s], ld_full, 0, h, pix0, 0, b);
                }
            }
            const uint32_t idesc_qk = make_idesc_f16(128, 128, DT<T>::kBf16, false, false);
            const uint32_t idesc_pv = make_idesc_f16(128, DP, DT<T>::kBf16, false, true);
            mbar_wait(ld_full, 0);
            tc_fence_after();
            const uint32_t q_addr = smem_u32(sQ), k_addr = smem_u32(sK), v_addr = smem_u32(sV), p_addr = smem_u32(sP);
            for (int ks = 0; ks < DP / 16; ++ks)
                umma_ss(tmem_base, make_smem_desc(q_addr + ks * 32, 16, 1024, kSwizzle128),
                        make_smem_desc(k_addr + ks * 32, 16, 1024, kSwizzle128), idesc_qk, ks > 0 ? 1u : 0u);
            umma_commit(s_full);
            mbar_wait(p_full, 0);
            tc_fence_after();
            for (int ks = 0; ks < 8; ++ks)
                umma_ss(tmem_base, make_smem_desc(p_addr + (ks >> 2) * TA_TILE_BYTES + (ks & 3) * 32, 16, 1024, kSwizzle128),
                        make_smem_desc(v_addr + ks * 16 * 128, TA_TILE_BYTES, 1024, kSwizzle128), idesc_pv, ks > 0 ? 1u : 0u);
            umma_commit(o_full);
        }
    } else {
        const int qd = warp & 3;
        const int r = qd * 32 + lane;
        const uint32_t lane_off = static_cast<uint32_t>(qd * 32) << 16;
        const int my_pix = PM ? r / p.frames : r % p.px;
        const int col_lo = my_pix * p.frames, col_hi = col_lo + p.frames;     // PM: this row's keys
        mbar_wait(s_full, 0);
        tc_fence_after();
        float mx = -INFINITY;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t raw[32];
            tmem_ld32(tmem_base + lane_off + c * 32, raw);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int col = c * 32 + i;
                const bool mine = PM ? (col >= col_lo && col < col_hi) : (col < R && (col % p.px) == my_pix);
                if (mine) mx = fmaxf(mx, __uint_as_float(raw[i]));
            }
        }
        mx *= p.scale_log2;
        const float neg_m = (mx == -INFINITY) ? 0.f : -mx;
        float l_sum = 0.f;
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
            uint32_t raw[32];
            tmem_ld32(tmem_base + lane_off + c * 32, raw);
            tmem_wait_ld();
            float pv[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int col = c * 32 + i;
                const bool mine = PM ? (col >= col_lo && col < col_hi) : (col < R && (col % p.px) == my_pix);
                pv[i] = mine ? ex2_approx_ta(fmaf(__uint_as_float(raw[i]), p.scale_log2, neg_m)) : 0.f;
                l_sum += pv[i];
            }
            uint8_t* slab = sP + (c >> 1) * TA_TILE_BYTES + r * 128;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                float f8[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) f8[i] = pv[8 * g + i];
                const int chunk16 = (c & 1) * 4 + g;
                *reinterpret_cast<uint4*>(slab + ((chunk16 ^ (r & 7)) << 4)) = pack8<T>(f8);
            }
        }
        fence_proxy_async();
        tc_fence_before();
        mbar_arrive(p_full);
        mbar_wait(o_full, 0);
        tc_fence_after();
        const float inv = l_sum > 0.f ? 1.f / l_sum : 0.f;
        const int f = PM ? r % p.frames : r / p.px;
        const int pix = pix0 + my_pix;
        const bool ok = r < R && pix < p.pix_end;
        const int src = ok ? f / p.frames_per_src : 0;
        const int fl = f - src * p.frames_per_src;
        T* orow = reinterpret_cast<T*>(p.out[src]) +
                  ((static_cast<long long>(b) * p.frames_per_src + fl) * p.hw + pix) * p.o_ls + static_cast<long long>(h) * D;
#pragma unroll 1
        for (int c = 0; c < DP; c += 16) {
            uint32_t raw[16];
            tmem_ld16(tmem_base + lane_off + c, raw);
            tmem_wait_ld();
            if (ok) {
#pragma unroll
                for (int g = 0; g < 2; ++g) {
                    if (c + 8 * g < D) {
                        float f8[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) f8[i] = __uint_as_float(raw[8 * g + i]) * inv;
                        *reinterpret_cast<uint4*>(orow + c + 8 * g) = pack8<T>(f8);
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<128>(tmem_base);
}

}  // namespace ab

extern "C" int ab_temporal_attn(const ab_temporal_attn_args* a, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a, "null args");
    const int nsrc = a->nsrc > 0 ? a->nsrc : 1;
    AB_REQUIRE(nsrc == 1 || nsrc == 2 || nsrc == 4 || nsrc == 8, "nsrc must be 1, 2, 4 or 8");
    AB_REQUIRE(a->frames >= 2 && a->frames <= 128 && a->frames % nsrc == 0, "frames must be in [2,128] and divide over the sources");
    AB_REQUIRE(a->head_dim % 8 == 0 && a->head_dim >= 16 && a->head_dim <= 64, "head_dim must be a multiple of 8 in [16,64]");
    AB_REQUIRE(a->row_stride % 8 == 0 && a->out_row_stride % 8 == 0, "row strides must be multiples of 8 elements");
    const int D = a->head_dim, F = a->frames, Fs = F / nsrc;
    const int px = 128 / F;
    AB_REQUIRE(px >= 1, "too many frames for one tile");
    TaParams p{};
    p.batch = a->batch; p.heads = a->heads; p.frames = F; p.frames_per_src = Fs; p.nsrc = nsrc; p.hw = a->hw; p.px = px;
    p.head_dim = D; p.o_ls = a->out_row_stride;
    const float scale = a->scale > 0.f ? a->scale : 1.0f / sqrtf(static_cast<float>(D));
    p.scale_log2 = scale * 1.4426950408889634f;
    // Pixel-major needs every pixel's block of F rows to start on a 1024-byte swizzle atom (F % 8 == 0); a source's
    // F_s rows inside that block only need the 128-byte row alignment: TMA swizzles by shared-memory address.
    const bool pm = (F % 8 == 0);
    const int pix_begin = a->pix_count > 0 ? a->pix_begin : 0;
    const int pix_count = a->pix_count > 0 ? a->pix_count : a->hw;
    AB_REQUIRE(pix_begin >= 0 && pix_begin + pix_count <= a->hw, "pixel band outside the frame");
    p.pix_begin = pix_begin; p.pix_end = pix_begin + pix_count;
    TaMaps maps;
    for (int s = 0; s < nsrc; ++s) {
        AB_REQUIRE(a->q[s] && a->k[s] && a->v[s] && a->out[s], "null source pointer");
        p.out[s] = a->out[s];
        const uint64_t ls = (uint64_t)a->row_stride * 2;
        uint64_t dims[5] = {(uint64_t)D, (uint64_t)a->heads, (uint64_t)a->hw, (uint64_t)Fs, (uint64_t)a->batch};
        uint64_t str[4] = {(uint64_t)D * 2, ls, ls * a->hw, ls * a->hw * Fs};
        uint32_t box[5] = {64, 1, (uint32_t)(pm ? 1 : px), (uint32_t)Fs, 1};
        if (ab_make_tensor_map(&maps.q[s], a->dtype, 5, a->q[s], dims, str, box, 128)) return -1;
        if (ab_make_tensor_map(&maps.k[s], a->dtype, 5, a->k[s], dims, str, box, 128)) return -1;
        if (ab_make_tensor_map(&maps.v[s], a->dtype, 5, a->v[s], dims, str, box, 128)) return -1;
    }
    for (int s = nsrc; s < 8; ++s) { maps.q[s] = maps.q[0]; maps.k[s] = maps.k[0]; maps.v[s] = maps.v[0]; }
    constexpr int smem = 3 * TA_TILE_BYTES + 8 * 8 + 16 + 1024;
    dim3 grid((pix_count + px - 1) / px, a->heads, a->batch);
#define AB_TA_LAUNCH(T, PMV)                                                                                      \
    do {                                                                                                         \
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(temporal_attn_kernel<T, PMV>), smem)) return -1;    \
        AB_CHECK_CUDA(launch_pdl(temporal_attn_kernel<T, PMV>, grid, dim3(TA_THREADS), smem, stream, maps, p));  \
    } while (0)
    if (a->dtype == AB_BF16) { if (pm) AB_TA_LAUNCH(__nv_bfloat16, true); else AB_TA_LAUNCH(__nv_bfloat16, false); }
    else { if (pm) AB_TA_LAUNCH(__half, true); else AB_TA_LAUNCH(__half, false); }
#undef AB_TA_LAUNCH
    ab_count_launch(1);
    return 0;
}
