// Shared by the tcgen05 GEMM kernels (gemm_tc.cu: one-SM tiles, every mode; gemm_2sm.cu: two-SM tiles for large plain
// GEMMs): tile constants, the kernel parameter block and the fused epilogue for one row x 32 accumulator columns.
#pragma once
#include <cuda.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

namespace ab {

constexpr int BM = 128;      // rows per tile == TMEM lanes
constexpr int BK = 64;       // 64 x 16-bit = one 128-byte swizzle row
constexpr int UK = 16;       // K per tcgen05.mma for 16-bit inputs
constexpr int GEMM_THREADS = 320;   // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two per TMEM lane quarter)

struct GemmParams {
    void* out;
    const void* bias;
    const void* rowbias;
    const void* residual;
    float* workspace;
    long long* dbg;             // optional per-CTA phase timestamps (debug/profiling only)
    int stages;                 // pipeline depth of this launch (set by launch_gemm_mt: deep for grids of one wave)
    int M, N, Nout;
    int k1_chunks, k2_chunks;   // 64-wide chunks from source 1 / source 2 (per tap in conv mode)
    int taps;                   // 1 (linear) or 9 (conv)
    int conv, nimg, H, W, halo;
    int bw_log2, bh_log2;       // conv tile: BW x BH x BI pixels, all powers of two, product 128
    int tiles_w, tiles_h;
    int rows_per_group;
    int act, geglu;
    float out_scale;
    int splitk, chunks_per_split;
    int ldo, ldr, ldrb;
    // ---- multi-GPU extensions (one NVSwitch box; pointers may be peer-mapped) ----
    // all-gather -> GEMM: rows of A live in `nsrc` buffers of `tiles_per_src` 128-row tiles each; the gather is
    // the TMA loads themselves (linear mode only).
    int nsrc, tiles_per_src;
    // GEMM -> reduce-scatter: the fp32 partial tile of split (z_base + blockIdx.z) for output row m goes to the
    // workspace of the rank that owns m; owners then sum the 8 canonical K shards in order.
    int rows_per_owner, z_base;
    float* ws_peer[8];
    // Output through TMA: the epilogue stages each 128-row x 32-column box in shared memory (reusing the drained
    // pipeline stages) and one thread issues a bulk tensor store — full-line writes instead of 16-byte pieces
    // scattered over 32 rows per warp instruction. Off for GEGLU, split-K partials and reduce-scatter.
    int tma_store;
    int mt;                     // 128-row M tiles per CTA (1 or 2)
    // cross-GPU hand-off folded into the kernel (patch parallelism): the producer thread waits for the ranks in
    // sync_wait_mask to publish *sync_tick in this GPU's copy of the site before its first load (halo rows arrived)
    const unsigned int* sync_tick;
    const unsigned int* sync_wait;
    unsigned int sync_wait_mask;
    unsigned int sync_top_mask;   // the ranks (of sync_wait_mask) whose rows fill the TOP halo row; the others fill the bottom
    // resampling folded into the conv's coordinates (ab_gemm_args.resample): 1 = stride 2 (H, W = output size, the A
    // map has traversal stride 2 along W and H), 2 = nearest 2x upsample (H, W = input size, one tile = one parity class)
    int resample, pad_lo;
    // M-persistent launches: the grid has fewer CTAs along x than there are (groups of MT) M tiles and CTA x walks tiles
    // x, x + gridDim.x, ... of its N column. Barrier init, TMEM allocation and the launch itself are paid once, and the
    // producer already loads the next tile's operands while the epilogue drains this tile's accumulator. For short K
    // loops (a 128-row tile of a K=320 GEMM is five K steps) those fixed costs were most of a CTA's life.
    int tiles_x;
    alignas(64) CUtensorMap amaps[8];
    alignas(64) CUtensorMap omap;
};

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&f)[8]) {
    uint4 v = *reinterpret_cast<const uint4*>(p);
    unpack8<T>(v, f);
}

// Epilogue modes are template parameters: a run-time switch per element made the epilogue ~100 KB of
// branchy code that thrashed the instruction cache (10k cycles per 32-column chunk, measured).
enum Epi : int { EPI_NONE = 0, EPI_SILU = 1, EPI_GELU = 2, EPI_QGELU = 3, EPI_GEGLU = 4, EPI_ANY = 5 };

template <int EPI>
__device__ __forceinline__ float epi_act(float x, int act) {
    if constexpr (EPI == EPI_NONE) return x;
    else if constexpr (EPI == EPI_SILU) return __fdividef(x, 1.0f + __expf(-x));
    else if constexpr (EPI == EPI_GELU) return act_gelu_erf(x);
    else if constexpr (EPI == EPI_QGELU) return __fdividef(x, 1.0f + __expf(-1.702f * x));
    else return apply_act(x, act);
}

// Epilogue for one row x 32 accumulator columns starting at global column n_base.
template <typename T, int EPI>
__device__ __forceinline__ void epilogue_chunk32(const GemmParams& p, long long out_row, int grp, int n_base,
                                                 float (&v)[32], const float* sb, bool rowbias_global,
                                                 uint8_t* stage_row = nullptr, int stage_swz = 0, int stage_slot0 = 0) {
    const T* bias = reinterpret_cast<const T*>(p.bias);
    const T* rowbias = reinterpret_cast<const T*>(p.rowbias);
    const T* residual = reinterpret_cast<const T*>(p.residual);
    T* out = reinterpret_cast<T*>(p.out);
    if (sb) {
        // bias (+ row bias when the tile sits in one group) was staged in shared memory during the main loop
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            const float4 b4 = *reinterpret_cast<const float4*>(sb + 4 * g);
            v[4 * g] += b4.x; v[4 * g + 1] += b4.y; v[4 * g + 2] += b4.z; v[4 * g + 3] += b4.w;
        }
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const int n = n_base + 8 * g;
        if (n < p.N) {
            float b[8];
            if (bias && !sb) {
                load8<T>(bias + n, b);
#pragma unroll
                for (int i = 0; i < 8; ++i) v[8 * g + i] += b[i];
            }
            if (rowbias && rowbias_global) {
                load8<T>(rowbias + static_cast<long long>(grp) * p.ldrb + n, b);
#pragma unroll
                for (int i = 0; i < 8; ++i) v[8 * g + i] += b[i];
            }
        }
    }
    if constexpr (EPI == EPI_GEGLU) {
        // columns [0,16) are values, [16,32) the matching gates (weights were interleaved at load).
        const int no = n_base >> 1;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
            const int n = no + 8 * g;
            if (n < p.Nout) {
                float o[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = v[8 * g + i] * act_gelu_erf(v[16 + 8 * g + i]) * p.out_scale;
                if (residual) {
                    float r[8];
                    load8<T>(residual + out_row * p.ldr + n, r);
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] += r[i];
                }
                if (stage_row) {
                    // two GEGLU chunks (16 output columns each) share one 64-byte staging row: slots 0-1 and 2-3
                    *reinterpret_cast<uint4*>(stage_row + (((stage_slot0 + g) ^ stage_swz) << 4)) = pack8<T>(o);
                } else {
                    *reinterpret_cast<uint4*>(out + out_row * p.ldo + n) = pack8<T>(o);
                }
            }
        }
    } else {
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            const int n = n_base + 8 * g;
            if (n < p.N) {
                float o[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = epi_act<EPI>(v[8 * g + i], p.act) * p.out_scale;
                if (residual) {
                    float r[8];
                    load8<T>(residual + out_row * p.ldr + n, r);
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] += r[i];
                }
                if (stage_row) {
                    // 64-byte rows in the SWIZZLE_64B pattern of the output tensor map: 16-byte slot g ^ ((row >> 1) & 3)
                    *reinterpret_cast<uint4*>(stage_row + ((g ^ stage_swz) << 4)) = pack8<T>(o);
                } else {
                    *reinterpret_cast<uint4*>(out + out_row * p.ldo + n) = pack8<T>(o);
                }
            }
        }
    }
}

// ---- host side: the 128-pixel conv tile (BW x BH x BI, powers of two) for an H x W image ----
inline int ilog2(int v) { int l = 0; while ((1 << l) < v) ++l; return l; }

struct ConvTile { int bw, bh, bi; };
inline ConvTile pick_conv_tile(int H, int W) {
    // BW: power of two <= 128 with the least padding waste (ties -> larger)
    int best_bw = 8; double best = 1e9;
    for (int bw = 128; bw >= 8; bw >>= 1) {
        double waste = double((W + bw - 1) / bw * bw) / W;
        if (waste < best - 1e-9) { best = waste; best_bw = bw; }
    }
    int bh_max = 128 / best_bw;
    int bh = 1;
    while (bh < bh_max && bh < H) bh <<= 1;   // smallest pow2 >= H, capped
    if (bh > bh_max) bh = bh_max;
    return {best_bw, bh, 128 / (best_bw * bh)};
}

}  // namespace ab
