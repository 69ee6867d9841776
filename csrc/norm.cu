// GroupNorm(+SiLU) and LayerNorm on channels-last tensors.
//
// GroupNorm statistics follow the canonical 8-shard rule: the HW axis is cut into 8 row shards, each
// shard into SUB blocks (SUB depends on HW only), every block reduces its rows in a fixed order to a
// (count, mean, M2) triple per group, and the last block to finish merges all partials (Chan's parallel
// variance update, fp32) in (shard, block) index order. Which block happens to be last does not matter: the combine order is
// fixed, so the statistics are bit-identical for any grid size — and, with shards living on
// different GPUs, for any GPU count.
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

constexpr int GN_SHARDS = 8;
constexpr int GN_MAX_SUB = 32;
constexpr int GN_THREADS = 512;
constexpr int GN_APPLY_THREADS = 256;   // apply is elementwise: smaller blocks, two or more per SM, so loads of one overlap stores of another
constexpr int GN_MAX_GROUPS = 64;

__host__ __device__ inline int gn_sub_blocks(int hw) {
    int rows = (hw + GN_SHARDS - 1) / GN_SHARDS;
    // 64 rows per block keeps small maps parallel; from 8192 positions up, 128 rows amortise the per-block
    // reduction and ticket (a function of HW only: the partition is part of the canonical reduction order)
    int sub = hw >= 8192 ? (rows + 127) / 128 : (rows + 63) / 64;
    return sub < 1 ? 1 : (sub > GN_MAX_SUB ? GN_MAX_SUB : sub);
}

struct GnParams {
    const void* x; const void* x2; void* y;
    const void* gamma; const void* beta;
    const void* scale; const void* shift;
    float4* partials;      // [n][8][sub][groups]  (count, mean, M2, -)
    float2* stats;         // [n][groups] (mean, rstd)
    unsigned int* counters;  // [n]
    int n, hw, c1, c2, groups, sub;
    float eps;
    int act;
    int ss;                  // row stride (elements) of scale/shift
    int apply_nb;            // batches of GN_APPLY_ROWS rows per apply thread (amortises the affine set-up on big maps)
    // sharded tensors (one sample spread over GPUs by rows): this GPU holds canonical shards
    // [shard_base, shard_base + ns_local) of every sample; hw is then the LOCAL row count. ns_local == 0: all 8.
    int ns_local, shard_base;
    long long y_bs;          // elements between consecutive samples of y (0: hw * C)
};

// workspace layout: counters (n * 4 B, padded to 256) | stats (n*groups*8 B, padded to 256) | partials
// (the two words after the n per-image counters are the fused kernel's grid barrier: arrivals, generation)
__host__ inline size_t gn_counters_bytes(int n) { return ((size_t)(n + 2) * 4 + 255) / 256 * 256; }
__host__ inline size_t gn_stats_bytes(int n, int groups) { return ((size_t)n * groups * 8 + 255) / 256 * 256; }

// Parallel-variance merge (Chan et al.): exact-order, fp32, no cancellation between blocks.
struct GnMoments { float cnt, mean, m2; };
__device__ __forceinline__ GnMoments gn_merge(const GnMoments a, const GnMoments b) {
    if (b.cnt == 0.f) return a;
    if (a.cnt == 0.f) return b;
    GnMoments r;
    r.cnt = a.cnt + b.cnt;
    const float delta = b.mean - a.mean;
    const float w = __fdividef(b.cnt, r.cnt);
    r.mean = fmaf(delta, w, a.mean);
    r.m2 = a.m2 + b.m2 + delta * delta * a.cnt * w;
    return r;
}

// Per-block group moments from the per-thread column sums in shared memory ([RL][C] sums and squares): sixteen lanes
// per group, lane j takes the values j, j+16, ... of the group's RL*cpg entries (index rl*cpg + k) and a fixed xor tree
// combines them — the same order in every kernel that calls this, so the two-kernel and the fused path agree bitwise.
__device__ __forceinline__ void gn_block_moments(const float* s_sum, const float* s_sq, int RL, int C, int cpg, int groups,
                                                 int rows, float4* dst) {
    const int j = threadIdx.x & 15;
    const int nval = RL * cpg;
    for (int g = threadIdx.x >> 4; g < ((groups + 1) & ~1); g += GN_THREADS >> 4) {
        float ds = 0.f, dq = 0.f;
        if (g < groups) {
            for (int idx = j; idx < nval; idx += 16) {
                const int rl = idx / cpg, k = idx - rl * cpg;
                ds += s_sum[rl * C + g * cpg + k];
                dq += s_sq[rl * C + g * cpg + k];
            }
        }
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            ds += __shfl_xor_sync(0xffffffffu, ds, o);
            dq += __shfl_xor_sync(0xffffffffu, dq, o);
        }
        if (j == 0 && g < groups) {
            const float cnt = static_cast<float>(rows) * cpg;
            const float mean = cnt > 0.f ? ds / cnt : 0.f;
            const float m2 = fmaxf(dq - ds * mean, 0.f);
            dst[g] = make_float4(cnt, mean, m2, 0.f);
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(GN_THREADS) gn_stats_kernel(const GnParams p) {
    extern __shared__ float sm[];   // [2][RL][C]
    __shared__ bool is_last;
    __shared__ GnMoments shard_mom[GN_SHARDS][GN_MAX_GROUPS + 1];
    const int C = p.c1 + p.c2;
    const int ncol = C / 8;
    const int cpg = C / p.groups;
    const int n = blockIdx.z, shard = blockIdx.y, sub = blockIdx.x;
    pdl_wait();
    pdl_launch_dependents();
    const int ns = p.ns_local > 0 ? p.ns_local : GN_SHARDS;
    const int s_begin = static_cast<int>((static_cast<long long>(shard) * p.hw) / ns);
    const int s_end = static_cast<int>((static_cast<long long>(shard + 1) * p.hw) / ns);
    const int len = s_end - s_begin;
    const int r0 = s_begin + static_cast<int>((static_cast<long long>(sub) * len) / p.sub);
    const int r1 = s_begin + static_cast<int>((static_cast<long long>(sub + 1) * len) / p.sub);

    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const int passes = (ncol + GN_THREADS - 1) / GN_THREADS;
    float* s_sum = sm;
    float* s_sq = sm + RL * C;
    const T* x1 = reinterpret_cast<const T*>(p.x) + static_cast<long long>(n) * p.hw * p.c1;
    const T* x2 = p.x2 ? reinterpret_cast<const T*>(p.x2) + static_cast<long long>(n) * p.hw * p.c2 : nullptr;

    for (int pass = 0; pass < passes; ++pass) {
        int vc, rl;
        bool active;
        if (passes == 1) { vc = threadIdx.x % ncol; rl = threadIdx.x / ncol; active = rl < RL; }
        else { vc = pass * GN_THREADS + threadIdx.x; rl = 0; active = vc < ncol; }
        if (active) {
            const int c = vc * 8;
            const T* src; int ld, cc;
            if (c < p.c1) { src = x1; ld = p.c1; cc = c; } else { src = x2; ld = p.c2; cc = c - p.c1; }
            float s[8], q[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
            const T* ptr = src + static_cast<long long>(r0 + rl) * ld + cc;
            const long long step = static_cast<long long>(RL) * ld;
            int r = r0 + rl;
            for (; r + 3 * RL < r1; r += 4 * RL) {          // 4 independent loads in flight
                uint4 v0 = *reinterpret_cast<const uint4*>(ptr);
                uint4 v1 = *reinterpret_cast<const uint4*>(ptr + step);
                uint4 v2 = *reinterpret_cast<const uint4*>(ptr + 2 * step);
                uint4 v3 = *reinterpret_cast<const uint4*>(ptr + 3 * step);
                ptr += 4 * step;
                float f[8];
                unpack8<T>(v0, f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                unpack8<T>(v1, f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                unpack8<T>(v2, f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                unpack8<T>(v3, f);
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
            }
            for (; r < r1; r += RL) {
                float f[8];
                unpack8<T>(*reinterpret_cast<const uint4*>(ptr), f);
                ptr += step;
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) { s_sum[rl * C + c + i] = s[i]; s_sq[rl * C + c + i] = q[i]; }
        }
    }
    __syncthreads();
    gn_block_moments(s_sum, s_sq, RL, C, cpg, p.groups, r1 - r0,
                     p.partials + ((static_cast<long long>(n) * GN_SHARDS + p.shard_base + shard) * p.sub + sub) * p.groups);
    if (p.ns_local > 0) return;   // sharded: gn_finalize_sharded_kernel merges every GPU's partials after the barrier
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned int ticket = atomicAdd(&p.counters[n], 1u);
        is_last = (ticket == static_cast<unsigned int>(GN_SHARDS * p.sub - 1));
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        // level 1: thread (shard, g) merges that shard's blocks in index order; level 2: shards 0..7 in order
        const volatile float4* part = p.partials + static_cast<long long>(n) * GN_SHARDS * p.sub * p.groups;
        for (int idx = threadIdx.x; idx < GN_SHARDS * p.groups; idx += GN_THREADS) {
            const int sh = idx / p.groups, g = idx % p.groups;
            GnMoments acc{0.f, 0.f, 0.f};
            for (int sb = 0; sb < p.sub; ++sb) {
                const volatile float4* q4 = &part[(sh * p.sub + sb) * p.groups + g];
                acc = gn_merge(acc, GnMoments{q4->x, q4->y, q4->z});
            }
            shard_mom[sh][g] = acc;
        }
        __syncthreads();
        if (threadIdx.x < p.groups) {
            const int g = threadIdx.x;
            GnMoments acc = shard_mom[0][g];
#pragma unroll
            for (int sh = 1; sh < GN_SHARDS; ++sh) acc = gn_merge(acc, shard_mom[sh][g]);
            const float var = acc.m2 / acc.cnt;
            p.stats[n * p.groups + g] = make_float2(acc.mean, rsqrtf(var + p.eps));
        }
        if (threadIdx.x == 0) p.counters[n] = 0;   // ready for the next call on this workspace
    }
}

// y = act((x - mean) * rstd * gamma + beta [ * (1 + scale) + shift ]): per-thread channel vector is fixed, so
// the affine is folded once into (a, b) and every row costs one 16-byte load, 8 FMAs and one 16-byte store.
constexpr int GN_APPLY_ROWS = 8;   // rows per thread
template <typename T>
__global__ void __launch_bounds__(GN_APPLY_THREADS, 2) gn_apply_kernel(const GnParams p) {
    const int C = p.c1 + p.c2;
    const int ncol = C / 8;
    const int cpg = C / p.groups;
    const int n = blockIdx.y;
    const int cols = ncol < GN_APPLY_THREADS ? ncol : GN_APPLY_THREADS;
    const int RL = GN_APPLY_THREADS / cols;
    const int passes = (ncol + cols - 1) / cols;
    const int rl = threadIdx.x / cols;
    pdl_wait();
    pdl_launch_dependents();
    if (rl >= RL) return;
    const int nb = p.apply_nb > 0 ? p.apply_nb : 1;
    const int r_block = blockIdx.x * (RL * GN_APPLY_ROWS * nb);
    const int r_end = min(p.hw, r_block + RL * GN_APPLY_ROWS * nb);
    const T* gamma = reinterpret_cast<const T*>(p.gamma);
    const T* beta = reinterpret_cast<const T*>(p.beta);
    const T* scale = reinterpret_cast<const T*>(p.scale);
    const T* shift = reinterpret_cast<const T*>(p.shift);
    T* y = reinterpret_cast<T*>(p.y) + (p.y_bs > 0 ? static_cast<long long>(n) * p.y_bs : static_cast<long long>(n) * p.hw * C);
    for (int pass = 0; pass < passes; ++pass) {
        const int vc = pass * cols + threadIdx.x % cols;
        if (vc >= ncol) continue;
        const int c = vc * 8;
        const T* src; int ld, cc;
        if (c < p.c1) { src = reinterpret_cast<const T*>(p.x) + static_cast<long long>(n) * p.hw * p.c1; ld = p.c1; cc = c; }
        else { src = reinterpret_cast<const T*>(p.x2) + static_cast<long long>(n) * p.hw * p.c2; ld = p.c2; cc = c - p.c1; }
        float a[8], b[8];
        {
            float ga[8], be[8], sc[8], sh[8];
            if (gamma) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(gamma + c)), ga);
            if (beta) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(beta + c)), be);
            if (scale) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(scale + static_cast<long long>(n) * p.ss + c)), sc);
            if (shift) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(shift + static_cast<long long>(n) * p.ss + c)), sh);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float2 st = __ldg(&p.stats[n * p.groups + (c + i) / cpg]);
                float ai = st.y, bi = -st.x * st.y;
                if (gamma) { ai *= ga[i]; bi *= ga[i]; }
                if (beta) bi += be[i];
                if (scale) { ai *= (1.0f + sc[i]); bi *= (1.0f + sc[i]); }
                if (shift) bi += sh[i];
                a[i] = ai; b[i] = bi;
            }
        }
        for (int bt = 0; bt < nb; ++bt) {
        const int r_begin = r_block + bt * RL * GN_APPLY_ROWS;
        // all of this thread's row vectors are requested before the first one is used (8 x 16 B in flight per thread)
        uint4 v[GN_APPLY_ROWS];
#pragma unroll
        for (int k = 0; k < GN_APPLY_ROWS; ++k) {
            const int r = r_begin + rl + k * RL;
            if (r < r_end) v[k] = __ldg(reinterpret_cast<const uint4*>(src + static_cast<long long>(r) * ld + cc));
        }
#pragma unroll
        for (int k = 0; k < GN_APPLY_ROWS; ++k) {
            const int r = r_begin + rl + k * RL;
            if (r >= r_end) break;
            float f[8], o[8];
            unpack8<T>(v[k], f);
            if (p.act == ACT_SILU) {
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = act_silu(fmaf(f[i], a[i], b[i]));
            } else if (p.act == ACT_NONE) {
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = fmaf(f[i], a[i], b[i]);
            } else {
#pragma unroll
                for (int i = 0; i < 8; ++i) o[i] = apply_act(fmaf(f[i], a[i], b[i]), p.act);
            }
            *reinterpret_cast<uint4*>(y + static_cast<long long>(r) * C + c) = pack8<T>(o);
        }
        }   // batches
    }
}

// Sharded GroupNorm, second step: every GPU merges ALL partial moments — its own and, through peer pointers, the other
// GPUs' — in the canonical order of gn_stats_kernel's last block (blocks of a shard in index order, then shards 0..7),
// so each GPU ends up with the statistics the undivided tensor gives, bit for bit. One block per sample.
struct GnShardSrcs { const float4* part[GN_SHARDS]; };   // partial-moment array of the GPU that owns canonical shard i
__global__ void __launch_bounds__(GN_THREADS) gn_finalize_sharded_kernel(const GnParams p, const GnShardSrcs srcs) {
    __shared__ GnMoments shard_mom[GN_SHARDS][GN_MAX_GROUPS + 1];
    const int n = blockIdx.x;
    for (int idx = threadIdx.x; idx < GN_SHARDS * p.groups; idx += GN_THREADS) {
        const int sh = idx / p.groups, g = idx % p.groups;
        const volatile float4* part = srcs.part[sh] + static_cast<long long>(n) * GN_SHARDS * p.sub * p.groups;
        GnMoments acc{0.f, 0.f, 0.f};
        for (int sb = 0; sb < p.sub; ++sb) {
            const volatile float4* q4 = &part[(sh * p.sub + sb) * p.groups + g];
            acc = gn_merge(acc, GnMoments{q4->x, q4->y, q4->z});
        }
        shard_mom[sh][g] = acc;
    }
    __syncthreads();
    if (threadIdx.x < p.groups) {
        const int g = threadIdx.x;
        GnMoments acc = shard_mom[0][g];
#pragma unroll
        for (int sh = 1; sh < GN_SHARDS; ++sh) acc = gn_merge(acc, shard_mom[sh][g]);
        const float var = acc.m2 / acc.cnt;
        p.stats[n * p.groups + g] = make_float2(acc.mean, rsqrtf(var + p.eps));
    }
}

// ---- fused GroupNorm: statistics + grid barrier + apply in ONE cooperative launch ----------------------
// Same canonical decomposition and the same arithmetic order as gn_stats_kernel + gn_apply_kernel (bit-identical
// output), but the rows a CTA reduced stay in its shared memory, so the tensor is read from HBM once, and the second
// launch (and its drain/ramp) is gone. Items (n, shard, sub) are dealt round-robin to at most one CTA per SM; all CTAs
// are co-resident (cooperative launch), so the barrier between the phases cannot deadlock.
__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
// bar[0] = arrivals, bar[1] = generation; self-resetting, so the same two words serve every call on a stream
__device__ __forceinline__ void gn_grid_barrier(unsigned int* bar, unsigned int nblocks) {
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int gen = ld_acquire_u32(&bar[1]);
        __threadfence();
        const unsigned int ticket = atomicAdd(&bar[0], 1u);
        if (ticket == nblocks - 1) {
            bar[0] = 0;
            __threadfence();
            st_release_u32(&bar[1], gen + 1);
        } else {
            while (ld_acquire_u32(&bar[1]) == gen) __nanosleep(32);
        }
    }
    __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(GN_THREADS) gn_fused_kernel(const GnParams p, unsigned int* bar, int cache_rows) {
    extern __shared__ __align__(16) unsigned char gn_smem[];
    __shared__ GnMoments shard_mom[GN_SHARDS][GN_MAX_GROUPS + 1];
    __shared__ float2 s_stats[GN_MAX_GROUPS];
    const int C = p.c1 + p.c2;
    const int ncol = C / 8;
    const int cpg = C / p.groups;
    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const int passes = (ncol + GN_THREADS - 1) / GN_THREADS;
    float* s_sum = reinterpret_cast<float*>(gn_smem);
    float* s_sq = s_sum + RL * C;
    uint4* cache = reinterpret_cast<uint4*>(gn_smem + static_cast<size_t>(2) * RL * C * sizeof(float));   // [rows][ncol]
    const int items_per_img = GN_SHARDS * p.sub;
    const int total_items = p.n * items_per_img;
    const bool use_cache = cache_rows > 0;

    auto item_rows = [&](int item, int& n, int& r0, int& r1) {
        n = item / items_per_img;
        const int rem = item - n * items_per_img;
        const int shard = rem / p.sub, sub = rem - shard * p.sub;
        const int s_begin = static_cast<int>((static_cast<long long>(shard) * p.hw) / GN_SHARDS);
        const int s_end = static_cast<int>((static_cast<long long>(shard + 1) * p.hw) / GN_SHARDS);
        const int len = s_end - s_begin;
        r0 = s_begin + static_cast<int>((static_cast<long long>(sub) * len) / p.sub);
        r1 = s_begin + static_cast<int>((static_cast<long long>(sub + 1) * len) / p.sub);
    };

    // ---------------- phase 1: per-item moments (+ keep the rows) ----------------
    int cached = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        int n, r0, r1;
        item_rows(item, n, r0, r1);
        const T* x1 = reinterpret_cast<const T*>(p.x) + static_cast<long long>(n) * p.hw * p.c1;
        const T* x2 = p.x2 ? reinterpret_cast<const T*>(p.x2) + static_cast<long long>(n) * p.hw * p.c2 : nullptr;
        for (int pass = 0; pass < passes; ++pass) {
            int vc, rl;
            bool active;
            if (passes == 1) { vc = threadIdx.x % ncol; rl = threadIdx.x / ncol; active = rl < RL; }
            else { vc = pass * GN_THREADS + threadIdx.x; rl = 0; active = vc < ncol; }
            if (active) {
                const int c = vc * 8;
                const T* src; int ld, cc;
                if (c < p.c1) { src = x1; ld = p.c1; cc = c; } else { src = x2; ld = p.c2; cc = c - p.c1; }
                float s[8], q[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
                const T* ptr = src + static_cast<long long>(r0 + rl) * ld + cc;
                const long long step = static_cast<long long>(RL) * ld;
                uint4* cp = cache + static_cast<long long>(cached + rl) * ncol + vc;
                const long long cstep = static_cast<long long>(RL) * ncol;
                int r = r0 + rl;
                for (; r + 3 * RL < r1; r += 4 * RL) {
                    uint4 v0 = *reinterpret_cast<const uint4*>(ptr);
                    uint4 v1 = *reinterpret_cast<const uint4*>(ptr + step);
                    uint4 v2 = *reinterpret_cast<const uint4*>(ptr + 2 * step);
                    uint4 v3 = *reinterpret_cast<const uint4*>(ptr + 3 * step);
                    ptr += 4 * step;
                    if (use_cache) { cp[0] = v0; cp[cstep] = v1; cp[2 * cstep] = v2; cp[3 * cstep] = v3; cp += 4 * cstep; }
                    float f[8];
                    unpack8<T>(v0, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                    unpack8<T>(v1, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                    unpack8<T>(v2, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                    unpack8<T>(v3, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                }
                for (; r < r1; r += RL) {
                    const uint4 v = *reinterpret_cast<const uint4*>(ptr);
                    ptr += step;
                    if (use_cache) { cp[0] = v; cp += cstep; }
                    float f[8];
                    unpack8<T>(v, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) { s_sum[rl * C + c + i] = s[i]; s_sq[rl * C + c + i] = q[i]; }
            }
        }
        __syncthreads();
        gn_block_moments(s_sum, s_sq, RL, C, cpg, p.groups, r1 - r0, p.partials + static_cast<long long>(item) * p.groups);
        __threadfence();
        __syncthreads();
        cached += r1 - r0;
    }

    // the affine parameters do not depend on the statistics: fetch them before the barrier, not after it
    const T* gamma = reinterpret_cast<const T*>(p.gamma);
    const T* beta = reinterpret_cast<const T*>(p.beta);
    uint4 h_gamma = make_uint4(0, 0, 0, 0), h_beta = make_uint4(0, 0, 0, 0);
    if (passes == 1 && threadIdx.x / ncol < RL) {
        const int c_h = (threadIdx.x % ncol) * 8;
        if (gamma) h_gamma = __ldg(reinterpret_cast<const uint4*>(gamma + c_h));
        if (beta) h_beta = __ldg(reinterpret_cast<const uint4*>(beta + c_h));
    }

    gn_grid_barrier(bar, gridDim.x);

    // ---------------- phase 2 + 3: merge the image's partials in canonical order, then normalise ----------------
    const T* scale = reinterpret_cast<const T*>(p.scale);
    const T* shift = reinterpret_cast<const T*>(p.shift);
    int stats_n = -1;
    cached = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        int n, r0, r1;
        item_rows(item, n, r0, r1);
        if (n != stats_n) {
            __syncthreads();
            const float4* part = p.partials + static_cast<long long>(n) * items_per_img * p.groups;
            for (int idx = threadIdx.x; idx < GN_SHARDS * p.groups; idx += GN_THREADS) {
                const int sh = idx / p.groups, g = idx % p.groups;
                GnMoments acc{0.f, 0.f, 0.f};
                for (int sb0 = 0; sb0 < p.sub; sb0 += 8) {           // eight L2 loads in flight, merged in index order
                    float4 q4[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (sb0 + u < p.sub) q4[u] = __ldcg(&part[(sh * p.sub + sb0 + u) * p.groups + g]);
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (sb0 + u < p.sub) acc = gn_merge(acc, GnMoments{q4[u].x, q4[u].y, q4[u].z});
                }
                shard_mom[sh][g] = acc;
            }
            __syncthreads();
            if (threadIdx.x < p.groups) {
                const int g = threadIdx.x;
                GnMoments acc = shard_mom[0][g];
#pragma unroll
                for (int sh = 1; sh < GN_SHARDS; ++sh) acc = gn_merge(acc, shard_mom[sh][g]);
                const float var = acc.m2 / acc.cnt;
                s_stats[g] = make_float2(acc.mean, rsqrtf(var + p.eps));
            }
            __syncthreads();
            stats_n = n;
        }
        T* y = reinterpret_cast<T*>(p.y) + static_cast<long long>(n) * p.hw * C;
        for (int pass = 0; pass < passes; ++pass) {
            int vc, rl;
            bool active;
            if (passes == 1) { vc = threadIdx.x % ncol; rl = threadIdx.x / ncol; active = rl < RL; }
            else { vc = pass * GN_THREADS + threadIdx.x; rl = 0; active = vc < ncol; }
            if (!active) continue;
            const int c = vc * 8;
            const T* src; int ld, cc;
            if (c < p.c1) { src = reinterpret_cast<const T*>(p.x) + static_cast<long long>(n) * p.hw * p.c1; ld = p.c1; cc = c; }
            else { src = reinterpret_cast<const T*>(p.x2) + static_cast<long long>(n) * p.hw * p.c2; ld = p.c2; cc = c - p.c1; }
            float a[8], b[8];
            {
                float ga[8], be[8], sc[8], sh[8];
                if (gamma) unpack8<T>(passes == 1 ? h_gamma : __ldg(reinterpret_cast<const uint4*>(gamma + c)), ga);
                if (beta) unpack8<T>(passes == 1 ? h_beta : __ldg(reinterpret_cast<const uint4*>(beta + c)), be);
                if (scale) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(scale + static_cast<long long>(n) * p.ss + c)), sc);
                if (shift) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(shift + static_cast<long long>(n) * p.ss + c)), sh);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float2 st = s_stats[(c + i) / cpg];
                    float ai = st.y, bi = -st.x * st.y;
                    if (gamma) { ai *= ga[i]; bi *= ga[i]; }
                    if (beta) bi += be[i];
                    if (scale) { ai *= (1.0f + sc[i]); bi *= (1.0f + sc[i]); }
                    if (shift) bi += sh[i];
                    a[i] = ai; b[i] = bi;
                }
            }
            for (int r = r0 + rl; r < r1; r += RL) {
                float f[8], o[8];
                const uint4 v = use_cache ? cache[static_cast<long long>(cached + (r - r0)) * ncol + vc]
                                          : *reinterpret_cast<const uint4*>(src + static_cast<long long>(r) * ld + cc);
                unpack8<T>(v, f);
                if (p.act == ACT_SILU) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = act_silu(fmaf(f[i], a[i], b[i]));
                } else if (p.act == ACT_NONE) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = fmaf(f[i], a[i], b[i]);
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = apply_act(fmaf(f[i], a[i], b[i]), p.act);
                }
                *reinterpret_cast<uint4*>(y + static_cast<long long>(r) * C + c) = pack8<T>(o);
            }
        }
        cached += r1 - r0;
    }
}

// ---- patch-parallel GroupNorm: statistics, the exchange between the GPUs, and apply in ONE launch ----------------
// One image's rows are spread over `world` GPUs in row bands; this GPU holds canonical shards
// [shard_base, shard_base + ns_local). Phase 1 is gn_fused_kernel's (same items, same arithmetic). Every item's moments
// then go to EVERY GPU's array (its own included) as self-validating 16-byte entries {mean, tick, M2, tick}: the flag
// travels inside the data (each 8-byte half carries the tick, so a torn 16-byte store cannot pass), which removes the
// fence + flag hop of a separate publish and the local grid barrier — a consumer simply polls the entries it needs
// until they carry this step's tick. The count of an entry is a function of the shapes alone and is recomputed by the
// reader with the writer's expression, so the merged statistics are bit-identical to the undivided kernel's.
// Phase 2/3: merge the 8 shards in canonical order, normalise the local rows; rows at the band's edges are also stored
// into the neighbours' halo slots, and the last CTA to finish publishes "rows written" (a separate flag: the halo rows
// are too large to carry their own) — the next conv's TMA producer warp, or a gather, waits on that.
struct GnPatch {
    uint4* part_peers[8];
    const unsigned int* tick;
    void* up_halo; void* down_halo;
    int halo_rows;
    unsigned int* done_sites[8];
    unsigned int* done_local;
    int world, rank, dry;
    // single-GPU use (ab_groupnorm): the tick is a counter in the caller's workspace that the last CTA of every launch
    // advances, so consecutive launches on a stream never see each other's entries as current; no flags go anywhere
    unsigned int* self_tick;
    unsigned int* self_ticket;
};

__device__ __forceinline__ void st_ll16(uint4* p, uint4 v) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_ll16(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
// poll one entry until both halves carry `tick` (bounded: a dead peer traps after ~60 s instead of hanging the box)
__device__ __forceinline__ uint4 gn_poll_entry(const uint4* p, unsigned int tick, bool dry) {
    uint4 v = ld_ll16(p);
    if (dry || (v.y == tick && v.w == tick)) return v;
    const long long t0 = clock64();
    for (;;) {
        v = ld_ll16(p);
        if (v.y == tick && v.w == tick) return v;
        __nanosleep(20);
        if (clock64() - t0 > 120000000000LL) {
            printf("gn_patch: moments entry never arrived (tick %u, have %u/%u) block %d\n", tick, v.y, v.w, blockIdx.x);
            __trap();
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(GN_THREADS) gn_patch_kernel(const GnParams p, const __grid_constant__ GnPatch x,
                                                              int cache_rows) {
    extern __shared__ __align__(16) unsigned char gn_smem[];
    __shared__ GnMoments shard_mom[GN_SHARDS][GN_MAX_GROUPS + 1];
    __shared__ float2 s_stats[GN_MAX_GROUPS];
    __shared__ float4 s_item[GN_MAX_GROUPS];
    const int C = p.c1 + p.c2;
    const int ncol = C / 8;
    const int cpg = C / p.groups;
    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const int passes = (ncol + GN_THREADS - 1) / GN_THREADS;
    float* s_sum = reinterpret_cast<float*>(gn_smem);
    float* s_sq = s_sum + RL * C;
    uint4* cache = reinterpret_cast<uint4*>(gn_smem + static_cast<size_t>(2) * RL * C * sizeof(float));   // [rows][ncol]
    const int ns = p.ns_local;
    const int items_per_img = ns * p.sub;
    const int total_items = p.n * items_per_img;
    const bool use_cache = cache_rows > 0;
    const unsigned int tick = x.self_tick ? *x.self_tick + 1u : (x.tick ? *x.tick : 0u);
    const int shard_len = p.hw / ns;                       // rows of one canonical shard (hw % ns == 0)

    auto item_rows = [&](int item, int& n, int& shard, int& sb, int& r0, int& r1) {
        n = item / items_per_img;
        const int rem = item - n * items_per_img;
        shard = rem / p.sub; sb = rem - shard * p.sub;
        const int s_begin = shard * shard_len;             // (callers guarantee hw % ns == 0)
        r0 = s_begin + static_cast<int>((static_cast<long long>(sb) * shard_len) / p.sub);
        r1 = s_begin + static_cast<int>((static_cast<long long>(sb + 1) * shard_len) / p.sub);
    };

    // ---------------- phase 1: per-item moments (+ keep the rows), stored to every GPU with the tick inside ----------------
    int cached = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        int n, shard, sb, r0, r1;
        item_rows(item, n, shard, sb, r0, r1);
        const T* x1 = reinterpret_cast<const T*>(p.x) + static_cast<long long>(n) * p.hw * p.c1;
        const T* x2 = p.x2 ? reinterpret_cast<const T*>(p.x2) + static_cast<long long>(n) * p.hw * p.c2 : nullptr;
        for (int pass = 0; pass < passes; ++pass) {
            int vc, rl;
            bool active;
            if (passes == 1) { vc = threadIdx.x % ncol; rl = threadIdx.x / ncol; active = rl < RL; }
            else { vc = pass * GN_THREADS + threadIdx.x; rl = 0; active = vc < ncol; }
            if (active) {
                const int c = vc * 8;
                const T* src; int ld, cc;
                if (c < p.c1) { src = x1; ld = p.c1; cc = c; } else { src = x2; ld = p.c2; cc = c - p.c1; }
                float s[8], q[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) { s[i] = 0.f; q[i] = 0.f; }
                const T* ptr = src + static_cast<long long>(r0 + rl) * ld + cc;
                const long long step = static_cast<long long>(RL) * ld;
                uint4* cp = cache + static_cast<long long>(cached + rl) * ncol + vc;
                const long long cstep = static_cast<long long>(RL) * ncol;
                int r = r0 + rl;
                for (; r + 3 * RL < r1; r += 4 * RL) {
                    uint4 v0 = *reinterpret_cast<const uint4*>(ptr);
                    uint4 v1 = *reinterpret_cast<const uint4*>(ptr + step);
                    uint4 v2 = *reinterpret_cast<const uint4*>(ptr + 2 * step);
                    uint4 v3 = *reinterpret_cast<const uint4*>(ptr + 3 * step);
                    ptr += 4 * step;
                    if (use_cache) { cp[0] = v0; cp[cstep] = v1; cp[2 * cstep] = v2; cp[3 * cstep] = v3; cp += 4 * cstep; }
                    float f[8];
                    unpack8<T>(v0, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                    unpack8<T>(v1, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                    unpack8<T>(v2, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                    unpack8<T>(v3, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                }
                for (; r < r1; r += RL) {
                    const uint4 v = *reinterpret_cast<const uint4*>(ptr);
                    ptr += step;
                    if (use_cache) { cp[0] = v; cp += cstep; }
                    float f[8];
                    unpack8<T>(v, f);
#pragma unroll
                    for (int i = 0; i < 8; ++i) { s[i] += f[i]; q[i] = fmaf(f[i], f[i], q[i]); }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) { s_sum[rl * C + c + i] = s[i]; s_sq[rl * C + c + i] = q[i]; }
            }
        }
        __syncthreads();
        gn_block_moments(s_sum, s_sq, RL, C, cpg, p.groups, r1 - r0, s_item);
        __syncthreads();
        const long long pidx = ((static_cast<long long>(n) * GN_SHARDS + p.shard_base + shard) * p.sub + sb) * p.groups;
        for (int t = threadIdx.x; t < p.groups * x.world; t += GN_THREADS) {
            const int peer = t / p.groups, g = t - peer * p.groups;
            const float4 m = s_item[g];
            st_ll16(x.part_peers[peer] + pidx + g, make_uint4(__float_as_uint(m.y), tick, __float_as_uint(m.z), tick));
        }
        __syncthreads();
        cached += r1 - r0;
    }

    const T* gamma = reinterpret_cast<const T*>(p.gamma);
    const T* beta = reinterpret_cast<const T*>(p.beta);
    uint4 h_gamma = make_uint4(0, 0, 0, 0), h_beta = make_uint4(0, 0, 0, 0);
    if (passes == 1 && threadIdx.x / ncol < RL) {
        const int c_h = (threadIdx.x % ncol) * 8;
        if (gamma) h_gamma = __ldg(reinterpret_cast<const uint4*>(gamma + c_h));
        if (beta) h_beta = __ldg(reinterpret_cast<const uint4*>(beta + c_h));
    }

    // ---------------- phase 2 + 3: poll + merge all 8 shards in canonical order, normalise, push the edge rows ----------------
    const T* scale = reinterpret_cast<const T*>(p.scale);
    const T* shift = reinterpret_cast<const T*>(p.shift);
    const long long y_bs = p.y_bs > 0 ? p.y_bs : static_cast<long long>(p.hw) * C;
    const int lo_edge = x.halo_rows, hi_edge = p.hw - x.halo_rows;
    const uint4* part_local = x.part_peers[x.rank];
    int stats_n = -1;
    cached = 0;
    for (int item = blockIdx.x; item < total_items; item += gridDim.x) {
        int n, shard, sb, r0, r1;
        item_rows(item, n, shard, sb, r0, r1);
        if (n != stats_n) {
            __syncthreads();
            const uint4* part = part_local + static_cast<long long>(n) * GN_SHARDS * p.sub * p.groups;
            for (int idx = threadIdx.x; idx < GN_SHARDS * p.groups; idx += GN_THREADS) {
                const int sh = idx / p.groups, g = idx % p.groups;
                GnMoments acc{0.f, 0.f, 0.f};
                for (int sb0 = 0; sb0 < p.sub; sb0 += 8) {           // eight loads in flight, merged in index order
                    uint4 e[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (sb0 + u < p.sub) e[u] = ld_ll16(&part[(sh * p.sub + sb0 + u) * p.groups + g]);
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        if (sb0 + u >= p.sub) break;
                        const int sbi = sb0 + u;
                        if (!x.dry && (e[u].y != tick || e[u].w != tick))
                            e[u] = gn_poll_entry(&part[(sh * p.sub + sbi) * p.groups + g], tick, false);
                        // the writer's count: rows of sub-block sbi of a shard, times channels per group (gn_block_moments)
                        const int rows = static_cast<int>((static_cast<long long>(sbi + 1) * shard_len) / p.sub) -
                                         static_cast<int>((static_cast<long long>(sbi) * shard_len) / p.sub);
                        acc = gn_merge(acc, GnMoments{static_cast<float>(rows) * cpg, __uint_as_float(e[u].x), __uint_as_float(e[u].z)});
                    }
                }
                shard_mom[sh][g] = acc;
            }
            __syncthreads();
            if (threadIdx.x < p.groups) {
                const int g = threadIdx.x;
                GnMoments acc = shard_mom[0][g];
#pragma unroll
                for (int sh = 1; sh < GN_SHARDS; ++sh) acc = gn_merge(acc, shard_mom[sh][g]);
                const float var = acc.m2 / acc.cnt;
                s_stats[g] = make_float2(acc.mean, rsqrtf(var + p.eps));
            }
            __syncthreads();
            stats_n = n;
        }
        T* y = reinterpret_cast<T*>(p.y) + static_cast<long long>(n) * y_bs;
        T* y_up = x.up_halo ? reinterpret_cast<T*>(x.up_halo) + static_cast<long long>(n) * y_bs : nullptr;
        T* y_dn = x.down_halo ? reinterpret_cast<T*>(x.down_halo) + static_cast<long long>(n) * y_bs : nullptr;
        for (int pass = 0; pass < passes; ++pass) {
            int vc, rl;
            bool active;
            if (passes == 1) { vc = threadIdx.x % ncol; rl = threadIdx.x / ncol; active = rl < RL; }
            else { vc = pass * GN_THREADS + threadIdx.x; rl = 0; active = vc < ncol; }
            if (!active) continue;
            const int c = vc * 8;
            const T* src; int ld, cc;
            if (c < p.c1) { src = reinterpret_cast<const T*>(p.x) + static_cast<long long>(n) * p.hw * p.c1; ld = p.c1; cc = c; }
            else { src = reinterpret_cast<const T*>(p.x2) + static_cast<long long>(n) * p.hw * p.c2; ld = p.c2; cc = c - p.c1; }
            float a[8], b[8];
            {
                float ga[8], be[8], sc[8], sh[8];
                if (gamma) unpack8<T>(passes == 1 ? h_gamma : __ldg(reinterpret_cast<const uint4*>(gamma + c)), ga);
                if (beta) unpack8<T>(passes == 1 ? h_beta : __ldg(reinterpret_cast<const uint4*>(beta + c)), be);
                if (scale) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(scale + static_cast<long long>(n) * p.ss + c)), sc);
                if (shift) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(shift + static_cast<long long>(n) * p.ss + c)), sh);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float2 st = s_stats[(c + i) / cpg];
                    float ai = st.y, bi = -st.x * st.y;
                    if (gamma) { ai *= ga[i]; bi *= ga[i]; }
                    if (beta) bi += be[i];
                    if (scale) { ai *= (1.0f + sc[i]); bi *= (1.0f + sc[i]); }
                    if (shift) bi += sh[i];
                    a[i] = ai; b[i] = bi;
                }
            }
            for (int r = r0 + rl; r < r1; r += RL) {
                float f[8], o[8];
                const uint4 v = use_cache ? cache[static_cast<long long>(cached + (r - r0)) * ncol + vc]
                                          : *reinterpret_cast<const uint4*>(src + static_cast<long long>(r) * ld + cc);
                unpack8<T>(v, f);
                if (p.act == ACT_SILU) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = act_silu(fmaf(f[i], a[i], b[i]));
                } else if (p.act == ACT_NONE) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = fmaf(f[i], a[i], b[i]);
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) o[i] = apply_act(fmaf(f[i], a[i], b[i]), p.act);
                }
                const uint4 packed = pack8<T>(o);
                *reinterpret_cast<uint4*>(y + static_cast<long long>(r) * C + c) = packed;
                if (y_up && r < lo_edge) *reinterpret_cast<uint4*>(y_up + static_cast<long long>(r) * C + c) = packed;
                if (y_dn && r >= hi_edge) *reinterpret_cast<uint4*>(y_dn + static_cast<long long>(r - hi_edge) * C + c) = packed;
            }
        }
        cached += r1 - r0;
    }
    if (x.self_tick) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned int ticket = atomicAdd(x.self_ticket, 1u);
            if (ticket == gridDim.x - 1) {        // every CTA has read what it needs: the next launch may use tick + 1
                *x.self_ticket = 0;
                *x.self_tick = tick;
            }
        }
    }
    if (x.done_local && !x.dry) {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();       // this CTA's rows (and any it stored into a neighbour) before its arrival counts
            const unsigned int ticket = atomicAdd(x.done_local + 8, 1u);
            if (ticket == gridDim.x - 1) {
                x.done_local[8] = 0;
                __threadfence_system();
                for (int d = 0; d < x.world; ++d)
                    if (x.done_sites[d]) st_release_sys_u32(x.done_sites[d] + x.rank, tick);
            }
        }
    }
}

// ---- LayerNorm: one warp per row, row held in registers -----------------------------------------
template <typename T, int MAXV>
__global__ void __launch_bounds__(256) layernorm_kernel(const T* __restrict__ x, const T* __restrict__ gamma,
                                                        const T* __restrict__ beta, T* __restrict__ y, int rows,
                                                        int cols, float eps) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    pdl_wait();
    pdl_launch_dependents();
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const int nvec = cols / 8;
    const T* xr = x + static_cast<long long>(row) * cols;
    float v[MAXV][8];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < MAXV; ++j) {
        const int vi = lane + 32 * j;
        if (vi < nvec) {
            unpack8<T>(*reinterpret_cast<const uint4*>(xr + vi * 8), v[j]);
#pragma unroll
            for (int i = 0; i < 8; ++i) s += v[j][i];
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    const float mean = s / cols;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < MAXV; ++j) {
        const int vi = lane + 32 * j;
        if (vi < nvec) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { const float d = v[j][i] - mean; q = fmaf(d, d, q); }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    const float rstd = rsqrtf(q / cols + eps);
    T* yr = y + static_cast<long long>(row) * cols;
#pragma unroll
    for (int j = 0; j < MAXV; ++j) {
        const int vi = lane + 32 * j;
        if (vi < nvec) {
            float ga[8], be[8], o[8];
            if (gamma) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(gamma + vi * 8)), ga);
            if (beta) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(beta + vi * 8)), be);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float t = (v[j][i] - mean) * rstd;
                if (gamma) t *= ga[i];
                if (beta) t += be[i];
                o[i] = t;
            }
            *reinterpret_cast<uint4*>(yr + vi * 8) = pack8<T>(o);
        }
    }
}

}  // namespace ab

extern "C" size_t ab_groupnorm_partials_bytes(int n, int groups) {
    return ab::gn_counters_bytes(n) + ab::gn_stats_bytes(n, groups) +
           static_cast<size_t>(n) * ab::GN_SHARDS * ab::GN_MAX_SUB * groups * sizeof(double2);
}

extern "C" int ab_groupnorm(const ab_groupnorm_args* a, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->x && a->y && a->partials, "null pointer (x, y and partials are required)");
    const int C = a->c1 + a->c2;
    AB_REQUIRE(a->n > 0 && a->hw > 0 && a->c1 > 0, "bad shape");
    AB_REQUIRE(a->c1 % 8 == 0 && a->c2 % 8 == 0, "channel counts must be multiples of 8");
    AB_REQUIRE(a->groups > 0 && a->groups <= GN_MAX_GROUPS && C % a->groups == 0, "bad group count (1..64, dividing C)");
    AB_REQUIRE(a->c2 == 0 || a->x2, "x2 missing");
    AB_REQUIRE(a->hw >= GN_SHARDS || a->hw > 0, "bad hw");
    GnParams p{};
    p.x = a->x; p.x2 = a->x2; p.y = a->y; p.gamma = a->gamma; p.beta = a->beta; p.scale = a->scale; p.shift = a->shift;
    uint8_t* ws = reinterpret_cast<uint8_t*>(a->partials);
    p.counters = reinterpret_cast<unsigned int*>(ws);
    p.stats = reinterpret_cast<float2*>(ws + gn_counters_bytes(a->n));
    p.partials = reinterpret_cast<float4*>(ws + gn_counters_bytes(a->n) + gn_stats_bytes(a->n, a->groups));
    p.n = a->n; p.hw = a->hw; p.c1 = a->c1; p.c2 = a->c2; p.groups = a->groups;
    p.sub = gn_sub_blocks(a->hw);
    p.eps = a->eps; p.act = a->act; p.ss = a->ss_stride > 0 ? a->ss_stride : C;
    const int ncol = C / 8;
    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const size_t smem = static_cast<size_t>(2) * RL * C * sizeof(float);
    AB_REQUIRE(smem <= 48 * 1024, "channel count too large for the stats kernel");
    static const bool fused_on = [] { const char* e = getenv("AB_GN_FUSED"); return !(e && e[0] == '0'); }();
    // The one-launch kernel pays off while a CTA's rows fit its shared memory (tensor read once, one CTA per SM is
    // enough). Larger tensors (video latents, VAE top levels) are bandwidth-bound and want every SM oversubscribed:
    // they keep the two-kernel path (ZeroScope GN: 360 us fused without cache vs ~2x the copy time with two kernels).
    const int n_sm = ab_sm_count();
    const int total_items = a->n * GN_SHARDS * p.sub;
    const int grid_f = total_items < n_sm ? total_items : n_sm;
    const int items_per_cta = (total_items + grid_f - 1) / grid_f;
    const int max_item_rows = ((a->hw + GN_SHARDS - 1) / GN_SHARDS + p.sub - 1) / p.sub + 1;
    const long long need_rows = static_cast<long long>(items_per_cta) * max_item_rows;
    const long long cap_rows = (200LL * 1024 - static_cast<long long>(smem)) / (static_cast<long long>(C) * 2);
    // Flag-in-data form of the one-launch kernel (gn_patch_kernel with a single rank): every CTA stores its moments as
    // self-validating entries and polls the ones it needs, which replaces the grid barrier between the statistics and
    // the apply phase. Same items, same arithmetic, same merge order: the bytes are gn_fused_kernel's.
    // Measured on the AnythingV3 task: 126.96 ms with it, 126.54 ms with the grid-barrier kernel — the barrier was not the
    // cost — so it is opt-in (AB_GN_LL=1); the multi-GPU form (ab_groupnorm_patch) is where the flag-in-data exchange pays.
    static const bool ll_on = [] { const char* e = getenv("AB_GN_LL"); return e && e[0] == '1'; }();
    if (fused_on && ll_on && need_rows <= cap_rows && a->hw % GN_SHARDS == 0 && a->n <= 56) {
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(gn_patch_kernel<__half>), 200 * 1024)) return -1;
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(gn_patch_kernel<__nv_bfloat16>), 200 * 1024)) return -1;
        GnPatch x{};
        x.part_peers[0] = reinterpret_cast<uint4*>(p.partials);
        x.world = 1; x.rank = 0;
        x.self_ticket = p.counters + a->n + 2;
        x.self_tick = p.counters + a->n + 3;
        p.ns_local = GN_SHARDS; p.shard_base = 0;
        const int cache_rows = static_cast<int>(need_rows);
        const size_t smem_f = smem + static_cast<size_t>(cache_rows) * C * 2;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid_f);
        cfg.blockDim = dim3(GN_THREADS);
        cfg.dynamicSmemBytes = smem_f;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (a->dtype == AB_BF16) AB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gn_patch_kernel<__nv_bfloat16>, p, x, cache_rows));
        else AB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gn_patch_kernel<__half>, p, x, cache_rows));
        ab_count_launch(1);
        return 0;
    }
    if (fused_on && need_rows <= cap_rows) {
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(gn_fused_kernel<__half>), 200 * 1024)) return -1;
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(gn_fused_kernel<__nv_bfloat16>), 200 * 1024)) return -1;
        const int cache_rows = static_cast<int>(need_rows);
        const size_t smem_f = smem + static_cast<size_t>(cache_rows) * C * 2;
        unsigned int* bar = p.counters + a->n;
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(grid_f);
        cfg.blockDim = dim3(GN_THREADS);
        cfg.dynamicSmemBytes = smem_f;
        cfg.stream = stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeCooperative;
        attr[0].val.cooperative = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        if (a->dtype == AB_BF16) AB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gn_fused_kernel<__nv_bfloat16>, p, bar, cache_rows));
        else AB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gn_fused_kernel<__half>, p, bar, cache_rows));
        ab_count_launch(1);
        return 0;
    }
    dim3 grid(p.sub, GN_SHARDS, a->n);
    const int apply_cols = ncol < GN_APPLY_THREADS ? ncol : GN_APPLY_THREADS;
    p.apply_nb = static_cast<long long>(a->n) * a->hw >= 65536 ? 4 : 1;   // elementwise: free to depend on n
    const int apply_rows = (GN_APPLY_THREADS / apply_cols) * GN_APPLY_ROWS * p.apply_nb;
    dim3 blocks((a->hw + apply_rows - 1) / apply_rows, a->n);
    if (a->dtype == AB_BF16) {
        AB_CHECK_CUDA(launch_pdl(gn_stats_kernel<__nv_bfloat16>, grid, dim3(GN_THREADS), smem, stream, p));
        AB_CHECK_CUDA(launch_pdl(gn_apply_kernel<__nv_bfloat16>, blocks, dim3(GN_APPLY_THREADS), 0, stream, p));
    } else {
        AB_CHECK_CUDA(launch_pdl(gn_stats_kernel<__half>, grid, dim3(GN_THREADS), smem, stream, p));
        AB_CHECK_CUDA(launch_pdl(gn_apply_kernel<__half>, blocks, dim3(GN_APPLY_THREADS), 0, stream, p));
    }
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(2);
    return 0;
}

// Sharded GroupNorm (one sample's rows spread over 2, 4 or 8 GPUs, e.g. video frames): two entry points with the GPUs'
// barrier between them.
//   ab_groupnorm_sharded_stats : partial moments of this GPU's canonical shards -> `partials_out` (peer-visible memory,
//                                layout [n][8][sub][groups] float4, only this GPU's shard rows are written)
//   ab_groupnorm_sharded_apply : merge all GPUs' partials (`part_srcs[i]` = array of the GPU owning shard i) in canonical
//                                order, then normalise the local rows. `a->partials` is the usual local workspace.
// a->hw is the LOCAL row count; hw_global = a->hw * world fixes the canonical sub-block count.
extern "C" int ab_groupnorm_sharded_stats(const ab_groupnorm_args* a, int world, int rank, void* partials_out, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->x && partials_out, "null pointer");
    AB_REQUIRE(world == 2 || world == 4 || world == 8, "world must be 2, 4 or 8");
    const int C = a->c1 + a->c2;
    AB_REQUIRE(a->n > 0 && a->hw > 0 && a->c1 > 0 && a->c1 % 8 == 0 && a->c2 % 8 == 0, "bad shape");
    AB_REQUIRE(a->groups > 0 && a->groups <= GN_MAX_GROUPS && C % a->groups == 0, "bad group count");
    const int ns = GN_SHARDS / world;
    AB_REQUIRE(a->hw % ns == 0, "local rows must split evenly into this GPU's canonical shards");
    GnParams p{};
    p.x = a->x; p.x2 = a->x2;
    p.partials = reinterpret_cast<float4*>(partials_out);
    p.n = a->n; p.hw = a->hw; p.c1 = a->c1; p.c2 = a->c2; p.groups = a->groups;
    p.sub = gn_sub_blocks(a->hw * world);
    p.ns_local = ns; p.shard_base = rank * ns;
    const int ncol = C / 8;
    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const size_t smem = static_cast<size_t>(2) * RL * C * sizeof(float);
    AB_REQUIRE(smem <= 48 * 1024, "channel count too large for the stats kernel");
    dim3 grid(p.sub, ns, a->n);
    if (a->dtype == AB_BF16) AB_CHECK_CUDA(launch_pdl(gn_stats_kernel<__nv_bfloat16>, grid, dim3(GN_THREADS), smem, stream, p));
    else AB_CHECK_CUDA(launch_pdl(gn_stats_kernel<__half>, grid, dim3(GN_THREADS), smem, stream, p));
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_groupnorm_sharded_apply(const ab_groupnorm_args* a, int world, void* const* part_srcs, long long y_batch_stride,
                                          void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->x && a->y && a->partials && part_srcs, "null pointer");
    AB_REQUIRE(world == 2 || world == 4 || world == 8, "world must be 2, 4 or 8");
    const int C = a->c1 + a->c2;
    AB_REQUIRE(a->groups > 0 && a->groups <= GN_MAX_GROUPS && C % a->groups == 0, "bad group count");
    GnParams p{};
    p.x = a->x; p.x2 = a->x2; p.y = a->y; p.gamma = a->gamma; p.beta = a->beta; p.scale = a->scale; p.shift = a->shift;
    uint8_t* ws = reinterpret_cast<uint8_t*>(a->partials);
    p.stats = reinterpret_cast<float2*>(ws + gn_counters_bytes(a->n));
    p.n = a->n; p.hw = a->hw; p.c1 = a->c1; p.c2 = a->c2; p.groups = a->groups;
    p.sub = gn_sub_blocks(a->hw * world);
    p.eps = a->eps; p.act = a->act; p.ss = a->ss_stride > 0 ? a->ss_stride : C;
    p.y_bs = y_batch_stride;
    GnShardSrcs srcs;
    const int ns = GN_SHARDS / world;
    for (int sh = 0; sh < GN_SHARDS; ++sh) {
        AB_REQUIRE(part_srcs[sh / ns], "null partials source");
        srcs.part[sh] = reinterpret_cast<const float4*>(part_srcs[sh / ns]);
    }
    gn_finalize_sharded_kernel<<<a->n, GN_THREADS, 0, stream>>>(p, srcs);
    AB_CHECK_CUDA(cudaGetLastError());
    const int ncol = C / 8;
    const int apply_cols = ncol < GN_APPLY_THREADS ? ncol : GN_APPLY_THREADS;
    p.apply_nb = static_cast<long long>(a->n) * a->hw >= 65536 ? 4 : 1;
    const int apply_rows = (GN_APPLY_THREADS / apply_cols) * GN_APPLY_ROWS * p.apply_nb;
    dim3 blocks((a->hw + apply_rows - 1) / apply_rows, a->n);
    if (a->dtype == AB_BF16) AB_CHECK_CUDA(launch_pdl(gn_apply_kernel<__nv_bfloat16>, blocks, dim3(GN_APPLY_THREADS), 0, stream, p));
    else AB_CHECK_CUDA(launch_pdl(gn_apply_kernel<__half>, blocks, dim3(GN_APPLY_THREADS), 0, stream, p));
    ab_count_launch(2);
    return 0;
}

extern "C" int ab_groupnorm_patch(const ab_groupnorm_args* a, const ab_gn_patch_args* pa, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && pa && a->x && a->y, "null pointer");
    AB_REQUIRE(pa->world == 2 || pa->world == 4 || pa->world == 8, "world must be 2, 4 or 8");
    AB_REQUIRE(pa->rank >= 0 && pa->rank < pa->world, "bad rank");
    const int C = a->c1 + a->c2;
    AB_REQUIRE(a->n > 0 && a->hw > 0 && a->c1 > 0 && a->c1 % 8 == 0 && a->c2 % 8 == 0, "bad shape");
    AB_REQUIRE(a->groups > 0 && a->groups <= GN_MAX_GROUPS && C % a->groups == 0, "bad group count");
    AB_REQUIRE(a->c2 == 0 || a->x2, "x2 missing");
    const int ns = GN_SHARDS / pa->world;
    AB_REQUIRE(a->hw % ns == 0, "local rows must split evenly into this GPU's canonical shards");
    AB_REQUIRE(pa->dry || pa->tick, "tick is required");
    AB_REQUIRE(pa->halo_rows >= 0 && pa->halo_rows <= a->hw, "bad halo row count");
    GnParams p{};
    p.x = a->x; p.x2 = a->x2; p.y = a->y; p.gamma = a->gamma; p.beta = a->beta; p.scale = a->scale; p.shift = a->shift;
    p.n = a->n; p.hw = a->hw; p.c1 = a->c1; p.c2 = a->c2; p.groups = a->groups;
    p.sub = gn_sub_blocks(a->hw * pa->world);
    p.ns_local = ns; p.shard_base = pa->rank * ns;
    p.eps = a->eps; p.act = a->act; p.ss = a->ss_stride > 0 ? a->ss_stride : C;
    p.y_bs = pa->y_batch_stride;
    GnPatch x{};
    for (int r = 0; r < pa->world; ++r) {
        AB_REQUIRE(pa->part_peers[r], "missing partial-moment array");
        x.part_peers[r] = reinterpret_cast<uint4*>(pa->part_peers[r]);
        x.done_sites[r] = pa->done_sites[r];
    }
    x.tick = pa->tick; x.up_halo = pa->up_halo; x.down_halo = pa->down_halo; x.halo_rows = pa->halo_rows;
    x.done_local = pa->done_local; x.world = pa->world; x.rank = pa->rank; x.dry = pa->dry;
    const int ncol = C / 8;
    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const size_t smem = static_cast<size_t>(2) * RL * C * sizeof(float);
    AB_REQUIRE(smem <= 48 * 1024, "channel count too large for the stats kernel");
    // One CTA per SM at most, all co-resident (cooperative launch): a CTA polls for moments that other CTAs of this
    // grid produce, so none of them may be waiting for an SM.
    const int n_sm = ab_sm_count();
    const int total_items = a->n * ns * p.sub;
    const int grid_f = total_items < n_sm ? total_items : n_sm;
    const int items_per_cta = (total_items + grid_f - 1) / grid_f;
    const int max_item_rows = (a->hw / ns + p.sub - 1) / p.sub + 1;
    const long long need_rows = static_cast<long long>(items_per_cta) * max_item_rows;
    const long long cap_rows = (200LL * 1024 - static_cast<long long>(smem)) / (static_cast<long long>(C) * 2);
    const int cache_rows = need_rows <= cap_rows ? static_cast<int>(need_rows) : 0;   // 0: phase 3 re-reads the rows
    if (ab_ensure_max_smem(reinterpret_cast<const void*>(gn_patch_kernel<__half>), 200 * 1024)) return -1;
    if (ab_ensure_max_smem(reinterpret_cast<const void*>(gn_patch_kernel<__nv_bfloat16>), 200 * 1024)) return -1;
    const size_t smem_f = smem + static_cast<size_t>(cache_rows) * C * 2;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid_f);
    cfg.blockDim = dim3(GN_THREADS);
    cfg.dynamicSmemBytes = smem_f;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (a->dtype == AB_BF16) AB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gn_patch_kernel<__nv_bfloat16>, p, x, cache_rows));
    else AB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, gn_patch_kernel<__half>, p, x, cache_rows));
    ab_count_launch(1);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// ab_groupnorm_static: GroupNorm(+activation) for callers that must not allocate device memory.
//
// Who needs it: a host that has already given every free byte of HBM to its own arena (weights,
// activations, KV) and has no allocator left to ask; a C or Rust caller that links this library
// without torch and so has no caching allocator at all; a serving loop that cannot afford the
// device-wide synchronisation a cudaMalloc/cudaFree implies while other tasks share the GPU.
// For them the statistics scratch that ab_groupnorm takes from the caller lives here instead, in
// a fixed array inside the library image. The price: every GPU that loads the library carries the
// array (kStaticScratchBytes, about 0.5 MB) until the process exits, and the problem size is capped
// at kStaticMaxImages images and kStaticMaxGroups groups per call.
//
// `g_gn_scratch` below is only a host-side handle when named in host code; the device address is
// looked up with cudaGetSymbolAddress the first time the entry point runs on a device (with that
// device current) and cached per device. Make that first call on each device outside any stream
// capture. All calls share the one array and it keeps whatever the last call left in it, so the
// block counters are cleared on the stream before every launch, and two calls that overlap on the
// same GPU (different streams, or different host threads) race on it: serialise them.
// ---------------------------------------------------------------------------------------------
#include <mutex>

namespace ab {
constexpr int kStaticMaxImages = 4;
constexpr int kStaticMaxGroups = 32;
constexpr size_t kStaticCountersBytes = 256;                                   // >= 4 B x kStaticMaxImages
constexpr size_t kStaticStatsBytes = 1024;                                     // 8 B x images x groups
constexpr size_t kStaticPartialsBytes =
    static_cast<size_t>(kStaticMaxImages) * GN_SHARDS * GN_MAX_SUB * kStaticMaxGroups * sizeof(double2);
constexpr size_t kStaticScratchBytes = kStaticCountersBytes + kStaticStatsBytes + kStaticPartialsBytes;
static_assert(kStaticCountersBytes >= kStaticMaxImages * sizeof(unsigned int), "counter region too small");
static_assert(kStaticStatsBytes >= kStaticMaxImages * kStaticMaxGroups * sizeof(float2), "stats region too small");
constexpr int kStaticMaxDevices = 64;
}  // namespace ab

static __device__ __align__(256) unsigned char g_gn_scratch[ab::kStaticScratchBytes];

static std::mutex g_gn_static_mu;
static unsigned char* g_gn_static_ptr[ab::kStaticMaxDevices] = {nullptr};

extern "C" int ab_groupnorm_static(const ab_groupnorm_args* a, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->x && a->y, "null pointer (x and y are required)");
    const int C = a->c1 + a->c2;
    AB_REQUIRE(a->n > 0 && a->hw > 0 && a->c1 > 0, "bad shape");
    AB_REQUIRE(a->n <= kStaticMaxImages,
               "ab_groupnorm_static handles at most 4 images per call: split the batch or use ab_groupnorm");
    AB_REQUIRE(a->groups > 0 && a->groups <= kStaticMaxGroups,
               "ab_groupnorm_static handles at most 32 groups: use ab_groupnorm with a caller workspace");
    AB_REQUIRE(a->c1 % 8 == 0 && a->c2 % 8 == 0, "channel counts must be multiples of 8");
    AB_REQUIRE(C % a->groups == 0, "bad group count");
    AB_REQUIRE(a->c2 == 0 || a->x2, "x2 missing");

    // one-time set-up per device: resolve the device address of the scratch array
    int dev = -1;
    AB_CHECK_CUDA(cudaGetDevice(&dev));
    AB_REQUIRE(dev >= 0 && dev < kStaticMaxDevices, "device index out of range");
    unsigned char* ws = nullptr;
    {
        std::lock_guard<std::mutex> lk(g_gn_static_mu);
        if (g_gn_static_ptr[dev] == nullptr) {
            unsigned char* sym = nullptr;
            AB_CHECK_CUDA(cudaGetSymbolAddress((void**)&sym, g_gn_scratch));
            AB_REQUIRE(sym != nullptr, "scratch symbol resolved to a null device address");
            g_gn_static_ptr[dev] = sym;
        }
        ws = g_gn_static_ptr[dev];
    }

    GnParams p{};
    p.x = a->x; p.x2 = a->x2; p.y = a->y; p.gamma = a->gamma; p.beta = a->beta; p.scale = a->scale; p.shift = a->shift;
    p.counters = reinterpret_cast<unsigned int*>(ws);
    p.stats = reinterpret_cast<float2*>(ws + kStaticCountersBytes);
    p.partials = reinterpret_cast<float4*>(ws + kStaticCountersBytes + kStaticStatsBytes);
    p.n = a->n; p.hw = a->hw; p.c1 = a->c1; p.c2 = a->c2; p.groups = a->groups;
    p.sub = gn_sub_blocks(a->hw);
    p.eps = a->eps; p.act = a->act; p.ss = a->ss_stride > 0 ? a->ss_stride : C;
    const int ncol = C / 8;
    const int RL = ncol >= GN_THREADS ? 1 : GN_THREADS / ncol;
    const size_t smem = static_cast<size_t>(2) * RL * C * sizeof(float);
    AB_REQUIRE(smem <= 48 * 1024, "channel count too large for the stats kernel");
    // the block tickets must start at zero whatever an earlier (possibly failed) call left behind
    AB_CHECK_CUDA(cudaMemsetAsync(ws, 0, kStaticCountersBytes, stream));
    dim3 grid(p.sub, GN_SHARDS, a->n);
    const int apply_cols = ncol < GN_APPLY_THREADS ? ncol : GN_APPLY_THREADS;
    p.apply_nb = static_cast<long long>(a->n) * a->hw >= 65536 ? 4 : 1;   // elementwise: free to depend on n
    const int apply_rows = (GN_APPLY_THREADS / apply_cols) * GN_APPLY_ROWS * p.apply_nb;
    dim3 blocks((a->hw + apply_rows - 1) / apply_rows, a->n);
    if (a->dtype == AB_BF16) {
        AB_CHECK_CUDA(launch_pdl(gn_stats_kernel<__nv_bfloat16>, grid, dim3(GN_THREADS), smem, stream, p));
        AB_CHECK_CUDA(launch_pdl(gn_apply_kernel<__nv_bfloat16>, blocks, dim3(GN_APPLY_THREADS), 0, stream, p));
    } else {
        AB_CHECK_CUDA(launch_pdl(gn_stats_kernel<__half>, grid, dim3(GN_THREADS), smem, stream, p));
        AB_CHECK_CUDA(launch_pdl(gn_apply_kernel<__half>, blocks, dim3(GN_APPLY_THREADS), 0, stream, p));
    }
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(2);
    return 0;
}

extern "C" int ab_layernorm(const void* x, const void* gamma, const void* beta, void* y, int rows, int cols, float eps,
                            int dtype, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && rows > 0 && cols > 0, "bad arguments");
    AB_REQUIRE(cols % 8 == 0 && cols <= 32 * 8 * 16, "cols must be a multiple of 8 and <= 4096");
    const int warps = 8;
    const int blocks = (rows + warps - 1) / warps;
    const int nvec = cols / 8;
#define AB_LN(T, MAXV)                                                                                              \
    launch_pdl(layernorm_kernel<T, MAXV>, dim3(blocks), dim3(warps * 32), 0, stream, reinterpret_cast<const T*>(x), \
               reinterpret_cast<const T*>(gamma), reinterpret_cast<const T*>(beta), reinterpret_cast<T*>(y), rows,  \
               cols, eps)
    if (dtype == AB_BF16) {
        if (nvec <= 64) AB_LN(__nv_bfloat16, 2); else if (nvec <= 160) AB_LN(__nv_bfloat16, 5); else if (nvec <= 256) AB_LN(__nv_bfloat16, 8); else AB_LN(__nv_bfloat16, 16);
    } else {
        if (nvec <= 64) AB_LN(__half, 2); else if (nvec <= 160) AB_LN(__half, 5); else if (nvec <= 256) AB_LN(__half, 8); else AB_LN(__half, 16);
    }
#undef AB_LN
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

// ---- row softmax / transpose / embedding (small helpers around the GEMM-based wide-head attention) ----
namespace ab {
template <typename T>
__global__ void __launch_bounds__(256) softmax_rows_kernel(const T* __restrict__ x, T* __restrict__ y, int rows, int cols,
                                                           float scale) {
    __shared__ float red[8];
    const int row = blockIdx.x;
    const T* xr = x + static_cast<long long>(row) * cols;
    T* yr = y + static_cast<long long>(row) * cols;
    const int nvec = cols / 8;
    float mx = -INFINITY;
    for (int v = threadIdx.x; v < nvec; v += 256) {
        float f[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(xr + v * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) mx = fmaxf(mx, f[i] * scale);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
    __syncthreads();
    mx = red[0];
#pragma unroll
    for (int i = 1; i < 8; ++i) mx = fmaxf(mx, red[i]);
    __syncthreads();
    float sum = 0.f;
    for (int v = threadIdx.x; v < nvec; v += 256) {
        float f[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(xr + v * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) sum += __expf(f[i] * scale - mx);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
    __syncthreads();
    sum = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) sum += red[i];     // fixed order
    const float inv = 1.f / sum;
    for (int v = threadIdx.x; v < nvec; v += 256) {
        float f[8], o[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(xr + v * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = __expf(f[i] * scale - mx) * inv;
        *reinterpret_cast<uint4*>(yr + v * 8) = pack8<T>(o);
    }
}

template <typename T>
__global__ void transpose2d_kernel(const T* __restrict__ x, T* __restrict__ y, int rows, int cols) {
    __shared__ T tile[32][33];
    const long long base = static_cast<long long>(blockIdx.z) * rows * cols;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += 8) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[i][threadIdx.x] = x[base + static_cast<long long>(r) * cols + c];
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += 8) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (r < rows && c < cols) y[base + static_cast<long long>(c) * rows + r] = tile[threadIdx.x][i];
    }
}

template <typename T>
__global__ void embedding_kernel(const long long* __restrict__ ids, const T* __restrict__ tok, const T* __restrict__ pos,
                                 T* __restrict__ out, int len, int dim) {
    const int bl = blockIdx.x;
    const int l = bl % len;
    const long long id = ids[bl];
    for (int v = threadIdx.x; v < dim / 8; v += blockDim.x) {
        float a[8], b[8], o[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(tok + id * dim + v * 8), a);
        unpack8<T>(*reinterpret_cast<const uint4*>(pos + static_cast<long long>(l) * dim + v * 8), b);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = a[i] + b[i];
        *reinterpret_cast<uint4*>(out + static_cast<long long>(bl) * dim + v * 8) = pack8<T>(o);
    }
}
}  // namespace ab

extern "C" int ab_softmax_rows(const void* x, void* y, int rows, int cols, float scale, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && rows > 0 && cols % 8 == 0, "bad arguments");
    if (dtype == AB_BF16) ab::softmax_rows_kernel<__nv_bfloat16><<<rows, 256, 0, stream>>>(reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(y), rows, cols, scale);
    else ab::softmax_rows_kernel<__half><<<rows, 256, 0, stream>>>(reinterpret_cast<const __half*>(x), reinterpret_cast<__half*>(y), rows, cols, scale);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_transpose2d(const void* x, void* y, int batch, int rows, int cols, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && rows > 0 && cols > 0 && batch > 0, "bad arguments");
    (void)dtype;   // 16-bit payload either way
    dim3 grid((cols + 31) / 32, (rows + 31) / 32, batch), block(32, 8);
    ab::transpose2d_kernel<__half><<<grid, block, 0, stream>>>(reinterpret_cast<const __half*>(x), reinterpret_cast<__half*>(y), rows, cols);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_embedding(const long long* ids, const void* tok, const void* pos, void* out, int batch, int len, int dim,
                            int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(ids && tok && pos && out && dim % 8 == 0, "bad arguments");
    if (dtype == AB_BF16) ab::embedding_kernel<__nv_bfloat16><<<batch * len, 128, 0, stream>>>(ids, reinterpret_cast<const __nv_bfloat16*>(tok), reinterpret_cast<const __nv_bfloat16*>(pos), reinterpret_cast<__nv_bfloat16*>(out), len, dim);
    else ab::embedding_kernel<__half><<<batch * len, 128, 0, stream>>>(ids, reinterpret_cast<const __half*>(tok), reinterpret_cast<const __half*>(pos), reinterpret_cast<__half*>(out), len, dim);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
