// Cross-GPU hand-offs for patch parallelism (one image's row bands on 2/4/8 GPUs of an NVSwitch box) without
// barrier launches: tick counter, flag publish / wait, and strided copies through peer memory with the flag wait in
// their prologue and the flag publish in their epilogue. The GroupNorm, conv and attention kernels carry the same
// wait / publish inside themselves (norm.cu, gemm_tc.cu, attention_v2.cuh); these small kernels cover the places where
// the neighbour of a hand-off is an elementwise op. SURVEY §2.8 "patch parallelism", §5.9 item 3.
// No counterpart in the reference (single GPU).
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

__global__ void sync_tick_kernel(unsigned int* tick) { *tick = *tick + 1; }

struct PeerSites { unsigned int* site[8]; };

__global__ void sync_signal_kernel(const unsigned int* tick, PeerSites s, int world, int rank) {
    const int t = threadIdx.x;
    if (t < world && s.site[t]) {
        __threadfence_system();
        st_release_sys_u32(s.site[t] + rank, *tick);
    }
}

__global__ void sync_wait_kernel(const unsigned int* tick, const unsigned int* site, unsigned int mask) {
    const int t = threadIdx.x;
    if (t < 8 && ((mask >> t) & 1u)) sync_wait_flags(site, 1u << t, *tick);
}

struct PatchCopyParams {
    ab_copy_seg seg[16];
    int nseg;
    const unsigned int* tick;
    const unsigned int* wait_site;
    unsigned int wait_mask;
    unsigned int* signal_sites[8];
    int world, rank;
    unsigned int* local_site;
};

__global__ void __launch_bounds__(256) patch_copy_kernel(const __grid_constant__ PatchCopyParams p) {
    if (p.wait_site && p.wait_mask) {
        if (threadIdx.x == 0) sync_wait_flags(p.wait_site, p.wait_mask, *p.tick);
        __syncthreads();
    }
    for (int s = 0; s < p.nseg; ++s) {
        const ab_copy_seg& g = p.seg[s];
        const long long w16 = g.width_bytes >> 4, dp = g.dst_pitch >> 4, sp = g.src_pitch >> 4;
        const long long total = w16 * g.rows;
        uint4* dst = reinterpret_cast<uint4*>(g.dst);
        const uint4* src = reinterpret_cast<const uint4*>(g.src);
        const long long stride = static_cast<long long>(gridDim.x) * 256;
        long long i = blockIdx.x * 256LL + threadIdx.x;
        for (; i + 3 * stride < total; i += 4 * stride) {            // four loads in flight per thread
            uint4 v[4];
            long long o[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long long j = i + u * stride, r = j / w16, c = j - r * w16;
                v[u] = src[r * sp + c];
                o[u] = r * dp + c;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) dst[o[u]] = v[u];
        }
        for (; i < total; i += stride) {
            const long long r = i / w16, c = i - r * w16;
            dst[r * dp + c] = src[r * sp + c];
        }
    }
    if (p.local_site) {
        // last CTA publishes: every CTA's stores are fenced to system scope before its arrival is counted
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();
            const unsigned int ticket = atomicAdd(p.local_site + 8, 1u);
            if (ticket == gridDim.x - 1) {
                p.local_site[8] = 0;
                __threadfence_system();
                const unsigned int t = *p.tick;
                for (int d = 0; d < p.world; ++d)
                    if (p.signal_sites[d]) st_release_sys_u32(p.signal_sites[d] + p.rank, t);
            }
        }
    }
}

}  // namespace ab

extern "C" int ab_sync_tick(unsigned int* tick, void* stream_) {
    AB_REQUIRE(tick, "null tick");
    ab::sync_tick_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(tick);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_sync_signal(const unsigned int* tick, unsigned int* const* peer_sites, int world, int rank, void* stream_) {
    AB_REQUIRE(tick && peer_sites && world >= 1 && world <= 8 && rank >= 0 && rank < world, "bad arguments");
    ab::PeerSites s{};
    for (int d = 0; d < world; ++d) s.site[d] = peer_sites[d];
    ab::sync_signal_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(tick, s, world, rank);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_sync_wait(const unsigned int* tick, const unsigned int* site, unsigned int mask, void* stream_) {
    AB_REQUIRE(tick && site && mask < 256, "bad arguments");
    ab::sync_wait_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(tick, site, mask);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_patch_copy(const ab_patch_copy_args* a, void* stream_) {
    using namespace ab;
    AB_REQUIRE(a && a->nseg >= 0 && a->nseg <= 16, "bad segment count");
    AB_REQUIRE((!a->wait_site && !a->local_site) || a->tick, "tick is required with wait or signal");
    AB_REQUIRE(a->world >= 1 && a->world <= 8 && a->rank >= 0 && a->rank < a->world, "bad world / rank");
    PatchCopyParams p{};
    long long total = 0;
    for (int s = 0; s < a->nseg; ++s) {
        const ab_copy_seg& g = a->seg[s];
        AB_REQUIRE(g.dst && g.src && g.rows > 0 && g.width_bytes > 0, "bad segment");
        AB_REQUIRE(g.width_bytes % 16 == 0 && g.dst_pitch % 16 == 0 && g.src_pitch % 16 == 0 &&
                   (reinterpret_cast<uintptr_t>(g.dst) & 15) == 0 && (reinterpret_cast<uintptr_t>(g.src) & 15) == 0,
                   "16-byte alignment required");
        p.seg[s] = g;
        total += g.width_bytes / 16 * g.rows;
    }
    p.nseg = a->nseg; p.tick = a->tick; p.wait_site = a->wait_site; p.wait_mask = a->wait_mask;
    for (int d = 0; d < 8; ++d) p.signal_sites[d] = d < a->world ? a->signal_sites[d] : nullptr;
    p.world = a->world; p.rank = a->rank; p.local_site = a->local_site;
    // 16 x 16 bytes per thread: enough CTAs to keep NVLink busy, few enough that the arrival count stays short
    long long blocks = total / (256 * 16) + 1;
    if (blocks > 148) blocks = 148;
    patch_copy_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(p);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

// ---- hand-off latency probe (profiles/: the numbers the patch-parallel design is sized by) --------------------------
// Two GPUs bounce a flag `iters` times. mode 0: flag only (st.release.sys / ld.acquire.sys). mode 1: a 16-byte
// flag-in-data entry (volatile v4 store / poll). mode 2: `bytes` of plain stores into the peer's buffer by the whole
// CTA, __threadfence_system, then the flag — the shape of a halo push. One CTA per GPU; rank 0 serves first.
// out[0] = cycles for all iterations (clock64 of this GPU), out[1] = iterations.
namespace ab {
__global__ void __launch_bounds__(256) pingpong_kernel(unsigned int* my_flags, unsigned int* peer_flags, uint4* peer_buf,
                                                       int rank, int iters, int mode, int bytes, unsigned int base,
                                                       long long* out) {
    __shared__ long long t0;
    if (threadIdx.x == 0) t0 = clock64();
    __syncthreads();
    for (int i = 1; i <= iters; ++i) {
        const unsigned int v = base + static_cast<unsigned int>(i);
        for (int turn = 0; turn < 2; ++turn) {
            if (turn == rank) {               // my turn to send
                if (mode == 2) {
                    for (int j = threadIdx.x; j < bytes / 16; j += 256) peer_buf[j] = make_uint4(v, v, v, v);
                    __threadfence_system();
                    __syncthreads();
                }
                if (threadIdx.x == 0) {
                    if (mode == 1) {
                        asm volatile("st.volatile.global.v4.u32 [%0], {%1, %1, %1, %1};" ::"l"(peer_flags + 16), "r"(v) : "memory");
                    } else {
                        st_release_sys_u32(peer_flags + rank, v);
                    }
                }
            } else {                          // wait for the peer's send
                if (threadIdx.x == 0) {
                    if (mode == 1) {
                        unsigned int a, b, c, d;
                        do {
                            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(my_flags + 16) : "memory");
                        } while (b != v || d != v);
                    } else {
                        while (static_cast<int>(ld_acquire_sys_u32(my_flags + (1 - rank)) - v) < 0) {}
                    }
                }
                __syncthreads();
            }
        }
    }
    if (threadIdx.x == 0) { out[0] = clock64() - t0; out[1] = iters; }
}
}  // namespace ab

extern "C" int ab_sync_pingpong(unsigned int* my_flags, unsigned int* peer_flags, void* peer_buf, int rank, int iters,
                                int mode, int bytes, unsigned int base, long long* out, void* stream_) {
    AB_REQUIRE(my_flags && peer_flags && out && (rank == 0 || rank == 1) && iters > 0 && bytes % 16 == 0, "bad arguments");
    ab::pingpong_kernel<<<1, 256, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(my_flags, peer_flags, reinterpret_cast<uint4*>(peer_buf),
                                                                                 rank, iters, mode, bytes, base, out);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

// ---- NVLS broadcast: one multimem.st per 16 bytes, the NVSwitch replicates it into every bound GPU's copy ------------
// dst_mc is the MULTICAST mapping of a buffer (parallel/nvls.py). Stores only: the switch copies, it never reduces.
namespace ab {
__global__ void __launch_bounds__(256) multimem_copy_kernel(float4* __restrict__ dst_mc, const float4* __restrict__ src,
                                                            long long n16) {
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n16; i += static_cast<long long>(gridDim.x) * 256) {
        const float4 v = src[i];
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst_mc + i), "f"(v.x), "f"(v.y),
                     "f"(v.z), "f"(v.w)
                     : "memory");
    }
}
}  // namespace ab

extern "C" int ab_multimem_copy(void* dst_mc, const void* src, long long nbytes, void* stream_) {
    AB_REQUIRE(dst_mc && src && nbytes > 0 && nbytes % 16 == 0, "bad arguments");
    AB_REQUIRE((reinterpret_cast<uintptr_t>(dst_mc) & 15) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0, "16-byte alignment required");
    const long long n16 = nbytes / 16;
    long long blocks = n16 / (256 * 8) + 1;
    if (blocks > 148) blocks = 148;
    ab::multimem_copy_kernel<<<static_cast<int>(blocks), 256, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(
        reinterpret_cast<float4*>(dst_mc), reinterpret_cast<const float4*>(src), n16);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
