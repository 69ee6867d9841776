// Peer-memory runtime for one NVSwitch box: a per-rank heap that every GPU of a subset maps into its own
// address space (CUDA IPC), and a device-side barrier over peer-mapped flags. SURVEY §5.9 / §8 row R1.
// Kernels then address remote tiles directly (TMA or ld/st through the peer mapping); no host round trip and no
// NCCL call sits between a compute step and the data it needs from another GPU.
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

// Heap memory comes straight from cudaMalloc (IPC handles cover whole allocations).
extern "C" int ab_heap_alloc(void** ptr, size_t bytes) {
    AB_REQUIRE(ptr && bytes > 0, "bad arguments");
    AB_CHECK_CUDA(cudaMalloc(ptr, bytes));
    AB_CHECK_CUDA(cudaMemset(*ptr, 0, bytes));
    return 0;
}
extern "C" int ab_heap_free(void* ptr) {
    AB_CHECK_CUDA(cudaFree(ptr));
    return 0;
}
extern "C" int ab_ipc_get_handle(void* ptr, unsigned char* handle64) {
    cudaIpcMemHandle_t h;
    AB_CHECK_CUDA(cudaIpcGetMemHandle(&h, ptr));
    static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
    memcpy(handle64, &h, 64);
    return 0;
}
extern "C" int ab_ipc_open_handle(const unsigned char* handle64, void** ptr) {
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    AB_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return 0;
}
extern "C" int ab_ipc_close_handle(void* ptr) {
    AB_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
    return 0;
}

namespace ab {
struct PeerFlags { unsigned int* flags[8]; };   // flags[r] = rank r's flag array (peer-mapped), [8] slots each

// Rank `rank` publishes `epoch` into slot [rank] of every peer's flag array, then waits until every peer has
// published the same epoch into its own array. System-scope release/acquire orders the data written before.
__global__ void peer_barrier_kernel(PeerFlags f, int rank, int world, unsigned int epoch) {
    const int t = threadIdx.x;
    if (t < world) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.flags[t] + rank), "r"(epoch) : "memory");
        long long t0 = clock64();
        unsigned int v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f.flags[rank] + t) : "memory");
            if (clock64() - t0 > 120000000000LL) { printf("peer barrier timeout rank %d waiting for %d\n", rank, t); __trap(); }
        } while (static_cast<int>(v - epoch) < 0);
    }
}
// Replayable form for CUDA graphs: the epoch is not a launch argument (a replayed graph would present the same value
// again and fall straight through) but a counter in this rank's own flag page, advanced by the kernel itself. Every
// rank executes the same sequence of these barriers, so the counters stay equal. It uses its own flag words
// (u32 [32..40) of the page, counter at [64]) and must not be mixed with the host-epoch form on the same words.
__global__ void peer_barrier_auto_kernel(PeerFlags f, int rank, int world) {
    const int t = threadIdx.x;
    unsigned int epoch = 0;
    if (t == 0) {
        unsigned int* cnt = f.flags[rank] + 64;
        epoch = *cnt + 1;
        *cnt = epoch;
    }
    epoch = __shfl_sync(0xffffffffu, epoch, 0);
    if (t < world) {
        __threadfence_system();
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.flags[t] + 32 + rank), "r"(epoch) : "memory");
        long long t0 = clock64();
        unsigned int v;
        do {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f.flags[rank] + 32 + t) : "memory");
            if (clock64() - t0 > 120000000000LL) { printf("peer barrier (auto) timeout rank %d waiting for %d\n", rank, t); __trap(); }
        } while (static_cast<int>(v - epoch) < 0);
    }
}
}  // namespace ab

extern "C" int ab_peer_barrier_auto(void* const* flag_ptrs, int rank, int world, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(flag_ptrs && world >= 1 && world <= 8 && rank >= 0 && rank < world, "bad arguments");
    ab::PeerFlags f{};
    for (int r = 0; r < world; ++r) f.flags[r] = reinterpret_cast<unsigned int*>(flag_ptrs[r]);
    ab::peer_barrier_auto_kernel<<<1, 32, 0, stream>>>(f, rank, world);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_peer_barrier(void* const* flag_ptrs, int rank, int world, unsigned int epoch, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(flag_ptrs && world >= 1 && world <= 8 && rank >= 0 && rank < world, "bad arguments");
    ab::PeerFlags f{};
    for (int r = 0; r < world; ++r) f.flags[r] = reinterpret_cast<unsigned int*>(flag_ptrs[r]);
    ab::peer_barrier_kernel<<<1, 32, 0, stream>>>(f, rank, world, epoch);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

// ---- strided copy between this GPU's memory and a peer's (halo rows of a sharded tensor) ------------------------
// rows x width_bytes, 16 bytes per thread and access; src or dst may be a peer-mapped address. src == nullptr writes
// zeros (the sequence ends, where the conv's padding is zero). Plain loads/stores: graph-capturable, stream-ordered.
namespace ab {
__global__ void __launch_bounds__(256) peer_copy2d_kernel(uint4* __restrict__ dst, long long dst_pitch16,
                                                          const uint4* __restrict__ src, long long src_pitch16,
                                                          long long width16, int rows) {
    const long long total = width16 * rows;
    for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * 256) {
        const long long r = i / width16, c = i - r * width16;
        dst[r * dst_pitch16 + c] = src ? src[r * src_pitch16 + c] : make_uint4(0, 0, 0, 0);
    }
}
}  // namespace ab

extern "C" int ab_peer_copy2d(void* dst, long long dst_pitch, const void* src, long long src_pitch, long long width_bytes,
                              int rows, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(dst && rows > 0 && width_bytes > 0, "bad arguments");
    AB_REQUIRE(width_bytes % 16 == 0 && dst_pitch % 16 == 0 && src_pitch % 16 == 0 &&
               (reinterpret_cast<uintptr_t>(dst) & 15) == 0 && (reinterpret_cast<uintptr_t>(src) & 15) == 0,
               "16-byte alignment required");
    const long long total = width_bytes / 16 * rows;
    const int blocks = static_cast<int>(total / 256 + 1 < 148 * 8 ? total / 256 + 1 : 148 * 8);
    ab::peer_copy2d_kernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<uint4*>(dst), dst_pitch / 16,
                                                       reinterpret_cast<const uint4*>(src), src_pitch / 16,
                                                       width_bytes / 16, rows);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
