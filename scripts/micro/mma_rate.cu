// Micro-benchmark: cycles per tcgen05.mma (M=128, K=16, fp16) issued back to back by one elected lane, operands
// resident in shared memory (no TMA): SS form (A and B from shared memory, K-major, 128-byte swizzle) and TS form
// (A from tensor memory), N = 32..256.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I csrc -o /tmp/mma_rate scripts/micro/mma_rate.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda.h>
extern "C" void ab_set_error(const char*) {}
extern "C" int ab_ensure_max_smem(const void*, int) { return 0; }
extern "C" int ab_sm_count(void) { return 148; }
extern "C" int ab_pdl_enabled(void) { return 0; }
#include "common.cuh"
using namespace ab;

template <int N, bool TS, bool SAME>
__global__ void __launch_bounds__(128) k(int iters, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (64 + 32) * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (warp == 0) tmem_alloc<512>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = slot;
    if (warp == 0) {
        const bool leader = elect_one();
        constexpr uint32_t idesc = make_idesc_f16(128, N, false, false, false);
        const uint64_t d0 = make_smem_desc(0, 16, 1024, kSwizzle128);
        const uint32_t sa = smem_u32(smem), sb = sa + 64 * 1024;     // A: 4 stages x 16 KB, B: 32 KB
        long long t0 = clock64();
        if (leader) {
            for (int i = 0; i < iters; ++i) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint32_t a = sa + (SAME ? 0 : (i & 3) * 16384) + ks * 32;
                    if (TS) umma_ts(tmem_base, tmem_base + 256 + ks * 8, smem_desc_at(d0, sb + ks * 32), idesc, 1u);
                    else umma_ss(tmem_base, smem_desc_at(d0, a), smem_desc_at(d0, sb + ks * 32), idesc, 1u);
                }
            }
            umma_commit(&bar);
        }
        __syncwarp();
        long long t1 = clock64();
        mbar_wait(&bar, 0);
        long long t2 = clock64();
        if (leader) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc<512>(tmem_base);
}

template <int N, bool TS, bool SAME>
void run(const char* name, long long* out) {
    const int iters = 512, smem = 100 * 1024;
    cudaFuncSetAttribute(k<N, TS, SAME>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    k<N, TS, SAME><<<1, 128, smem>>>(iters, out);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: error %s\n", name, cudaGetErrorString(e)); return; }
    printf("%-34s N=%3d: issue %.1f cycles/MMA, complete %.1f cycles/MMA (%.0f%% of the N/2 floor rate)\n", name, N,
           (double)out[0] / (4 * iters), (double)out[1] / (4 * iters), 100.0 * (N / 2.0) / ((double)out[1] / (4 * iters)));
}

int main() {
    long long* out;
    cudaMallocManaged(&out, 16);
    run<32, false, false>("SS, A rotates over 4 stages", out);
    run<64, false, false>("SS, A rotates over 4 stages", out);
    run<128, false, false>("SS, A rotates over 4 stages", out);
    run<256, false, false>("SS, A rotates over 4 stages", out);
    run<64, false, true>("SS, same A tile", out);
    run<256, false, true>("SS, same A tile", out);
    run<32, true, false>("TS (A in tensor memory)", out);
    run<64, true, false>("TS (A in tensor memory)", out);
    run<128, true, false>("TS (A in tensor memory)", out);
    run<256, true, false>("TS (A in tensor memory)", out);
    return 0;
}
