// Micro-benchmark: tcgen05.ld (32x32b.x32) throughput with 1, 2, 4 (one per SM sub-partition) and 8 warps reading
// tensor memory at once, and MUFU.EX2 throughput for the same warp counts. Prints cycles per instruction.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tmem_ld_bw scripts/micro/tmem_ld_bw.cu && /tmp/tmem_ld_bw
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}

__global__ void __launch_bounds__(256) k(int nwarps, int mode, int iters, long long* out, float* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;\n" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
    }
    asm volatile("tcgen05.fence::before_thread_sync;\n");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;\n");
    const uint32_t base = slot;
    float acc = 0.f;
    long long t0 = 0, t1 = 0;
    if (warp < nwarps) {
        const uint32_t addr = base + ((uint32_t)((warp & 3) * 32) << 16) + (warp >> 2) * 128;
        __syncwarp();
        t0 = clock64();
        if (mode == 0) {
            uint32_t r[32];
            for (int i = 0; i < iters; ++i) {
                tmem_ld32(addr + (i & 3) * 32, r);
                asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
                acc += __uint_as_float(r[i & 31]);
            }
        } else if (mode == 1) {  // two loads in flight
            uint32_t r[2][32];
            tmem_ld32(addr, r[0]);
            for (int i = 0; i < iters; i += 2) {
                asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
                tmem_ld32(addr + 32, r[1]);
                acc += __uint_as_float(r[0][i & 31]);
                asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
                tmem_ld32(addr + 64, r[0]);
                acc += __uint_as_float(r[1][i & 31]);
            }
            asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
        } else if (mode == 2) {  // MUFU.EX2
            float x[8];
            for (int j = 0; j < 8; ++j) x[j] = -0.001f * (threadIdx.x + j);
            for (int i = 0; i < iters; ++i) {
#pragma unroll
                for (int j = 0; j < 8; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
            }
            for (int j = 0; j < 8; ++j) acc += x[j];
        } else if (mode == 3) {  // F2FP: cvt.rn.f16x2.f32 (two floats -> packed halves)
            float x[8];
            uint32_t y[4] = {0, 0, 0, 0};
            for (int j = 0; j < 8; ++j) x[j] = 0.001f * (threadIdx.x + j);
            for (int i = 0; i < iters; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t h;
                    asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(x[2 * j]), "f"(x[2 * j + 1]));
                    y[j] ^= h;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t h;
                    asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(x[2 * j + 1]), "f"(x[2 * j]));
                    y[j] ^= h;
                }
            }
            for (int j = 0; j < 4; ++j) acc += __uint_as_float(y[j]);
        } else if (mode == 4) {  // the softmax mix: 2 MUFU.EX2 per 1 F2FP
            float x[8];
            uint32_t y[4] = {0, 0, 0, 0};
            for (int j = 0; j < 8; ++j) x[j] = -0.001f * (threadIdx.x + j);
            for (int i = 0; i < iters; ++i) {
#pragma unroll
                for (int j = 0; j < 8; ++j) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(x[j]));
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    uint32_t h;
                    asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(h) : "f"(x[2 * j]), "f"(x[2 * j + 1]));
                    y[j] ^= h;
                }
            }
            for (int j = 0; j < 4; ++j) acc += __uint_as_float(y[j]);
            for (int j = 0; j < 8; ++j) acc += x[j];
        }
        t1 = clock64();
    }
    if ((threadIdx.x & 31) == 0 && warp < nwarps) out[warp] = t1 - t0;
    sink[threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;\n");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;\n" ::"r"(base));
}

int main() {
    long long* out;
    float* sink;
    cudaMallocManaged(&out, 8 * sizeof(long long));
    cudaMalloc(&sink, 256 * sizeof(float));
    const int iters = 4096;
    for (int mode = 0; mode < 5; ++mode)
        for (int nw : {1, 2, 4, 8}) {
            k<<<1, 256>>>(nw, mode, iters, out, sink);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
            long long mx = 0;
            for (int w = 0; w < nw; ++w) mx = out[w] > mx ? out[w] : mx;
            const double per = (double)mx / iters / (mode >= 2 ? 8 : 1);
            printf("%s warps=%d: %.1f cycles per warp-instruction; aggregate %.1f B/clk/SM\n",
                   mode == 0 ? "tcgen05.ld x32 (serial)   " : mode == 1 ? "tcgen05.ld x32 (2 in flight)" : mode == 2 ? "MUFU.EX2                  " : mode == 3 ? "F2FP (cvt.rn.f16x2.f32)   " : "8 MUFU.EX2 + 4 F2FP (per 8)",
                   nw, per, mode >= 2 ? nw * 32 / per : nw * 4096.0 / per);
        }
    return 0;
}
