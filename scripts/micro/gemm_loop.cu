// Micro-benchmark of the GEMM main loop in isolation: one CTA per SM (grid = nctas), a TMA producer warp and an MMA
// warp around a STAGES-deep ring, K steps of 64. Prints cycles per K step for the variants:
//   mode 0: MMAs only (operands already in shared memory), one tcgen05.commit per K step
//   mode 1: TMA loads only (producer waits for a software "consumed" arrive instead of MMAs)
//   mode 2: the full loop (TMA -> full barrier -> 4 MMAs -> commit -> empty barrier)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I csrc -o /tmp/gemm_loop scripts/micro/gemm_loop.cu -lcuda
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

template <int BN>
__global__ void __launch_bounds__(128) k(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                         int stages, int ksteps_, int mode, long long* out) {
    int ksteps = ksteps_;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int STAGE = 16384 + BN * 128;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + stages * STAGE);
    uint64_t* empty = full + stages;
    uint64_t* done = empty + stages;
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) { mbar_init(&full[s], mode >= 10 ? 2 : 1); mbar_init(&empty[s], 1); }
        mbar_init(done, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<256>(&slot);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = slot;
    const long long t0 = clock64();
    if ((warp == 0 || warp == 2) && mode >= 10) {
        // two producer lanes: warp 0 loads A, warp 2 loads B (each arrives once on the stage's full barrier)
        if (elect_one()) {
            int stage = 0; uint32_t phase = 0;
            for (int kc = 0; kc < ksteps; ++kc) {
                mbar_wait(&empty[stage], phase ^ 1);
                if (warp == 0) {
                    mbar_expect_tx(&full[stage], 16384);
                    tma_load_2d(smem + stage * STAGE, &tmA, &full[stage], kc * 64, blockIdx.x * 128);
                } else {
                    mbar_expect_tx(&full[stage], BN * 128);
                    tma_load_2d(smem + stage * STAGE + 16384, &tmB, &full[stage], kc * 64, 0);
                }
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == 0 && mode >= 6) {
        if (elect_one()) {
            int stage = 0; uint32_t phase = 0;
            for (int kc = 0; kc < ksteps; ++kc) {
                if (mode < 8) mbar_wait(&empty[stage], phase ^ 1);
                mbar_expect_tx(&full[stage], mode == 9 ? 16384 : STAGE);
                tma_load_2d(smem + stage * STAGE, &tmA, &full[stage], kc * 64, blockIdx.x * 128);
                if (mode != 9) tma_load_2d(smem + stage * STAGE + 16384, &tmB, &full[stage], kc * 64, 0);
                if (++stage == stages) { stage = 0; phase ^= 1; }
            }
        }
        __syncwarp();
    } else if (warp == 0 && (mode == 1 || mode == 2)) {
        const bool leader = elect_one();
        int stage = 0; uint32_t phase = 0;
        for (int kc = 0; kc < ksteps; ++kc) {
            mbar_wait(&empty[stage], phase ^ 1);
            if (leader) {
                mbar_expect_tx(&full[stage], STAGE);
                tma_load_2d(smem + stage * STAGE, &tmA, &full[stage], kc * 64, blockIdx.x * 128);
                tma_load_2d(smem + stage * STAGE + 16384, &tmB, &full[stage], kc * 64, 0);
            }
            __syncwarp();
            if (++stage == stages) { stage = 0; phase ^= 1; }
        }
    } else if (warp == 1) {
        const bool leader = elect_one();
        constexpr uint32_t idesc = make_idesc_f16(128, BN, false, false, false);
        const uint64_t d0 = make_smem_desc(0, 16, 1024, kSwizzle128);
        const uint32_t base = smem_u32(smem);
        int stage = 0; uint32_t phase = 0;
        if (mode == 8 || mode == 9) {
            // producer-rate probe: nobody recycles stages; wait for the LAST load only (the ring is overwritten freely)
            if (leader) {
                const int last = (ksteps - 1) % stages;
                mbar_wait(&full[last], ((ksteps - 1) / stages) & 1);
            }
            __syncwarp();
            ksteps = 0;
        } else if (mode >= 6) {
            if (leader) {
                for (int kc = 0; kc < ksteps; ++kc) {
                    mbar_wait(&full[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = base + stage * STAGE, sb = sa + 16384;
                    if (mode == 6 || mode == 10) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
                            umma_ss(tmem_base, smem_desc_at(d0, sa + ks * 32), smem_desc_at(d0, sb + ks * 32), idesc, 1u);
                        umma_commit(&empty[stage]);
                    } else {
                        mbar_arrive(&empty[stage]);
                    }
                    if (++stage == stages) { stage = 0; phase ^= 1; }
                }
            }
            __syncwarp();
            ksteps = 0;
        } else if (mode >= 3) {
            // the whole loop under the elected lane: 3 = commit per K step, 4 = no commit, 5 = commit to a rotating
            // barrier but descriptors of one fixed stage
            if (leader) {
                for (int kc = 0; kc < ksteps; ++kc) {
                    const uint32_t sa = base + (mode == 5 ? 0 : stage) * STAGE, sb = sa + 16384;
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_ss(tmem_base, smem_desc_at(d0, sa + ks * 32), smem_desc_at(d0, sb + ks * 32), idesc, 1u);
                    if (mode != 4) umma_commit(&empty[stage]);
                    if (++stage == stages) stage = 0;
                }
            }
            __syncwarp();
            ksteps = 0;
        }
        for (int kc = 0; kc < ksteps; ++kc) {
            if (mode != 0) { mbar_wait(&full[stage], phase); tc_fence_after(); }
            const uint32_t sa = base + stage * STAGE, sb = sa + 16384;
            if (leader) {
                if (mode != 1) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks)
                        umma_ss(tmem_base, smem_desc_at(d0, sa + ks * 32), smem_desc_at(d0, sb + ks * 32), idesc, 1u);
                    umma_commit(&empty[stage]);
                } else {
                    mbar_arrive(&empty[stage]);
                }
            }
            __syncwarp();
            if (++stage == stages) { stage = 0; phase ^= 1; }
        }
        if (leader) umma_commit(done);
        __syncwarp();
        mbar_wait(done, 0);
        if (leader) out[blockIdx.x] = clock64() - t0;
        ksteps = ksteps_;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc<256>(tmem_base);
}

static CUtensorMap make_map(void* ptr, int rows, int cols, int box_rows) {
    CUtensorMap m;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t es[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) printf("cuTensorMapEncodeTiled failed %d\n", (int)r);
    return m;
}

template <int BN>
void run(int nctas, int stages, int K, long long* out, __half* A, __half* B) {
    const int M = nctas * 128, ksteps = K / 64;
    CUtensorMap tA = make_map(A, M, K, 128), tB = make_map(B, BN, K, BN);
    const int smem = stages * (16384 + BN * 128) + (2 * stages + 1) * 8 + 1024;
    cudaFuncSetAttribute(k<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    for (int mode = 6; mode < 12; ++mode) {
        if (mode == 8 || mode == 9) continue;
        for (int rep = 0; rep < 2; ++rep) k<BN><<<nctas, 128, smem>>>(tA, tB, stages, ksteps, mode, out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
        long long mx = 0;
        for (int i = 0; i < nctas; ++i) mx = out[i] > mx ? out[i] : mx;
        printf("BN=%3d ctas=%3d stages=%2d K=%5d %-12s %7.1f cycles per K step (slowest CTA)\n", BN, nctas, stages, K,
               mode == 0 ? "MMA+commit" : mode == 1 ? "TMA only" : mode == 2 ? "full loop" : mode == 3 ? "leader-loop MMA+commit" : mode == 4 ? "leader-loop MMA" : mode == 5 ? "leader-loop fixed-stage" : mode == 6 ? "elected-lane full loop" : mode == 7 ? "elected-lane TMA only" : mode == 10 ? "two producers, full loop" : "two producers, TMA only", (double)mx / ksteps);

    }
}

int main() {
    cuInit(0);
    long long* out;
    cudaMallocManaged(&out, 1024 * 8);
    __half *A, *B;
    const int K = 6912;
    cudaMalloc(&A, (size_t)148 * 128 * K * 2);
    cudaMalloc(&B, (size_t)256 * K * 2);
    cudaMemset(A, 0, (size_t)148 * 128 * K * 2);
    cudaMemset(B, 0, (size_t)256 * K * 2);
    run<32>(30, 4, K, out, A, B);
    run<128>(30, 3, K, out, A, B);
    run<256>(30, 4, K, out, A, B);
    run<32>(30, 10, K, out, A, B);
    run<64>(30, 9, K, out, A, B);
    run<128>(30, 7, K, out, A, B);
    run<160>(30, 6, K, out, A, B);
    run<160>(148, 6, K, out, A, B);
    run<256>(148, 4, K, out, A, B);
    return 0;
}
