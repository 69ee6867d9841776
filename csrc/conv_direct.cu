// Direct convolutions for layers whose channel count is too small for a tensor-core tile
// (conv_in 4->320, conv_out 320->4, VAE 4->512 / 128->3, RVM RGB stem). SURVEY §2.9 row K1.
// Two shapes: "few inputs" (Cin <= 8, one thread per pixel x 8 output channels) and
// "few outputs" (Cout <= 8, one thread per pixel, weights in shared memory).
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

template <typename T>
__global__ void conv_few_in_kernel(const T* __restrict__ x, const T* __restrict__ w, const T* __restrict__ bias,
                                   T* __restrict__ out, int n, int h, int wd, int cin, int cout, int ks, int stride,
                                   int ho, int wo, int act) {
    extern __shared__ float sw[];   // [ks*ks*cin][cout]
    const int kk = ks * ks * cin;
    for (int i = threadIdx.x; i < kk * cout; i += blockDim.x) {
        const int co = i % cout, k = i / cout;
        sw[i] = DT<T>::to_f(w[static_cast<long long>(co) * kk + k]);
    }
    __syncthreads();
    const int ncol = cout / 8;
    const int pad = ks / 2;
    const long long total = static_cast<long long>(n) * ho * wo * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        long long t = idx / ncol;
        const int xo = static_cast<int>(t % wo); t /= wo;
        const int yo = static_cast<int>(t % ho);
        const int ni = static_cast<int>(t / ho);
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = bias ? DT<T>::to_f(bias[vc * 8 + j]) : 0.f;
        for (int ky = 0; ky < ks; ++ky) {
            const int yi = yo * stride + ky - pad;
            if (yi < 0 || yi >= h) continue;
            for (int kx = 0; kx < ks; ++kx) {
                const int xi = xo * stride + kx - pad;
                if (xi < 0 || xi >= wd) continue;
                const T* px = x + ((static_cast<long long>(ni) * h + yi) * wd + xi) * cin;
                const float* pw = sw + ((ky * ks + kx) * cin) * cout + vc * 8;
                for (int ci = 0; ci < cin; ++ci) {
                    const float xv = DT<T>::to_f(px[ci]);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] = fmaf(xv, pw[ci * cout + j], acc[j]);
                }
            }
        }
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = apply_act(acc[j], act);
        *reinterpret_cast<uint4*>(out + ((static_cast<long long>(ni) * ho + yo) * wo + xo) * cout + vc * 8) = pack8<T>(o);
    }
}

template <typename T, int MAXCO>
__global__ void conv_few_out_kernel(const T* __restrict__ x, const T* __restrict__ w, const T* __restrict__ bias,
                                    T* __restrict__ out, int n, int h, int wd, int cin, int cout, int ks, int stride,
                                    int ho, int wo, int act) {
    extern __shared__ float sw[];   // [cout][ks*ks*cin]
    const int kk = ks * ks * cin;
    for (int i = threadIdx.x; i < kk * cout; i += blockDim.x) sw[i] = DT<T>::to_f(w[i]);
    __syncthreads();
    const int pad = ks / 2;
    const long long total = static_cast<long long>(n) * ho * wo;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        long long t = idx;
        const int xo = static_cast<int>(t % wo); t /= wo;
        const int yo = static_cast<int>(t % ho);
        const int ni = static_cast<int>(t / ho);
        float acc[MAXCO];
#pragma unroll
        for (int j = 0; j < MAXCO; ++j) acc[j] = (bias && j < cout) ? DT<T>::to_f(bias[j]) : 0.f;
        for (int ky = 0; ky < ks; ++ky) {
            const int yi = yo * stride + ky - pad;
            if (yi < 0 || yi >= h) continue;
            for (int kx = 0; kx < ks; ++kx) {
                const int xi = xo * stride + kx - pad;
                if (xi < 0 || xi >= wd) continue;
                const T* px = x + ((static_cast<long long>(ni) * h + yi) * wd + xi) * cin;
                const float* pw = sw + (ky * ks + kx) * cin;
                for (int c8 = 0; c8 < cin; c8 += 8) {
                    float f[8];
                    unpack8<T>(*reinterpret_cast<const uint4*>(px + c8), f);
#pragma unroll
                    for (int j = 0; j < MAXCO; ++j) {
                        if (j < cout) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) acc[j] = fmaf(f[i], pw[j * kk + c8 + i], acc[j]);
                        }
                    }
                }
            }
        }
        T* po = out + idx * cout;
#pragma unroll
        for (int j = 0; j < MAXCO; ++j)
            if (j < cout) po[j] = DT<T>::from_f(apply_act(acc[j], act));
    }
}

}  // namespace ab

extern "C" int ab_conv_direct(const void* x, const void* w, const void* bias, void* out, int n, int h, int wd, int cin,
                              int cout, int ksize, int stride, int act, int dtype, void* stream_) {
    using namespace ab;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && w && out, "null pointer");
    AB_REQUIRE((ksize == 1 || ksize == 3) && (stride == 1 || stride == 2), "ksize 1|3, stride 1|2");
    const int pad = ksize / 2;
    const int ho = (h + 2 * pad - ksize) / stride + 1, wo = (wd + 2 * pad - ksize) / stride + 1;
    const size_t smem = static_cast<size_t>(ksize) * ksize * cin * cout * sizeof(float);
    AB_REQUIRE(smem <= 200 * 1024, "weights do not fit in shared memory: use ab_gemm");
    const bool few_out = cout <= 8 && cin % 8 == 0;
    const bool few_in = !few_out && cin <= 16 && cout % 8 == 0;
    AB_REQUIRE(few_out || few_in, "ab_conv_direct handles (Cout<=8, Cin%8==0) or (Cin<=16, Cout%8==0)");
    const long long work = few_out ? static_cast<long long>(n) * ho * wo : static_cast<long long>(n) * ho * wo * (cout / 8);
    const int threads = 256;
    int blocks = static_cast<int>(std::min<long long>((work + threads - 1) / threads, 148LL * 8));
#define AB_LAUNCH(T)                                                                                                 \
    do {                                                                                                             \
        if (few_out) {                                                                                               \
            auto k = conv_few_out_kernel<T, 8>;                                                                      \
            if (smem > 48 * 1024) AB_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            k<<<blocks, threads, smem, stream>>>(reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(w),      \
                                                 reinterpret_cast<const T*>(bias), reinterpret_cast<T*>(out), n, h, \
                                                 wd, cin, cout, ksize, stride, ho, wo, act);                        \
        } else {                                                                                                     \
            auto k = conv_few_in_kernel<T>;                                                                          \
            if (smem > 48 * 1024) AB_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
            k<<<blocks, threads, smem, stream>>>(reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(w),      \
                                                 reinterpret_cast<const T*>(bias), reinterpret_cast<T*>(out), n, h, \
                                                 wd, cin, cout, ksize, stride, ho, wo, act);                        \
        }                                                                                                            \
    } while (0)
    if (dtype == AB_BF16) AB_LAUNCH(__nv_bfloat16); else AB_LAUNCH(__half);
#undef AB_LAUNCH
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
