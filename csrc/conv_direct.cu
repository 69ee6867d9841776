// Direct convolutions for layers whose channel count is too small for a tensor-core tile
// (conv_in 4->320, conv_out 320->4, VAE 4->512 / 128->3, RVM RGB stem). SURVEY §2.9 row K1.
//   few outputs (Cout <= 8, Cin % 8 == 0): one WARP per output pixel; lanes split the (tap, 8-channel)
//       vectors, weights [Cout][kh][kw][Cin] are read through L1, partial sums are combined with a
//       fixed xor-shuffle tree (deterministic).
//   few inputs  (Cin <= 16, Cout % 8 == 0): one thread per (pixel, 8 output channels); weights are
//       expected pre-transposed as [kh][kw][Cin][Cout] so each (tap, cin) is one 16-byte load.
// No shared memory, no per-block weight staging: both shapes are latency/bandwidth trivial once the
// work is spread over the whole chip.
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

template <typename T>
__global__ void __launch_bounds__(256) conv_few_in_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                          const T* __restrict__ bias, T* __restrict__ out, int n, int h,
                                                          int wd, int cin, int cout, int ks, int stride, int ho, int wo,
                                                          int act) {
    const int ncol = cout / 8;
    const int pad = ks / 2;
    const long long total = static_cast<long long>(n) * ho * wo * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        int t = static_cast<int>(idx / ncol);
        const int xo = t % wo; t /= wo;
        const int yo = t % ho;
        const int ni = t / ho;
        float acc[8];
        if (bias) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(bias + vc * 8)), acc);
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = 0.f;
        }
        for (int ky = 0; ky < ks; ++ky) {
            const int yi = yo * stride + ky - pad;
            if (yi < 0 || yi >= h) continue;
            for (int kx = 0; kx < ks; ++kx) {
                const int xi = xo * stride + kx - pad;
                if (xi < 0 || xi >= wd) continue;
                const T* px = x + (static_cast<long long>(ni * h + yi) * wd + xi) * cin;
                const T* pw = w + static_cast<long long>((ky * ks + kx) * cin) * cout + vc * 8;
                for (int ci = 0; ci < cin; ++ci) {
                    const float xv = DT<T>::to_f(px[ci]);
                    float wv[8];
                    unpack8<T>(__ldg(reinterpret_cast<const uint4*>(pw + static_cast<long long>(ci) * cout)), wv);
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc[j] = fmaf(xv, wv[j], acc[j]);
                }
            }
        }
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = apply_act(acc[j], act);
        *reinterpret_cast<uint4*>(out + (static_cast<long long>(ni * ho + yo) * wo + xo) * cout + vc * 8) = pack8<T>(o);
    }
}

template <typename T, int MAXCO, bool SMEM_W>
__global__ void __launch_bounds__(256) conv_few_out_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                           const T* __restrict__ bias, T* __restrict__ out, int n, int h,
                                                           int wd, int cin, int cout, int ks, int stride, int ho, int wo,
                                                           int act) {
    // The whole weight tensor (cout x ks*ks*cin, <= 48 KB for the UNet/VAE output convs) is staged in shared memory once
    // per block when it fits: every pixel re-reads all of it, and from global memory that re-read is the kernel's
    // traffic. Same values, same order of operations: the result bits do not change.
    extern __shared__ uint4 w_smem[];
    const int lane = threadIdx.x & 31;
    const int warps_per_block = blockDim.x >> 5;
    const int pad = ks / 2;
    const int vec_per_tap = cin / 8;
    const int kk = ks * ks * cin;
    if (SMEM_W) {
        const uint4* wg = reinterpret_cast<const uint4*>(w);
        for (int i = threadIdx.x; i < cout * kk / 8; i += blockDim.x) w_smem[i] = __ldg(wg + i);
        __syncthreads();
        w = reinterpret_cast<const T*>(w_smem);
    }
    const long long total = static_cast<long long>(n) * ho * wo;
    for (long long pix = static_cast<long long>(blockIdx.x) * warps_per_block + (threadIdx.x >> 5); pix < total;
         pix += static_cast<long long>(gridDim.x) * warps_per_block) {
        int t = static_cast<int>(pix);
        const int xo = t % wo; t /= wo;
        const int yo = t % ho;
        const int ni = t / ho;
        float acc[MAXCO];
#pragma unroll
        for (int j = 0; j < MAXCO; ++j) acc[j] = 0.f;
        for (int tap = 0; tap < ks * ks; ++tap) {
            const int yi = yo * stride + tap / ks - pad, xi = xo * stride + tap % ks - pad;
            if (yi < 0 || yi >= h || xi < 0 || xi >= wd) continue;      // warp-uniform
            const T* px = x + (static_cast<long long>(ni * h + yi) * wd + xi) * cin;
            const T* pw = w + tap * cin;
            for (int v = lane; v < vec_per_tap; v += 32) {
                float f[8];
                unpack8<T>(*reinterpret_cast<const uint4*>(px + v * 8), f);
#pragma unroll
                for (int j = 0; j < MAXCO; ++j) {
                    if (j < cout) {
                        float wv[8];
                        const uint4* wp = reinterpret_cast<const uint4*>(pw + static_cast<long long>(j) * kk + v * 8);
                        unpack8<T>(SMEM_W ? *wp : __ldg(wp), wv);
#pragma unroll
                        for (int i = 0; i < 8; ++i) acc[j] = fmaf(f[i], wv[i], acc[j]);
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < MAXCO; ++j) {
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
        }
        if (lane == 0) {
            T* po = out + pix * cout;
#pragma unroll
            for (int j = 0; j < MAXCO; ++j)
                if (j < cout) po[j] = DT<T>::from_f(apply_act(acc[j] + (bias ? DT<T>::to_f(bias[j]) : 0.f), act));
        }
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
    const bool few_out = cout <= 8 && cin % 8 == 0;
    const bool few_in = !few_out && cin <= 16 && cout % 8 == 0;
    AB_REQUIRE(few_out || few_in, "ab_conv_direct handles (Cout<=8, Cin%8==0) or (Cin<=16, Cout%8==0)");
    const int threads = 256;
    const long long work = few_out ? (static_cast<long long>(n) * ho * wo + 7) / 8
                                   : (static_cast<long long>(n) * ho * wo * (cout / 8) + threads - 1) / threads;
    const int blocks = static_cast<int>(std::min<long long>(std::max<long long>(work, 1), 148LL * 8));
    const size_t w_bytes = static_cast<size_t>(cout) * ksize * ksize * cin * 2;
#define AB_LAUNCH(T)                                                                                               \
    do {                                                                                                           \
        if (few_out && w_bytes <= 48 * 1024)                                                                       \
            conv_few_out_kernel<T, 8, true><<<blocks, threads, w_bytes, stream>>>(                                 \
                reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(w), reinterpret_cast<const T*>(bias),    \
                reinterpret_cast<T*>(out), n, h, wd, cin, cout, ksize, stride, ho, wo, act);                       \
        else if (few_out)                                                                                          \
            conv_few_out_kernel<T, 8, false><<<blocks, threads, 0, stream>>>(                                      \
                reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(w), reinterpret_cast<const T*>(bias),    \
                reinterpret_cast<T*>(out), n, h, wd, cin, cout, ksize, stride, ho, wo, act);                       \
        else                                                                                                       \
            conv_few_in_kernel<T><<<blocks, threads, 0, stream>>>(                                                 \
                reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(w), reinterpret_cast<const T*>(bias),    \
                reinterpret_cast<T*>(out), n, h, wd, cin, cout, ksize, stride, ho, wo, act);                       \
    } while (0)
    if (dtype == AB_BF16) AB_LAUNCH(__nv_bfloat16); else AB_LAUNCH(__half);
#undef AB_LAUNCH
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
