// More bandwidth-bound helpers: 2x2 average pooling (guided-diffusion "resblock down"), MoVQ SpatialNorm
// modulation, RVM ConvGRU gate fusion and box filters. SURVEY §2.9 rows K11, K14, K16.
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

static inline int grid2(long long work, int threads) {
    long long b = (work + threads - 1) / threads;
    return static_cast<int>(b < 1 ? 1 : (b > 148LL * 16 ? 148LL * 16 : b));
}

template <typename T>
__global__ void avgpool2x_kernel(const T* __restrict__ x, T* __restrict__ y, int n, int h, int w, int c) {
    const int ncol = c / 8, ho = h / 2, wo = w / 2;
    const long long total = static_cast<long long>(n) * ho * wo * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        long long t = idx / ncol;
        const int xo = static_cast<int>(t % wo); t /= wo;
        const int yo = static_cast<int>(t % ho);
        const int ni = static_cast<int>(t / ho);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int dy = 0; dy < 2; ++dy)
#pragma unroll
            for (int dx = 0; dx < 2; ++dx) {
                float f[8];
                unpack8<T>(*reinterpret_cast<const uint4*>(x + ((static_cast<long long>(ni) * h + 2 * yo + dy) * w + 2 * xo + dx) * c + vc * 8), f);
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] += f[i];          // fixed order: (0,0) (0,1) (1,0) (1,1)
            }
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] *= 0.25f;
        *reinterpret_cast<uint4*>(y + ((static_cast<long long>(ni) * ho + yo) * wo + xo) * c + vc * 8) = pack8<T>(acc);
    }
}

// out = act(x * ymap[up] + bmap[up]); ymap/bmap are [N, H>>s, W>>s, C] (nearest-neighbour upsampling by 2^s)
template <typename T>
__global__ void spatial_modulate_kernel(const T* __restrict__ x, const T* __restrict__ ym, const T* __restrict__ bm,
                                        T* __restrict__ out, int n, int h, int w, int c, int lh, int lw, int act) {
    const int ncol = c / 8;
    const long long total = static_cast<long long>(n) * h * w * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        long long t = idx / ncol;
        const int xx = static_cast<int>(t % w); t /= w;
        const int yy = static_cast<int>(t % h);
        const int ni = static_cast<int>(t / h);
        const int ly = yy * lh / h, lx = xx * lw / w;             // torch 'nearest': floor(dst * in / out)
        const long long lo = ((static_cast<long long>(ni) * lh + ly) * lw + lx) * c + vc * 8;
        float f[8], a[8], b[8], o[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(x + ((static_cast<long long>(ni) * h + yy) * w + xx) * c + vc * 8), f);
        unpack8<T>(*reinterpret_cast<const uint4*>(ym + lo), a);
        unpack8<T>(*reinterpret_cast<const uint4*>(bm + lo), b);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = apply_act(fmaf(f[i], a[i], b[i]), act);
        *reinterpret_cast<uint4*>(out + ((static_cast<long long>(ni) * h + yy) * w + xx) * c + vc * 8) = pack8<T>(o);
    }
}

// ConvGRU tail (RVM): h_new = (1 - z) * h + z * tanh(c), with z already sigmoid-activated. All [n] elements.
template <typename T>
__global__ void gru_update_kernel(const T* __restrict__ z, const T* __restrict__ cand, const T* __restrict__ hprev,
                                  T* __restrict__ out, long long nvec) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float fz[8], fc[8], fh[8], o[8];
        unpack8<T>(reinterpret_cast<const uint4*>(z)[i], fz);
        unpack8<T>(reinterpret_cast<const uint4*>(cand)[i], fc);
        unpack8<T>(reinterpret_cast<const uint4*>(hprev)[i], fh);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (1.f - fz[j]) * fh[j] + fz[j] * tanhf(fc[j]);
        reinterpret_cast<uint4*>(out)[i] = pack8<T>(o);
    }
}

// Depthwise KxK conv (MobileNetV3), NHWC, stride 1|2, pad K/2, weights [K*K, C], fused bias + activation.
template <typename T>
__global__ void dwconv_kernel(const T* __restrict__ x, const T* __restrict__ w, const T* __restrict__ bias,
                              T* __restrict__ y, int n, int h, int wd, int c, int ks, int stride, int dil, int ho, int wo, int act) {
    const int ncol = c / 8;
    const int pad = (ks / 2) * dil;
    const long long total = static_cast<long long>(n) * ho * wo * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        long long t = idx / ncol;
        const int xo = static_cast<int>(t % wo); t /= wo;
        const int yo = static_cast<int>(t % ho);
        const int ni = static_cast<int>(t / ho);
        float acc[8];
        if (bias) unpack8<T>(__ldg(reinterpret_cast<const uint4*>(bias + vc * 8)), acc);
        else {
#pragma unroll
            for (int i = 0; i < 8; ++i) acc[i] = 0.f;
        }
        for (int ky = 0; ky < ks; ++ky) {
            const int yi = yo * stride + ky * dil - pad;
            if (yi < 0 || yi >= h) continue;
            for (int kx = 0; kx < ks; ++kx) {
                const int xi = xo * stride + kx * dil - pad;
                if (xi < 0 || xi >= wd) continue;
                float f[8], wv[8];
                unpack8<T>(*reinterpret_cast<const uint4*>(x + ((static_cast<long long>(ni) * h + yi) * wd + xi) * c + vc * 8), f);
                unpack8<T>(__ldg(reinterpret_cast<const uint4*>(w + static_cast<long long>(ky * ks + kx) * c + vc * 8)), wv);
#pragma unroll
                for (int i = 0; i < 8; ++i) acc[i] = fmaf(f[i], wv[i], acc[i]);
            }
        }
        float o[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = apply_act(acc[i], act);
        *reinterpret_cast<uint4*>(y + ((static_cast<long long>(ni) * ho + yo) * wo + xo) * c + vc * 8) = pack8<T>(o);
    }
}

// Global average pool over HW -> [N, C] (squeeze-excite, LR-ASPP): one block per (n, 8-channel vector), fixed tree.
template <typename T>
__global__ void __launch_bounds__(256) global_avgpool_kernel(const T* __restrict__ x, T* __restrict__ y, int hw, int c) {
    __shared__ float red[8][256 / 32 * 8 / 8 * 1 + 32];
    const int ncol = c / 8;
    const int ni = blockIdx.x / ncol, vc = blockIdx.x % ncol;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int r = threadIdx.x; r < hw; r += 256) {
        float f[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(x + (static_cast<long long>(ni) * hw + r) * c + vc * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[i] += f[i];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], o);
    if ((threadIdx.x & 31) == 0)
#pragma unroll
        for (int i = 0; i < 8; ++i) red[i][threadIdx.x >> 5] = acc[i];
    __syncthreads();
    if (threadIdx.x < 8) {
        float s = 0.f;
        for (int wv = 0; wv < 8; ++wv) s += red[threadIdx.x][wv];       // fixed order
        y[static_cast<long long>(ni) * c + vc * 8 + threadIdx.x] = DT<T>::from_f(s / hw);
    }
}

// y[n, r, c] = x[n, r, c] * s[n, c]  (squeeze-excite scaling), optional activation on s done by the producer
template <typename T>
__global__ void channel_scale_kernel(const T* __restrict__ x, const T* __restrict__ s, T* __restrict__ y, int n, int hw, int c) {
    const int ncol = c / 8;
    const long long total = static_cast<long long>(n) * hw * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        const long long row = idx / ncol;
        const int ni = static_cast<int>(row / hw);
        float f[8], sv[8], o[8];
        unpack8<T>(*reinterpret_cast<const uint4*>(x + row * c + vc * 8), f);
        unpack8<T>(__ldg(reinterpret_cast<const uint4*>(s + static_cast<long long>(ni) * c + vc * 8)), sv);
#pragma unroll
        for (int i = 0; i < 8; ++i) o[i] = f[i] * sv[i];
        *reinterpret_cast<uint4*>(y + row * c + vc * 8) = pack8<T>(o);
    }
}

}  // namespace ab

#define AB_DISPATCH2(dtype, ...)                                           \
    do {                                                                   \
        if ((dtype) == AB_BF16) { using T = __nv_bfloat16; __VA_ARGS__; }  \
        else { using T = __half; __VA_ARGS__; }                            \
    } while (0)

extern "C" int ab_avgpool2x(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && c % 8 == 0 && h % 2 == 0 && w % 2 == 0, "bad arguments (c % 8, even h/w)");
    const long long total = static_cast<long long>(n) * (h / 2) * (w / 2) * (c / 8);
    AB_DISPATCH2(dtype, (ab::avgpool2x_kernel<T><<<ab::grid2(total, 256), 256, 0, stream>>>(reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), n, h, w, c)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_spatial_modulate(const void* x, const void* ymap, const void* bmap, void* out, int n, int h, int w, int c,
                                   int lh, int lw, int act, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && ymap && bmap && out && c % 8 == 0 && lh > 0 && lw > 0, "bad arguments");
    const long long total = static_cast<long long>(n) * h * w * (c / 8);
    AB_DISPATCH2(dtype, (ab::spatial_modulate_kernel<T><<<ab::grid2(total, 256), 256, 0, stream>>>(
                            reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(ymap), reinterpret_cast<const T*>(bmap),
                            reinterpret_cast<T*>(out), n, h, w, c, lh, lw, act)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_gru_update(const void* z, const void* cand, const void* hprev, void* out, long long n, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(z && cand && hprev && out && n % 8 == 0, "bad arguments (n % 8)");
    AB_DISPATCH2(dtype, (ab::gru_update_kernel<T><<<ab::grid2(n / 8, 256), 256, 0, stream>>>(
                            reinterpret_cast<const T*>(z), reinterpret_cast<const T*>(cand), reinterpret_cast<const T*>(hprev),
                            reinterpret_cast<T*>(out), n / 8)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_dwconv(const void* x, const void* w, const void* bias, void* y, int n, int h, int wd, int c, int ksize,
                         int stride, int dilation, int act, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && w && y && c % 8 == 0 && (ksize == 3 || ksize == 5) && (stride == 1 || stride == 2) && dilation >= 1, "bad arguments");
    const int pad = (ksize / 2) * dilation;
    const int ho = (h + 2 * pad - dilation * (ksize - 1) - 1) / stride + 1, wo = (wd + 2 * pad - dilation * (ksize - 1) - 1) / stride + 1;
    const long long total = static_cast<long long>(n) * ho * wo * (c / 8);
    AB_DISPATCH2(dtype, (ab::dwconv_kernel<T><<<ab::grid2(total, 256), 256, 0, stream>>>(
                            reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(w), reinterpret_cast<const T*>(bias),
                            reinterpret_cast<T*>(y), n, h, wd, c, ksize, stride, dilation, ho, wo, act)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_global_avgpool(const void* x, void* y, int n, int hw, int c, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && c % 8 == 0 && hw > 0, "bad arguments");
    AB_DISPATCH2(dtype, (ab::global_avgpool_kernel<T><<<n * (c / 8), 256, 0, stream>>>(reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), hw, c)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_channel_scale(const void* x, const void* s, void* y, int n, int hw, int c, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && s && y && c % 8 == 0, "bad arguments");
    const long long total = static_cast<long long>(n) * hw * (c / 8);
    AB_DISPATCH2(dtype, (ab::channel_scale_kernel<T><<<ab::grid2(total, 256), 256, 0, stream>>>(
                            reinterpret_cast<const T*>(x), reinterpret_cast<const T*>(s), reinterpret_cast<T*>(y), n, hw, c)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
