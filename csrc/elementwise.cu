// Elementwise / data-movement kernels (§8 rows S11, G1-G3).
// All are bandwidth-bound: 16-byte vector accesses, grid-stride loops sized to 148 SMs.
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

static inline int grid_for(long long work, int threads) {
    long long b = (work + threads - 1) / threads;
    long long cap = 148LL * 16;
    return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

// ---- Philox4x32-10 --------------------------------------------------------------------------------
__host__ __device__ inline void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = static_cast<uint64_t>(M0) * c[0];
        const uint64_t p1 = static_cast<uint64_t>(M1) * c[2];
        const uint32_t n0 = static_cast<uint32_t>(p1 >> 32) ^ c[1] ^ k0;
        const uint32_t n1 = static_cast<uint32_t>(p1);
        const uint32_t n2 = static_cast<uint32_t>(p0 >> 32) ^ c[3] ^ k1;
        const uint32_t n3 = static_cast<uint32_t>(p0);
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += W0; k1 += W1;
    }
}

__global__ void philox_normal_kernel(float* __restrict__ out, long long n, unsigned long long seed,
                                     unsigned long long stream_id, long long offset) {
    // one thread per aligned group of 4 global indices [4q, 4q+4)
    const long long first_q = offset >> 2;
    const long long last_q = (offset + n - 1) >> 2;
    for (long long q = first_q + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; q <= last_q;
         q += static_cast<long long>(gridDim.x) * blockDim.x) {
        uint32_t c[4] = {static_cast<uint32_t>(q), static_cast<uint32_t>(static_cast<unsigned long long>(q) >> 32),
                         static_cast<uint32_t>(stream_id), static_cast<uint32_t>(stream_id >> 32)};
        philox4x32_10(c, static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
        float z[4];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float u1 = (static_cast<float>(c[2 * h] >> 8) + 0.5f) * (1.0f / 16777216.0f);      // (0,1)
            const float u2 = (static_cast<float>(c[2 * h + 1] >> 8) + 0.5f) * (1.0f / 16777216.0f);
            const float r = sqrtf(-2.0f * logf(u1));
            float sn, cs;
            sincospif(2.0f * u2, &sn, &cs);
            z[2 * h] = r * cs;
            z[2 * h + 1] = r * sn;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const long long gi = 4 * q + j;
            if (gi >= offset && gi < offset + n) out[gi - offset] = z[j];
        }
    }
}

// ---- CFG combine + generic linear scheduler update -------------------------------------------------
template <typename T>
__global__ void cfg_step_kernel(const T* __restrict__ ec, const T* __restrict__ eu, const float* __restrict__ x,
                                const float* __restrict__ aux, const float* __restrict__ noise,
                                float* __restrict__ x_out, T* __restrict__ xm_out, float* __restrict__ eps_out,
                                long long n, float g, float ca, float cb, float cc, float cn) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const float c = DT<T>::to_f(ec[i]);
        float e = c;
        if (eu) { const float u = DT<T>::to_f(eu[i]); e = u + g * (c - u); }
        if (eps_out) eps_out[i] = e;
        float v = cb * e;
        if (x) v = fmaf(ca, x[i], v);
        if (aux) v = fmaf(cc, aux[i], v);
        if (noise) v = fmaf(cn, noise[i], v);
        if (x_out) x_out[i] = v;
        if (xm_out) xm_out[i] = DT<T>::from_f(v);
    }
}

// ---- image finalisation --------------------------------------------------------------------------
template <typename T>
__global__ void finalize_u8_kernel(const T* __restrict__ x, unsigned char* __restrict__ out, long long n) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float v = DT<T>::to_f(x[i]) * 0.5f + 0.5f;
        v = fminf(fmaxf(v, 0.0f), 1.0f) * 255.0f;
        out[i] = static_cast<unsigned char>(__float2int_rn(v));
    }
}

template <typename T>
__global__ void upsample2x_kernel(const T* __restrict__ x, T* __restrict__ y, int n, int h, int w, int c) {
    const int ncol = c / 8;
    const long long total = static_cast<long long>(n) * (2 * h) * (2 * w) * ncol;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        long long t = idx / ncol;
        const int ow = static_cast<int>(t % (2 * w)); t /= (2 * w);
        const int oh = static_cast<int>(t % (2 * h));
        const int ni = static_cast<int>(t / (2 * h));
        const uint4 v = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(ni) * h + (oh >> 1)) * w + (ow >> 1)) * c + vc * 8);
        *reinterpret_cast<uint4*>(y + ((static_cast<long long>(ni) * 2 * h + oh) * 2 * w + ow) * c + vc * 8) = v;
    }
}

template <typename T>
__global__ void resize_bilinear_kernel(const T* __restrict__ x, T* __restrict__ y, int n, int h, int w, int c, int oh,
                                       int ow, int align) {
    const long long total = static_cast<long long>(n) * oh * ow * c;
    const float sh = align ? (oh > 1 ? float(h - 1) / float(oh - 1) : 0.f) : float(h) / float(oh);
    const float sw = align ? (ow > 1 ? float(w - 1) / float(ow - 1) : 0.f) : float(w) / float(ow);
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int ch = static_cast<int>(idx % c);
        long long t = idx / c;
        const int x_o = static_cast<int>(t % ow); t /= ow;
        const int y_o = static_cast<int>(t % oh);
        const int ni = static_cast<int>(t / oh);
        float fy = align ? y_o * sh : fmaxf((y_o + 0.5f) * sh - 0.5f, 0.f);
        float fx = align ? x_o * sw : fmaxf((x_o + 0.5f) * sw - 0.5f, 0.f);
        int y0 = min(static_cast<int>(fy), h - 1), x0 = min(static_cast<int>(fx), w - 1);
        int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
        const float ly = fy - y0, lx = fx - x0;
        const T* base = x + static_cast<long long>(ni) * h * w * c + ch;
        const float v00 = DT<T>::to_f(base[(static_cast<long long>(y0) * w + x0) * c]);
        const float v01 = DT<T>::to_f(base[(static_cast<long long>(y0) * w + x1) * c]);
        const float v10 = DT<T>::to_f(base[(static_cast<long long>(y1) * w + x0) * c]);
        const float v11 = DT<T>::to_f(base[(static_cast<long long>(y1) * w + x1) * c]);
        const float top = v00 + (v01 - v00) * lx, bot = v10 + (v11 - v10) * lx;
        y[idx] = DT<T>::from_f(top + (bot - top) * ly);
    }
}

// pad_mode 0: symmetric pad 1 (SD Downsample2D padding=1); 1: pad right/bottom only (VAE encoder style)
template <typename T>
__global__ void im2col3x3_kernel(const T* __restrict__ x, T* __restrict__ col, int n, int h, int w, int c, int stride,
                                 int ho, int wo, int pad_mode) {
    const int ncol = c / 8;
    const long long total = static_cast<long long>(n) * ho * wo * 9 * ncol;
    const int off = pad_mode == 0 ? 1 : 0;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int vc = static_cast<int>(idx % ncol);
        long long t = idx / ncol;
        const int tap = static_cast<int>(t % 9); t /= 9;
        const int xo = static_cast<int>(t % wo); t /= wo;
        const int yo = static_cast<int>(t % ho);
        const int ni = static_cast<int>(t / ho);
        const int yi = yo * stride + tap / 3 - off, xi = xo * stride + tap % 3 - off;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (yi >= 0 && yi < h && xi >= 0 && xi < w)
            v = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(ni) * h + yi) * w + xi) * c + vc * 8);
        *reinterpret_cast<uint4*>(col + (((static_cast<long long>(ni) * ho + yo) * wo + xo) * 9 + tap) * c + vc * 8) = v;
    }
}

template <typename T>
__global__ void binary_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y, long long nvec, int op,
                              float s, int act) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float fa[8], fb[8], o[8];
        unpack8<T>(reinterpret_cast<const uint4*>(a)[i], fa);
        if (b) unpack8<T>(reinterpret_cast<const uint4*>(b)[i], fb);
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) fb[j] = 0.f;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float v = op == 0 ? fa[j] + fb[j] : op == 1 ? fa[j] * fb[j] : fmaf(fb[j], s, fa[j]);
            o[j] = apply_act(v, act);
        }
        reinterpret_cast<uint4*>(y)[i] = pack8<T>(o);
    }
}

template <typename T>
__global__ void timestep_embedding_kernel(const float* __restrict__ t, T* __restrict__ out, int batch, int dim,
                                          int flip, float freq_shift, float max_period) {
    const int half = dim / 2;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= batch * half) return;
    const int b = idx / half, i = idx % half;
    const float freq = expf(-logf(max_period) * static_cast<float>(i) / (static_cast<float>(half) - freq_shift));
    const float arg = t[b] * freq;
    const float sn = sinf(arg), cs = cosf(arg);
    T* o = out + static_cast<long long>(b) * dim;
    if (flip) { o[i] = DT<T>::from_f(cs); o[half + i] = DT<T>::from_f(sn); }
    else { o[i] = DT<T>::from_f(sn); o[half + i] = DT<T>::from_f(cs); }
}

// mode 0: green screen, 1: alpha mask, 2: foreground mask (fgr * pha on black)
template <typename T>
__global__ void matting_composite_kernel(const T* __restrict__ fgr, const T* __restrict__ pha,
                                         unsigned char* __restrict__ out, long long npix, int mode) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < npix;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const float a = fminf(fmaxf(DT<T>::to_f(pha[i]), 0.f), 1.f);
        const float bg[3] = {120.f / 255.f, 255.f / 255.f, 155.f / 255.f};
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            const float f = fminf(fmaxf(DT<T>::to_f(fgr[3 * i + ch]), 0.f), 1.f);
            float v;
            if (mode == 0) v = f * a + bg[ch] * (1.f - a);
            else if (mode == 1) v = a;
            else v = f * a;
            out[3 * i + ch] = static_cast<unsigned char>(__float2int_rn(v * 255.f));
        }
    }
}

// dst = sum_k coef[k] * src[k] (fp32, left-to-right), optional model-dtype copy scaled by m_scale.
struct LinCombArgs { const float* src[6]; float coef[6]; int n; };
template <typename T>
__global__ void lincomb_kernel(float* __restrict__ dst, T* __restrict__ dst_m, float m_scale, LinCombArgs a, long long n) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float acc = a.coef[0] * a.src[0][i];
#pragma unroll
        for (int k = 1; k < 6; ++k)
            if (k < a.n) acc = acc + a.coef[k] * a.src[k][i];
        if (dst) dst[i] = acc;
        if (dst_m) dst_m[i] = DT<T>::from_f(acc * m_scale);
    }
}

}  // namespace ab

#define AB_DISPATCH(dtype, ...)                                 \
    do {                                                        \
        if ((dtype) == AB_BF16) { using T = __nv_bfloat16; __VA_ARGS__; } \
        else { using T = __half; __VA_ARGS__; }                 \
    } while (0)

extern "C" int ab_philox_normal(float* out, long long n, unsigned long long seed, unsigned long long stream_id,
                                long long offset, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(out && n > 0 && offset >= 0, "bad arguments");
    ab::philox_normal_kernel<<<ab::grid_for(n / 4 + 2, 256), 256, 0, stream>>>(out, n, seed, stream_id, offset);
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_cfg_step(const void* ec, const void* eu, const float* x, const float* aux, const float* noise,
                           float* x_out, void* xm_out, float* eps_out, long long n, float g, float ca, float cb,
                           float cc, float cn, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(ec && n > 0, "bad arguments");
    AB_DISPATCH(dtype, (ab::cfg_step_kernel<T><<<ab::grid_for(n, 256), 256, 0, stream>>>(
                           reinterpret_cast<const T*>(ec), reinterpret_cast<const T*>(eu), x, aux, noise, x_out,
                           reinterpret_cast<T*>(xm_out), eps_out, n, g, ca, cb, cc, cn)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_finalize_u8(const void* x, unsigned char* out, long long n, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && out && n > 0, "bad arguments");
    AB_DISPATCH(dtype, (ab::finalize_u8_kernel<T><<<ab::grid_for(n, 256), 256, 0, stream>>>(reinterpret_cast<const T*>(x), out, n)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_upsample_nearest2x(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && c % 8 == 0, "bad arguments (c % 8)");
    const long long total = static_cast<long long>(n) * 4 * h * w * (c / 8);
    AB_DISPATCH(dtype, (ab::upsample2x_kernel<T><<<ab::grid_for(total, 256), 256, 0, stream>>>(
                           reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), n, h, w, c)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_resize_bilinear(const void* x, void* y, int n, int h, int w, int c, int oh, int ow, int align,
                                  int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && oh > 0 && ow > 0, "bad arguments");
    const long long total = static_cast<long long>(n) * oh * ow * c;
    AB_DISPATCH(dtype, (ab::resize_bilinear_kernel<T><<<ab::grid_for(total, 256), 256, 0, stream>>>(
                           reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), n, h, w, c, oh, ow, align)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_im2col3x3(const void* x, void* col, int n, int h, int w, int c, int stride, int pad_mode, int dtype,
                            void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && col && c % 8 == 0 && stride >= 1, "bad arguments");
    const int ho = pad_mode == 0 ? (h + 2 - 3) / stride + 1 : (h + 1 - 3) / stride + 1;
    const int wo = pad_mode == 0 ? (w + 2 - 3) / stride + 1 : (w + 1 - 3) / stride + 1;
    const long long total = static_cast<long long>(n) * ho * wo * 9 * (c / 8);
    AB_DISPATCH(dtype, (ab::im2col3x3_kernel<T><<<ab::grid_for(total, 256), 256, 0, stream>>>(
                           reinterpret_cast<const T*>(x), reinterpret_cast<T*>(col), n, h, w, c, stride, ho, wo, pad_mode)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_binary(const void* a, const void* b, void* y, long long n, int op, float s, int act, int dtype,
                         void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && y && n > 0 && n % 8 == 0, "bad arguments (n % 8)");
    AB_DISPATCH(dtype, (ab::binary_kernel<T><<<ab::grid_for(n / 8, 256), 256, 0, stream>>>(
                           reinterpret_cast<const T*>(a), reinterpret_cast<const T*>(b), reinterpret_cast<T*>(y), n / 8, op, s, act)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_timestep_embedding(const float* t, void* out, int batch, int dim, int flip, float freq_shift,
                                     float max_period, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(t && out && dim % 2 == 0, "bad arguments");
    const int total = batch * dim / 2;
    AB_DISPATCH(dtype, (ab::timestep_embedding_kernel<T><<<(total + 127) / 128, 128, 0, stream>>>(
                           t, reinterpret_cast<T*>(out), batch, dim, flip, freq_shift, max_period)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_matting_composite(const void* fgr, const void* pha, unsigned char* out, long long npix, int mode,
                                    int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(fgr && pha && out && npix > 0, "bad arguments");
    AB_DISPATCH(dtype, (ab::matting_composite_kernel<T><<<ab::grid_for(npix, 256), 256, 0, stream>>>(
                           reinterpret_cast<const T*>(fgr), reinterpret_cast<const T*>(pha), out, npix, mode)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_lincomb(float* dst, void* dst_m, float m_scale, const float* const* srcs, const float* coefs, int nsrc,
                          long long n, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(nsrc >= 1 && nsrc <= 6 && n > 0 && (dst || dst_m), "bad arguments");
    ab::LinCombArgs a{};
    a.n = nsrc;
    for (int k = 0; k < nsrc; ++k) { a.src[k] = srcs[k]; a.coef[k] = coefs[k]; }
    AB_DISPATCH(dtype, (ab::lincomb_kernel<T><<<ab::grid_for(n, 256), 256, 0, stream>>>(dst, reinterpret_cast<T*>(dst_m), m_scale, a, n)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
