// Robust Video Matting kernels: small-channel direct convolution with virtual concat,
// gates and a fused ConvGRU epilogue; explicit-step bilinear resize; ceil-mode pooling; deep guided filter.
// These shapes (3/35/59/107/171 input channels, 4..80 outputs, <= 270x480 pixels) do not tile onto the tcgen05
// path; the 1x1 convolutions of the MobileNetV3 encoder do and go through ab_gemm.
#include <stdio.h>

#include "../include/arbius_b200.h"
#include "common.cuh"

extern "C" void ab_count_launch(int n);

namespace ab {

static inline int grid_m(long long work, int threads) {
    long long b = (work + threads - 1) / threads;
    return static_cast<int>(b < 1 ? 1 : (b > 148LL * 32 ? 148LL * 32 : b));
}

constexpr int CG_THREADS = 128;   // one output pixel per thread
constexpr int CG_CO = 16;         // output channels per thread
constexpr int CG_CK = 32;         // input channels staged per step

struct ConvGenParams {
    const void* src[3];
    const void* gate[3];
    int c[3], ld[3], gate_ld[3];
    int nsrc, ctot;
    const void* w;
    const void* bias;
    const float* in_scale;
    const float* in_shift;
    void* out;
    int out_ld;
    const void* gru_z;
    const void* gru_h;
    int gru_z_ld, gru_h_ld;
    int n, h, w_in, ho, wo, ks, stride, dilation, pad, cout, cout_pad, act;
};

// ALLW: the block's whole weight slice [ks*ks*ctot][16] (fp32) is staged once in dynamic shared memory — every RVM
// layer fits (<= 98.5 KB) — so the tap/channel loops run without barriers; otherwise 32-channel slabs are staged per step.
template <typename T, bool ALLW>
__global__ void __launch_bounds__(CG_THREADS) conv_generic_kernel(const __grid_constant__ ConvGenParams p) {
    __shared__ __align__(16) float ws[ALLW ? 1 : CG_CK][CG_CO];
    extern __shared__ __align__(16) float wall[];
    const long long npix = static_cast<long long>(p.n) * p.ho * p.wo;
    const long long pix = static_cast<long long>(blockIdx.x) * CG_THREADS + threadIdx.x;
    const bool live = pix < npix;
    const int co0 = blockIdx.y * CG_CO;
    int xo = 0, yo = 0, ni = 0;
    if (live) {
        long long t = pix;
        xo = static_cast<int>(t % p.wo); t /= p.wo;
        yo = static_cast<int>(t % p.ho);
        ni = static_cast<int>(t / p.ho);
    }
    float acc[CG_CO];
#pragma unroll
    for (int j = 0; j < CG_CO; ++j) acc[j] = 0.f;
    const T* wt = reinterpret_cast<const T*>(p.w);
    if constexpr (ALLW) {
        const int total = p.ks * p.ks * p.ctot * CG_CO;
        for (int i = threadIdx.x; i < total; i += CG_THREADS)
            wall[i] = DT<T>::to_f(wt[static_cast<long long>(i / CG_CO) * p.cout_pad + co0 + (i % CG_CO)]);
        __syncthreads();
    }

    int coff = 0;
    for (int s = 0; s < p.nsrc; ++s) {
        const T* src = reinterpret_cast<const T*>(p.src[s]);
        const T* gate = reinterpret_cast<const T*>(p.gate[s]);
        const int cs = p.c[s], ld = p.ld[s], gld = p.gate_ld[s];
        const bool vec = (cs % 8 == 0) && (ld % 8 == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                         (!gate || ((gld % 8 == 0) && ((reinterpret_cast<uintptr_t>(gate) & 15) == 0)));
        const bool affine = (s == 0) && p.in_scale;
        const bool vec_s = vec && !affine;
        for (int tap = 0; tap < p.ks * p.ks; ++tap) {
            const int ky = tap / p.ks, kx = tap % p.ks;
            const int yi = yo * p.stride - p.pad + ky * p.dilation;
            const int xi = xo * p.stride - p.pad + kx * p.dilation;
            const bool inb = live && yi >= 0 && yi < p.h && xi >= 0 && xi < p.w_in;
            const long long pin = (static_cast<long long>(ni) * p.h + yi) * p.w_in + xi;
            for (int c0 = 0; c0 < cs; c0 += CG_CK) {
                const int ck = min(CG_CK, cs - c0);
                const float* wbase;
                if constexpr (ALLW) {
                    wbase = wall + (static_cast<long long>(tap) * p.ctot + coff + c0) * CG_CO;
                } else {
                    __syncthreads();
                    // stage weights [ck x 16] as fp32
                    for (int i = threadIdx.x; i < CG_CK * CG_CO; i += CG_THREADS) {
                        const int ci = i / CG_CO, j = i % CG_CO;
                        float v = 0.f;
                        if (ci < ck)
                            v = DT<T>::to_f(wt[(static_cast<long long>(tap) * p.ctot + coff + c0 + ci) * p.cout_pad + co0 + j]);
                        ws[ci][j] = v;
                    }
                    __syncthreads();
                    wbase = &ws[0][0];
                }
                if (!inb) continue;
                if (vec_s) {
                    for (int ci = 0; ci < ck; ci += 8) {
                        float f[8];
                        unpack8<T>(*reinterpret_cast<const uint4*>(src + pin * ld + c0 + ci), f);
                        if (gate) {
                            float g[8];
                            unpack8<T>(*reinterpret_cast<const uint4*>(gate + pin * gld + c0 + ci), g);
#pragma unroll
                            for (int u = 0; u < 8; ++u) f[u] *= g[u];
                        }
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const float4* wr = reinterpret_cast<const float4*>(wbase + (ci + u) * CG_CO);
#pragma unroll
                            for (int q = 0; q < CG_CO / 4; ++q) {
                                const float4 wv = wr[q];
                                acc[4 * q + 0] = fmaf(f[u], wv.x, acc[4 * q + 0]);
                                acc[4 * q + 1] = fmaf(f[u], wv.y, acc[4 * q + 1]);
                                acc[4 * q + 2] = fmaf(f[u], wv.z, acc[4 * q + 2]);
                                acc[4 * q + 3] = fmaf(f[u], wv.w, acc[4 * q + 3]);
                            }
                        }
                    }
                } else {
                    for (int ci = 0; ci < ck; ++ci) {
                        float f = DT<T>::to_f(src[pin * ld + c0 + ci]);
                        if (gate) f *= DT<T>::to_f(gate[pin * gld + c0 + ci]);
                        if (affine) f = f * p.in_scale[c0 + ci] + p.in_shift[c0 + ci];
                        const float4* wr = reinterpret_cast<const float4*>(wbase + ci * CG_CO);
#pragma unroll
                        for (int q = 0; q < CG_CO / 4; ++q) {
                            const float4 wv = wr[q];
                            acc[4 * q + 0] = fmaf(f, wv.x, acc[4 * q + 0]);
                            acc[4 * q + 1] = fmaf(f, wv.y, acc[4 * q + 1]);
                            acc[4 * q + 2] = fmaf(f, wv.z, acc[4 * q + 2]);
                            acc[4 * q + 3] = fmaf(f, wv.w, acc[4 * q + 3]);
                        }
                    }
                }
            }
        }
        coff += cs;
    }
    if (!live) return;
    const T* bias = reinterpret_cast<const T*>(p.bias);
    const T* gz = reinterpret_cast<const T*>(p.gru_z);
    const T* gh = reinterpret_cast<const T*>(p.gru_h);
    T* out = reinterpret_cast<T*>(p.out) + pix * p.out_ld;
#pragma unroll
    for (int j = 0; j < CG_CO; ++j) {
        const int co = co0 + j;
        if (co >= p.cout) break;
        float v = acc[j] + (bias ? DT<T>::to_f(bias[co]) : 0.f);
        v = apply_act(v, p.act);
        if (gz) {
            const float z = DT<T>::to_f(gz[pix * p.gru_z_ld + co]);
            const float hp = DT<T>::to_f(gh[pix * p.gru_h_ld + co]);
            v = (1.f - z) * hp + z * v;
        }
        out[co] = DT<T>::from_f(v);
    }
}

template <typename T, bool U8>
__global__ void resize_bilinear_ex_kernel(const void* __restrict__ x_, int ldx, T* __restrict__ y, int ldy, int n, int h,
                                          int w, int c, int oh, int ow, float sh, float sw) {
    const long long total = static_cast<long long>(n) * oh * ow * c;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int ch = static_cast<int>(idx % c);
        long long t = idx / c;
        const int x_o = static_cast<int>(t % ow); t /= ow;
        const int y_o = static_cast<int>(t % oh);
        const int ni = static_cast<int>(t / oh);
        const float fy = fmaxf((y_o + 0.5f) * sh - 0.5f, 0.f), fx = fmaxf((x_o + 0.5f) * sw - 0.5f, 0.f);
        const int y0 = min(static_cast<int>(fy), h - 1), x0 = min(static_cast<int>(fx), w - 1);
        const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
        const float ly = fy - y0, lx = fx - x0;
        const long long b = static_cast<long long>(ni) * h * w;
        float v00, v01, v10, v11;
        if (U8) {
            const unsigned char* x = reinterpret_cast<const unsigned char*>(x_);
            const float k = 1.0f / 255.0f;
            v00 = x[(b + static_cast<long long>(y0) * w + x0) * ldx + ch] * k;
            v01 = x[(b + static_cast<long long>(y0) * w + x1) * ldx + ch] * k;
            v10 = x[(b + static_cast<long long>(y1) * w + x0) * ldx + ch] * k;
            v11 = x[(b + static_cast<long long>(y1) * w + x1) * ldx + ch] * k;
        } else {
            const T* x = reinterpret_cast<const T*>(x_);
            v00 = DT<T>::to_f(x[(b + static_cast<long long>(y0) * w + x0) * ldx + ch]);
            v01 = DT<T>::to_f(x[(b + static_cast<long long>(y0) * w + x1) * ldx + ch]);
            v10 = DT<T>::to_f(x[(b + static_cast<long long>(y1) * w + x0) * ldx + ch]);
            v11 = DT<T>::to_f(x[(b + static_cast<long long>(y1) * w + x1) * ldx + ch]);
        }
        // PyTorch's order: weights (1-l), l on each axis
        const float top = (1.f - lx) * v00 + lx * v01, bot = (1.f - lx) * v10 + lx * v11;
        y[((static_cast<long long>(ni) * oh + y_o) * ow + x_o) * ldy + ch] = DT<T>::from_f((1.f - ly) * top + ly * bot);
    }
}

template <typename T>
__global__ void avgpool2x_ceil_kernel(const T* __restrict__ x, T* __restrict__ y, int n, int h, int w, int c, int ho, int wo) {
    const long long total = static_cast<long long>(n) * ho * wo * c;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int ch = static_cast<int>(idx % c);
        long long t = idx / c;
        const int xo = static_cast<int>(t % wo); t /= wo;
        const int yo = static_cast<int>(t % ho);
        const int ni = static_cast<int>(t / ho);
        float acc = 0.f;
        int cnt = 0;
        for (int dy = 0; dy < 2; ++dy)
            for (int dx = 0; dx < 2; ++dx) {
                const int yi = 2 * yo + dy, xi = 2 * xo + dx;
                if (yi < h && xi < w) {
                    acc += DT<T>::to_f(x[((static_cast<long long>(ni) * h + yi) * w + xi) * c + ch]);
                    ++cnt;
                }
            }
        y[idx] = DT<T>::from_f(acc / static_cast<float>(cnt));
    }
}

template <typename T>
__global__ void copy_cols_kernel(const T* __restrict__ src, int ls, T* __restrict__ d1, int l1, T* __restrict__ d2, int l2,
                                 long long rows, int cols) {
    const long long total = rows * cols;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = idx / cols;
        const int cc = static_cast<int>(idx % cols);
        const T v = src[r * ls + cc];
        d1[r * l1 + cc] = v;
        if (d2) d2[r * l2 + cc] = v;
    }
}

// one thread per low-res pixel: 3x3 zero-padded box means of x = [src, mean(src)], y = proj, x*y and x*x
template <typename T>
__global__ void guided_stats_kernel(const T* __restrict__ bs, const T* __restrict__ proj, T* __restrict__ stats,
                                    T* __restrict__ mean_x, T* __restrict__ mean_y, int n, int h, int w) {
    const long long total = static_cast<long long>(n) * h * w;
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int xo = static_cast<int>(idx % w);
        const int yo = static_cast<int>((idx / w) % h);
        const long long nb = idx / (static_cast<long long>(w) * h) * h * w;
        float sx[4] = {0, 0, 0, 0}, sy[4] = {0, 0, 0, 0}, sxy[4] = {0, 0, 0, 0}, sxx[4] = {0, 0, 0, 0};
        for (int dy = -1; dy <= 1; ++dy)
            for (int dx = -1; dx <= 1; ++dx) {
                const int yi = yo + dy, xi = xo + dx;
                if (yi < 0 || yi >= h || xi < 0 || xi >= w) continue;
                const long long pp = nb + static_cast<long long>(yi) * w + xi;
                float xv[4], yv[4];
                xv[0] = DT<T>::to_f(bs[pp * 3 + 0]);
                xv[1] = DT<T>::to_f(bs[pp * 3 + 1]);
                xv[2] = DT<T>::to_f(bs[pp * 3 + 2]);
                xv[3] = (xv[0] + xv[1] + xv[2]) * (1.f / 3.f);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    yv[k] = DT<T>::to_f(proj[pp * 4 + k]);
                    sx[k] += xv[k];
                    sy[k] += yv[k];
                    sxy[k] += xv[k] * yv[k];
                    sxx[k] += xv[k] * xv[k];
                }
            }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float mx = sx[k] * (1.f / 9.f), my = sy[k] * (1.f / 9.f);
            mean_x[idx * 4 + k] = DT<T>::from_f(mx);
            mean_y[idx * 4 + k] = DT<T>::from_f(my);
            stats[idx * 8 + k] = DT<T>::from_f(sxy[k] * (1.f / 9.f) - mx * my);
            stats[idx * 8 + 4 + k] = DT<T>::from_f(sxx[k] * (1.f / 9.f) - mx * mx);
        }
    }
}

template <typename T>
__global__ void guided_apply_kernel(const unsigned char* __restrict__ src, const T* __restrict__ A, const T* __restrict__ mx,
                                    const T* __restrict__ my, T* __restrict__ fgr, T* __restrict__ pha,
                                    unsigned char* __restrict__ out_u8, int n, int H, int W, int h, int w, int refine, int mode) {
    const long long total = static_cast<long long>(n) * H * W;
    const float sh = static_cast<float>(h) / static_cast<float>(H), sw = static_cast<float>(w) / static_cast<float>(W);
    for (long long idx = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; idx < total;
         idx += static_cast<long long>(gridDim.x) * blockDim.x) {
        const int xo = static_cast<int>(idx % W);
        const int yo = static_cast<int>((idx / W) % H);
        const long long ni = idx / (static_cast<long long>(W) * H);
        float s[4];
        s[0] = src[idx * 3 + 0] * (1.f / 255.f);
        s[1] = src[idx * 3 + 1] * (1.f / 255.f);
        s[2] = src[idx * 3 + 2] * (1.f / 255.f);
        s[3] = (s[0] + s[1] + s[2]) * (1.f / 3.f);
        float o[4];
        if (refine) {
            const float fy = fmaxf((yo + 0.5f) * sh - 0.5f, 0.f), fx = fmaxf((xo + 0.5f) * sw - 0.5f, 0.f);
            const int y0 = min(static_cast<int>(fy), h - 1), x0 = min(static_cast<int>(fx), w - 1);
            const int y1 = min(y0 + 1, h - 1), x1 = min(x0 + 1, w - 1);
            const float ly = fy - y0, lx = fx - x0;
            const long long nb = ni * h * w;
            const long long q[4] = {nb + static_cast<long long>(y0) * w + x0, nb + static_cast<long long>(y0) * w + x1,
                                    nb + static_cast<long long>(y1) * w + x0, nb + static_cast<long long>(y1) * w + x1};
            const float wq[4] = {(1.f - ly) * (1.f - lx), (1.f - ly) * lx, ly * (1.f - lx), ly * lx};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float av = 0.f, bv = 0.f;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float a = DT<T>::to_f(A[q[j] * 4 + k]);
                    const float b = DT<T>::to_f(my[q[j] * 4 + k]) - a * DT<T>::to_f(mx[q[j] * 4 + k]);
                    av += wq[j] * a;
                    bv += wq[j] * b;
                }
                o[k] = av * s[k] + bv;
            }
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) o[k] = DT<T>::to_f(A[idx * 4 + k]);
        }
        const float a = fminf(fmaxf(o[3], 0.f), 1.f);
        float f[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) f[k] = fminf(fmaxf(o[k] + s[k], 0.f), 1.f);
        if (fgr) {
#pragma unroll
            for (int k = 0; k < 3; ++k) fgr[idx * 3 + k] = DT<T>::from_f(f[k]);
        }
        if (pha) pha[idx] = DT<T>::from_f(a);
        if (out_u8) {
            const float bg[3] = {120.f / 255.f, 255.f / 255.f, 155.f / 255.f};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float v;
                if (mode == 0) v = f[k] * a + bg[k] * (1.f - a);
                else if (mode == 1) v = a;
                else v = f[k] * a;
                out_u8[idx * 3 + k] = static_cast<unsigned char>(__float2int_rn(v * 255.f));
            }
        }
    }
}

}  // namespace ab

#define AB_DISPATCH_M(dtype, ...)                                          \
    do {                                                                   \
        if ((dtype) == AB_BF16) { using T = __nv_bfloat16; __VA_ARGS__; }  \
        else { using T = __half; __VA_ARGS__; }                            \
    } while (0)

extern "C" int ab_conv_generic(const ab_conv_generic_args* a, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(a && a->nsrc >= 1 && a->nsrc <= 3 && a->w && a->out, "bad arguments");
    AB_REQUIRE((a->ks == 1 || a->ks == 3) && a->stride >= 1 && a->dilation >= 1 && a->cout >= 1, "bad conv geometry");
    AB_REQUIRE(!a->gru_z == !a->gru_h, "ConvGRU epilogue needs both z and h");
    AB_REQUIRE(!a->in_scale == !a->in_shift, "input affine needs both scale and shift");
    ab::ConvGenParams p{};
    p.nsrc = a->nsrc;
    for (int s = 0; s < a->nsrc; ++s) {
        AB_REQUIRE(a->src[s] && a->c[s] > 0 && a->ld[s] >= a->c[s], "bad source");
        p.src[s] = a->src[s]; p.gate[s] = a->gate[s]; p.c[s] = a->c[s]; p.ld[s] = a->ld[s]; p.gate_ld[s] = a->gate_ld[s];
        p.ctot += a->c[s];
    }
    p.w = a->w; p.bias = a->bias; p.in_scale = a->in_scale; p.in_shift = a->in_shift;
    p.out = a->out; p.out_ld = a->out_ld;
    p.gru_z = a->gru_z; p.gru_h = a->gru_h; p.gru_z_ld = a->gru_z_ld; p.gru_h_ld = a->gru_h_ld;
    p.n = a->n; p.h = a->h; p.w_in = a->w_in; p.ks = a->ks; p.stride = a->stride; p.dilation = a->dilation;
    p.pad = (a->ks / 2) * a->dilation;
    p.ho = (a->h + 2 * p.pad - a->dilation * (a->ks - 1) - 1) / a->stride + 1;
    p.wo = (a->w_in + 2 * p.pad - a->dilation * (a->ks - 1) - 1) / a->stride + 1;
    p.cout = a->cout; p.cout_pad = (a->cout + ab::CG_CO - 1) / ab::CG_CO * ab::CG_CO; p.act = a->act;
    AB_REQUIRE(p.out_ld >= p.cout && p.ho > 0 && p.wo > 0, "bad output geometry");
    const long long npix = static_cast<long long>(p.n) * p.ho * p.wo;
    dim3 grid(static_cast<unsigned>((npix + ab::CG_THREADS - 1) / ab::CG_THREADS), p.cout_pad / ab::CG_CO);
    const size_t wbytes = static_cast<size_t>(p.ks) * p.ks * p.ctot * ab::CG_CO * sizeof(float);
    if (wbytes <= 100 * 1024) {
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(ab::conv_generic_kernel<__half, true>), 100 * 1024)) return -1;
        if (ab_ensure_max_smem(reinterpret_cast<const void*>(ab::conv_generic_kernel<__nv_bfloat16, true>), 100 * 1024)) return -1;
        AB_DISPATCH_M(a->dtype, (ab::conv_generic_kernel<T, true><<<grid, ab::CG_THREADS, wbytes, stream>>>(p)));
    } else {
        AB_DISPATCH_M(a->dtype, (ab::conv_generic_kernel<T, false><<<grid, ab::CG_THREADS, 0, stream>>>(p)));
    }
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_resize_bilinear_ex(const void* x, int ldx, void* y, int ldy, int n, int h, int w, int c, int oh, int ow,
                                     float step_h, float step_w, int in_u8, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && oh > 0 && ow > 0 && c > 0 && ldx >= c && ldy >= c, "bad arguments");
    const long long total = static_cast<long long>(n) * oh * ow * c;
    if (in_u8) {
        AB_DISPATCH_M(dtype, (ab::resize_bilinear_ex_kernel<T, true><<<ab::grid_m(total, 256), 256, 0, stream>>>(
                                 x, ldx, reinterpret_cast<T*>(y), ldy, n, h, w, c, oh, ow, step_h, step_w)));
    } else {
        AB_DISPATCH_M(dtype, (ab::resize_bilinear_ex_kernel<T, false><<<ab::grid_m(total, 256), 256, 0, stream>>>(
                                 x, ldx, reinterpret_cast<T*>(y), ldy, n, h, w, c, oh, ow, step_h, step_w)));
    }
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_avgpool2x_ceil(const void* x, void* y, int n, int h, int w, int c, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(x && y && n > 0 && h > 0 && w > 0 && c > 0, "bad arguments");
    const int ho = (h + 1) / 2, wo = (w + 1) / 2;
    const long long total = static_cast<long long>(n) * ho * wo * c;
    AB_DISPATCH_M(dtype, (ab::avgpool2x_ceil_kernel<T><<<ab::grid_m(total, 256), 256, 0, stream>>>(
                             reinterpret_cast<const T*>(x), reinterpret_cast<T*>(y), n, h, w, c, ho, wo)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_copy_cols(const void* src, int ld_src, void* dst1, int ld1, void* dst2, int ld2, long long rows, int cols,
                            int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(src && dst1 && rows > 0 && cols > 0, "bad arguments");
    AB_DISPATCH_M(dtype, (ab::copy_cols_kernel<T><<<ab::grid_m(rows * cols, 256), 256, 0, stream>>>(
                             reinterpret_cast<const T*>(src), ld_src, reinterpret_cast<T*>(dst1), ld1,
                             reinterpret_cast<T*>(dst2), ld2, rows, cols)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_guided_stats(const void* base_src, const void* proj, void* stats, void* mean_x, void* mean_y, int n, int h,
                               int w, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(base_src && proj && stats && mean_x && mean_y && n > 0 && h > 0 && w > 0, "bad arguments");
    const long long total = static_cast<long long>(n) * h * w;
    AB_DISPATCH_M(dtype, (ab::guided_stats_kernel<T><<<ab::grid_m(total, 128), 128, 0, stream>>>(
                             reinterpret_cast<const T*>(base_src), reinterpret_cast<const T*>(proj), reinterpret_cast<T*>(stats),
                             reinterpret_cast<T*>(mean_x), reinterpret_cast<T*>(mean_y), n, h, w)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}

extern "C" int ab_guided_apply(const unsigned char* src, const void* a_or_proj, const void* mean_x, const void* mean_y,
                               void* fgr, void* pha, unsigned char* out_u8, int n, int H, int W, int h, int w, int refine,
                               int mode, int dtype, void* stream_) {
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
    AB_REQUIRE(src && a_or_proj && n > 0 && H > 0 && W > 0, "bad arguments");
    AB_REQUIRE(!refine || (mean_x && mean_y && h > 0 && w > 0), "refine needs the low-res means");
    AB_REQUIRE(fgr || pha || out_u8, "no output requested");
    const long long total = static_cast<long long>(n) * H * W;
    AB_DISPATCH_M(dtype, (ab::guided_apply_kernel<T><<<ab::grid_m(total, 256), 256, 0, stream>>>(
                             src, reinterpret_cast<const T*>(a_or_proj), reinterpret_cast<const T*>(mean_x),
                             reinterpret_cast<const T*>(mean_y), reinterpret_cast<T*>(fgr), reinterpret_cast<T*>(pha), out_u8,
                             n, H, W, h, w, refine, mode)));
    AB_CHECK_CUDA(cudaGetLastError());
    ab_count_launch(1);
    return 0;
}
