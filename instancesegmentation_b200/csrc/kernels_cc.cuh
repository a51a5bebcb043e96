// CUDA-core (fp32 FMA) kernels of the Segment forward.  They implement ISG_MODE_FP32
// completely and the non-GEMM layers of ISG_MODE_BF16 (T = bf16 storage, fp32 math).
// Layout: activations NHWC with channel stride Cs (multiple of 8, pad channels are zero).
// Every conv applies its bias/BatchNorm as an fp32 epilogue affine y = acc*scale + shift
// (reference op order act(bn(conv(x))), model/segment.py:44-45) followed by the activation.
#pragma once
#include "common.cuh"

namespace isg {

// ---------------------------------------------------------------------------------
// stem layer1: 5x5 stride 2 pad 2, Cin -> 16, reads the NCHW fp32 network input
// (model/segment.py:23-24).  One thread = one output pixel x 16 channels.
// ---------------------------------------------------------------------------------
constexpr int S1_TX = 32, S1_TY = 8, S1_TR = 2 * S1_TY + 3, S1_TC = 2 * S1_TX + 3 + 1, S1_CC = 4;

template <typename T>
__global__ void __launch_bounds__(S1_TX* S1_TY)
stem1_kernel(const float* __restrict__ x, int Cin, int H, int W, T* __restrict__ out,
             const float* __restrict__ w, const float* __restrict__ scale,
             const float* __restrict__ shift, const float* __restrict__ slope) {
    extern __shared__ float smem[];
    float* sw = smem;                      // [Cin*25][16]
    float* tile = smem + Cin * 25 * 16;    // [S1_CC][S1_TR][S1_TC]
    const int Ho = H >> 1, Wo = W >> 1;
    const int tx = threadIdx.x % S1_TX, ty = threadIdx.x / S1_TX;
    const int ox0 = blockIdx.x * S1_TX, oy0 = blockIdx.y * S1_TY, b = blockIdx.z;
    for (int i = threadIdx.x; i < Cin * 25 * 16; i += blockDim.x) sw[i] = w[i];
    float acc[16];
#pragma unroll
    for (int o = 0; o < 16; ++o) acc[o] = 0.f;
    for (int c0 = 0; c0 < Cin; c0 += S1_CC) {
        const int nc = min(S1_CC, Cin - c0);
        __syncthreads();
        for (int i = threadIdx.x; i < nc * S1_TR * S1_TC; i += blockDim.x) {
            int c = i / (S1_TR * S1_TC), r = (i / S1_TC) % S1_TR, col = i % S1_TC;
            int iy = 2 * oy0 - 2 + r, ix = 2 * ox0 - 2 + col;
            float v = 0.f;
            if (iy >= 0 && iy < H && ix >= 0 && ix < W)
                v = __ldg(x + (((size_t)b * Cin + c0 + c) * H + iy) * W + ix);
            tile[i] = v;
        }
        __syncthreads();
        for (int c = 0; c < nc; ++c) {
#pragma unroll
            for (int ky = 0; ky < 5; ++ky) {
#pragma unroll
                for (int kx = 0; kx < 5; ++kx) {
                    float v = tile[(c * S1_TR + 2 * ty + ky) * S1_TC + 2 * tx + kx];
                    const float4* wr = reinterpret_cast<const float4*>(sw + ((c0 + c) * 25 + ky * 5 + kx) * 16);
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        float4 ww = wr[j];
                        acc[4 * j + 0] = fmaf(v, ww.x, acc[4 * j + 0]);
                        acc[4 * j + 1] = fmaf(v, ww.y, acc[4 * j + 1]);
                        acc[4 * j + 2] = fmaf(v, ww.z, acc[4 * j + 2]);
                        acc[4 * j + 3] = fmaf(v, ww.w, acc[4 * j + 3]);
                    }
                }
            }
        }
    }
    const int oy = oy0 + ty, ox = ox0 + tx;
    if (oy < Ho && ox < Wo) {
        T* o = out + (((size_t)b * Ho + oy) * Wo + ox) * 16;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float r[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                int c = 4 * j + e;
                r[e] = apply_act(fmaf(acc[c], scale[c], shift[c]), ACT_PRELU, slope[c]);
            }
            st4(o + 4 * j, make_float4(r[0], r[1], r[2], r[3]));
        }
    }
}

// ---------------------------------------------------------------------------------
// stem max-pool 4x4 stride 4 of the raw NCHW input into channels [0,Cin) of init_down
// (model/segment.py:29,31 -- pooled input comes FIRST in the concat).
// ---------------------------------------------------------------------------------
template <typename T>
__global__ void pool4_kernel(const float* __restrict__ x, int Cin, int H, int W, T* __restrict__ out,
                             int Cs_out, int total /* B*Cin*Ho*Wo */) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int Ho = H >> 2, Wo = W >> 2;
    int ox = i % Wo, oy = (i / Wo) % Ho, c = (i / (Wo * Ho)) % Cin, b = i / (Wo * Ho * Cin);
    const float* p = x + (((size_t)b * Cin + c) * H + 4 * oy) * W + 4 * ox;
    float m = -INFINITY;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        float4 v = *reinterpret_cast<const float4*>(p + (size_t)r * W);
        m = fmaxf(m, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
    }
    st1(out + (((size_t)b * Ho + oy) * Wo + ox) * Cs_out + c, m);
}

// ---------------------------------------------------------------------------------
// generic small dense conv, NHWC, Cout = CO (16 or 4): stem layer2 (5x5 s2), the 2x2 s2
// convs of BottleneckDown2 (segment.py:121) and the dense 3x3 of BottleneckDim (:242).
// Weights [(ky,kx,ci)][CO] in shared memory; one thread = one output pixel x CO channels.
// ---------------------------------------------------------------------------------
struct ConvKParams {
    const void* in; int Hi, Wi, Cs_in;
    void* out; int Ho, Wo, Cs_out, co_off;
    int KH, KW, stride, pad;
    const float* w; const float* scale; const float* shift; const float* slope;
    int act, total_pix;
};

template <typename T, int CO>
__global__ void __launch_bounds__(128) convk_kernel(ConvKParams p) {
    extern __shared__ float smem[];
    const int K = p.KH * p.KW * p.Cs_in;
    for (int i = threadIdx.x; i < K * CO; i += blockDim.x) smem[i] = p.w[i];
    __syncthreads();
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= p.total_pix) return;
    const int ox = pix % p.Wo, oy = (pix / p.Wo) % p.Ho, b = pix / (p.Wo * p.Ho);
    const T* in = static_cast<const T*>(p.in);
    float acc[CO];
#pragma unroll
    for (int o = 0; o < CO; ++o) acc[o] = 0.f;
    for (int ky = 0; ky < p.KH; ++ky) {
        int iy = oy * p.stride - p.pad + ky;
        if (iy < 0 || iy >= p.Hi) continue;
        for (int kx = 0; kx < p.KW; ++kx) {
            int ix = ox * p.stride - p.pad + kx;
            if (ix < 0 || ix >= p.Wi) continue;
            const T* src = in + (((size_t)b * p.Hi + iy) * p.Wi + ix) * p.Cs_in;
            const float* wt = smem + (size_t)(ky * p.KW + kx) * p.Cs_in * CO;
            for (int c = 0; c < p.Cs_in; c += 4) {
                float4 xv = ld4(src + c);
                const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float4* wr = reinterpret_cast<const float4*>(wt + (c + e) * CO);
#pragma unroll
                    for (int j = 0; j < CO / 4; ++j) {
                        float4 ww = wr[j];
                        acc[4 * j + 0] = fmaf(xs[e], ww.x, acc[4 * j + 0]);
                        acc[4 * j + 1] = fmaf(xs[e], ww.y, acc[4 * j + 1]);
                        acc[4 * j + 2] = fmaf(xs[e], ww.z, acc[4 * j + 2]);
                        acc[4 * j + 3] = fmaf(xs[e], ww.w, acc[4 * j + 3]);
                    }
                }
            }
        }
    }
    T* o = static_cast<T*>(p.out) + (size_t)pix * p.Cs_out + p.co_off;
    const bool vec = (p.co_off % 4 == 0);
#pragma unroll
    for (int j = 0; j < CO / 4; ++j) {
        float r[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int c = 4 * j + e;
            r[e] = apply_act(fmaf(acc[c], p.scale[c], p.shift[c]), p.act, p.slope ? p.slope[c] : 0.f);
        }
        if (vec) st4(o + 4 * j, make_float4(r[0], r[1], r[2], r[3]));
        else {
#pragma unroll
            for (int e = 0; e < 4; ++e) st1(o + 4 * j + e, r[e]);
        }
    }
}

}  // namespace isg
