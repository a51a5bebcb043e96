// Depthwise stencils, transposed convolutions and the fused network tail (CUDA cores).
#pragma once
#include "common.cuh"

namespace isg {

// ---------------------------------------------------------------------------------
// depthwise KHxKW conv with dilation (3x3 d=1/2/4, (5,1), (1,5)) + affine + activation
// (segment.py:64, :91-97, :127, :167).  One thread = one pixel x 4 channels.
// Weights [tap][C]; the bare (5,1) conv has scale 1 / shift = bias / no activation.
// ---------------------------------------------------------------------------------
struct DwParams {
    const void* in; void* out; int H, W, Cs, total /* B*H*W*(Cs/4) */;
    int KH, KW, dil_h, dil_w, pad_h, pad_w;
    const float* w; const float* scale; const float* shift; const float* slope; int act;
};

template <typename T>
__global__ void __launch_bounds__(256) dw_kernel(DwParams p) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.total) return;
    const int cg = p.Cs >> 2;
    const int c = (i % cg) * 4, pix = i / cg;
    const int x = pix % p.W, y = (pix / p.W) % p.H, b = pix / (p.W * p.H);
    const T* in = static_cast<const T*>(p.in);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int ky = 0; ky < p.KH; ++ky) {
        int iy = y - p.pad_h + ky * p.dil_h;
        if (iy < 0 || iy >= p.H) continue;
        for (int kx = 0; kx < p.KW; ++kx) {
            int ix = x - p.pad_w + kx * p.dil_w;
            if (ix < 0 || ix >= p.W) continue;
            float4 v = ld4(in + (((size_t)b * p.H + iy) * p.W + ix) * p.Cs + c);
            float4 ww = __ldg(reinterpret_cast<const float4*>(p.w + (ky * p.KW + kx) * p.Cs + c));
            acc.x = fmaf(v.x, ww.x, acc.x); acc.y = fmaf(v.y, ww.y, acc.y);
            acc.z = fmaf(v.z, ww.z, acc.z); acc.w = fmaf(v.w, ww.w, acc.w);
        }
    }
    float4 sc = __ldg(reinterpret_cast<const float4*>(p.scale + c));
    float4 sh = __ldg(reinterpret_cast<const float4*>(p.shift + c));
    float4 sl = p.slope ? __ldg(reinterpret_cast<const float4*>(p.slope + c)) : make_float4(0, 0, 0, 0);
    float4 r;
    r.x = apply_act(fmaf(acc.x, sc.x, sh.x), p.act, sl.x);
    r.y = apply_act(fmaf(acc.y, sc.y, sh.y), p.act, sl.y);
    r.z = apply_act(fmaf(acc.z, sc.z, sh.z), p.act, sl.z);
    r.w = apply_act(fmaf(acc.w, sc.w, sh.w), p.act, sl.w);
    st4(static_cast<T*>(p.out) + (size_t)pix * p.Cs + c, r);
}

// ---------------------------------------------------------------------------------
// 2x2 max-pool, NHWC (F.max_pool2d, segment.py:145): materialises `residual_1`, which both
// the block's own 1x1 skip conv and the decoder's concat (segment.py:331) consume.
// ---------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) pool2_kernel(const T* __restrict__ in, T* __restrict__ out, int Ho, int Wo,
                                                    int Cs, int total /* B*Ho*Wo*(Cs/4) */) {
    asm volatile("griddepcontrol.wait;" ::: "memory");   // no-op unless launched with programmatic serialization
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int cg = Cs >> 2, c = (i % cg) * 4, pix = i / cg;
    const int x = pix % Wo, y = (pix / Wo) % Ho, b = pix / (Wo * Ho);
    const T* s = in + (((size_t)b * 2 * Ho + 2 * y) * 2 * Wo + 2 * x) * Cs + c;
    const size_t row = (size_t)2 * Wo * Cs;
    st4(out + (size_t)pix * Cs + c, max4(max4(ld4(s), ld4(s + Cs)), max4(ld4(s + row), ld4(s + row + Cs))));
}

// ---------------------------------------------------------------------------------
// ConvTranspose2d k4 s2 p1, C -> C (C = 16 or 4) + BatchNorm + ReLU (segment.py:305-308).
// Gather form: output (oy,ox) sees the 2x2 inputs with ky = (oy+1)%2 + {0,2}.
// Weights [ky][kx][ci(CS)][co(C)], CS = storage stride of the input (pad rows are zero).
// ---------------------------------------------------------------------------------
struct ConvT4Params {
    const void* in; void* out; int Hi, Wi, Cs_in, Cs_out, total_pix /* B*2Hi*2Wi */;
    const float* w; const float* scale; const float* shift;
};

template <typename T, int C>
__global__ void __launch_bounds__(128) convt4_kernel(ConvT4Params p) {
    extern __shared__ float smem[];  // [16][Cs_in][C]
    for (int i = threadIdx.x; i < 16 * p.Cs_in * C; i += blockDim.x) smem[i] = p.w[i];
    __syncthreads();
    int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= p.total_pix) return;
    const int Ho = 2 * p.Hi, Wo = 2 * p.Wi;
    const int ox = pix % Wo, oy = (pix / Wo) % Ho, b = pix / (Wo * Ho);
    const T* in = static_cast<const T*>(p.in);
    float acc[C];
#pragma unroll
    for (int o = 0; o < C; ++o) acc[o] = 0.f;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
        int ky = ((oy + 1) & 1) + 2 * a, iy = (oy + 1 - ky) >> 1;
        if (iy < 0 || iy >= p.Hi) continue;
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
            int kx = ((ox + 1) & 1) + 2 * bb, ix = (ox + 1 - kx) >> 1;
            if (ix < 0 || ix >= p.Wi) continue;
            const T* src = in + (((size_t)b * p.Hi + iy) * p.Wi + ix) * p.Cs_in;
            const float* wt = smem + (size_t)(ky * 4 + kx) * p.Cs_in * C;
            for (int c = 0; c < p.Cs_in; c += 4) {
                float4 xv = ld4(src + c);
                const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float4* wr = reinterpret_cast<const float4*>(wt + (c + e) * C);
#pragma unroll
                    for (int j = 0; j < C / 4; ++j) {
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
    T* o = static_cast<T*>(p.out) + (size_t)pix * p.Cs_out;
#pragma unroll
    for (int j = 0; j < C / 4; ++j) {
        float r[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            int c = 4 * j + e;
            r[e] = fmaxf(fmaf(acc[c], p.scale[c], p.shift[c]), 0.f);
        }
        st4(o + 4 * j, make_float4(r[0], r[1], r[2], r[3]));
    }
}

// ---------------------------------------------------------------------------------
// network tail, fused: ConvTranspose2d k8 s4 p2 16->4 (+bias) followed by Conv2d 3x3 p1
// 4->1 (+bias) (segment.py:435-438, :504-505).  The [4,H,W] map lives only in shared
// memory.  CTA = 32x32 logits; phase 1 builds the 34x34x4 transposed-conv tile (zero outside
// the image = the 3x3's padding), phase 2 applies the 3x3.  Output fp32 [B,1,H,W].
// ---------------------------------------------------------------------------------
constexpr int TL = 32, TLH = TL + 2, TIN = TL / 4 + 3;  // input tile 11x11 covers (34+7)/4
struct TailParams {
    const void* in; int Hq, Wq, Cs_in;  // [B,Hq,Wq,16]
    float* logits;                      // [B,4Hq,4Wq]
    const float* wt;                    // [ky(8)][kx(8)][ci(16)][co(4)]
    const float* bt;                    // [4]
    const float* w3;                    // [ky(3)][kx(3)][ci(4)]
    float b3;
};

template <typename T>
__global__ void __launch_bounds__(256) tail_kernel(TailParams p) {
    __shared__ float4 swt[64 * 16];           // 16 KB
    __shared__ float sin_[TIN * TIN * 16];    // 7.7 KB
    __shared__ float4 mid[TLH * TLH];         // 18.5 KB
    __shared__ float sw3[36];
    const int H = 4 * p.Hq, W = 4 * p.Wq;
    const int x0 = blockIdx.x * TL, y0 = blockIdx.y * TL, b = blockIdx.z;
    const int tid = threadIdx.x;
    for (int i = tid; i < 64 * 16; i += blockDim.x) swt[i] = reinterpret_cast<const float4*>(p.wt)[i];
    if (tid < 36) sw3[tid] = p.w3[tid];
    // input tile origin: smallest iy feeding output row y0-1: iy = floor((y0-1+2-7)/4) = y0/4 - 2
    const int iy0 = y0 / 4 - 2, ix0 = x0 / 4 - 2;
    const T* in = static_cast<const T*>(p.in);
    for (int i = tid; i < TIN * TIN * 4; i += blockDim.x) {
        int c4 = (i & 3) * 4, cell = i >> 2, r = cell / TIN, cidx = cell % TIN;
        int iy = iy0 + r, ix = ix0 + cidx;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (iy >= 0 && iy < p.Hq && ix >= 0 && ix < p.Wq)
            v = ld4(in + (((size_t)b * p.Hq + iy) * p.Wq + ix) * p.Cs_in + c4);
        *reinterpret_cast<float4*>(&sin_[cell * 16 + c4]) = v;
    }
    __syncthreads();
    const float4 bias = make_float4(p.bt[0], p.bt[1], p.bt[2], p.bt[3]);
    for (int i = tid; i < TLH * TLH; i += blockDim.x) {
        int my = i / TLH, mx = i % TLH;
        int oy = y0 - 1 + my, ox = x0 - 1 + mx;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (oy >= 0 && oy < H && ox >= 0 && ox < W) {
            acc = bias;
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                int ky = ((oy + 2) & 3) + 4 * a, iy = (oy + 2 - ky) >> 2;  // exact division
                if (iy < 0 || iy >= p.Hq) continue;
#pragma unroll
                for (int bb = 0; bb < 2; ++bb) {
                    int kx = ((ox + 2) & 3) + 4 * bb, ix = (ox + 2 - kx) >> 2;
                    if (ix < 0 || ix >= p.Wq) continue;
                    const float* src = &sin_[((iy - iy0) * TIN + (ix - ix0)) * 16];
                    const float4* wr = &swt[(ky * 8 + kx) * 16];
#pragma unroll
                    for (int c = 0; c < 16; ++c) {
                        float v = src[c];
                        float4 ww = wr[c];
                        acc.x = fmaf(v, ww.x, acc.x); acc.y = fmaf(v, ww.y, acc.y);
                        acc.z = fmaf(v, ww.z, acc.z); acc.w = fmaf(v, ww.w, acc.w);
                    }
                }
            }
        }
        mid[i] = acc;
    }
    __syncthreads();
    for (int i = tid; i < TL * TL; i += blockDim.x) {
        int ty = i / TL, tx = i % TL;
        int oy = y0 + ty, ox = x0 + tx;
        if (oy >= H || ox >= W) continue;
        float acc = p.b3;
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                float4 m = mid[(ty + ky) * TLH + tx + kx];
                const float* w = &sw3[(ky * 3 + kx) * 4];
                acc = fmaf(m.x, w[0], acc); acc = fmaf(m.y, w[1], acc);
                acc = fmaf(m.z, w[2], acc); acc = fmaf(m.w, w[3], acc);
            }
        p.logits[((size_t)b * H + oy) * W + ox] = acc;
    }
}

}  // namespace isg
