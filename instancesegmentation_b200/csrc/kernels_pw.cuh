// Generalised pointwise (1x1) convolution on CUDA cores with fused prologue/epilogue.
//   out = act_out( act_a(affine_a(Wa . gather_a(p)))  +  branch_b(p) )
// gather_* concatenates the channels of up to two NHWC tensors (torch.cat,
// segment.py:331,485,494) read either directly, through nearest-x2 upsampling
// (UpsamplingNearest2d, :322) or through a 2x2 max-pool (F.max_pool2d, :145);
// branch_b is nothing, an identity residual (`out += residual`, :76) or a second 1x1
// conv with its own affine (convm :137, resconv :192, uppool.1 :323).  The concatenated /
// upsampled / pooled tensors are never written to memory (except the pooled map that the
// decoder needs later, `residual_1` :150).
#pragma once
#include "common.cuh"

namespace isg {

enum SrcMode { SRC_DIRECT = 0, SRC_UP2 = 1, SRC_POOL2 = 2 };
enum BranchB { BR_NONE = 0, BR_IDENTITY = 1, BR_GEMM = 2 };

struct PwSrc {
    const void* p0; int Cs0;
    const void* p1; int Cs1;  // second concat operand (nullptr / 0 when absent)
    int mode;
};

struct PwParams {
    int total_pix, Ho, Wo;
    PwSrc a; const float* wa; const float* scale_a; const float* shift_a; const float* slope_a; int act_a;
    int b_kind; PwSrc b; const float* wb; const float* scale_b; const float* shift_b;
    const void* res; int res_Cs;
    int act_out; const float* slope_out;
    void* out; int out_Cs, Cout, CoutPad;
    void* pool_out; int pool_Cs;  // SRC_POOL2 on branch b only: materialise the pooled map
};

// accumulate one source tensor row (C channels at `src`) against weights rows [k0, k0+C)
template <typename T, int CO>
__device__ __forceinline__ void pw_accum(const T* __restrict__ src, int C, const float* __restrict__ sw,
                                         float (&acc)[CO]) {
    for (int c = 0; c < C; c += 4) {
        float4 xv = ld4(src + c);
        const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float4* wr = reinterpret_cast<const float4*>(sw + (c + e) * CO);
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

template <typename T, int CO>
__device__ __forceinline__ void pw_pool_accum(const T* __restrict__ src, int C, int rowstride,
                                              const float* __restrict__ sw, float (&acc)[CO],
                                              T* __restrict__ pool_dst) {
    for (int c = 0; c < C; c += 4) {
        float4 xv = max4(max4(ld4(src + c), ld4(src + C + c)),
                         max4(ld4(src + rowstride + c), ld4(src + rowstride + C + c)));
        if (pool_dst) st4(pool_dst + c, xv);
        const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float4* wr = reinterpret_cast<const float4*>(sw + (c + e) * CO);
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

template <typename T, int CO>
__device__ __forceinline__ void pw_branch(const PwSrc& s, const float* sw, int b, int oy, int ox, int Ho,
                                          int Wo, float (&acc)[CO], T* pool_dst) {
    if (s.mode == SRC_POOL2) {
        const int Wi = 2 * Wo;
        const T* src = static_cast<const T*>(s.p0) + (((size_t)b * 2 * Ho + 2 * oy) * Wi + 2 * ox) * s.Cs0;
        pw_pool_accum<T, CO>(src, s.Cs0, Wi * s.Cs0, sw, acc, pool_dst);
        return;
    }
    size_t pix;
    if (s.mode == SRC_UP2) pix = ((size_t)b * (Ho >> 1) + (oy >> 1)) * (Wo >> 1) + (ox >> 1);
    else pix = ((size_t)b * Ho + oy) * Wo + ox;
    pw_accum<T, CO>(static_cast<const T*>(s.p0) + pix * s.Cs0, s.Cs0, sw, acc);
    if (s.p1) pw_accum<T, CO>(static_cast<const T*>(s.p1) + pix * s.Cs1, s.Cs1, sw + s.Cs0 * CO, acc);
}

// grid: (ceil(total_pix/128), CoutPad/CO); weights of this channel chunk live in smem.
template <typename T, int CO>
__global__ void __launch_bounds__(128) pw_kernel(PwParams p) {
    extern __shared__ float smem[];
    const int Ka = p.a.Cs0 + p.a.Cs1;
    const int Kb = p.b_kind == BR_GEMM ? p.b.Cs0 + p.b.Cs1 : 0;
    const int co0 = blockIdx.y * CO;
    float* swa = smem;
    float* swb = smem + Ka * CO;
    for (int i = threadIdx.x; i < Ka * CO; i += blockDim.x)
        swa[i] = p.wa[(size_t)(i / CO) * p.CoutPad + co0 + (i % CO)];
    for (int i = threadIdx.x; i < Kb * CO; i += blockDim.x)
        swb[i] = p.wb[(size_t)(i / CO) * p.CoutPad + co0 + (i % CO)];
    __syncthreads();
    const int pix = blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= p.total_pix) return;
    const int ox = pix % p.Wo, oy = (pix / p.Wo) % p.Ho, b = pix / (p.Wo * p.Ho);

    float acc[CO];
#pragma unroll
    for (int o = 0; o < CO; ++o) acc[o] = 0.f;
    pw_branch<T, CO>(p.a, swa, b, oy, ox, p.Ho, p.Wo, acc, nullptr);
#pragma unroll
    for (int o = 0; o < CO; ++o) {
        int c = co0 + o;
        acc[o] = apply_act(fmaf(acc[o], p.scale_a[c], p.shift_a[c]), p.act_a, p.slope_a ? p.slope_a[c] : 0.f);
    }
    if (p.b_kind == BR_GEMM) {
        float accb[CO];
#pragma unroll
        for (int o = 0; o < CO; ++o) accb[o] = 0.f;
        T* pool_dst = (p.pool_out && blockIdx.y == 0)
                          ? static_cast<T*>(p.pool_out) + (size_t)pix * p.pool_Cs : nullptr;
        pw_branch<T, CO>(p.b, swb, b, oy, ox, p.Ho, p.Wo, accb, pool_dst);
#pragma unroll
        for (int o = 0; o < CO; ++o) acc[o] += fmaf(accb[o], p.scale_b[co0 + o], p.shift_b[co0 + o]);
    } else if (p.b_kind == BR_IDENTITY) {
        const T* r = static_cast<const T*>(p.res) + (size_t)pix * p.res_Cs + co0;
#pragma unroll
        for (int j = 0; j < CO / 4; ++j) {
            if (co0 + 4 * j < p.Cout) {
                float4 v = ld4(r + 4 * j);
                acc[4 * j + 0] += v.x; acc[4 * j + 1] += v.y; acc[4 * j + 2] += v.z; acc[4 * j + 3] += v.w;
            }
        }
    }
    T* o = static_cast<T*>(p.out) + (size_t)pix * p.out_Cs + co0;
#pragma unroll
    for (int j = 0; j < CO / 4; ++j) {
        if (co0 + 4 * j < p.Cout) {  // Cout is a multiple of 4 on every layer of this network
            float r[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                int c = co0 + 4 * j + e;
                r[e] = apply_act(acc[4 * j + e], p.act_out, p.slope_out ? p.slope_out[c] : 0.f);
            }
            st4(o + 4 * j, make_float4(r[0], r[1], r[2], r[3]));
        }
    }
}

}  // namespace isg
