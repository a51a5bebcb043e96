// Mask decode: logits -> sigmoid -> bilinear resize to each instance's box -> threshold ->
// per-instance 0/1 masks (+ optional soft bytes), area and bbox by warp/block reductions,
// and the per-image id map.  Memory-bound by design: coalesced 16-byte logit loads, 4-byte
// mask stores, one atomic set per CTA.  Semantics: DESIGN.md section 3; sigmoid =
// model/segment.py:534, soft byte = train_instance.py:398-399.
#include <climits>

#include "decode_math.cuh"

namespace isg {
namespace {

constexpr int DEC_THREADS = 256, DEC_ITERS = 4, DEC_PPB = DEC_THREADS * 4 * DEC_ITERS;

__global__ void decode_init_kernel(int n_inst, int32_t* area, int32_t* bbox) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_inst) return;
    area[n] = 0;
    bbox[4 * n] = INT_MAX; bbox[4 * n + 1] = INT_MAX; bbox[4 * n + 2] = -1; bbox[4 * n + 3] = -1;
}

__global__ void decode_finalize_kernel(int n_inst, const int32_t* boxes, const int32_t* area, int32_t* bbox) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_inst) return;
    if (area[n] == 0) {
        bbox[4 * n] = bbox[4 * n + 1] = bbox[4 * n + 2] = bbox[4 * n + 3] = 0;
    } else {
        int x0 = boxes[4 * n], y0 = boxes[4 * n + 1];
        bbox[4 * n] += x0; bbox[4 * n + 1] += y0; bbox[4 * n + 2] += x0 + 1; bbox[4 * n + 3] += y0 + 1;
    }
}

// One CTA = DEC_PPB consecutive pixels of one instance's flat hk*wk mask.
template <typename TL, bool SOFT>
__global__ void __launch_bounds__(DEC_THREADS)
decode_kernel(const TL* __restrict__ logits, int h, int w, const int32_t* __restrict__ boxes,
              const int64_t* __restrict__ mask_off, float thr, uint8_t* __restrict__ masks,
              uint8_t* __restrict__ soft, int32_t* __restrict__ area, int32_t* __restrict__ bbox) {
    const int n = blockIdx.y;
    const int wk = boxes[4 * n + 2], hk = boxes[4 * n + 3];
    const int npix = wk * hk;
    const int start = blockIdx.x * DEC_PPB;
    if (start >= npix) return;
    const TL* L = logits + (size_t)n * h * w;
    uint8_t* dst = masks + mask_off[n];
    uint8_t* sdst = SOFT ? soft + mask_off[n] : nullptr;
    const bool identity = (wk == w && hk == h);
    const float sy = __fdiv_rn((float)h, (float)hk), sx = __fdiv_rn((float)w, (float)wk);

    int cnt = 0, xmin = INT_MAX, ymin = INT_MAX, xmax = -1, ymax = -1;
#pragma unroll
    for (int it = 0; it < DEC_ITERS; ++it) {
        const int idx = start + (it * DEC_THREADS + threadIdx.x) * 4;
        if (idx >= npix) break;
        const int nv = min(4, npix - idx);
        int oy = idx / wk, ox = idx - oy * wk;
        unsigned mbits = 0, sbits = 0;
        float lv[4];
        bool have = false;
        if (identity && sizeof(TL) == 4 && nv == 4 && ((reinterpret_cast<uintptr_t>(L + idx) & 15) == 0)) {
            float4 v = __ldcs(reinterpret_cast<const float4*>(L + idx));
            lv[0] = v.x; lv[1] = v.y; lv[2] = v.z; lv[3] = v.w;
            have = true;
        }
        Tap ty = make_tap(oy, h, sy);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (e < nv) {
                float p;
                bool decided = false, m = false;
                if (identity) {
                    float x = have ? lv[e] : ldl(L + idx + e);
                    if (!SOFT) {
                        p = fast_sigmoid(x);
                        if (fabsf(p - thr) > DECIDE_MARGIN) { m = p > thr; decided = true; }
                    }
                    if (!decided) { p = exact_sigmoid(x); m = p > thr; }
                } else {
                    Tap tx = make_tap(ox, w, sx);
                    if (!SOFT) {
                        p = fast_prob(L, w, ty, tx);
                        if (fabsf(p - thr) > DECIDE_MARGIN) { m = p > thr; decided = true; }
                    }
                    if (!decided) { p = exact_prob(L, w, ty, tx); m = p > thr; }
                }
                if (m) {
                    mbits |= 1u << (8 * e);
                    ++cnt;
                    xmin = min(xmin, ox); xmax = max(xmax, ox);
                    ymin = min(ymin, oy); ymax = max(ymax, oy);
                }
                if (SOFT) sbits |= soft_byte(p) << (8 * e);
                if (++ox == wk) { ox = 0; ++oy; ty = make_tap(oy, h, sy); }
            }
        }
        if (nv == 4 && ((reinterpret_cast<uintptr_t>(dst + idx) & 3) == 0)) {
            __stcs(reinterpret_cast<unsigned*>(dst + idx), mbits);
            if (SOFT) __stcs(reinterpret_cast<unsigned*>(sdst + idx), sbits);
        } else {
            for (int e = 0; e < nv; ++e) {
                dst[idx + e] = (uint8_t)((mbits >> (8 * e)) & 0xff);
                if (SOFT) sdst[idx + e] = (uint8_t)((sbits >> (8 * e)) & 0xff);
            }
        }
    }
    // warp reduction, then one set of atomics per CTA
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    xmin = __reduce_min_sync(0xffffffffu, xmin); ymin = __reduce_min_sync(0xffffffffu, ymin);
    xmax = __reduce_max_sync(0xffffffffu, xmax); ymax = __reduce_max_sync(0xffffffffu, ymax);
    __shared__ int red[DEC_THREADS / 32][5];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { red[warp][0] = cnt; red[warp][1] = xmin; red[warp][2] = ymin; red[warp][3] = xmax; red[warp][4] = ymax; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < DEC_THREADS / 32; ++i) {
            cnt += red[i][0];
            xmin = min(xmin, red[i][1]); ymin = min(ymin, red[i][2]);
            xmax = max(xmax, red[i][3]); ymax = max(ymax, red[i][4]);
        }
        if (cnt) {
            atomicAdd(area + n, cnt);
            atomicMin(bbox + 4 * n, xmin); atomicMin(bbox + 4 * n + 1, ymin);
            atomicMax(bbox + 4 * n + 2, xmax); atomicMax(bbox + 4 * n + 3, ymax);
        }
    }
}

// one thread per canvas pixel; exact probabilities so that comparisons between instances
// are bit-identical to the specification
template <typename TL>
__global__ void __launch_bounds__(256)
idmap_kernel(const TL* __restrict__ logits, int h, int w, const int32_t* __restrict__ boxes,
             const int32_t* __restrict__ img_first, int Hs, int Ws, float thr, int32_t* __restrict__ idmap) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, img = blockIdx.z;
    if (x >= Ws) return;
    int best = 0;
    float bp = 0.f;
    const int n1 = img_first[img + 1];
    for (int n = img_first[img]; n < n1; ++n) {
        const int4 bx = __ldg(reinterpret_cast<const int4*>(boxes) + n);
        const int ox = x - bx.x, oy = y - bx.y;
        if (ox < 0 || oy < 0 || ox >= bx.z || oy >= bx.w) continue;
        const TL* L = logits + (size_t)n * h * w;
        Tap ty = make_tap(oy, h, __fdiv_rn((float)h, (float)bx.w));
        Tap tx = make_tap(ox, w, __fdiv_rn((float)w, (float)bx.z));
        float pf = fast_prob(L, w, ty, tx);
        if (pf < thr - DECIDE_MARGIN) continue;  // certainly not set (NaN falls through to exact)
        float p = exact_prob(L, w, ty, tx);
        if (p > thr && (best == 0 || p > bp)) { best = n + 1; bp = p; }
    }
    idmap[((size_t)img * Hs + y) * Ws + x] = best;
}

__global__ void iou_counts_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int64_t n,
                                  unsigned long long* out) {
    unsigned inter = 0, uni = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        bool pa = a[i] != 0, pb = b[i] != 0;
        inter += pa && pb; uni += pa || pb;
    }
    inter = __reduce_add_sync(0xffffffffu, inter);
    uni = __reduce_add_sync(0xffffffffu, uni);
    if ((threadIdx.x & 31) == 0) {
        if (inter) atomicAdd(out, (unsigned long long)inter);
        if (uni) atomicAdd(out + 1, (unsigned long long)uni);
    }
}

// train_instance.py:297,380 -- torch.optim.Adam defaults, applied to the flat bucket
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, float gs, float lr, float b1, float b2, float eps,
                            float bc1, float bc2_sqrt) {
    int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float grad = g[i] * gs;
    float mi = m[i] + (grad - m[i]) * (1.f - b1);            // lerp form used by torch
    float vi = v[i] * b2 + grad * grad * (1.f - b2);
    m[i] = mi; v[i] = vi;
    float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] -= (lr / bc1) * (mi / denom);
}

}  // namespace
}  // namespace isg

using namespace isg;

extern "C" {

int isg_decode(const void* logits_dev, int logits_bf16, int n_inst, int h, int w, const int32_t* boxes_dev,
               const int64_t* mask_off_dev, int max_wk, int max_hk, float threshold, uint8_t* masks_dev,
               uint8_t* soft_dev, int32_t* area_dev, int32_t* bbox_dev, void* stream) {
    if (n_inst == 0) return ISG_OK;
    if (!logits_dev || !boxes_dev || !mask_off_dev || !masks_dev || !area_dev || !bbox_dev)
        return fail(ISG_ERR_ARG, "isg_decode: null argument");
    if (n_inst < 0 || n_inst > 65535 || h < 1 || w < 1 || max_wk < 1 || max_hk < 1)
        return fail(ISG_ERR_ARG, "isg_decode: bad sizes (n_inst=%d h=%d w=%d max box %dx%d)", n_inst, h, w, max_wk, max_hk);
    if ((int64_t)max_wk * max_hk > INT_MAX) return fail(ISG_ERR_ARG, "isg_decode: box too large");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    decode_init_kernel<<<cdiv(n_inst, 128), 128, 0, st>>>(n_inst, area_dev, bbox_dev);
    dim3 grid(cdiv(max_wk * max_hk, DEC_PPB), n_inst);
#define ISG_DEC(TL, SOFT)                                                                                     \
    decode_kernel<TL, SOFT><<<grid, DEC_THREADS, 0, st>>>(static_cast<const TL*>(logits_dev), h, w, boxes_dev, \
                                                           mask_off_dev, threshold, masks_dev, soft_dev,      \
                                                           area_dev, bbox_dev)
    if (logits_bf16) { if (soft_dev) ISG_DEC(bf16, true); else ISG_DEC(bf16, false); }
    else { if (soft_dev) ISG_DEC(float, true); else ISG_DEC(float, false); }
#undef ISG_DEC
    decode_finalize_kernel<<<cdiv(n_inst, 128), 128, 0, st>>>(n_inst, boxes_dev, area_dev, bbox_dev);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_idmap(const void* logits_dev, int logits_bf16, int h, int w, const int32_t* boxes_dev,
              const int32_t* img_first_dev, int n_img, int canvas_h, int canvas_w, float threshold,
              int32_t* idmap_dev, void* stream) {
    if (n_img == 0) return ISG_OK;
    if (!logits_dev || !boxes_dev || !img_first_dev || !idmap_dev) return fail(ISG_ERR_ARG, "isg_idmap: null argument");
    if (n_img < 0 || n_img > 65535 || canvas_h < 1 || canvas_h > 65535 || canvas_w < 1)
        return fail(ISG_ERR_ARG, "isg_idmap: bad sizes");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    dim3 grid(cdiv(canvas_w, 256), canvas_h, n_img);
    if (logits_bf16)
        idmap_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(logits_dev), h, w, boxes_dev, img_first_dev,
                                                 canvas_h, canvas_w, threshold, idmap_dev);
    else
        idmap_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(logits_dev), h, w, boxes_dev,
                                                  img_first_dev, canvas_h, canvas_w, threshold, idmap_dev);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_mask_iou_counts(const uint8_t* a_dev, const uint8_t* b_dev, int64_t n, unsigned long long* inter_union_dev,
                        void* stream) {
    if (n <= 0) return ISG_OK;
    if (!a_dev || !b_dev || !inter_union_dev) return fail(ISG_ERR_ARG, "isg_mask_iou_counts: null argument");
    int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    iou_counts_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a_dev, b_dev, n, inter_union_dev);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_adam_step(float* param_dev, const float* grad_dev, float* exp_avg_dev, float* exp_avg_sq_dev, int64_t n,
                  float grad_scale, float lr, float beta1, float beta2, float eps, int step, void* stream) {
    if (n <= 0) return ISG_OK;
    if (!param_dev || !grad_dev || !exp_avg_dev || !exp_avg_sq_dev || step < 1)
        return fail(ISG_ERR_ARG, "isg_adam_step: bad argument");
    float bc1 = 1.f - powf(beta1, (float)step), bc2s = sqrtf(1.f - powf(beta2, (float)step));
    adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        param_dev, grad_dev, exp_avg_dev, exp_avg_sq_dev, n, grad_scale, lr, beta1, beta2, eps, bc1, bc2s);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

}  // extern "C"
