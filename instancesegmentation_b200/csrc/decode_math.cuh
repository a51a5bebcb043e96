// Device-side arithmetic of the mask decode.
//
// "exact_*": every fp32 operation is a single IEEE round-to-nearest add/sub/mul/div issued
// through __f*_rn intrinsics (never contracted to FMA), in the operation order fixed by the
// decode specification, so results are bit-identical to any IEEE host evaluation of the same
// sequence.  "fast_*": MUFU-based approximations (abs error < 2e-6 on a probability) used
// only to classify pixels that are further than DECIDE_MARGIN from a decision boundary;
// everything closer is re-evaluated with the exact sequence, so the emitted masks / soft
// values / ids equal the all-exact result.
#pragma once
#include "common.cuh"

namespace isg {

constexpr float DECIDE_MARGIN = 1e-4f;

__device__ __forceinline__ float exact_exp_nonpos(float y) {
    if (y < -87.0f) return 0.0f;
    float t = __fmul_rn(y, 1.44269504088896341f);
    float n = floorf(__fadd_rn(t, 0.5f));
    float r = __fsub_rn(y, __fmul_rn(n, 0.693359375f));
    r = __fsub_rn(r, __fmul_rn(n, -2.12194440e-4f));
    float p = 1.9875691500e-4f;
    p = __fadd_rn(__fmul_rn(p, r), 1.3981999507e-3f);
    p = __fadd_rn(__fmul_rn(p, r), 8.3334519073e-3f);
    p = __fadd_rn(__fmul_rn(p, r), 4.1665795894e-2f);
    p = __fadd_rn(__fmul_rn(p, r), 1.6666665459e-1f);
    p = __fadd_rn(__fmul_rn(p, r), 5.0000001201e-1f);
    float e = __fmul_rn(p, __fmul_rn(r, r));
    e = __fadd_rn(e, r);
    e = __fadd_rn(e, 1.0f);
    float s = __int_as_float(((int)n + 127) << 23);
    return __fmul_rn(e, s);
}

__device__ __forceinline__ float exact_sigmoid(float x) {
    if (x != x) return x;
    float e = exact_exp_nonpos(-fabsf(x));
    float d = __fadd_rn(1.0f, e);
    return x >= 0.0f ? __fdiv_rn(1.0f, d) : __fdiv_rn(e, d);
}

__device__ __forceinline__ float fast_sigmoid(float x) {
    return __fdividef(1.0f, 1.0f + __expf(-x));
}

struct Tap { int i0, i1; float l0, l1; };

// align_corners=False source index rule; `scale` = exact fp32 quotient in/out
__device__ __forceinline__ Tap make_tap(int dst, int in, float scale) {
    Tap t;
    float s = __fsub_rn(__fmul_rn(scale, __fadd_rn((float)dst, 0.5f)), 0.5f);
    if (s < 0.0f) s = 0.0f;
    t.i0 = (int)s;
    if (t.i0 > in - 1) t.i0 = in - 1;
    t.i1 = t.i0 + (t.i0 < in - 1 ? 1 : 0);
    t.l1 = __fsub_rn(s, (float)t.i0);
    t.l0 = __fsub_rn(1.0f, t.l1);
    return t;
}

__device__ __forceinline__ float exact_lerp2(float a, float b, float c, float d, const Tap& ty, const Tap& tx) {
    float top = __fadd_rn(__fmul_rn(tx.l0, a), __fmul_rn(tx.l1, b));
    float bot = __fadd_rn(__fmul_rn(tx.l0, c), __fmul_rn(tx.l1, d));
    return __fadd_rn(__fmul_rn(ty.l0, top), __fmul_rn(ty.l1, bot));
}

__device__ __forceinline__ float ldl(const float* p) { return __ldg(p); }
__device__ __forceinline__ float ldl(const bf16* p) {
    return __uint_as_float((uint32_t)__ldg(reinterpret_cast<const unsigned short*>(p)) << 16);
}

// exact probability of map L (h x w) at output (oy, ox)
template <typename TL>
__device__ __forceinline__ float exact_prob(const TL* L, int w, const Tap& ty, const Tap& tx) {
    float a = exact_sigmoid(ldl(L + (size_t)ty.i0 * w + tx.i0));
    float b = exact_sigmoid(ldl(L + (size_t)ty.i0 * w + tx.i1));
    float c = exact_sigmoid(ldl(L + (size_t)ty.i1 * w + tx.i0));
    float d = exact_sigmoid(ldl(L + (size_t)ty.i1 * w + tx.i1));
    return exact_lerp2(a, b, c, d, ty, tx);
}

template <typename TL>
__device__ __forceinline__ float fast_prob(const TL* L, int w, const Tap& ty, const Tap& tx) {
    float a = fast_sigmoid(ldl(L + (size_t)ty.i0 * w + tx.i0));
    float b = fast_sigmoid(ldl(L + (size_t)ty.i0 * w + tx.i1));
    float c = fast_sigmoid(ldl(L + (size_t)ty.i1 * w + tx.i0));
    float d = fast_sigmoid(ldl(L + (size_t)ty.i1 * w + tx.i1));
    return ty.l0 * (tx.l0 * a + tx.l1 * b) + ty.l1 * (tx.l0 * c + tx.l1 * d);
}

// train_instance.py:398-399 soft mask byte: trunc(p*255), NaN -> 0
__device__ __forceinline__ unsigned soft_byte(float p) {
    float v = __fmul_rn(p, 255.0f);
    if (!(v >= 0.0f)) return 0u;
    if (v > 255.0f) v = 255.0f;
    return (unsigned)(int)v;
}

}  // namespace isg
