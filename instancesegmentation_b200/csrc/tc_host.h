// Host helpers for the tcgen05 path: CUtensorMap construction (driver entry point resolved
// through the runtime, so libisg.so does not link libcuda) and bf16 conversion.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstring>

#include "common.cuh"

namespace isg {
namespace tc {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// 5-D bf16 map.  dims/box: fastest first; strides_bytes: dims 1..4.  Returns 0 or an isg error.
inline int make_map5(CUtensorMap* out, void* base, const uint64_t dims[5], const uint64_t strides_bytes[4],
                     const uint32_t box[5], const char* what) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(ISG_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    cuuint64_t d[5], s[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < 5; ++i) { d[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i < 4; ++i) s[i] = strides_bytes[i];
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, base, d, s, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(ISG_ERR_CUDA,
                    "cuTensorMapEncodeTiled(%s) failed with %d: dims %llu,%llu,%llu,%llu,%llu strides %llu,%llu,%llu,%llu "
                    "box %u,%u,%u,%u,%u", what, (int)r, (unsigned long long)d[0], (unsigned long long)d[1],
                    (unsigned long long)d[2], (unsigned long long)d[3], (unsigned long long)d[4],
                    (unsigned long long)s[0], (unsigned long long)s[1], (unsigned long long)s[2],
                    (unsigned long long)s[3], bx[0], bx[1], bx[2], bx[3], bx[4]);
    return 0;
}

// NHWC bf16 activation [B,H,W,Cs] viewed as (8ch, x, y, chunk, b); box (8, bw, bh, Cs/8, 1)
inline int make_act_map(CUtensorMap* out, void* base, int B, int H, int W, int Cs, int bw, int bh, const char* what) {
    const uint64_t dims[5] = {8, (uint64_t)W, (uint64_t)H, (uint64_t)Cs / 8, (uint64_t)B};
    const uint64_t st[4] = {(uint64_t)Cs * 2, (uint64_t)W * Cs * 2, 16, (uint64_t)H * W * Cs * 2};
    const uint32_t box[5] = {8, (uint32_t)bw, (uint32_t)bh, (uint32_t)Cs / 8, 1};
    return make_map5(out, base, dims, st, box, what);
}

// planar activation [B][Cs/8][H][W][8 ch] as (8-byte elements: 2 per pixel along x, y, chunk, b); box (2 bw, bh, nchunks, 1).
// Same shared-memory image as the 5-D pixel-major box, with rows of bw * 16 contiguous bytes instead of 16-byte elements.
inline int make_act_map_planar(CUtensorMap* out, void* base, int B, int H, int W, int Cs, int bw, int bh, int nchunks,
                               const char* what) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(ISG_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
    if (2 * bw > 256) return fail(ISG_ERR_ARG, "%s: planar box of %d pixels exceeds 256 elements", what, bw);
    cuuint64_t d[4] = {(cuuint64_t)2 * W, (cuuint64_t)H, (cuuint64_t)Cs / 8, (cuuint64_t)B};
    cuuint64_t st[3] = {(cuuint64_t)W * 16, (cuuint64_t)H * W * 16, (cuuint64_t)(Cs / 8) * H * W * 16};
    cuuint32_t bx[4] = {(cuuint32_t)(2 * bw), (cuuint32_t)bh, (cuuint32_t)nchunks, 1}, es[4] = {1, 1, 1, 1};
    CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT64, 4, base, d, st, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(ISG_ERR_CUDA, "cuTensorMapEncodeTiled(%s, planar) failed with %d: %dx%dx%d box %dx%dx%d", what, (int)r, W, H,
                    Cs, bw, bh, nchunks);
    return 0;
}

// same view, but the box covers only `nchunks` 8-channel chunks (channel subsets, e.g. the conv half
// of init_down or the two halves of an up block's low-resolution tensor)
inline int make_act_map_chunks(CUtensorMap* out, void* base, int B, int H, int W, int Cs, int bw, int bh, int nchunks,
                               const char* what, bool planar = false) {
    if (planar) return make_act_map_planar(out, base, B, H, W, Cs, bw, bh, nchunks, what);
    const uint64_t dims[5] = {8, (uint64_t)W, (uint64_t)H, (uint64_t)Cs / 8, (uint64_t)B};
    const uint64_t st[4] = {(uint64_t)Cs * 2, (uint64_t)W * Cs * 2, 16, (uint64_t)H * W * Cs * 2};
    const uint32_t box[5] = {8, (uint32_t)bw, (uint32_t)bh, (uint32_t)nchunks, 1};
    return make_map5(out, base, dims, st, box, what);
}

// 2x2 stride-2 patches of [B,H,W,Cs] (segment.py:121): (8ch, ox, b*Ho+oy, chunk of the pixel PAIR, row a)
inline int make_s2d_map(CUtensorMap* out, void* base, int B, int H, int W, int Cs, int bw, int bh, const char* what) {
    const uint64_t dims[5] = {8, (uint64_t)W / 2, (uint64_t)B * (H / 2), (uint64_t)2 * Cs / 8, 2};
    const uint64_t st[4] = {(uint64_t)2 * Cs * 2, (uint64_t)2 * W * Cs * 2, 16, (uint64_t)W * Cs * 2};
    const uint32_t box[5] = {8, (uint32_t)bw, (uint32_t)bh, (uint32_t)(2 * Cs / 8), 2};
    return make_map5(out, base, dims, st, box, what);
}

inline uint16_t f2bf(float f) {  // round-to-nearest-even, like __float2bfloat16_rn
    uint32_t u;
    std::memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

inline float bf2f(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}

}  // namespace tc
}  // namespace isg
