// Mask decode: logits -> sigmoid -> bilinear resize to each instance's box -> threshold ->
// per-instance 0/1 masks (+ optional soft bytes), area and bbox by warp/block reductions,
// and the per-image id map.  Memory-bound by design: coalesced 16-byte logit loads, 4-byte
// mask stores, one atomic set per CTA.  Semantics: DESIGN.md section 3; sigmoid =
// model/segment.py:534, soft byte = train_instance.py:398-399.
#include <algorithm>
#include <climits>
#include <cmath>

#include "decode_math.cuh"

namespace isg {
namespace {

constexpr int DEC_THREADS = 256, DEC_VEC = 8, DEC_ITERS = 2, DEC_PPB = DEC_THREADS * DEC_VEC * DEC_ITERS;
constexpr int DEC_MAX_INST = 65536;

// Per-instance accumulators of the launch in flight: {area, enc(xmin), enc(ymin), xmax+1, ymax+1}
// with enc(v) = INT_MAX - v, so that zero is the identity of every field (area: add, others: max)
// and the last CTA of an instance can publish the result and hand the slot back zeroed -- no
// init or finalize launch.  One decode per device may be in flight at a time (stream order).
__device__ int g_dec_acc[DEC_MAX_INST][5];
__device__ unsigned g_dec_ticket[DEC_MAX_INST];
__device__ unsigned g_dt[4];   // tiled kernel: [0] next tile, [1] CTAs that ran dry

// Accumulators of one decode in flight: per-instance {area, xmin, ymin, xmax, ymax} + arrival tickets + the tile counters
// of the tiled kernel.  Default = the arrays above (one decode in flight per device, <= DEC_MAX_INST instances); a caller
// workspace (isg_decode_bind_workspace: [4 x u32 counters][max_inst x u32 tickets][max_inst x 5 x i32], zeroed once, left
// zeroed by every launch) gives every stream / pipeline slot its own.
struct DecWs { unsigned* dt; unsigned* ticket; int (*acc)[5]; };
__device__ __forceinline__ DecWs dec_ws(unsigned* ws, int ws_inst) {
    DecWs r;
    if (ws) { r.dt = ws; r.ticket = ws + 4; r.acc = reinterpret_cast<int(*)[5]>(ws + 4 + ws_inst); }
    else { r.dt = g_dt; r.ticket = g_dec_ticket; r.acc = g_dec_acc; }
    return r;
}

// rarely taken exact re-evaluations stay out of line: the hot loop must fit the instruction cache
__device__ __noinline__ float exact_sigmoid_nl(float x) { return exact_sigmoid(x); }
template <typename TL>
__device__ __noinline__ float exact_prob_nl(const TL* L, int w, int y0, int y1, float ly0, float ly1, int ox, float sx) {
    Tap ty; ty.i0 = y0; ty.i1 = y1; ty.l0 = ly0; ty.l1 = ly1;
    Tap tx = make_tap(ox, w, sx);
    return exact_prob(L, w, ty, tx);
}

// same, from output coordinates (id-map tie-breaks)
template <typename TL>
__device__ __noinline__ float exact_prob_at_nl(const TL* L, int h, int w, int oy, float sy, int ox, float sx) {
    return exact_prob(L, w, make_tap(oy, h, sy), make_tap(ox, w, sx));
}

struct DecAcc { int cnt, xmin, ymin, xmax, ymax; };

// launch with programmatic stream serialization (the kernels start with griddepcontrol.wait)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, int block, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(block); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// pixels [idx, idx+nv) of one instance; IDENT: the mask has the size of the logit map
template <typename TL, bool SOFT, bool IDENT>
__device__ __forceinline__ void decode_run(const TL* __restrict__ L, int h, int w, int wk, int hk, int npix, int idx, int nv,
                                           bool have, const float4& p0, const float4& p1, float thr, float thr_logit,
                                           int use_logit_cmp, float sy, float sx, bool vec_out, uint8_t* dst,
                                           uint8_t* sdst, DecAcc& A) {
    int oy = idx / wk, ox = idx - oy * wk;
    unsigned mb[2] = {0u, 0u}, sb[2] = {0u, 0u};
    const float lv[DEC_VEC] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    if (IDENT && !SOFT && have && use_logit_cmp && (wk & (DEC_VEC - 1)) == 0) {
        // hot path: the 8 pixels share a row; decide on the logit, keep bits, update stats once
        unsigned m8 = 0, u8 = 0;
#pragma unroll
        for (int e = 0; e < DEC_VEC; ++e) {
            const float d = lv[e] - thr_logit;
            m8 |= (d > 0.f ? 1u : 0u) << e;
            u8 |= (fabsf(d) <= 1e-3f ? 1u : 0u) << e;
        }
        while (u8) {   // within 1e-3 of the boundary: the specification's exact sequence decides
            const int e = __ffs(u8) - 1;
            u8 &= u8 - 1;
            const bool m = exact_sigmoid_nl(lv[e]) > thr;
            m8 = (m8 & ~(1u << e)) | ((m ? 1u : 0u) << e);
        }
        if (m8) {
            A.cnt += __popc(m8);
            A.xmin = min(A.xmin, ox + __ffs(m8) - 1); A.xmax = max(A.xmax, ox + 31 - __clz(m8));
            A.ymin = min(A.ymin, oy); A.ymax = max(A.ymax, oy);
        }
        const unsigned lo = ((m8 & 0xfu) * 0x00204081u) & 0x01010101u, hi = ((m8 >> 4) * 0x00204081u) & 0x01010101u;
        if (vec_out) __stcs(reinterpret_cast<uint2*>(dst + idx), make_uint2(lo, hi));
        else {
            for (int e = 0; e < DEC_VEC; ++e) dst[idx + e] = (uint8_t)((m8 >> e) & 1u);
        }
        return;
    }
    Tap ty = make_tap(oy, h, sy);
#pragma unroll
    for (int e = 0; e < DEC_VEC; ++e) {
        if (e < nv) {
            float p = 0.f;
            bool decided = false, m = false;
            if (IDENT) {
                const float x = have ? lv[e] : ldl(L + idx + e);
                if (!SOFT) {
                    // sigmoid is monotone: far from logit(thr) the comparison is made on the logit itself
                    if (use_logit_cmp) {
                        if (fabsf(x - thr_logit) > 1e-3f) { m = x > thr_logit; decided = true; }
                    } else {
                        p = fast_sigmoid(x);
                        if (fabsf(p - thr) > DECIDE_MARGIN) { m = p > thr; decided = true; }
                    }
                }
                if (!decided) { p = exact_sigmoid_nl(x); m = p > thr; }
            } else {
                if (!SOFT) {
                    Tap tx = make_tap(ox, w, sx);
                    p = fast_prob(L, w, ty, tx);
                    if (fabsf(p - thr) > DECIDE_MARGIN) { m = p > thr; decided = true; }
                }
                if (!decided) { p = exact_prob_nl(L, w, ty.i0, ty.i1, ty.l0, ty.l1, ox, sx); m = p > thr; }
            }
            if (m) {
                mb[e >> 2] |= 1u << (8 * (e & 3));
                ++A.cnt;
                A.xmin = min(A.xmin, ox); A.xmax = max(A.xmax, ox);
                A.ymin = min(A.ymin, oy); A.ymax = max(A.ymax, oy);
            }
            if (SOFT) sb[e >> 2] |= soft_byte(p) << (8 * (e & 3));
            if (++ox == wk) { ox = 0; ++oy; if (!IDENT) ty = make_tap(oy, h, sy); }
        }
    }
    if (nv == DEC_VEC && vec_out) {
        __stcs(reinterpret_cast<uint2*>(dst + idx), make_uint2(mb[0], mb[1]));
        if (SOFT) __stcs(reinterpret_cast<uint2*>(sdst + idx), make_uint2(sb[0], sb[1]));
    } else {
        for (int e = 0; e < nv; ++e) {
            dst[idx + e] = (uint8_t)((mb[e >> 2] >> (8 * (e & 3))) & 0xff);
            if (SOFT) sdst[idx + e] = (uint8_t)((sb[e >> 2] >> (8 * (e & 3))) & 0xff);
        }
    }
}

// ---- lean path: identity size, fp32 logits, 0/1 masks only, rows and pointers 8-pixel aligned ----
// exact decision of 8 pixels (specification sequence), out of line: taken only near the boundary
__device__ __noinline__ unsigned exact_m8(float4 a, float4 b, float thr) {
    const float lv[DEC_VEC] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    unsigned m8 = 0;
#pragma unroll
    for (int e = 0; e < DEC_VEC; ++e) m8 |= (exact_sigmoid(lv[e]) > thr ? 1u : 0u) << e;
    return m8;
}

// 8 pixels of one row: decide on the logit (sigmoid is monotone; `margin` in logit units covers the
// rounding of the exact sequence and of thr_logit), keep bits, update the statistics once
__device__ __forceinline__ void lean_item(const float4& p0, const float4& p1, int idx, int oy, int ox, float thr,
                                          float thr_logit, float margin, uint8_t* dst, DecAcc& A) {
    const float lv[DEC_VEC] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
    unsigned m8 = 0;
    float near = 3.0e38f;
#pragma unroll
    for (int e = 0; e < DEC_VEC; ++e) {
        const float d = lv[e] - thr_logit;
        m8 |= (d > 0.f ? 1u : 0u) << e;
        near = fminf(near, fabsf(d));     // NaN logits drop out here and give 0 above, as the exact sequence does
    }
    if (near <= margin) m8 = exact_m8(p0, p1, thr);
    if (m8) {
        A.cnt += __popc(m8);
        A.xmin = min(A.xmin, ox + __ffs(m8) - 1); A.xmax = max(A.xmax, ox + 31 - __clz(m8));
        A.ymin = min(A.ymin, oy); A.ymax = max(A.ymax, oy);
    }
    const unsigned lo = ((m8 & 0xfu) * 0x00204081u) & 0x01010101u, hi = ((m8 >> 4) * 0x00204081u) & 0x01010101u;
    __stcs(reinterpret_cast<uint2*>(dst + idx), make_uint2(lo, hi));
}

__device__ __forceinline__ void decode_lean(const float* __restrict__ L, int wk, int npix, int ngroups, float thr,
                                            float thr_logit, float margin, uint8_t* dst, DecAcc& A) {
    static_assert(DEC_ITERS == 2, "lean path is written for two items per group");
    constexpr int HALF = DEC_THREADS * DEC_VEC;
    int q = blockIdx.x;
    int idx = q * DEC_PPB + threadIdx.x * DEC_VEC;
    // (row, column) of the thread's item follow the two strides of its walk without further divisions
    int oy = idx / wk, ox = idx - oy * wk;
    const int dq1 = HALF / wk, dr1 = HALF - dq1 * wk;
    const int step2 = (int)gridDim.x * DEC_PPB - HALF;
    const int dq2 = step2 / wk, dr2 = step2 - dq2 * wk;
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float4 c0 = z4, c1 = z4, c2 = z4, c3 = z4, n0 = z4, n1 = z4, n2 = z4, n3 = z4;
    if (idx < npix) { c0 = __ldcs(reinterpret_cast<const float4*>(L + idx)); c1 = __ldcs(reinterpret_cast<const float4*>(L + idx) + 1); }
    if (idx + HALF < npix) { c2 = __ldcs(reinterpret_cast<const float4*>(L + idx + HALF)); c3 = __ldcs(reinterpret_cast<const float4*>(L + idx + HALF) + 1); }
    while (true) {
        const int qn = q + gridDim.x;
        const bool more = qn < ngroups;
        const int idn = qn * DEC_PPB + threadIdx.x * DEC_VEC;
        if (more) {   // the next group's loads are in flight while this one is decided
            if (idn < npix) { n0 = __ldcs(reinterpret_cast<const float4*>(L + idn)); n1 = __ldcs(reinterpret_cast<const float4*>(L + idn) + 1); }
            if (idn + HALF < npix) { n2 = __ldcs(reinterpret_cast<const float4*>(L + idn + HALF)); n3 = __ldcs(reinterpret_cast<const float4*>(L + idn + HALF) + 1); }
        }
        if (idx < npix) lean_item(c0, c1, idx, oy, ox, thr, thr_logit, margin, dst, A);
        ox += dr1; oy += dq1;
        if (ox >= wk) { ox -= wk; ++oy; }
        if (idx + HALF < npix) lean_item(c2, c3, idx + HALF, oy, ox, thr, thr_logit, margin, dst, A);
        if (!more) break;
        ox += dr2; oy += dq2;
        if (ox >= wk) { ox -= wk; ++oy; }
        q = qn; idx = idn;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    }
}

// gridDim.x CTAs share one instance's flat hk*wk mask in groups of DEC_PPB consecutive pixels
// (group q = blockIdx.x, blockIdx.x + gridDim.x, ...); one thread-iteration = 8 consecutive pixels:
// two 16-byte logit loads in, one 8-byte mask store out.  The loads of the next group are in flight
// while the current one is decided.
template <typename TL, bool SOFT>
__global__ void __launch_bounds__(DEC_THREADS, 4)
decode_kernel(const TL* __restrict__ logits, int h, int w, const int32_t* __restrict__ boxes,
              const int64_t* __restrict__ mask_off, float thr, float thr_logit, float margin, int use_logit_cmp,
              uint8_t* __restrict__ masks, uint8_t* __restrict__ soft, int32_t* __restrict__ area,
              int32_t* __restrict__ bbox, unsigned* ws, int ws_inst) {
    // launched with programmatic stream serialization: the grid is staged while the previous kernel of the
    // stream drains; nothing is read or written before that kernel has completed
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const DecWs W = dec_ws(ws, ws_inst);
    const int n = blockIdx.y;
    const int4 bx = __ldg(reinterpret_cast<const int4*>(boxes) + n);
    const int wk = bx.z, hk = bx.w;
    const int npix = wk * hk;
    if (npix <= 0 && blockIdx.x == 0 && threadIdx.x == 0) {   // degenerate box: empty result
        area[n] = 0;
        bbox[4 * n] = bbox[4 * n + 1] = bbox[4 * n + 2] = bbox[4 * n + 3] = 0;
    }
    const int ngroups = (npix + DEC_PPB - 1) / DEC_PPB;
    if ((int)blockIdx.x >= ngroups) return;
    const TL* L = logits + (size_t)n * h * w;
    const int64_t moff = mask_off[n];
    uint8_t* dst = masks + moff;
    uint8_t* sdst = SOFT ? soft + moff : nullptr;
    const bool identity = (wk == w && hk == h);
    const bool vec_in = identity && sizeof(TL) == 4 && ((reinterpret_cast<uintptr_t>(L) & 15) == 0);
    const bool vec_out = ((reinterpret_cast<uintptr_t>(dst) & 7) == 0) && (!SOFT || (reinterpret_cast<uintptr_t>(sdst) & 7) == 0);
    const float sy = __fdiv_rn((float)h, (float)hk), sx = __fdiv_rn((float)w, (float)wk);
    DecAcc A = {0, INT_MAX, INT_MAX, -1, -1};
    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!SOFT && sizeof(TL) == 4 && vec_in && vec_out && use_logit_cmp && (wk & (DEC_VEC - 1)) == 0) {
        decode_lean(reinterpret_cast<const float*>(L), wk, npix, ngroups, thr, thr_logit, margin, dst, A);
    } else if (identity) {
        float4 cur[DEC_ITERS][2], nxt[DEC_ITERS][2];
        auto load = [&](float4 (&buf)[DEC_ITERS][2], int q) {
#pragma unroll
            for (int it = 0; it < DEC_ITERS; ++it) {
                const int idx = q * DEC_PPB + (it * DEC_THREADS + threadIdx.x) * DEC_VEC;
                buf[it][0] = z4; buf[it][1] = z4;
                if (vec_in && idx + DEC_VEC <= npix) {
                    buf[it][0] = __ldcs(reinterpret_cast<const float4*>(L + idx));
                    buf[it][1] = __ldcs(reinterpret_cast<const float4*>(L + idx) + 1);
                }
            }
        };
        int q = blockIdx.x;
        load(cur, q);
        while (true) {
            const int qn = q + gridDim.x;
            const bool more = qn < ngroups;
            if (more) load(nxt, qn);
#pragma unroll
            for (int it = 0; it < DEC_ITERS; ++it) {
                const int idx = q * DEC_PPB + (it * DEC_THREADS + threadIdx.x) * DEC_VEC;
                if (idx < npix) {
                    const int nv = min(DEC_VEC, npix - idx);
                    decode_run<TL, SOFT, true>(L, h, w, wk, hk, npix, idx, nv, vec_in && nv == DEC_VEC, cur[it][0], cur[it][1],
                                               thr, thr_logit, use_logit_cmp, sy, sx, vec_out, dst, sdst, A);
                }
            }
            if (!more) break;
            q = qn;
#pragma unroll
            for (int it = 0; it < DEC_ITERS; ++it) { cur[it][0] = nxt[it][0]; cur[it][1] = nxt[it][1]; }
        }
    } else {
        for (int q = blockIdx.x; q < ngroups; q += gridDim.x) {
#pragma unroll 1
            for (int it = 0; it < DEC_ITERS; ++it) {
                const int idx = q * DEC_PPB + (it * DEC_THREADS + threadIdx.x) * DEC_VEC;
                if (idx >= npix) break;
                decode_run<TL, SOFT, false>(L, h, w, wk, hk, npix, idx, min(DEC_VEC, npix - idx), false, z4, z4, thr,
                                            thr_logit, use_logit_cmp, sy, sx, vec_out, dst, sdst, A);
            }
        }
    }
    int cnt = A.cnt, xmin = A.xmin, ymin = A.ymin, xmax = A.xmax, ymax = A.ymax;
    // warp reduction, one set of atomics per CTA, last CTA of the instance publishes
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
        int* acc = W.acc[n];
        if (cnt) {
            atomicAdd(acc, cnt);
            atomicMax(acc + 1, INT_MAX - xmin); atomicMax(acc + 2, INT_MAX - ymin);
            atomicMax(acc + 3, xmax + 1); atomicMax(acc + 4, ymax + 1);
        }
        __threadfence();
        const unsigned nblk = (unsigned)min((int)gridDim.x, ngroups);   // CTAs of this instance that take part
        if (atomicAdd(&W.ticket[n], 1u) == nblk - 1) {
            __threadfence();
            const int a = atomicExch(acc, 0);
            const int e1 = atomicExch(acc + 1, 0), e2 = atomicExch(acc + 2, 0);
            const int e3 = atomicExch(acc + 3, 0), e4 = atomicExch(acc + 4, 0);
            W.ticket[n] = 0;
            area[n] = a;
            if (a == 0) { bbox[4 * n] = bbox[4 * n + 1] = bbox[4 * n + 2] = bbox[4 * n + 3] = 0; }
            else {
                bbox[4 * n] = INT_MAX - e1 + bx.x; bbox[4 * n + 1] = INT_MAX - e2 + bx.y;
                bbox[4 * n + 2] = e3 + bx.x; bbox[4 * n + 3] = e4 + bx.y;
            }
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

// ---- tiled resize decode (any box size, 0/1 masks only) with fused id-map claims ------------------
// Work unit = one output ROW per warp: the vertical tap, the two source-row pointers and the row's
// statistics are warp-uniform, lanes take consecutive output pixels (the four bilinear taps of a warp
// fall into ~scale*128 bytes per source row), no per-pixel division.  A tile = DT_ROWS consecutive
// rows of one instance (two adjacent rows per warp); `tile_first` is the exclusive prefix of the
// instances' tile counts.  Tiles are handed out by an atomic counter (boxes of U(96,640) pixels make
// tiles of very different cost; the next tile's index and instance are fetched while the current one
// is computed), so the resident wave stays busy until the list is empty.
// Decision ladder per pixel: (1) all four logits on one side of logit(thr) by more than `margin`
// -> decided without any sigmoid (a convex mix of probabilities stays on that side); (2) MUFU
// probability further than DECIDE_MARGIN from thr; (3) the exact sequence, out of line.
// Statistics: counts / x-extent come from the ballots, y-extent from the row index; one set of
// atomics per warp and tile, the warp that completes the instance's last row publishes.
// With `claims` (the id-map buffer, zeroed by the caller) every set pixel ORs the instance's bit
// (index within its image, < 32) into the canvas word; idmap_finalize_kernel turns the bit sets into
// ids and compares probabilities only where two or more instances claim the same pixel.
constexpr int DT_THREADS = 256, DT_WARPS = DT_THREADS / 32, DT_RPW = 2, DT_ROWS = DT_WARPS * DT_RPW;


// source indices of the align_corners=False rule; returns the clamped source coordinate s, from which the
// weights follow as make_tap derives them: l1 = s - i0, l0 = 1 - l1
__device__ __forceinline__ float tap_index(int dst, int in, float scale, int& i0, int& i1) {
    float s = __fsub_rn(__fmul_rn(scale, __fadd_rn((float)dst, 0.5f)), 0.5f);
    if (s < 0.0f) s = 0.0f;
    i0 = min((int)s, in - 1);
    i1 = i0 + (i0 < in - 1 ? 1 : 0);
    return s;
}

template <typename TL>
__global__ void __launch_bounds__(DT_THREADS, 4)
decode_tiled_kernel(const TL* __restrict__ logits, int n_inst, int h, int w, const int32_t* __restrict__ boxes,
                    const int64_t* __restrict__ mask_off, const int32_t* __restrict__ tile_first, int n_tiles,
                    float thr, float thr_logit, float margin, uint8_t* __restrict__ masks,
                    int32_t* __restrict__ area, int32_t* __restrict__ bbox,
                    const int32_t* __restrict__ inst_img, const int32_t* __restrict__ img_first,
                    int Hs, int Ws, unsigned* __restrict__ claims, unsigned* ws, int ws_inst) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const DecWs W = dec_ws(ws, ws_inst);
    __shared__ int s_tile[2][2];   // {tile, instance} of the current and the next tile
    asm volatile("griddepcontrol.wait;" ::: "memory");   // see decode_kernel
    auto fetch = [&](int slot) {   // thread 0: next tile index and the instance it belongs to
        const int t = (int)atomicAdd(&W.dt[0], 1u);
        int lo = 0;
        if (t < n_tiles) {
            int hi = n_inst - 1;
            while (lo < hi) {   // last n with tile_first[n] <= t (instances without tiles are skipped: <=)
                const int mid = (lo + hi + 1) >> 1;
                if (__ldg(tile_first + mid) <= t) lo = mid; else hi = mid - 1;
            }
        }
        s_tile[slot][0] = t; s_tile[slot][1] = lo;
    };
    if (blockIdx.x == 0) {   // boxes without pixels publish an empty result (nobody owns a tile of them)
        for (int n = threadIdx.x; n < n_inst; n += DT_THREADS)
            if (tile_first[n + 1] == tile_first[n]) {
                area[n] = 0;
                bbox[4 * n] = bbox[4 * n + 1] = bbox[4 * n + 2] = bbox[4 * n + 3] = 0;
            }
    }
    if (threadIdx.x == 0) fetch(0);
    __syncthreads();
    for (int it = 0;; ++it) {
        const int t = s_tile[it & 1][0], n = s_tile[it & 1][1];
        if (threadIdx.x == 0) fetch((it + 1) & 1);
        if (t >= n_tiles) break;
        const int4 bx = __ldg(reinterpret_cast<const int4*>(boxes) + n);
        const int wk = bx.z, hk = bx.w;
        const TL* L = logits + (size_t)n * h * w;
        uint8_t* dst = masks + mask_off[n];
        const float sy = __fdiv_rn((float)h, (float)hk), sx = __fdiv_rn((float)w, (float)wk);
        unsigned* cl = nullptr;
        unsigned bit = 0;
        if (claims) {
            const int img = __ldg(inst_img + n);
            cl = claims + (size_t)img * Hs * Ws;
            bit = 1u << (n - __ldg(img_first + img));
        }
        const int row0 = (t - __ldg(tile_first + n)) * DT_ROWS + warp * DT_RPW;
        int cnt = 0, xmin = INT_MAX, xmax = -1, ymin = INT_MAX, ymax = -1, rows = 0;
#pragma unroll 1
        for (int oy = row0; oy < min(row0 + DT_RPW, hk); ++oy) {
            ++rows;
            int y0, y1;
            const float ys = tap_index(oy, h, sy, y0, y1);
            const float ly1 = __fsub_rn(ys, (float)y0), ly0 = __fsub_rn(1.0f, ly1);
            const TL* r0 = L + (size_t)y0 * w;
            const TL* r1 = L + (size_t)y1 * w;
            uint8_t* drow = dst + (size_t)oy * wk;
            const bool word_out = (reinterpret_cast<uintptr_t>(drow) & 3) == 0;
            const int cy = bx.y + oy;
            unsigned* crow = (cl && cy >= 0 && cy < Hs) ? cl + (size_t)cy * Ws : nullptr;
            int rcnt = 0;
            for (int x0 = 0; x0 < wk; x0 += 128) {
                float a[4], b[4], c[4], d[4], fx[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int ox = x0 + 32 * r + lane;
                    a[r] = b[r] = c[r] = d[r] = fx[r] = 0.f;
                    if (ox < wk) {
                        int i0, i1;
                        fx[r] = __fsub_rn(tap_index(ox, w, sx, i0, i1), (float)i0);   // lx1
                        a[r] = ldl(r0 + i0); b[r] = ldl(r0 + i1); c[r] = ldl(r1 + i0); d[r] = ldl(r1 + i1);
                    }
                }
                unsigned bal[4];
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int ox = x0 + 32 * r + lane;
                    bool m = false;
                    if (ox < wk) {
                        const float lo4 = fminf(fminf(a[r], b[r]), fminf(c[r], d[r]));
                        const float hi4 = fmaxf(fmaxf(a[r], b[r]), fmaxf(c[r], d[r]));
                        // fminf/fmaxf drop NaN operands: a NaN tap (or inf - inf) shows in the sum and goes down the ladder
                        const float s4 = (a[r] + b[r]) + (c[r] + d[r]);
                        if (s4 == s4 && lo4 - thr_logit > margin) m = true;
                        else if (s4 == s4 && thr_logit - hi4 > margin) m = false;
                        else {
                            const float lx1 = fx[r], lx0 = 1.0f - lx1;
                            const float p = ly0 * (lx0 * fast_sigmoid(a[r]) + lx1 * fast_sigmoid(b[r])) +
                                            ly1 * (lx0 * fast_sigmoid(c[r]) + lx1 * fast_sigmoid(d[r]));
                            if (fabsf(p - thr) > DECIDE_MARGIN) m = p > thr;
                            else m = exact_prob_at_nl(L, h, w, oy, sy, ox, sx) > thr;
                        }
                        if (m && crow) {
                            const int cx = bx.x + ox;
                            if (cx >= 0 && cx < Ws) atomicOr(crow + cx, bit);
                        }
                    }
                    bal[r] = __ballot_sync(0xffffffffu, m);
                    if (bal[r]) {
                        rcnt += __popc(bal[r]);
                        xmin = min(xmin, x0 + 32 * r + __ffs(bal[r]) - 1);
                        xmax = max(xmax, x0 + 32 * r + 31 - __clz(bal[r]));
                    }
                }
                if (word_out && x0 + 128 <= wk) {
                    // lane i owns pixels x0 + 4i .. x0 + 4i + 3: nibble (i & 7) of ballot (i >> 3)
                    const unsigned bsel = lane < 16 ? (lane < 8 ? bal[0] : bal[1]) : (lane < 24 ? bal[2] : bal[3]);
                    const unsigned nib = (bsel >> (4 * (lane & 7))) & 0xfu;
                    __stcs(reinterpret_cast<unsigned*>(drow + x0) + lane, (nib * 0x00204081u) & 0x01010101u);
                } else {
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const int ox = x0 + 32 * r + lane;
                        if (ox < wk) drow[ox] = (uint8_t)((bal[r] >> lane) & 1u);
                    }
                }
            }
            if (rcnt) { cnt += rcnt; ymin = min(ymin, oy); ymax = max(ymax, oy); }
        }
        if (lane == 0 && rows) {
            int* acc = W.acc[n];
            if (cnt) {
                atomicAdd(acc, cnt);
                atomicMax(acc + 1, INT_MAX - xmin); atomicMax(acc + 2, INT_MAX - ymin);
                atomicMax(acc + 3, xmax + 1); atomicMax(acc + 4, ymax + 1);
            }
            __threadfence();
            if (atomicAdd(&W.ticket[n], (unsigned)rows) + (unsigned)rows == (unsigned)hk) {   // all rows are in
                __threadfence();
                const int ar = atomicExch(acc, 0);
                const int e1 = atomicExch(acc + 1, 0), e2 = atomicExch(acc + 2, 0);
                const int e3 = atomicExch(acc + 3, 0), e4 = atomicExch(acc + 4, 0);
                W.ticket[n] = 0;
                area[n] = ar;
                if (ar == 0) { bbox[4 * n] = bbox[4 * n + 1] = bbox[4 * n + 2] = bbox[4 * n + 3] = 0; }
                else {
                    bbox[4 * n] = INT_MAX - e1 + bx.x; bbox[4 * n + 1] = INT_MAX - e2 + bx.y;
                    bbox[4 * n + 2] = e3 + bx.x; bbox[4 * n + 3] = e4 + bx.y;
                }
            }
        }
        __syncthreads();   // the next tile's {index, instance} is in place; this one's slot may be refilled
    }
    if (threadIdx.x == 0) {   // the last CTA to run dry re-arms the tile counter for the next launch
        __threadfence();
        if (atomicAdd(&W.dt[1], 1u) == gridDim.x - 1) { W.dt[0] = 0; W.dt[1] = 0; __threadfence(); }
    }
}

// claims word -> id: one claimant needs no arithmetic; contested pixels compare exact probabilities
// (largest wins, ties -> lowest index), exactly the rule idmap_kernel applies to every pixel
template <typename TL>
__global__ void __launch_bounds__(256)
idmap_finalize_kernel(const TL* __restrict__ logits, int h, int w, const int32_t* __restrict__ boxes,
                      const int32_t* __restrict__ img_first, int Hs, int Ws, int64_t total, int32_t* __restrict__ idmap) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= total) return;
    unsigned bits = (unsigned)idmap[i];
    if (!bits) return;
    const int64_t per = (int64_t)Hs * Ws;
    const int img = (int)(i / per);
    const int rem = (int)(i - img * per);
    const int y = rem / Ws, x = rem - y * Ws;
    const int first = __ldg(img_first + img);
    int best = first + __ffs(bits);
    if (bits & (bits - 1)) {
        // MUFU probabilities (abs error < 2e-6) name the winner unless the two best are closer than 1e-5;
        // only then the exact sequence is evaluated for every claimant
        float p1 = -1.f, p2 = -1.f;
        for (unsigned bb = bits; bb; bb &= bb - 1) {
            const int n = first + __ffs(bb) - 1;
            const int4 bx = __ldg(reinterpret_cast<const int4*>(boxes) + n);
            const TL* L = logits + (size_t)n * h * w;
            const Tap ty = make_tap(y - bx.y, h, __fdiv_rn((float)h, (float)bx.w));
            const Tap tx = make_tap(x - bx.x, w, __fdiv_rn((float)w, (float)bx.z));
            const float p = fast_prob(L, w, ty, tx);
            if (p > p1) { p2 = p1; p1 = p; best = n + 1; }
            else if (p > p2) p2 = p;
        }
        if (!(p1 - p2 > 1e-5f)) {
            float bp = 0.f;
            best = 0;
            for (unsigned bb = bits; bb; bb &= bb - 1) {
                const int n = first + __ffs(bb) - 1;
                const int4 bx = __ldg(reinterpret_cast<const int4*>(boxes) + n);
                const float p = exact_prob_at_nl(logits + (size_t)n * h * w, h, w, y - bx.y, __fdiv_rn((float)h, (float)bx.w),
                                                 x - bx.x, __fdiv_rn((float)w, (float)bx.z));
                if (best == 0 || p > bp) { best = n + 1; bp = p; }
            }
        }
    }
    idmap[i] = best;
}

// a byte counts as set when it is >= its threshold (1: any nonzero byte, numpy's logical_and / logical_or view of the
// reference's 0-255 soft masks, train_instance.py:398-403)
__global__ void iou_counts_kernel(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int64_t n,
                                  unsigned long long* out, int thr_a, int thr_b) {
    unsigned inter = 0, uni = 0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        bool pa = a[i] >= thr_a, pb = b[i] >= thr_b;
        inter += pa && pb; uni += pa || pb;
    }
    inter = __reduce_add_sync(0xffffffffu, inter);
    uni = __reduce_add_sync(0xffffffffu, uni);
    if ((threadIdx.x & 31) == 0) {
        if (inter) atomicAdd(out, (unsigned long long)inter);
        if (uni) atomicAdd(out + 1, (unsigned long long)uni);
    }
}

// 0/1 mask bytes -> bits, little-endian within a byte (bit i of byte j = mask[8 j + i], numpy packbits(bitorder="little")):
// one thread packs 16 mask bytes (one 16-byte load) into 2 output bytes; a warp reads 512 and writes 64 contiguous bytes
__global__ void __launch_bounds__(256) pack_bits_kernel(const uint8_t* __restrict__ m, int64_t n, uint8_t* __restrict__ bits) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;   // group of 16 mask bytes
    const int64_t base = i * 16;
    if (base >= n) return;
    unsigned out = 0;
    if (base + 16 <= n) {
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(m) + i);
        const unsigned w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // bytes are 0/1: gather bit 0 of the four bytes of a word into a nibble
            const unsigned v = w[k] & 0x01010101u;
            out |= ((v | (v >> 7) | (v >> 14) | (v >> 21)) & 0xfu) << (4 * k);
        }
        *reinterpret_cast<uint16_t*>(bits + 2 * i) = (uint16_t)out;
    } else {
        for (int k = 0; base + k < n; ++k) out |= (unsigned)(m[base + k] & 1u) << k;
        bits[2 * i] = (uint8_t)out;
        if (base + 8 < n) bits[2 * i + 1] = (uint8_t)(out >> 8);
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

namespace {
// workspace bound by isg_decode_bind_workspace for the decode calls of THIS host thread (null: the default arrays)
thread_local unsigned* t_ws = nullptr;
thread_local int t_ws_inst = 0;
}  // namespace

extern "C" {

int64_t isg_decode_workspace_bytes(int max_inst) {
    return max_inst < 1 ? 0 : 16 + (int64_t)max_inst * 4 + (int64_t)max_inst * 20;
}

int isg_decode_bind_workspace(void* ws_dev, int max_inst) {
    if (ws_dev && (max_inst < 1 || (reinterpret_cast<uintptr_t>(ws_dev) & 15)))
        return fail(ISG_ERR_ARG, "isg_decode_bind_workspace: workspace must be 16-byte aligned with max_inst >= 1");
    t_ws = static_cast<unsigned*>(ws_dev);
    t_ws_inst = ws_dev ? max_inst : 0;
    return ISG_OK;
}

int isg_decode(const void* logits_dev, int logits_bf16, int n_inst, int h, int w, const int32_t* boxes_dev,
               const int64_t* mask_off_dev, int max_wk, int max_hk, float threshold, uint8_t* masks_dev,
               uint8_t* soft_dev, int32_t* area_dev, int32_t* bbox_dev, void* stream) {
    if (n_inst == 0) return ISG_OK;
    if (!logits_dev || !boxes_dev || !mask_off_dev || !masks_dev || !area_dev || !bbox_dev)
        return fail(ISG_ERR_ARG, "isg_decode: null argument");
    if (n_inst < 0 || n_inst > 65535 || (t_ws && n_inst > t_ws_inst) || h < 1 || w < 1 || max_wk < 1 || max_hk < 1)
        return fail(ISG_ERR_ARG, "isg_decode: bad sizes (n_inst=%d h=%d w=%d max box %dx%d)", n_inst, h, w, max_wk, max_hk);
    if ((int64_t)max_wk * max_hk > INT_MAX) return fail(ISG_ERR_ARG, "isg_decode: box too large");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // identity-size pixels far from the boundary are decided on the logit itself (sigmoid is monotone);
    // only when thr is so close to 0/1 that 1e-3 in logit units is not well separated in probability we
    // keep the probability-space test
    const double t = threshold;
    const int use_logit_cmp = (t > 1e-3 && t < 1.0 - 1e-3) ? 1 : 0;
    const float thr_logit = use_logit_cmp ? (float)log(t / (1.0 - t)) : 0.f;
    // lean path: |logit - thr_logit| beyond this is decided on the logit.  The exact sequence is within
    // 3e-7 of the true sigmoid (Cephes expf, two roundings), i.e. 3e-7 / (t (1 - t)) in logit units, and
    // thr_logit carries one rounding (<= 5e-7): six-fold cover
    const float margin = use_logit_cmp ? (float)(2e-6 / (t * (1.0 - t)) + 2e-6) : 0.f;
    // one resident wave: as many CTAs as fit on the device at once, shared evenly by the instances
    // (more would queue behind the first wave; env ISG_DEC_SLOTS overrides)
    static int slots_tab[4] = {0, 0, 0, 0};
    int& slots = slots_tab[(logits_bf16 ? 2 : 0) + (soft_dev ? 1 : 0)];
    if (!slots) {
        int dev = 0, sms = 148, occ = 2;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (logits_bf16) {
            if (soft_dev) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, decode_kernel<bf16, true>, DEC_THREADS, 0);
            else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, decode_kernel<bf16, false>, DEC_THREADS, 0);
        } else {
            if (soft_dev) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, decode_kernel<float, true>, DEC_THREADS, 0);
            else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, decode_kernel<float, false>, DEC_THREADS, 0);
        }
        slots = getenv("ISG_DEC_SLOTS") ? atoi(getenv("ISG_DEC_SLOTS")) : sms * std::max(occ, 1);
    }
    const int groups = cdiv(max_wk * max_hk, DEC_PPB);
    dim3 grid(std::max(1, std::min(groups, slots / n_inst)), n_inst);
#define ISG_DEC(TL, SOFT)                                                                                        \
    ISG_CUDA(launch_pdl(decode_kernel<TL, SOFT>, grid, DEC_THREADS, st, static_cast<const TL*>(logits_dev), h, w, \
                        boxes_dev, mask_off_dev, threshold, thr_logit, margin, use_logit_cmp, masks_dev, soft_dev, \
                        area_dev, bbox_dev, t_ws, t_ws_inst))
    if (logits_bf16) { if (soft_dev) ISG_DEC(bf16, true); else ISG_DEC(bf16, false); }
    else { if (soft_dev) ISG_DEC(float, true); else ISG_DEC(float, false); }
#undef ISG_DEC
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_decode_tile_rows(void) { return DT_ROWS; }

int isg_decode_tiled(const void* logits_dev, int logits_bf16, int n_inst, int h, int w, const int32_t* boxes_dev,
                     const int64_t* mask_off_dev, const int32_t* tile_first_dev, int n_tiles, float threshold,
                     uint8_t* masks_dev, int32_t* area_dev, int32_t* bbox_dev, const int32_t* inst_img_dev,
                     const int32_t* img_first_dev, int n_img, int canvas_h, int canvas_w, int32_t* idmap_dev,
                     void* stream) {
    if (n_inst == 0) return ISG_OK;
    if (!logits_dev || !boxes_dev || !mask_off_dev || !tile_first_dev || !masks_dev || !area_dev || !bbox_dev)
        return fail(ISG_ERR_ARG, "isg_decode_tiled: null argument");
    if (n_inst < 0 || n_inst > (t_ws ? t_ws_inst : 65535) || h < 1 || w < 1 || n_tiles < 0)
        return fail(ISG_ERR_ARG, "isg_decode_tiled: bad sizes (n_inst=%d h=%d w=%d n_tiles=%d)", n_inst, h, w, n_tiles);
    if (idmap_dev && (!inst_img_dev || !img_first_dev || n_img < 1 || canvas_h < 1 || canvas_w < 1))
        return fail(ISG_ERR_ARG, "isg_decode_tiled: id map requested without image table / canvas size");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const double t = threshold;
    const bool logit_cmp = t > 1e-3 && t < 1.0 - 1e-3;
    const float thr_logit = logit_cmp ? (float)log(t / (1.0 - t)) : 0.f;
    // step 1 of the ladder decides |logit - logit(thr)| > margin for all four taps: 2e-5 in probability,
    // against < 1e-6 of rounding in the exact bilinear sequence; thresholds next to 0/1 skip the step
    const float margin = logit_cmp ? (float)(2e-5 / (t * (1.0 - t)) + 4e-6 * (1.0 + fabs((double)thr_logit))) : INFINITY;
    const int64_t total = idmap_dev ? (int64_t)n_img * canvas_h * canvas_w : 0;
    if (idmap_dev) ISG_CUDA(cudaMemsetAsync(idmap_dev, 0, (size_t)total * 4, st));
    static int slots_tab[2] = {0, 0};
    int& slots = slots_tab[logits_bf16 ? 1 : 0];
    if (!slots) {
        int dev = 0, sms = 148, occ = 2;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (logits_bf16) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, decode_tiled_kernel<bf16>, DT_THREADS, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, decode_tiled_kernel<float>, DT_THREADS, 0);
        slots = getenv("ISG_DEC_SLOTS") ? atoi(getenv("ISG_DEC_SLOTS")) : sms * std::max(occ, 1);
    }
    const int grid = std::max(1, std::min(n_tiles, slots));
    unsigned* claims = reinterpret_cast<unsigned*>(idmap_dev);
    if (logits_bf16)
        ISG_CUDA(launch_pdl(decode_tiled_kernel<bf16>, dim3(grid), DT_THREADS, st, static_cast<const bf16*>(logits_dev), n_inst,
                            h, w, boxes_dev, mask_off_dev, tile_first_dev, n_tiles, threshold, thr_logit, margin, masks_dev,
                            area_dev, bbox_dev, inst_img_dev, img_first_dev, canvas_h, canvas_w, claims, t_ws, t_ws_inst));
    else
        ISG_CUDA(launch_pdl(decode_tiled_kernel<float>, dim3(grid), DT_THREADS, st, static_cast<const float*>(logits_dev), n_inst,
                            h, w, boxes_dev, mask_off_dev, tile_first_dev, n_tiles, threshold, thr_logit, margin, masks_dev,
                            area_dev, bbox_dev, inst_img_dev, img_first_dev, canvas_h, canvas_w, claims, t_ws, t_ws_inst));
    ISG_CUDA(cudaGetLastError());
    if (idmap_dev) {
        const unsigned blocks = (unsigned)((total + 255) / 256);
        if (logits_bf16)
            idmap_finalize_kernel<bf16><<<blocks, 256, 0, st>>>(static_cast<const bf16*>(logits_dev), h, w, boxes_dev,
                                                                 img_first_dev, canvas_h, canvas_w, total, idmap_dev);
        else
            idmap_finalize_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(logits_dev), h, w, boxes_dev,
                                                                  img_first_dev, canvas_h, canvas_w, total, idmap_dev);
        ISG_CUDA(cudaGetLastError());
    }
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

int isg_pack_mask_bits(const uint8_t* masks_dev, int64_t n, uint8_t* bits_dev, void* stream) {
    if (n <= 0) return ISG_OK;
    if (!masks_dev || !bits_dev) return fail(ISG_ERR_ARG, "isg_pack_mask_bits: null argument");
    if ((reinterpret_cast<uintptr_t>(masks_dev) & 15) || (reinterpret_cast<uintptr_t>(bits_dev) & 1))
        return fail(ISG_ERR_ARG, "isg_pack_mask_bits: masks must be 16-byte and bits 2-byte aligned");
    const int64_t groups = (n + 15) / 16;
    pack_bits_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(masks_dev, n, bits_dev);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_mask_iou_counts(const uint8_t* a_dev, const uint8_t* b_dev, int64_t n, unsigned long long* inter_union_dev,
                        void* stream) {
    if (n <= 0) return ISG_OK;
    if (!a_dev || !b_dev || !inter_union_dev) return fail(ISG_ERR_ARG, "isg_mask_iou_counts: null argument");
    int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    iou_counts_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a_dev, b_dev, n, inter_union_dev, 1, 1);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_mask_iou_counts_thr(const uint8_t* a_dev, const uint8_t* b_dev, int64_t n, int thr_a, int thr_b,
                            unsigned long long* inter_union_dev, void* stream) {
    if (n <= 0) return ISG_OK;
    if (!a_dev || !b_dev || !inter_union_dev) return fail(ISG_ERR_ARG, "isg_mask_iou_counts_thr: null argument");
    if (thr_a < 1 || thr_a > 255 || thr_b < 1 || thr_b > 255)
        return fail(ISG_ERR_ARG, "isg_mask_iou_counts_thr: thresholds must be in [1,255], got %d and %d", thr_a, thr_b);
    int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
    iou_counts_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(a_dev, b_dev, n, inter_union_dev, thr_a, thr_b);
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
