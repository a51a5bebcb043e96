// fk -- the fused block kernel of ISG_MODE_BF16 (sm_100a: TMA + tcgen05 + TMEM).
//
// One CTA produces a TH x TW tile of "q pixels" of one image through up to two chained
// implicit GEMMs whose 128-row M tiles are pixels:
//
//   loads     TMA boxes of NHWC bf16 activations into [chunk][row][col][8ch] planes (the
//             canonical no-swizzle K-major UMMA layout, see tc_common.cuh); TMA zero fill
//             outside the image is the convolutions' zero padding; the packed weights and
//             epilogue vectors arrive by one bulk copy on the same mbarrier.
//   stencil   (optional) depthwise 3x3 (dilation d) or the composed (5,1)x(1,5) pair
//             (segment.py:64,91-97) + folded BN + PReLU on CUDA cores, from a haloed TMA tile
//             straight into A-operand planes: the depthwise output never exists in HBM.
//   stage 0   D0[128 x N0] (TMEM, fp32) = sum over taps / K slices of A * W0 (tcgen05.mma,
//             one thread issues); conv taps are shifted views of the same planes.
//             epilogue: + shift, + residual tile (TMA-staged, optional nearest-x2 broadcast),
//             per-column PReLU slope (0 = ReLU, 1 = identity), bf16 -> staging planes.
//   stage 1   (optional) D1 = staging0 * W1: the staging planes of stage 0 ARE the A operand
//             of the next 1x1 convolution (next block's first conv, or the block-diagonal 1x1
//             after a transposed conv), then the same epilogue.
//   output    staging planes leave through TMA stores; or direct stores for the layouts TMA
//             cannot express in 5 dims: space-to-depth (stem), pixel shuffle x2 (transposed
//             conv k4 s2, segment.py:305) and pixel shuffle x4 fp32 (the logits).
#pragma once
#include "tc_common.cuh"

namespace isg {
namespace fk {

constexpr int THREADS = 256;
constexpr int MAX_LOADS = 5;
constexpr int MAX_KS = 20;      // K <= 320 per tap
constexpr int MAX_MMA_TAPS = 9;
constexpr int MAX_ST_TAPS = 25;

enum OutMode { OUT_STAGE = 0, OUT_S2D = 1, OUT_SHUF_BF16 = 2, OUT_SHUF4_F32 = 3 };

struct Load {
    int map;                  // index of the tensor map
    int smem_off;             // destination (128-byte aligned)
    int dx, dy;               // box origin relative to the tile origin
    int chunk0;               // first 8-channel chunk
    int folded;               // 1: map folds the batch into y (2x2 stride-2 patches), 5th coordinate 0
    int rows_per_img;         // folded maps: rows of one image
    int step;                 // tile origin multiplier (1; loads of a x2-upsampled consumer would use 1/2)
};

struct Stage {
    int N, ntaps, nks;
    int tap_off[MAX_MMA_TAPS];     // pixel offsets dy*pitch + dx
    int ks_off[MAX_KS];            // smem offset of the first plane of K slice (16 channels = 2 planes)
    int ks_lbo[MAX_KS];            // plane stride of that slice
    int w_off;                     // packed B: [tap][ks][2][N][8] bf16
    int vec_off;                   // fp32 shift[N], slope[N]
    int pitch;                     // pitch of the M index space (TW + 2*halo of the tap planes)
    int res_on, res_off, res_plane_bytes, res_mod;
    int out_mode;
    int stage_off, stage_plane_bytes, stage_compact;
    int store;                     // 1: TMA-store the staging planes through the stage's out map
    int store_chunk0;
    void* gptr; int Cs_out, shuf, Cp;   // direct-store modes
};

struct Stencil {
    int on;
    int in_off, in_pitch, in_plane_bytes, halo;
    int nchunks, ntaps;
    int tap_off[MAX_ST_TAPS];      // (dy*in_pitch + dx) pixel offsets relative to the centre
    int w_off;                     // fp32 [tap][C]
    int vec_off;                   // fp32 shift[C], slope[C], then bw[5][C] when sep5
    int sep5;
    int out_off, out_plane_bytes;
};

struct Params {
    int TH, TW, H, W, tiles_x, tiles_y;
    int n_mt, tmem_cols;
    int nloads; Load loads[MAX_LOADS];
    uint32_t tx_bytes;             // all TMA boxes + the constant blob
    const void* blob; int blob_off, blob_bytes;
    Stencil st;
    int nstages; Stage stage[2];
    int bar_off;                   // 3 mbarriers + tmem slot
};

__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(tc::smem_u32(dst)), "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(tc::smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void st_global_16(void* p, uint4 v) {
    asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void unpack8(uint4 q, float* f) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        f[2 * i] = __uint_as_float(w[i] << 16);
        f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
}

// ---- depthwise stencil: haloed input planes -> A operand planes ----
__device__ __forceinline__ void run_stencil(uint8_t* smem, const Params& p, int x0) {
    const Stencil& s = p.st;
    const int npix = p.TH * p.TW, C = s.nchunks * 8;
    const float* wt = reinterpret_cast<const float*>(smem + s.w_off);
    const float* vec = reinterpret_cast<const float*>(smem + s.vec_off);
    for (int i = threadIdx.x; i < s.nchunks * npix; i += THREADS) {
        const int k = i / npix, pix = i - k * npix;
        const int r = pix / p.TW, c = pix - r * p.TW;
        const uint8_t* in = smem + s.in_off + (size_t)k * s.in_plane_bytes +
                            (size_t)((r + s.halo) * s.in_pitch + c + s.halo) * 16;
        float acc[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] = 0.f;
        for (int t = 0; t < s.ntaps; ++t) {
            float v[8];
            unpack8(*reinterpret_cast<const uint4*>(in + s.tap_off[t] * 16), v);
            const float4 w0 = *reinterpret_cast<const float4*>(wt + t * C + k * 8);
            const float4 w1 = *reinterpret_cast<const float4*>(wt + t * C + k * 8 + 4);
            acc[0] = fmaf(v[0], w0.x, acc[0]); acc[1] = fmaf(v[1], w0.y, acc[1]);
            acc[2] = fmaf(v[2], w0.z, acc[2]); acc[3] = fmaf(v[3], w0.w, acc[3]);
            acc[4] = fmaf(v[4], w1.x, acc[4]); acc[5] = fmaf(v[5], w1.y, acc[5]);
            acc[6] = fmaf(v[6], w1.z, acc[6]); acc[7] = fmaf(v[7], w1.w, acc[7]);
        }
        float sh[8], sl[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) { sh[e] = vec[k * 8 + e]; sl[e] = vec[C + k * 8 + e]; }
        if (s.sep5) {
            // bias of the bare (5,1) conv seen through the (1,5) conv: only taps inside the image
            const int xa = x0 + c;
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                const int xx = xa + kx - 2;
                if (xx >= 0 && xx < p.W) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) sh[e] += vec[(2 + kx) * C + k * 8 + e];
                }
            }
        }
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            float a = acc[2 * e] + sh[2 * e], b = acc[2 * e + 1] + sh[2 * e + 1];
            a = a >= 0.f ? a : a * sl[2 * e];
            b = b >= 0.f ? b : b * sl[2 * e + 1];
            o[e] = pack_bf16(a, b);
        }
        *reinterpret_cast<uint4*>(smem + s.out_off + (size_t)k * s.out_plane_bytes + (size_t)pix * 16) =
            make_uint4(o[0], o[1], o[2], o[3]);
    }
}

__device__ __forceinline__ void issue_mma(const Params& p, const Stage& s, uint32_t smem_base, uint32_t tmem_base) {
    const uint32_t idesc = tc::umma_idesc_bf16(128, s.N);
    const uint32_t b_block = (uint32_t)s.N * 32;
    for (int mt = 0; mt < p.n_mt; ++mt) {
        uint32_t acc = 0;
        for (int t = 0; t < s.ntaps; ++t) {
            const uint32_t a_shift = (uint32_t)(mt * 128 + s.tap_off[t]) * 16;
            for (int ks = 0; ks < s.nks; ++ks) {
                const uint64_t ad = tc::umma_smem_desc(smem_base + s.ks_off[ks] + a_shift, s.ks_lbo[ks], 128);
                const uint64_t bd = tc::umma_smem_desc(smem_base + s.w_off + (uint32_t)(t * s.nks + ks) * b_block,
                                                       s.N * 16, 128);
                tc::umma_bf16(tmem_base + mt * s.N, ad, bd, idesc, acc);
                acc = 1;
            }
        }
    }
}

__device__ __forceinline__ void run_epilogue(uint8_t* smem, const Params& p, const Stage& s, uint32_t tmem_base,
                                             int b, int x0, int y0) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3, grp = warp >> 2;
    const int ncc = s.N >> 4;
    const float* vec = reinterpret_cast<const float*>(smem + s.vec_off);
    for (int item = grp; item < p.n_mt * ncc; item += 2) {
        const int mt = item / ncc, cc = item - mt * ncc;
        const int m = mt * 128 + quad * 32 + lane;
        const int r = m / s.pitch, c = m - r * s.pitch;
        const bool geom = r < p.TH && c < p.TW;
        const int y = y0 + r, x = x0 + c;
        const bool inimg = geom && y < p.H && x < p.W;
        const int pc = r * p.TW + c;
        float v[16];
        __syncwarp();
        tc::tmem_ld16(tmem_base + ((uint32_t)(quad * 32) << 16) + mt * s.N + cc * 16, v);
        const int col0 = cc * 16;
        if (s.res_on && geom) {
            const int ch0 = s.res_mod ? col0 % s.res_mod : col0;
            const uint8_t* rp = smem + s.res_off + (size_t)(ch0 >> 3) * s.res_plane_bytes + (size_t)pc * 16;
            float f[16];
            unpack8(*reinterpret_cast<const uint4*>(rp), f);
            unpack8(*reinterpret_cast<const uint4*>(rp + s.res_plane_bytes), f + 8);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += f[i];
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 sh = *reinterpret_cast<const float4*>(vec + col0 + 4 * j);
            const float4 sl = *reinterpret_cast<const float4*>(vec + s.N + col0 + 4 * j);
            float a;
            a = v[4 * j + 0] + sh.x; v[4 * j + 0] = a >= 0.f ? a : a * sl.x;
            a = v[4 * j + 1] + sh.y; v[4 * j + 1] = a >= 0.f ? a : a * sl.y;
            a = v[4 * j + 2] + sh.z; v[4 * j + 2] = a >= 0.f ? a : a * sl.z;
            a = v[4 * j + 3] + sh.w; v[4 * j + 3] = a >= 0.f ? a : a * sl.w;
        }
        if (s.out_mode == OUT_SHUF4_F32) {
            if (inimg) {
                float* g = static_cast<float*>(s.gptr) + ((size_t)b * 4 * p.H + 4 * y) * (4 * (size_t)p.W) + 4 * x;
#pragma unroll
                for (int ry = 0; ry < 4; ++ry)
                    *reinterpret_cast<float4*>(g + (size_t)ry * 4 * p.W) =
                        make_float4(v[4 * ry], v[4 * ry + 1], v[4 * ry + 2], v[4 * ry + 3]);
            }
            continue;
        }
        uint4 q0 = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
        uint4 q1 = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]),
                              pack_bf16(v[14], v[15]));
        if (s.out_mode == OUT_STAGE) {
            if (s.stage_compact && !geom) continue;
            const int si = s.stage_compact ? pc : m;
            uint8_t* d = smem + s.stage_off + (size_t)(col0 >> 3) * s.stage_plane_bytes + (size_t)si * 16;
            *reinterpret_cast<uint4*>(d) = q0;
            *reinterpret_cast<uint4*>(d + s.stage_plane_bytes) = q1;
        } else if (s.out_mode == OUT_S2D) {
            if (inimg) {
                bf16* g = static_cast<bf16*>(s.gptr) +
                          (((size_t)b * (p.H >> 1) + (y >> 1)) * (p.W >> 1) + (x >> 1)) * (4 * (size_t)s.N) +
                          ((y & 1) * 2 + (x & 1)) * s.N + col0;
                st_global_16(g, q0);
                st_global_16(g + 8, q1);
            }
        } else {  // OUT_SHUF_BF16
            if (inimg) {
                const int phase = col0 / s.Cp, ch = col0 - phase * s.Cp;
                const int py = phase / s.shuf, px = phase - py * s.shuf;
                bf16* g = static_cast<bf16*>(s.gptr) +
                          (((size_t)b * p.H * s.shuf + (size_t)y * s.shuf + py) * ((size_t)p.W * s.shuf) +
                           (size_t)x * s.shuf + px) * s.Cs_out + ch;
                st_global_16(g, q0);
                st_global_16(g + 8, q1);
            }
        }
    }
}

__global__ void __launch_bounds__(THREADS, 1)
fk_kernel(const __grid_constant__ CUtensorMap m0, const __grid_constant__ CUtensorMap m1,
          const __grid_constant__ CUtensorMap m2, const __grid_constant__ CUtensorMap m3,
          const __grid_constant__ CUtensorMap mo0, const __grid_constant__ CUtensorMap mo1,
          const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);
    const int tid = threadIdx.x, warp = tid >> 5;
    int bid = blockIdx.x;
    const int tx = bid % p.tiles_x; bid /= p.tiles_x;
    const int ty = bid % p.tiles_y;
    const int b = bid / p.tiles_y;
    const int x0 = tx * p.TW, y0 = ty * p.TH;

    if (tid == 0) {
        tc::mbar_init(&bars[0], 1);
        tc::mbar_init(&bars[1], 1);
        tc::mbar_init(&bars[2], 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, p.tmem_cols);
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (tid == 0) {
        tc::mbar_expect_tx(&bars[0], p.tx_bytes);
        const CUtensorMap* maps[4] = {&m0, &m1, &m2, &m3};
        for (int i = 0; i < p.nloads; ++i) {
            const Load& l = p.loads[i];
            const int cx = x0 * l.step + l.dx;
            const int cy = y0 * l.step + l.dy + (l.folded ? b * l.rows_per_img : 0);
            tc::tma_load_5d(smem + l.smem_off, maps[l.map], &bars[0], 0, cx, cy, l.chunk0, l.folded ? 0 : b);
        }
        bulk_g2s(smem + p.blob_off, p.blob, p.blob_bytes, &bars[0]);
    }
    tc::mbar_wait(&bars[0], 0);

    if (p.st.on) {
        run_stencil(smem, p, x0);
        tc::fence_proxy_async();
        __syncthreads();
    }
    const uint32_t smem_base = tc::smem_u32(smem);
    if (tid == 0) {
        tc::tc_fence_after();
        issue_mma(p, p.stage[0], smem_base, tmem_base);
        tc::umma_commit(&bars[1]);
    }
    __syncwarp();
    tc::mbar_wait(&bars[1], 0);
    tc::tc_fence_after();
    run_epilogue(smem, p, p.stage[0], tmem_base, b, x0, y0);
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    if (tid == 0 && p.stage[0].out_mode == OUT_STAGE && p.stage[0].store) {
        tc::tma_store_5d(&mo0, smem + p.stage[0].stage_off, 0, x0, y0, p.stage[0].store_chunk0, b);
        tc::tma_store_commit();
    }
    if (p.nstages > 1) {
        if (tid == 0) {
            tc::tc_fence_after();
            issue_mma(p, p.stage[1], smem_base, tmem_base);
            tc::umma_commit(&bars[2]);
        }
        __syncwarp();
        tc::mbar_wait(&bars[2], 0);
        tc::tc_fence_after();
        run_epilogue(smem, p, p.stage[1], tmem_base, b, x0, y0);
        tc::fence_proxy_async();
        tc::tc_fence_before();
        __syncthreads();
        if (tid == 0 && p.stage[1].out_mode == OUT_STAGE && p.stage[1].store) {
            tc::tma_store_5d(&mo1, smem + p.stage[1].stage_off, 0, x0, y0, p.stage[1].store_chunk0, b);
            tc::tma_store_commit();
        }
    }
    if (tid == 0) tc::tma_store_wait_all();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, p.tmem_cols);
    }
}

static_assert(sizeof(Params) + 6 * sizeof(CUtensorMap) <= 4000, "kernel parameter space");

}  // namespace fk
}  // namespace isg
