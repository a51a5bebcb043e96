// Implicit-GEMM convolution on tcgen05 tensor cores (ISG_MODE_BF16).
//
//   D[128 pixels x N] (fp32, TMEM) += A_tap[128 pixels x 16 ch] (bf16, smem) * W_tap[16 ch x N]
//
// One CTA = one output tile of TH rows x full width of one image.  The input tile (with halo)
// is staged by ONE TMA box per source tensor into [chunk][row][x][8ch] planes; TMA zero-fills
// outside the image, which is the convolution's zero padding.  Every conv tap (dy,dx) is the
// same plane read from a shifted start address (see tc_common.cuh), so 1x1, 2x2-stride-2
// (space-to-depth folded into the tensor map), dense 3x3 and concatenated inputs (torch.cat,
// segment.py:331,485,494: two tensor maps, one K loop) all run through this kernel.  BatchNorm
// scale is folded into the bf16 weights; the epilogue adds the fp32 shift, an optional residual
// (identity, or nearest-x2 upsampled: segment.py:76,322) and the activation, converts to bf16
// and leaves through one TMA store (out-of-image junk columns are clipped by TMA).
#pragma once
#include "tc_common.cuh"

namespace isg {
namespace tc {

constexpr int TC_THREADS = 128;
constexpr int TC_MAX_TAPS = 9;
enum ResMode { RES_NONE = 0, RES_DIRECT = 1, RES_UP2 = 2 };

struct TcParams {
    // input staging
    int n_src;          // 1..3 tensor maps (concatenated along K)
    int planes[3];      // 16-byte chunk planes delivered per source (= K/8 of that source)
    int src_off[3];     // byte offset of each source's first plane inside the A region (128-aligned)
    int a_off[24];      // byte offset of the first plane of K-slice kk (16 channels = 2 planes)
    int Wp, R;          // plane pitch (pixels) and rows per plane: plane = R*Wp pixels
    int hx, hy;         // halo: the box starts at (x = -hx, y = y0*sy - hy)
    int batch_in_y;     // 1: source maps fold the batch into y (coordinate y = b*Hin + ...), 5th coord = 0
    int Hin;            // rows per image of the source maps' y dimension
    // MMA schedule
    int ntaps, tap_off[TC_MAX_TAPS];  // pixel offsets dy*Wp + dx
    int nk16;           // K/16 over all sources
    int TH, n_mt, N;    // output rows per tile, 128-pixel M tiles, MMA N
    // output / epilogue
    int H, W;           // output image size
    int tiles_per_img;
    const float* shift; const float* slope; int act;
    const bf16* res; int res_Cs, res_mode, res_H, res_W;
    const bf16* wpack; int w_bytes;   // pre-packed B operand blocks [tap][k16][2][N][8]
    int stage_pix;      // pixels per staging plane (= TH*Wp)
    int a_bytes;        // bytes reserved for the A region (planes + overrun slack)
};

__global__ void __launch_bounds__(TC_THREADS)
tc_conv_kernel(const __grid_constant__ CUtensorMap map_a0, const __grid_constant__ CUtensorMap map_a1,
               const __grid_constant__ CUtensorMap map_a2, const __grid_constant__ CUtensorMap map_out,
               const TcParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint8_t* s_w = smem;                              // weights
    uint8_t* s_a = s_w + p.w_bytes;                   // A planes (source 0 then source 1)
    uint8_t* s_o = s_a + p.a_bytes;                   // output staging planes
    uint64_t* bars = reinterpret_cast<uint64_t*>(s_o + (size_t)(p.N / 8) * p.stage_pix * 16);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2);

    const int tid = threadIdx.x, warp = tid >> 5;
    const int b = blockIdx.x / p.tiles_per_img, y0 = (blockIdx.x % p.tiles_per_img) * p.TH;
    const uint32_t ncols = tmem_cols_pow2(p.n_mt * p.N);

    if (tid == 0) {
        tma_prefetch_desc(&map_a0);
        if (p.n_src > 1) tma_prefetch_desc(&map_a1);
        if (p.n_src > 2) tma_prefetch_desc(&map_a2);
        tma_prefetch_desc(&map_out);
        mbar_init(&bars[0], 1);   // TMA landed
        mbar_init(&bars[1], 1);   // MMAs complete
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(tmem_slot, ncols);
    // weights: plain 16-byte copies (L2-resident after the first CTAs), then hand over to the async proxy
    for (int i = tid; i < p.w_bytes / 16; i += TC_THREADS)
        reinterpret_cast<uint4*>(s_w)[i] = __ldg(reinterpret_cast<const uint4*>(p.wpack) + i);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const int plane_bytes = p.R * p.Wp * 16;
    if (tid == 0) {
        // ---- TMA producer ----
        uint32_t tx = 0;
        for (int s = 0; s < p.n_src; ++s) tx += (uint32_t)p.planes[s] * plane_bytes;
        mbar_expect_tx(&bars[0], tx);
        const int ycoord = (p.batch_in_y ? b * p.Hin : 0) + y0 - p.hy;
        const int bcoord = p.batch_in_y ? 0 : b;
        tma_load_5d(s_a + p.src_off[0], &map_a0, &bars[0], 0, -p.hx, ycoord, 0, bcoord);
        if (p.n_src > 1) tma_load_5d(s_a + p.src_off[1], &map_a1, &bars[0], 0, -p.hx, ycoord, 0, bcoord);
        if (p.n_src > 2) tma_load_5d(s_a + p.src_off[2], &map_a2, &bars[0], 0, -p.hx, ycoord, 0, bcoord);
        // ---- MMA issuer ----
        mbar_wait(&bars[0], 0);
        tc_fence_after();
        const uint32_t idesc = umma_idesc_bf16(128, p.N);
        const uint32_t a_base = smem_u32(s_a), w_base = smem_u32(s_w);
        const uint32_t b_block = (uint32_t)p.N * 32;   // bytes of one [2][N][8] weight block
        for (int mt = 0; mt < p.n_mt; ++mt) {
            uint32_t acc = 0;
            for (int t = 0; t < p.ntaps; ++t) {
                const uint32_t a_tap = a_base + (uint32_t)(mt * 128 + p.tap_off[t]) * 16;
                for (int kk = 0; kk < p.nk16; ++kk) {
                    uint64_t ad = umma_smem_desc(a_tap + (uint32_t)p.a_off[kk], plane_bytes, 128);
                    uint64_t bd = umma_smem_desc(w_base + (uint32_t)(t * p.nk16 + kk) * b_block, p.N * 16, 128);
                    umma_bf16(tmem_base + mt * p.N, ad, bd, idesc, acc);
                    acc = 1;
                }
            }
        }
        umma_commit(&bars[1]);
    }
    __syncwarp();
    // ---- epilogue: all 4 warps, thread = TMEM lane = pixel row of the M tile ----
    mbar_wait(&bars[1], 0);
    tc_fence_after();
    for (int mt = 0; mt < p.n_mt; ++mt) {
        const int m = mt * 128 + tid;
        const int r = m / p.Wp, c = m - r * p.Wp;
        const int y = y0 + r;
        const bool in_stage = m < p.stage_pix;
        const bool valid = in_stage && c < p.W && y < p.H;
        const bf16* resp = nullptr;
        if (valid && p.res_mode == RES_DIRECT) resp = p.res + (((size_t)b * p.res_H + y) * p.res_W + c) * p.res_Cs;
        if (valid && p.res_mode == RES_UP2)
            resp = p.res + (((size_t)b * p.res_H + (y >> 1)) * p.res_W + (c >> 1)) * p.res_Cs;
        for (int cc = 0; cc < p.N / 16; ++cc) {
            float v[16];
            tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + mt * p.N + cc * 16, v);
            if (!in_stage) continue;
            float rs[16];
            if (resp) {
                uint4 q0 = __ldg(reinterpret_cast<const uint4*>(resp + cc * 16));
                uint4 q1 = __ldg(reinterpret_cast<const uint4*>(resp + cc * 16 + 8));
                const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    rs[2 * i] = __uint_as_float(w[i] << 16);
                    rs[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
                }
            }
            uint32_t packed[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float o[2];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int ch = cc * 16 + 2 * i + e;
                    float x = v[2 * i + e] + __ldg(p.shift + ch);
                    if (resp) x += rs[2 * i + e];
                    o[e] = apply_act(x, p.act, p.act == ACT_PRELU ? __ldg(p.slope + ch) : 0.f);
                }
                __nv_bfloat162 h = __floats2bfloat162_rn(o[0], o[1]);
                packed[i] = *reinterpret_cast<uint32_t*>(&h);
            }
            uint4* d0 = reinterpret_cast<uint4*>(s_o + ((size_t)(2 * cc) * p.stage_pix + m) * 16);
            uint4* d1 = reinterpret_cast<uint4*>(s_o + ((size_t)(2 * cc + 1) * p.stage_pix + m) * 16);
            *d0 = make_uint4(packed[0], packed[1], packed[2], packed[3]);
            *d1 = make_uint4(packed[4], packed[5], packed[6], packed[7]);
        }
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
        tma_store_5d(&map_out, s_o, 0, 0, y0, 0, b);
        tma_store_commit();
        tma_store_wait_all();
    }
    if (warp == 0) {
        tc_fence_after();
        tmem_dealloc(tmem_base, ncols);
    }
}

}  // namespace tc
}  // namespace isg
