// ISG_MODE_BF16 plan: Segment.forward (model/segment.py:466-508) lowered to fused block
// kernels (fk.cuh).  Every residual block is one or two launches; BatchNorm is folded into
// bf16 weights + fp32 shift, activations into the epilogue, torch.cat into concatenated K
// slices, depthwise convs into the consumer's prologue, the next block's first 1x1 conv
// into the producer's second GEMM, nearest-x2 + 1x1 into a low-resolution GEMM whose result
// is the pixel-shuffled consumer's residual, and bottle6_1 o bottle6_2 into one composed
// transposed convolution.
#include <algorithm>
#include <cstdlib>
#include <memory>

#include "engine.h"
#include "fk.cuh"
#include "mega.cuh"
#include "mega3.cuh"
#include "kernels_misc.cuh"
#include "pack_util.h"
#include "tc_host.h"

namespace isg {
namespace {

// ---------------------------------------------------------------------------------
// CUDA-core helpers of the fused path
// ---------------------------------------------------------------------------------
// network input NCHW fp32 -> X2 [B,H/2,W/2,Cs2] bf16, channel (ry*2+rx)*Cin + c (space-to-depth
// of the stride-2 stem conv, segment.py:23-24), plus max_pool2d(x,4,4) (segment.py:29) into
// channels [16,16+Cin) of init_down.  One thread = one 4x4 input patch.
// channels [0,8) of pixel (y, x) of image b in the space-to-depth input x2 ([B][H2][W2][Cs2] or planar
// [B][Cs2/8][H2][W2][8]); `half` = element distance to channels [8,16)
__device__ __forceinline__ bf16* x2_pixel(bf16* x2, int b, int y, int x, int H2, int W2, int Cs2, int planar, int& half) {
    if (planar) { half = H2 * W2 * 8; return x2 + ((size_t)b * (Cs2 >> 3) * H2 * W2 + (size_t)y * W2 + x) * 8; }
    half = 8;
    return x2 + (((size_t)b * H2 + y) * W2 + x) * Cs2;
}

// 16 channels of one pixel: two 8-channel groups `half` elements apart; split: the lo halves follow 2 groups later
__device__ __forceinline__ void store16(bf16* d, int half, const float* f, int split) {
    if (split) {
        uint32_t h[8], l[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) fk::split_bf16(f[2 * i], f[2 * i + 1], h[i], l[i]);
        fk::st_global_16(d, make_uint4(h[0], h[1], h[2], h[3]));
        fk::st_global_16(d + half, make_uint4(h[4], h[5], h[6], h[7]));
        fk::st_global_16(d + 2 * half, make_uint4(l[0], l[1], l[2], l[3]));
        fk::st_global_16(d + 3 * half, make_uint4(l[4], l[5], l[6], l[7]));
    } else {
        fk::st_global_16(d, make_uint4(fk::pack_bf16(f[0], f[1]), fk::pack_bf16(f[2], f[3]),
                                       fk::pack_bf16(f[4], f[5]), fk::pack_bf16(f[6], f[7])));
        fk::st_global_16(d + half, make_uint4(fk::pack_bf16(f[8], f[9]), fk::pack_bf16(f[10], f[11]),
                                           fk::pack_bf16(f[12], f[13]), fk::pack_bf16(f[14], f[15])));
    }
}
// max-pooled input channels (<= 8) of one init_down pixel at storage channel 16: 8 values + 8 zeros; split: lo at + lo_off
__device__ __forceinline__ void store_pooled(bf16* d, const float* m, int split, int lo_off) {
    if (split) {
        uint32_t h[4], l[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) fk::split_bf16(m[2 * i], m[2 * i + 1], h[i], l[i]);
        fk::st_global_16(d, make_uint4(h[0], h[1], h[2], h[3]));
        fk::st_global_16(d + 8, make_uint4(0, 0, 0, 0));
        fk::st_global_16(d + lo_off, make_uint4(l[0], l[1], l[2], l[3]));
        fk::st_global_16(d + lo_off + 8, make_uint4(0, 0, 0, 0));
    } else {
        fk::st_global_16(d, make_uint4(fk::pack_bf16(m[0], m[1]), fk::pack_bf16(m[2], m[3]), fk::pack_bf16(m[4], m[5]),
                                       fk::pack_bf16(m[6], m[7])));
        fk::st_global_16(d + 8, make_uint4(0, 0, 0, 0));
    }
}

template <int CIN>
__global__ void __launch_bounds__(128)
prep_kernel(const float* __restrict__ x, int H, int W, bf16* __restrict__ x2, int Cs2, int planar2,
            bf16* __restrict__ init, int Cs_init, int total, int split, int init_lo) {
    // launched with programmatic stream serialization: x2 / init_down may still be read by the
    // previous forward's kernels until the wait returns
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int W4 = W >> 2, H4 = H >> 2;
    const int x4 = i % W4, y4 = (i / W4) % H4, b = i / (W4 * H4);
    float v[CIN][4][4];
#pragma unroll
    for (int c = 0; c < CIN; ++c)
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const float4 q = __ldg(reinterpret_cast<const float4*>(
                x + (((size_t)b * CIN + c) * H + 4 * y4 + r) * W + 4 * x4));
            v[c][r][0] = q.x; v[c][r][1] = q.y; v[c][r][2] = q.z; v[c][r][3] = q.w;
        }
    // four space-to-depth pixels (sy, sx) = (2*y4 + a, 2*x4 + bb)
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
            float f[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) f[k] = 0.f;
#pragma unroll
            for (int ry = 0; ry < 2; ++ry)
#pragma unroll
                for (int rx = 0; rx < 2; ++rx)
#pragma unroll
                    for (int c = 0; c < CIN; ++c) f[(ry * 2 + rx) * CIN + c] = v[c][2 * a + ry][2 * bb + rx];
            int half;
            bf16* d = x2_pixel(x2, b, 2 * y4 + a, 2 * x4 + bb, H >> 1, W >> 1, Cs2, planar2, half);
            store16(d, half, f, split);
        }
    float m[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) m[k] = 0.f;
#pragma unroll
    for (int c = 0; c < CIN; ++c) {
        float mx = v[c][0][0];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int q = 0; q < 4; ++q) mx = fmaxf(mx, v[c][r][q]);
        m[c] = mx;
    }
    store_pooled(init + (((size_t)b * H4 + y4) * W4 + x4) * Cs_init + 16, m, split, init_lo);
}

// the same from uint8 NHWC [B,H,W,3] (isg_forward_u8): ToTensor + Normalize of the reference's data
// path (train_instance.py:80-85) in fp32 with IEEE division, then exactly prep_kernel<3>.
// One thread = one 4x4 patch = four 12-byte rows.
struct NormU8 { float mean[3], std[3]; int exact; };   // exact: the space-to-depth planes get v - 255*mean[c] as single bf16 values
__global__ void __launch_bounds__(128)
prep_u8_kernel(const uint8_t* __restrict__ x, int H, int W, bf16* __restrict__ x2, int Cs2, int planar2,
               bf16* __restrict__ init, int Cs_init, int total, NormU8 nm, int split, int init_lo) {
    // a byte has 256 values: the IEEE sequence is evaluated once per (channel, value) and CTA
    __shared__ float lut[3][256];
    for (int e = threadIdx.x; e < 768; e += blockDim.x) {
        const int c = e >> 8;
        lut[c][e & 255] = __fdiv_rn(__fdiv_rn((float)(e & 255), 255.f) - nm.mean[c], nm.std[c]);
    }
    const float ctr[3] = {255.f * nm.mean[0], 255.f * nm.mean[1], 255.f * nm.mean[2]};
    __syncthreads();
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int W4 = W >> 2, H4 = H >> 2;
    const int x4 = i % W4, y4 = (i / W4) % H4, b = i / (W4 * H4);
    float v[3][4][4];
    uint32_t wsrc[4][3];   // the 4 x 12 source bytes (exact path: the planes get v - 255*mean[c])
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(x + (((size_t)b * H + 4 * y4 + r) * W + 4 * x4) * 3);
        const uint32_t w[3] = {__ldg(src), __ldg(src + 1), __ldg(src + 2)};
        wsrc[r][0] = w[0]; wsrc[r][1] = w[1]; wsrc[r][2] = w[2];
#pragma unroll
        for (int k = 0; k < 12; ++k) v[k % 3][r][k / 3] = lut[k % 3][(w[k >> 2] >> (8 * (k & 3))) & 0xffu];
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int bb = 0; bb < 2; ++bb) {
            float f[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) f[k] = 0.f;
#pragma unroll
            for (int ry = 0; ry < 2; ++ry)
#pragma unroll
                for (int rx = 0; rx < 2; ++rx)
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const int row = 2 * a + ry, k = (2 * bb + rx) * 3 + c;
                        f[(ry * 2 + rx) * 3 + c] = nm.exact ? (float)((wsrc[row][k >> 2] >> (8 * (k & 3))) & 0xffu) - ctr[c]
                                                            : v[c][row][2 * bb + rx];
                    }
            int half;
            bf16* d = x2_pixel(x2, b, 2 * y4 + a, 2 * x4 + bb, H >> 1, W >> 1, Cs2, planar2, half);
            store16(d, half, f, nm.exact ? 0 : split);   // exact: bf16-exact values, hi planes only
        }
    float m[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        float mx = v[c][0][0];
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int q = 0; q < 4; ++q) mx = fmaxf(mx, v[c][r][q]);
        m[c] = mx;
    }
    store_pooled(init + (((size_t)b * H4 + y4) * W4 + x4) * Cs_init + 16, m, split, init_lo);
}

// generic (any Cin) variant: one thread = one (patch, channel)
__global__ void __launch_bounds__(128)
prep_generic_kernel(const float* __restrict__ x, int Cin, int H, int W, bf16* __restrict__ x2, int Cs2, int planar2,
                    bf16* __restrict__ init, int Cs_init, int total, int split, int x2_lo, int init_lo) {
    // Cs2 / Cs_init: storage channels per pixel (hi + lo when split); x2_lo / init_lo: channels from hi to lo
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int W4 = W >> 2, H4 = H >> 2;
    const int x4 = i % W4, y4 = (i / W4) % H4, c = (i / (W4 * H4)) % Cin, b = i / (W4 * H4 * Cin);
    float mx = -INFINITY;
    for (int r = 0; r < 4; ++r) {
        const float4 q = __ldg(reinterpret_cast<const float4*>(x + (((size_t)b * Cin + c) * H + 4 * y4 + r) * W + 4 * x4));
        const float f[4] = {q.x, q.y, q.z, q.w};
        for (int k = 0; k < 4; ++k) {
            mx = fmaxf(mx, f[k]);
            const int sy = 2 * y4 + (r >> 1), sx = 2 * x4 + (k >> 1);
            const int sc = ((r & 1) * 2 + (k & 1)) * Cin + c, H2 = H >> 1, W2 = W >> 1;
            const size_t e = planar2 ? (((size_t)b * (Cs2 >> 3) + (sc >> 3)) * H2 * W2 + (size_t)sy * W2 + sx) * 8 + (sc & 7)
                                     : (((size_t)b * H2 + sy) * W2 + sx) * Cs2 + sc;
            const bf16 hi = __float2bfloat16_rn(f[k]);
            x2[e] = hi;
            if (split) {
                const int sl = sc + x2_lo;
                const size_t el = planar2 ? (((size_t)b * (Cs2 >> 3) + (sl >> 3)) * H2 * W2 + (size_t)sy * W2 + sx) * 8 + (sl & 7)
                                          : (((size_t)b * H2 + sy) * W2 + sx) * Cs2 + sl;
                x2[el] = __float2bfloat16_rn(f[k] - __bfloat162float(hi));
            }
        }
    }
    bf16* dp = init + (((size_t)b * H4 + y4) * W4 + x4) * Cs_init + 16 + c;
    const bf16 hi = __float2bfloat16_rn(mx);
    *dp = hi;
    if (split) dp[init_lo] = __float2bfloat16_rn(mx - __bfloat162float(hi));
}

// The composed tail GEMM treats the transposed conv's output as if it extended one pixel
// beyond the image; the reference zero-pads it there (Conv2d padding=1, segment.py:437).
// This kernel subtracts, for the border ring of the logits, the contributions of those
// outside positions.  One thread = one border pixel.
struct TailFixParams {
    const bf16* in; int Hq, Wq, Cs, planar;    // [B,Hq,Wq,16] or planar [B,2,Hq,Wq,8]; Cs = storage channels per pixel
    int lo;                            // split precision: 1 = add the lo half (16 channels after the hi half)
    float* logits;                     // [B,4Hq,4Wq]
    const float* g;                    // [ky(8)][kx(8)][d(9)][ci(16)]: transposed-conv tap (ky,kx) seen through 3x3 tap d
    const float* gb;                   // [d(9)]: the transposed conv's bias seen through 3x3 tap d
    int total;                         // B * ring
};
__device__ __forceinline__ int floordiv4(int a) { return a >= 0 ? a >> 2 : -((3 - a) >> 2); }
__global__ void __launch_bounds__(128) tail_fix_kernel(TailFixParams p) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.total) return;
    const int H = 4 * p.Hq, W = 4 * p.Wq, ring = 2 * W + 2 * (H - 2);
    const int b = i / ring;
    int j = i - b * ring, oy, ox;
    if (j < W) { oy = 0; ox = j; }
    else if (j < 2 * W) { oy = H - 1; ox = j - W; }
    else { j -= 2 * W; oy = 1 + (j >> 1); ox = (j & 1) ? W - 1 : 0; }
    float corr = 0.f;
#pragma unroll
    for (int d = 0; d < 9; ++d) {
        const int ty = oy + d / 3 - 1, tx = ox + d % 3 - 1;
        if (ty >= 0 && ty < H && tx >= 0 && tx < W) continue;
        corr += __ldg(p.gb + d);
        // transposed conv k8 s4 p2 (segment.py:436): output t gets input i through tap k = t + 2 - 4 i
        const int iy1 = floordiv4(ty + 2), ix1 = floordiv4(tx + 2);
#pragma unroll
        for (int a = 0; a < 2; ++a) {
            const int iy = iy1 - 1 + a, ky = ty + 2 - 4 * iy;
            if (iy < 0 || iy >= p.Hq) continue;
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int ix = ix1 - 1 + e, kx = tx + 2 - 4 * ix;
                if (ix < 0 || ix >= p.Wq) continue;
                const uint4* src = reinterpret_cast<const uint4*>(
                    p.planar ? p.in + ((size_t)b * (p.Cs >> 3) * p.Hq * p.Wq + (size_t)iy * p.Wq + ix) * 8
                             : p.in + (((size_t)b * p.Hq + iy) * p.Wq + ix) * p.Cs);
                const int half = p.planar ? p.Hq * p.Wq : 1;   // uint4 units between channels [0,8) and [8,16)
                const float4* g = reinterpret_cast<const float4*>(p.g + ((size_t)(ky * 8 + kx) * 9 + d) * 16);
                float f[16];
                fk::unpack8(__ldg(src), f);
                fk::unpack8(__ldg(src + half), f + 8);
                if (p.lo) {
                    float l[16];
                    fk::unpack8(__ldg(src + 2 * half), l);
                    fk::unpack8(__ldg(src + 3 * half), l + 8);
#pragma unroll
                    for (int q = 0; q < 16; ++q) f[q] += l[q];
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float4 w = __ldg(g + q);
                    corr = fmaf(f[4 * q], w.x, corr); corr = fmaf(f[4 * q + 1], w.y, corr);
                    corr = fmaf(f[4 * q + 2], w.z, corr); corr = fmaf(f[4 * q + 3], w.w, corr);
                }
            }
        }
    }
    p.logits[((size_t)b * H + oy) * W + ox] -= corr;
}

// ---------------------------------------------------------------------------------
// specification of one fused launch
// ---------------------------------------------------------------------------------
struct LoadSpec { int tensor; int chunk0 = 0; int nchunks = -1; int halo = 0; bool s2d = false;
                  bool hi_only = false; };   // split plans: only the hi planes are loaded (the operand has no lo part)
enum PartKind { PART_LOAD = 0, PART_STENCIL = 1, PART_PREV = 2 };
struct Part { int kind; int load; };
typedef std::function<void(const HostWeights&, float* B, float* shift, float* slope)> StageFill;
typedef std::function<void(const HostWeights&, float* w, float* shift, float* slope, float* bw)> StencilFill;
struct StageSpec {
    std::vector<Part> parts;
    bool taps9 = false;
    int N = 0;
    StageFill fill;                 // B[tap][K][N] (zeroed), shift[N] (zeroed), slope[N] (ones)
    std::function<bool(int tap, int ks)> used;   // optional: false = that 16-row K slice of the tap is structurally zero
    int res_load = -1, res_mod = 0;
    int out_mode = fk::OUT_STAGE, out_tensor = -1, out_chunk0 = 0;
    bool store = true;
    int shuf = 1, Cp = 0;
    bool to_logits = false;
    // split precision: MMA passes of this stage (bit 0 A_hi.W_hi, bit 1 A_lo.W_hi, bit 2 A_hi.W_lo); -1 = all three
    int passes = -1;
};
struct StencilSpec { int kind = 0; int load = -1; int dil = 1; StencilFill fill; };  // 1 dw3x3, 2 sep5
struct Spec {
    std::string name;
    int qH = 0, qW = 0;             // q-pixel grid of one image
    std::vector<LoadSpec> loads;
    StencilSpec st;
    std::vector<StageSpec> stages;
    int pool_tensor = -1;           // single-stage launch over a folded s2d load: also emit max_pool2d(x, 2, 2) here
};

struct Lay {
    int TH = 0, TW = 0, pitch = 0, n_mt = 0, cols = 0;
    int ctas = 1, nlbuf = 2;   // CTAs per SM (1: 512 threads, 2: 256 threads), load buffers of the generic loop
    int pipe = 0;              // > 0: warp-specialised launch (fk_pipe_kernel) with this many load slots
    int groups = 3;            // ... and this many epilogue warp groups (the other warps issue MMAs)
    size_t smem = 0;
    int load_off[fk::MAX_LOADS], load_plane[fk::MAX_LOADS], load_planes[fk::MAX_LOADS];   // planes = hi planes (logical chunks)
    int load_lo[fk::MAX_LOADS];   // split precision: bytes from a hi plane of the load to its lo plane (0 = plain)
    int st_out_lo = 0, stage_lo[2] = {0, 0}, pool_lo = 0, tab_n = 0;
    int st_out_off = 0, st_out_plane = 0, st_in_pitch = 0;
    int lbuf_stride = 0, stage_in_load[2] = {0, 0};
    int pool_off = 0, pool_plane = 0;
    int stage_off[2] = {0, 0}, stage_plane[2] = {0, 0}, stage_compact[2] = {1, 1};
    int K[2] = {0, 0};
    int b_off[2] = {0, 0}, vec_off[2] = {0, 0}, stw_off = 0, stvec_off = 0, blob_off = 0, blob_bytes = 0, bar_off = 0;
    uint32_t tx = 0;
    double score = -1;
};

inline int r128(int v) { return (v + 127) / 128 * 128; }

struct FkMaps { CUtensorMap in[4], out[2]; };

// debug (ISG_FK_PROF=1): run one launch with per-CTA phase stamps and print their means
// plain kernels: launched with programmatic stream serialization (they start with griddepcontrol.wait)
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), int grid, int block, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// programmatic dependent launch: the grid may be scheduled (prologue + constants) while the previous
// kernel of the stream drains; fk_kernel orders its activation traffic with griddepcontrol.wait
inline cudaError_t launch_fk(int ctas, int grid, size_t smem, const FkMaps& m, const fk::Params& p, cudaStream_t stream) {
    if (p.pipe_slots) ctas = 1;
    static const bool pdl = getenv("ISG_FK_NO_PDL") == nullptr;
    cudaLaunchConfig_t cfg;
    std::memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(ctas == 4 ? 128 : ctas == 2 ? 256 : 512);
    cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    if (p.split) {   // split-precision instantiations (ISG_MODE_BF16X3)
        if (p.pipe_slots)
            return cudaLaunchKernelEx(&cfg, fk::fk_pipe_kernel<true>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
        if (ctas == 4)
            return cudaLaunchKernelEx(&cfg, fk::fk_kernel<128, true>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
        if (ctas == 2)
            return cudaLaunchKernelEx(&cfg, fk::fk_kernel<256, true>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
        return cudaLaunchKernelEx(&cfg, fk::fk_kernel<512, true>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
    }
    if (p.pipe_slots)
        return cudaLaunchKernelEx(&cfg, fk::fk_pipe_kernel<false>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
    if (ctas == 4)
        return cudaLaunchKernelEx(&cfg, fk::fk_kernel<128>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
    if (ctas == 2)
        return cudaLaunchKernelEx(&cfg, fk::fk_kernel<256>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
    return cudaLaunchKernelEx(&cfg, fk::fk_kernel<512>, m.in[0], m.in[1], m.in[2], m.in[3], m.out[0], m.out[1], p);
}

inline cudaError_t launch_profiled(const std::string& name, fk::Params p, int ctas, int grid, size_t smem, const FkMaps& m,
                                   cudaStream_t stream) {
    long long* d = nullptr;
    cudaMalloc(&d, (size_t)grid * 16 * sizeof(long long));
    cudaMemset(d, 0, (size_t)grid * 16 * sizeof(long long));
    p.prof = d;
    cudaStreamSynchronize(stream);
    launch_fk(ctas, grid, smem, m, p, stream);
    cudaError_t e = cudaStreamSynchronize(stream);
    std::vector<long long> h((size_t)grid * 16);
    cudaMemcpy(h.data(), d, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
    cudaFree(d);
    double ph[9] = {0};
    long long t0 = h[14], t1 = h[15];
    double life = 0;
    for (int b = 0; b < grid; ++b) {
        const long long* r = &h[(size_t)b * 16];
        long long prev = r[0];
        for (int k = 1; k <= 8; ++k) {
            if (r[k] == 0) continue;
            ph[k] += (double)(r[k] - prev);
            prev = r[k];
        }
        t0 = std::min(t0, r[14]); t1 = std::max(t1, r[15]);
        life += (double)(r[15] - r[14]);
    }
    fprintf(stderr, "prof %-40s grid %5d span %7.1f us  cta life %6.2f us | cycles: alloc %6.0f load %6.0f stencil %6.0f "
            "mma0 %6.0f epi0 %6.0f mma1 %6.0f epi1 %6.0f store %6.0f\n", name.c_str(), grid, (t1 - t0) * 1e-3,
            life / grid * 1e-3, ph[1] / grid, ph[2] / grid, ph[3] / grid, ph[4] / grid, ph[5] / grid, ph[6] / grid,
            ph[7] / grid, ph[8] / grid);
    double ss[5] = {0};
    for (int b = 0; b < grid; ++b)
        for (int k = 0; k < 5; ++k) ss[k] += (double)h[(size_t)b * 16 + 9 + k];
    const double nt = std::max(1, p.total_tiles - grid);
    fprintf(stderr, "     steady state per tile (%d tiles, %.1f per CTA): wait+stencil %6.0f mma0 %6.0f epi0 %6.0f mma1 %6.0f "
            "epi1 %6.0f | sum %6.0f cycles\n", p.total_tiles, (double)p.total_tiles / grid, ss[0] / nt, ss[1] / nt,
            ss[2] / nt, ss[3] / nt, ss[4] / nt, (ss[0] + ss[1] + ss[2] + ss[3] + ss[4]) / nt);
    return e;
}

struct FkBuilder {
    isg_plan* P;
    explicit FkBuilder(isg_plan* p) : P(p) {}

    // Global layout: planar [B][Cs/8][H][W][8] unless the tensor is read through a 2x2 stride-2 (space-to-depth)
    // map, whose pixel-pair chunks only exist pixel-major: the stem output and the output of the bottle1 run.
    // ISG_FK_NO_PLANAR=1 keeps every tensor pixel-major (A/B runs).
    static bool planar_for(const std::string& name) {
        static const bool off = getenv("ISG_FK_NO_PLANAR") != nullptr;
        return !off && name != "init_down" && name != "bottle1_5";
    }
    int tensor(const std::string& name, int C, int H, int W, bool tap = false) {
        TensorD t{name, C, round_up(C, 16), H, W, P->ws_bytes, tap};
        t.planar = planar_for(name);
        t.split = P->split;
        size_t bytes = (size_t)P->max_batch * H * W * t.Ct() * 2;
        P->ws_bytes += (bytes + 255) / 256 * 256;
        P->tensors.push_back(t);
        if (tap) P->taps.push_back((int)P->tensors.size() - 1);
        return (int)P->tensors.size() - 1;
    }
    int view(const std::string& name, int of, int C, int Cs, int H, int W, bool tap, int layout) {
        TensorD t{name, C, Cs, H, W, P->tensors[of].off, tap};
        t.layout = layout;
        t.planar = P->tensors[of].planar;
        t.split = P->tensors[of].split;
        P->tensors.push_back(t);
        if (tap) P->taps.push_back((int)P->tensors.size() - 1);
        return (int)P->tensors.size() - 1;
    }
    const TensorD& td(int id) const { return P->tensors[id]; }
    size_t arena(size_t nfloats) {
        P->arena_floats = (P->arena_floats + 63) / 64 * 64;  // 256-byte aligned blobs
        size_t o = P->arena_floats;
        P->arena_floats += nfloats;
        return o;
    }
    void macs(int64_t cout, int64_t cin_g, int kh, int kw, int Ho, int Wo) {
        P->macs_per_image += cout * cin_g * kh * kw * Ho * Wo;
    }

    // ---- layout of one candidate tile ----
    // single stage, no stencil / pooled output / residual: eligible for the warp-specialised kernel
    static bool pipe_ok(const Spec& s) {
        return s.stages.size() == 1 && !s.st.kind && s.pool_tensor < 0 && s.stages[0].res_load < 0 && !getenv("ISG_FK_NO_PIPE");
    }
    // MMA passes of a stage: 0 = plain bf16 plan; split plans default to all three
    int passes_of(const StageSpec& st) const { return !P->split ? 0 : (st.passes < 0 ? 7 : st.passes); }
    static int npasses(int passes) { return passes ? (passes & 1) + ((passes >> 1) & 1) + ((passes >> 2) & 1) : 1; }
    // narrow three-pass stages run pass A_hi as one MMA over [W_hi | W_lo] (fk::Stage::nmerge); ISG_X3_NO_NMERGE=1 disables
    bool nmerge_of(const StageSpec& st) const {
        static const bool off = getenv("ISG_X3_NO_NMERGE") != nullptr;
        return !off && P->split && (passes_of(st) == 7 || passes_of(st) == 5) && st.N <= 32;
    }

    bool layout(const Spec& s, int TH, int TW, int ctas, int nlbuf, Lay& L, int pipe = 0) const {
        const int hx = s.stages[0].taps9 ? 1 : 0;
        const int mult = P->split ? 2 : 1;   // hi + lo plane groups
        L.TH = TH; L.TW = TW; L.pitch = TW + 2 * hx; L.ctas = ctas; L.nlbuf = nlbuf; L.pipe = pipe;
        if (pipe) nlbuf = pipe;
        const int mrows = TH * L.pitch;
        L.n_mt = cdiv(mrows, 128);
        int maxN = 0;
        for (auto& st : s.stages) maxN = std::max(maxN, nmerge_of(st) ? 2 * st.N : st.N);
        L.cols = L.n_mt * maxN;
        if (pipe && 2 * L.cols > 512) return false;   // two accumulators
        if (L.cols > 512 || (int)tc::tmem_cols_pow2(L.cols) > 512 / ctas) return false;
        const int a_rows = L.n_mt * 128 + (hx ? 2 * L.pitch + 2 : 0);
        const int npix = TH * TW;
        const bool chained = s.stages.size() > 1;
        int off = 0;
        L.tx = 0;
        for (size_t i = 0; i < s.loads.size(); ++i) {
            const LoadSpec& l = s.loads[i];
            const TensorD& t = td(l.tensor);
            int bw, bh, planes;
            if (l.s2d) { bw = TW; bh = TH; planes = 4 * t.Cs / 8; }
            else { bw = TW + 2 * l.halo; bh = TH + 2 * l.halo; planes = l.nchunks < 0 ? t.Cs / 8 : l.nchunks; }
            if (bw > 256 || bh > 256 || planes * (l.s2d ? mult : 1) > 256) return false;
            L.load_off[i] = off; L.load_plane[i] = bw * bh * 16; L.load_planes[i] = planes;
            // split: [hi planes | lo planes]; a folded 2x2 load interleaves them per sub-pixel: [hi Cs/8 | lo Cs/8] x 4
            // (the lo box is a TMA destination of its own: 128-byte aligned)
            L.load_lo[i] = P->split ? (l.s2d ? (t.Cs / 8) * L.load_plane[i] : r128(planes * L.load_plane[i])) : 0;
            const int load_bytes = (P->split && !l.s2d) ? L.load_lo[i] + planes * L.load_plane[i] : planes * mult * L.load_plane[i];
            bool is_a0 = false;
            for (auto& pt : s.stages[0].parts) is_a0 |= pt.kind == PART_LOAD && pt.load == (int)i;
            if (is_a0 && hx && l.halo != 1) return false;
            if (is_a0 && !hx && l.halo != 0) return false;
            int slack = 0;
            if (is_a0) slack = std::max(0, a_rows * 16 - L.load_plane[i]);
            if ((int)i == s.stages[0].res_load && chained) slack = std::max(slack, L.n_mt * 128 * 16 - L.load_plane[i]);
            L.tx += (uint32_t)(planes * (l.hi_only ? 1 : mult)) * L.load_plane[i];
            off = r128(off + load_bytes + slack);
        }
        // persistent CTAs double-buffer the whole load area: tile i+1 streams in while tile i computes
        L.lbuf_stride = off;
        off *= nlbuf;
        if (s.st.kind) {
            const LoadSpec& l = s.loads[s.st.load];
            const int nch = l.nchunks < 0 ? td(l.tensor).Cs / 8 : l.nchunks;
            L.st_in_pitch = TW + 2 * l.halo;
            L.st_out_off = off; L.st_out_plane = npix * 16;
            L.st_out_lo = P->split ? r128(nch * L.st_out_plane) : 0;
            off = r128(off + L.st_out_lo + nch * L.st_out_plane + std::max(0, L.n_mt * 128 * 16 - L.st_out_plane));
        }
        for (size_t k = 0; k < s.stages.size(); ++k) {
            const StageSpec& st = s.stages[k];
            int K = 0;
            for (auto& pt : st.parts) {
                if (pt.kind == PART_LOAD) K += L.load_planes[pt.load] * 8;
                else if (pt.kind == PART_STENCIL) K += (s.loads[s.st.load].nchunks < 0 ? td(s.loads[s.st.load].tensor).Cs / 8
                                                                                         : s.loads[s.st.load].nchunks) * 8;
                else K += s.stages[0].N;
            }
            L.K[k] = K;
            if (K % 16 || K / 16 > fk::MAX_KS) return false;
            if (st.out_mode != fk::OUT_STAGE) continue;
            const bool is_chain_src = k == 0 && chained;
            L.stage_compact[k] = (hx == 0) || !is_chain_src;
            if (!L.stage_compact[k] && st.store) return false;
            const int pix = L.stage_compact[k] ? npix : L.n_mt * 128;
            L.stage_plane[k] = pix * 16;
            L.stage_lo[k] = P->split ? r128((st.N / 8) * L.stage_plane[k]) : 0;   // TMA store source of its own
            const bool alias = st.res_load >= 0 && st.res_mod == 0 && L.stage_compact[k] &&
                               L.load_planes[st.res_load] == st.N / 8 && s.loads[st.res_load].halo == 0;
            L.stage_in_load[k] = alias ? 1 : 0;
            if (alias) { L.stage_off[k] = L.load_off[st.res_load]; continue; }
            L.stage_off[k] = off;
            int slack = is_chain_src ? std::max(0, L.n_mt * 128 * 16 - L.stage_plane[k]) : 0;
            off = r128(off + L.stage_lo[k] + (st.N / 8) * L.stage_plane[k] + slack);
        }
        if (s.pool_tensor >= 0) {
            L.pool_off = off; L.pool_plane = npix * 16;
            L.pool_lo = P->split ? r128((td(s.pool_tensor).Cs / 8) * L.pool_plane) : 0;
            off = r128(off + L.pool_lo + (td(s.pool_tensor).Cs / 8) * L.pool_plane);
        }
        // constants blob
        L.blob_off = off;
        int rel = 0;
        for (size_t k = 0; k < s.stages.size(); ++k) {
            L.b_off[k] = rel;
            // W_hi blocks, then W_lo blocks when the stage runs the A_hi.W_lo pass
            rel = r128(rel + (s.stages[k].taps9 ? 9 : 1) * (L.K[k] / 16) * s.stages[k].N * 32 * ((passes_of(s.stages[k]) & 4) ? 2 : 1));
        }
        for (size_t k = 0; k < s.stages.size(); ++k) { L.vec_off[k] = rel; rel = r128(rel + 2 * s.stages[k].N * 4); }
        if (s.st.kind) {
            const LoadSpec& l = s.loads[s.st.load];
            const int C = (l.nchunks < 0 ? td(l.tensor).Cs / 8 : l.nchunks) * 8;
            const int ntaps = s.st.kind == 2 ? 25 : 9;
            L.stw_off = rel; rel = r128(rel + ntaps * C * 4);
            L.stvec_off = rel; rel = r128(rel + (s.st.kind == 2 ? 7 : 2) * C * 4);
        }
        L.blob_bytes = rel;
        off = r128(off + rel);
        L.bar_off = off;
        L.tab_n = 0;
        if (P->split) {   // (A, B) descriptor pairs per table: MMAs per M tile over all passes
            for (size_t k = 0; k < s.stages.size(); ++k) {
                const StageSpec& st = s.stages[k];
                int n = 0;
                for (int t = 0; t < (st.taps9 ? 9 : 1); ++t)
                    for (int ks = 0; ks < L.K[k] / 16; ++ks) n += (!st.used || st.used(t, ks)) ? 1 : 0;
                L.tab_n = std::max(L.tab_n, n * (nmerge_of(st) ? ((passes_of(st) & 2) ? 2 : 1) : npasses(passes_of(st))));
            }
            if (L.tab_n > fk::TAB_N_SPLIT) return false;
            L.smem = (size_t)off + 128 + 2 * 3 * (size_t)L.tab_n * 16;
        } else
        L.smem = (size_t)off + 128 + 2 * 3 * 40 * 8;   // barriers + A-descriptor tables
        return L.smem <= (ctas == 4 ? 55 * 1024 : ctas == 2 ? 112 * 1024 : 224 * 1024);
    }

    // candidate layouts, best heuristic score first.  The plan times them on the device
    // (autotune) unless ISG_FK_NO_TUNE is set; ISG_FK_CTAS / ISG_FK_LBUF pin the mode.
    std::vector<Lay> candidates(const Spec& s) const {
        const int hx = s.stages[0].taps9 ? 1 : 0;
        int max_halo = hx;
        for (auto& l : s.loads) max_halo = std::max(max_halo, l.halo);
        // tile width: the image width split evenly so that every TMA box fits 256 columns
        int nx = 1;
        while (cdiv(s.qW, nx) + 2 * max_halo > 128) ++nx;   // a planar box row holds at most 256 8-byte elements = 128 pixels
        const int TW0 = cdiv(s.qW, nx);
        const char* cap_env = getenv("ISG_FK_MAX_MT");
        const int mt_cap = cap_env ? atoi(cap_env) : 16;
        const bool tune = !getenv("ISG_FK_NO_TUNE");
        const int env_ctas = getenv("ISG_FK_CTAS") ? atoi(getenv("ISG_FK_CTAS")) : 0;
        const int env_lbuf = getenv("ISG_FK_LBUF") ? atoi(getenv("ISG_FK_LBUF")) : 0;
        const int per_mode = getenv("ISG_FK_TUNE_TOP") ? atoi(getenv("ISG_FK_TUNE_TOP")) : 3;
        std::vector<Lay> out;
        for (int mode = 0; mode < 10; ++mode) {
            // modes 0-3: generic kernel (mode 0 = one double-buffered CTA per SM); 4-7: warp-specialised,
            // 2 / 3 load slots x 3 / 2 epilogue warp groups; 8-9: generic kernel, four 128-thread CTAs per SM
            const bool quad = mode >= 8;
            const int pipe = (mode >= 4 && !quad) ? 2 + (mode & 1) : 0;
            const int groups = mode >= 6 ? 2 : 3;
            if (quad && getenv("ISG_FK_NO_QUAD")) continue;
            if (pipe && (!pipe_ok(s) || env_ctas || env_lbuf)) continue;
            if (!pipe && getenv("ISG_FK_FORCE_PIPE") && pipe_ok(s)) continue;   // tests: every eligible launch warp-specialised
            const int c = quad ? 4 : pipe ? 1 : (mode & 1 ? 2 : 1), nb = quad ? (mode & 1 ? 1 : 2) : pipe ? 2 : (mode & 2 ? 1 : 2);
            if (!tune && mode && !(env_ctas || env_lbuf)) break;
            if (env_ctas && c != env_ctas) continue;
            if (env_lbuf && nb != env_lbuf) continue;
            const int slots = 148 * c;
            std::vector<Lay> found;
            // half-width tiles trade halo columns for more rows per tile (better M-tile fill when few rows fit)
            for (int split = 1; split <= (tune && !getenv("ISG_FK_NO_SPLIT") ? 2 : 1); ++split)
            for (int TH = 1; TH <= s.qH && TH + 2 * max_halo <= 256; ++TH) {
                const int TW = cdiv(s.qW, nx * split);
                if (split > 1 && TW < 32) continue;
                Lay L;
                if (!layout(s, TH, TW, c, nb, L, pipe)) continue;
                L.groups = groups;
                if (L.n_mt > mt_cap && TH > 1) continue;
                const int tiles_y = cdiv(s.qH, TH);
                const double eff_m = (double)(TH * TW) / (L.n_mt * 128);
                const double eff_tail = (double)s.qH / (tiles_y * TH);
                const double eff_halo = (double)TH / (TH + 2 * max_halo) * TW / (TW + 2 * max_halo);
                const long total = (long)P->max_batch * tiles_y * cdiv(s.qW, TW);
                const double eff_wave = (double)total / (double)(cdiv((int)total, slots) * (long)slots);
                L.score = eff_m * eff_tail * (0.5 + 0.5 * eff_halo) * (0.3 + 0.7 * eff_wave) + 1e-4 * TH;
                found.push_back(L);
            }
            std::stable_sort(found.begin(), found.end(), [](const Lay& a, const Lay& b) { return a.score > b.score; });
            const int keep = tune ? per_mode : 1;
            for (int i = 0; i < (int)found.size() && i < keep; ++i) out.push_back(found[i]);
        }
        if (out.empty()) {
            // the requested mode (pinned by the environment, or the single default mode when tuning is off) has no
            // feasible tile for this launch at this image size: search every generic mode and narrower tiles
            for (int nb = 2; nb >= 1 && out.empty(); --nb)
                for (int c = 1; c <= 2 && out.empty(); ++c)
                    for (int split = 1; split <= 8; ++split) {
                        const int TW = cdiv(s.qW, nx * split);
                        if (split > 1 && TW < 16) break;
                        for (int TH = 1; TH <= s.qH && TH + 2 * max_halo <= 256; ++TH) {
                            Lay L;
                            if (!layout(s, TH, TW, c, nb, L)) continue;
                            L.score = (double)(TH * TW) / (L.n_mt * 128) + 1e-4 * TH;
                            if (out.empty() || L.score > out[0].score) { out.clear(); out.push_back(L); }
                        }
                        if (!out.empty()) break;
                    }
        }
        // the heuristic's favourite of the default mode stays first (it is what runs without tuning)
        return out;
    }

    // ---- kernel parameters of one layout ----
    fk::Params make_params(const Spec& s, const Lay& L, int& st_C) const {
        fk::Params q;
        std::memset(&q, 0, sizeof q);
        q.TH = L.TH; q.TW = L.TW; q.H = s.qH; q.W = s.qW;
        q.tiles_x = cdiv(s.qW, L.TW); q.tiles_y = cdiv(s.qH, L.TH);
        q.n_mt = L.n_mt; q.tmem_cols = tc::tmem_cols_pow2(L.pipe ? 2 * L.cols : L.cols);
        q.pipe_slots = L.pipe; q.acc_stride = L.cols;
        q.pipe_groups = L.groups;
        auto magic = [](int d) -> uint32_t { return d > 1 ? (uint32_t)(((1ull << 32) + d - 1) / d) : 0u; };
        q.tw_magic = magic(L.TW); q.npix_magic = magic(L.TH * L.TW);
        q.nlbuf = L.nlbuf;
        q.split = P->split ? 1 : 0; q.tab_n = L.tab_n;
        q.nloads = 0;
        for (int i = 0; i < (int)s.loads.size(); ++i) {
            const LoadSpec& l = s.loads[i];
            // split tensors arrive as two boxes through the same map: hi chunks, then the lo chunks Cs/8 later
            // (a folded 2x2 load takes every chunk of its pixel pairs in one box)
            for (int part = 0; part < (P->split && !l.s2d && !l.hi_only ? 2 : 1); ++part) {
                if (q.nloads >= fk::MAX_LOADS) { fprintf(stderr, "libisg: %s: too many TMA loads\n", s.name.c_str()); abort(); }
                fk::Load& d = q.loads[q.nloads++];
                d.map = i; d.smem_off = L.load_off[i] + part * L.load_lo[i]; d.dx = -l.halo; d.dy = -l.halo;
                d.chunk0 = l.chunk0 + part * td(l.tensor).Cs / 8;
                d.folded = l.s2d ? 1 : 0; d.rows_per_img = s.qH; d.step = 1;
                d.planar = td(l.tensor).planar ? 1 : 0;
                if (l.s2d && d.planar) { fprintf(stderr, "libisg: %s: space-to-depth load of a planar tensor\n", s.name.c_str()); abort(); }
            }
        }
        q.lbuf_stride = L.lbuf_stride;
        q.tx_bytes = L.tx; q.blob_off = L.blob_off; q.blob_bytes = L.blob_bytes; q.bar_off = L.bar_off;
        st_C = 0;
        if (s.st.kind) {
            const LoadSpec& l = s.loads[s.st.load];
            fk::Stencil& t = q.st;
            t.on = 1; t.in_off = L.load_off[s.st.load]; t.in_pitch = L.st_in_pitch;
            t.in_plane_bytes = L.load_plane[s.st.load]; t.halo = l.halo; t.nchunks = L.load_planes[s.st.load];
            st_C = t.nchunks * 8;
            t.dil = s.st.kind == 1 ? s.st.dil : 1;
            t.sep5 = s.st.kind == 2 ? 1 : 0;
            t.w_off = L.blob_off + L.stw_off; t.vec_off = L.blob_off + L.stvec_off;
            t.out_off = L.st_out_off; t.out_plane_bytes = L.st_out_plane;
            t.in_lo = L.load_lo[s.st.load]; t.out_lo = L.st_out_lo;
        }
        q.nstages = (int)s.stages.size();
        for (int k = 0; k < q.nstages; ++k) {
            const StageSpec& st = s.stages[k];
            fk::Stage& d = q.stage[k];
            d.N = st.N; d.ntaps = st.taps9 ? 9 : 1; d.pitch = L.pitch; d.pitch_magic = magic(L.pitch);
            for (int t = 0; t < d.ntaps; ++t) d.tap_off[t] = st.taps9 ? (t / 3) * L.pitch + (t % 3) : 0;
            int nks = 0;
            for (auto& pt : st.parts) {
                int base, plane, planes, lo = 0, sub_planes = 0;   // sub_planes > 0: folded split load, hi planes per sub-pixel
                bool in_load = false;
                if (pt.kind == PART_LOAD) {
                    base = L.load_off[pt.load]; plane = L.load_plane[pt.load]; planes = L.load_planes[pt.load]; in_load = true;
                    lo = L.load_lo[pt.load];
                    if (P->split && s.loads[pt.load].s2d) sub_planes = td(s.loads[pt.load].tensor).Cs / 8;
                }
                else if (pt.kind == PART_STENCIL) { base = L.st_out_off; plane = L.st_out_plane; planes = st_C / 8; lo = L.st_out_lo; }
                else { base = L.stage_off[0]; plane = L.stage_plane[0]; planes = s.stages[0].N / 8; in_load = L.stage_in_load[0]; lo = L.stage_lo[0]; }
                for (int j = 0; j < planes / 2; ++j) {
                    if (in_load) d.a_buf_mask |= 1u << nks;
                    int pl = 2 * j;   // first of the two hi planes of this K slice
                    if (sub_planes) pl = (pl / sub_planes) * 2 * sub_planes + pl % sub_planes;   // [hi | lo] per sub-pixel
                    d.a_lo[nks] = (uint32_t)lo >> 4;
                    d.a_desc[nks++] = tc::umma_smem_desc(base + pl * plane, plane, 128);
                }
            }
            d.stage_in_load = L.stage_in_load[k];
            d.nks = nks;
            d.nent = 0;
            for (int t = 0; t < d.ntaps; ++t) {
                d.tap_mask[t] = 0;
                for (int ks = 0; ks < nks; ++ks)
                    if (!st.used || st.used(t, ks)) { d.tap_mask[t] |= 1u << ks; ++d.nent; }
            }
            d.passes = passes_of(st);
            d.nmerge = nmerge_of(st) ? 1 : 0;
            d.nent_tab = d.nent * (d.nmerge ? ((d.passes & 2) ? 2 : 1) : npasses(d.passes));
            d.b_lo = (uint32_t)d.nent * st.N * 2;
            if (P->split ? d.nent_tab > L.tab_n : d.nent > fk::TAB_N) { fprintf(stderr, "libisg: %s: %d MMAs per M tile exceed the descriptor table\n", s.name.c_str(), d.nent); abort(); }
            // merged blocks are [2][2N][8]: the K chunks of a block are 2N rows apart
            d.b_desc = tc::umma_smem_desc(L.blob_off + L.b_off[k], (d.nmerge ? 2 : 1) * st.N * 16, 128);
            d.vec_off = L.blob_off + L.vec_off[k];
            if (st.res_load >= 0) {
                d.res_on = 1; d.res_off = L.load_off[st.res_load]; d.res_plane_bytes = L.load_plane[st.res_load];
                d.res_lo = L.load_lo[st.res_load];
                d.res_mod = st.res_mod;
                if (st.res_mod % 16) { fprintf(stderr, "libisg: res_mod must be a multiple of 16\n"); abort(); }
                d.res_magic = magic(st.res_mod / 16);
            }
            d.out_mode = st.out_mode; d.stage_off = L.stage_off[k]; d.stage_plane_bytes = L.stage_plane[k];
            d.stage_compact = L.stage_compact[k]; d.store = st.store && st.out_mode == fk::OUT_STAGE;
            d.store_chunk0 = st.out_chunk0; d.shuf = st.shuf; d.Cp = st.Cp;
            if (st.out_mode == fk::OUT_SHUF_BF16) {
                if (st.shuf != 2 || st.Cp % 16) { fprintf(stderr, "libisg: pixel shuffle needs shuf 2, Cp %% 16 == 0\n"); abort(); }
                d.cp_magic = magic(st.Cp / 16);
            }
            if (st.out_tensor >= 0) { d.Cs_out = td(st.out_tensor).Cs; d.out_planar = td(st.out_tensor).planar ? 1 : 0; }
            d.split_out = (P->split && st.out_mode != fk::OUT_SHUF4_F32) ? 1 : 0;
            d.stage_lo = L.stage_lo[k];
            if (st.out_tensor >= 0) { d.out_lo_chunks = td(st.out_tensor).Cs / 8; d.Ct_out = td(st.out_tensor).Ct(); }
        }
        if (s.pool_tensor >= 0) {
            q.pool_on = 1; q.pool_in_off = L.load_off[0]; q.pool_in_plane_bytes = L.load_plane[0];
            q.pool_nchunks = td(s.pool_tensor).Cs / 8; q.pool_off = L.pool_off; q.pool_plane_bytes = L.pool_plane;
            q.pool_planar = td(s.pool_tensor).planar ? 1 : 0;
            q.pool_grp_planes = q.pool_nchunks * (P->split ? 2 : 1);
            q.pool_lo_in = L.load_lo[0]; q.pool_lo_out = L.pool_lo; q.pool_lo_chunks = q.pool_nchunks;
        }
        // single load buffer: may the next tile's boxes be issued as soon as the stage-0 MMAs completed?
        q.early_load = s.pool_tensor >= 0 ? 0 : 1;
        for (int k = 0; k < q.nstages; ++k)
            if (q.stage[k].res_on || q.stage[k].stage_in_load || (k > 0 && q.stage[k].a_buf_mask)) q.early_load = 0;
        return q;
    }

    // one way to run a fused launch: tile, CTAs per SM, load buffering
    struct Variant { Lay L; fk::Params q; std::shared_ptr<FkMaps> maps; };

    // ---- emit one fused launch ----
    void emit(const Spec& s) {
        isg_plan* P = this->P;
        std::vector<Lay> cands = candidates(s);
        if (cands.empty()) {
            fprintf(stderr, "libisg: no feasible tile for %s\n", s.name.c_str());
            abort();
        }
        auto vars = std::make_shared<std::vector<Variant>>();
        int st_C = 0;
        for (const Lay& c : cands) vars->push_back(Variant{c, make_params(s, c, st_C), std::make_shared<FkMaps>()});
        const Lay L = cands[0];   // blob layout (tile independent)
        auto chosen = std::make_shared<int>(0);
        // constants blob in the arena
        const size_t blob = arena((size_t)L.blob_bytes / 4);
        const Spec sp = s;
        std::vector<int> pass_bits;
        std::vector<int> merge_bits;
        for (auto& st : s.stages) { pass_bits.push_back(passes_of(st)); merge_bits.push_back(nmerge_of(st) ? 1 : 0); }
        Op op;
        op.name = s.name + " [fk]";
        op.pack = [=](const HostWeights& w, float* A) {
            uint8_t* base = reinterpret_cast<uint8_t*>(A + blob);
            std::memset(base, 0, L.blob_bytes);
            for (size_t k = 0; k < sp.stages.size(); ++k) {
                const StageSpec& st = sp.stages[k];
                const int ntaps = st.taps9 ? 9 : 1, K = L.K[k], N = st.N, nks = K / 16;
                std::vector<float> B((size_t)ntaps * K * N, 0.f), shift(N, 0.f), slope(N, 1.f);
                st.fill(w, B.data(), shift.data(), slope.data());
                uint16_t* wb = reinterpret_cast<uint16_t*>(base + L.b_off[k]);
                // blocks of structurally zero K slices are not stored: the kernel skips their MMAs
                std::vector<int> slot((size_t)ntaps * nks, -1);
                int nslot = 0;
                for (int t = 0; t < ntaps; ++t)
                    for (int ks = 0; ks < nks; ++ks)
                        if (!st.used || st.used(t, ks)) slot[(size_t)t * nks + ks] = nslot++;
                const bool lo_blocks = (pass_bits[k] & 4) != 0;
                for (int t = 0; t < ntaps; ++t)
                    for (int kk = 0; kk < K; ++kk)
                        for (int n = 0; n < N; ++n) {
                            const float v = B[((size_t)t * K + kk) * N + n];
                            const int blk = slot[(size_t)t * nks + kk / 16];
                            if (blk < 0) {
                                if (v != 0.f) { fprintf(stderr, "libisg: %s: weight in a K slice declared unused\n", sp.name.c_str()); abort(); }
                                continue;
                            }
                            const uint16_t hi = tc::f2bf(v);
                            if (merge_bits[k]) {   // [W_hi | W_lo] side by side: blocks of [2][2N][8]
                                const size_t idx = (size_t)blk * N * 32 + (size_t)((kk % 16) / 8) * N * 16 + (size_t)n * 8 + (kk % 8);
                                wb[idx] = hi;
                                wb[idx + (size_t)N * 8] = tc::f2bf(v - tc::bf2f(hi));
                                continue;
                            }
                            const size_t idx = (size_t)blk * N * 16 + (size_t)((kk % 16) / 8) * N * 8 + (size_t)n * 8 + (kk % 8);
                            wb[idx] = hi;
                            // split precision: W_lo = bf16(w - W_hi) in a second run of blocks (pass A_hi.W_lo)
                            if (lo_blocks) wb[(size_t)nslot * N * 16 + idx] = tc::f2bf(v - tc::bf2f(hi));
                        }
                float* v = reinterpret_cast<float*>(base + L.vec_off[k]);
                std::memcpy(v, shift.data(), sizeof(float) * N);
                std::memcpy(v + N, slope.data(), sizeof(float) * N);
            }
            if (sp.st.kind) {
                const int C = st_C, ntaps = sp.st.kind == 2 ? 25 : 9;
                std::vector<float> wt((size_t)ntaps * C, 0.f), shift(C, 0.f), slope(C, 1.f), bw(5 * (size_t)C, 0.f);
                sp.st.fill(w, wt.data(), shift.data(), slope.data(), bw.data());
                std::memcpy(base + L.stw_off, wt.data(), sizeof(float) * wt.size());
                float* v = reinterpret_cast<float*>(base + L.stvec_off);
                std::memcpy(v, shift.data(), sizeof(float) * C);
                std::memcpy(v + C, slope.data(), sizeof(float) * C);
                if (sp.st.kind == 2) std::memcpy(v + 2 * C, bw.data(), sizeof(float) * 5 * C);
            }
        };
        const int maxB = P->max_batch;
        P->finalizers.push_back([=]() -> int {
            for (Variant& v : *vars) {
                const Lay& L = v.L;
                FkMaps* maps = v.maps.get();
                for (size_t i = 0; i < sp.loads.size(); ++i) {
                    const LoadSpec& l = sp.loads[i];
                    const TensorD& t = P->tensors[l.tensor];
                    int rc;
                    if (l.s2d) rc = tc::make_s2d_map(&maps->in[i], P->ws + t.off, maxB, t.H, t.W, t.Ct(), L.TW, L.TH, sp.name.c_str());
                    else rc = tc::make_act_map_chunks(&maps->in[i], P->ws + t.off, maxB, t.H, t.W, t.Ct(), L.TW + 2 * l.halo,
                                                      L.TH + 2 * l.halo, L.load_planes[i], sp.name.c_str(), t.planar);
                    if (rc) return rc;
                }
                for (size_t i = sp.loads.size(); i < 4; ++i) maps->in[i] = maps->in[0];
                for (size_t k = 0; k < 2; ++k) {
                    if (k < sp.stages.size() && sp.stages[k].out_mode == fk::OUT_STAGE && sp.stages[k].store) {
                        const TensorD& t = P->tensors[sp.stages[k].out_tensor];
                        int rc = tc::make_act_map_chunks(&maps->out[k], P->ws + t.off, maxB, t.H, t.W, t.Ct(), L.TW, L.TH,
                                                         sp.stages[k].N / 8, sp.name.c_str(), t.planar);
                        if (rc) return rc;
                    } else maps->out[k] = maps->in[0];
                }
                if (sp.pool_tensor >= 0) {
                    const TensorD& t = P->tensors[sp.pool_tensor];
                    int rc = tc::make_act_map_chunks(&maps->out[1], P->ws + t.off, maxB, t.H, t.W, t.Ct(), L.TW, L.TH, t.Cs / 8,
                                                     sp.name.c_str(), t.planar);
                    if (rc) return rc;
                }
            }
            return 0;
        });
        cudaFuncSetAttribute(fk::fk_kernel<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        cudaFuncSetAttribute(fk::fk_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024);
        cudaFuncSetAttribute(fk::fk_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 56 * 1024);
        cudaFuncSetAttribute(fk::fk_pipe_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        if (P->split) {
            cudaFuncSetAttribute(fk::fk_kernel<512, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
            cudaFuncSetAttribute(fk::fk_kernel<256, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 113 * 1024);
            cudaFuncSetAttribute(fk::fk_kernel<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 56 * 1024);
            cudaFuncSetAttribute(fk::fk_pipe_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        }
        int num_sms = 148;
        cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, P->device);
        auto run = [=](int which, int n, float* logits, cudaStream_t stream, bool prof) -> cudaError_t {
            const Variant& v = (*vars)[which];
            fk::Params p = v.q;
            p.blob = P->arena + blob;
            for (int k = 0; k < p.nstages; ++k) {
                const StageSpec& st = sp.stages[k];
                if (st.to_logits) p.stage[k].gptr = logits;
                else if (st.out_mode != fk::OUT_STAGE) p.stage[k].gptr = P->ws + P->tensors[st.out_tensor].off;
            }
            p.total_tiles = n * p.tiles_x * p.tiles_y;
            // persistent: one 512-thread CTA or two 256-thread CTAs per SM
            const int grid = std::min(p.total_tiles, num_sms * v.L.ctas);
            if (prof) return launch_profiled(sp.name, p, v.L.ctas, grid, v.L.smem, *v.maps, stream);
            return launch_fk(v.L.ctas, grid, v.L.smem, *v.maps, p, stream);
        };
        op.launch = [=](int n, const float*, float* logits, cudaStream_t stream) {
            static const bool prof_on = getenv("ISG_FK_PROF") != nullptr;
            return run(*chosen, n, logits, stream, prof_on);
        };
        // autotune: time every variant at the plan's batch size on the device, keep the fastest
        const size_t site = P->tune_sites.size();
        {
            isg_plan::TuneSite ts;
            ts.name = sp.name;
            ts.iso_ms.assign(vars->size(), 0.f);
            ts.set = [=](int i) { *chosen = i; };
            ts.get = [=]() { return *chosen; };
            P->tune_sites.push_back(ts);
        }
        P->tuners.push_back([=](float* logits, cudaStream_t stream) -> int {
            const bool verbose = getenv("ISG_FK_VERBOSE") != nullptr;
            float best = 1e30f;
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            for (int i = 0; i < (int)vars->size() && vars->size() > 1; ++i) {
                const int reps = 3;
                if (run(i, maxB, logits, stream, false) != cudaSuccess) return ISG_ERR_CUDA;
                cudaEventRecord(e0, stream);
                for (int r = 0; r < reps; ++r) run(i, maxB, logits, stream, false);
                cudaEventRecord(e1, stream);
                if (cudaEventSynchronize(e1) != cudaSuccess) return ISG_ERR_CUDA;
                float ms = 0;
                cudaEventElapsedTime(&ms, e0, e1);
                ms /= reps;
                const Lay& L = (*vars)[i].L;
                if (verbose)
                    fprintf(stderr, "tune %-40s tile %3dx%-3d n_mt %2d ctas %d lbuf %d pipe %d smem %6zu: %7.1f us\n", sp.name.c_str(),
                            L.TH, L.TW, L.n_mt, L.ctas, L.nlbuf, L.pipe * 10 + (L.pipe ? L.groups : 0), L.smem, ms * 1e3);
                P->tune_sites[site].iso_ms[i] = ms;
                if (ms < best) { best = ms; *chosen = i; }
            }
            cudaEventDestroy(e0); cudaEventDestroy(e1);
            const Lay& L = (*vars)[*chosen].L;
            if (verbose) {
                // issue floor: every MMA of this shape costs >= 32 clk (4 KB A operand from shared memory)
                const fk::Params& q = (*vars)[*chosen].q;
                long mma = 0;
                for (int k = 0; k < q.nstages; ++k) mma += (long)q.stage[k].nent * q.n_mt;
                mma *= (long)maxB * q.tiles_x * q.tiles_y;
                fprintf(stderr, "fk %-34s q %4dx%-4d tile %3dx%-3d n_mt %2d cols %3d smem %6zu K %d,%d ctas %d lbuf %d pipe %d"
                        "  best %6.1f us  mma %ld (floor %.1f us)\n",
                        sp.name.c_str(), sp.qH, sp.qW, L.TH, L.TW, L.n_mt, L.cols, L.smem, L.K[0], L.K[1], L.ctas, L.nlbuf,
                        L.pipe * 10 + (L.pipe ? L.groups : 0), best * 1e3, mma, mma * 32.0 / 148 / 1965.0);
            }
            return 0;
        });
        P->ops.push_back(op);
    }

    // ---- plain CUDA-core launches ----
    void emit_prep(int t_x2, int t_init) {
        isg_plan* P = this->P;
        const int cin = P->cin, H = P->H, W = P->W;
        const TensorD x2 = td(t_x2), init = td(t_init);
        Op op;
        op.name = "prep: NCHW fp32 -> space-to-depth bf16 + max_pool4";
        op.pack = [](const HostWeights&, float*) {};
        op.launch = [=](int n, const float* x, float*, cudaStream_t st) {
            bf16* px2 = reinterpret_cast<bf16*>(P->ws + x2.off);
            bf16* pin = reinterpret_cast<bf16*>(P->ws + init.off);
            if (cin == 3 && P->u8_src) {
                NormU8 nm;
                for (int c = 0; c < 3; ++c) { nm.mean[c] = P->u8_mean[c]; nm.std[c] = P->u8_std[c]; }
                nm.exact = P->exact_u8 ? 1 : 0;
                const int total = n * (H / 4) * (W / 4);
                return launch_pdl(prep_u8_kernel, cdiv(total, 128), 128, st, P->u8_src, H, W, px2, x2.Ct(), (int)x2.planar, pin, init.Ct(), total, nm,
                                  (int)x2.split, init.Cs);
            }
            if (cin == 3) {
                const int total = n * (H / 4) * (W / 4);
                return launch_pdl(prep_kernel<3>, cdiv(total, 128), 128, st, x, H, W, px2, x2.Ct(), (int)x2.planar, pin, init.Ct(), total,
                                  (int)x2.split, init.Cs);
            }
            const int total = n * cin * (H / 4) * (W / 4);
            return launch_pdl(prep_generic_kernel, cdiv(total, 128), 128, st, x, cin, H, W, px2, x2.Ct(), (int)x2.planar, pin, init.Ct(), total,
                              (int)x2.split, x2.Cs, init.Cs);
        };
        P->ops.push_back(op);
    }
    void emit_pool2(const std::string& name, int tin, int tout) {
        isg_plan* P = this->P;
        const TensorD ti = td(tin), to = td(tout);
        Op op;
        op.name = name;
        op.pack = [](const HostWeights&, float*) {};
        op.launch = [=](int n, const float*, float*, cudaStream_t st) {
            int total = n * to.H * to.W * (to.Cs / 4);
            return launch_pdl(pool2_kernel<bf16>, cdiv(total, 256), 256, st, reinterpret_cast<const bf16*>(P->ws + ti.off),
                              reinterpret_cast<bf16*>(P->ws + to.off), to.H, to.W, to.Cs, total);
        };
        P->ops.push_back(op);
    }

    // ---- fill helpers ----
    // 1x1 conv weights Wt[Cout][cin_total] (columns cin_off.. of the logical input) times
    // scale[o] into rows [k0, k0 + 8*nchunks) (storage channels of tensor t) and columns
    // [n0, n0 + Cout) of B[K][N]
    static void put_pw(float* B, int N, int k0, const TensorD& t, int chunk0, int nchunks, const float* Wt,
                       int cin_total, int cin_off, const float* scale, int Cout, int n0) {
        for (int j = 0; j < nchunks * 8; ++j) {
            const int l = t.logical_of(chunk0 * 8 + j);
            if (l < 0) continue;
            for (int o = 0; o < Cout; ++o)
                B[(size_t)(k0 + j) * N + n0 + o] = Wt[(size_t)o * cin_total + cin_off + l] * (scale ? scale[o] : 1.f);
        }
    }

    // next block's first 1x1 conv (Conv wrapper `unit`) as a chained stage over the producer's staging
    struct Next { bool on = false; std::string unit; int mid = 0; int act = ACT_PRELU; int out = -1; };
    StageSpec chain_stage(const Next& nx, int Cprev /*logical channels of staging*/, int Nprev) {
        StageSpec st;
        st.parts = {Part{PART_PREV, -1}};
        st.N = round_up(nx.mid, 16);
        st.out_tensor = nx.out;
        const std::string unit = nx.unit;
        const int mid = nx.mid, act = nx.act, N = st.N;
        (void)Nprev;
        st.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(mid), sh(mid);
            fold_unit(w, unit + ".conv", unit + ".bn", mid, sc.data(), sh.data());
            const float* Wt = w.get(unit + ".conv.weight");  // [mid][Cprev]
            for (int k = 0; k < Cprev; ++k)
                for (int o = 0; o < mid; ++o) B[(size_t)k * N + o] = Wt[(size_t)o * Cprev + k] * sc[o];
            for (int o = 0; o < mid; ++o) {
                shift[o] = sh[o];
                slope[o] = act == ACT_PRELU ? w.get(unit + ".act.weight")[o] : 0.f;
            }
        };
        return st;
    }
    Next next_of(const std::string& unit, int mid, int act, int H, int W) {
        Next n;
        n.on = true; n.unit = unit; n.mid = mid; n.act = act;
        n.out = tensor(unit + ".out", mid, H, W);
        return n;
    }

    StencilSpec dw3_stencil(const std::string& unit, int load, int mid, int dil) {
        StencilSpec s;
        s.kind = 1; s.load = load; s.dil = dil;
        s.fill = [=](const HostWeights& w, float* wt, float* shift, float* slope, float*) {
            const int C = round_up(mid, 16);
            std::vector<float> sc(mid), sh(mid);
            fold_unit(w, unit + ".conv", unit + ".bn", mid, sc.data(), sh.data());
            const float* k = w.get(unit + ".conv.weight");  // [mid][1][3][3]
            for (int c = 0; c < mid; ++c) {
                for (int t = 0; t < 9; ++t) wt[(size_t)t * C + c] = k[(size_t)c * 9 + t] * sc[c];
                shift[c] = sh[c];
                slope[c] = w.get(unit + ".act.weight")[c];
            }
        };
        return s;
    }
    // Bottleneck5x5 (segment.py:91-97): bare depthwise (5,1) + bias, then (1,5) + BN + PReLU
    StencilSpec sep5_stencil(const std::string& p, int load, int mid) {
        StencilSpec s;
        s.kind = 2; s.load = load;
        s.fill = [=](const HostWeights& w, float* wt, float* shift, float* slope, float* bw) {
            const int C = round_up(mid, 16);
            std::vector<float> sc(mid), sh(mid);
            fold_unit(w, p + ".convs.2.conv", p + ".convs.2.bn", mid, sc.data(), sh.data());
            const float* w1 = w.get(p + ".convs.1.weight");        // [mid][1][5][1]
            const float* b1 = w.get(p + ".convs.1.bias");
            const float* w2 = w.get(p + ".convs.2.conv.weight");   // [mid][1][1][5]
            for (int c = 0; c < mid; ++c) {
                for (int ky = 0; ky < 5; ++ky)
                    for (int kx = 0; kx < 5; ++kx)
                        wt[(size_t)(ky * 5 + kx) * C + c] = sc[c] * w2[c * 5 + kx] * w1[c * 5 + ky];
                for (int kx = 0; kx < 5; ++kx) bw[(size_t)kx * C + c] = sc[c] * w2[c * 5 + kx] * b1[c];
                shift[c] = sh[c];
                slope[c] = w.get(p + ".convs.2.act.weight")[c];
            }
        };
        return s;
    }

    // standalone 1x1 Conv wrapper over the concat of `srcs`
    int pw_unit(const std::string& unit, const std::vector<int>& srcs, int cout, int act, const std::string& outname,
                int out_existing = -1) {
        const TensorD t0 = td(srcs[0]);
        Spec s;
        s.name = unit;
        s.qH = t0.H; s.qW = t0.W;
        StageSpec st;
        std::vector<TensorD> tin;
        int cin_total = 0;
        for (size_t i = 0; i < srcs.size(); ++i) {
            s.loads.push_back(LoadSpec{srcs[i]});
            st.parts.push_back(Part{PART_LOAD, (int)i});
            tin.push_back(td(srcs[i]));
            cin_total += td(srcs[i]).C;
        }
        st.N = round_up(cout, 16);
        const int out = out_existing >= 0 ? out_existing : tensor(outname, cout, t0.H, t0.W);
        st.out_tensor = out;
        macs(cout, cin_total, 1, 1, t0.H, t0.W);
        const int N = st.N;
        st.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(cout), sh(cout);
            fold_unit(w, unit + ".conv", unit + ".bn", cout, sc.data(), sh.data());
            int k0 = 0, c0 = 0;
            for (auto& t : tin) {
                put_pw(B, N, k0, t, 0, t.Cs / 8, w.get(unit + ".conv.weight"), cin_total, c0, sc.data(), cout, 0);
                k0 += t.Cs; c0 += t.C;
            }
            for (int o = 0; o < cout; ++o) {
                shift[o] = sh[o];
                slope[o] = act == ACT_PRELU ? w.get(unit + ".act.weight")[o] : (act == ACT_RELU ? 0.f : 1.f);
            }
        };
        s.stages.push_back(st);
        emit(s);
        return out;
    }

    // Emit a block launch whose last stage may be followed by the next block's first 1x1 conv as a chained stage.
    // Split-precision plans hold twice the planes and weights per tile: when the fused launch has no feasible tile the
    // chained conv becomes a launch of its own (same arithmetic, one more round trip of `out`).
    void emit_chained(Spec& s, const Next& nx, int Cprev, int Nprev, int out, int H, int W) {
        if (nx.on) {
            s.stages.push_back(chain_stage(nx, Cprev, Nprev));
            if (P->split && candidates(s).empty()) {
                s.stages.pop_back();
                emit(s);
                pw_unit(nx.unit, {out}, nx.mid, nx.act, "", nx.out);
                return;
            }
            macs(nx.mid, Cprev, 1, 1, H, W);
        }
        emit(s);
    }

    // one K range of a 1x1 Conv wrapper as its own launch: out = W[:, cin_off : cin_off + C(src)] . src (+ res) (+ shift);
    // no activation.  Used to split a convolution over a concatenation along K when the whole K does not fit a tile.
    int pw_partial(const std::string& name, const std::string& unit, int src, int cin_total, int cin_off, int cout, int res,
                   bool with_shift) {
        const TensorD t0 = td(src);
        Spec s;
        s.name = name;
        s.qH = t0.H; s.qW = t0.W;
        s.loads.push_back(LoadSpec{src});
        StageSpec st;
        st.parts = {Part{PART_LOAD, 0}};
        st.N = round_up(cout, 16);
        if (res >= 0) { s.loads.push_back(LoadSpec{res}); st.res_load = 1; }
        const int out = tensor(name, cout, t0.H, t0.W);
        st.out_tensor = out;
        const int N = st.N;
        st.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(cout), sh(cout);
            fold_unit(w, unit + ".conv", unit + ".bn", cout, sc.data(), sh.data());
            put_pw(B, N, 0, t0, 0, t0.Cs / 8, w.get(unit + ".conv.weight"), cin_total, cin_off, sc.data(), cout, 0);
            for (int o = 0; o < cout; ++o) { shift[o] = with_shift ? sh[o] : 0.f; slope[o] = 1.f; }
        };
        s.stages.push_back(st);
        emit(s);
        return out;
    }

    // Bottleneck3x3 / Bottleneck5x5 (segment.py:52-111) given the block's first conv output m1
    int block_res(const std::string& p, int x, int m1, int mid, int dil, bool five, const Next& nx,
                  const std::string& outname, bool tap) {
        const TensorD tx = td(x);
        const int C = tx.C;
        const std::string last = five ? p + ".convs.3" : p + ".convs.2";
        Spec s;
        s.name = p;
        s.qH = tx.H; s.qW = tx.W;
        s.loads.push_back(LoadSpec{m1, 0, -1, five ? 2 : dil});
        s.loads.push_back(LoadSpec{x});
        s.st = five ? sep5_stencil(p, 0, mid) : dw3_stencil(p + ".convs.1", 0, mid, dil);
        if (five) { macs(mid, 1, 5, 1, tx.H, tx.W); macs(mid, 1, 1, 5, tx.H, tx.W); }
        else macs(mid, 1, 3, 3, tx.H, tx.W);
        macs(C, mid, 1, 1, tx.H, tx.W);
        StageSpec st;
        st.parts = {Part{PART_STENCIL, -1}};
        st.N = tx.Cs;
        st.res_load = 1;
        const int out = tensor(outname, C, tx.H, tx.W, tap);
        st.out_tensor = out;
        const int N = st.N;
        st.fill = res_out_fill(p, last, C, mid, N);
        s.stages.push_back(st);
        emit_chained(s, nx, C, N, out, tx.H, tx.W);
        return out;
    }
    // last 1x1 conv + BN of a residual bottleneck and the block's PReLU: B[K = mid][N], shift, slope
    static StageFill res_out_fill(const std::string& p, const std::string& last, int C, int mid, int N) {
        return [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(C), sh(C);
            fold_unit(w, last + ".conv", last + ".bn", C, sc.data(), sh.data());
            const float* Wt = w.get(last + ".conv.weight");  // [C][mid]
            for (int k = 0; k < mid; ++k)
                for (int o = 0; o < C; ++o) B[(size_t)k * N + o] = Wt[(size_t)o * mid + k] * sc[o];
            for (int o = 0; o < C; ++o) { shift[o] = sh[o]; slope[o] = w.get(p + ".prelu.weight")[o]; }
        };
    }

    // ---- a run of residual bottlenecks as ONE cluster launch (csrc/mega.cuh) ----
    struct MegaLay { mega::Params q; size_t smem = 0; bool ok = false; };
    MegaLay mega_layout(int H, int W, int C, int mid, int nblocks) const {
        MegaLay L;
        mega::Params& q = L.q;
        std::memset(&q, 0, sizeof q);
        // split-precision plans keep hi + lo of x, m1 and d: twice the shared memory a 4-CTA cluster has -> per-block launches
        if (P->split || getenv("ISG_FK_NO_MEGA") || nblocks > mega::MAX_BLOCKS || H % mega::CLUSTER) return L;
        q.H = H; q.W = W; q.R = H / mega::CLUSTER;
        q.Cs = round_up(C, 16); q.Cm = round_up(mid, 16);
        q.npx = q.R * W;
        if (q.npx % 128 || q.R < mega::HALO || W > 128 || q.R > 256) return L;   // W <= 128: one planar box row
        q.n_mt = q.npx / 128;
        q.pitch = W + 2 * mega::HALO; q.rows_p = q.R + 2 * mega::HALO;
        auto magic = [](int d) -> uint32_t { return d > 1 ? (uint32_t)(((1ull << 32) + d - 1) / d) : 0u; };
        q.w_magic = magic(W); q.npx_magic = magic(q.npx);
        q.nblocks = nblocks;
        const int cols = q.n_mt * std::max(q.Cs, q.Cm);
        if (cols > 512) return L;
        q.tmem_cols = tc::tmem_cols_pow2(cols);
        int rel = 0;
        q.w0_off = rel; rel = r128(rel + q.Cs * q.Cm * 2);
        q.v0_off = rel; rel = r128(rel + 2 * q.Cm * 4);
        q.dw_off = rel; rel = r128(rel + 25 * q.Cm * 4);
        q.dv_off = rel; rel = r128(rel + 7 * q.Cm * 4);
        q.w2_off = rel; rel = r128(rel + q.Cm * q.Cs * 2);
        q.v2_off = rel; rel = r128(rel + 2 * q.Cs * 4);
        q.blob_bytes = rel; q.blob_stride = rel;
        int off = 0;
        q.x_off = off; off = r128(off + (q.Cs / 8) * q.npx * 16);
        q.m1_off = off; off = r128(off + (q.Cm / 8) * q.rows_p * q.pitch * 16);
        q.d_off = off; off = r128(off + (q.Cm / 8) * q.npx * 16);
        q.wbuf_off[0] = off; off = r128(off + rel);
        q.wbuf_off[1] = off; off = r128(off + rel);
        q.bar_off = off; off += 128;
        L.smem = off;
        L.ok = L.smem <= 227 * 1024;
        return L;
    }
    struct MegaBlock { std::string p; int dil; bool five; };
    int mega_run(const std::string& name, int x, int mid, const std::vector<MegaBlock>& blocks, const std::string& outname) {
        isg_plan* P = this->P;
        const TensorD tx = td(x);
        const int C = tx.C;
        MegaLay L = mega_layout(tx.H, tx.W, C, mid, (int)blocks.size());
        mega::Params q = L.q;
        const int out = tensor(outname, C, tx.H, tx.W, true);
        struct Fills { StageFill f0, f2; StencilFill fs; };
        auto fills = std::make_shared<std::vector<Fills>>();
        for (size_t i = 0; i < blocks.size(); ++i) {
            const MegaBlock& b = blocks[i];
            q.blk[i].dil = b.five ? 1 : b.dil; q.blk[i].sep5 = b.five ? 1 : 0;
            q.blk[i].sr = 4 * q.blk[i].dil <= q.R ? 4 : 2;   // stencil strip height (rows `dil` apart) within the CTA's rows
            Next n0; n0.on = true; n0.unit = b.p + ".convs.0"; n0.mid = mid; n0.act = ACT_PRELU;
            Fills f;
            f.f0 = chain_stage(n0, C, q.Cs).fill;
            f.fs = b.five ? sep5_stencil(b.p, 0, mid).fill : dw3_stencil(b.p + ".convs.1", 0, mid, b.dil).fill;
            f.f2 = res_out_fill(b.p, b.five ? b.p + ".convs.3" : b.p + ".convs.2", C, mid, q.Cs);
            fills->push_back(f);
            macs(mid, C, 1, 1, tx.H, tx.W);
            if (b.five) { macs(mid, 1, 5, 1, tx.H, tx.W); macs(mid, 1, 1, 5, tx.H, tx.W); }
            else macs(mid, 1, 3, 3, tx.H, tx.W);
            macs(C, mid, 1, 1, tx.H, tx.W);
        }
        const size_t blob = arena((size_t)q.blob_stride * blocks.size() / 4);
        Op op;
        op.name = name + " (" + std::to_string(blocks.size()) + " residual bottlenecks, cluster-resident) [mega]";
        const int nb = (int)blocks.size();
        op.pack = [=](const HostWeights& w, float* A) {
            auto pack_b = [](uint16_t* wb, const float* B, int K, int N) {   // B[K][N] -> blocked [k16][2][N][8] bf16
                for (int kk = 0; kk < K; ++kk)
                    for (int n = 0; n < N; ++n)
                        wb[(size_t)(kk / 16) * N * 16 + (size_t)((kk % 16) / 8) * N * 8 + (size_t)n * 8 + (kk % 8)] =
                            tc::f2bf(B[(size_t)kk * N + n]);
            };
            for (int i = 0; i < nb; ++i) {
                uint8_t* base = reinterpret_cast<uint8_t*>(A + blob) + (size_t)i * q.blob_stride;
                std::memset(base, 0, q.blob_bytes);
                const Fills& f = (*fills)[i];
                {
                    std::vector<float> B((size_t)q.Cs * q.Cm, 0.f), sh(q.Cm, 0.f), sl(q.Cm, 1.f);
                    f.f0(w, B.data(), sh.data(), sl.data());
                    pack_b(reinterpret_cast<uint16_t*>(base + q.w0_off), B.data(), q.Cs, q.Cm);
                    float* v = reinterpret_cast<float*>(base + q.v0_off);
                    std::memcpy(v, sh.data(), 4 * q.Cm); std::memcpy(v + q.Cm, sl.data(), 4 * q.Cm);
                }
                {
                    std::vector<float> wt((size_t)25 * q.Cm, 0.f), sh(q.Cm, 0.f), sl(q.Cm, 1.f), bw((size_t)5 * q.Cm, 0.f);
                    f.fs(w, wt.data(), sh.data(), sl.data(), bw.data());
                    std::memcpy(base + q.dw_off, wt.data(), 4 * wt.size());
                    float* v = reinterpret_cast<float*>(base + q.dv_off);
                    std::memcpy(v, sh.data(), 4 * q.Cm); std::memcpy(v + q.Cm, sl.data(), 4 * q.Cm);
                    std::memcpy(v + 2 * q.Cm, bw.data(), 4 * bw.size());
                }
                {
                    std::vector<float> B((size_t)q.Cm * q.Cs, 0.f), sh(q.Cs, 0.f), sl(q.Cs, 1.f);
                    f.f2(w, B.data(), sh.data(), sl.data());
                    pack_b(reinterpret_cast<uint16_t*>(base + q.w2_off), B.data(), q.Cm, q.Cs);
                    float* v = reinterpret_cast<float*>(base + q.v2_off);
                    std::memcpy(v, sh.data(), 4 * q.Cs); std::memcpy(v + q.Cs, sl.data(), 4 * q.Cs);
                }
            }
        };
        auto maps = std::make_shared<FkMaps>();
        const int maxB = P->max_batch;
        const TensorD to = td(out);
        P->finalizers.push_back([=]() -> int {
            int rc = tc::make_act_map_chunks(&maps->in[0], P->ws + tx.off, maxB, tx.H, tx.W, tx.Cs, tx.W, q.R, tx.Cs / 8, name.c_str(),
                                             tx.planar);
            if (rc) return rc;
            return tc::make_act_map_chunks(&maps->out[0], P->ws + to.off, maxB, to.H, to.W, to.Cs, to.W, q.R, to.Cs / 8, name.c_str(),
                                           to.planar);
        });
        cudaFuncSetAttribute(mega::mega_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        cudaFuncSetAttribute(mega::mega_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        const size_t smem = L.smem;
        auto nclusters = std::make_shared<int>(0);
        op.launch = [=](int n, const float*, float*, cudaStream_t stream) -> cudaError_t {
            mega::Params p = q;
            p.in_planar = tx.planar ? 1 : 0; p.out_planar = to.planar ? 1 : 0;
            p.blob = reinterpret_cast<const uint8_t*>(P->arena + blob);
            p.nimg = n;
            cudaLaunchConfig_t cfg;
            std::memset(&cfg, 0, sizeof cfg);
            cfg.blockDim = dim3(mega::THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
            cudaLaunchAttribute at[2];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = mega::CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[1].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 2;
            if (*nclusters == 0) {   // co-resident clusters of this footprint (GPC boundaries matter): once
                cfg.gridDim = dim3(mega::CLUSTER * 64);
                int nc = 0;
                cudaError_t e = cudaOccupancyMaxActiveClusters(&nc, mega::mega_kernel, &cfg);
                if (e != cudaSuccess) return e;
                if (nc < 1) return cudaErrorLaunchOutOfResources;
                *nclusters = nc;
                if (getenv("ISG_FK_VERBOSE")) fprintf(stderr, "mega %s: %d co-resident clusters of %d, smem %zu\n", name.c_str(), nc, mega::CLUSTER, smem);
            }
            cfg.gridDim = dim3(mega::CLUSTER * std::min(*nclusters, std::max(n, 1)));
            static const bool prof_on = getenv("ISG_MEGA_PROF") != nullptr;
            if (prof_on) {   // debug: per-phase cycle sums of every CTA
                const int grid = cfg.gridDim.x;
                cudaMalloc(&p.prof, (size_t)grid * 8 * sizeof(long long));
                cudaMemset(p.prof, 0, (size_t)grid * 8 * sizeof(long long));
                cudaError_t e = cudaLaunchKernelEx(&cfg, mega::mega_kernel, maps->in[0], maps->out[0], p);
                cudaStreamSynchronize(stream);
                std::vector<long long> h((size_t)grid * 8);
                cudaMemcpy(h.data(), p.prof, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
                cudaFree(p.prof);
                double ph[8] = {0};
                for (int b = 0; b < grid; ++b) for (int k = 0; k < 8; ++k) ph[k] += (double)h[(size_t)b * 8 + k];
                const double steps = (double)n * nb * mega::CLUSTER;   // (image, block) steps summed over CTAs
                int n5 = 0;
                for (int k = 0; k < nb; ++k) n5 += p.blk[k].sep5;
                const double s3 = steps * (nb - n5) / nb, s5 = steps * n5 / nb;
                fprintf(stderr, "megaprof %-12s cycles per block and CTA: wgt+gemm0 %6.0f epi_m1 %6.0f halo-barrier %6.0f stencil 3x3 %6.0f "
                        "5x5 %6.0f gemm2 %6.0f epi_x %6.0f | x load per image %6.0f\n", name.c_str(), ph[1] / steps, ph[2] / steps,
                        ph[3] / steps, s3 > 0 ? ph[4] / s3 : 0.0, s5 > 0 ? ph[7] / s5 : 0.0, ph[5] / steps, ph[6] / steps,
                        ph[0] / ((double)n * mega::CLUSTER));
                return e;
            }
            return cudaLaunchKernelEx(&cfg, mega::mega_kernel, maps->in[0], maps->out[0], p);
        };
        P->ops.push_back(op);
        return out;
    }


    // ---- split-precision plans: a run of residual bottlenecks as ONE launch of 8-CTA clusters (csrc/mega3.cuh) ----
    struct Mega3Lay { mega3::Params q; size_t smem = 0; bool ok = false; };
    Mega3Lay mega3_layout(int H, int W, int C, int mid, int nblocks) const {
        Mega3Lay L;
        mega3::Params& q = L.q;
        std::memset(&q, 0, sizeof q);
        if (!P->split || getenv("ISG_FK_NO_MEGA") || nblocks > mega3::MAX_BLOCKS || H % mega3::CLUSTER) return L;
        q.H = H; q.W = W; q.R = H / mega3::CLUSTER;
        q.Cs = round_up(C, 16); q.Cm = round_up(mid, 16);
        q.npx = q.R * W;
        if (q.npx % 128 || q.R < mega3::HALO || W > 128 || q.R > 256) return L;
        q.n_mt = q.npx / 128;
        q.pitch = W + 2 * mega3::HALO; q.rows_p = q.R + 2 * mega3::HALO;
        q.w_magic = W > 1 ? (uint32_t)(((1ull << 32) + W - 1) / W) : 0u;
        q.nblocks = nblocks;
        const int cols = q.n_mt * std::max(q.Cs, q.Cm);
        if (cols > 512) return L;
        q.tmem_cols = tc::tmem_cols_pow2(cols);
        // global blob of one block
        q.w0_bytes = q.Cs * q.Cm * 2 * 2; q.w2_bytes = q.Cm * q.Cs * 2 * 2;   // hi + lo
        int rel = 0;
        q.v0_off = rel; rel = r128(rel + 2 * q.Cm * 4);
        q.dw_off = rel; rel = r128(rel + 25 * q.Cm * 4);
        q.dv_off = rel; rel = r128(rel + 7 * q.Cm * 4);
        q.v2_off = rel; rel = r128(rel + 2 * q.Cs * 4);
        q.c_bytes = rel;
        q.w0_goff = 0; q.w2_goff = r128(q.w0_bytes); q.c_goff = q.w2_goff + r128(q.w2_bytes);
        q.blob_stride = (q.c_goff + q.c_bytes + 255) / 256 * 256;
        // shared memory
        int off = 0;
        const int xh = (q.Cs / 8) * q.npx * 16, mh = (q.Cm / 8) * q.rows_p * q.pitch * 16, dh = (q.Cm / 8) * q.npx * 16;
        q.x_off = off; q.x_lo = r128(xh); off = r128(off + q.x_lo + xh);
        q.m1_off = off; q.m1_lo = 0; off = r128(off + 2 * mh);   // m1 as fp32 planes (32 bytes per pixel and chunk)
        q.d_off = off; q.d_lo = r128(dh); off = r128(off + q.d_lo + dh);
        q.wg_off = off; off = r128(off + std::max(q.w0_bytes, q.w2_bytes));
        q.wc_off[0] = off; off = r128(off + q.c_bytes);
        q.wc_off[1] = off; off = r128(off + q.c_bytes);
        q.bar_off = off; off += 128;
        L.smem = off;
        L.ok = L.smem <= 227 * 1024;
        return L;
    }
    int mega3_run(const std::string& name, int x, int mid, const std::vector<MegaBlock>& blocks, const std::string& outname) {
        isg_plan* P = this->P;
        const TensorD tx = td(x);
        const int C = tx.C;
        Mega3Lay L = mega3_layout(tx.H, tx.W, C, mid, (int)blocks.size());
        mega3::Params q = L.q;
        const int out = tensor(outname, C, tx.H, tx.W, true);
        struct Fills { StageFill f0, f2; StencilFill fs; };
        auto fills = std::make_shared<std::vector<Fills>>();
        for (size_t i = 0; i < blocks.size(); ++i) {
            const MegaBlock& b = blocks[i];
            q.blk[i].dil = b.five ? 1 : b.dil; q.blk[i].sep5 = b.five ? 1 : 0;
            const int d = q.blk[i].dil;
            q.blk[i].sr = 4 * d <= q.R ? 4 : (2 * d <= q.R ? 2 : 1);   // stencil strip height (rows `dil` apart) within the CTA's rows
            if (b.five && q.blk[i].sr == 1) q.blk[i].sr = 2;
            Next n0; n0.on = true; n0.unit = b.p + ".convs.0"; n0.mid = mid; n0.act = ACT_PRELU;
            Fills f;
            f.f0 = chain_stage(n0, C, q.Cs).fill;
            f.fs = b.five ? sep5_stencil(b.p, 0, mid).fill : dw3_stencil(b.p + ".convs.1", 0, mid, b.dil).fill;
            f.f2 = res_out_fill(b.p, b.five ? b.p + ".convs.3" : b.p + ".convs.2", C, mid, q.Cs);
            fills->push_back(f);
            macs(mid, C, 1, 1, tx.H, tx.W);
            if (b.five) { macs(mid, 1, 5, 1, tx.H, tx.W); macs(mid, 1, 1, 5, tx.H, tx.W); }
            else macs(mid, 1, 3, 3, tx.H, tx.W);
            macs(C, mid, 1, 1, tx.H, tx.W);
        }
        const size_t blob = arena((size_t)q.blob_stride * blocks.size() / 4);
        Op op;
        op.name = name + " (" + std::to_string(blocks.size()) + " residual bottlenecks, cluster-resident, split precision) [mega3]";
        const int nb = (int)blocks.size();
        op.pack = [=](const HostWeights& w, float* A) {
            // B[K][N] -> blocked [k16][2][N][8] bf16: W_hi blocks, then W_lo = bf16(w - W_hi) blocks `half` bytes later
            auto pack_b = [](uint8_t* dst, int half, const float* B, int K, int N) {
                uint16_t* hi = reinterpret_cast<uint16_t*>(dst);
                uint16_t* lo = reinterpret_cast<uint16_t*>(dst + half);
                for (int kk = 0; kk < K; ++kk)
                    for (int n = 0; n < N; ++n) {
                        const size_t idx = (size_t)(kk / 16) * N * 16 + (size_t)((kk % 16) / 8) * N * 8 + (size_t)n * 8 + (kk % 8);
                        const float v = B[(size_t)kk * N + n];
                        hi[idx] = tc::f2bf(v);
                        lo[idx] = tc::f2bf(v - tc::bf2f(hi[idx]));
                    }
            };
            for (int i = 0; i < nb; ++i) {
                uint8_t* base = reinterpret_cast<uint8_t*>(A + blob) + (size_t)i * q.blob_stride;
                std::memset(base, 0, q.blob_stride);
                const Fills& f = (*fills)[i];
                float* c = reinterpret_cast<float*>(base + q.c_goff);
                {
                    std::vector<float> B((size_t)q.Cs * q.Cm, 0.f), sh(q.Cm, 0.f), sl(q.Cm, 1.f);
                    f.f0(w, B.data(), sh.data(), sl.data());
                    pack_b(base + q.w0_goff, q.w0_bytes / 2, B.data(), q.Cs, q.Cm);
                    float* v = c + q.v0_off / 4;
                    std::memcpy(v, sh.data(), 4 * q.Cm); std::memcpy(v + q.Cm, sl.data(), 4 * q.Cm);
                }
                {
                    std::vector<float> wt((size_t)25 * q.Cm, 0.f), sh(q.Cm, 0.f), sl(q.Cm, 1.f), bw((size_t)5 * q.Cm, 0.f);
                    f.fs(w, wt.data(), sh.data(), sl.data(), bw.data());
                    std::memcpy(c + q.dw_off / 4, wt.data(), 4 * wt.size());
                    float* v = c + q.dv_off / 4;
                    std::memcpy(v, sh.data(), 4 * q.Cm); std::memcpy(v + q.Cm, sl.data(), 4 * q.Cm);
                    std::memcpy(v + 2 * q.Cm, bw.data(), 4 * bw.size());
                }
                {
                    std::vector<float> B((size_t)q.Cm * q.Cs, 0.f), sh(q.Cs, 0.f), sl(q.Cs, 1.f);
                    f.f2(w, B.data(), sh.data(), sl.data());
                    pack_b(base + q.w2_goff, q.w2_bytes / 2, B.data(), q.Cm, q.Cs);
                    float* v = c + q.v2_off / 4;
                    std::memcpy(v, sh.data(), 4 * q.Cs); std::memcpy(v + q.Cs, sl.data(), 4 * q.Cs);
                }
            }
        };
        auto maps = std::make_shared<FkMaps>();
        const int maxB = P->max_batch;
        const TensorD to = td(out);
        P->finalizers.push_back([=]() -> int {
            int rc = tc::make_act_map_chunks(&maps->in[0], P->ws + tx.off, maxB, tx.H, tx.W, tx.Ct(), tx.W, q.R, tx.Cs / 8, name.c_str(),
                                             tx.planar);
            if (rc) return rc;
            return tc::make_act_map_chunks(&maps->out[0], P->ws + to.off, maxB, to.H, to.W, to.Ct(), to.W, q.R, to.Cs / 8, name.c_str(),
                                           to.planar);
        });
        cudaFuncSetAttribute(mega3::mega3_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
        cudaFuncSetAttribute(mega3::mega3_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
        const size_t smem = L.smem;
        auto nclusters = std::make_shared<int>(0);
        op.launch = [=](int n, const float*, float*, cudaStream_t stream) -> cudaError_t {
            mega3::Params p = q;
            p.in_planar = tx.planar ? 1 : 0; p.out_planar = to.planar ? 1 : 0;
            p.blob = reinterpret_cast<const uint8_t*>(P->arena + blob);
            p.nimg = n;
            cudaLaunchConfig_t cfg;
            std::memset(&cfg, 0, sizeof cfg);
            cfg.blockDim = dim3(mega3::THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
            cudaLaunchAttribute at[2];
            at[0].id = cudaLaunchAttributeClusterDimension;
            at[0].val.clusterDim.x = mega3::CLUSTER; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            at[1].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = at; cfg.numAttrs = 2;
            if (*nclusters == 0) {   // co-resident clusters of this footprint: once
                cfg.gridDim = dim3(mega3::CLUSTER * 32);
                int nc = 0;
                cudaError_t e = cudaOccupancyMaxActiveClusters(&nc, mega3::mega3_kernel, &cfg);
                if (e != cudaSuccess) return e;
                if (nc < 1) return cudaErrorLaunchOutOfResources;
                *nclusters = nc;
                if (getenv("ISG_FK_VERBOSE")) fprintf(stderr, "mega3 %s: %d co-resident clusters of %d, smem %zu\n", name.c_str(), nc, mega3::CLUSTER, smem);
            }
            cfg.gridDim = dim3(mega3::CLUSTER * std::min(*nclusters, std::max(n, 1)));
            static const bool prof_on = getenv("ISG_MEGA_PROF") != nullptr;
            if (prof_on) {   // debug: per-phase cycle sums of every CTA
                const int grid = cfg.gridDim.x;
                cudaMalloc(&p.prof, (size_t)grid * 8 * sizeof(long long));
                cudaMemset(p.prof, 0, (size_t)grid * 8 * sizeof(long long));
                cudaError_t e = cudaLaunchKernelEx(&cfg, mega3::mega3_kernel, maps->in[0], maps->out[0], p);
                cudaStreamSynchronize(stream);
                std::vector<long long> h((size_t)grid * 8);
                cudaMemcpy(h.data(), p.prof, h.size() * sizeof(long long), cudaMemcpyDeviceToHost);
                cudaFree(p.prof);
                double ph[8] = {0};
                for (int b = 0; b < grid; ++b) for (int k = 0; k < 8; ++k) ph[k] += (double)h[(size_t)b * 8 + k];
                const double steps = (double)n * nb * mega3::CLUSTER;   // (image, block) steps summed over CTAs
                fprintf(stderr, "mega3prof %-10s cycles per block and CTA: wgt+gemm0 %6.0f epi_m1+push %6.0f halo-barrier %6.0f stencil %6.0f "
                        "w2+gemm2 %6.0f epi_x %6.0f | x load per image %6.0f | sum per block %6.0f\n", name.c_str(), ph[1] / steps,
                        ph[2] / steps, ph[3] / steps, ph[4] / steps, ph[5] / steps, ph[6] / steps, ph[0] / ((double)n * mega3::CLUSTER),
                        (ph[1] + ph[2] + ph[3] + ph[4] + ph[5] + ph[6]) / steps);
                {
                    int n5 = 0, n4 = 0;
                    for (int k = 0; k < nb; ++k) { n5 += p.blk[k].sep5; n4 += (!p.blk[k].sep5 && p.blk[k].dil == 4); }
                    const double per = (double)n * mega3::CLUSTER;
                    fprintf(stderr, "mega3prof %-10s stencil cycles per block by kind: dil 1/2 %6.0f | dil 4 (+ x load) %6.0f | 5x5 %6.0f\n", name.c_str(),
                            nb - n5 - n4 > 0 ? ph[4] / (per * (nb - n5 - n4)) : 0.0, n4 ? ph[0] / (per * n4) : 0.0, n5 ? ph[7] / (per * n5) : 0.0);
                }
                return e;
            }
            return cudaLaunchKernelEx(&cfg, mega3::mega3_kernel, maps->in[0], maps->out[0], p);
        };
        P->ops.push_back(op);
        return out;
    }

    // BottleneckDown2 (segment.py:114-150)
    int block_down(const std::string& p, int x, int mid, int cout, int* keep, const std::string& outname,
                   const std::string& keepname, const Next& nx) {
        const TensorD tx = td(x);
        const int h = tx.H / 2, w = tx.W / 2;
        // convs.0: 2x2 stride-2 conv as a GEMM over the space-to-depth view
        const int m1 = tensor(p + ".m1", mid, h, w);
        // residual_1 = max_pool2d(x) (segment.py:145), also the decoder's skip: pooled in the same launch
        // from the sub-pixel planes the 2x2 conv already holds in shared memory
        *keep = tensor(keepname, tx.C, h, w, true);
        P->tensors[*keep].perm = tx.perm;
        {
            Spec s;
            s.name = p + ".convs.0 (2x2 s2) + maxpool2";
            s.pool_tensor = *keep;
            s.qH = h; s.qW = w;
            LoadSpec l{x};
            l.s2d = true;
            s.loads.push_back(l);
            StageSpec st;
            st.parts = {Part{PART_LOAD, 0}};
            st.N = round_up(mid, 16);
            st.out_tensor = m1;
            macs(mid, tx.C, 2, 2, h, w);
            const int N = st.N, Cs = tx.Cs, Cin = tx.C;
            const std::string unit = p + ".convs.0";
            const TensorD txc = tx;
            st.fill = [=](const HostWeights& wv, float* B, float* shift, float* slope) {
                std::vector<float> sc(mid), sh(mid);
                fold_unit(wv, unit + ".conv", unit + ".bn", mid, sc.data(), sh.data());
                const float* Wt = wv.get(unit + ".conv.weight");  // [mid][Cin][2][2]
                for (int a = 0; a < 2; ++a)
                    for (int bb = 0; bb < 2; ++bb)
                        for (int sidx = 0; sidx < Cs; ++sidx) {
                            const int l = txc.logical_of(sidx);
                            if (l < 0) continue;
                            const int k = a * 2 * Cs + bb * Cs + sidx;
                            for (int o = 0; o < mid; ++o)
                                B[(size_t)k * N + o] = Wt[(((size_t)o * Cin + l) * 2 + a) * 2 + bb] * sc[o];
                        }
                for (int o = 0; o < mid; ++o) { shift[o] = sh[o]; slope[o] = wv.get(unit + ".act.weight")[o]; }
            };
            s.stages.push_back(st);
            emit(s);
        }
        // dw3x3 -> [convs.2 | convm] -> +, PReLU  (-> next block's first conv)
        const TensorD tk = td(*keep);
        Spec s;
        s.name = p + ".dw+convs.2+convm";
        s.qH = h; s.qW = w;
        s.loads.push_back(LoadSpec{m1, 0, -1, 1});
        s.loads.push_back(LoadSpec{*keep});
        s.st = dw3_stencil(p + ".convs.1", 0, mid, 1);
        macs(mid, 1, 3, 3, h, w);
        macs(cout, mid, 1, 1, h, w);
        macs(cout, tx.C, 1, 1, h, w);
        StageSpec st;
        st.parts = {Part{PART_STENCIL, -1}, Part{PART_LOAD, 1}};
        st.N = round_up(cout, 16);
        const int out = tensor(outname, cout, h, w, true);
        st.out_tensor = out;
        const int N = st.N, midp = round_up(mid, 16), Cin = tx.C;
        st.fill = [=](const HostWeights& wv, float* B, float* shift, float* slope) {
            std::vector<float> sc(cout), sh(cout), scm(cout), shm(cout);
            fold_unit(wv, p + ".convs.2.conv", p + ".convs.2.bn", cout, sc.data(), sh.data());
            fold_unit(wv, p + ".convm.0.conv", p + ".convm.0.bn", cout, scm.data(), shm.data());
            const float* W2 = wv.get(p + ".convs.2.conv.weight");  // [cout][mid]
            for (int k = 0; k < mid; ++k)
                for (int o = 0; o < cout; ++o) B[(size_t)k * N + o] = W2[(size_t)o * mid + k] * sc[o];
            put_pw(B, N, midp, tk, 0, tk.Cs / 8, wv.get(p + ".convm.0.conv.weight"), Cin, 0, scm.data(), cout, 0);
            for (int o = 0; o < cout; ++o) { shift[o] = sh[o] + shm[o]; slope[o] = wv.get(p + ".prelu.weight")[o]; }
        };
        s.stages.push_back(st);
        emit_chained(s, nx, cout, N, out, h, w);
        return out;
    }

    // BottleneckDim_Res (segment.py:153-209) on cat(x0, x1); m1 = convs.0 output
    int block_proj(const std::string& p, int x0, int x1, int m1, int mid, int cout, bool last_prelu, const Next& nx) {
        const TensorD t0 = td(x0), t1 = td(x1);
        Spec s;
        s.name = p + ".dw+convs.2+resconv";
        s.qH = t0.H; s.qW = t0.W;
        s.loads.push_back(LoadSpec{m1, 0, -1, 1});
        s.loads.push_back(LoadSpec{x0});
        s.loads.push_back(LoadSpec{x1});
        s.st = dw3_stencil(p + ".convs.1", 0, mid, 1);
        const int Cin = t0.C + t1.C;
        macs(mid, 1, 3, 3, t0.H, t0.W);
        macs(cout, mid, 1, 1, t0.H, t0.W);
        macs(cout, Cin, 1, 1, t0.H, t0.W);
        StageSpec st;
        st.parts = {Part{PART_STENCIL, -1}, Part{PART_LOAD, 1}, Part{PART_LOAD, 2}};
        st.N = round_up(cout, 16);
        const int out = tensor(p, cout, t0.H, t0.W, true);
        st.out_tensor = out;
        const int N = st.N, midp = round_up(mid, 16);
        st.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(cout), sh(cout), scr(cout), shr(cout);
            fold_unit(w, p + ".convs.2.conv", p + ".convs.2.bn", cout, sc.data(), sh.data());
            fold_unit(w, p + ".resconv.0.conv", p + ".resconv.0.bn", cout, scr.data(), shr.data());
            const float* W2 = w.get(p + ".convs.2.conv.weight");
            for (int k = 0; k < mid; ++k)
                for (int o = 0; o < cout; ++o) B[(size_t)k * N + o] = W2[(size_t)o * mid + k] * sc[o];
            const float* Wr = w.get(p + ".resconv.0.conv.weight");  // [cout][Cin]
            put_pw(B, N, midp, t0, 0, t0.Cs / 8, Wr, Cin, 0, scr.data(), cout, 0);
            put_pw(B, N, midp + t0.Cs, t1, 0, t1.Cs / 8, Wr, Cin, t0.C, scr.data(), cout, 0);
            for (int o = 0; o < cout; ++o) {
                shift[o] = sh[o] + shr[o];
                slope[o] = last_prelu ? w.get(p + ".prelu.weight")[o] : 0.f;
            }
        };
        s.stages.push_back(st);
        bool ksplit = false;
        if (P->split) {   // no tile at all, or only tiles that fill less than 3/4 of a 128-pixel M tile
            const std::vector<Lay> c = candidates(s);
            int best_px = 0;
            for (const Lay& l : c) best_px = std::max(best_px, l.TH * l.TW);
            ksplit = c.empty() || best_px < 96;
        }
        if (ksplit) {
            // the whole K (mid + C0 + C1 channels, hi + lo) and its weights do not fit one tile (bottle3_1: 304 x 128):
            // resconv runs split along K over the two concatenated inputs, r = W[:, :C0] . x0 + shift, r += W[:, C0:] . x1,
            // and the block launch takes r as its residual
            int r = pw_partial(p + ".resconv[x0]", p + ".resconv.0", x0, Cin, 0, cout, -1, true);
            r = pw_partial(p + ".resconv[x1]+", p + ".resconv.0", x1, Cin, t0.C, cout, r, false);
            Spec c;
            c.name = p + ".dw+convs.2+res";
            c.qH = t0.H; c.qW = t0.W;
            c.loads.push_back(LoadSpec{m1, 0, -1, 1});
            c.loads.push_back(LoadSpec{r});
            c.st = s.st;
            StageSpec st2;
            st2.parts = {Part{PART_STENCIL, -1}};
            st2.N = N;
            st2.res_load = 1;
            st2.out_tensor = out;
            st2.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
                std::vector<float> sc(cout), sh(cout);
                fold_unit(w, p + ".convs.2.conv", p + ".convs.2.bn", cout, sc.data(), sh.data());
                const float* W2 = w.get(p + ".convs.2.conv.weight");
                for (int k = 0; k < mid; ++k)
                    for (int o = 0; o < cout; ++o) B[(size_t)k * N + o] = W2[(size_t)o * mid + k] * sc[o];
                for (int o = 0; o < cout; ++o) { shift[o] = sh[o]; slope[o] = last_prelu ? w.get(p + ".prelu.weight")[o] : 0.f; }
            };
            c.stages.push_back(st2);
            emit_chained(c, nx, cout, N, out, t0.H, t0.W);
            return out;
        }
        emit_chained(s, nx, cout, N, out, t0.H, t0.W);
        return out;
    }

    // BottleneckDim(usePrelu=False) (segment.py:232-261): dense 3x3 then 1x1 + x, ReLU; m1 = convs.0 output
    int block_plain(const std::string& p, int x, int m1, int mid) {
        const TensorD tx = td(x);
        const int C = tx.C;
        Spec s;
        s.name = p + ".convs.1(3x3)+convs.2+res";
        s.qH = tx.H; s.qW = tx.W;
        s.loads.push_back(LoadSpec{m1, 0, -1, 1});
        s.loads.push_back(LoadSpec{x});
        macs(mid, mid, 3, 3, tx.H, tx.W);
        macs(C, mid, 1, 1, tx.H, tx.W);
        StageSpec a;
        a.parts = {Part{PART_LOAD, 0}};
        a.taps9 = true;
        a.N = round_up(mid, 16);
        a.store = false;
        const int Na = a.N, Ka = td(m1).Cs;
        a.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(mid), sh(mid);
            fold_unit(w, p + ".convs.1.conv", p + ".convs.1.bn", mid, sc.data(), sh.data());
            const float* Wt = w.get(p + ".convs.1.conv.weight");  // [mid][mid][3][3]
            for (int t = 0; t < 9; ++t)
                for (int k = 0; k < mid; ++k)
                    for (int o = 0; o < mid; ++o)
                        B[((size_t)t * Ka + k) * Na + o] = Wt[((size_t)o * mid + k) * 9 + t] * sc[o];
            for (int o = 0; o < mid; ++o) { shift[o] = sh[o]; slope[o] = 0.f; }
        };
        s.stages.push_back(a);
        StageSpec b;
        b.parts = {Part{PART_PREV, -1}};
        b.N = tx.Cs;
        b.res_load = 1;
        const int out = tensor(p, C, tx.H, tx.W, true);
        b.out_tensor = out;
        const int Nb = b.N;
        b.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(C), sh(C);
            fold_unit(w, p + ".convs.2.conv", p + ".convs.2.bn", C, sc.data(), sh.data());
            const float* Wt = w.get(p + ".convs.2.conv.weight");  // [C][mid]
            for (int k = 0; k < mid; ++k)
                for (int o = 0; o < C; ++o) B[(size_t)k * Nb + o] = Wt[(size_t)o * mid + k] * sc[o];
            for (int o = 0; o < C; ++o) { shift[o] = sh[o]; slope[o] = 0.f; }
        };
        s.stages.push_back(b);
        emit(s);
        return out;
    }

    // BottleneckUp_Res / _Other (segment.py:296-344)
    int block_up(const std::string& p, int x, int skip, int mid, int cout, const std::string& outname) {
        const TensorD tx = td(x), ts = td(skip);
        const int h = tx.H, w = tx.W;
        // (a) one GEMM over (x | skip): columns [0,16) = convs.0 (+BN, ReLU); columns [16,16+cout) =
        //     uppool.1(cat(conv2(x), skip)) evaluated at LOW resolution (nearest-x2 commutes with a 1x1 conv)
        const int u = tensor(p + ".u", 16 + cout, h, w);
        {
            Spec s;
            s.name = p + ".convs.0|uppool(conv2|skip)";
            s.qH = h; s.qW = w;
            s.loads.push_back(LoadSpec{x});
            s.loads.push_back(LoadSpec{skip});
            StageSpec st;
            st.parts = {Part{PART_LOAD, 0}, Part{PART_LOAD, 1}};
            st.N = 16 + round_up(cout, 16);
            st.out_tensor = u;
            macs(mid, tx.C, 1, 1, h, w);
            macs(cout, tx.C, 1, 1, h, w);
            macs(cout, cout + ts.C, 1, 1, 2 * h, 2 * w);
            const int N = st.N, Cx = tx.C, Cskip = ts.C;
            st.fill = [=](const HostWeights& wv, float* B, float* shift, float* slope) {
                std::vector<float> s0(mid), h0(mid), s2(cout), h2(cout);
                fold_unit(wv, p + ".convs.0.conv", p + ".convs.0.bn", mid, s0.data(), h0.data());
                fold_unit(wv, p + ".conv2.0.conv", p + ".conv2.0.bn", cout, s2.data(), h2.data());
                put_pw(B, N, 0, tx, 0, tx.Cs / 8, wv.get(p + ".convs.0.conv.weight"), Cx, 0, s0.data(), mid, 0);
                for (int o = 0; o < mid; ++o) { shift[o] = h0[o]; slope[o] = 0.f; }
                for (int o = mid; o < 16; ++o) slope[o] = 0.f;
                const float* Wu = wv.get(p + ".uppool.1.weight");   // [cout][cout + Cskip]
                const float* bu = wv.get(p + ".uppool.1.bias");
                const float* Wc = wv.get(p + ".conv2.0.conv.weight");  // [cout][Cx]
                const int cu = cout + Cskip;
                std::vector<float> comp((size_t)cout * Cx);
                for (int o = 0; o < cout; ++o)
                    for (int k = 0; k < Cx; ++k) {
                        double a = 0;
                        for (int j = 0; j < cout; ++j) a += (double)Wu[(size_t)o * cu + j] * s2[j] * Wc[(size_t)j * Cx + k];
                        comp[(size_t)o * Cx + k] = (float)a;
                    }
                put_pw(B, N, 0, tx, 0, tx.Cs / 8, comp.data(), Cx, 0, nullptr, cout, 16);
                put_pw(B, N, tx.Cs, ts, 0, ts.Cs / 8, Wu, cu, cout, nullptr, cout, 16);
                for (int o = 0; o < cout; ++o) {
                    double a = bu[o];
                    for (int j = 0; j < cout; ++j) a += (double)Wu[(size_t)o * cu + j] * h2[j];
                    shift[16 + o] = (float)a;
                    slope[16 + o] = 1.f;
                }
            };
            s.stages.push_back(st);
            emit(s);
        }
        // (b) ConvTranspose2d k4 s2 p1 (+BN, ReLU) as a 9-tap GEMM with 4 output phases in N, then the
        //     block-diagonal convs.4 (+BN), + low-res side branch, ReLU, pixel-shuffle store
        const int Cp1 = mid <= 8 ? 8 : 16, Cp2 = round_up(cout, 16);
        Spec s;
        s.name = p + ".convT4+convs.4+side";
        s.qH = h; s.qW = w;
        s.loads.push_back(LoadSpec{u, 0, 2, 1});
        s.loads.push_back(LoadSpec{u, 2, Cp2 / 8, 0});
        macs(mid, mid, 4, 4, h, w);
        macs(cout, mid, 1, 1, 2 * h, 2 * w);
        StageSpec a;
        a.parts = {Part{PART_LOAD, 0}};
        a.taps9 = true;
        a.N = 4 * Cp1;
        a.store = false;
        const int Na = a.N;
        a.fill = [=](const HostWeights& wv, float* B, float* shift, float* slope) {
            std::vector<float> sc(mid), sh(mid);
            fold_unit(wv, p + ".convs.1", p + ".convs.2", mid, sc.data(), sh.data());
            const float* Wt = wv.get(p + ".convs.1.weight");  // [mid(in)][mid(out)][4][4]
            for (int t = 0; t < 9; ++t) {
                const int dy = t / 3 - 1, dx = t % 3 - 1;
                for (int ph = 0; ph < 4; ++ph) {
                    const int ky = (ph >> 1) - 2 * dy + 1, kx = (ph & 1) - 2 * dx + 1;
                    if (ky < 0 || ky > 3 || kx < 0 || kx > 3) continue;
                    for (int k = 0; k < mid; ++k)
                        for (int c = 0; c < mid; ++c)
                            B[((size_t)t * 16 + k) * Na + ph * Cp1 + c] = Wt[(((size_t)k * mid + c) * 4 + ky) * 4 + kx] * sc[c];
                }
            }
            for (int ph = 0; ph < 4; ++ph)
                for (int c = 0; c < Cp1; ++c) {
                    shift[ph * Cp1 + c] = c < mid ? sh[c] : 0.f;
                    slope[ph * Cp1 + c] = 0.f;
                }
        };
        s.stages.push_back(a);
        StageSpec b;
        b.parts = {Part{PART_PREV, -1}};
        b.N = 4 * Cp2;
        b.res_load = 1; b.res_mod = Cp2;
        b.out_mode = fk::OUT_SHUF_BF16; b.shuf = 2; b.Cp = Cp2;
        const int out = tensor(outname, cout, 2 * h, 2 * w, true);
        b.out_tensor = out;
        const int Nb = b.N;
        b.fill = [=](const HostWeights& wv, float* B, float* shift, float* slope) {
            std::vector<float> sc(cout), sh(cout);
            fold_unit(wv, p + ".convs.4.conv", p + ".convs.4.bn", cout, sc.data(), sh.data());
            const float* Wt = wv.get(p + ".convs.4.conv.weight");  // [cout][mid]
            for (int ph = 0; ph < 4; ++ph)
                for (int c = 0; c < mid; ++c)
                    for (int o = 0; o < cout; ++o)
                        B[(size_t)(ph * Cp1 + c) * Nb + ph * Cp2 + o] = Wt[(size_t)o * mid + c] * sc[o];
            for (int ph = 0; ph < 4; ++ph)
                for (int o = 0; o < Cp2; ++o) {
                    shift[ph * Cp2 + o] = o < cout ? sh[o] : 0.f;
                    slope[ph * Cp2 + o] = 0.f;
                }
        };
        s.stages.push_back(b);
        emit(s);
        return out;
    }

    // 5x5 stride-2 pad-2 Conv wrapper as a 9-tap GEMM over a space-to-depth input (segment.py:23-26)
    // exact (split plans, isg_forward_u8 with bf16-exact centred pixels): the input planes hold v - 255*mean[c] as single
    // bf16 values (no lo planes, no A_lo pass) and the weights carry 1 / (255*std[c]); runs instead of the generic launch
    void stem_conv(const std::string& unit, int src, int cin, int out_tensor, int out_mode, int qH, int qW, bool exact = false) {
        isg_plan* P = this->P;
        Spec s;
        s.name = unit + (exact ? " (5x5 s2 via space-to-depth, exact uint8 input)" : " (5x5 s2 via space-to-depth)");
        s.qH = qH; s.qW = qW;
        s.loads.push_back(LoadSpec{src, 0, -1, 1});
        s.loads[0].hi_only = exact;
        StageSpec st;
        st.parts = {Part{PART_LOAD, 0}};
        st.taps9 = true;
        st.N = 16;
        if (exact) st.passes = 5;   // A_hi.W_hi + A_hi.W_lo
        st.out_mode = out_mode;
        st.out_tensor = out_tensor;
        if (!exact) macs(16, cin, 5, 5, qH, qW);
        const int K = td(src).Cs;
        // a K slice of 16 rows is one sub-pixel (ry, rx) of the space-to-depth input when cin == 16: the taps
        // that would need kernel row / column 5 do not exist (25 of the 36 slices carry weights)
        if (cin == 16)
            st.used = [](int t, int ks) { return 2 * (t / 3) + (ks >> 1) <= 4 && 2 * (t % 3) + (ks & 1) <= 4; };
        st.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            std::vector<float> sc(16), sh(16);
            fold_unit(w, unit + ".conv", unit + ".bn", 16, sc.data(), sh.data());
            const float* Wt = w.get(unit + ".conv.weight");  // [16][cin][5][5]
            for (int t = 0; t < 9; ++t)
                for (int ph = 0; ph < 4; ++ph) {
                    const int ky = 2 * (t / 3) + (ph >> 1), kx = 2 * (t % 3) + (ph & 1);
                    if (ky > 4 || kx > 4) continue;
                    for (int c = 0; c < cin; ++c) {
                        // exact: normalised = (v - 255 mean) / (255 std); the input planes hold v - 255 mean
                        const double in_scale = !exact ? 1.0 : (P->exact_std[c] != 0.f ? 1.0 / (255.0 * (double)P->exact_std[c]) : 0.0);
                        for (int o = 0; o < 16; ++o)
                            B[((size_t)t * K + ph * cin + c) * 16 + o] =
                                (float)((double)Wt[(((size_t)o * cin + c) * 5 + ky) * 5 + kx] * (double)sc[o] * in_scale);
                    }
                }
            for (int o = 0; o < 16; ++o) { shift[o] = sh[o]; slope[o] = w.get(unit + ".act.weight")[o]; }
        };
        s.stages.push_back(st);
        emit(s);
        if (exact) P->ops.back().when = 1;
    }

    // bottle6_1 (ConvTranspose2d k8 s4 p2, 16->4) composed with bottle6_2 (Conv2d 3x3 p1, 4->1):
    // no activation in between (segment.py:504-505), so the pair is one 16->1 transposed conv with a
    // 10x10 kernel, stride 4: a 9-tap GEMM with the 16 output phases in N + a border correction
    void tail(int y) {
        isg_plan* P = this->P;
        const TensorD ty = td(y);
        macs(4, 16, 8, 8, ty.H, ty.W);
        macs(1, 4, 3, 3, 4 * ty.H, 4 * ty.W);
        Spec s;
        s.name = "bottle6_1*bottle6_2 (composed k10 s4 transposed conv)";
        s.qH = ty.H; s.qW = ty.W;
        s.loads.push_back(LoadSpec{y, 0, -1, 1});
        StageSpec st;
        st.parts = {Part{PART_LOAD, 0}};
        st.taps9 = true;
        st.N = 16;
        st.out_mode = fk::OUT_SHUF4_F32;
        st.to_logits = true;
        if (getenv("ISG_X3_TAIL_PASSES")) st.passes = atoi(getenv("ISG_X3_TAIL_PASSES"));   // experiments: 1 / 3 / 5 / 7
        const int tail_passes = passes_of(st);
        st.fill = [=](const HostWeights& w, float* B, float* shift, float* slope) {
            const float* W1 = w.get("bottle6_1.weight");  // [16][4][8][8]
            const float* b1 = w.get("bottle6_1.bias");
            const float* W3 = w.get("bottle6_2.weight");  // [1][4][3][3]
            const float b3 = w.get("bottle6_2.bias")[0];
            std::vector<double> Wc(16 * 100, 0.0);        // [ci][ku][kv]
            for (int ci = 0; ci < 16; ++ci)
                for (int ku = 0; ku < 10; ++ku)
                    for (int kv = 0; kv < 10; ++kv) {
                        double a = 0;
                        for (int c = 0; c < 4; ++c)
                            for (int dy = -1; dy <= 1; ++dy)
                                for (int dx = -1; dx <= 1; ++dx) {
                                    const int ky = ku - 1 + dy, kx = kv - 1 + dx;
                                    if (ky < 0 || ky > 7 || kx < 0 || kx > 7) continue;
                                    a += (double)W3[c * 9 + (dy + 1) * 3 + dx + 1] * W1[(((size_t)ci * 4 + c) * 8 + ky) * 8 + kx];
                                }
                        Wc[(ci * 10 + ku) * 10 + kv] = a;
                    }
            for (int t = 0; t < 9; ++t) {
                const int dy = t / 3 - 1, dx = t % 3 - 1;
                for (int ph = 0; ph < 16; ++ph) {
                    const int ku = (ph >> 2) - 4 * dy + 3, kv = (ph & 3) - 4 * dx + 3;
                    if (ku < 0 || ku > 9 || kv < 0 || kv > 9) continue;
                    for (int ci = 0; ci < 16; ++ci) B[((size_t)t * 16 + ci) * 16 + ph] = (float)Wc[(ci * 10 + ku) * 10 + kv];
                }
            }
            double bias = b3;
            for (int c = 0; c < 4; ++c)
                for (int k = 0; k < 9; ++k) bias += (double)W3[c * 9 + k] * b1[c];
            for (int ph = 0; ph < 16; ++ph) { shift[ph] = (float)bias; slope[ph] = 1.f; }
        };
        s.stages.push_back(st);
        emit(s);
        // border correction: fp32 tables composed on the host
        const size_t og = arena(64 * 9 * 16), ogb = arena(12);
        Op op;
        op.name = "logits border correction";
        op.pack = [=](const HostWeights& w, float* A) {
            const float* W1 = w.get("bottle6_1.weight");   // [ci(16)][co(4)][8][8]
            const float* w3 = w.get("bottle6_2.weight");   // [1][c(4)][3][3]
            const float* b1 = w.get("bottle6_1.bias");
            for (int t = 0; t < 64; ++t)
                for (int d = 0; d < 9; ++d)
                    for (int ci = 0; ci < 16; ++ci) {
                        double a = 0;
                        for (int co = 0; co < 4; ++co) a += (double)W1[((size_t)ci * 4 + co) * 64 + t] * w3[co * 9 + d];
                        A[og + ((size_t)t * 9 + d) * 16 + ci] = (float)a;
                    }
            for (int d = 0; d < 9; ++d) {
                double a = 0;
                for (int co = 0; co < 4; ++co) a += (double)b1[co] * w3[co * 9 + d];
                A[ogb + d] = (float)a;
            }
        };
        op.launch = [=](int n, const float*, float* logits, cudaStream_t stream) {
            TailFixParams p;
            p.in = reinterpret_cast<const bf16*>(P->ws + ty.off); p.Hq = ty.H; p.Wq = ty.W; p.Cs = ty.Ct(); p.planar = ty.planar ? 1 : 0;
            p.lo = (ty.split && (tail_passes & 2)) ? 1 : 0;   // the same operand value the GEMM saw
            p.logits = logits; p.g = P->arena + og; p.gb = P->arena + ogb;
            const int H = 4 * ty.H, W = 4 * ty.W;
            p.total = n * (2 * W + 2 * (H - 2));
            return launch_pdl(tail_fix_kernel, cdiv(p.total, 128), 128, stream, p);
        };
        P->ops.push_back(op);
    }

    // Segment.forward (segment.py:466-508)
    void build() {
        const int H = P->H, W = P->W, cin = P->cin;
        // stem (init_head_s4, segment.py:19-31)
        const int x2 = tensor("x2", 4 * cin, H / 2, W / 2);
        const int s1 = tensor("stem1.s2d", 64, H / 4, W / 4);          // layer1 output, space-to-depth storage
        view("stem1", s1, 16, 16, H / 2, W / 2, true, 1);
        const int init = tensor("init_down", 16 + cin, H / 4, W / 4, true);
        {
            std::vector<int> perm(16 + cin);                            // logical: pooled input first (segment.py:31)
            for (int c = 0; c < cin; ++c) perm[c] = 16 + c;
            for (int c = 0; c < 16; ++c) perm[cin + c] = c;
            P->tensors[init].perm = perm;
        }
        emit_prep(x2, init);
        stem_conv("init_conv.layer1", x2, cin, s1, fk::OUT_S2D, H / 2, W / 2);
        if (P->split && cin == 3) {   // the two first-layer variants: exactly one of them runs (Op::when)
            P->ops.back().when = 2;
            stem_conv("init_conv.layer1", x2, cin, s1, fk::OUT_S2D, H / 2, W / 2, true);
        }
        stem_conv("init_conv.layer2", s1, 16, init, fk::OUT_STAGE, H / 4, W / 4);

        const int h8 = H / 8, w8 = W / 8, h16 = H / 16, w16 = W / 16;
        int keep1, keep2;
        const bool split = P->split;
        const bool mega8 = (split ? mega3_layout(h8, w8, 48, 16, 4).ok : mega_layout(h8, w8, 48, 16, 4).ok) && !getenv("ISG_FK_NO_MEGA8");
        Next nx = mega8 ? Next() : next_of("bottle1_x.0.convs.0", 16, ACT_PRELU, h8, w8);
        const int d1 = block_down("bottle1_1", init, 16, 48, &keep1, "bottle1_down", "keep1", nx);
        int y = d1, m1 = nx.out;
        if (mega8) {
            std::vector<MegaBlock> blocks;
            for (int i = 0; i < 4; ++i) blocks.push_back(MegaBlock{"bottle1_x." + std::to_string(i), 1, false});
            y = split ? mega3_run("bottle1_x", d1, 16, blocks, "bottle1_5") : mega_run("bottle1_x", d1, 16, blocks, "bottle1_5");
        }
        for (int i = 0; i < 4 && !mega8; ++i) {
            const std::string p = "bottle1_x." + std::to_string(i);
            Next n2;
            if (i < 3) n2 = next_of("bottle1_x." + std::to_string(i + 1) + ".convs.0", 16, ACT_PRELU, h8, w8);
            y = block_res(p, y, m1, 16, 1, false, n2, i == 3 ? "bottle1_5" : p, true);   // every block output is a test tap
            m1 = n2.out;
        }
        const bool mega16 = split ? mega3_layout(h16, w16, 128, 48, 5).ok : mega_layout(h16, w16, 128, 48, 5).ok;
        nx = mega16 ? Next() : next_of("bottle2_x.0.convs.0", 48, ACT_PRELU, h16, w16);
        const int d2 = block_down("bottle2_1", y, 16, 128, &keep2, "bottle2_down", "keep2", nx);
        const int dils[4] = {1, 2, 1, 4};
        auto wide = [&](const std::string& pre, int t, int m, const Next& after, const char* outname) {
            if (mega16) {
                std::vector<MegaBlock> blocks;
                for (int i = 0; i < 5; ++i) blocks.push_back(MegaBlock{pre + "." + std::to_string(i), i < 4 ? dils[i] : 1, i == 4});
                return split ? mega3_run(pre, t, 48, blocks, outname) : mega_run(pre, t, 48, blocks, outname);
            }
            for (int i = 0; i < 5; ++i) {
                const std::string p = pre + "." + std::to_string(i);
                Next n2 = after;
                if (i < 4) n2 = next_of(pre + "." + std::to_string(i + 1) + ".convs.0", 48, ACT_PRELU, h16, w16);
                t = block_res(p, t, m, 48, i < 4 ? dils[i] : 1, i == 4, n2, i == 4 ? outname : p, true);
                m = n2.out;
            }
            return t;
        };
        y = wide("bottle2_x", d2, nx.out, Next(), "bottle2_8");
        m1 = pw_unit("bottle3_1.convs.0", {y, d2}, 48, ACT_PRELU, "bottle3_1.m1");
        nx = mega16 ? Next() : next_of("bottle3_x.0.convs.0", 48, ACT_PRELU, h16, w16);
        y = block_proj("bottle3_1", y, d2, m1, 48, 128, true, nx);
        y = wide("bottle3_x", y, nx.out, Next(), "bottle3_8");
        y = block_up("bottle4_1up", y, keep2, 16, 48, "bottle4_1");
        m1 = pw_unit("bottle4_2.convs.0", {d1, y}, 16, ACT_PRELU, "bottle4_2.m1");
        nx = next_of("bottle4_3.convs.0", 16, ACT_RELU, h8, w8);
        y = block_proj("bottle4_2", d1, y, m1, 16, 48, false, nx);
        y = block_plain("bottle4_3", y, nx.out, 16);
        y = block_up("bottle5_1up", y, keep1, 4, 16, "bottle5_1");
        m1 = pw_unit("bottle5_2.convs.0", {y}, 4, ACT_RELU, "bottle5_2.m1");
        y = block_plain("bottle5_2", y, m1, 4);
        tail(y);
    }
};

}  // namespace

void build_fk_plan(isg_plan* p) { FkBuilder(p).build(); }

}  // namespace isg
