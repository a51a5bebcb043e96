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

constexpr int THREADS = 512;    // one CTA per SM; the 256-thread instantiation runs two CTAs per SM
constexpr int MAX_LOADS = 10;   // split precision: every tensor arrives as a hi and a lo box
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
    int planar;               // 1: the tensor is stored planar [B][Cs/8][H][W][8] (4-D map), see tc::tma_load_act
};

struct Stage {
    int N, ntaps, nks;
    int tap_off[MAX_MMA_TAPS];     // pixel offsets dy*pitch + dx
    uint32_t tap_mask[MAX_MMA_TAPS];   // bit ks set: K slice ks of this tap has weights (structural zeros are skipped)
    int nent;                      // MMAs per M tile and pass = set bits; the packed B blocks follow the same order
    int nent_tab;                  // split precision: entries of the (A, B) table = nent * number of passes
    uint64_t a_desc[MAX_KS];       // UMMA descriptor of K slice ks (16 channels = 2 planes) relative to the
                                   // dynamic shared memory base: the kernel adds (base + row offset) >> 4
    uint64_t b_desc;               // descriptor of the first packed B block [tap][ks][2][N][8] bf16
    uint32_t a_buf_mask;           // bit ks set: that K slice lives in the double-buffered load area
    // split precision (ISG_MODE_BF16X3): every operand element is a bf16 pair hi + lo (hi = bf16(v), lo = bf16(v - hi)).
    // passes: bit 0 = A_hi.W_hi, bit 1 = A_lo.W_hi, bit 2 = A_hi.W_lo accumulate into the same TMEM tile (0 = plain bf16)
    int passes;
    // nmerge (narrow stages, N <= 32, whose MMAs are bound by reading the 4 KB A tile, not by math): the weight blocks are
    // [W_hi | W_lo] side by side (2N columns), pass A_hi is ONE MMA of width 2N into TMEM columns [0,N) and [N,2N), pass
    // A_lo uses the first N columns of the same blocks; the epilogue adds the two column halves.  2 MMAs per K slice, not 3.
    int nmerge;
    uint32_t a_lo[MAX_KS];         // K slice ks: distance from its hi planes to its lo planes, 16-byte units
    uint32_t b_lo;                 // distance from the W_hi blocks to the W_lo blocks, 16-byte units
    int res_lo;                    // bytes from a residual hi plane to its lo plane (0: residual is plain bf16)
    int split_out;                 // 1: the output is written as hi + lo
    int stage_lo;                  // bytes from a staging hi plane to its lo plane
    int out_lo_chunks;             // 8-channel chunks from a channel's hi half to its lo half in the output tensor
    int Ct_out;                    // storage channels of the output tensor (direct-store modes)
    int stage_in_load;             // staging planes alias a load region (follow the load buffer)
    int vec_off;                   // fp32 shift[N], slope[N]
    int pitch;                     // pitch of the M index space (TW + 2*halo of the tap planes)
    uint32_t pitch_magic;          // ceil(2^32 / pitch) (0 when pitch == 1): m / pitch = umulhi(m, magic)
    int res_on, res_off, res_plane_bytes, res_mod;
    int out_mode;
    int stage_off, stage_plane_bytes, stage_compact;
    int store;                     // 1: TMA-store the staging planes through the stage's out map
    int store_chunk0;
    int out_planar;                // the output tensor (TMA store or direct stores) is stored planar
    void* gptr; int Cs_out, shuf, Cp;   // direct-store modes
    // 16-column chunk cc -> (cc / (res_mod/16), cc / (Cp/16)) by magic multiplication (0: divisor 1 or unused)
    uint32_t res_magic, cp_magic;
};

struct Stencil {
    int on;
    int in_off, in_pitch, in_plane_bytes, halo;
    int nchunks, dil;
    int w_off;                     // fp32 [tap][C]
    int vec_off;                   // fp32 shift[C], slope[C], then bw[5][C] when sep5
    int sep5;
    int out_off, out_plane_bytes;
    int in_lo, out_lo;             // split precision: bytes from a hi plane to its lo plane (input / output planes)
};

struct Params {
    int TH, TW, H, W, tiles_x, tiles_y;
    int n_mt, tmem_cols;
    uint32_t tw_magic, npix_magic; // ceil(2^32 / TW), ceil(2^32 / (TH*TW)) (0 when the divisor is 1)
    int nloads; Load loads[MAX_LOADS];
    uint32_t tx_bytes;             // all TMA boxes of one tile
    int lbuf_stride;               // byte distance between the two load buffers (persistent double buffering)
    int total_tiles;               // batch * tiles_y * tiles_x of this launch
    int pipe_slots;                // > 0: warp-specialised single-stage launch (fk_pipe_kernel) with this many load slots
    int acc_stride;                // ... and two TMEM accumulators this many columns apart
    int pipe_groups;               // ... and this many epilogue warp groups (of 4 warps); the rest issue MMAs
    int split;                     // 1: split-precision launch (the *<true> kernel instantiations)
    int tab_n;                     // split: (A, B) pairs per (stage, load buffer) table
    int nlbuf;                     // load buffers of the generic loop: 2 = double buffered; 1 = the co-resident
                                   // CTA hides the load latency instead (two 256-thread CTAs per SM)
    int early_load;                // nlbuf == 1: the load buffer is dead once the stage-0 MMAs completed
    const void* blob; int blob_off, blob_bytes;
    Stencil st;
    // optional 2x2 max-pool of a folded space-to-depth load (BottleneckDown2's residual, segment.py:145):
    // max over the four sub-pixel plane groups -> pool planes -> TMA store through the second out map
    int pool_on, pool_in_off, pool_in_plane_bytes, pool_nchunks, pool_off, pool_plane_bytes, pool_planar;
    int pool_grp_planes;           // planes between the sub-pixel groups of the folded load (pool_nchunks, or twice that when split)
    int pool_lo_in, pool_lo_out, pool_lo_chunks;   // split precision: hi -> lo distances (bytes, bytes, output chunks)
    int nstages; Stage stage[2];
    int bar_off;                   // 3 mbarriers + tmem slot
    long long* prof;               // debug (ISG_FK_PROF=1): per-CTA phase time stamps, else null
};

// x / d for x * d < 2^32 with magic = ceil(2^32 / d); magic == 0 encodes d == 1
__device__ __forceinline__ int fastdiv(int x, uint32_t magic) {
    return magic ? (int)__umulhi((uint32_t)x, magic) : x;
}

__device__ __forceinline__ long long globaltimer_ns() {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define FK_STAMP(i) do { if (p.prof && threadIdx.x == 0) p.prof[(size_t)blockIdx.x * 16 + (i)] = clock64(); } while (0)
// steady state (tiles after the first): per-CTA sums of phase cycles in slots 9..13
#define FK_ACC(i) do { if (p.prof && threadIdx.x == 0 && it > 0) { const long long t_ = clock64(); \
    p.prof[(size_t)blockIdx.x * 16 + (i)] += t_ - tprev; tprev = t_; } } while (0)

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

// packed fp32x2 FMA (Blackwell FFMA2): d = a * b + c on both halves
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b),
                       rc = *reinterpret_cast<unsigned long long*>(&c), rd;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    return *reinterpret_cast<float2*>(&rd);
}
// packed fp32x2 add (same IEEE result per half as a scalar add)
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long ra = *reinterpret_cast<unsigned long long*>(&a), rb = *reinterpret_cast<unsigned long long*>(&b), rd;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    return *reinterpret_cast<float2*>(&rd);
}
__device__ __forceinline__ float2 bf2_to_f2(uint32_t w) {
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}

// split precision: v = hi + lo with hi = bf16(v), lo = bf16(v - hi) (two channels per word)
__device__ __forceinline__ void split_bf16(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = pack_bf16(a, b);
    const float2 h = bf2_to_f2(hi);
    lo = pack_bf16(a - h.x, b - h.y);
}

// ---- depthwise stencil: haloed input planes -> A operand planes ----
// SEP5 = false: 3x3 with dilation; SEP5 = true: the composed (5,1)x(1,5) pair (25 taps).
// A thread owns one channel pair of one column and a strip of ST_ROWS output rows DIL apart, so every
// input row it loads feeds up to KS outputs from registers, and its KS*KS weight pairs live in registers.
// The four lanes of a pixel read the four 4-byte quarters of its 16-byte chunk: a warp's load is eight
// neighbouring columns = 128 contiguous bytes, one shared-memory wavefront.  Each output accumulates its
// taps in (ty, tx) order from zero, whatever the strip it falls in.
// Output rows [rb, TH) of the tile; ST_ROWS = rows per strip.
// SPLIT: input and output planes are hi + lo pairs (s.in_lo / s.out_lo bytes apart); the arithmetic is the same fp32.
// F32IN (with SPLIT): the input planes hold fp32 values as planes of FOUR channels, [4-channel group][row][col][4 x fp32]
// (16 bytes per pixel, s.in_plane_bytes = one such plane) -- for an operand that only this stencil reads (mega3's m1)
// that is the same bytes as a hi + lo pair and a quarter of the instructions per tap (one 8-byte load instead of two
// loads, two unpacks and an add); 16-byte pixels keep the producer's 16-byte stores free of bank conflicts.
template <int KS, int DIL, int ST_ROWS, int NT, bool SPLIT = false, bool F32IN = false>
__device__ __forceinline__ void stencil_strips(uint8_t* smem, const Stencil& s, int rb, int TH, int TW, int Wimg, int x0, int boff) {
    constexpr int NW = NT / 32, R = KS / 2;
    const int C = s.nchunks * 8, pitch16 = s.in_pitch * 16;
    const float* wt = reinterpret_cast<const float*>(smem + s.w_off);
    const float* vec = reinterpret_cast<const float*>(smem + s.vec_off);
    const int lane = threadIdx.x & 31, sub = lane & 3, cl = lane >> 2;
    const int ncg = (TW + 7) >> 3;                         // groups of 8 columns
    const int ngroups = (TH - rb + ST_ROWS * DIL - 1) / (ST_ROWS * DIL);   // row groups; each holds DIL interleaved strips
    const int last_row = TH + 2 * s.halo - 1;
    // warp item = (chunk k, row group g, strip sd within the group, column group cg), cg fastest
    int cg = threadIdx.x >> 5, sd = 0, g = 0, k = 0;
    for (;;) {
        while (cg >= ncg) {
            cg -= ncg;
            if (++sd == DIL) { sd = 0; if (++g == ngroups) { g = 0; ++k; } }
        }
        if (k >= s.nchunks) break;
        const int c = cg * 8 + cl, cld = c < TW ? c : TW - 1;
        const int r0 = rb + g * (ST_ROWS * DIL) + sd;
        const int ch = k * 8 + sub * 2;
        float2 w[KS * KS];
#pragma unroll
        for (int t = 0; t < KS * KS; ++t) w[t] = *reinterpret_cast<const float2*>(wt + t * C + ch);
        const int row0 = r0 + s.halo - R * DIL;            // plane row of the first input row
        const uint8_t* in = F32IN
            ? smem + s.in_off + boff + (size_t)(2 * k + (sub >> 1)) * s.in_plane_bytes + (sub & 1) * 8 + (cld + s.halo - R * DIL) * 16
            : smem + s.in_off + boff + (size_t)k * s.in_plane_bytes + sub * 4 + (cld + s.halo - R * DIL) * 16;
        const int rmax = last_row - row0;                  // rows past the planes only feed outputs past the tile: clamp
        float2 acc[ST_ROWS];
#pragma unroll
        for (int j = 0; j < ST_ROWS; ++j) acc[j] = make_float2(0.f, 0.f);
#pragma unroll
        for (int iy = 0; iy < ST_ROWS + KS - 1; ++iy) {
            const uint8_t* row = in + (row0 + min(iy * DIL, rmax)) * pitch16;
            float2 v[KS];
#pragma unroll
            for (int tx = 0; tx < KS; ++tx) {
                if (F32IN) { v[tx] = *reinterpret_cast<const float2*>(row + tx * DIL * 16); continue; }
                v[tx] = bf2_to_f2(*reinterpret_cast<const uint32_t*>(row + tx * DIL * 16));
                if (SPLIT) {
                    const float2 l = bf2_to_f2(*reinterpret_cast<const uint32_t*>(row + s.in_lo + tx * DIL * 16));
                    v[tx].x += l.x; v[tx].y += l.y;
                }
            }
#pragma unroll
            for (int j = 0; j < ST_ROWS; ++j) {
                const int ty = iy - j;
                if (ty >= 0 && ty < KS) {
#pragma unroll
                    for (int tx = 0; tx < KS; ++tx) acc[j] = ffma2(v[tx], w[ty * KS + tx], acc[j]);
                }
            }
        }
        float2 sh = *reinterpret_cast<const float2*>(vec + ch);
        const float2 sl = *reinterpret_cast<const float2*>(vec + C + ch);
        if (KS == 5) {
            // bias of the bare (5,1) conv seen through the (1,5) conv: only taps inside the image
            const int xa = x0 + c;
#pragma unroll
            for (int kx = 0; kx < 5; ++kx) {
                const int xx = xa + kx - 2;
                if (xx >= 0 && xx < Wimg) {
                    const float2 b = *reinterpret_cast<const float2*>(vec + (2 + kx) * C + ch);
                    sh.x += b.x; sh.y += b.y;
                }
            }
        }
        uint8_t* out = smem + s.out_off + (size_t)k * s.out_plane_bytes + sub * 4 + (r0 * TW + c) * 16;
        const int nrows = c < TW ? TH - r0 : 0;            // outputs j with j * DIL < nrows exist
#pragma unroll
        for (int j = 0; j < ST_ROWS; ++j) {
            float a0 = acc[j].x + sh.x, a1 = acc[j].y + sh.y;
            a0 = a0 >= 0.f ? a0 : a0 * sl.x; a1 = a1 >= 0.f ? a1 : a1 * sl.y;
            if (SPLIT) {
                uint32_t oh, ol;
                split_bf16(a0, a1, oh, ol);
                if (j * DIL < nrows) {
                    *reinterpret_cast<uint32_t*>(out + j * DIL * TW * 16) = oh;
                    *reinterpret_cast<uint32_t*>(out + s.out_lo + j * DIL * TW * 16) = ol;
                }
            } else {
                const uint32_t o = pack_bf16(a0, a1);
                if (j * DIL < nrows) *reinterpret_cast<uint32_t*>(out + j * DIL * TW * 16) = o;
            }
        }
        cg += NW;
    }
}

// SEP5 = false: 3x3 with dilation 1 / 2 / 4 (segment.py:399-418); SEP5 = true: the composed 5x5
template <bool SEP5, int NT, bool SPLIT = false>
__device__ __forceinline__ void run_stencil(uint8_t* smem, const Stencil& s, int TH, int TW, int Wimg, int x0, int boff) {
    if (SEP5) stencil_strips<5, 1, 4, NT, SPLIT>(smem, s, 0, TH, TW, Wimg, x0, boff);
    else if (s.dil == 1) stencil_strips<3, 1, 4, NT, SPLIT>(smem, s, 0, TH, TW, Wimg, x0, boff);
    else if (s.dil == 2) stencil_strips<3, 2, 4, NT, SPLIT>(smem, s, 0, TH, TW, Wimg, x0, boff);
    else stencil_strips<3, 4, 4, NT, SPLIT>(smem, s, 0, TH, TW, Wimg, x0, boff);
}

__device__ __forceinline__ uint32_t bf2max(uint32_t a, uint32_t b) {
    __nv_bfloat162 r = __hmax2(*reinterpret_cast<__nv_bfloat162*>(&a), *reinterpret_cast<__nv_bfloat162*>(&b));
    return *reinterpret_cast<uint32_t*>(&r);
}
__device__ __forceinline__ uint4 bf8max(uint4 a, uint4 b) {
    return make_uint4(bf2max(a.x, b.x), bf2max(a.y, b.y), bf2max(a.z, b.z), bf2max(a.w, b.w));
}

// 2x2 max-pool from the four sub-pixel plane groups of a folded space-to-depth load
// value of 8 channels stored as a hi and a lo chunk (the sum is exact in fp32)
__device__ __forceinline__ void load_split8(const uint8_t* hi, int lo_off, float* f) {
    float l[8];
    unpack8(*reinterpret_cast<const uint4*>(hi), f);
    unpack8(*reinterpret_cast<const uint4*>(hi + lo_off), l);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] += l[i];
}

template <int NT, bool SPLIT = false>
__device__ __forceinline__ void run_pool(uint8_t* smem, const Params& p, int boff) {
    const int npix = p.TH * p.TW;
    const size_t grp = (size_t)p.pool_grp_planes * p.pool_in_plane_bytes;
    for (int item = threadIdx.x; item < npix * p.pool_nchunks; item += NT) {
        const int k = fastdiv(item, p.npix_magic), pix = item - k * npix;
        const uint8_t* src = smem + p.pool_in_off + boff + (size_t)k * p.pool_in_plane_bytes + (size_t)pix * 16;
        uint8_t* dst = smem + p.pool_off + (size_t)k * p.pool_plane_bytes + (size_t)pix * 16;
        if (SPLIT) {
            float m[8], f[8];
            load_split8(src, p.pool_lo_in, m);
#pragma unroll
            for (int g = 1; g < 4; ++g) {
                load_split8(src + g * grp, p.pool_lo_in, f);
#pragma unroll
                for (int i = 0; i < 8; ++i) m[i] = fmaxf(m[i], f[i]);
            }
            uint32_t h[4], l[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) split_bf16(m[2 * i], m[2 * i + 1], h[i], l[i]);
            *reinterpret_cast<uint4*>(dst) = make_uint4(h[0], h[1], h[2], h[3]);
            *reinterpret_cast<uint4*>(dst + p.pool_lo_out) = make_uint4(l[0], l[1], l[2], l[3]);
        } else {
            const uint4 a = *reinterpret_cast<const uint4*>(src), b = *reinterpret_cast<const uint4*>(src + grp);
            const uint4 c = *reinterpret_cast<const uint4*>(src + 2 * grp), d = *reinterpret_cast<const uint4*>(src + 3 * grp);
            *reinterpret_cast<uint4*>(dst) = bf8max(bf8max(a, b), bf8max(c, d));
        }
    }
}

constexpr int TAB_N = 40;          // A descriptors per (stage, load buffer): ntaps * nks <= 40
constexpr int TAB_N_SPLIT = 128;   // split precision: (A, B) descriptor pairs per table, passes * ntaps * nks <= 128

// A-descriptor table of one stage and one load buffer, built once per CTA in shared memory:
// entry (t * nks + ks) = descriptor of K slice ks shifted by conv tap t, absolute shared address.
// SPLIT: the table holds (A, B) descriptor pairs, one run of entries per pass of Stage::passes:
//   A_hi . W_hi, then A_lo . W_hi (the same weight blocks again), then A_hi . W_lo.
template <int NT, bool SPLIT = false>
__device__ __forceinline__ void build_desc_table(const Params& p, uint64_t* tabs, uint32_t smem_base) {
    const uint64_t base16 = smem_base >> 4;
    const int stride = SPLIT ? p.tab_n * 2 : TAB_N;
    for (int i = threadIdx.x; i < p.nstages * 3; i += NT) {   // one thread per (stage, load buffer) table
        const int k = i / 3, buf = i - 3 * k;
        const Stage& s = p.stage[k];
        uint64_t* tab = tabs + i * stride;
        int j = 0;
        for (int pass = 0; pass < (SPLIT ? 3 : 1); ++pass) {
            if (SPLIT && !((s.passes >> pass) & 1)) continue;
            if (SPLIT && s.nmerge && pass == 2) continue;      // W_lo rides in pass 0's wide blocks
            uint64_t bd = s.b_desc + base16 + (pass == 2 ? s.b_lo : 0u);
            for (int t = 0; t < s.ntaps; ++t)
                for (int ks = 0; ks < s.nks; ++ks) {
                    if (!((s.tap_mask[t] >> ks) & 1u)) continue;
                    uint64_t d = s.a_desc[ks] + base16 + (uint64_t)(int64_t)s.tap_off[t];
                    if ((s.a_buf_mask >> ks) & 1u) d += (uint64_t)((buf * p.lbuf_stride) >> 4);
                    if (SPLIT) {
                        if (pass == 1) d += s.a_lo[ks];
                        tab[2 * j] = d; tab[2 * j + 1] = bd;
                        bd += (uint32_t)s.N * (s.nmerge ? 4 : 2);
                        ++j;
                    } else {
                        tab[j++] = d;
                    }
                }
        }
    }
}

// MMAs of the M tiles mt = first, first + step, ...; called by ONE thread per issuing warp
template <bool SPLIT = false>
__device__ __forceinline__ void issue_mma(const Params& p, const Stage& s, const uint64_t* tab, uint32_t smem_base,
                                          uint32_t tmem_base, int first, int step) {
    const uint32_t idesc = tc::umma_idesc_bf16(128, s.N);
    if (SPLIT) {
        const int n = s.nent_tab;
        if (s.nmerge) {   // entries [0, nent): A_hi . [W_hi | W_lo] (width 2N); [nent, 2 nent): A_lo . W_hi (width N)
            const uint32_t idesc2 = tc::umma_idesc_bf16(128, 2 * s.N);
            const int nh = s.nent;
            for (int mt = first; mt < p.n_mt; mt += step) {
                const uint64_t row = (uint64_t)(mt * 128);
                const uint32_t d = tmem_base + mt * 2 * s.N;
                tc::umma_bf16(d, tab[0] + row, tab[1], idesc2, 0u);
#pragma unroll 4
                for (int j = 1; j < nh; ++j) tc::umma_bf16(d, tab[2 * j] + row, tab[2 * j + 1], idesc2, 1u);
#pragma unroll 4
                for (int j = nh; j < n; ++j) tc::umma_bf16(d, tab[2 * j] + row, tab[2 * j + 1], idesc, 1u);
            }
            return;
        }
        for (int mt = first; mt < p.n_mt; mt += step) {
            const uint64_t row = (uint64_t)(mt * 128);
            tc::umma_bf16(tmem_base + mt * s.N, tab[0] + row, tab[1], idesc, 0u);
#pragma unroll 4
            for (int j = 1; j < n; ++j) tc::umma_bf16(tmem_base + mt * s.N, tab[2 * j] + row, tab[2 * j + 1], idesc, 1u);
        }
        return;
    }
    const uint64_t bd0 = s.b_desc + (smem_base >> 4);
    const uint32_t b_step = (uint32_t)s.N * 2;  // 16-byte units per [2][N][8] block
    const int n = s.nent;
    for (int mt = first; mt < p.n_mt; mt += step) {
        const uint64_t row = (uint64_t)(mt * 128);
        uint64_t bd = bd0;
        tc::umma_bf16(tmem_base + mt * s.N, tab[0] + row, bd, idesc, 0u);
#pragma unroll 4
        for (int j = 1; j < n; ++j) {
            bd += b_step;
            tc::umma_bf16(tmem_base + mt * s.N, tab[j] + row, bd, idesc, 1u);
        }
    }
}

__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// wait for the TMEM loads in flight; the registers they fill pass through the statement so that no use
// of them can be scheduled before it
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&r)[16]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                   "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15])
                 :
                 : "memory");
}

// ---- epilogue of one stage: TMEM accumulators -> + shift (+ residual) -> activation -> output ----
// Warp w may only touch TMEM lanes [32*(w%4), +32): the NT/128 warp groups alternate over the
// (M tile, 16-column chunk) items.  Everything that depends on the tile only (image base pointers) is
// computed once per call, everything that depends on the M tile only (pixel geometry) once per M tile;
// the TMEM load of the next item is issued as soon as the current accumulators have been consumed.
struct EpiGeom { int r, c, y, x, pc, m; bool geom, inimg; };

__device__ __forceinline__ EpiGeom epi_geom(const Params& p, const Stage& s, int mt, int quad, int lane, int x0, int y0) {
    EpiGeom g;
    g.m = mt * 128 + quad * 32 + lane;
    g.r = fastdiv(g.m, s.pitch_magic); g.c = g.m - g.r * s.pitch;
    g.geom = g.r < p.TH && g.c < p.TW;
    g.y = y0 + g.r; g.x = x0 + g.c;
    g.inimg = g.geom && g.y < p.H && g.x < p.W;
    g.pc = g.r * p.TW + g.c;
    return g;
}

template <bool SPLIT = false>
__device__ __forceinline__ void run_epilogue(uint8_t* smem, const Params& p, const Stage& s, uint32_t tmem_base,
                                             int b, int x0, int y0, int boff, int grp, int NGROUPS) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = warp & 3;
    const int ncc = s.N >> 4, nitems = p.n_mt * ncc;
    int item = grp;
    if (item >= nitems) return;
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
    const float* vec = reinterpret_cast<const float*>(smem + s.vec_off);
    const int out_mode = s.out_mode, N = s.N;
    // per-tile bases: one 64-bit product per call, 32-bit offsets inside the image afterwards
    uint8_t* gimg = static_cast<uint8_t*>(s.gptr);
    if (out_mode == OUT_S2D) gimg += (size_t)b * (p.H >> 1) * (p.W >> 1) * (SPLIT ? s.Ct_out : 4 * N) * sizeof(bf16);
    else if (out_mode == OUT_SHUF_BF16) gimg += (size_t)b * p.H * p.W * (s.shuf * s.shuf * (SPLIT ? s.Ct_out : s.Cs_out)) * sizeof(bf16);
    else if (out_mode == OUT_SHUF4_F32) gimg += (size_t)b * p.H * p.W * 16 * sizeof(float);
    uint8_t* stage = smem + s.stage_off + (s.stage_in_load ? boff : 0);
    const uint8_t* res = smem + s.res_off + boff;
    int mt = 0, cc = item;
    while (cc >= ncc) { cc -= ncc; ++mt; }
    int mt_geom = -1;
    EpiGeom g = {};
    uint32_t ra[16];
    const bool merged = SPLIT && s.nmerge;      // accumulator columns of M tile mt: [mt*2N, +N) and [mt*2N + N, +N)
    auto tcol = [&](int it_) { return merged ? (it_ / ncc) * 2 * N + (it_ % ncc) * 16 : it_ * 16; };   // plain: mt*N + cc*16 = item*16
    __syncwarp();
    tmem_ld16_nowait(lane_base + tcol(item), ra);
    while (true) {
        tmem_ld_wait_dep(ra);
        if (mt != mt_geom) { g = epi_geom(p, s, mt, quad, lane, x0, y0); mt_geom = mt; }
        const int col0 = cc * 16;
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(ra[i]);
        if (merged) {   // + the A_hi . W_lo half
            __syncwarp();
            tmem_ld16_nowait(lane_base + tcol(item) + N, ra);
            tmem_ld_wait_dep(ra);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += __uint_as_float(ra[i]);
        }
        // the accumulators are in v: the next item's TMEM load overlaps the arithmetic below
        const int nxt = item + NGROUPS;
        __syncwarp();
        if (nxt < nitems) tmem_ld16_nowait(lane_base + tcol(nxt), ra);
        if (s.res_on && g.geom) {
            const int ch0 = s.res_mod ? col0 - fastdiv(cc, s.res_magic) * s.res_mod : col0;
            const uint8_t* rp = res + (size_t)(ch0 >> 3) * s.res_plane_bytes + (size_t)g.pc * 16;
            float f[16];
            unpack8(*reinterpret_cast<const uint4*>(rp), f);
            unpack8(*reinterpret_cast<const uint4*>(rp + s.res_plane_bytes), f + 8);
            if (SPLIT && s.res_lo) {   // residual value = hi + lo (exact in fp32), then one add as in the reference
                float l[16];
                unpack8(*reinterpret_cast<const uint4*>(rp + s.res_lo), l);
                unpack8(*reinterpret_cast<const uint4*>(rp + s.res_lo + s.res_plane_bytes), l + 8);
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] += l[i];
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += f[i];
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 sh = *reinterpret_cast<const float4*>(vec + col0 + 4 * j);
            const float4 sl = *reinterpret_cast<const float4*>(vec + N + col0 + 4 * j);
            float a;
            a = v[4 * j + 0] + sh.x; v[4 * j + 0] = a >= 0.f ? a : a * sl.x;
            a = v[4 * j + 1] + sh.y; v[4 * j + 1] = a >= 0.f ? a : a * sl.y;
            a = v[4 * j + 2] + sh.z; v[4 * j + 2] = a >= 0.f ? a : a * sl.z;
            a = v[4 * j + 3] + sh.w; v[4 * j + 3] = a >= 0.f ? a : a * sl.w;
        }
        if (out_mode == OUT_SHUF4_F32) {
            if (g.inimg) {   // 16 columns = the 4x4 sub-pixels of this q pixel, fp32 logits
                float* o = reinterpret_cast<float*>(gimg) + (4 * g.y) * (4 * p.W) + 4 * g.x;
#pragma unroll
                for (int ry = 0; ry < 4; ++ry)
                    *reinterpret_cast<float4*>(o + ry * 4 * p.W) = make_float4(v[4 * ry], v[4 * ry + 1], v[4 * ry + 2], v[4 * ry + 3]);
            }
        } else {
            uint4 q0, q1, l0, l1;
            if (SPLIT && s.split_out) {
                uint32_t h[8], l[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) split_bf16(v[2 * i], v[2 * i + 1], h[i], l[i]);
                q0 = make_uint4(h[0], h[1], h[2], h[3]); q1 = make_uint4(h[4], h[5], h[6], h[7]);
                l0 = make_uint4(l[0], l[1], l[2], l[3]); l1 = make_uint4(l[4], l[5], l[6], l[7]);
            } else {
                q0 = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
                q1 = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                l0 = l1 = make_uint4(0, 0, 0, 0);
            }
            if (out_mode == OUT_STAGE) {
                if (!s.stage_compact || g.geom) {
                    uint8_t* d = stage + (size_t)(col0 >> 3) * s.stage_plane_bytes + (size_t)(s.stage_compact ? g.pc : g.m) * 16;
                    *reinterpret_cast<uint4*>(d) = q0;
                    *reinterpret_cast<uint4*>(d + s.stage_plane_bytes) = q1;
                    if (SPLIT && s.split_out) {
                        *reinterpret_cast<uint4*>(d + s.stage_lo) = l0;
                        *reinterpret_cast<uint4*>(d + s.stage_lo + s.stage_plane_bytes) = l1;
                    }
                }
            } else if (g.inimg) {
                bf16* o;
                int plane = 8;   // elements between the two 8-channel halves: adjacent (pixel-major) or one plane apart
                if (out_mode == OUT_S2D) {
                    const int sc = ((g.y & 1) * 2 + (g.x & 1)) * N + col0, pix = (g.y >> 1) * (p.W >> 1) + (g.x >> 1);
                    if (s.out_planar) { plane = (p.H >> 1) * (p.W >> 1) * 8; o = reinterpret_cast<bf16*>(gimg) + (sc >> 3) * plane + pix * 8; }
                    else o = reinterpret_cast<bf16*>(gimg) + pix * (SPLIT ? s.Ct_out : 4 * N) + sc;
                } else {  // OUT_SHUF_BF16: column = phase * Cp + channel, phase = (py, px) of the x`shuf` upsampling
                    const int phase = fastdiv(cc, s.cp_magic), ch = col0 - phase * s.Cp;
                    const int py = phase >> 1, px = phase & 1;   // shuf == 2 (ConvT k4 s2)
                    const int pix = (g.y * s.shuf + py) * (p.W * s.shuf) + g.x * s.shuf + px;
                    if (s.out_planar) { plane = p.H * p.W * s.shuf * s.shuf * 8; o = reinterpret_cast<bf16*>(gimg) + (ch >> 3) * plane + pix * 8; }
                    else o = reinterpret_cast<bf16*>(gimg) + pix * (SPLIT ? s.Ct_out : s.Cs_out) + ch;
                }
                st_global_16(o, q0);
                st_global_16(o + plane, q1);
                if (SPLIT && s.split_out) {   // the lo half sits out_lo_chunks chunks after the hi half (planes or 8-channel groups)
                    o += (size_t)s.out_lo_chunks * plane;
                    st_global_16(o, l0);
                    st_global_16(o + plane, l1);
                }
            }
        }
        if (nxt >= nitems) break;
        item = nxt;
        cc += NGROUPS;
        while (cc >= ncc) { cc -= ncc; ++mt; }
    }
}

template <int NT, bool SPLIT = false>
__global__ void __launch_bounds__(NT, NT == 512 ? 1 : (NT == 256 ? 2 : 4))
fk_kernel(const __grid_constant__ CUtensorMap m0, const __grid_constant__ CUtensorMap m1,
          const __grid_constant__ CUtensorMap m2, const __grid_constant__ CUtensorMap m3,
          const __grid_constant__ CUtensorMap mo0, const __grid_constant__ CUtensorMap mo1,
          const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // bars: [0],[1] tile loads of buffer 0/1; [2],[3] MMAs of stage 0/1; [4] constants blob
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
    uint64_t* tabs = bars + 16;   // [stage][load buffer][TAB_N] A descriptors
    constexpr int NWARPS = NT / 32;
    const int tid = threadIdx.x, warp = tid >> 5;
    const CUtensorMap* maps[4] = {&m0, &m1, &m2, &m3};

    if (p.prof && tid == 0) p.prof[(size_t)blockIdx.x * 16 + 14] = globaltimer_ns();
    FK_STAMP(0);
    const int n_issuers = p.n_mt < NWARPS ? p.n_mt : NWARPS;   // warp w issues the MMAs of M tiles w, w+NWARPS, ...
    if (tid == 0) {
        tc::mbar_init(&bars[0], 1);
        tc::mbar_init(&bars[1], 1);
        tc::mbar_init(&bars[2], n_issuers);
        tc::mbar_init(&bars[3], n_issuers);
        tc::mbar_init(&bars[4], 1);
        tc::mbar_init(&bars[8], 1);
        tc::mbar_init(&bars[9], 1);
        tc::mbar_init(&bars[10], 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, p.tmem_cols);
    build_desc_table<NT, SPLIT>(p, tabs, tc::smem_u32(smem));
    const int tab_stride = SPLIT ? p.tab_n * 2 : TAB_N;
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t smem_base = tc::smem_u32(smem);
    FK_STAMP(1);

    // one elected thread: TMA boxes of tile `t` into load buffer `buf`
    auto issue_loads = [&](int t, int buf) {
        const int txi = t % p.tiles_x, r1 = t / p.tiles_x;
        const int tyi = r1 % p.tiles_y, bi = r1 / p.tiles_y;
        tc::mbar_expect_tx(&bars[buf], p.tx_bytes);
        for (int i = 0; i < p.nloads; ++i) {
            const Load& l = p.loads[i];
            const int cx = txi * p.TW * l.step + l.dx;
            const int cy = tyi * p.TH * l.step + l.dy + (l.folded ? bi * l.rows_per_img : 0);
            tc::tma_load_act(smem + l.smem_off + buf * p.lbuf_stride, maps[l.map], &bars[buf], l.planar, cx, cy, l.chunk0,
                             l.folded ? 0 : bi);
        }
    };
    // programmatic dependent launch: this grid may have been scheduled while its predecessor in the
    // stream still runs; everything above (barriers, TMEM, descriptor tables) and the constants blob do
    // not depend on it.  Activations are only touched after griddepcontrol.wait.
    if (tid == 0) {
        tc::mbar_expect_tx(&bars[4], p.blob_bytes);
        bulk_g2s(smem + p.blob_off, p.blob, p.blob_bytes, &bars[4]);
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (tid == 0) issue_loads(blockIdx.x, 0);
    tc::mbar_wait(&bars[4], 0);

    int it = 0;
    long long tprev = 0;
    const bool dbl = p.nlbuf == 2;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int buf = dbl ? (it & 1) : 0, boff = buf * p.lbuf_stride;
        const int txi = tile % p.tiles_x, r1 = tile / p.tiles_x;
        const int tyi = r1 % p.tiles_y, b = r1 / p.tiles_y;
        const int x0 = txi * p.TW, y0 = tyi * p.TH;
        const int next = tile + (int)gridDim.x;
        if (p.prof) tprev = clock64();
        if (tid == 0 && dbl) {
            // the other buffer (and the single-buffered staging planes) may still be read by the TMA
            // stores of the previous tile
            tc::tma_store_wait_read();
            if (next < p.total_tiles) issue_loads(next, buf ^ 1);
        }
        tc::mbar_wait(&bars[buf], (dbl ? (it >> 1) : it) & 1);
        if (it == 0) FK_STAMP(2);

        if (p.st.on) {
            if (p.st.sep5) run_stencil<true, NT, SPLIT>(smem, p.st, p.TH, p.TW, p.W, x0, boff);
            else run_stencil<false, NT, SPLIT>(smem, p.st, p.TH, p.TW, p.W, x0, boff);
            tc::fence_proxy_async();
            __syncthreads();
        }
        if (it == 0) FK_STAMP(3);
        FK_ACC(9);
        for (int k = 0; k < p.nstages; ++k) {
            const Stage& s = p.stage[k];
            if ((tid & 31) == 0 && warp < n_issuers) {
                tc::tc_fence_after();
                issue_mma<SPLIT>(p, s, tabs + (3 * k + buf) * tab_stride, smem_base, tmem_base, warp, NWARPS);
                tc::umma_commit(&bars[2 + k]);
            }
            __syncwarp();
            tc::mbar_wait(&bars[2 + k], it & 1);
            if (it == 0) FK_STAMP(4 + 2 * k);
            FK_ACC(10 + 2 * k);
            tc::tc_fence_after();
            if (k == 0 && p.pool_on) run_pool<NT, SPLIT>(smem, p, boff);   // before the load buffer may be refilled
            if (k == 0 && tid == 0 && !dbl && p.early_load && next < p.total_tiles) issue_loads(next, 0);
            run_epilogue<SPLIT>(smem, p, s, tmem_base, b, x0, y0, boff, warp >> 2, NT / 128);
            tc::fence_proxy_async();
            tc::tc_fence_before();
            __syncthreads();
            if (it == 0) FK_STAMP(5 + 2 * k);
            FK_ACC(11 + 2 * k);
            if (tid == 0 && s.out_mode == OUT_STAGE && s.store) {
                const uint8_t* src = smem + s.stage_off + (s.stage_in_load ? boff : 0);
                tc::tma_store_act(k ? &mo1 : &mo0, src, s.out_planar, x0, y0, s.store_chunk0, b);
                if (SPLIT && s.split_out)
                    tc::tma_store_act(k ? &mo1 : &mo0, src + s.stage_lo, s.out_planar, x0, y0, s.store_chunk0 + s.out_lo_chunks, b);
                tc::tma_store_commit();
            }
            if (k == 0 && tid == 0 && p.pool_on) {
                tc::tma_store_act(&mo1, smem + p.pool_off, p.pool_planar, x0, y0, 0, b);
                if (SPLIT) tc::tma_store_act(&mo1, smem + p.pool_off + p.pool_lo_out, p.pool_planar, x0, y0, p.pool_lo_chunks, b);
                tc::tma_store_commit();
            }
        }
        if (tid == 0 && !dbl) {
            // single load buffer: every reader of this tile is done (the __syncthreads above); staging
            // planes (which may alias the load buffer) must have left through their TMA stores
            tc::tma_store_wait_read();
            if (!p.early_load && next < p.total_tiles) issue_loads(next, 0);
        }
    }
    if (tid == 0) tc::tma_store_wait_read();
    FK_STAMP(8);
    if (p.prof && tid == 0) p.prof[(size_t)blockIdx.x * 16 + 15] = globaltimer_ns();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, p.tmem_cols);
    }
}

// ---- warp-specialised variant for single-stage launches without stencil / residual ----
// warps 0, 2, 3: MMA issuers (one lane each, M tiles interleaved -- a single thread cannot issue fast
// enough to keep the tensor pipe busy); warp 1: TMA producer (one lane); warps 4..15: epilogue (three
// groups of four TMEM lane quadrants).  Two TMEM accumulators and `pipe_slots` load slots: the tensor pipe works on
// tile i+1 (and TMA on tile i+2) while the CUDA cores are in the epilogue of tile i.
constexpr int PIPE_THREADS = 512;

__device__ __forceinline__ void epi_bar_sync(int nthreads) { asm volatile("bar.sync 1, %0;" ::"r"(nthreads) : "memory"); }

template <bool SPLIT = false>
__global__ void __launch_bounds__(PIPE_THREADS, 1)
fk_pipe_kernel(const __grid_constant__ CUtensorMap m0, const __grid_constant__ CUtensorMap m1,
               const __grid_constant__ CUtensorMap m2, const __grid_constant__ CUtensorMap m3,
               const __grid_constant__ CUtensorMap mo0, const __grid_constant__ CUtensorMap mo1,
               const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // bars: [0..2] slot full, [3..5] slot free, [6],[7] accumulator full, [8],[9] accumulator free, [10] blob
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);
    uint64_t* tabs = bars + 16;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const CUtensorMap* maps[4] = {&m0, &m1, &m2, &m3};
    (void)mo1;
    const Stage& s = p.stage[0];
    const int nslots = p.pipe_slots;
    // epilogue: the last 4 * pipe_groups warps; before them warp 1 produces, the others issue MMAs
    const int PIPE_EPI_WARPS = 4 * p.pipe_groups, PIPE_EPI_WARP0 = PIPE_THREADS / 32 - PIPE_EPI_WARPS;
    const int PIPE_ISSUERS = PIPE_EPI_WARP0 - 1;
    if (tid == 0) {
        for (int i = 0; i < 3; ++i) { tc::mbar_init(&bars[i], 1); tc::mbar_init(&bars[3 + i], PIPE_ISSUERS); }
        tc::mbar_init(&bars[6], PIPE_ISSUERS); tc::mbar_init(&bars[7], PIPE_ISSUERS);
        tc::mbar_init(&bars[8], PIPE_EPI_WARPS); tc::mbar_init(&bars[9], PIPE_EPI_WARPS);
        tc::mbar_init(&bars[10], 1);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, p.tmem_cols);
    build_desc_table<PIPE_THREADS, SPLIT>(p, tabs, tc::smem_u32(smem));
    const int tab_stride = SPLIT ? p.tab_n * 2 : TAB_N;
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t smem_base = tc::smem_u32(smem);
    if (tid == 0) {
        tc::mbar_expect_tx(&bars[10], p.blob_bytes);
        bulk_g2s(smem + p.blob_off, p.blob, p.blob_bytes, &bars[10]);
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int step = gridDim.x;

    if (warp == 1) {
        if (lane == 0) {   // ---- TMA producer ----
            int it = 0;
            for (int tile = blockIdx.x; tile < p.total_tiles; tile += step, ++it) {
                const int slot = it % nslots, use = it / nslots;
                if (use > 0) tc::mbar_wait_backoff(&bars[3 + slot], (use - 1) & 1);   // the MMAs that read this slot completed
                const int txi = tile % p.tiles_x, r1 = tile / p.tiles_x;
                const int tyi = r1 % p.tiles_y, bi = r1 / p.tiles_y;
                tc::mbar_expect_tx(&bars[slot], p.tx_bytes);
                for (int i = 0; i < p.nloads; ++i) {
                    const Load& l = p.loads[i];
                    const int cx = txi * p.TW * l.step + l.dx;
                    const int cy = tyi * p.TH * l.step + l.dy + (l.folded ? bi * l.rows_per_img : 0);
                    tc::tma_load_act(smem + l.smem_off + slot * p.lbuf_stride, maps[l.map], &bars[slot], l.planar, cx, cy, l.chunk0,
                                    l.folded ? 0 : bi);
                }
            }
        }
    } else if (warp < PIPE_EPI_WARP0) {
        if (lane == 0) {   // ---- MMA issuers: M tiles first, first + PIPE_ISSUERS, ... ----
            // (descriptors from the shared-memory table; a converged-warp variant with warp-uniform descriptor
            // arithmetic was measured slower: ~19 dependent uniform-datapath instructions per MMA)
            const int first = warp == 0 ? 0 : warp - 1;
            tc::mbar_wait(&bars[10], 0);
            int it = 0;
            for (int tile = blockIdx.x; tile < p.total_tiles; tile += step, ++it) {
                const int slot = it % nslots, use = it / nslots, acc = it & 1;
                tc::mbar_wait_backoff(&bars[slot], use & 1);
                if (it >= 2) tc::mbar_wait_backoff(&bars[8 + acc], ((it >> 1) - 1) & 1);   // the epilogue has drained this accumulator
                tc::tc_fence_after();
                issue_mma<SPLIT>(p, s, tabs + slot * tab_stride, smem_base, tmem_base + acc * p.acc_stride, first, PIPE_ISSUERS);
                tc::umma_commit(&bars[3 + slot]);
                tc::umma_commit(&bars[6 + acc]);
            }
        }
    } else if (warp >= PIPE_EPI_WARP0) {   // ---- epilogue ----
        tc::mbar_wait(&bars[10], 0);
        const bool stores = s.out_mode == OUT_STAGE && s.store;
        const bool leader = tid == PIPE_EPI_WARP0 * 32;
        int it = 0;
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += step, ++it) {
            const int acc = it & 1;
            const int txi = tile % p.tiles_x, r1 = tile / p.tiles_x;
            const int tyi = r1 % p.tiles_y, b = r1 / p.tiles_y;
            const int x0 = txi * p.TW, y0 = tyi * p.TH;
            tc::mbar_wait(&bars[6 + acc], (it >> 1) & 1);
            tc::tc_fence_after();
            if (stores) {   // the previous tile's TMA store has read the (single) staging planes
                if (leader) tc::tma_store_wait_read();
                epi_bar_sync(PIPE_EPI_WARPS * 32);
            }
            run_epilogue<SPLIT>(smem, p, s, tmem_base + acc * p.acc_stride, b, x0, y0, 0, (warp - PIPE_EPI_WARP0) >> 2, p.pipe_groups);
            tc::tc_fence_before();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&bars[8 + acc]);
            if (stores) {
                tc::fence_proxy_async();
                epi_bar_sync(PIPE_EPI_WARPS * 32);
                if (leader) {
                    tc::tma_store_act(&mo0, smem + s.stage_off, s.out_planar, x0, y0, s.store_chunk0, b);
                    if (SPLIT && s.split_out)
                        tc::tma_store_act(&mo0, smem + s.stage_off + s.stage_lo, s.out_planar, x0, y0, s.store_chunk0 + s.out_lo_chunks, b);
                    tc::tma_store_commit();
                }
            }
        }
        if (leader) tc::tma_store_wait_read();
    }
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, p.tmem_cols);
    }
}

static_assert(sizeof(Params) + 6 * sizeof(CUtensorMap) <= 4000, "kernel parameter space");

}  // namespace fk
}  // namespace isg
