// mega3 -- the split-precision (ISG_MODE_BF16X3) version of mega.cuh: a whole run of residual bottlenecks
// (Bottleneck3x3 / Bottleneck5x5, segment.py:52-111) in ONE launch with the activation resident in shared memory,
// every value a bf16 pair hi + lo and every GEMM three passes A_hi.W_hi + A_lo.W_hi + A_hi.W_lo into one TMEM tile.
//
// Twice the planes of mega.cuh do not fit a 4-CTA cluster, so a cluster of 8 CTAs owns one image; CTA `rank` owns
// R = H/8 rows.  The weights of the two GEMMs stream through ONE buffer (w0 of block k while GEMM0 runs, then w2 of
// block k under the m1 epilogue + halo exchange + stencil, then w0 of block k+1 under the x epilogue); the small
// per-block constants (shifts, slopes, depthwise taps) are double buffered.  Per block and CTA:
//   m1 = PReLU(BN(W0 . x))        GEMM (K = C, N = mid) -> hi/lo into the zero-haloed m1 planes, edge rows pushed into
//                                 the neighbours' halo rows over distributed shared memory
//   d  = PReLU(BN(dw(m1)))        depthwise stencil on hi + lo (fp32), re-split
//   x  = PReLU(BN(W2 . d) + x)    GEMM (K = mid, N = C), residual hi + lo from and result hi/lo into the resident x planes
#pragma once
#include "mega.cuh"

namespace isg {
namespace mega3 {

constexpr int THREADS = 512, NWARPS = THREADS / 32, NGROUPS = THREADS / 128;
constexpr int CLUSTER = 8;
constexpr int HALO = 4;
constexpr int MAX_BLOCKS = 5;

struct Block { int dil, sep5, sr; };   // sr: rows per stencil strip (4 / 2 / 1)

struct Params {
    int H, W, R;                  // rows / columns of the image at this resolution, rows per CTA (H = 8 R)
    int Cs, Cm;                   // logical channels of x (multiple of 16) and of m1 / d
    int npx, n_mt;                // R * W (multiple of 128), M tiles
    int pitch, rows_p;            // m1 plane: W + 2 HALO columns, R + 2 HALO rows
    uint32_t w_magic;
    int nimg, nblocks;
    int in_planar, out_planar;
    Block blk[MAX_BLOCKS];
    // per block in global memory: [w0 hi | w0 lo] [w2 hi | w2 lo] [consts: v0 | dw taps | dv | v2]
    const uint8_t* blob;
    int blob_stride, w0_goff, w2_goff, c_goff, w0_bytes, w2_bytes, c_bytes;
    int v0_off, dw_off, dv_off, v2_off;              // offsets inside the constants
    int x_off, x_lo, m1_off, m1_lo, d_off, d_lo;      // shared memory: planes and hi -> lo distances (bytes)
    int wg_off, wc_off[2], bar_off;
    int tmem_cols;
    long long* prof;              // debug (ISG_MEGA_PROF=1): per-CTA sums of phase cycles [8], else null
};
#define M3_ACC(i) do { if (p.prof && threadIdx.x == 0) { const long long t_ = clock64(); \
    p.prof[(size_t)blockIdx.x * 8 + (i)] += t_ - tprev; tprev = t_; } } while (0)

// three passes of one GEMM for the CTA's M tiles: A planes hi (+ a_lo for the lo planes), B blocks hi (+ b_lo)
__device__ __forceinline__ void issue_gemm3(uint32_t a_addr, uint32_t a_lo, int a_plane_bytes, uint32_t b_addr, uint32_t b_lo,
                                            int N, int nks, int n_mt, uint32_t tmem_base, int first) {
    const uint32_t idesc = tc::umma_idesc_bf16(128, N);
    for (int mt = first; mt < n_mt; mt += NWARPS) {
        uint32_t acc = 0u;
#pragma unroll 1
        for (int pass = 0; pass < 3; ++pass) {
            const uint32_t aa = a_addr + (pass == 1 ? a_lo : 0u) + mt * 128 * 16;
            const uint32_t bb = b_addr + (pass == 2 ? b_lo : 0u);
            for (int ks = 0; ks < nks; ++ks) {
                const uint64_t ad = tc::umma_smem_desc(aa + 2 * ks * a_plane_bytes, a_plane_bytes, 128);
                const uint64_t bd = tc::umma_smem_desc(bb + ks * (2 * N * 16), N * 16, 128);
                tc::umma_bf16(tmem_base + mt * N, ad, bd, idesc, acc);
                acc = 1u;
            }
        }
    }
}

// m1 = PReLU(acc + shift) -> fp32 into the zero-haloed m1 planes [4-channel group][row][col][4 x fp32] (+ the neighbours' halo rows).
// m1 is read by the depthwise stencil only, never by an MMA: fp32 costs the bytes of a hi + lo pair and no split.
__device__ __forceinline__ void epi_m1(uint8_t* smem, const Params& p, const uint8_t* wc, uint32_t tmem_base, int reach,
                                       uint32_t rank) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = warp & 3, grp = warp >> 2;
    const int ncc = p.Cm >> 4, nitems = p.n_mt * ncc;
    const float* vec = reinterpret_cast<const float*>(wc + p.v0_off);
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int plane_bytes = p.rows_p * p.pitch * 16;   // one plane = 4 channels, 16 bytes per pixel
    for (int item = grp; item < nitems; item += NGROUPS) {
        int mt = 0, cc = item;
        while (cc >= ncc) { cc -= ncc; ++mt; }
        uint32_t ra[16];
        __syncwarp();
        fk::tmem_ld16_nowait(lane_base + item * 16, ra);
        fk::tmem_ld_wait_dep(ra);
        const int px = mt * 128 + quad * 32 + lane;
        const int r = fk::fastdiv(px, p.w_magic), c = px - r * p.W;
        const int col0 = cc * 16;
        uint4 q[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 sh = *reinterpret_cast<const float4*>(vec + col0 + 4 * j);
            const float4 sl = *reinterpret_cast<const float4*>(vec + p.Cm + col0 + 4 * j);
            float a0 = __uint_as_float(ra[4 * j]) + sh.x, a1 = __uint_as_float(ra[4 * j + 1]) + sh.y;
            float a2 = __uint_as_float(ra[4 * j + 2]) + sh.z, a3 = __uint_as_float(ra[4 * j + 3]) + sh.w;
            a0 = a0 >= 0.f ? a0 : a0 * sl.x; a1 = a1 >= 0.f ? a1 : a1 * sl.y;
            a2 = a2 >= 0.f ? a2 : a2 * sl.z; a3 = a3 >= 0.f ? a3 : a3 * sl.w;
            q[j] = make_uint4(__float_as_uint(a0), __float_as_uint(a1), __float_as_uint(a2), __float_as_uint(a3));
        }
        if (px < p.npx) {
            // channels col0 + 4j .. col0 + 4j + 3 -> plane col0/4 + j
            uint8_t* d = smem + p.m1_off + (size_t)(col0 >> 2) * plane_bytes + (size_t)((r + HALO) * p.pitch + c + HALO) * 16;
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(d + j * plane_bytes) = q[j];
            // edge rows are the neighbours' halo: the CTA above sees my row r as its row R + r, the one below as r - R
            if (r < reach && rank > 0) {
                const uint32_t la = tc::smem_u32(d) + (uint32_t)(p.R * p.pitch * 16);
#pragma unroll
                for (int j = 0; j < 4; ++j) mega::st_remote_16(la + j * plane_bytes, rank - 1, q[j]);
            }
            if (r >= p.R - reach && rank < CLUSTER - 1) {
                const uint32_t la = tc::smem_u32(d) - (uint32_t)(p.R * p.pitch * 16);
#pragma unroll
                for (int j = 0; j < 4; ++j) mega::st_remote_16(la + j * plane_bytes, rank + 1, q[j]);
            }
        }
    }
}

// x = PReLU((acc + (x_hi + x_lo)) + shift), hi / lo in place in the resident x planes
__device__ __forceinline__ void epi_x(uint8_t* smem, const Params& p, const uint8_t* wc, uint32_t tmem_base) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = warp & 3, grp = warp >> 2;
    const int ncc = p.Cs >> 4, nitems = p.n_mt * ncc;
    const float* vec = reinterpret_cast<const float*>(wc + p.v2_off);
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16);
    const int plane_bytes = p.npx * 16;
    uint8_t* xbase = smem + p.x_off + (size_t)(quad * 32 + lane) * 16;
    for (int item = grp; item < nitems; item += NGROUPS) {
        int mt = 0, cc = item;
        while (cc >= ncc) { cc -= ncc; ++mt; }
        uint32_t ra[16];
        __syncwarp();
        fk::tmem_ld16_nowait(lane_base + item * 16, ra);
        uint8_t* xp = xbase + (size_t)(cc * 2) * plane_bytes + (size_t)mt * (128 * 16);
        float f[16], g[16];
        fk::unpack8(*reinterpret_cast<const uint4*>(xp), f);
        fk::unpack8(*reinterpret_cast<const uint4*>(xp + plane_bytes), f + 8);
        fk::unpack8(*reinterpret_cast<const uint4*>(xp + p.x_lo), g);
        fk::unpack8(*reinterpret_cast<const uint4*>(xp + p.x_lo + plane_bytes), g + 8);
        fk::tmem_ld_wait_dep(ra);
        const int col0 = cc * 16;
        uint32_t h[8], l[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float2 sh = *reinterpret_cast<const float2*>(vec + col0 + 2 * j);
            const float2 sl = *reinterpret_cast<const float2*>(vec + p.Cs + col0 + 2 * j);
            float a = (__uint_as_float(ra[2 * j]) + (f[2 * j] + g[2 * j])) + sh.x;
            float b = (__uint_as_float(ra[2 * j + 1]) + (f[2 * j + 1] + g[2 * j + 1])) + sh.y;
            a = a >= 0.f ? a : a * sl.x; b = b >= 0.f ? b : b * sl.y;
            fk::split_bf16(a, b, h[j], l[j]);
        }
        *reinterpret_cast<uint4*>(xp) = make_uint4(h[0], h[1], h[2], h[3]);
        *reinterpret_cast<uint4*>(xp + plane_bytes) = make_uint4(h[4], h[5], h[6], h[7]);
        *reinterpret_cast<uint4*>(xp + p.x_lo) = make_uint4(l[0], l[1], l[2], l[3]);
        *reinterpret_cast<uint4*>(xp + p.x_lo + plane_bytes) = make_uint4(l[4], l[5], l[6], l[7]);
    }
}

// depthwise stencil of the CTA's rows with the strip height the host chose for this block (split planes)
__device__ __forceinline__ void stencil(uint8_t* smem, const fk::Stencil& st, const Block& blk, int R, int W) {
    if (blk.sep5) {
        if (blk.sr == 4) fk::stencil_strips<5, 1, 4, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
        else fk::stencil_strips<5, 1, 2, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
    } else if (blk.dil == 1) {
        if (blk.sr == 4) fk::stencil_strips<3, 1, 4, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
        else fk::stencil_strips<3, 1, 2, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
    } else if (blk.dil == 2) {
        if (blk.sr == 4) fk::stencil_strips<3, 2, 4, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
        else fk::stencil_strips<3, 2, 2, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
    } else {
        if (blk.sr == 4) fk::stencil_strips<3, 4, 4, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
        else if (blk.sr == 2) fk::stencil_strips<3, 4, 2, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
        else fk::stencil_strips<3, 4, 1, THREADS, true, true>(smem, st, 0, R, W, W, 0, 0);
    }
}

__global__ void __launch_bounds__(THREADS, 1)
mega3_kernel(const __grid_constant__ CUtensorMap min, const __grid_constant__ CUtensorMap mout,
             const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // bars: [0] x tile, [1] GEMM weights (one buffer, phase flips per load), [2],[3] constants, [4] MMAs
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t rank = mega::cluster_rank();
    const int cluster_id = blockIdx.x / CLUSTER, nclusters = gridDim.x / CLUSTER;
    const int n_issuers = p.n_mt < NWARPS ? p.n_mt : NWARPS;
    if (tid == 0) {
        tc::mbar_init(&bars[0], 1);
        tc::mbar_init(&bars[1], 1);
        tc::mbar_init(&bars[2], 1);
        tc::mbar_init(&bars[3], 1);
        tc::mbar_init(&bars[4], n_issuers);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, p.tmem_cols);
    // m1 planes (fp32) start all zero: halo columns and the halo rows beyond the image stay zero for the launch
    {
        const int n16 = ((p.Cm >> 3) * p.rows_p * p.pitch * 32) >> 4;
        uint4* z = reinterpret_cast<uint4*>(smem + p.m1_off);
        for (int i = tid; i < n16; i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t smem_base = tc::smem_u32(smem);
    mega::cluster_arrive();   // no CTA may push into a neighbour that is still zeroing
    mega::cluster_wait();

    const int my_images = cluster_id < p.nimg ? (p.nimg - cluster_id + nclusters - 1) / nclusters : 0;
    const int total_seq = my_images * p.nblocks;     // (image, block) steps of this CTA
    // GEMM weights: load number 2 * seq = w0 of step seq, 2 * seq + 1 = its w2; all through the one buffer
    auto load_wg = [&](int n) {
        const int k = (n >> 1) % p.nblocks, second = n & 1;
        const int bytes = second ? p.w2_bytes : p.w0_bytes;
        tc::mbar_expect_tx(&bars[1], bytes);
        fk::bulk_g2s(smem + p.wg_off, p.blob + (size_t)k * p.blob_stride + (second ? p.w2_goff : p.w0_goff), bytes, &bars[1]);
    };
    auto load_wc = [&](int seq) {
        const int k = seq % p.nblocks, b = seq & 1;
        tc::mbar_expect_tx(&bars[2 + b], p.c_bytes);
        fk::bulk_g2s(smem + p.wc_off[b], p.blob + (size_t)k * p.blob_stride + p.c_goff, p.c_bytes, &bars[2 + b]);
    };
    if (tid == 0 && total_seq > 0) {
        load_wg(0);
        load_wc(0);
        if (total_seq > 1) load_wc(1);
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    const int x_half_bytes = (p.Cs >> 3) * p.npx * 16;
    int seq = 0, mma_n = 0, it = 0, wg_n = 0;
    long long tprev = p.prof ? clock64() : 0;
    bool pending2 = false;      // a cluster barrier arrival (m1 planes read) not yet waited for
    for (int img = cluster_id; img < p.nimg; img += nclusters, ++it) {
        if (tid == 0) {
            tc::tma_store_wait_read();       // the previous image's stores have read the x planes
            tc::mbar_expect_tx(&bars[0], 2 * x_half_bytes);
            tc::tma_load_act(smem + p.x_off, &min, &bars[0], p.in_planar, 0, (int)rank * p.R, 0, img);
            tc::tma_load_act(smem + p.x_off + p.x_lo, &min, &bars[0], p.in_planar, 0, (int)rank * p.R, p.Cs >> 3, img);
        }
        tc::mbar_wait(&bars[0], it & 1);
        M3_ACC(0);   // x load (+ previous store drained)
        for (int k = 0; k < p.nblocks; ++k, ++seq) {
            const int wbi = seq & 1;
            const uint8_t* wc = smem + p.wc_off[wbi];
            const Block blk = p.blk[k];
            tc::mbar_wait(&bars[2 + wbi], (seq >> 1) & 1);
            // ---- m1 = PReLU(W0 . x) ----
            tc::mbar_wait(&bars[1], wg_n & 1); ++wg_n;
            if ((tid & 31) == 0 && warp < n_issuers) {
                tc::tc_fence_after();
                issue_gemm3(smem_base + p.x_off, p.x_lo, p.npx * 16, smem_base + p.wg_off, p.w0_bytes >> 1, p.Cm, p.Cs >> 4, p.n_mt,
                            tmem_base, warp);
                tc::umma_commit(&bars[4]);
            }
            __syncwarp();
            tc::mbar_wait(&bars[4], mma_n & 1); ++mma_n;
            tc::tc_fence_after();
            M3_ACC(1);   // weights wait + GEMM0
            if (tid == 0) load_wg(2 * seq + 1);     // w0 has been consumed: w2 streams in under the epilogue + stencil
            if (pending2) { mega::cluster_wait(); pending2 = false; }   // neighbours have read their halo of the previous block
            epi_m1(smem, p, wc, tmem_base, blk.sep5 ? 2 : blk.dil, rank);
            tc::tc_fence_before();
            M3_ACC(2);   // m1 epilogue + pushes (+ wait for the neighbours' previous reads)
            mega::cluster_arrive();       // release: my rows (local and pushed) are written ...
            mega::cluster_wait();         // ... acquire: so are the halo rows the neighbours pushed to me
            M3_ACC(3);   // halo exchange barrier
            // ---- d = PReLU(dw(m1)) ----
            {
                fk::Stencil st;
                st.on = 1; st.in_off = p.m1_off; st.in_pitch = p.pitch; st.in_plane_bytes = p.rows_p * p.pitch * 16;   // 4-channel fp32 plane
                st.halo = HALO; st.nchunks = p.Cm >> 3; st.dil = blk.dil;
                st.w_off = p.wc_off[wbi] + p.dw_off; st.vec_off = p.wc_off[wbi] + p.dv_off; st.sep5 = blk.sep5;
                st.out_off = p.d_off; st.out_plane_bytes = p.npx * 16;
                st.in_lo = 0; st.out_lo = p.d_lo;
                stencil(smem, st, blk, p.R, p.W);
            }
            tc::fence_proxy_async();
            __syncthreads();
            M3_ACC(blk.sep5 ? 7 : (blk.dil == 4 ? 0 : 4));   // stencil (debug slots: 7 = 5x5, 0 = dilation 4 [shared with x load], 4 = dilation 1/2)
            mega::cluster_arrive_relaxed();   // my reads of the m1 planes are done; waited for before the next block's pushes
            pending2 = true;
            // ---- x = PReLU(W2 . d + x) ----
            tc::mbar_wait(&bars[1], wg_n & 1); ++wg_n;
            if ((tid & 31) == 0 && warp < n_issuers) {
                tc::tc_fence_after();
                issue_gemm3(smem_base + p.d_off, p.d_lo, p.npx * 16, smem_base + p.wg_off, p.w2_bytes >> 1, p.Cs, p.Cm >> 4, p.n_mt,
                            tmem_base, warp);
                tc::umma_commit(&bars[4]);
            }
            __syncwarp();
            tc::mbar_wait(&bars[4], mma_n & 1); ++mma_n;
            tc::tc_fence_after();
            M3_ACC(5);   // w2 wait + GEMM2
            if (tid == 0 && seq + 1 < total_seq) load_wg(2 * (seq + 1));   // next block's w0 under the x epilogue
            epi_x(smem, p, wc, tmem_base);
            tc::fence_proxy_async();   // x planes: next read by tcgen05.mma / the TMA stores
            tc::tc_fence_before();
            __syncthreads();
            M3_ACC(6);   // x epilogue
            if (tid == 0 && seq + 2 < total_seq) load_wc(seq + 2);   // this constants buffer is free
        }
        if (tid == 0) {
            tc::tma_store_act(&mout, smem + p.x_off, p.out_planar, 0, (int)rank * p.R, 0, img);
            tc::tma_store_act(&mout, smem + p.x_off + p.x_lo, p.out_planar, 0, (int)rank * p.R, p.Cs >> 3, img);
            tc::tma_store_commit();
        }
    }
    if (pending2) mega::cluster_wait();
    if (tid == 0) tc::tma_store_wait_read();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, p.tmem_cols);
    }
    mega::cluster_arrive();   // nobody leaves while a neighbour could still address its shared memory
    mega::cluster_wait();
}

}  // namespace mega3
}  // namespace isg
