// mega -- a whole run of residual bottlenecks (Bottleneck3x3 / Bottleneck5x5, segment.py:52-111) in ONE
// launch, the activation resident in shared memory across the run.
//
// A thread-block cluster of 4 CTAs owns one image; CTA `rank` owns R = H/4 rows (all W columns, all C
// channels) as bf16 A-operand planes [chunk][pixel][8ch].  Per block, per CTA:
//   m1 = PReLU(BN(W0 . x))          tcgen05 GEMM (K = C, N = mid) -> epilogue writes m1 into a zero-haloed
//                                   plane and PUSHES its edge rows into the neighbours' halo rows over
//                                   distributed shared memory (st.shared::cluster)
//   d  = PReLU(BN(dw(m1)))          depthwise 3x3 (dilation 1/2/4) or the composed (5,1)x(1,5) pair, CUDA cores
//   x  = PReLU(BN(W2 . d) + x)      tcgen05 GEMM (K = mid, N = C), residual read from and result written to
//                                   the resident x planes
// Only the run's input and output cross HBM (one TMA load / store per image and CTA); weights of block
// k+2 stream in while block k computes.  Two cluster barriers per block order the halo exchange.
#pragma once
#include "fk.cuh"

namespace isg {
namespace mega {

constexpr int THREADS = 512, NWARPS = THREADS / 32, NGROUPS = THREADS / 128;
constexpr int CLUSTER = 4;
constexpr int HALO = 4;           // rows / columns of zero (or neighbour) halo around m1: max dilation
constexpr int MAX_BLOCKS = 5;

struct Block { int dil, sep5, sr; };   // sr: output rows per stencil strip (4 / 2): 4 unless sr * dil would exceed the CTA's rows

struct Params {
    int H, W, R;                  // image rows / columns at this resolution, rows per CTA (H = 4 R)
    int Cs, Cm;                   // channels of x (multiple of 16), of m1 / d (multiple of 16)
    int npx, n_mt;                // R * W (multiple of 128), M tiles
    int pitch, rows_p;            // m1 plane: W + 2 HALO columns, R + 2 HALO rows
    uint32_t w_magic, npx_magic;  // fastdiv by W, by npx
    int nimg, nblocks;
    int in_planar, out_planar;    // global layout of the run's input / output tensor (tc::tma_load_act)
    Block blk[MAX_BLOCKS];
    const uint8_t* blob;          // per block: [W0 | shift0,slope0 | dw taps | dw shift,slope(,bw) | W2 | shift2,slope2]
    int blob_stride, blob_bytes;
    int w0_off, v0_off, dw_off, dv_off, w2_off, v2_off;   // byte offsets inside one block's blob
    int x_off, m1_off, d_off, wbuf_off[2], bar_off;        // shared memory layout
    int tmem_cols;
    long long* prof;              // debug (ISG_MEGA_PROF=1): per-CTA sums of phase cycles [8], else null
};
#define MEGA_ACC(i) do { if (p.prof && threadIdx.x == 0) { const long long t_ = clock64(); \
    p.prof[(size_t)blockIdx.x * 8 + (i)] += t_ - tprev; tprev = t_; } } while (0)

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// 16-byte store into the shared memory of CTA `rank` of the cluster at the same offset as local address `laddr`
__device__ __forceinline__ void st_remote_16(uint32_t laddr, uint32_t rank, uint4 v) {
    uint32_t ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(laddr), "r"(rank));
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(ra), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// MMAs D[mt] (+)= A(ks) . B(ks) for the CTA's M tiles: A = compact planes [chunk][pixel][8], K slice ks = chunks
// 2ks, 2ks+1; B = blocked [ks][2][N][8].  One lane per issuing warp; warp w takes M tiles w, w+NWARPS, ...
__device__ __forceinline__ void issue_gemm(uint32_t a_addr, int a_plane_bytes, uint32_t b_addr, int N, int nks, int n_mt,
                                           uint32_t tmem_base, int first) {
    const uint32_t idesc = tc::umma_idesc_bf16(128, N);
    for (int mt = first; mt < n_mt; mt += NWARPS) {
        for (int ks = 0; ks < nks; ++ks) {
            const uint64_t ad = tc::umma_smem_desc(a_addr + 2 * ks * a_plane_bytes + mt * 128 * 16, a_plane_bytes, 128);
            const uint64_t bd = tc::umma_smem_desc(b_addr + ks * (2 * N * 16), N * 16, 128);
            tc::umma_bf16(tmem_base + mt * N, ad, bd, idesc, ks ? 1u : 0u);
        }
    }
}

// ---- epilogue of the first GEMM: m1 = PReLU(acc + shift) -> zero-haloed m1 planes (+ neighbours' halo rows) ----
// (M tiles mt0 .. mt0 + nmt - 1)
__device__ __forceinline__ void epi_m1(uint8_t* smem, const Params& p, const uint8_t* wb, uint32_t tmem_base, int reach,
                                       uint32_t rank, int mt0, int nmt) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = warp & 3, grp = warp >> 2;
    const int ncc = p.Cm >> 4, nitems = nmt * ncc;
    const float* vec = reinterpret_cast<const float*>(wb + p.v0_off);
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16) + mt0 * p.Cm;
    const int plane_bytes = p.rows_p * p.pitch * 16;
    for (int item = grp; item < nitems; item += NGROUPS) {
        int mt = mt0, cc = item;
        while (cc >= ncc) { cc -= ncc; ++mt; }
        uint32_t ra[16];
        __syncwarp();
        fk::tmem_ld16_nowait(lane_base + item * 16, ra);
        fk::tmem_ld_wait_dep(ra);
        const int px = mt * 128 + quad * 32 + lane;
        const int r = fk::fastdiv(px, p.w_magic), c = px - r * p.W;
        const int col0 = cc * 16;
        float v[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 sh = *reinterpret_cast<const float4*>(vec + col0 + 4 * j);
            const float4 sl = *reinterpret_cast<const float4*>(vec + p.Cm + col0 + 4 * j);
            const float2 a = fk::fadd2(make_float2(__uint_as_float(ra[4 * j]), __uint_as_float(ra[4 * j + 1])), make_float2(sh.x, sh.y));
            const float2 b = fk::fadd2(make_float2(__uint_as_float(ra[4 * j + 2]), __uint_as_float(ra[4 * j + 3])), make_float2(sh.z, sh.w));
            v[4 * j + 0] = a.x >= 0.f ? a.x : a.x * sl.x; v[4 * j + 1] = a.y >= 0.f ? a.y : a.y * sl.y;
            v[4 * j + 2] = b.x >= 0.f ? b.x : b.x * sl.z; v[4 * j + 3] = b.y >= 0.f ? b.y : b.y * sl.w;
        }
        const uint4 q0 = make_uint4(fk::pack_bf16(v[0], v[1]), fk::pack_bf16(v[2], v[3]), fk::pack_bf16(v[4], v[5]),
                                    fk::pack_bf16(v[6], v[7]));
        const uint4 q1 = make_uint4(fk::pack_bf16(v[8], v[9]), fk::pack_bf16(v[10], v[11]), fk::pack_bf16(v[12], v[13]),
                                    fk::pack_bf16(v[14], v[15]));
        if (px < p.npx) {
            uint8_t* d = smem + p.m1_off + (size_t)(col0 >> 3) * plane_bytes + (size_t)((r + HALO) * p.pitch + c + HALO) * 16;
            *reinterpret_cast<uint4*>(d) = q0;
            *reinterpret_cast<uint4*>(d + plane_bytes) = q1;
            // edge rows are the neighbours' halo: the CTA above sees my row r as its row R + r, the one below as r - R
            if (r < reach && rank > 0) {
                const uint32_t la = tc::smem_u32(d) + (uint32_t)(p.R * p.pitch * 16);
                st_remote_16(la, rank - 1, q0);
                st_remote_16(la + plane_bytes, rank - 1, q1);
            }
            if (r >= p.R - reach && rank < CLUSTER - 1) {
                const uint32_t la = tc::smem_u32(d) - (uint32_t)(p.R * p.pitch * 16);
                st_remote_16(la, rank + 1, q0);
                st_remote_16(la + plane_bytes, rank + 1, q1);
            }
        }
    }
}

// ---- epilogue of the second GEMM: x = PReLU(acc + shift + x), in place in the resident x planes ----
// one item = 16 accumulator columns of one pixel per lane
__device__ __forceinline__ void epi_x_item(uint8_t* xp, int plane_bytes, const float* vec, int Cs, int col0,
                                           const uint32_t (&ra)[16]) {
    const uint4 q0 = *reinterpret_cast<const uint4*>(xp), q1 = *reinterpret_cast<const uint4*>(xp + plane_bytes);
    const uint32_t w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
    uint32_t o[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const float4 sh = *reinterpret_cast<const float4*>(vec + col0 + 4 * j);
        const float4 sl = *reinterpret_cast<const float4*>(vec + Cs + col0 + 4 * j);
        // (acc + x) + shift, two columns per instruction
        float2 a = fk::fadd2(fk::fadd2(make_float2(__uint_as_float(ra[4 * j]), __uint_as_float(ra[4 * j + 1])),
                                       fk::bf2_to_f2(w[2 * j])), make_float2(sh.x, sh.y));
        float2 b = fk::fadd2(fk::fadd2(make_float2(__uint_as_float(ra[4 * j + 2]), __uint_as_float(ra[4 * j + 3])),
                                       fk::bf2_to_f2(w[2 * j + 1])), make_float2(sh.z, sh.w));
        a.x = a.x >= 0.f ? a.x : a.x * sl.x; a.y = a.y >= 0.f ? a.y : a.y * sl.y;
        b.x = b.x >= 0.f ? b.x : b.x * sl.z; b.y = b.y >= 0.f ? b.y : b.y * sl.w;
        o[2 * j] = fk::pack_bf16(a.x, a.y); o[2 * j + 1] = fk::pack_bf16(b.x, b.y);
    }
    *reinterpret_cast<uint4*>(xp) = make_uint4(o[0], o[1], o[2], o[3]);
    *reinterpret_cast<uint4*>(xp + plane_bytes) = make_uint4(o[4], o[5], o[6], o[7]);
}

// (M tiles mt0 .. mt0 + nmt - 1)  Two register sets alternate so that the TMEM load of the next item is in
// flight during the arithmetic of the current one without copying the accumulators.
__device__ __forceinline__ void epi_x(uint8_t* smem, const Params& p, const uint8_t* wb, uint32_t tmem_base, int mt0, int nmt) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, quad = warp & 3, grp = warp >> 2;
    const int ncc = p.Cs >> 4, nitems = nmt * ncc;
    const float* vec = reinterpret_cast<const float*>(wb + p.v2_off);
    const uint32_t lane_base = tmem_base + ((uint32_t)(quad * 32) << 16) + mt0 * p.Cs;
    const int plane_bytes = p.npx * 16;
    int item = grp;
    if (item >= nitems) return;
    int mt = mt0, cc = item;
    while (cc >= ncc) { cc -= ncc; ++mt; }
    uint8_t* xbase = smem + p.x_off + (size_t)(quad * 32 + lane) * 16;
    auto xptr = [&]() { return xbase + (size_t)(cc * 2) * plane_bytes + (size_t)mt * (128 * 16); };
    auto advance = [&]() { cc += NGROUPS; while (cc >= ncc) { cc -= ncc; ++mt; } };
    uint32_t ra[16], rb[16];
    __syncwarp();
    fk::tmem_ld16_nowait(lane_base + item * 16, ra);
    while (true) {
        fk::tmem_ld_wait_dep(ra);
        int nxt = item + NGROUPS;
        __syncwarp();
        if (nxt < nitems) fk::tmem_ld16_nowait(lane_base + nxt * 16, rb);
        epi_x_item(xptr(), plane_bytes, vec, p.Cs, cc * 16, ra);
        if (nxt >= nitems) break;
        advance();
        fk::tmem_ld_wait_dep(rb);
        item = nxt + NGROUPS;
        __syncwarp();
        if (item < nitems) fk::tmem_ld16_nowait(lane_base + item * 16, ra);
        epi_x_item(xptr(), plane_bytes, vec, p.Cs, cc * 16, rb);
        if (item >= nitems) break;
        advance();
    }
}

// depthwise stencil of output rows [rb, re) with the strip height the host chose for this block
__device__ __forceinline__ void stencil(uint8_t* smem, const fk::Stencil& st, const Block& blk, int rb, int re, int W) {
    if (blk.sep5) {
        if (blk.sr == 4) fk::stencil_strips<5, 1, 4, THREADS>(smem, st, rb, re, W, W, 0, 0);
        else fk::stencil_strips<5, 1, 2, THREADS>(smem, st, rb, re, W, W, 0, 0);
    } else if (blk.dil == 1) {
        if (blk.sr == 4) fk::stencil_strips<3, 1, 4, THREADS>(smem, st, rb, re, W, W, 0, 0);
        else fk::stencil_strips<3, 1, 2, THREADS>(smem, st, rb, re, W, W, 0, 0);
    } else if (blk.dil == 2) {
        if (blk.sr == 4) fk::stencil_strips<3, 2, 4, THREADS>(smem, st, rb, re, W, W, 0, 0);
        else fk::stencil_strips<3, 2, 2, THREADS>(smem, st, rb, re, W, W, 0, 0);
    } else {
        if (blk.sr == 4) fk::stencil_strips<3, 4, 4, THREADS>(smem, st, rb, re, W, W, 0, 0);
        else fk::stencil_strips<3, 4, 2, THREADS>(smem, st, rb, re, W, W, 0, 0);
    }
}

__global__ void __launch_bounds__(THREADS, 1)
mega_kernel(const __grid_constant__ CUtensorMap min, const __grid_constant__ CUtensorMap mout,
            const __grid_constant__ Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    // bars: [0] x tile, [1],[2] weight buffers, [3] MMAs
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t rank = cluster_rank();
    const int cluster_id = blockIdx.x / CLUSTER, nclusters = gridDim.x / CLUSTER;
    const int n_issuers = p.n_mt < NWARPS ? p.n_mt : NWARPS;
    if (tid == 0) {
        tc::mbar_init(&bars[0], 1);
        tc::mbar_init(&bars[1], 1);
        tc::mbar_init(&bars[2], 1);
        tc::mbar_init(&bars[3], n_issuers);
        tc::fence_barrier_init();
    }
    if (warp == 1) tc::tmem_alloc(tmem_slot, p.tmem_cols);
    // the m1 planes start all zero: left/right halo columns and the halo rows beyond the image stay zero for
    // the whole launch (conv zero padding); interior rows and neighbour halo rows are rewritten every block
    {
        const int n16 = (p.Cm >> 3) * p.rows_p * p.pitch;
        uint4* z = reinterpret_cast<uint4*>(smem + p.m1_off);
        for (int i = tid; i < n16; i += THREADS) z[i] = make_uint4(0, 0, 0, 0);
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t smem_base = tc::smem_u32(smem);
    // no CTA may push into a neighbour that is still zeroing
    cluster_arrive();
    cluster_wait();

    const int my_images = cluster_id < p.nimg ? (p.nimg - cluster_id + nclusters - 1) / nclusters : 0;
    const int total_seq = my_images * p.nblocks;       // (image, block) steps of this CTA: weights of step s in buffer s & 1
    auto load_weights = [&](int seq) {
        const int k = seq % p.nblocks, wb = seq & 1;
        tc::mbar_expect_tx(&bars[1 + wb], p.blob_bytes);
        fk::bulk_g2s(smem + p.wbuf_off[wb], p.blob + (size_t)k * p.blob_stride, p.blob_bytes, &bars[1 + wb]);
    };
    if (tid == 0) {
        if (total_seq > 0) load_weights(0);
        if (total_seq > 1) load_weights(1);
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    const int x_bytes = (p.Cs >> 3) * p.npx * 16;
    int seq = 0, mma_n = 0, it = 0;
    long long tprev = p.prof ? clock64() : 0;
    bool pending2 = false;      // a cluster barrier arrival (m1 planes read) not yet waited for
    for (int img = cluster_id; img < p.nimg; img += nclusters, ++it) {
        if (tid == 0) {
            tc::tma_store_wait_read();       // the previous image's store has read the x planes
            tc::mbar_expect_tx(&bars[0], x_bytes);
            tc::tma_load_act(smem + p.x_off, &min, &bars[0], p.in_planar, 0, (int)rank * p.R, 0, img);
        }
        tc::mbar_wait(&bars[0], it & 1);
        MEGA_ACC(0);   // x load
        for (int k = 0; k < p.nblocks; ++k, ++seq) {
            const int wbi = seq & 1;
            const uint8_t* wb = smem + p.wbuf_off[wbi];
            tc::mbar_wait(&bars[1 + wbi], (seq >> 1) & 1);
            const Block blk = p.blk[k];
            // ---- m1 = PReLU(W0 . x) ----
            if ((tid & 31) == 0 && warp < n_issuers) {
                tc::tc_fence_after();
                issue_gemm(smem_base + p.x_off, p.npx * 16, smem_base + p.wbuf_off[wbi] + p.w0_off, p.Cm, p.Cs >> 4, p.n_mt,
                           tmem_base, warp);
                tc::umma_commit(&bars[3]);
            }
            __syncwarp();
            tc::mbar_wait(&bars[3], mma_n & 1); ++mma_n;
            tc::tc_fence_after();
            MEGA_ACC(1);   // weights wait + first GEMM
            if (pending2) { cluster_wait(); pending2 = false; }   // neighbours have read their halo of the previous block
            epi_m1(smem, p, wb, tmem_base, blk.sep5 ? 2 : blk.dil, rank, 0, p.n_mt);
            tc::tc_fence_before();
            MEGA_ACC(2);   // m1 epilogue (+ wait for the neighbours' previous reads)
            cluster_arrive();       // release: my rows (local and pushed) are written ...
            cluster_wait();         // ... acquire: so are the halo rows the neighbours pushed to me
            MEGA_ACC(3);   // halo exchange barrier
            // ---- d = PReLU(dw(m1)) ----
            {
                fk::Stencil st;
                st.on = 1; st.in_off = p.m1_off; st.in_pitch = p.pitch; st.in_plane_bytes = p.rows_p * p.pitch * 16;
                st.halo = HALO; st.nchunks = p.Cm >> 3; st.dil = blk.dil;
                st.w_off = p.wbuf_off[wbi] + p.dw_off; st.vec_off = p.wbuf_off[wbi] + p.dv_off; st.sep5 = blk.sep5;
                st.out_off = p.d_off; st.out_plane_bytes = p.npx * 16;
                stencil(smem, st, blk, 0, p.R, p.W);
            }
            tc::fence_proxy_async();
            __syncthreads();
            MEGA_ACC(blk.sep5 ? 7 : 4);   // depthwise stencil (slot 7: the 25-tap kind)
            // my reads of the m1 planes are done (their values are consumed): no memory ordering to publish, so
            // the arrival is relaxed -- a releasing one drains every outstanding store first.  Waited for before
            // the next block's pushes.
            cluster_arrive_relaxed();
            pending2 = true;
            // ---- x = PReLU(W2 . d + x) ----
            if ((tid & 31) == 0 && warp < n_issuers) {
                tc::tc_fence_after();
                issue_gemm(smem_base + p.d_off, p.npx * 16, smem_base + p.wbuf_off[wbi] + p.w2_off, p.Cs, p.Cm >> 4, p.n_mt,
                           tmem_base, warp);
                tc::umma_commit(&bars[3]);
            }
            __syncwarp();
            tc::mbar_wait(&bars[3], mma_n & 1); ++mma_n;
            tc::tc_fence_after();
            MEGA_ACC(5);   // second GEMM
            epi_x(smem, p, wb, tmem_base, 0, p.n_mt);
            tc::fence_proxy_async();   // x planes: next read by tcgen05.mma / the TMA store
            tc::tc_fence_before();
            __syncthreads();
            MEGA_ACC(6);   // x epilogue
            // this weight buffer is free: fetch the weights of step seq + 2
            if (tid == 0 && seq + 2 < total_seq) load_weights(seq + 2);
        }
        if (tid == 0) {
            tc::tma_store_act(&mout, smem + p.x_off, p.out_planar, 0, (int)rank * p.R, 0, img);
            tc::tma_store_commit();
        }
    }
    if (pending2) cluster_wait();
    if (tid == 0) tc::tma_store_wait_read();
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc::tc_fence_after();
        tc::tmem_dealloc(tmem_base, p.tmem_cols);
    }
    // nobody leaves while a neighbour could still address its shared memory
    cluster_arrive();
    cluster_wait();
}

}  // namespace mega
}  // namespace isg
