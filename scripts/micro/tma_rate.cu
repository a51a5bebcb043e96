// Microbenchmark: TMA load rate per SM as a function of the box's inner (contiguous) extent.
// The product's activation layout ([chunk][pixel][8 ch] planes = canonical no-swizzle K-major UMMA operand)
// makes every TMA box element 16 bytes wide.  This measures what that costs against 32/64/128-byte
// inner extents (what 32/64/128-byte swizzled operand rows would use) for the same bytes per tile.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I instancesegmentation_b200/csrc \
//        scripts/micro/tma_rate.cu -o gpurun_out/tma_rate
#include <cstdio>
#include <vector>

#include "tc_common.cuh"
#include "tc_host.h"

using namespace isg;

// tensor [B][H][W][C] bf16 viewed as (inner ch, x, y, chunk, b), box (inner, TW, TH, C/inner, 1)
__global__ void __launch_bounds__(128, 1)
tma_rate_kernel(const __grid_constant__ CUtensorMap m, int tiles_x, int tiles_y, int B, int TW, int TH, int bytes,
                int nbuf, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bars[4];
    if (threadIdx.x == 0) {
        for (int i = 0; i < nbuf; ++i) tc::mbar_init(&bars[i], 1);
        tc::fence_barrier_init();
    }
    __syncthreads();
    const int total = tiles_x * tiles_y * B;
    auto issue = [&](int t, int buf) {
        const int tx = t % tiles_x, r = t / tiles_x, ty = r % tiles_y, b = r / tiles_y;
        tc::mbar_expect_tx(&bars[buf], bytes);
        tc::tma_load_5d(smem + (size_t)buf * bytes, &m, &bars[buf], 0, tx * TW, ty * TH, 0, b);
    };
    long long t0 = clock64();
    if (threadIdx.x == 0) {
        // nbuf loads in flight per SM; a buffer is re-armed as soon as its load has landed
        int issued = 0, it = 0;
        for (int t = blockIdx.x; t < total && issued < nbuf; t += gridDim.x, ++issued) issue(t, issued);
        int next = blockIdx.x + nbuf * gridDim.x;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            const int buf = it % nbuf;
            tc::mbar_wait(&bars[buf], (it / nbuf) & 1);
            if (next < total) { issue(next, buf); next += gridDim.x; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

__global__ void __launch_bounds__(128, 1)
tma_rate4_kernel(const __grid_constant__ CUtensorMap m, int tiles_x, int tiles_y, int B, int TWe, int TH, int bytes,
                 int nbuf, long long* cycles) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bars[4];
    if (threadIdx.x == 0) {
        for (int i = 0; i < nbuf; ++i) tc::mbar_init(&bars[i], 1);
        tc::fence_barrier_init();
    }
    __syncthreads();
    const int total = tiles_x * tiles_y * B;
    auto issue = [&](int t, int buf) {
        const int tx = t % tiles_x, r = t / tiles_x, ty = r % tiles_y, b = r / tiles_y;
        tc::mbar_expect_tx(&bars[buf], bytes);
        tc::tma_load_4d(smem + (size_t)buf * bytes, &m, &bars[buf], tx * TWe, ty * TH, 0, b);
    };
    long long t0 = clock64();
    if (threadIdx.x == 0) {
        int issued = 0, it = 0;
        for (int t = blockIdx.x; t < total && issued < nbuf; t += gridDim.x, ++issued) issue(t, issued);
        int next = blockIdx.x + nbuf * gridDim.x;
        for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
            const int buf = it % nbuf;
            tc::mbar_wait(&bars[buf], (it / nbuf) & 1);
            if (next < total) { issue(next, buf); next += gridDim.x; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
    const int B = 64, H = 128, W = 128, C = 64;           // 134 MB of bf16 activations (> L2)
    const size_t n = (size_t)B * H * W * C;
    __nv_bfloat16* d;
    cudaMalloc(&d, n * 2);
    cudaMemset(d, 0, n * 2);
    long long* cyc;
    cudaMalloc(&cyc, 148 * sizeof(long long));
    const int TW = 32, TH = 8;                            // tile = 256 pixels x 64 ch = 32 KB
    const int bytes = TW * TH * C * 2;
    cudaFuncSetAttribute(tma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(tma_rate4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    printf("tile %dx%d pixels x %d ch = %d bytes, %d tiles, tensor %.0f MB\n", TW, TH, C, bytes, (W / TW) * (H / TH) * B, n * 2 / 1e6);
    for (int inner : {8, 16, 32, 64}) {                   // channels per box element: 16 / 32 / 64 / 128 bytes
        for (int nbuf : {1, 2, 4}) {
            CUtensorMap m;
            const uint64_t dims[5] = {(uint64_t)inner, (uint64_t)W, (uint64_t)H, (uint64_t)C / inner, (uint64_t)B};
            const uint64_t st[4] = {(uint64_t)C * 2, (uint64_t)W * C * 2, (uint64_t)inner * 2, (uint64_t)H * W * C * 2};
            const uint32_t box[5] = {(uint32_t)inner, (uint32_t)TW, (uint32_t)TH, (uint32_t)(C / inner), 1};
            if (tc::make_map5(&m, d, dims, st, box, "bench")) { printf("map failed: %s\n", err_slot().c_str()); return 1; }
            float best = 1e30f;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0); cudaEventCreate(&e1);
                cudaEventRecord(e0);
                tma_rate_kernel<<<148, 128, (size_t)nbuf * bytes>>>(m, W / TW, H / TH, B, TW, TH, bytes, nbuf, cyc);
                cudaEventRecord(e1);
                if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                best = ms < best ? ms : best;
            }
            std::vector<long long> h(148);
            cudaMemcpy(h.data(), cyc, 148 * sizeof(long long), cudaMemcpyDeviceToHost);
            long long mx = 0;
            for (auto v : h) mx = v > mx ? v : mx;
            const double per_sm_bytes = (double)n * 2 / 148;
            printf("inner %3d B  loads in flight %d: %7.1f us  %6.0f GB/s  %5.1f B/clk/SM  %5.2f box elements/clk/SM\n", inner * 2, nbuf,
                   best * 1e3, n * 2 / best / 1e6, per_sm_bytes / mx, per_sm_bytes / (inner * 2) / mx);
        }
    }
    // planar-by-chunk global layout [B][C/8][H][W][8 ch]: the same (8 ch, x, y, chunk, b) box, but x is now
    // contiguous after the 8 channels.  (a) described as before (5-D, 16-byte inner dim, x stride 16 B);
    // (b) x and the channels merged into one dimension of 8-byte elements (box inner extent = TW * 16 B)
    printf("planar-by-chunk global layout [B][C/8][H][W][8]:\n");
    for (int variant = 0; variant < 2; ++variant)
        for (int nbuf : {1, 2, 4}) {
            CUtensorMap m;
            cuuint64_t dd[5], ss[4];
            cuuint32_t bx[5], es[5] = {1, 1, 1, 1, 1};
            CUtensorMapDataType dt;
            int rank;
            if (variant == 0) {
                dt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16; rank = 5;
                dd[0] = 8; dd[1] = W; dd[2] = H; dd[3] = C / 8; dd[4] = B;
                ss[0] = 16; ss[1] = (cuuint64_t)W * 16; ss[2] = (cuuint64_t)H * W * 16; ss[3] = (cuuint64_t)H * W * C * 2;
                bx[0] = 8; bx[1] = TW; bx[2] = TH; bx[3] = C / 8; bx[4] = 1;
            } else {
                dt = CU_TENSOR_MAP_DATA_TYPE_UINT64; rank = 4;
                dd[0] = 2 * W; dd[1] = H; dd[2] = C / 8; dd[3] = B;
                ss[0] = (cuuint64_t)W * 16; ss[1] = (cuuint64_t)H * W * 16; ss[2] = (cuuint64_t)H * W * C * 2;
                bx[0] = 2 * TW; bx[1] = TH; bx[2] = C / 8; bx[3] = 1;
            }
            CUresult r = tc::encode_fn()(&m, dt, rank, d, dd, ss, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) { printf("encode failed %d\n", (int)r); return 1; }
            float best = 1e30f;
            for (int rep = 0; rep < 3; ++rep) {
                cudaEvent_t e0, e1;
                cudaEventCreate(&e0); cudaEventCreate(&e1);
                cudaEventRecord(e0);
                if (variant == 0) tma_rate_kernel<<<148, 128, (size_t)nbuf * bytes>>>(m, W / TW, H / TH, B, TW, TH, bytes, nbuf, cyc);
                else tma_rate4_kernel<<<148, 128, (size_t)nbuf * bytes>>>(m, W / TW, H / TH, B, 2 * TW, TH, bytes, nbuf, cyc);
                cudaEventRecord(e1);
                if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                float ms;
                cudaEventElapsedTime(&ms, e0, e1);
                best = ms < best ? ms : best;
            }
            printf("%s  loads in flight %d: %7.1f us  %6.0f GB/s\n", variant ? "4-D, 8-byte elements, inner extent TW*16 B" : "5-D, 16-byte inner dim, x stride 16 B    ",
                   nbuf, best * 1e3, n * 2 / best / 1e6);
        }
    return 0;
}
