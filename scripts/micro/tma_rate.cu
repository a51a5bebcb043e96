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
    return 0;
}
