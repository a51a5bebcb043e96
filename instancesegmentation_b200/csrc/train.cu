// Loss side of the training step (first slice): reference train_instance.py:299 + :377-378 is
// BCELoss()(sigmoid(forward(x)), mask) followed by loss.backward().  One launch computes the mean loss AND
// d(loss)/d(logit) from the raw logits, so the step keeps no probability map and runs no separate sigmoid /
// loss / loss-backward kernels.  Semantics are torch's, restated in oracle/bce_oracle.py:
//   p = 1 / (1 + exp(-x));  l = -(y * max(log p, -100) + (1 - y) * max(log(1 - p), -100));  loss = mean(l)
//   dloss/dx = (p - y) / max(p * (1 - p), 1e-12) * p * (1 - p) / n
// The sum is deterministic: a fixed grid, every block adds its elements in a fixed order in double, the last
// block to finish (ticket) adds the per-block partials in index order and resets the ticket.
#include <cmath>

#include "common.cuh"

namespace isg {
namespace {

constexpr int BCE_THREADS = 256;
constexpr int BCE_MAX_BLOCKS = 1024;   // partials live in the caller's workspace: BCE_MAX_BLOCKS doubles + ticket

__device__ __forceinline__ float load_logit(const float* p, int64_t i) { return __ldg(p + i); }
__device__ __forceinline__ float load_logit(const __nv_bfloat16* p, int64_t i) {
    return __bfloat162float(__ldg(p + i));
}
__device__ __forceinline__ void store_grad(float* p, int64_t i, float v) { p[i] = v; }
__device__ __forceinline__ void store_grad(__nv_bfloat16* p, int64_t i, float v) { p[i] = __float2bfloat16_rn(v); }

template <typename T>
__global__ void __launch_bounds__(BCE_THREADS)
bce_sigmoid_kernel(const T* __restrict__ logits, const float* __restrict__ target, int64_t n, float inv_n,
                   T* __restrict__ dlogits, double* __restrict__ partial, unsigned* __restrict__ ticket,
                   float* __restrict__ loss_out) {
    double acc = 0.0;
    const int64_t stride = (int64_t)gridDim.x * BCE_THREADS;
    for (int64_t i = blockIdx.x * (int64_t)BCE_THREADS + threadIdx.x; i < n; i += stride) {
        const float x = load_logit(logits, i);
        const float y = __ldg(target + i);
        const float p = 1.f / (1.f + expf(-x));
        const float q = 1.f - p;
        const float lp = fmaxf(logf(p), -100.f), lq = fmaxf(logf(q), -100.f);
        acc += (double)(-(y * lp + (1.f - y) * lq));
        if (dlogits) {
            const float pq = p * q;
            store_grad(dlogits, i, (p - y) / fmaxf(pq, 1e-12f) * pq * inv_n);
        }
    }
    // block sum in a fixed order: lanes by xor shuffles, then thread 0 over the 8 warp sums
    __shared__ double warp_sum[BCE_THREADS / 32];
    __shared__ bool last;
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) warp_sum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < BCE_THREADS / 32; ++w) s += warp_sum[w];
        partial[blockIdx.x] = s;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        double s = 0.0;
        for (unsigned b = 0; b < gridDim.x; ++b) s += ((volatile double*)partial)[b];
        *loss_out = (float)(s * (double)inv_n);
        *ticket = 0u;   // ready for the next launch on the same workspace
    }
}

}  // namespace
}  // namespace isg

using namespace isg;

extern "C" {

int64_t isg_bce_workspace_bytes(void) { return (int64_t)BCE_MAX_BLOCKS * 8 + 64; }

int isg_bce_sigmoid(const void* logits_dev, int logits_bf16, const float* target_dev, int64_t n, void* dlogits_dev,
                    void* workspace_dev, float* loss_dev, void* stream) {
    if (n <= 0) return fail(ISG_ERR_ARG, "isg_bce_sigmoid: the mean of an empty batch is undefined (n=%lld)", (long long)n);
    if (!logits_dev || !target_dev || !workspace_dev || !loss_dev) return fail(ISG_ERR_ARG, "isg_bce_sigmoid: null argument");
    if (reinterpret_cast<uintptr_t>(workspace_dev) & 7) return fail(ISG_ERR_ARG, "isg_bce_sigmoid: workspace must be 8-byte aligned");
    int64_t blocks = (n + BCE_THREADS * 4 - 1) / (BCE_THREADS * 4);   // >= 4 elements per thread before the grid grows
    if (blocks > BCE_MAX_BLOCKS) blocks = BCE_MAX_BLOCKS;
    double* partial = static_cast<double*>(workspace_dev);
    unsigned* ticket = reinterpret_cast<unsigned*>(partial + BCE_MAX_BLOCKS);   // zero before the first use (caller memsets once)
    const float inv_n = (float)(1.0 / (double)n);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (logits_bf16)
        bce_sigmoid_kernel<__nv_bfloat16><<<(unsigned)blocks, BCE_THREADS, 0, st>>>(
            static_cast<const __nv_bfloat16*>(logits_dev), target_dev, n, inv_n, static_cast<__nv_bfloat16*>(dlogits_dev),
            partial, ticket, loss_dev);
    else
        bce_sigmoid_kernel<float><<<(unsigned)blocks, BCE_THREADS, 0, st>>>(
            static_cast<const float*>(logits_dev), target_dev, n, inv_n, static_cast<float*>(dlogits_dev), partial, ticket,
            loss_dev);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

}  // extern "C"
