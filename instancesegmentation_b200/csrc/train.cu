// Loss side of the training step (first slice): reference train_instance.py:299 + :377-378 is
// BCELoss()(sigmoid(forward(x)), mask) followed by loss.backward().  One launch computes the mean loss AND
// d(loss)/d(logit) from the raw logits, so the step keeps no probability map and runs no separate sigmoid /
// loss / loss-backward kernels.  Semantics are torch's, restated in oracle/bce_oracle.py:
//   p = 1 / (1 + exp(-x));  l = -(y * max(log p, -100) + (1 - y) * max(log(1 - p), -100));  loss = mean(l)
//   dloss/dx = (p - y) / max(p * (1 - p), 1e-12) * p * (1 - p) / n
// The sum is deterministic: a fixed grid, every block adds its elements in a fixed order in double, the last
// block to finish (ticket) adds the per-block partials in index order and resets the ticket.
#include <cmath>
#include <cstdlib>
#include <cstring>

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


// ---- train-mode BatchNorm2d (reference segment.py:41 nn.BatchNorm2d inside every Conv, train_instance.py:372) ----
// NCHW, bf16 or fp32 activations, fp32 statistics.  The library kernels behind the same op take 44 % of the
// reference's training step on this GPU (scripts/train_profile.py); these are plain streaming passes:
//   forward : bn_stats (sum, sum of squares about a per-channel shift)  ->  bn_apply (y = x*scale + shift', running stats)
//   backward: bn_bwd_reduce (sum dy, sum dy*(x-mean))                   ->  bn_bwd_apply (dx, dgamma, dbeta)
// grid = (splits, channels): every block owns a contiguous range of 16-byte vectors of one channel; partial sums are
// combined in double in a fixed order, so results are deterministic.
constexpr int BN_THREADS = 256;
constexpr int BN_MAX_SPLIT = 64;

template <typename T, int V> struct VecIO;
template <> struct VecIO<float, 4> {
    typedef float4 Raw;
    static __device__ __forceinline__ Raw ldraw(const float* p) { return *reinterpret_cast<const float4*>(p); }
    static __device__ __forceinline__ void unpack(const Raw& a, float (&v)[4]) { v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; }
    static __device__ __forceinline__ void store(float* p, const float (&v)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <> struct VecIO<float, 1> {
    typedef float Raw;
    static __device__ __forceinline__ Raw ldraw(const float* p) { return *p; }
    static __device__ __forceinline__ void unpack(const Raw& a, float (&v)[1]) { v[0] = a; }
    static __device__ __forceinline__ void store(float* p, const float (&v)[1]) { *p = v[0]; }
};
template <> struct VecIO<__nv_bfloat16, 8> {
    typedef uint4 Raw;
    static __device__ __forceinline__ Raw ldraw(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }
    static __device__ __forceinline__ void unpack(const Raw& a, float (&v)[8]) {
        const unsigned w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
        unsigned w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            w[i] = *reinterpret_cast<const unsigned*>(&h);
        }
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};
template <> struct VecIO<__nv_bfloat16, 1> {
    typedef __nv_bfloat16 Raw;
    static __device__ __forceinline__ Raw ldraw(const __nv_bfloat16* p) { return *p; }
    static __device__ __forceinline__ void unpack(const Raw& a, float (&v)[1]) { v[0] = __bfloat162float(a); }
    static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[1]) { *p = __float2bfloat16_rn(v[0]); }
};

__device__ __forceinline__ float to_float(float v) { return v; }
__device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }

// this block's [lo, hi) range of V-element vectors of one channel (n_vec = B*HW/V of them)
__device__ __forceinline__ void bn_range(unsigned n_vec, unsigned& lo, unsigned& hi) {
    const unsigned per = (n_vec + gridDim.x - 1) / gridDim.x;
    lo = min(n_vec, blockIdx.x * per);
    hi = min(n_vec, lo + per);
}

// element offset of vector i of channel c: planes of HW elements, batch stride C*HW (hw_shift >= 0: HW is a power of two)
template <int V>
__device__ __forceinline__ size_t bn_at(unsigned i, int c, int C, unsigned HW, int hw_shift) {
    const unsigned e = i * V;
    const unsigned b = hw_shift >= 0 ? e >> hw_shift : e / HW;
    return ((size_t)b * C + c) * HW + (e - b * HW);
}
constexpr int BN_U1 = 4;   // independent 16-byte loads in flight per thread, one-tensor kernels
constexpr int BN_U2 = 2;   // two-tensor kernels (dy and x)

// block-wide sums of N doubles in a fixed order; valid in thread 0
template <int N>
__device__ __forceinline__ void block_sum(double (&a)[N]) {
    __shared__ double sh[N][BN_THREADS / 32];
#pragma unroll
    for (int k = 0; k < N; ++k)
        for (int o = 16; o; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) sh[k][threadIdx.x >> 5] = a[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) {
            a[k] = 0.0;
            for (int w = 0; w < BN_THREADS / 32; ++w) a[k] += sh[k][w];
        }
    }
}

// partials are stored [channel][split][N]; total over the gridDim.x splits of channel c (every thread gets it)
template <int N>
__device__ __forceinline__ void partial_total(const double* __restrict__ part, int c, double (&a)[N]) {
#pragma unroll
    for (int k = 0; k < N; ++k) a[k] = 0.0;
    for (unsigned s = 0; s < gridDim.x; ++s) {
#pragma unroll
        for (int k = 0; k < N; ++k) a[k] += part[((size_t)c * gridDim.x + s) * N + k];
    }
}
template <int N>
__device__ __forceinline__ void partial_store(double* __restrict__ part, int c, const double (&a)[N]) {
#pragma unroll
    for (int k = 0; k < N; ++k) part[((size_t)c * gridDim.x + blockIdx.x) * N + k] = a[k];
}

// activation that follows the BatchNorm inside the reference's Conv block (segment.py:42-45): ISG_ACT_* of isg.h
__device__ __forceinline__ float act_fwd(float z, int act, float slope) {
    if (act == 1) return z > 0.f ? z : 0.f;
    if (act == 2) return z > 0.f ? z : z * slope;
    return z;
}
// d act / d z times the incoming gradient (torch's conventions at z == 0: ReLU 0, PReLU slope)
__device__ __forceinline__ float act_bwd(float z, float g, int act, float slope) {
    if (act == 1) return z > 0.f ? g : 0.f;
    if (act == 2) return z > 0.f ? g : g * slope;
    return g;
}

template <typename T, int V>
__global__ void __launch_bounds__(BN_THREADS)
bn_stats_kernel(const T* __restrict__ x, int C, unsigned HW, unsigned n_vec, int hw_shift, double* __restrict__ part) {
    const int c = blockIdx.y;
    const float shift = to_float(x[(size_t)c * HW]);   // sums about the channel's first element: no cancellation in E[d^2]-E[d]^2
    unsigned lo, hi;
    bn_range(n_vec, lo, hi);
    float s1 = 0.f, s2 = 0.f;
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += BN_THREADS * BN_U1) {
        typename VecIO<T, V>::Raw r[BN_U1];
#pragma unroll
        for (int u = 0; u < BN_U1; ++u)
            if (i0 + u * BN_THREADS < hi) r[u] = VecIO<T, V>::ldraw(x + bn_at<V>(i0 + u * BN_THREADS, c, C, HW, hw_shift));
#pragma unroll
        for (int u = 0; u < BN_U1; ++u) {
            if (i0 + u * BN_THREADS >= hi) break;
            float v[V];
            VecIO<T, V>::unpack(r[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) { const float d = v[k] - shift; s1 += d; s2 = fmaf(d, d, s2); }
        }
    }
    double a[2] = {s1, s2};
    block_sum<2>(a);
    if (threadIdx.x == 0) partial_store<2>(part, c, a);
}

template <typename T, int V>
__global__ void __launch_bounds__(BN_THREADS)
bn_apply_kernel(const T* __restrict__ x, int C, unsigned HW, unsigned n_vec, int hw_shift, const double* __restrict__ part,
                const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean_offset,
                float* __restrict__ running_mean, float* __restrict__ running_var, float momentum, float eps, int act,
                const float* __restrict__ slope, int slope_stride, T* __restrict__ y, float* __restrict__ save_mean,
                float* __restrict__ save_invstd) {
    const int c = blockIdx.y;
    const double n = (double)n_vec * V;
    double s[2];
    partial_total<2>(part, c, s);
    const double dm = s[0] / n;
    const double var = fmax(s[1] / n - dm * dm, 0.0);
    const float mean = (float)((double)to_float(x[(size_t)c * HW]) + dm);
    const float invstd = (float)(1.0 / sqrt(var + (double)eps));
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        save_mean[c] = mean;
        save_invstd[c] = invstd;
        // mean_offset: a per-channel constant the caller left out of x (the conv bias, which the mean subtraction cancels)
        const float m_full = mean + (mean_offset ? mean_offset[c] : 0.f);
        if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m_full;
        if (running_var) {
            const float unbiased = (float)(n > 1.0 ? var * n / (n - 1.0) : var);
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
    }
    const float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
    const float scale = g * invstd;
    const float sl = (act == 2) ? slope[c * slope_stride] : 0.f;
    unsigned lo, hi;
    bn_range(n_vec, lo, hi);
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += BN_THREADS * BN_U1) {
        typename VecIO<T, V>::Raw r[BN_U1];
        size_t at[BN_U1];
#pragma unroll
        for (int u = 0; u < BN_U1; ++u)
            if (i0 + u * BN_THREADS < hi) {
                at[u] = bn_at<V>(i0 + u * BN_THREADS, c, C, HW, hw_shift);
                r[u] = VecIO<T, V>::ldraw(x + at[u]);
            }
#pragma unroll
        for (int u = 0; u < BN_U1; ++u) {
            if (i0 + u * BN_THREADS >= hi) break;
            float v[V];
            VecIO<T, V>::unpack(r[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) v[k] = act_fwd(fmaf(v[k] - mean, scale, bt), act, sl);
            VecIO<T, V>::store(y + at[u], v);
        }
    }
}

// sums over one channel of: gz, gz * (x - mean), and (PReLU) dy * z where z <= 0;  gz = dy * act'(z), z = bn(x)
template <typename T, int V>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ x, int C, unsigned HW, unsigned n_vec, int hw_shift,
                     const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ save_mean,
                     const float* __restrict__ save_invstd, int act, const float* __restrict__ slope, int slope_stride,
                     double* __restrict__ part) {
    const int c = blockIdx.y;
    const float mean = save_mean[c];
    const float scale = (gamma ? gamma[c] : 1.f) * save_invstd[c], bt = beta ? beta[c] : 0.f;
    const float sl = (act == 2) ? slope[c * slope_stride] : 0.f;
    unsigned lo, hi;
    bn_range(n_vec, lo, hi);
    float s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += BN_THREADS * BN_U2) {
        typename VecIO<T, V>::Raw rg[BN_U2], rx[BN_U2];
#pragma unroll
        for (int u = 0; u < BN_U2; ++u)
            if (i0 + u * BN_THREADS < hi) {
                const size_t at = bn_at<V>(i0 + u * BN_THREADS, c, C, HW, hw_shift);
                rg[u] = VecIO<T, V>::ldraw(dy + at);
                rx[u] = VecIO<T, V>::ldraw(x + at);
            }
#pragma unroll
        for (int u = 0; u < BN_U2; ++u) {
            if (i0 + u * BN_THREADS >= hi) break;
            float g[V], v[V];
            VecIO<T, V>::unpack(rg[u], g);
            VecIO<T, V>::unpack(rx[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const float xm = v[k] - mean;
                float gz = g[k];
                if (act) {
                    const float z = fmaf(xm, scale, bt);
                    gz = act_bwd(z, g[k], act, sl);
                    if (act == 2 && !(z > 0.f)) s3 = fmaf(g[k], z, s3);
                }
                s1 += gz;
                s2 = fmaf(gz, xm, s2);
            }
        }
    }
    double a[3] = {s1, s2, s3};
    block_sum<3>(a);
    if (threadIdx.x == 0) partial_store<3>(part, c, a);
}

template <typename T, int V>
__global__ void __launch_bounds__(BN_THREADS)
bn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ x, int C, unsigned HW, unsigned n_vec, int hw_shift,
                    const double* __restrict__ part, const float* __restrict__ gamma, const float* __restrict__ beta,
                    const float* __restrict__ save_mean, const float* __restrict__ save_invstd, int act,
                    const float* __restrict__ slope, int slope_stride, T* __restrict__ dx, float* __restrict__ dgamma,
                    float* __restrict__ dbeta, float* __restrict__ dslope) {
    const int c = blockIdx.y;
    const double n = (double)n_vec * V;
    double s[3];
    partial_total<3>(part, c, s);
    const float mean = save_mean[c], invstd = save_invstd[c];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (dgamma) dgamma[c] = (float)(s[1] * (double)invstd);
        if (dbeta) dbeta[c] = (float)s[0];
        if (dslope) dslope[c] = (float)s[2];
    }
    if (!dx) return;
    const float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
    const float k1 = (float)(s[0] / n);
    const float k2 = (float)(s[1] * (double)invstd * (double)invstd / n);
    const float gs = g * invstd;
    const float sl = (act == 2) ? slope[c * slope_stride] : 0.f;
    unsigned lo, hi;
    bn_range(n_vec, lo, hi);
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += BN_THREADS * BN_U2) {
        typename VecIO<T, V>::Raw rg[BN_U2], rx[BN_U2];
        size_t at[BN_U2];
#pragma unroll
        for (int u = 0; u < BN_U2; ++u)
            if (i0 + u * BN_THREADS < hi) {
                at[u] = bn_at<V>(i0 + u * BN_THREADS, c, C, HW, hw_shift);
                rg[u] = VecIO<T, V>::ldraw(dy + at[u]);
                rx[u] = VecIO<T, V>::ldraw(x + at[u]);
            }
#pragma unroll
        for (int u = 0; u < BN_U2; ++u) {
            if (i0 + u * BN_THREADS >= hi) break;
            float gy[V], v[V];
            VecIO<T, V>::unpack(rg[u], gy);
            VecIO<T, V>::unpack(rx[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const float xm = v[k] - mean;
                const float gz = act ? act_bwd(fmaf(xm, gs, bt), gy[k], act, sl) : gy[k];
                v[k] = (gz - k1 - xm * k2) * gs;
            }
            VecIO<T, V>::store(dx + at[u], v);
        }
    }
}

// ---- one launch per op: a thread-block cluster per channel --------------------------------------------------------
// grid (CL, channels), cluster (CL,1,1): the CL CTAs of a cluster split one channel, leave their partial sums in their own
// shared memory, meet at a cluster barrier, read each other's partials through distributed shared memory (fixed rank
// order -> deterministic, every CTA gets bit-identical totals) and go straight on to the normalisation pass, whose
// re-read of x (and dy) is served by L2 for the layer sizes of this network.  No global partials, no second launch.
constexpr int BN_CL_MAX = 8;
constexpr int BN_CL_MAX_THREADS = 1024;

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ double ld_dsmem_f64(const double* local, unsigned rank) {
    unsigned a = (unsigned)__cvta_generic_to_shared(local), ra;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
    double v;
    asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(v) : "r"(ra) : "memory");
    return v;
}

// sums of N per-thread values over the whole cluster (all CTAs, all threads get the same totals)
template <int N>
__device__ __forceinline__ void cluster_sum(double (&a)[N]) {
    __shared__ double warp_part[N][BN_CL_MAX_THREADS / 32];
    __shared__ double cta_part[N];
    const int nwarp = blockDim.x >> 5;
#pragma unroll
    for (int k = 0; k < N; ++k)
        for (int o = 16; o; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) warp_part[k][threadIdx.x >> 5] = a[k];
    }
    __syncthreads();
    if (threadIdx.x < N) {
        double t = 0.0;
        for (int w = 0; w < nwarp; ++w) t += warp_part[threadIdx.x][w];
        cta_part[threadIdx.x] = t;
    }
    cluster_sync_all();                       // every CTA's cta_part is written and visible cluster-wide
#pragma unroll
    for (int k = 0; k < N; ++k) {
        double t = 0.0;
        for (unsigned r = 0; r < gridDim.x; ++r) t += ld_dsmem_f64(&cta_part[k], r);
        a[k] = t;
    }
    cluster_sync_all();                       // nobody leaves (or reuses cta_part) while a peer still reads it
}

template <typename T, int V>
__global__ void __launch_bounds__(BN_CL_MAX_THREADS)
bn_fwd_cluster_kernel(const T* __restrict__ x, int C, unsigned HW, unsigned n_vec, int hw_shift,
                      const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean_offset,
                      float* __restrict__ running_mean, float* __restrict__ running_var, float momentum, float eps, int act,
                      const float* __restrict__ slope, int slope_stride, T* __restrict__ y, float* __restrict__ save_mean,
                      float* __restrict__ save_invstd) {
    const int c = blockIdx.y;
    const unsigned nthr = blockDim.x;
    const float shift = to_float(x[(size_t)c * HW]);
    unsigned lo, hi;
    bn_range(n_vec, lo, hi);
    float s1 = 0.f, s2 = 0.f;
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += nthr * BN_U1) {
        typename VecIO<T, V>::Raw r[BN_U1];
#pragma unroll
        for (int u = 0; u < BN_U1; ++u)
            if (i0 + u * nthr < hi) r[u] = VecIO<T, V>::ldraw(x + bn_at<V>(i0 + u * nthr, c, C, HW, hw_shift));
#pragma unroll
        for (int u = 0; u < BN_U1; ++u) {
            if (i0 + u * nthr >= hi) break;
            float v[V];
            VecIO<T, V>::unpack(r[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) { const float d = v[k] - shift; s1 += d; s2 = fmaf(d, d, s2); }
        }
    }
    double s[2] = {s1, s2};
    cluster_sum<2>(s);
    const double n = (double)n_vec * V;
    const double dm = s[0] / n;
    const double var = fmax(s[1] / n - dm * dm, 0.0);
    const float mean = (float)((double)shift + dm);
    const float invstd = (float)(1.0 / sqrt(var + (double)eps));
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        save_mean[c] = mean;
        save_invstd[c] = invstd;
        const float m_full = mean + (mean_offset ? mean_offset[c] : 0.f);
        if (running_mean) running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m_full;
        if (running_var) {
            const float unbiased = (float)(n > 1.0 ? var * n / (n - 1.0) : var);
            running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
    }
    const float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
    const float scale = g * invstd;
    const float sl = (act == 2) ? slope[c * slope_stride] : 0.f;
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += nthr * BN_U1) {
        typename VecIO<T, V>::Raw r[BN_U1];
        size_t at[BN_U1];
#pragma unroll
        for (int u = 0; u < BN_U1; ++u)
            if (i0 + u * nthr < hi) {
                at[u] = bn_at<V>(i0 + u * nthr, c, C, HW, hw_shift);
                r[u] = VecIO<T, V>::ldraw(x + at[u]);
            }
#pragma unroll
        for (int u = 0; u < BN_U1; ++u) {
            if (i0 + u * nthr >= hi) break;
            float v[V];
            VecIO<T, V>::unpack(r[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) v[k] = act_fwd(fmaf(v[k] - mean, scale, bt), act, sl);
            VecIO<T, V>::store(y + at[u], v);
        }
    }
}

template <typename T, int V>
__global__ void __launch_bounds__(BN_CL_MAX_THREADS)
bn_bwd_cluster_kernel(const T* __restrict__ dy, const T* __restrict__ x, int C, unsigned HW, unsigned n_vec, int hw_shift,
                      const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ save_mean,
                      const float* __restrict__ save_invstd, int act, const float* __restrict__ slope, int slope_stride,
                      T* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dslope) {
    const int c = blockIdx.y;
    const unsigned nthr = blockDim.x;
    const float mean = save_mean[c], invstd = save_invstd[c];
    const float g = gamma ? gamma[c] : 1.f, bt = beta ? beta[c] : 0.f;
    const float gs = g * invstd;
    const float sl = (act == 2) ? slope[c * slope_stride] : 0.f;
    unsigned lo, hi;
    bn_range(n_vec, lo, hi);
    float s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += nthr * BN_U2) {
        typename VecIO<T, V>::Raw rg[BN_U2], rx[BN_U2];
#pragma unroll
        for (int u = 0; u < BN_U2; ++u)
            if (i0 + u * nthr < hi) {
                const size_t at = bn_at<V>(i0 + u * nthr, c, C, HW, hw_shift);
                rg[u] = VecIO<T, V>::ldraw(dy + at);
                rx[u] = VecIO<T, V>::ldraw(x + at);
            }
#pragma unroll
        for (int u = 0; u < BN_U2; ++u) {
            if (i0 + u * nthr >= hi) break;
            float gy[V], v[V];
            VecIO<T, V>::unpack(rg[u], gy);
            VecIO<T, V>::unpack(rx[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const float xm = v[k] - mean;
                float gz = gy[k];
                if (act) {
                    const float z = fmaf(xm, gs, bt);
                    gz = act_bwd(z, gy[k], act, sl);
                    if (act == 2 && !(z > 0.f)) s3 = fmaf(gy[k], z, s3);
                }
                s1 += gz;
                s2 = fmaf(gz, xm, s2);
            }
        }
    }
    double s[3] = {s1, s2, s3};
    cluster_sum<3>(s);
    const double n = (double)n_vec * V;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (dgamma) dgamma[c] = (float)(s[1] * (double)invstd);
        if (dbeta) dbeta[c] = (float)s[0];
        if (dslope) dslope[c] = (float)s[2];
    }
    if (!dx) return;
    const float k1 = (float)(s[0] / n);
    const float k2 = (float)(s[1] * (double)invstd * (double)invstd / n);
    for (unsigned i0 = lo + threadIdx.x; i0 < hi; i0 += nthr * BN_U2) {
        typename VecIO<T, V>::Raw rg[BN_U2], rx[BN_U2];
        size_t at[BN_U2];
#pragma unroll
        for (int u = 0; u < BN_U2; ++u)
            if (i0 + u * nthr < hi) {
                at[u] = bn_at<V>(i0 + u * nthr, c, C, HW, hw_shift);
                rg[u] = VecIO<T, V>::ldraw(dy + at[u]);
                rx[u] = VecIO<T, V>::ldraw(x + at[u]);
            }
#pragma unroll
        for (int u = 0; u < BN_U2; ++u) {
            if (i0 + u * nthr >= hi) break;
            float gy[V], v[V];
            VecIO<T, V>::unpack(rg[u], gy);
            VecIO<T, V>::unpack(rx[u], v);
#pragma unroll
            for (int k = 0; k < V; ++k) {
                const float xm = v[k] - mean;
                const float gz = act ? act_bwd(fmaf(xm, gs, bt), gy[k], act, sl) : gy[k];
                v[k] = (gz - k1 - xm * k2) * gs;
            }
            VecIO<T, V>::store(dx + at[u], v);
        }
    }
}

// cluster size and CTA size for n_vec vectors per channel: aim at >= 4 vectors per thread, then as wide as allowed
inline void bn_cluster_shape(unsigned n_vec, int channels, int* cl, int* threads) {
    int t = 256;
    int c = 1;
    while (c < BN_CL_MAX && (unsigned)(c * t * 4) < n_vec) c *= 2;
    while (t < BN_CL_MAX_THREADS && (unsigned)(c * t * 8) < n_vec) t *= 2;
    // few channels: keep the machine busy with more, smaller CTAs rather than one fat CTA per SM
    (void)channels;
    *cl = c; *threads = t;
}

template <typename K, typename... Args>
inline cudaError_t launch_cluster(K kernel, int cl, int channels, int threads, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    cfg.gridDim = dim3(cl, channels);
    cfg.blockDim = dim3(threads);
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, args...);
}

inline int bn_split(int channels, unsigned n_vec) {
    int s = (148 * 8 + channels - 1) / channels;                     // ~8 blocks per SM over all channels
    const unsigned by_work = (n_vec + BN_THREADS * 2 - 1) / (BN_THREADS * 2);   // at least two vectors per thread
    if ((unsigned)s > by_work) s = (int)by_work;
    if (s > BN_MAX_SPLIT) s = BN_MAX_SPLIT;
    return s < 1 ? 1 : s;
}


// ---- depthwise convolution of the bottleneck blocks in train mode (segment.py:58, :93-96, :121: groups == channels) ----
// NCHW, stride 1, output size == input size (2*pad == dil*(k-1)), any kh x kw <= 25 taps, dilation: the 3x3 (dil 1/2/4),
// (5,1) and (1,5) convs of the reference.  The library splits these over three code paths, one of which (grouped-direct
// cuDNN dgrad/wgrad for the non-square kernels) takes 0.45 ms per call on a 1.5 MB tensor.  Here: one stencil kernel
// (forward; dgrad = the same kernel with flipped taps) and one weight/bias-gradient kernel, a cluster per channel.
// Activations bf16 or fp32, weights / bias / their gradients fp32, accumulation fp32 (double across threads).
constexpr int DW_TH = 16, DW_TW = 64, DW_THREADS = 256, DW_MAX_TAPS = 25;
struct DwGeom { int C, H, W, kh, kw, ph, pw, dh, dw, SH, SW, tiles_h, tiles_w; };

template <typename T>
__device__ __forceinline__ void dw_load_tile(const T* __restrict__ plane, const DwGeom& g, int h0, int w0, int ph, int pw,
                                             float* __restrict__ sm) {
    for (int idx = threadIdx.x; idx < g.SH * g.SW; idx += blockDim.x) {
        const int r = idx / g.SW, q = idx - r * g.SW;
        const int hh = h0 + r - ph, ww = w0 + q - pw;
        sm[idx] = (hh >= 0 && hh < g.H && ww >= 0 && ww < g.W) ? to_float(plane[(size_t)hh * g.W + ww]) : 0.f;
    }
}
__device__ __forceinline__ void store_out(float* p, float v) { *p = v; }
__device__ __forceinline__ void store_out(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// grid (tiles_h * tiles_w, C, B).  flip: taps reversed and padding (k-1)*d - p, i.e. the gradient w.r.t. the input
template <typename T>
__global__ void __launch_bounds__(DW_THREADS)
dw_stencil_kernel(const T* __restrict__ in, const float* __restrict__ weight, const float* __restrict__ bias, DwGeom g,
                  int flip, T* __restrict__ out) {
    extern __shared__ float dw_sm[];
    __shared__ float wt[DW_MAX_TAPS];
    __shared__ int toff[DW_MAX_TAPS];
    const int c = blockIdx.y, b = blockIdx.z;
    const int th = blockIdx.x / g.tiles_w, tw = blockIdx.x - th * g.tiles_w;
    const int h0 = th * DW_TH, w0 = tw * DW_TW;
    const int ntaps = g.kh * g.kw;
    if (threadIdx.x < ntaps) {
        const int t = threadIdx.x, i = t / g.kw, j = t - i * g.kw;
        wt[t] = weight[(size_t)c * ntaps + (flip ? ntaps - 1 - t : t)];
        toff[t] = i * g.dh * g.SW + j * g.dw;
    }
    const int ph = flip ? (g.kh - 1) * g.dh - g.ph : g.ph, pw = flip ? (g.kw - 1) * g.dw - g.pw : g.pw;
    const size_t plane = ((size_t)b * g.C + c) * g.H * g.W;
    dw_load_tile(in + plane, g, h0, w0, ph, pw, dw_sm);
    __syncthreads();
    const float bv = bias ? bias[c] : 0.f;
    for (int idx = threadIdx.x; idx < DW_TH * DW_TW; idx += DW_THREADS) {
        const int r = idx / DW_TW, q = idx - r * DW_TW;
        if (h0 + r >= g.H || w0 + q >= g.W) continue;
        const float* s0 = dw_sm + r * g.SW + q;
        float acc = bv;
        for (int t = 0; t < ntaps; ++t) acc = fmaf(wt[t], s0[toff[t]], acc);
        store_out(out + plane + (size_t)(h0 + r) * g.W + (w0 + q), acc);
    }
}

// grid (CL, C), cluster (CL,1,1): dweight[c][t] = sum dy * x(shifted by tap t), dbias[c] = sum dy, over batch and plane
template <typename T, int K>
__global__ void __launch_bounds__(DW_THREADS)
dw_wgrad_kernel(const T* __restrict__ dy, const T* __restrict__ x, DwGeom g, int batch, float* __restrict__ dweight,
                float* __restrict__ dbias, float* __restrict__ partial) {
    extern __shared__ float dw_sm[];
    __shared__ int toff[K];
    const int c = blockIdx.y;
    const int ntaps = g.kh * g.kw;
    if (threadIdx.x < K) {
        const int t = threadIdx.x < ntaps ? threadIdx.x : 0, i = t / g.kw, j = t - i * g.kw;
        toff[threadIdx.x] = i * g.dh * g.SW + j * g.dw;
    }
    float acc[K + 1];
#pragma unroll
    for (int t = 0; t <= K; ++t) acc[t] = 0.f;
    const int tiles = g.tiles_h * g.tiles_w, total = batch * tiles;
    // gridDim.z > 1: several clusters share a channel, each leaves its sums in partial[z][c][K+1] for dw_wgrad_final_kernel
    for (int it = blockIdx.z * gridDim.x + blockIdx.x; it < total; it += gridDim.x * gridDim.z) {
        const int b = it / tiles, tt = it - b * tiles;
        const int th = tt / g.tiles_w, tw = tt - th * g.tiles_w;
        const int h0 = th * DW_TH, w0 = tw * DW_TW;
        const size_t plane = ((size_t)b * g.C + c) * g.H * g.W;
        __syncthreads();                                   // previous tile fully consumed (and toff written)
        dw_load_tile(x + plane, g, h0, w0, g.ph, g.pw, dw_sm);
        __syncthreads();
        for (int idx = threadIdx.x; idx < DW_TH * DW_TW; idx += blockDim.x) {
            const int r = idx / DW_TW, q = idx - r * DW_TW;
            if (h0 + r >= g.H || w0 + q >= g.W) continue;
            const float gy = to_float(dy[plane + (size_t)(h0 + r) * g.W + (w0 + q)]);
            const float* s0 = dw_sm + r * g.SW + q;
#pragma unroll
            for (int t = 0; t < K; ++t)
                if (t < ntaps) acc[t] = fmaf(gy, s0[toff[t]], acc[t]);
            acc[K] += gy;
        }
    }
    double a[K + 1];
#pragma unroll
    for (int t = 0; t <= K; ++t) a[t] = acc[t];
    cluster_sum<K + 1>(a);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (gridDim.z > 1) {
            float* dst = partial + ((size_t)blockIdx.z * g.C + c) * (DW_MAX_TAPS + 1);
            for (int t = 0; t < ntaps; ++t) dst[t] = (float)a[t];
            dst[DW_MAX_TAPS] = (float)a[K];
        } else {
            for (int t = 0; t < ntaps; ++t) dweight[(size_t)c * ntaps + t] = (float)a[t];
            if (dbias) dbias[c] = (float)a[K];
        }
    }
}

// grid (C), 32 threads: sums the Z partial rows of a channel in z order
__global__ void dw_wgrad_final_kernel(const float* __restrict__ partial, int C, int Z, int ntaps, float* __restrict__ dweight,
                                      float* __restrict__ dbias) {
    const int c = blockIdx.x, t = threadIdx.x;
    if (t > ntaps) return;
    const int col = t < ntaps ? t : DW_MAX_TAPS;
    float s = 0.f;
    for (int z = 0; z < Z; ++z) s += partial[((size_t)z * C + c) * (DW_MAX_TAPS + 1) + col];
    if (t < ntaps) dweight[(size_t)c * ntaps + t] = s;
    else if (dbias) dbias[c] = s;
}


// ---- ConvTranspose2d with kernel == 2 * stride (segment.py:435 bottle6_1: 16 -> 4, k8 s4 p2; the k4 s2 p1 of the up blocks) --
// forward only (the library's grouped-direct dgrad kernel takes 0.47 ms for bottle6_1 on 16 images; its backward stays
// with the library).  Every output pixel has exactly 2 x 2 contributing input pixels: iy = (y+p)/s - {0,1}, ky = (y+p)%s + {0,s}.
// One thread = one output pixel, all output channels (<= 8) in registers; weights [cin][cout][k][k] fp32 in shared memory.
constexpr int CT_MAX_COUT = 8;
template <typename T>
__global__ void __launch_bounds__(256)
convt_k2s_kernel(const T* __restrict__ in, const float* __restrict__ weight, const float* __restrict__ bias, int cin, int cout,
                 int hin, int win, int k, int s, int p, int hout, int wout, T* __restrict__ out) {
    extern __shared__ float ct_w[];
    for (int i = threadIdx.x; i < cin * cout * k * k; i += blockDim.x) ct_w[i] = weight[i];
    __syncthreads();
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y, b = blockIdx.z;
    if (x >= wout) return;
    const int iy0 = (y + p) / s, ky0 = (y + p) - iy0 * s, ix0 = (x + p) / s, kx0 = (x + p) - ix0 * s;
    float acc[CT_MAX_COUT];
#pragma unroll
    for (int co = 0; co < CT_MAX_COUT; ++co) acc[co] = (bias && co < cout) ? bias[co] : 0.f;
    const size_t in_plane = (size_t)hin * win;
    const T* inb = in + (size_t)b * cin * in_plane;
#pragma unroll
    for (int ty = 0; ty < 2; ++ty) {
        const int iy = iy0 - ty, ky = ky0 + ty * s;
        if (iy < 0 || iy >= hin) continue;
#pragma unroll
        for (int tx = 0; tx < 2; ++tx) {
            const int ix = ix0 - tx, kx = kx0 + tx * s;
            if (ix < 0 || ix >= win) continue;
            const T* src = inb + (size_t)iy * win + ix;
            const float* wk = ct_w + ky * k + kx;
            for (int ci = 0; ci < cin; ++ci) {
                const float v = to_float(src[ci * in_plane]);
#pragma unroll
                for (int co = 0; co < CT_MAX_COUT; ++co)
                    if (co < cout) acc[co] = fmaf(v, wk[(ci * cout + co) * k * k], acc[co]);
            }
        }
    }
    const size_t out_plane = (size_t)hout * wout;
    T* dst = out + (size_t)b * cout * out_plane + (size_t)y * wout + x;
#pragma unroll
    for (int co = 0; co < CT_MAX_COUT; ++co)
        if (co < cout) store_out(dst + co * out_plane, acc[co]);
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

// ---- isg_bce_sigmoid without a caller workspace and without any allocation ----
// The per-block partial sums and the ticket live in a file-scope device array that every GPU loading the library
// holds until exit (8.3 KB).  In host code `g_bce_static` is only a stand-in for the symbol: its device address is
// looked up once per device with cudaGetSymbolAddress (that device current) and handed to the kernel.
}  // extern "C"
namespace {
constexpr int64_t BCE_STATIC_MAX_N = (int64_t)1 << 26;   // elements per call (256 logit maps of 512 x 512)
constexpr int BCE_STATIC_WORDS = BCE_MAX_BLOCKS + 8;     // doubles: partials, then the ticket
static __device__ __align__(256) double g_bce_static[BCE_STATIC_WORDS];
constexpr int BCE_STATIC_MAX_DEV = 64;
double* g_bce_static_addr[BCE_STATIC_MAX_DEV] = {nullptr};   // device address of the array on each device, once known
}  // namespace
extern "C" {

int64_t isg_bce_sigmoid_static_max_n(void) { return BCE_STATIC_MAX_N; }

int isg_bce_sigmoid_static(const void* logits_dev, int logits_bf16, const float* target_dev, int64_t n, void* dlogits_dev,
                           float* loss_dev, void* stream) {
    if (n <= 0) return fail(ISG_ERR_ARG, "isg_bce_sigmoid_static: the mean of an empty batch is undefined (n=%lld)", (long long)n);
    if (n > BCE_STATIC_MAX_N)
        return fail(ISG_ERR_ARG, "isg_bce_sigmoid_static: n=%lld exceeds the static-scratch limit of %lld elements per call; "
                    "use isg_bce_sigmoid with a caller workspace", (long long)n, (long long)BCE_STATIC_MAX_N);
    if (!logits_dev || !target_dev || !loss_dev) return fail(ISG_ERR_ARG, "isg_bce_sigmoid_static: null argument");
    int dev = -1;
    ISG_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= BCE_STATIC_MAX_DEV) return fail(ISG_ERR_ARG, "isg_bce_sigmoid_static: device %d out of range", dev);
    if (!g_bce_static_addr[dev]) {   // set-up, once per device, with that device current
        double* p = nullptr;
        ISG_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_bce_static));
        if (!p) return fail(ISG_ERR_CUDA, "isg_bce_sigmoid_static: cudaGetSymbolAddress returned a null address");
        g_bce_static_addr[dev] = p;
    }
    double* partial = g_bce_static_addr[dev];
    unsigned* ticket = reinterpret_cast<unsigned*>(partial + BCE_MAX_BLOCKS);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // the array keeps the previous call's contents: the ticket must start at zero (the partials are overwritten)
    ISG_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned), st));
    int64_t blocks = (n + BCE_THREADS * 4 - 1) / (BCE_THREADS * 4);   // the grid of isg_bce_sigmoid: same sum order
    if (blocks > BCE_MAX_BLOCKS) blocks = BCE_MAX_BLOCKS;
    const float inv_n = (float)(1.0 / (double)n);
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

int isg_bn_max_split(void) { return BN_MAX_SPLIT; }

static int bn_check(const char* who, int bf16, int batch, int channels, int64_t hw, const void* a, const void* b,
                    const void* ws) {
    if (!a || !b || !ws) return fail(ISG_ERR_ARG, "%s: null argument", who);
    if (batch < 1 || channels < 1 || channels > 65535 || hw < 1 || (int64_t)batch * hw >= (1ll << 31))
        return fail(ISG_ERR_ARG, "%s: bad sizes (batch=%d channels=%d hw=%lld; batch*hw must be below 2^31)", who, batch,
                    channels, (long long)hw);
    (void)bf16;
    return ISG_OK;
}

static int bn_act_check(const char* who, int act, const float* slope, int slope_count, int channels) {
    if (act < 0 || act > 2) return fail(ISG_ERR_ARG, "%s: act must be 0 (none), 1 (ReLU) or 2 (PReLU)", who);
    if (act == 2 && (!slope || (slope_count != 1 && slope_count != channels)))
        return fail(ISG_ERR_ARG, "%s: PReLU needs 1 or %d slopes, got %d", who, channels, slope_count);
    return ISG_OK;
}

// 16-byte vectors when every channel plane starts 16-byte aligned and holds a whole number of them
static bool bn_wide(int bf16, int64_t hw, const void* p0, const void* p1, const void* p2) {
    const int v = bf16 ? 8 : 4;
    auto al = [](const void* p) { return !p || (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    return hw % v == 0 && al(p0) && al(p1) && al(p2);
}

int isg_bn_train_forward(const void* x_dev, int bf16, int batch, int channels, int64_t hw, const float* gamma_dev,
                         const float* beta_dev, const float* mean_offset_dev, float* running_mean_dev,
                         float* running_var_dev, float momentum, float eps, int act, const float* slope_dev,
                         int slope_count, void* y_dev, float* save_mean_dev, float* save_invstd_dev, void* workspace_dev,
                         void* stream) {
    if (int rc = bn_check("isg_bn_train_forward", bf16, batch, channels, hw, x_dev, y_dev, workspace_dev)) return rc;
    if (!save_mean_dev || !save_invstd_dev) return fail(ISG_ERR_ARG, "isg_bn_train_forward: null argument");
    if (!(eps >= 0.f) || !(momentum >= 0.f && momentum <= 1.f))
        return fail(ISG_ERR_ARG, "isg_bn_train_forward: eps must be >= 0 and momentum in [0,1]");
    if (int rc = bn_act_check("isg_bn_train_forward", act, slope_dev, slope_count, channels)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    double* part = static_cast<double*>(workspace_dev);
    const bool wide = bn_wide(bf16, hw, x_dev, y_dev, nullptr);
    const unsigned HW = (unsigned)hw;
    // one cluster launch, except for planes so large that 8 CTAs per channel cannot stream them as fast as the
    // many-block two-launch kernels do (measured: 33.6 MB / 16 channels 36 vs 27 us); ISG_BN_TWO_PASS / ISG_BN_CLUSTER force
    const bool two_pass = getenv("ISG_BN_TWO_PASS") || (!getenv("ISG_BN_CLUSTER") && (int64_t)batch * hw > (bf16 ? 65536 * 8 : 65536 * 4));
    int hs = -1;
    if ((hw & (hw - 1)) == 0) { hs = 0; while ((1ll << hs) < hw) ++hs; }
    const int sstride = slope_count == 1 ? 0 : 1;
#define ISG_BN_FWD(T, V)                                                                                              \
    do {                                                                                                              \
        const unsigned n_vec = (unsigned)((int64_t)batch * hw / V);                                                   \
        if (!two_pass) {                                                                                              \
            int cl, threads;                                                                                          \
            bn_cluster_shape(n_vec, channels, &cl, &threads);                                                         \
            ISG_CUDA(launch_cluster(bn_fwd_cluster_kernel<T, V>, cl, channels, threads, st,                           \
                                    static_cast<const T*>(x_dev), channels, HW, n_vec, hs, gamma_dev, beta_dev,       \
                                    mean_offset_dev, running_mean_dev, running_var_dev, momentum, eps, act, slope_dev,\
                                    sstride, static_cast<T*>(y_dev), save_mean_dev, save_invstd_dev));                \
            break;                                                                                                    \
        }                                                                                                             \
        dim3 grid(bn_split(channels, n_vec), channels);                                                               \
        bn_stats_kernel<T, V><<<grid, BN_THREADS, 0, st>>>(static_cast<const T*>(x_dev), channels, HW, n_vec, hs, part);   \
        bn_apply_kernel<T, V><<<grid, BN_THREADS, 0, st>>>(static_cast<const T*>(x_dev), channels, HW, n_vec, hs, part,   \
                                                           gamma_dev, beta_dev, mean_offset_dev, running_mean_dev,    \
                                                           running_var_dev, momentum, eps, act, slope_dev, sstride,   \
                                                           static_cast<T*>(y_dev), save_mean_dev, save_invstd_dev);   \
    } while (0)
    if (bf16) { if (wide) ISG_BN_FWD(__nv_bfloat16, 8); else ISG_BN_FWD(__nv_bfloat16, 1); }
    else      { if (wide) ISG_BN_FWD(float, 4); else ISG_BN_FWD(float, 1); }
#undef ISG_BN_FWD
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int isg_bn_train_backward(const void* dy_dev, const void* x_dev, int bf16, int batch, int channels, int64_t hw,
                          const float* gamma_dev, const float* beta_dev, const float* save_mean_dev,
                          const float* save_invstd_dev, int act, const float* slope_dev, int slope_count, void* dx_dev,
                          float* dgamma_dev, float* dbeta_dev, float* dslope_dev, void* workspace_dev, void* stream) {
    if (int rc = bn_check("isg_bn_train_backward", bf16, batch, channels, hw, dy_dev, x_dev, workspace_dev)) return rc;
    if (!save_mean_dev || !save_invstd_dev) return fail(ISG_ERR_ARG, "isg_bn_train_backward: null argument");
    if (int rc = bn_act_check("isg_bn_train_backward", act, slope_dev, slope_count, channels)) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    double* part = static_cast<double*>(workspace_dev);
    const bool wide = bn_wide(bf16, hw, dy_dev, x_dev, dx_dev);
    const unsigned HW = (unsigned)hw;
    // one cluster launch, except for planes so large that 8 CTAs per channel cannot stream them as fast as the
    // many-block two-launch kernels do (measured: 33.6 MB / 16 channels 36 vs 27 us); ISG_BN_TWO_PASS / ISG_BN_CLUSTER force
    const bool two_pass = getenv("ISG_BN_TWO_PASS") || (!getenv("ISG_BN_CLUSTER") && (int64_t)batch * hw > (bf16 ? 65536 * 8 : 65536 * 4));
    int hs = -1;
    if ((hw & (hw - 1)) == 0) { hs = 0; while ((1ll << hs) < hw) ++hs; }
    const int sstride = slope_count == 1 ? 0 : 1;
#define ISG_BN_BWD(T, V)                                                                                              \
    do {                                                                                                              \
        const unsigned n_vec = (unsigned)((int64_t)batch * hw / V);                                                   \
        if (!two_pass) {                                                                                              \
            int cl, threads;                                                                                          \
            bn_cluster_shape(n_vec, channels, &cl, &threads);                                                         \
            ISG_CUDA(launch_cluster(bn_bwd_cluster_kernel<T, V>, cl, channels, threads, st,                           \
                                    static_cast<const T*>(dy_dev), static_cast<const T*>(x_dev), channels, HW, n_vec, \
                                    hs, gamma_dev, beta_dev, save_mean_dev, save_invstd_dev, act, slope_dev, sstride, \
                                    static_cast<T*>(dx_dev), dgamma_dev, dbeta_dev, dslope_dev));                     \
            break;                                                                                                    \
        }                                                                                                             \
        dim3 grid(bn_split(channels, n_vec), channels);                                                               \
        bn_bwd_reduce_kernel<T, V><<<grid, BN_THREADS, 0, st>>>(static_cast<const T*>(dy_dev),                        \
                                                                static_cast<const T*>(x_dev), channels, HW, n_vec, hs,\
                                                                gamma_dev, beta_dev, save_mean_dev, save_invstd_dev,  \
                                                                act, slope_dev, sstride, part);                       \
        bn_bwd_apply_kernel<T, V><<<grid, BN_THREADS, 0, st>>>(static_cast<const T*>(dy_dev),                         \
                                                               static_cast<const T*>(x_dev), channels, HW, n_vec, hs, \
                                                               part, gamma_dev, beta_dev, save_mean_dev,              \
                                                               save_invstd_dev, act, slope_dev, sstride,              \
                                                               static_cast<T*>(dx_dev), dgamma_dev, dbeta_dev,        \
                                                               dslope_dev);                                           \
    } while (0)
    if (bf16) { if (wide) ISG_BN_BWD(__nv_bfloat16, 8); else ISG_BN_BWD(__nv_bfloat16, 1); }
    else      { if (wide) ISG_BN_BWD(float, 4); else ISG_BN_BWD(float, 1); }
#undef ISG_BN_BWD
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

static int dw_geom(const char* who, int batch, int channels, int height, int width, int kh, int kw, int pad_h, int pad_w,
                   int dil_h, int dil_w, DwGeom* g, size_t* smem) {
    if (batch < 1 || channels < 1 || channels > 65535 || batch > 65535 || height < 1 || width < 1)
        return fail(ISG_ERR_ARG, "%s: bad sizes (batch=%d channels=%d %dx%d)", who, batch, channels, height, width);
    if (kh < 1 || kw < 1 || kh * kw > DW_MAX_TAPS || dil_h < 1 || dil_w < 1 || 2 * pad_h != dil_h * (kh - 1) ||
        2 * pad_w != dil_w * (kw - 1))
        return fail(ISG_ERR_ARG, "%s: kernel %dx%d pad %d,%d dilation %d,%d: need <= %d taps and 2*pad == dil*(k-1) "
                    "(stride 1, output size == input size)", who, kh, kw, pad_h, pad_w, dil_h, dil_w, DW_MAX_TAPS);
    g->C = channels; g->H = height; g->W = width; g->kh = kh; g->kw = kw; g->ph = pad_h; g->pw = pad_w; g->dh = dil_h;
    g->dw = dil_w; g->SH = DW_TH + (kh - 1) * dil_h; g->SW = DW_TW + (kw - 1) * dil_w;
    g->tiles_h = (height + DW_TH - 1) / DW_TH; g->tiles_w = (width + DW_TW - 1) / DW_TW;
    *smem = (size_t)g->SH * g->SW * sizeof(float);
    if (*smem > 40 * 1024) return fail(ISG_ERR_ARG, "%s: dilated window of %zu bytes does not fit the tile buffer", who, *smem);
    return ISG_OK;
}

int isg_dwconv_forward(const void* in_dev, int bf16, int batch, int channels, int height, int width, const float* weight_dev,
                       const float* bias_dev, int kh, int kw, int pad_h, int pad_w, int dil_h, int dil_w, int flip,
                       void* out_dev, void* stream) {
    if (!in_dev || !weight_dev || !out_dev) return fail(ISG_ERR_ARG, "isg_dwconv_forward: null argument");
    DwGeom g; size_t smem;
    if (int rc = dw_geom("isg_dwconv_forward", batch, channels, height, width, kh, kw, pad_h, pad_w, dil_h, dil_w, &g, &smem)) return rc;
    dim3 grid(g.tiles_h * g.tiles_w, channels, batch);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (bf16)
        dw_stencil_kernel<__nv_bfloat16><<<grid, DW_THREADS, smem, st>>>(static_cast<const __nv_bfloat16*>(in_dev), weight_dev,
                                                                        bias_dev, g, flip ? 1 : 0, static_cast<__nv_bfloat16*>(out_dev));
    else
        dw_stencil_kernel<float><<<grid, DW_THREADS, smem, st>>>(static_cast<const float*>(in_dev), weight_dev, bias_dev, g,
                                                                flip ? 1 : 0, static_cast<float*>(out_dev));
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

int64_t isg_dwconv_wgrad_workspace_bytes(int channels) {
    return (int64_t)BN_CL_MAX * (channels > 0 ? channels : 0) * (DW_MAX_TAPS + 1) * (int64_t)sizeof(float);
}

int isg_dwconv_wgrad(const void* dy_dev, const void* x_dev, int bf16, int batch, int channels, int height, int width, int kh,
                     int kw, int pad_h, int pad_w, int dil_h, int dil_w, float* dweight_dev, float* dbias_dev,
                     void* workspace_dev, void* stream) {
    if (!dy_dev || !x_dev || !dweight_dev) return fail(ISG_ERR_ARG, "isg_dwconv_wgrad: null argument");
    DwGeom g; size_t smem;
    if (int rc = dw_geom("isg_dwconv_wgrad", batch, channels, height, width, kh, kw, pad_h, pad_w, dil_h, dil_w, &g, &smem)) return rc;
    const int total = batch * g.tiles_h * g.tiles_w;
    int cl = 1;
    while (cl < BN_CL_MAX && cl * 2 <= total) cl *= 2;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof cfg);
    // few channels: several clusters per channel (z) so that ~2 CTAs per SM are in flight; needs the workspace
    int zs = 1;
    if (workspace_dev)
        while (zs < BN_CL_MAX && cl * channels * zs < 148 * 2 && cl * zs * 2 <= total) zs *= 2;
    float* partial = static_cast<float*>(workspace_dev);
    cfg.gridDim = dim3(cl, channels, zs);
    cfg.blockDim = dim3(DW_THREADS);   // wider CTAs measured slower (81 vs 35 us per call at 1024 threads)
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = cl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    const bool small = kh * kw <= 9;
#define ISG_DW_WGRAD(T, K)                                                                                        \
    ISG_CUDA(cudaLaunchKernelEx(&cfg, dw_wgrad_kernel<T, K>, static_cast<const T*>(dy_dev), static_cast<const T*>(x_dev), \
                                g, batch, dweight_dev, dbias_dev, partial))
    if (bf16) { if (small) ISG_DW_WGRAD(__nv_bfloat16, 9); else ISG_DW_WGRAD(__nv_bfloat16, 25); }
    else      { if (small) ISG_DW_WGRAD(float, 9); else ISG_DW_WGRAD(float, 25); }
#undef ISG_DW_WGRAD
    if (zs > 1) {
        dw_wgrad_final_kernel<<<channels, 32, 0, st>>>(partial, channels, zs, kh * kw, dweight_dev, dbias_dev);
        ISG_CUDA(cudaGetLastError());
    }
    return ISG_OK;
}

int isg_convt_k2s_forward(const void* in_dev, int bf16, int batch, int cin, int cout, int hin, int win, int k, int s, int p,
                          const float* weight_dev, const float* bias_dev, void* out_dev, void* stream) {
    if (!in_dev || !weight_dev || !out_dev) return fail(ISG_ERR_ARG, "isg_convt_k2s_forward: null argument");
    if (batch < 1 || batch > 65535 || cin < 1 || cout < 1 || cout > CT_MAX_COUT || hin < 1 || win < 1 || s < 1 || k != 2 * s ||
        p < 0 || p >= s || (size_t)cin * cout * k * k * 4 > 40 * 1024)
        return fail(ISG_ERR_ARG, "isg_convt_k2s_forward: need kernel == 2*stride, 0 <= pad < stride, cout <= %d and weights "
                    "<= 40 KB (batch=%d cin=%d cout=%d %dx%d k=%d s=%d p=%d)", CT_MAX_COUT, batch, cin, cout, hin, win, k, s, p);
    const int hout = (hin - 1) * s - 2 * p + k, wout = (win - 1) * s - 2 * p + k;
    if (hout > 65535) return fail(ISG_ERR_ARG, "isg_convt_k2s_forward: output height %d too large", hout);
    dim3 grid((wout + 255) / 256, hout, batch);
    const size_t smem = (size_t)cin * cout * k * k * 4;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (bf16)
        convt_k2s_kernel<__nv_bfloat16><<<grid, 256, smem, st>>>(static_cast<const __nv_bfloat16*>(in_dev), weight_dev, bias_dev,
                                                                cin, cout, hin, win, k, s, p, hout, wout,
                                                                static_cast<__nv_bfloat16*>(out_dev));
    else
        convt_k2s_kernel<float><<<grid, 256, smem, st>>>(static_cast<const float*>(in_dev), weight_dev, bias_dev, cin, cout, hin,
                                                        win, k, s, p, hout, wout, static_cast<float*>(out_dev));
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

}  // extern "C"
