// Input side of the hot path: the per-instance crop / pad / resize to the network
// size (geometry of reference train_instance.py:160-196) and the keypoint heat maps of the Segment(20)
// variant (train_instance.py:33-68), both written straight into the tensors the first conv reads.
#include <cmath>

#include "common.cuh"

namespace isg {
namespace {

// ---- keypoint heat maps ---------------------------------------------------------------------------
// One thread = 4 consecutive pixels of one (image, part) plane, one 16-byte store; pixels outside the
// part's disk box (almost all) cost two integer compares.  Inside, the reference's float64 sequence:
// e = exp(-((x - kx)^2 + (y - ky)^2) / sigma^2), kept where e > threshold, rounded to fp32 on store.
struct PartBox { int x0, x1, y0, y1; };   // [x0, x1) x [y0, y1), empty when the part is not drawn

__global__ void __launch_bounds__(256)
heatmap_kernel(const float* __restrict__ kp, int n_parts, int H, int W, double sigma2, double threshold, double r,
               float* __restrict__ out, int64_t batch_stride, int64_t total4) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= total4) return;
    const int w4 = W >> 2;
    const int xq = (int)(i % w4);
    const int64_t rest = i / w4;
    const int y = (int)(rest % H);
    const int64_t plane = rest / H;
    const int part = (int)(plane % n_parts);
    const int64_t b = plane / n_parts;
    const float* k = kp + (b * n_parts + part) * 3;
    const float kx = __ldg(k), ky = __ldg(k + 1), flag = __ldg(k + 2);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (flag == 1.f) {
        const double x = kx, yy = ky;
        // int() of the reference truncates toward zero; coordinates far outside are clamped before the cast
        const double lim = 1e9;
        const int x_min = max(0, (int)fmin(fmax(x - r, -lim), lim)), x_max = min(W - 1, (int)fmin(fmax(x + r + 1.0, -lim), lim));
        const int y_min = max(0, (int)fmin(fmax(yy - r, -lim), lim)), y_max = min(H - 1, (int)fmin(fmax(yy + r + 1.0, -lim), lim));
        const int xa = xq * 4;
        if (y >= y_min && y < y_max && xa + 4 > x_min && xa < x_max) {
            const double dy = (double)y - yy, dy2 = dy * dy;
            float e4[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int xx = xa + e;
                float val = 0.f;
                if (xx >= x_min && xx < x_max) {
                    const double dx = (double)xx - x;
                    const double ev = exp(-(dx * dx + dy2) / sigma2);
                    if (ev > threshold) val = (float)ev;
                }
                e4[e] = val;
            }
            v = make_float4(e4[0], e4[1], e4[2], e4[3]);
        }
    }
    float* dst = out + b * batch_stride + ((int64_t)part * H + y) * W + xq * 4;
    *reinterpret_cast<float4*>(dst) = v;
}

}  // namespace
}  // namespace isg

using namespace isg;

extern "C" {

int isg_keypoint_heatmaps(const float* kp_dev, int batch, int n_parts, int height, int width, float sigma,
                          float threshold, float* out_dev, int64_t out_batch_stride, void* stream) {
    if (batch == 0) return ISG_OK;
    if (!kp_dev || !out_dev) return fail(ISG_ERR_ARG, "isg_keypoint_heatmaps: null argument");
    if (batch < 0 || n_parts < 1 || height < 1 || width < 4 || (width & 3))
        return fail(ISG_ERR_ARG, "isg_keypoint_heatmaps: bad sizes (batch=%d parts=%d %dx%d; width must be a multiple of 4)",
                    batch, n_parts, height, width);
    if (!(sigma > 0.f) || !(threshold > 0.f && threshold < 1.f))
        return fail(ISG_ERR_ARG, "isg_keypoint_heatmaps: sigma must be > 0 and threshold in (0,1)");
    if (out_batch_stride < (int64_t)n_parts * height * width || (out_batch_stride & 3) ||
        (reinterpret_cast<uintptr_t>(out_dev) & 15))
        return fail(ISG_ERR_ARG, "isg_keypoint_heatmaps: output must be 16-byte aligned with a batch stride >= parts*H*W, multiple of 4");
    const double s2 = (double)sigma * (double)sigma;
    const double r = sqrt(log((double)threshold) * (-s2));   // train_instance.py:35
    const int64_t total4 = (int64_t)batch * n_parts * height * (width / 4);
    heatmap_kernel<<<(unsigned)((total4 + 255) / 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        kp_dev, n_parts, height, width, s2, (double)threshold, r, out_dev, out_batch_stride, total4);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

}  // extern "C"
