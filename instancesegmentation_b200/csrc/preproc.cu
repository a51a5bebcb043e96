// Input side of the hot path: the per-instance crop / pad / resize to the network
// size (geometry of reference train_instance.py:160-196) and the keypoint heat maps of the Segment(20)
// variant (train_instance.py:33-68), both written straight into the tensors the first conv reads.
#include <climits>
#include <cmath>

#include "decode_math.cuh"

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


// ---- per-instance crop / pad / resize to the network size -----------------------------------------
// One warp = CR_ROWS consecutive output rows, lanes = consecutive output pixels.  Windows are usually
// upsampled (instance boxes are smaller than the network input), so neighbouring output rows share
// their pair of source rows: the horizontal part of the bilinear form (lx0*a + lx1*b per source row and
// channel) is computed once per distinct source-row pair and only the vertical mix runs per output row.
// The 3 bytes of a pixel go to a per-warp 32-pixel chunk in shared memory and every chunk leaves as
// 16-byte streaming stores of whole 32-byte sectors (the network input is written once; the source
// windows are small and stay in L2).  Arithmetic = the decode's exact bilinear sequence on the window
// (decode_math.cuh make_tap / exact_lerp2 order), rounded half-to-even: bit-identical to
// oracle_crop_resize.
constexpr int CR_THREADS = 256, CR_WARPS = CR_THREADS / 32, CR_ROWS = 4;

__device__ __forceinline__ void ld_px3(const uint8_t* __restrict__ row, int x, int Ws, float (&v)[3]) {
    if (row && x >= 0 && x < Ws) {
        const uint8_t* q = row + 3 * x;
        // byte -> float through the mantissa of 2^23 (exact; keeps the conversions off the quarter-rate pipe)
        v[0] = __uint_as_float(0x4B000000u | __ldg(q)) - 8388608.0f;
        v[1] = __uint_as_float(0x4B000000u | __ldg(q + 1)) - 8388608.0f;
        v[2] = __uint_as_float(0x4B000000u | __ldg(q + 2)) - 8388608.0f;
    } else { v[0] = v[1] = v[2] = 0.f; }
}

// round half to even for v in [0, 255.5): adding 1.5 * 2^23 leaves the rounded integer in the low mantissa
// bits (the fp32 add itself rounds to nearest-even at unit spacing); the byte store keeps the low 8 bits
__device__ __forceinline__ unsigned sat_u8(float v) { return __float_as_uint(__fadd_rn(v, 12582912.0f)); }

__global__ void __launch_bounds__(CR_THREADS)
crop_resize3_kernel(const uint8_t* __restrict__ src, int Hs, int Ws, const int32_t* __restrict__ win,
                    const int32_t* __restrict__ inst_img, int S_h, int S_w, uint8_t* __restrict__ out, int n_img) {
    __shared__ __align__(16) uint8_t chunk[CR_WARPS][CR_ROWS][96];   // 32 pixels x 3 bytes of each of the warp's rows
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n = blockIdx.y, oy0 = (blockIdx.x * CR_WARPS + warp) * CR_ROWS;
    if (oy0 >= S_h) return;
    const int4 wn = __ldg(reinterpret_cast<const int4*>(win) + n);
    // an image index outside [0, n_img) is clamped: never a read outside the source batch
    const uint8_t* img = src + (size_t)min(max(__ldg(inst_img + n), 0), n_img - 1) * Hs * Ws * 3;
    const float sy = __fdiv_rn((float)wn.w, (float)S_h), sx = __fdiv_rn((float)wn.z, (float)S_w);
    const int nrows = min(CR_ROWS, S_h - oy0);
    Tap ty[CR_ROWS];
#pragma unroll
    for (int j = 0; j < CR_ROWS; ++j) ty[j] = make_tap(min(oy0 + j, S_h - 1), wn.w, sy);
    const int row_bytes = S_w * 3;
    uint8_t* orow = out + ((size_t)n * S_h + oy0) * row_bytes;
    // the 16-byte piece of the chunk this lane stores: row sj, bytes [16 sq, 16 sq + 16)
    const int sj = lane / 6, sq = lane - sj * 6;
    for (int x0 = 0; x0 < S_w; x0 += 32) {
        const int ox = x0 + lane;
        if (ox < S_w) {
            const Tap tx = make_tap(ox, wn.z, sx);
            const int xa = wn.x + tx.i0, xb = wn.x + tx.i1;
            float top[3] = {0.f, 0.f, 0.f}, bot[3] = {0.f, 0.f, 0.f};
            int pa = INT_MIN, pb = INT_MIN;
#pragma unroll
            for (int j = 0; j < CR_ROWS; ++j) {
                if (ty[j].i0 != pa || ty[j].i1 != pb) {      // warp-uniform: a new pair of source rows
                    pa = ty[j].i0; pb = ty[j].i1;
                    const int ya = wn.y + pa, yb = wn.y + pb;
                    const uint8_t* r0 = (ya >= 0 && ya < Hs) ? img + (size_t)ya * Ws * 3 : nullptr;
                    const uint8_t* r1 = (yb >= 0 && yb < Hs) ? img + (size_t)yb * Ws * 3 : nullptr;
                    float a[3], b[3], c[3], d[3];
                    ld_px3(r0, xa, Ws, a); ld_px3(r0, xb, Ws, b); ld_px3(r1, xa, Ws, c); ld_px3(r1, xb, Ws, d);
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch) {
                        top[ch] = __fadd_rn(__fmul_rn(tx.l0, a[ch]), __fmul_rn(tx.l1, b[ch]));
                        bot[ch] = __fadd_rn(__fmul_rn(tx.l0, c[ch]), __fmul_rn(tx.l1, d[ch]));
                    }
                }
                uint8_t* o = &chunk[warp][j][3 * lane];
#pragma unroll
                for (int ch = 0; ch < 3; ++ch)
                    o[ch] = (uint8_t)sat_u8(__fadd_rn(__fmul_rn(ty[j].l0, top[ch]), __fmul_rn(ty[j].l1, bot[ch])));
            }
        }
        __syncwarp();
        // 32 pixels = 96 bytes = three full 32-byte sectors of every row (rows and chunks are 32-byte aligned)
        if (sj < nrows && 3 * x0 + 16 * sq < row_bytes && lane < 6 * CR_ROWS)
            __stcs(reinterpret_cast<uint4*>(orow + (size_t)sj * row_bytes + 3 * x0 + 16 * sq),
                   *reinterpret_cast<const uint4*>(&chunk[warp][sj][16 * sq]));
        __syncwarp();
    }
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


int isg_crop_resize_u8(const uint8_t* src_dev, int n_img, int src_h, int src_w, const int32_t* win_dev,
                       const int32_t* inst_img_dev, int n_inst, int out_h, int out_w, uint8_t* out_dev, void* stream) {
    if (n_inst == 0) return ISG_OK;
    if (!src_dev || !win_dev || !inst_img_dev || !out_dev) return fail(ISG_ERR_ARG, "isg_crop_resize_u8: null argument");
    if (n_img < 1 || src_h < 1 || src_w < 1 || n_inst < 0 || n_inst > 65535 || out_h < 1 || out_w < 16 || (out_w & 15))
        return fail(ISG_ERR_ARG, "isg_crop_resize_u8: bad sizes (n_img=%d src %dx%d n_inst=%d out %dx%d; out_w must be a multiple of 16)",
                    n_img, src_h, src_w, n_inst, out_h, out_w);
    if (reinterpret_cast<uintptr_t>(out_dev) & 15) return fail(ISG_ERR_ARG, "isg_crop_resize_u8: output must be 16-byte aligned");
    dim3 grid(cdiv(out_h, CR_WARPS * CR_ROWS), n_inst);
    crop_resize3_kernel<<<grid, CR_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(src_dev, src_h, src_w, win_dev,
                                                                                      inst_img_dev, out_h, out_w, out_dev, n_img);
    ISG_CUDA(cudaGetLastError());
    return ISG_OK;
}

}  // extern "C"
