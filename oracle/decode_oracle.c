/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the mask decode.
 *
 * PARITY UNPINNED for rows D2/D3: the reference has no decode code (infer.py:32-36
 * is a stub) and no tests; the only reference fragments are torch.sigmoid
 * (model/segment.py:534) and the soft-mask cast (train_instance.py:398-399).  The
 * resize follows the published F.interpolate(mode="bilinear", align_corners=False)
 * index rule; area / bbox / id-map conventions are defined in DESIGN.md section 3.
 * tests/test_oracle.py cross-checks this file against torch.sigmoid and
 * F.interpolate (mismatch counts reported, masks must agree away from ulp ties).
 *
 * Every fp32 operation below is a single IEEE-754 round-to-nearest +,-,*,/ (build
 * with -ffp-contract=off, no -ffast-math) so a CUDA kernel that issues the same
 * sequence with __f*_rn intrinsics is bit-identical by construction.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library; the product never does.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* exp(y) for y <= 0, Cephes-style: y = n ln2 + r, degree-6 polynomial, exact 2^n scale */
static float exp_nonpos(float y) {
    if (y < -87.0f) return 0.0f;
    float t = y * 1.44269504088896341f;
    float n = floorf(t + 0.5f);
    float r = y - n * 0.693359375f;
    r = r - n * -2.12194440e-4f;
    float p = 1.9875691500e-4f;
    p = p * r + 1.3981999507e-3f;
    p = p * r + 8.3334519073e-3f;
    p = p * r + 4.1665795894e-2f;
    p = p * r + 1.6666665459e-1f;
    p = p * r + 5.0000001201e-1f;
    float rr = r * r;
    float e = p * rr;
    e = e + r;
    e = e + 1.0f;
    uint32_t bits = (uint32_t)((int32_t)n + 127) << 23;
    float s;
    memcpy(&s, &bits, 4);
    return e * s;
}

/* torch.sigmoid (segment.py:534): 1/(1+exp(-x)), evaluated on |x| for stability */
float oracle_sigmoid(float x) {
    if (x != x) return x;
    float e = exp_nonpos(-fabsf(x));
    float d = 1.0f + e;
    return x >= 0.0f ? 1.0f / d : e / d;
}

typedef struct { int i0, i1; float l0, l1; } tap_t;

/* align_corners=False source index: max(0, scale*(dst+0.5)-0.5), scale = in/out */
static tap_t tap(int dst, int in, int out) {
    tap_t t;
    float scale = (float)in / (float)out;
    float s = scale * ((float)dst + 0.5f) - 0.5f;
    if (s < 0.0f) s = 0.0f;
    t.i0 = (int)s;
    if (t.i0 > in - 1) t.i0 = in - 1;
    t.i1 = t.i0 + (t.i0 < in - 1 ? 1 : 0);
    t.l1 = s - (float)t.i0;
    t.l0 = 1.0f - t.l1;
    return t;
}

/* probability of instance map L (h x w) resampled at (oy, ox) of an hk x wk grid */
float oracle_prob(const float* L, int h, int w, int hk, int wk, int oy, int ox) {
    tap_t ty = tap(oy, h, hk), tx = tap(ox, w, wk);
    float a = oracle_sigmoid(L[(int64_t)ty.i0 * w + tx.i0]);
    float b = oracle_sigmoid(L[(int64_t)ty.i0 * w + tx.i1]);
    float c = oracle_sigmoid(L[(int64_t)ty.i1 * w + tx.i0]);
    float d = oracle_sigmoid(L[(int64_t)ty.i1 * w + tx.i1]);
    float top = tx.l0 * a + tx.l1 * b;
    float bot = tx.l0 * c + tx.l1 * d;
    return ty.l0 * top + ty.l1 * bot;
}

/* train_instance.py:398-399: (p*255) -> truncating uint8 cast; NaN -> 0 by definition */
static uint8_t soft_u8(float p) {
    float v = p * 255.0f;
    if (!(v >= 0.0f)) return 0;
    if (v > 255.0f) v = 255.0f;
    return (uint8_t)(int)v;
}

/* Per-instance decode.  boxes[n] = {x0, y0, wk, hk} (placement in the source canvas and
 * output size); masks of instance n start at masks + mask_off[n], tightly packed
 * hk x wk bytes (0/1).  soft may be NULL.  bbox[n] = {xmin, ymin, xmax+1, ymax+1} in
 * canvas coordinates, {0,0,0,0} when the mask is empty. */
void oracle_decode(const float* logits, int N, int h, int w, const int32_t* boxes,
                   const int64_t* mask_off, float thr, uint8_t* masks, uint8_t* soft,
                   int32_t* area, int32_t* bbox) {
    for (int n = 0; n < N; ++n) {
        const float* L = logits + (int64_t)n * h * w;
        int x0 = boxes[4 * n], y0 = boxes[4 * n + 1], wk = boxes[4 * n + 2], hk = boxes[4 * n + 3];
        int a = 0, xmin = INT32_MAX, ymin = INT32_MAX, xmax = -1, ymax = -1;
        for (int oy = 0; oy < hk; ++oy)
            for (int ox = 0; ox < wk; ++ox) {
                float p = oracle_prob(L, h, w, hk, wk, oy, ox);
                int m = p > thr;
                int64_t o = mask_off[n] + (int64_t)oy * wk + ox;
                masks[o] = (uint8_t)m;
                if (soft) soft[o] = soft_u8(p);
                if (m) {
                    ++a;
                    if (ox < xmin) xmin = ox;
                    if (ox > xmax) xmax = ox;
                    if (oy < ymin) ymin = oy;
                    if (oy > ymax) ymax = oy;
                }
            }
        area[n] = a;
        if (a) {
            bbox[4 * n] = x0 + xmin; bbox[4 * n + 1] = y0 + ymin;
            bbox[4 * n + 2] = x0 + xmax + 1; bbox[4 * n + 3] = y0 + ymax + 1;
        } else {
            bbox[4 * n] = bbox[4 * n + 1] = bbox[4 * n + 2] = bbox[4 * n + 3] = 0;
        }
    }
}

/* Id map per source image: instances img_first[i] .. img_first[i+1]-1 belong to image i.
 * A canvas pixel takes id n+1 of the instance whose box covers it, whose mask is set
 * there and whose probability is largest (ties -> lowest n); 0 = background.  Parts of
 * a box outside the Hs x Ws canvas are ignored. */
void oracle_idmap(const float* logits, int h, int w, const int32_t* boxes,
                  const int32_t* img_first, int n_img, int Hs, int Ws, float thr, int32_t* idmap) {
    for (int i = 0; i < n_img; ++i)
        for (int y = 0; y < Hs; ++y)
            for (int x = 0; x < Ws; ++x) {
                int best = 0;
                float bp = 0.0f;
                for (int n = img_first[i]; n < img_first[i + 1]; ++n) {
                    int ox = x - boxes[4 * n], oy = y - boxes[4 * n + 1];
                    int wk = boxes[4 * n + 2], hk = boxes[4 * n + 3];
                    if (ox < 0 || oy < 0 || ox >= wk || oy >= hk) continue;
                    float p = oracle_prob(logits + (int64_t)n * h * w, h, w, hk, wk, oy, ox);
                    if (p > thr && (best == 0 || p > bp)) { best = n + 1; bp = p; }
                }
                idmap[((int64_t)i * Hs + y) * Ws + x] = best;
            }
}

/* mask_iou analogue used by the reference's validation (train_instance.py:402-403; ymlib
 * source unavailable): intersection / union of two 0/1 byte masks, 0 when both empty. */
double oracle_mask_iou(const uint8_t* a, const uint8_t* b, int64_t n) {
    int64_t inter = 0, uni = 0;
    for (int64_t i = 0; i < n; ++i) {
        inter += (a[i] && b[i]);
        uni += (a[i] || b[i]);
    }
    return uni ? (double)inter / (double)uni : 0.0;
}

/* Input side, PARITY UNPINNED (imgaug is absent): per-instance crop window
 * win[n] = {x0, y0, wk, hk} of source image inst_img[n] (uint8 HWC, C channels), zero outside
 * the image (iaa.CropAndPad pads with 0, train_instance.py:177-181), resized to S_h x S_w
 * (iaa.Resize, :182-183) with the same align_corners=False bilinear rule as the decode, the
 * window being the resize input (edge taps clamp at the window, not at the image); value =
 * ly0*(lx0*a + lx1*b) + ly1*(lx0*c + lx1*d) in single IEEE operations, rounded half-to-even
 * to uint8. */
static float src_px(const uint8_t* img, int Hs, int Ws, int C, int y, int x, int c) {
    if (x < 0 || y < 0 || x >= Ws || y >= Hs) return 0.0f;
    return (float)img[((int64_t)y * Ws + x) * C + c];
}

void oracle_crop_resize(const uint8_t* src, int Hs, int Ws, int C, const int32_t* win,
                        const int32_t* inst_img, int n_inst, int S_h, int S_w, uint8_t* out) {
    for (int n = 0; n < n_inst; ++n) {
        const uint8_t* img = src + (int64_t)inst_img[n] * Hs * Ws * C;
        int x0 = win[4 * n], y0 = win[4 * n + 1], wk = win[4 * n + 2], hk = win[4 * n + 3];
        for (int oy = 0; oy < S_h; ++oy) {
            tap_t ty = tap(oy, hk, S_h);
            for (int ox = 0; ox < S_w; ++ox) {
                tap_t tx = tap(ox, wk, S_w);
                for (int c = 0; c < C; ++c) {
                    float a = src_px(img, Hs, Ws, C, y0 + ty.i0, x0 + tx.i0, c);
                    float b = src_px(img, Hs, Ws, C, y0 + ty.i0, x0 + tx.i1, c);
                    float cc = src_px(img, Hs, Ws, C, y0 + ty.i1, x0 + tx.i0, c);
                    float d = src_px(img, Hs, Ws, C, y0 + ty.i1, x0 + tx.i1, c);
                    float top = tx.l0 * a + tx.l1 * b;
                    float bot = tx.l0 * cc + tx.l1 * d;
                    float v = ty.l0 * top + ty.l1 * bot;
                    float r = nearbyintf(v);          /* default rounding mode: half to even */
                    if (r > 255.0f) r = 255.0f;
                    if (r < 0.0f) r = 0.0f;
                    out[(((int64_t)n * S_h + oy) * S_w + ox) * C + c] = (uint8_t)r;
                }
            }
        }
    }
}
