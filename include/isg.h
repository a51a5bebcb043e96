/* libisg -- C ABI of the B200-native (sm_100a) hot path of YanMiaoW/instanceSegmentation:
 * the forward of the network in model/segment.py and the mask decode meant for infer.py.
 *
 * The reference has no FFI seam (pure Python over torch library ops); the entry points
 * below are what a binding for this path would call, each citing the reference code it
 * replaces.  INTEGRATION.md shows the ctypes stub.
 *
 * Conventions
 *  - every pointer named *_dev is a caller-owned device pointer on the plan's device;
 *    pointers named *_host are host pointers; nothing crosses the ABI by ownership except
 *    the opaque plan handle;
 *  - `stream` is a cudaStream_t passed as void*; all work is asynchronous on it;
 *  - return value: 0 = ok, negative = isg_status; isg_last_error() gives the message of
 *    the calling thread's last failure; no C++ exception crosses the boundary;
 *  - a plan is used by one host thread at a time; plans on different devices are
 *    independent (one process or thread per GPU, no collective on the inference path).
 */
#ifndef ISG_H_
#define ISG_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define ISG_API __attribute__((visibility("default")))
#else
#define ISG_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

typedef struct isg_plan isg_plan;

enum isg_status {
    ISG_OK = 0,
    ISG_ERR_ARG = -1,      /* bad argument (size rule, null pointer, count mismatch) */
    ISG_ERR_CUDA = -2,     /* a CUDA runtime / driver call failed */
    ISG_ERR_STATE = -3,    /* call order (forward before weights were loaded, ...) */
    ISG_ERR_ARCH = -4      /* device is not sm_100 */
};

enum isg_mode {
    ISG_MODE_FP32 = 0,     /* fp32 storage + fp32 FMA, BatchNorm as fp32 epilogue affine (<=1e-4 vs reference) */
    ISG_MODE_BF16 = 1,     /* NHWC bf16 activations; GEMM-shaped convs on tcgen05 with fp32 TMEM accumulators
                              (logits 0.3-0.8 max-abs from the reference: bf16 operand rounding) */
    ISG_MODE_BF16X3 = 2    /* the same tcgen05 kernels with split-precision operands: every activation and weight is a bf16
                              pair hi + lo, every GEMM accumulates A_hi.W_hi + A_lo.W_hi + A_hi.W_lo in fp32 TMEM
                              (~16 significand bits per operand; logits within max-abs 2e-2 / mean-abs 2e-3) */
};

ISG_API const char* isg_last_error(void);
ISG_API int isg_version(void);

/* ---- network description: the 588 state_dict entries of Segment(in_channel) ----------
 * (model/segment.py:347-440; key grammar in SURVEY.md section 2.3).  Index order equals
 * the reference's state_dict() order. */
ISG_API int isg_state_count(int in_channel);
ISG_API const char* isg_state_name(int in_channel, int index);
ISG_API int64_t isg_state_numel(int in_channel, int index);
ISG_API int isg_state_is_int64(int in_channel, int index); /* 1 for *.num_batches_tracked */

/* ---- plan: replaces Segment.__init__ + .to(device) + .eval() (train_instance.py:294,350,396) */
ISG_API int isg_plan_create(int device, int in_channel, int height, int width, int max_batch,
                    int mode, isg_plan** out);
ISG_API int isg_plan_destroy(isg_plan* plan);
ISG_API size_t isg_plan_workspace_bytes(const isg_plan* plan);

/* replaces model.load_state_dict(checkpoint["state_dict"]) (train_instance.py:325):
 * tensors_host[i] points at the contiguous host data of state entry i (fp32, or int64 for
 * num_batches_tracked, which is ignored); numels[i] is checked against isg_state_numel.
 * BatchNorm folding / weight packing for the kernels happens here. */
ISG_API int isg_plan_load_weights(isg_plan* plan, const void* const* tensors_host,
                          const int64_t* numels, int count);

/* replaces Segment.forward (model/segment.py:466-508): x_dev is fp32 NCHW
 * [batch,in_channel,H,W]; logits_dev receives fp32 [batch,1,H,W] raw logits.  Any batch
 * size is accepted; it is processed in groups of max_batch. */
ISG_API int isg_forward(isg_plan* plan, const float* x_dev, int batch, float* logits_dev, void* stream);

/* the same forward from what an image decoder produces: x_u8_dev is uint8 NHWC [batch,H,W,in_channel];
 * ToTensor + Normalize(mean, std) of the reference's data path (train_instance.py:80-85:
 * (v / 255 - mean[c]) / std[c], fp32, IEEE division) run inside the first kernel, so a batch costs a
 * quarter of the PCIe bytes of the fp32 tensor.  mean_host / std_host: in_channel floats on the host. */
ISG_API int isg_forward_u8(isg_plan* plan, const uint8_t* x_u8_dev, int batch, const float* mean_host,
                           const float* std_host, float* logits_dev, void* stream);

/* test hook: lowers Segment(in_channel) at this size and mode to its launch list WITHOUT touching a device (tile
 * feasibility of every fused launch: shared memory, TMEM columns, descriptor tables); returns the number of launches
 * per forward or a negative isg_status.  Runs on a machine without a GPU. */
ISG_API int isg_plan_dry_run(int in_channel, int height, int width, int max_batch, int mode);

/* number of kernels isg_forward launches per group of <= max_batch images */
ISG_API int isg_plan_launches_per_forward(const isg_plan* plan);
/* conv/convT multiply-accumulates per image of the reference graph (SURVEY 8d constant) */
ISG_API int64_t isg_plan_macs_per_image(const isg_plan* plan);

/* test hooks: named intermediate activations of the last forward group, copied out as
 * fp32 NCHW [n,C,h,w] (blocking).  Names follow oracle/segment_oracle.py taps. */
ISG_API int isg_plan_tap_count(const isg_plan* plan);
ISG_API const char* isg_plan_tap_name(const isg_plan* plan, int index);
ISG_API int isg_plan_tap_shape(const isg_plan* plan, int index, int* channels, int* height, int* width);
ISG_API int isg_plan_tap_read(isg_plan* plan, int index, int n, float* out_host);

/* ---- mask decode (the body infer.py never got; sigmoid = model/segment.py:534, soft mask
 * = train_instance.py:398-399; semantics in DESIGN.md section 3) -------------------------
 * logits_dev: [n_inst,h,w] fp32 (logits_bf16 = 0) or bf16 (= 1).
 * boxes_dev:  [n_inst,4] int32 {x0,y0,wk,hk}: placement in the source canvas / output size.
 * mask_off_dev: [n_inst] int64 byte offsets of each instance's tightly packed hk*wk mask.
 * masks_dev:  0/1 bytes.  soft_dev (nullable): trunc(p*255) bytes, same layout.
 * area_dev [n_inst] int32, bbox_dev [n_inst,4] int32 {xmin,ymin,xmax+1,ymax+1} canvas
 * coordinates, zeros when empty.  max_wk/max_hk: upper bounds of wk/hk (grid sizing). */
ISG_API int isg_decode(const void* logits_dev, int logits_bf16, int n_inst, int h, int w,
               const int32_t* boxes_dev, const int64_t* mask_off_dev, int max_wk, int max_hk,
               float threshold, uint8_t* masks_dev, uint8_t* soft_dev, int32_t* area_dev,
               int32_t* bbox_dev, void* stream);

/* Accumulator workspace of the decode kernels (per-instance area / bbox partials, arrival tickets, tile counters).
 * By default they live in arrays of the library, which allows ONE decode in flight per device and <= 65535 instances
 * per call.  A caller that decodes on several streams at once (pipeline slots), or more instances with
 * isg_decode_tiled, binds its own: ws_dev = isg_decode_workspace_bytes(max_inst) bytes of device memory, 16-byte
 * aligned, zeroed ONCE (every launch leaves it zeroed); isg_decode_bind_workspace(ws_dev, max_inst) makes the
 * isg_decode / isg_decode_tiled calls of the calling host thread use it until rebound (NULL: back to the default).
 * The binding is host-thread-local state; the memory stays caller-owned; one workspace per concurrent decode. */
ISG_API int64_t isg_decode_workspace_bytes(int max_inst);
ISG_API int isg_decode_bind_workspace(void* ws_dev, int max_inst);

/* Tiled decode for ragged box sizes (0/1 masks only), optionally fused with the id map.
 * Same inputs/outputs as isg_decode; the work is cut into tiles of isg_decode_tile_rows()
 * consecutive mask rows of one instance, handed out dynamically so that boxes of very
 * different sizes keep every SM busy.  tile_first_dev [n_inst+1] int32 = exclusive prefix of
 * ceil(hk / rows) over the instances (0 tiles for an empty box), n_tiles = its last entry.
 * idmap_dev (nullable) [n_img,canvas_h,canvas_w] int32 with the rule of isg_idmap; needs
 * inst_img_dev [n_inst] (image of each instance), img_first_dev [n_img+1] and at most 32
 * instances per image (set pixels OR their instance bit into the canvas word; contested
 * pixels are resolved by comparing probabilities, exact ones when they are close). */
ISG_API int isg_decode_tile_rows(void);
ISG_API int isg_decode_tiled(const void* logits_dev, int logits_bf16, int n_inst, int h, int w,
                     const int32_t* boxes_dev, const int64_t* mask_off_dev,
                     const int32_t* tile_first_dev, int n_tiles, float threshold,
                     uint8_t* masks_dev, int32_t* area_dev, int32_t* bbox_dev,
                     const int32_t* inst_img_dev, const int32_t* img_first_dev, int n_img,
                     int canvas_h, int canvas_w, int32_t* idmap_dev, void* stream);

/* id map [n_img,canvas_h,canvas_w] int32: per pixel the 1-based index of the covering
 * instance with set mask and largest probability (ties -> lowest index), 0 = background.
 * img_first_dev [n_img+1] int32: instances of image i are img_first[i]..img_first[i+1]-1. */
ISG_API int isg_idmap(const void* logits_dev, int logits_bf16, int h, int w, const int32_t* boxes_dev,
              const int32_t* img_first_dev, int n_img, int canvas_h, int canvas_w,
              float threshold, int32_t* idmap_dev, void* stream);

/* 0/1 mask bytes (the masks_dev layout of isg_decode / isg_decode_tiled, alignment padding included) -> 1 bit per pixel,
 * little-endian within a byte: bit i of bits_dev[j] = masks_dev[8 j + i] (numpy.packbits(bitorder="little")).  bits_dev
 * holds (n + 7) / 8 bytes.  For serving paths whose device -> host copy matters: 8x fewer bytes per mask. */
ISG_API int isg_pack_mask_bits(const uint8_t* masks_dev, int64_t n, uint8_t* bits_dev, void* stream);

/* intersection/union counts of two byte masks (train_instance.py:402-403 mask_iou, applied there to tensor2mask
 * outputs, i.e. 0-255 SOFT bytes of prediction and label; ymlib's source is absent, parity unpinned): a byte counts
 * as set when it is nonzero -- what np.logical_and / np.logical_or, the usual implementation, do with such bytes
 * (soft byte >= 1  <=>  p >= 1/255).  inter_union_dev[0] += |a&b|, [1] += |a|b| (caller zeroes them).
 * isg_mask_iou_counts_thr: a byte of a (b) is set when >= thr_a (thr_b), thresholds in [1,255]; 128 = p >= 0.5 on soft
 * bytes, 1 = the rule above. */
ISG_API int isg_mask_iou_counts(const uint8_t* a_dev, const uint8_t* b_dev, int64_t n,
                        unsigned long long* inter_union_dev, void* stream);
ISG_API int isg_mask_iou_counts_thr(const uint8_t* a_dev, const uint8_t* b_dev, int64_t n, int thr_a, int thr_b,
                            unsigned long long* inter_union_dev, void* stream);

/* ---- input side ------------------------------------------------------------
 * Keypoint heat maps of the Segment(20) variant (train_instance.py:33-68 keypoint2heatmaps,
 * called at :202): kp_dev [batch,n_parts,3] fp32 rows (x, y, flag), flag 1 = present and
 * visible (drawn), anything else = zero map.  Writes fp32 [batch,n_parts,height,width] with
 * the given batch stride (elements), e.g. channels 3.. of the [B,20,H,W] network input, so
 * the torch.cat of model/segment.py:531-533 is never materialised separately.  The value is
 * the reference's float64 exp(-((x-kx)^2+(y-ky)^2)/sigma^2) where > threshold, inside the
 * reference's clipped box (last row/column excluded as np.arange does), rounded to fp32. */
ISG_API int isg_keypoint_heatmaps(const float* kp_dev, int batch, int n_parts, int height, int width,
                          float sigma, float threshold, float* out_dev, int64_t out_batch_stride,
                          void* stream);

/* Per-instance crop / pad / resize of the source images to the network input (geometry of
 * train_instance.py:160-196: window = instance box + 16 px pad, iaa.CropAndPad with zeros,
 * iaa.Resize to the network size).  src_dev uint8 [n_img,src_h,src_w,3] (HWC, what cv2.imread
 * gives), win_dev [n_inst,4] int32 {x0,y0,wk,hk} in source coordinates (may reach outside the
 * image: zeros), inst_img_dev [n_inst] int32.  out_dev uint8 [n_inst,out_h,out_w,3] = the
 * input of isg_forward_u8.  Bilinear with the decode's align_corners=False rule on the window,
 * rounded half-to-even; the same window passed to isg_decode as the box pastes the mask back. */
ISG_API int isg_crop_resize_u8(const uint8_t* src_dev, int n_img, int src_h, int src_w,
                       const int32_t* win_dev, const int32_t* inst_img_dev, int n_inst,
                       int out_h, int out_w, uint8_t* out_dev, void* stream);

/* ---- training step exchange (train_instance.py:379-380) ---------------------
 * fused "unscale + Adam" over the flat fp32 gradient bucket after the NCCL all-reduce:
 * g = grad * grad_scale; standard Adam (lr, beta1, beta2, eps), bias-corrected by `step`. */
ISG_API int isg_adam_step(float* param_dev, const float* grad_dev, float* exp_avg_dev,
                  float* exp_avg_sq_dev, int64_t n, float grad_scale, float lr, float beta1,
                  float beta2, float eps, int step, void* stream);

/* ---- training step loss (train_instance.py:299 BCELoss(), :377-378 sigmoid -> loss -> backward) ----
 * One launch: loss = mean(BCE(sigmoid(logit), target)) with torch's rules (log clamped at -100) and, when
 * dlogits_dev is not NULL, d(loss)/d(logit) in the logits' type (the product of BCELoss's and sigmoid's
 * backward, 1e-12 clamp included).  logits fp32 or bf16 [n], target fp32 [n]; loss_dev: one float.
 * workspace_dev: isg_bce_workspace_bytes() bytes, 8-byte aligned, zeroed once by the caller before its first
 * use (the kernel leaves it ready for the next launch).  The sum order is fixed: results are deterministic. */
ISG_API int64_t isg_bce_workspace_bytes(void);
ISG_API int isg_bce_sigmoid(const void* logits_dev, int logits_bf16, const float* target_dev, int64_t n,
                    void* dlogits_dev, void* workspace_dev, float* loss_dev, void* stream);

/* ---- isg_bce_sigmoid for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.
 * Same inputs, outputs and arithmetic as isg_bce_sigmoid (same grid, same fixed sum order: the loss and the gradient
 * are bitwise equal), but the per-block partial sums and the ticket live in a file-scope `static __device__` array of
 * the library (8.3 KB, held by every GPU that loads libisg until exit) instead of a caller workspace.  The call
 * allocates nothing (no cudaMalloc / cudaFree, no allocating library routine): the array's device address is looked
 * up once per device (cudaGetSymbolAddress with the caller's current device) and the ticket is cleared on `stream`
 * before every launch.  Limit: n <= isg_bce_sigmoid_static_max_n() = 2^26 elements per call; larger n is rejected
 * with ISG_ERR_ARG (use isg_bce_sigmoid).  All calls on one device share the array: calls that overlap on the GPU
 * (different streams, no ordering) race on it -- order them, or use isg_bce_sigmoid with separate workspaces. */
ISG_API int64_t isg_bce_sigmoid_static_max_n(void);
ISG_API int isg_bce_sigmoid_static(const void* logits_dev, int logits_bf16, const float* target_dev, int64_t n,
                           void* dlogits_dev, float* loss_dev, void* stream);

/* ---- train-mode BatchNorm2d (+ the activation of the Conv block) (segment.py:41-45, one per Conv;
 * train_instance.py:372 model.train()) -------------------------------------------------------------------
 * NCHW activations [batch, channels, hw] contiguous, bf16 or fp32 (bf16 != 0); gamma/beta/statistics fp32 [channels]
 * (gamma, beta, running_* may be NULL).  forward: batch statistics (biased variance for y, unbiased into
 * running_var, running = (1-momentum)*running + momentum*batch), z = (x-mean)*invstd*gamma+beta, y = act(z)
 * with act 0 none / 1 ReLU / 2 PReLU (slope_dev: 1 or `channels` slopes), and the mean / invstd the backward
 * needs.  mean_offset_dev (may be NULL): a per-channel constant the caller left out of x -- the conv bias, which
 * the mean subtraction cancels exactly -- added to the batch mean that goes into running_mean.
 * backward: given dy = d loss / d y and the same x: dx (may be NULL), dgamma, dbeta, dslope [channels] (may be NULL;
 * a 1-slope PReLU sums dslope over channels itself); z is recomputed, no activation input is kept.
 * workspace_dev: channels * isg_bn_max_split() * 24 bytes, 8-byte aligned, contents irrelevant.  Sums are
 * combined in a fixed order: deterministic. */
ISG_API int isg_bn_max_split(void);
ISG_API int isg_bn_train_forward(const void* x_dev, int bf16, int batch, int channels, int64_t hw,
                         const float* gamma_dev, const float* beta_dev, const float* mean_offset_dev,
                         float* running_mean_dev, float* running_var_dev, float momentum, float eps,
                         int act, const float* slope_dev, int slope_count, void* y_dev,
                         float* save_mean_dev, float* save_invstd_dev, void* workspace_dev, void* stream);
ISG_API int isg_bn_train_backward(const void* dy_dev, const void* x_dev, int bf16, int batch, int channels,
                          int64_t hw, const float* gamma_dev, const float* beta_dev,
                          const float* save_mean_dev, const float* save_invstd_dev, int act,
                          const float* slope_dev, int slope_count, void* dx_dev, float* dgamma_dev,
                          float* dbeta_dev, float* dslope_dev, void* workspace_dev, void* stream);

/* ---- depthwise convolution in train mode (segment.py:58, :93-96, :121: Conv2d with groups == channels) ----------
 * NCHW activations [batch, channels, height, width] bf16 or fp32; weight fp32 [channels, kh, kw], bias fp32 [channels]
 * or NULL.  Stride 1 and output size == input size only (2*pad == dil*(k-1)), kh*kw <= 25.
 * isg_dwconv_forward: out = conv(in) (+ bias); flip != 0: taps reversed = the gradient w.r.t. the conv's input when
 * `in` is d loss / d out.  isg_dwconv_wgrad: dweight [channels, kh, kw] and dbias [channels] (may be NULL);
 * workspace_dev: isg_dwconv_wgrad_workspace_bytes(channels) bytes or NULL (then one cluster per channel only). */
ISG_API int isg_dwconv_forward(const void* in_dev, int bf16, int batch, int channels, int height, int width,
                       const float* weight_dev, const float* bias_dev, int kh, int kw, int pad_h, int pad_w,
                       int dil_h, int dil_w, int flip, void* out_dev, void* stream);
ISG_API int isg_dwconv_wgrad(const void* dy_dev, const void* x_dev, int bf16, int batch, int channels, int height,
                     int width, int kh, int kw, int pad_h, int pad_w, int dil_h, int dil_w,
                     float* dweight_dev, float* dbias_dev, void* workspace_dev, void* stream);
ISG_API int64_t isg_dwconv_wgrad_workspace_bytes(int channels);

/* ---- ConvTranspose2d forward with kernel == 2*stride, 0 <= pad < stride, cout <= 8 (segment.py:435 bottle6_1: 16 -> 4,
 * k8 s4 p2) in train mode.  in [batch,cin,hin,win] bf16 or fp32, weight fp32 [cin,cout,k,k] (torch's layout), bias fp32
 * [cout] or NULL, out [batch,cout,(hin-1)*s-2p+k,(win-1)*s-2p+k].  The backward of this layer stays with the library. */
ISG_API int isg_convt_k2s_forward(const void* in_dev, int bf16, int batch, int cin, int cout, int hin, int win,
                          int k, int s, int p, const float* weight_dev, const float* bias_dev,
                          void* out_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ISG_H_ */
