// C ABI of libisg (include/isg.h): network description, plan lifetime, forward, test taps.
#include <algorithm>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

#include "engine.h"

using namespace isg;

namespace {
std::mutex g_spec_mu;
std::map<int, StateSpec> g_specs;
const StateSpec& spec_for(int cin) {
    std::lock_guard<std::mutex> lk(g_spec_mu);
    auto it = g_specs.find(cin);
    if (it == g_specs.end()) it = g_specs.emplace(cin, build_state_spec(cin)).first;
    return it->second;
}

// storage element -> float, used only by the tap reader
inline float raw_to_float(const void* raw, size_t i, bool is_bf16) {
    if (!is_bf16) return static_cast<const float*>(raw)[i];
    uint32_t u = (uint32_t)static_cast<const uint16_t*>(raw)[i] << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}
}  // namespace

namespace {
int tune_in_chain(isg_plan* p, float* logits);   // defined at the end of the file
}  // namespace

extern "C" {

const char* isg_last_error(void) { return err_slot().c_str(); }
int isg_version(void) { return 100; }

int isg_state_count(int in_channel) { return (int)spec_for(in_channel).entries.size(); }
const char* isg_state_name(int in_channel, int index) {
    const StateSpec& s = spec_for(in_channel);
    return index >= 0 && index < (int)s.entries.size() ? s.entries[index].name.c_str() : nullptr;
}
int64_t isg_state_numel(int in_channel, int index) {
    const StateSpec& s = spec_for(in_channel);
    return index >= 0 && index < (int)s.entries.size() ? s.entries[index].numel : -1;
}
int isg_state_is_int64(int in_channel, int index) {
    const StateSpec& s = spec_for(in_channel);
    return index >= 0 && index < (int)s.entries.size() ? (int)s.entries[index].is_i64 : -1;
}

int isg_plan_create(int device, int in_channel, int height, int width, int max_batch, int mode,
                    isg_plan** out) {
    if (!out) return fail(ISG_ERR_ARG, "isg_plan_create: out is null");
    *out = nullptr;
    if (in_channel < 1 || in_channel > 64) return fail(ISG_ERR_ARG, "in_channel %d out of range [1,64]", in_channel);
    if (height < 16 || width < 16 || height % 16 || width % 16)
        return fail(ISG_ERR_ARG, "H and W must be multiples of 16 (>=16), got %dx%d", height, width);
    if (max_batch < 1) return fail(ISG_ERR_ARG, "max_batch must be >= 1");
    if (mode != ISG_MODE_FP32 && mode != ISG_MODE_BF16 && mode != ISG_MODE_BF16X3) return fail(ISG_ERR_ARG, "unknown mode %d", mode);
    ISG_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    ISG_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(ISG_ERR_ARCH, "libisg is built for sm_100a only; device %d is sm_%d%d", device, prop.major,
                    prop.minor);
    std::unique_ptr<isg_plan> p(new isg_plan);
    p->device = device; p->cin = in_channel; p->H = height; p->W = width;
    p->max_batch = max_batch; p->mode = mode;
    p->spec = build_state_spec(in_channel);
    build_plan(p.get());
    ISG_CUDA(cudaMalloc(&p->ws, p->ws_bytes));
    ISG_CUDA(cudaMemset(p->ws, 0, p->ws_bytes));  // pad channels stay zero for the plan's lifetime
    ISG_CUDA(cudaMalloc(&p->arena, p->arena_floats * sizeof(float)));
    ISG_CUDA(cudaMemset(p->arena, 0, p->arena_floats * sizeof(float)));
    p->arena_host.assign(p->arena_floats, 0.f);
    for (auto& fin : p->finalizers) {
        int rc = fin();
        if (rc) { cudaFree(p->ws); cudaFree(p->arena); return rc; }
    }
    if (!p->tuners.empty()) {
        // autotune on all-zero activations and constants: the timing of TMA, tcgen05 and the CUDA-core
        // phases does not depend on the data, and the workspace (pad channels included) stays zero
        float* scratch = nullptr;
        ISG_CUDA(cudaMalloc(&scratch, (size_t)max_batch * height * width * sizeof(float)));
        int rc = 0;
        for (auto& tune : p->tuners)
            if ((rc = tune(scratch, nullptr))) break;
        cudaDeviceSynchronize();
        if (!rc && !getenv("ISG_FK_NO_CHAIN_TUNE")) rc = tune_in_chain(p.get(), scratch);
        cudaFree(scratch);
        if (rc) { cudaFree(p->ws); cudaFree(p->arena); return fail(rc, "autotune launch failed: %s", cudaGetErrorString(cudaGetLastError())); }
    }
    ISG_CUDA(cudaDeviceSynchronize());
    *out = p.release();
    return ISG_OK;
}

int isg_plan_dry_run(int in_channel, int height, int width, int max_batch, int mode) {
    if (in_channel < 1 || in_channel > 64 || height < 16 || width < 16 || height % 16 || width % 16 || max_batch < 1 ||
        (mode != ISG_MODE_FP32 && mode != ISG_MODE_BF16 && mode != ISG_MODE_BF16X3))
        return fail(ISG_ERR_ARG, "isg_plan_dry_run: bad argument");
    isg_plan p;
    p.device = 0; p.cin = in_channel; p.H = height; p.W = width; p.max_batch = max_batch; p.mode = mode;
    p.spec = build_state_spec(in_channel);
    build_plan(&p);            // aborts with a message when a fused launch has no feasible tile
    cudaGetLastError();        // attribute calls fail without a device: irrelevant here
    if (getenv("ISG_FK_VERBOSE"))
        for (auto& op : p.ops) fprintf(stderr, "dry-run launch: %s\n", op.name.c_str());
    return (int)p.ops.size();
}

int isg_plan_destroy(isg_plan* plan) {
    if (!plan) return ISG_OK;
    cudaSetDevice(plan->device);
    cudaFree(plan->ws);
    cudaFree(plan->arena);
    cudaFree(plan->u8_scratch);
    delete plan;
    return ISG_OK;
}

size_t isg_plan_workspace_bytes(const isg_plan* plan) { return plan ? plan->ws_bytes : 0; }
int isg_plan_launches_per_forward(const isg_plan* plan) {
    if (!plan) return 0;
    int n = 0;
    for (auto& op : plan->ops) n += op.when != 1;   // of the two first-layer variants exactly one runs
    return n;
}
int64_t isg_plan_macs_per_image(const isg_plan* plan) { return plan ? plan->macs_per_image : 0; }

int isg_plan_load_weights(isg_plan* plan, const void* const* tensors_host, const int64_t* numels, int count) {
    if (!plan || !tensors_host || !numels) return fail(ISG_ERR_ARG, "isg_plan_load_weights: null argument");
    const auto& e = plan->spec.entries;
    if (count != (int)e.size())
        return fail(ISG_ERR_ARG, "expected %d state tensors, got %d", (int)e.size(), count);
    for (int i = 0; i < count; ++i) {
        if (numels[i] != e[i].numel)
            return fail(ISG_ERR_ARG, "state tensor %d (%s): expected %lld elements, got %lld", i,
                        e[i].name.c_str(), (long long)e[i].numel, (long long)numels[i]);
        if (!tensors_host[i]) return fail(ISG_ERR_ARG, "state tensor %d (%s) is null", i, e[i].name.c_str());
    }
    // keep a copy of the float tensors: the exact-uint8 stem variant is packed later, when mean / std are known
    plan->state_copy.assign(count, std::vector<float>());
    plan->state_ptrs.assign(count, nullptr);
    for (int i = 0; i < count; ++i) {
        if (!e[i].is_i64) {
            const float* src = static_cast<const float*>(tensors_host[i]);
            plan->state_copy[i].assign(src, src + e[i].numel);
        }
        plan->state_ptrs[i] = e[i].is_i64 ? nullptr : static_cast<const void*>(plan->state_copy[i].data());
    }
    HostWeights hw{&plan->spec, tensors_host};
    // (the exact-uint8 stem variant, Op::when == 1, is packed for the mean / std of the last isg_forward_u8: when those are
    // known it is re-packed here with the new weights, so that a captured CUDA graph replays with current weights too)
    for (auto& op : plan->ops) op.pack(hw, plan->arena_host.data());
    ISG_CUDA(cudaSetDevice(plan->device));
    ISG_CUDA(cudaDeviceSynchronize());  // no forward may still be reading the old weights
    ISG_CUDA(cudaMemcpy(plan->arena, plan->arena_host.data(), plan->arena_floats * sizeof(float),
                        cudaMemcpyHostToDevice));
    plan->weights_loaded = true;
    return ISG_OK;
}

namespace {

// uint8 NHWC -> fp32 NCHW with ToTensor + Normalize (generic path of isg_forward_u8)
__global__ void __launch_bounds__(256) u8_to_f32_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, int C,
                                                        int HW, int total, const float* __restrict__ mean_std) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;   // (b, pixel)
    if (i >= total) return;
    const int b = i / HW, pix = i - b * HW;
    for (int c = 0; c < C; ++c) {
        const float v = __fdiv_rn((float)src[(size_t)i * C + c], 255.f);
        dst[((size_t)b * C + c) * HW + pix] = __fdiv_rn(v - mean_std[c], mean_std[8 + c]);
    }
}

// one group of <= max_batch images through the plan's launches
int run_group(isg_plan* plan, const float* x, const uint8_t* u8, int n, float* logits, cudaStream_t st) {
    static const bool op_timing = getenv("ISG_OP_TIMING") != nullptr;   // debug: per-op device time
    std::vector<cudaEvent_t> evs;
    if (op_timing) {
        evs.resize(plan->ops.size() + 1);
        for (auto& e : evs) cudaEventCreate(&e);
        cudaEventRecord(evs[0], st);
    }
    plan->u8_src = u8;
    size_t oi = 0;
    for (auto& op : plan->ops) {
        if (op.when && (op.when == 1) != plan->exact_u8) {   // variant of the other first-layer path
            if (op_timing) cudaEventRecord(evs[++oi], st);
            continue;
        }
        cudaError_t e = op.launch(n, x, logits, st);
        if (e != cudaSuccess) {
            plan->u8_src = nullptr;
            return fail(ISG_ERR_CUDA, "launch of %s failed: %s", op.name.c_str(), cudaGetErrorString(e));
        }
        if (op_timing) cudaEventRecord(evs[++oi], st);
    }
    plan->u8_src = nullptr;
    if (op_timing) {
        cudaStreamSynchronize(st);
        float tot = 0;
        for (size_t i = 0; i < plan->ops.size(); ++i) {
            float ms = 0;
            cudaEventElapsedTime(&ms, evs[i], evs[i + 1]);
            tot += ms;
            fprintf(stderr, "optime %7.1f us  %s\n", ms * 1e3, plan->ops[i].name.c_str());
        }
        fprintf(stderr, "optime total %.1f us\n", tot * 1e3);
        for (auto& e : evs) cudaEventDestroy(e);
    }
    plan->last_n = n;
    return ISG_OK;
}

}  // namespace

int isg_forward(isg_plan* plan, const float* x_dev, int batch, float* logits_dev, void* stream) {
    if (!plan || !x_dev || !logits_dev) return fail(ISG_ERR_ARG, "isg_forward: null argument");
    if (!plan->weights_loaded) return fail(ISG_ERR_STATE, "isg_forward before isg_plan_load_weights");
    if (batch < 0) return fail(ISG_ERR_ARG, "negative batch");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t in_img = (size_t)plan->cin * plan->H * plan->W, out_img = (size_t)plan->H * plan->W;
    for (int b0 = 0; b0 < batch; b0 += plan->max_batch) {
        const int n = batch - b0 < plan->max_batch ? batch - b0 : plan->max_batch;
        int rc = run_group(plan, x_dev + b0 * in_img, nullptr, n, logits_dev + b0 * out_img, st);
        if (rc) return rc;
    }
    return ISG_OK;
}

int isg_forward_u8(isg_plan* plan, const uint8_t* x_u8_dev, int batch, const float* mean_host, const float* std_host,
                   float* logits_dev, void* stream) {
    if (!plan || !x_u8_dev || !logits_dev || !mean_host || !std_host)
        return fail(ISG_ERR_ARG, "isg_forward_u8: null argument");
    if (!plan->weights_loaded) return fail(ISG_ERR_STATE, "isg_forward_u8 before isg_plan_load_weights");
    if (batch < 0) return fail(ISG_ERR_ARG, "negative batch");
    if (plan->cin > 8) return fail(ISG_ERR_ARG, "isg_forward_u8: at most 8 input channels");
    for (int c = 0; c < plan->cin; ++c) {
        if (!(std_host[c] != 0.f)) return fail(ISG_ERR_ARG, "isg_forward_u8: std[%d] is zero", c);
        plan->u8_mean[c] = mean_host[c];
        plan->u8_std[c] = std_host[c];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t in_img = (size_t)plan->cin * plan->H * plan->W, out_img = (size_t)plan->H * plan->W;
    const bool direct = plan->use_tc && plan->cin == 3;   // the bf16 plan's first kernel reads uint8 itself
    // exact first layer: every centred pixel v - 255*mean[c] must be a bf16 value (see isg_plan::exact_u8)
    bool exact = direct && plan->split && !getenv("ISG_X3_NO_EXACT_U8");
    for (int c = 0; c < 3 && exact; ++c) {
        const float center = 255.f * mean_host[c];
        if ((double)center != 255.0 * (double)mean_host[c]) exact = false;
        for (int v = 0; v < 256 && exact; ++v) {
            const float u = (float)v - center;
            uint32_t bits;
            memcpy(&bits, &u, 4);
            if (bits & 0xffffu) exact = false;
        }
    }
    if (exact) {
        bool same = plan->exact_packed;
        for (int c = 0; c < 3; ++c) same = same && plan->exact_mean[c] == mean_host[c] && plan->exact_std[c] == std_host[c];
        if (!same) {   // (re)pack the variant's weights for this mean / std: rare, synchronous
            for (int c = 0; c < 3; ++c) { plan->exact_mean[c] = mean_host[c]; plan->exact_std[c] = std_host[c]; }
            HostWeights hw{&plan->spec, plan->state_ptrs.data()};
            for (auto& op : plan->ops) if (op.when == 1) op.pack(hw, plan->arena_host.data());
            ISG_CUDA(cudaDeviceSynchronize());
            ISG_CUDA(cudaMemcpy(plan->arena, plan->arena_host.data(), plan->arena_floats * sizeof(float), cudaMemcpyHostToDevice));
            plan->exact_packed = true;
        }
    }
    struct Flag { isg_plan* p; ~Flag() { p->exact_u8 = false; } } flag{plan};
    plan->exact_u8 = exact;
    if (!direct && !plan->u8_scratch) {
        ISG_CUDA(cudaMalloc(&plan->u8_scratch, ((size_t)plan->max_batch * in_img + 16) * sizeof(float)));
    }
    for (int b0 = 0; b0 < batch; b0 += plan->max_batch) {
        const int n = batch - b0 < plan->max_batch ? batch - b0 : plan->max_batch;
        const uint8_t* src = x_u8_dev + b0 * in_img;
        int rc;
        if (direct) {
            rc = run_group(plan, nullptr, src, n, logits_dev + b0 * out_img, st);
        } else {
            // mean / std ride at the end of the scratch buffer (stream ordered)
            float ms[16];
            for (int c = 0; c < 8; ++c) { ms[c] = plan->u8_mean[c]; ms[8 + c] = plan->u8_std[c]; }
            float* msd = plan->u8_scratch + (size_t)plan->max_batch * in_img;
            ISG_CUDA(cudaMemcpyAsync(msd, ms, sizeof ms, cudaMemcpyHostToDevice, st));
            const int total = n * plan->H * plan->W;
            u8_to_f32_kernel<<<(total + 255) / 256, 256, 0, st>>>(src, plan->u8_scratch, plan->cin, plan->H * plan->W,
                                                                  total, msd);
            ISG_CUDA(cudaGetLastError());
            rc = run_group(plan, plan->u8_scratch, nullptr, n, logits_dev + b0 * out_img, st);
        }
        if (rc) return rc;
    }
    return ISG_OK;
}

int isg_plan_tap_count(const isg_plan* plan) { return plan ? (int)plan->taps.size() : 0; }
const char* isg_plan_tap_name(const isg_plan* plan, int index) {
    if (!plan || index < 0 || index >= (int)plan->taps.size()) return nullptr;
    return plan->tensors[plan->taps[index]].name.c_str();
}
int isg_plan_tap_shape(const isg_plan* plan, int index, int* channels, int* height, int* width) {
    if (!plan || index < 0 || index >= (int)plan->taps.size()) return fail(ISG_ERR_ARG, "bad tap index");
    const TensorD& t = plan->tensors[plan->taps[index]];
    if (channels) *channels = t.C;
    if (height) *height = t.H;
    if (width) *width = t.W;
    return ISG_OK;
}
int isg_plan_tap_read(isg_plan* plan, int index, int n, float* out_host) {
    if (!plan || !out_host || index < 0 || index >= (int)plan->taps.size())
        return fail(ISG_ERR_ARG, "isg_plan_tap_read: bad argument");
    if (n < 1 || n > plan->max_batch) return fail(ISG_ERR_ARG, "tap read of %d images (max_batch %d)", n, plan->max_batch);
    const TensorD& t = plan->tensors[plan->taps[index]];
    const bool is_bf16 = plan->mode != ISG_MODE_FP32;
    const int Ct = t.Ct();   // storage channels (twice Cs when the tensor is stored as hi + lo)
    const size_t per_img = (size_t)t.H * t.W * Ct, es = is_bf16 ? 2 : 4;
    std::vector<uint8_t> raw(per_img * n * es);
    ISG_CUDA(cudaSetDevice(plan->device));
    ISG_CUDA(cudaDeviceSynchronize());
    ISG_CUDA(cudaMemcpy(raw.data(), plan->ws + t.off, raw.size(), cudaMemcpyDeviceToHost));
    for (int b = 0; b < n; ++b)
        for (int c = 0; c < t.C; ++c) {
            const int sc = t.storage_of(c);
            for (int y = 0; y < t.H; ++y)
                for (int x = 0; x < t.W; ++x) {
                    size_t i;   // element index in the workspace tensor (image b included)
                    float v = 0.f;
                    for (int part = 0; part < (t.split ? 2 : 1); ++part) {   // hi (+ lo)
                        if (t.layout == 1) {
                            const int sub = ((y & 1) * 2 + (x & 1)) * t.Cs + sc + part * 4 * t.Cs;
                            i = t.index(b, y >> 1, x >> 1, sub, t.H >> 1, t.W >> 1, 4 * Ct);
                        } else {
                            i = t.index(b, y, x, sc + part * t.Cs, t.H, t.W, Ct);
                        }
                        v += raw_to_float(raw.data(), i, is_bf16);
                    }
                    out_host[(((size_t)b * t.C + c) * t.H + y) * t.W + x] = v;
                }
        }
    return ISG_OK;
}

}  // extern "C"

namespace {

// Second tuning pass: for every launch whose first-pass (stand-alone) timings left several candidates within 25 % of
// the best, time the whole forward of max_batch images with each of them (all other launches at their current
// choice) and keep the one that makes the FORWARD fastest.  All-zero input and weights: timing is data independent.
int tune_in_chain(isg_plan* p, float* logits) {
    const bool verbose = getenv("ISG_FK_VERBOSE") != nullptr;
    float* x = nullptr;
    const size_t in_floats = (size_t)p->max_batch * p->cin * p->H * p->W;
    if (cudaMalloc(&x, in_floats * sizeof(float)) != cudaSuccess) { cudaGetLastError(); return ISG_OK; }   // no room: keep pass 1
    cudaMemset(x, 0, in_floats * sizeof(float));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    int rc = ISG_OK;
    auto forward_ms = [&]() -> float {   // best of 3 measurements of 3 forwards each
        float best = 1e30f;
        for (int m = 0; m < 3 && rc == ISG_OK; ++m) {
            cudaEventRecord(e0, nullptr);
            for (int r = 0; r < 3 && rc == ISG_OK; ++r) rc = run_group(p, x, nullptr, p->max_batch, logits, nullptr);
            cudaEventRecord(e1, nullptr);
            if (cudaEventSynchronize(e1) != cudaSuccess) rc = ISG_ERR_CUDA;
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            best = ms / 3 < best ? ms / 3 : best;
        }
        return best;
    };
    forward_ms();   // warm up
    for (int pass = 0; pass < 2 && rc == ISG_OK; ++pass)
        for (auto& site : p->tune_sites) {
            float iso_best = 1e30f;
            for (float v : site.iso_ms) if (v > 0.f && v < iso_best) iso_best = v;
            std::vector<int> cand;
            for (size_t i = 0; i < site.iso_ms.size(); ++i)
                if (site.iso_ms[i] > 0.f && site.iso_ms[i] <= 1.25f * iso_best) cand.push_back((int)i);
            if (cand.size() < 2) continue;
            std::sort(cand.begin(), cand.end(), [&](int a, int b) { return site.iso_ms[a] < site.iso_ms[b]; });
            if (cand.size() > 6) cand.resize(6);
            const int before = site.get();
            int best_i = before;
            float best_ms = 1e30f;
            for (int i : cand) {
                site.set(i);
                const float ms = forward_ms();
                if (rc != ISG_OK) break;
                if (ms < best_ms) { best_ms = ms; best_i = i; }
            }
            site.set(best_i);
            if (verbose)
                fprintf(stderr, "chain-tune pass %d %-44s %zu candidates: variant %d -> %d, forward %.1f us\n", pass,
                        site.name.c_str(), cand.size(), before, best_i, best_ms * 1e3);
        }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    cudaFree(x);
    return rc;
}
}  // namespace
