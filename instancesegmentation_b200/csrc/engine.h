// Plan = topology of Segment(in_channel) at a fixed HxW, lowered to a list of kernel
// launches over a private activation workspace (NHWC).
#pragma once
#include <functional>
#include <memory>
#include <string>
#include <vector>

#include "common.cuh"
#include "netspec.h"

namespace isg {

struct TensorD {
    std::string name;
    int C, Cs, H, W;      // logical channels, storage stride, spatial size (per image)
    size_t off;           // byte offset in the workspace
    bool tap;             // exposed through isg_plan_tap_*
    // storage variants of the fused bf16 path (engine_fk.cu):
    //  layout 1: space-to-depth -- logical (c,y,x) lives at [(y/2)][(x/2)][((y&1)*2+(x&1))*Cs + c]
    //  perm: logical channel -> storage channel (empty = identity)
    int layout = 0;
    std::vector<int> perm;
    // global memory order of the bf16 path: false = pixel-major [B][H][W][Cs]; true = planar [B][Cs/8][H][W][8]
    // (TMA boxes of planar tensors have rows of tile-width x 16 contiguous bytes, see tc_common.cuh tma_load_act)
    bool planar = false;
    // split precision (ISG_MODE_BF16X3): channel c is stored twice, hi = bf16(v) at storage channel sc and
    // lo = bf16(v - hi) at Cs + sc; the tensor holds Ct() = 2 Cs storage channels
    bool split = false;
    int Ct() const { return split ? 2 * Cs : Cs; }
    // element index of storage channel sc of pixel (y, x) of image b (Hs x Ws = stored spatial size)
    size_t index(int b, int y, int x, int sc, int Hs, int Ws, int Ct) const {
        if (planar) return (((size_t)b * (Ct / 8) + sc / 8) * Hs * Ws + (size_t)y * Ws + x) * 8 + sc % 8;
        return (((size_t)b * Hs + y) * Ws + x) * Ct + sc;
    }
    int storage_of(int c) const { return perm.empty() ? c : perm[c]; }
    int logical_of(int s) const {
        if (perm.empty()) return s < C ? s : -1;
        for (int c = 0; c < C; ++c) if (perm[c] == s) return c;
        return -1;
    }
};

// host view of the caller's state_dict during isg_plan_load_weights
struct HostWeights {
    const StateSpec* spec;
    const void* const* ptrs;
    const float* get(const std::string& name) const {
        auto it = spec->index.find(name);
        return it == spec->index.end() ? nullptr : static_cast<const float*>(ptrs[it->second]);
    }
};

struct Op {
    std::string name;
    // fills this op's slice of the host staging copy of the weight arena
    std::function<void(const HostWeights&, float* arena_host)> pack;
    // enqueues the kernel for `n` images
    std::function<cudaError_t(int n, const float* x, float* logits, cudaStream_t)> launch;
    // 0: always runs; 1: only on the exact-uint8 first-layer path (isg_plan::exact_u8); 2: only on the generic path
    int when = 0;
};

}  // namespace isg

struct isg_plan {
    int device = 0, cin = 3, H = 0, W = 0, max_batch = 0, mode = 0;
    bool weights_loaded = false;
    isg::StateSpec spec;
    std::vector<isg::TensorD> tensors;
    std::vector<isg::Op> ops;
    size_t ws_bytes = 0, arena_floats = 0;
    char* ws = nullptr;        // activation workspace (device)
    float* arena = nullptr;    // packed weights (device)
    std::vector<float> arena_host;
    int64_t macs_per_image = 0;
    int last_n = 0;
    std::vector<int> taps;     // tensor ids
    // isg_forward_u8: uint8 NHWC source of the running forward (null: fp32 NCHW input) and its
    // normalisation; the bf16 plan's first kernel reads it directly, the fp32 plan converts into u8_scratch
    const uint8_t* u8_src = nullptr;
    float u8_mean[8] = {0}, u8_std[8] = {1, 1, 1, 1, 1, 1, 1, 1};
    float* u8_scratch = nullptr;
    bool use_tc = false;       // bf16 modes: GEMM-shaped convs on tcgen05 (ISG_NO_TC=1 disables)
    bool split = false;        // ISG_MODE_BF16X3: split-precision operands (hi + lo bf16 pairs, 3 MMA passes)
    // Exact first layer (split plans, 3 input channels, isg_forward_u8): when v - 255*mean[c] is exactly representable
    // in bf16 for every byte v (the reference's Normalize(0.5, 0.5): v - 127.5), the stem reads those centred pixels as
    // single bf16 values (no lo plane, no A_lo pass) and 1 / (255*std[c]) is folded into layer1's packed weights.
    bool exact_u8 = false;             // the running forward takes that path
    bool exact_packed = false;         // the exact variant's weights are packed for exact_mean / exact_std
    float exact_mean[3] = {0, 0, 0}, exact_std[3] = {0, 0, 0};
    std::vector<std::vector<float>> state_copy;   // the state_dict tensors of the last load (re-packing needs them)
    std::vector<const void*> state_ptrs;
    std::vector<std::function<int()>> finalizers;  // run once after the workspace is allocated (tensor maps)
    // run once after the finalizers: each fused launch times its candidate tilings on the device
    // (scratch logits buffer of max_batch images, stream) and keeps the fastest
    std::vector<std::function<int(float*, cudaStream_t)>> tuners;
    // second pass (isg_plan_create): the choice of every launch with several close candidates is re-made on the time
    // of the WHOLE forward, because a launch timed alone (hot caches, nothing before or after it) ranks its variants
    // differently than the chain does
    struct TuneSite {
        std::string name;
        std::vector<float> iso_ms;          // time of every variant alone (first pass)
        std::function<void(int)> set;       // select a variant
        std::function<int()> get;
    };
    std::vector<TuneSite> tune_sites;

    size_t esize() const { return mode != ISG_MODE_FP32 ? 2 : 4; }
    void* tptr(int id) const { return ws + tensors[id].off; }
};

namespace isg {
int build_plan(isg_plan* p);     // engine_build.cu
void build_fk_plan(isg_plan* p);  // engine_fk.cu (ISG_MODE_BF16)
}
