// Lowers the Segment topology (model/segment.py:347-508) to a launch list.  Each emit_*
// function registers (a) a packer that rewrites the reference-layout weights into the
// kernel's layout inside the weight arena and (b) a launch closure.
#include <cmath>
#include <algorithm>
#include <cstdlib>
#include <cstring>

#include "engine.h"
#include "kernels_cc.cuh"
#include "kernels_misc.cuh"
#include "kernels_pw.cuh"

namespace isg {
namespace {

constexpr double BN_EPS = 1e-5;  // nn.BatchNorm2d default (segment.py:41)

struct Src {  // one operand of a (possibly concatenated) pointwise gather
    int t0 = -1, t1 = -1, mode = SRC_DIRECT;
};

template <typename T>
struct Builder {
    isg_plan* P;
    explicit Builder(isg_plan* p) : P(p) {}

    // ---- tensors / arena ----
    int tensor(const std::string& name, int C, int H, int W, bool tap = false) {
        TensorD t{name, C, round_up(C, sizeof(T) == 2 ? 16 : 8), H, W, P->ws_bytes, tap};
        size_t bytes = (size_t)P->max_batch * H * W * t.Cs * sizeof(T);
        P->ws_bytes += (bytes + 255) / 256 * 256;
        P->tensors.push_back(t);
        if (tap) P->taps.push_back((int)P->tensors.size() - 1);
        return (int)P->tensors.size() - 1;
    }
    const TensorD& td(int id) const { return P->tensors[id]; }
    size_t arena(size_t n) {  // returns float offset, 16-byte aligned
        size_t o = P->arena_floats;
        P->arena_floats += (n + 3) / 4 * 4;
        return o;
    }
    // pointers are resolved lazily (workspace / arena are allocated after building)
    T* tp(int id) const { return reinterpret_cast<T*>(P->ws + P->tensors[id].off); }
    const float* ap(size_t off) const { return P->arena + off; }

    void macs(int64_t cout, int64_t cin_g, int kh, int kw, int Ho, int Wo) {
        P->macs_per_image += cout * cin_g * kh * kw * Ho * Wo;
    }

    // ---- packers ----
    // scale/shift of a Conv wrapper: bn(conv(x)+b) = acc*scale + shift
    static void affine_unit(const HostWeights& w, const std::string& conv, const std::string& bn, int C,
                            float* scale, float* shift) {
        const float* b = w.get(conv + ".bias");
        const float *g = w.get(bn + ".weight"), *be = w.get(bn + ".bias");
        const float *mu = w.get(bn + ".running_mean"), *var = w.get(bn + ".running_var");
        for (int c = 0; c < C; ++c) {
            double s = (double)g[c] / std::sqrt((double)var[c] + BN_EPS);
            scale[c] = (float)s;
            shift[c] = (float)((double)be[c] + ((double)b[c] - (double)mu[c]) * s);
        }
    }
    static void affine_bias(const HostWeights& w, const std::string& bias, int C, float* scale, float* shift) {
        const float* b = w.get(bias);
        for (int c = 0; c < C; ++c) { scale[c] = 1.f; shift[c] = b[c]; }
    }
    static void copy_vec(const HostWeights& w, const std::string& name, int C, float* dst) {
        std::memcpy(dst, w.get(name), sizeof(float) * C);
    }
    // 1x1 weights [Cout][C0+C1] -> rows over storage channels [Cs0 + Cs1][CoutPad]
    static void pack_pw(const float* wt, int Cout, int CoutPad, int C0, int Cs0, int C1, int Cs1, float* dst) {
        std::memset(dst, 0, sizeof(float) * (size_t)(Cs0 + Cs1) * CoutPad);
        const int Cin = C0 + C1;
        for (int o = 0; o < Cout; ++o)
            for (int i = 0; i < Cin; ++i) {
                int row = i < C0 ? i : Cs0 + (i - C0);
                dst[(size_t)row * CoutPad + o] = wt[(size_t)o * Cin + i];
            }
    }
    // dense conv [Cout][Cin][KH][KW] -> [(ky,kx,ci_storage)][Cout]
    static void pack_convk(const float* wt, int Cout, int Cin, int Cs, int KH, int KW, float* dst) {
        std::memset(dst, 0, sizeof(float) * (size_t)KH * KW * Cs * Cout);
        for (int o = 0; o < Cout; ++o)
            for (int i = 0; i < Cin; ++i)
                for (int t = 0; t < KH * KW; ++t)
                    dst[((size_t)t * Cs + i) * Cout + o] = wt[((size_t)o * Cin + i) * KH * KW + t];
    }
    // transposed conv [Cin][Cout][K][K] -> [(ky,kx)][ci_storage][Cout]
    static void pack_convt(const float* wt, int Cin, int Cout, int Cs, int K, float* dst) {
        std::memset(dst, 0, sizeof(float) * (size_t)K * K * Cs * Cout);
        for (int i = 0; i < Cin; ++i)
            for (int o = 0; o < Cout; ++o)
                for (int t = 0; t < K * K; ++t)
                    dst[((size_t)t * Cs + i) * Cout + o] = wt[((size_t)i * Cout + o) * K * K + t];
    }

    static int pw_chunk(int Cout) { return Cout >= 128 ? 32 : (Cout >= 16 ? 16 : 4); }

    void emit_stem(int t_s1, int t_init);
    void emit_convk(const std::string& name, const std::string& unit, int tin, int tout, int co_off, int Cout,
                    int KH, int KW, int stride, int pad, int act);
    void emit_dw(const std::string& name, const std::string& conv, const std::string& bn,
                 const std::string& slope, int tin, int tout, int KH, int KW, int dil, int act);
    struct PwSpec {
        std::string name;
        Src a; std::string wa, bna, slope_a; int act_a = ACT_NONE; bool a_bias_only = false;
        int b_kind = BR_NONE; Src b; std::string wb, bnb; bool b_bias_only = false; int res = -1;
        int act_out = ACT_NONE; std::string slope_out;
        int out = -1; int pool_out = -1;
    };
    void emit_pw(const PwSpec& s);
    void emit_pool2(const std::string& name, int tin, int tout);
    void emit_convt4(const std::string& name, const std::string& prefix, int tin, int tout);
    void emit_tail(int tin);

    int block_down(const std::string& p, int x, int mid, int cout, int* keep, const std::string& tapname);
    int block_res3(const std::string& p, int x, int mid, int dil);
    int block_res5(const std::string& p, int x, int mid);
    int block_proj(const std::string& p, int x0, int x1, int mid, int cout, bool last_prelu);
    int block_plain(const std::string& p, int x, int mid);
    int block_up(const std::string& p, int x, int skip, int mid, int cout);
    void build();
};

#include "engine_emit.inl"
#include "engine_blocks.inl"

}  // namespace

int build_plan(isg_plan* p) {
    // bf16 mode = fused tcgen05 block kernels (engine_fk.cu); ISG_NO_TC=1 keeps the CUDA-core
    // kernels with bf16 storage, which exist only as a debugging cross-check
    p->split = p->mode == ISG_MODE_BF16X3;
    p->use_tc = p->split || (p->mode == ISG_MODE_BF16 && !getenv("ISG_NO_TC"));
    if (p->use_tc) build_fk_plan(p);
    else if (p->mode == ISG_MODE_BF16) Builder<bf16>(p).build();
    else Builder<float>(p).build();
    return 0;
}

}  // namespace isg
