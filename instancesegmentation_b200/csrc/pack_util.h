// Host-side helpers shared by the plan builders: BatchNorm folding of the reference's
// Conv wrapper (model/segment.py:34-45) into a per-channel affine.
#pragma once
#include <cmath>
#include <cstring>
#include <string>

#include "engine.h"

namespace isg {

constexpr double kBnEps = 1e-5;  // nn.BatchNorm2d default (segment.py:41)

// bn(conv(x) + b) = acc * scale + shift
inline void fold_unit(const HostWeights& w, const std::string& conv, const std::string& bn, int C, float* scale,
                      float* shift) {
    const float* b = w.get(conv + ".bias");
    const float *g = w.get(bn + ".weight"), *be = w.get(bn + ".bias");
    const float *mu = w.get(bn + ".running_mean"), *var = w.get(bn + ".running_var");
    for (int c = 0; c < C; ++c) {
        double s = (double)g[c] / std::sqrt((double)var[c] + kBnEps);
        scale[c] = (float)s;
        shift[c] = (float)((double)be[c] + ((double)b[c] - (double)mu[c]) * s);
    }
}
inline void fold_bias(const HostWeights& w, const std::string& bias, int C, float* scale, float* shift) {
    const float* b = w.get(bias);
    for (int c = 0; c < C; ++c) { scale[c] = 1.f; shift[c] = b[c]; }
}

}  // namespace isg
