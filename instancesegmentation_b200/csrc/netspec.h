// The 588 state_dict entries of Segment(in_channel), in the reference's order
// (model/segment.py:347-440; key grammar: SURVEY.md section 2.3).  Host-only code.
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <vector>

namespace isg {

struct StateEntry {
    std::string name;
    int64_t numel;
    bool is_i64;
};

struct StateSpec {
    std::vector<StateEntry> entries;
    std::unordered_map<std::string, int> index;
};

class SpecBuilder {
public:
    std::vector<StateEntry> e;
    void add(const std::string& n, int64_t numel, bool i64 = false) { e.push_back({n, numel, i64}); }
    void bn(const std::string& p, int c) {
        add(p + ".weight", c); add(p + ".bias", c);
        add(p + ".running_mean", c); add(p + ".running_var", c);
        add(p + ".num_batches_tracked", 1, true);
    }
    // Conv wrapper (segment.py:34-45): conv.{weight,bias}, bn.*, act.weight iff PReLU
    void unit(const std::string& p, int cout, int cin_per_group, int kh, int kw, bool prelu) {
        add(p + ".conv.weight", (int64_t)cout * cin_per_group * kh * kw);
        add(p + ".conv.bias", cout);
        bn(p + ".bn", cout);
        if (prelu) add(p + ".act.weight", cout);
    }
    void res3(const std::string& p, int c, int m) {  // Bottleneck3x3 :52-79
        unit(p + ".convs.0", m, c, 1, 1, true);
        unit(p + ".convs.1", m, 1, 3, 3, true);
        unit(p + ".convs.2", c, m, 1, 1, false);
        add(p + ".prelu.weight", c);
    }
    void res5(const std::string& p, int c, int m) {  // Bottleneck5x5 :82-111
        unit(p + ".convs.0", m, c, 1, 1, true);
        add(p + ".convs.1.weight", (int64_t)m * 5); add(p + ".convs.1.bias", m);
        unit(p + ".convs.2", m, 1, 1, 5, true);
        unit(p + ".convs.3", c, m, 1, 1, false);
        add(p + ".prelu.weight", c);
    }
    void down(const std::string& p, int ci, int m, int co) {  // BottleneckDown2 :114-150
        unit(p + ".convs.0", m, ci, 2, 2, true);
        unit(p + ".convs.1", m, 1, 3, 3, true);
        unit(p + ".convs.2", co, m, 1, 1, false);
        unit(p + ".convm.0", co, ci, 1, 1, false);
        add(p + ".prelu.weight", co);
    }
    void proj(const std::string& p, int ci, int m, int co) {  // BottleneckDim_Res :153-209
        unit(p + ".convs.0", m, ci, 1, 1, true);
        unit(p + ".convs.1", m, 1, 3, 3, true);
        unit(p + ".convs.2", co, m, 1, 1, false);
        unit(p + ".resconv.0", co, ci, 1, 1, false);
        add(p + ".prelu.weight", co);
    }
    void plain(const std::string& p, int c, int m) {  // BottleneckDim(usePrelu=False) :232-261
        unit(p + ".convs.0", m, c, 1, 1, false);
        unit(p + ".convs.1", m, m, 3, 3, false);
        unit(p + ".convs.2", c, m, 1, 1, false);
        add(p + ".prelu.weight", c);
    }
    void up(const std::string& p, int ci, int m, int co, int skip) {  // BottleneckUp_Res(_Other) :296-344
        unit(p + ".convs.0", m, ci, 1, 1, false);
        add(p + ".convs.1.weight", (int64_t)m * m * 16); add(p + ".convs.1.bias", m);
        bn(p + ".convs.2", m);
        unit(p + ".convs.4", co, m, 1, 1, false);
        unit(p + ".conv2.0", co, ci, 1, 1, false);
        add(p + ".uppool.1.weight", (int64_t)co * (co + skip)); add(p + ".uppool.1.bias", co);
    }
    void wide(const std::string& p) {
        for (int i = 0; i < 4; ++i) res3(p + "." + std::to_string(i), 128, 48);
        res5(p + ".4", 128, 48);
    }
};

inline StateSpec build_state_spec(int cin) {
    SpecBuilder b;
    const int ci = 16 + cin;
    b.unit("init_conv.layer1", 16, cin, 5, 5, true);
    b.unit("init_conv.layer2", 16, 16, 5, 5, true);
    b.down("bottle1_1", ci, 16, 48);
    for (int i = 0; i < 4; ++i) b.res3("bottle1_x." + std::to_string(i), 48, 16);
    b.down("bottle2_1", 48, 16, 128);
    b.wide("bottle2_x");
    b.proj("bottle3_1", 256, 48, 128);
    b.wide("bottle3_x");
    b.up("bottle4_1up", 128, 16, 48, 48);
    b.proj("bottle4_2", 96, 16, 48);
    b.plain("bottle4_3", 48, 16);
    b.up("bottle5_1up", 48, 4, 16, ci);
    b.plain("bottle5_2", 16, 4);
    b.add("bottle6_1.weight", 16 * 4 * 64); b.add("bottle6_1.bias", 4);
    b.add("bottle6_2.weight", 36); b.add("bottle6_2.bias", 1);
    StateSpec s;
    s.entries = std::move(b.e);
    for (int i = 0; i < (int)s.entries.size(); ++i) s.index[s.entries[i].name] = i;
    return s;
}

}  // namespace isg
