// Builder<T>::block_* -- the reference's residual blocks lowered to emit_* launches.
// Included inside namespace isg::{anon}.  `p` is the state_dict prefix of the block.

template <typename T>
void Builder<T>::emit_pool2(const std::string& name, int tin, int tout) {
    isg_plan* P = this->P;
    const TensorD ti = td(tin), to = td(tout);
    Op op;
    op.name = name;
    op.pack = [](const HostWeights&, float*) {};
    op.launch = [=](int n, const float*, float*, cudaStream_t st) {
        int total = n * to.H * to.W * (to.Cs / 4);
        pool2_kernel<T><<<cdiv(total, 256), 256, 0, st>>>(reinterpret_cast<const T*>(P->ws + ti.off),
                                                         reinterpret_cast<T*>(P->ws + to.off), to.H, to.W, to.Cs, total);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
}

#define ISG_PW(spec) emit_pw(spec)

// BottleneckDown2 (segment.py:114-150): returns the block output; *keep = max-pooled input.
template <typename T>
int Builder<T>::block_down(const std::string& p, int x, int mid, int cout, int* keep,
                           const std::string& tapname) {
    const TensorD tx = td(x);
    const int h = tx.H / 2, w = tx.W / 2;
    int m1 = tensor(p + ".m1", mid, h, w), m2 = tensor(p + ".m2", mid, h, w);
    *keep = tensor(p + ".keep", tx.C, h, w, true);
    int out = tensor(tapname, cout, h, w, true);
    emit_convk(p + ".convs.0", p + ".convs.0", x, m1, 0, mid, 2, 2, 2, 0, ACT_PRELU);
    emit_dw(p + ".convs.1", p + ".convs.1.conv", p + ".convs.1.bn", p + ".convs.1.act.weight", m1, m2, 3, 3, 1,
            ACT_PRELU);
    PwSpec s;
    s.name = p + ".convs.2+convm+prelu";
    s.a.t0 = m2; s.wa = p + ".convs.2.conv"; s.bna = p + ".convs.2.bn";
    s.b_kind = BR_GEMM; s.b.t0 = x; s.b.mode = SRC_POOL2; s.wb = p + ".convm.0.conv"; s.bnb = p + ".convm.0.bn";
    s.act_out = ACT_PRELU; s.slope_out = p + ".prelu.weight";
    s.out = out; s.pool_out = *keep;
    emit_pw(s);
    return out;
}

// Bottleneck3x3 (segment.py:52-79)
template <typename T>
int Builder<T>::block_res3(const std::string& p, int x, int mid, int dil) {
    const TensorD tx = td(x);
    int m1 = tensor(p + ".m1", mid, tx.H, tx.W), m2 = tensor(p + ".m2", mid, tx.H, tx.W);
    int out = tensor(p, tx.C, tx.H, tx.W);
    PwSpec a;
    a.name = p + ".convs.0";
    a.a.t0 = x; a.wa = p + ".convs.0.conv"; a.bna = p + ".convs.0.bn";
    a.slope_a = p + ".convs.0.act.weight"; a.act_a = ACT_PRELU; a.out = m1;
    ISG_PW(a);
    emit_dw(p + ".convs.1", p + ".convs.1.conv", p + ".convs.1.bn", p + ".convs.1.act.weight", m1, m2, 3, 3, dil,
            ACT_PRELU);
    PwSpec c;
    c.name = p + ".convs.2+res+prelu";
    c.a.t0 = m2; c.wa = p + ".convs.2.conv"; c.bna = p + ".convs.2.bn";
    c.b_kind = BR_IDENTITY; c.res = x;
    c.act_out = ACT_PRELU; c.slope_out = p + ".prelu.weight"; c.out = out;
    ISG_PW(c);
    return out;
}

// Bottleneck5x5 (segment.py:82-111): bare depthwise (5,1) with bias, then (1,5)+BN+PReLU
template <typename T>
int Builder<T>::block_res5(const std::string& p, int x, int mid) {
    const TensorD tx = td(x);
    int m1 = tensor(p + ".m1", mid, tx.H, tx.W), m2 = tensor(p + ".m2", mid, tx.H, tx.W);
    int m3 = tensor(p + ".m3", mid, tx.H, tx.W), out = tensor(p, tx.C, tx.H, tx.W);
    PwSpec a;
    a.name = p + ".convs.0";
    a.a.t0 = x; a.wa = p + ".convs.0.conv"; a.bna = p + ".convs.0.bn";
    a.slope_a = p + ".convs.0.act.weight"; a.act_a = ACT_PRELU; a.out = m1;
    ISG_PW(a);
    emit_dw(p + ".convs.1", p + ".convs.1", "", "", m1, m2, 5, 1, 1, ACT_NONE);
    emit_dw(p + ".convs.2", p + ".convs.2.conv", p + ".convs.2.bn", p + ".convs.2.act.weight", m2, m3, 1, 5, 1,
            ACT_PRELU);
    PwSpec c;
    c.name = p + ".convs.3+res+prelu";
    c.a.t0 = m3; c.wa = p + ".convs.3.conv"; c.bna = p + ".convs.3.bn";
    c.b_kind = BR_IDENTITY; c.res = x;
    c.act_out = ACT_PRELU; c.slope_out = p + ".prelu.weight"; c.out = out;
    ISG_PW(c);
    return out;
}

// BottleneckDim_Res (segment.py:153-209) on the concat (x0 | x1) (torch.cat :485, :494)
template <typename T>
int Builder<T>::block_proj(const std::string& p, int x0, int x1, int mid, int cout, bool last_prelu) {
    const TensorD tx = td(x0);
    int m1 = tensor(p + ".m1", mid, tx.H, tx.W), m2 = tensor(p + ".m2", mid, tx.H, tx.W);
    int out = tensor(p, cout, tx.H, tx.W, true);
    PwSpec a;
    a.name = p + ".convs.0";
    a.a.t0 = x0; a.a.t1 = x1; a.wa = p + ".convs.0.conv"; a.bna = p + ".convs.0.bn";
    a.slope_a = p + ".convs.0.act.weight"; a.act_a = ACT_PRELU; a.out = m1;
    ISG_PW(a);
    emit_dw(p + ".convs.1", p + ".convs.1.conv", p + ".convs.1.bn", p + ".convs.1.act.weight", m1, m2, 3, 3, 1,
            ACT_PRELU);
    PwSpec c;
    c.name = p + ".convs.2+resconv";
    c.a.t0 = m2; c.wa = p + ".convs.2.conv"; c.bna = p + ".convs.2.bn";
    c.b_kind = BR_GEMM; c.b.t0 = x0; c.b.t1 = x1; c.wb = p + ".resconv.0.conv"; c.bnb = p + ".resconv.0.bn";
    if (last_prelu) { c.act_out = ACT_PRELU; c.slope_out = p + ".prelu.weight"; }
    else c.act_out = ACT_RELU;
    c.out = out;
    ISG_PW(c);
    return out;
}

// BottleneckDim(usePrelu=False) (segment.py:232-261): dense 3x3 in the middle, ReLU everywhere
template <typename T>
int Builder<T>::block_plain(const std::string& p, int x, int mid) {
    const TensorD tx = td(x);
    int m1 = tensor(p + ".m1", mid, tx.H, tx.W), m2 = tensor(p + ".m2", mid, tx.H, tx.W);
    int out = tensor(p, tx.C, tx.H, tx.W, true);
    PwSpec a;
    a.name = p + ".convs.0";
    a.a.t0 = x; a.wa = p + ".convs.0.conv"; a.bna = p + ".convs.0.bn"; a.act_a = ACT_RELU; a.out = m1;
    ISG_PW(a);
    emit_convk(p + ".convs.1", p + ".convs.1", m1, m2, 0, mid, 3, 3, 1, 1, ACT_RELU);
    PwSpec c;
    c.name = p + ".convs.2+res+relu";
    c.a.t0 = m2; c.wa = p + ".convs.2.conv"; c.bna = p + ".convs.2.bn";
    c.b_kind = BR_IDENTITY; c.res = x; c.act_out = ACT_RELU; c.out = out;
    ISG_PW(c);
    return out;
}

// BottleneckUp_Res / _Other (segment.py:296-344); `skip` is the pooled map kept by a down block
template <typename T>
int Builder<T>::block_up(const std::string& p, int x, int skip, int mid, int cout) {
    const TensorD tx = td(x);
    const int h = tx.H, w = tx.W;
    int m1 = tensor(p + ".m1", mid, h, w), m2 = tensor(p + ".m2", mid, 2 * h, 2 * w);
    int r0 = tensor(p + ".r0", cout, h, w), out = tensor(p, cout, 2 * h, 2 * w, true);
    PwSpec a;
    a.name = p + ".convs.0";
    a.a.t0 = x; a.wa = p + ".convs.0.conv"; a.bna = p + ".convs.0.bn"; a.act_a = ACT_RELU; a.out = m1;
    ISG_PW(a);
    emit_convt4(p + ".convs.1+bn+relu", p, m1, m2);
    PwSpec r;
    r.name = p + ".conv2";
    r.a.t0 = x; r.wa = p + ".conv2.0.conv"; r.bna = p + ".conv2.0.bn"; r.out = r0;
    ISG_PW(r);
    PwSpec c;
    c.name = p + ".convs.4+uppool+relu";
    c.a.t0 = m2; c.wa = p + ".convs.4.conv"; c.bna = p + ".convs.4.bn";
    c.b_kind = BR_GEMM; c.b.t0 = r0; c.b.t1 = skip; c.b.mode = SRC_UP2;
    c.wb = p + ".uppool.1"; c.b_bias_only = true;
    c.act_out = ACT_RELU; c.out = out;
    ISG_PW(c);
    return out;
}

// Segment.forward (segment.py:466-508)
template <typename T>
void Builder<T>::build() {
    const int H = P->H, W = P->W, cin = P->cin;
    int s1 = tensor("stem1", 16, H / 2, W / 2, true);
    int init = tensor("init_down", 16 + cin, H / 4, W / 4, true);
    emit_stem(s1, init);
    int keep1, keep2;
    int d1 = block_down("bottle1_1", init, 16, 48, &keep1, "bottle1_down");
    P->tensors[keep1].name = "keep1";
    int y = d1;
    for (int i = 0; i < 4; ++i) y = block_res3("bottle1_x." + std::to_string(i), y, 16, 1);
    P->tensors[y].name = "bottle1_5"; P->tensors[y].tap = true; P->taps.push_back(y);
    int d2 = block_down("bottle2_1", y, 16, 128, &keep2, "bottle2_down");
    P->tensors[keep2].name = "keep2";
    const int dils[4] = {1, 2, 1, 4};
    auto wide = [&](const std::string& p, int t, const char* tapname) {
        for (int i = 0; i < 4; ++i) t = block_res3(p + "." + std::to_string(i), t, 48, dils[i]);
        t = block_res5(p + ".4", t, 48);
        P->tensors[t].name = tapname; P->tensors[t].tap = true; P->taps.push_back(t);
        return t;
    };
    y = wide("bottle2_x", d2, "bottle2_8");
    y = block_proj("bottle3_1", y, d2, 48, 128, true);
    y = wide("bottle3_x", y, "bottle3_8");
    y = block_up("bottle4_1up", y, keep2, 16, 48);
    P->tensors[y].name = "bottle4_1";
    y = block_proj("bottle4_2", d1, y, 16, 48, false);
    y = block_plain("bottle4_3", y, 16);
    y = block_up("bottle5_1up", y, keep1, 4, 16);
    P->tensors[y].name = "bottle5_1";
    y = block_plain("bottle5_2", y, 4);
    emit_tail(y);
}
