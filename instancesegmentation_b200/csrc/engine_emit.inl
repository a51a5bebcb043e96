// Builder<T>::emit_* -- one kernel launch each.  Included inside namespace isg::{anon}.

template <typename T>
void Builder<T>::emit_stem(int t_s1, int t_init) {
    isg_plan* P = this->P;
    const int cin = P->cin, H = P->H, W = P->W;
    // layer1 (segment.py:23-24)
    size_t ow = arena((size_t)cin * 25 * 16), oa = arena(48);
    macs(16, cin, 5, 5, H / 2, W / 2);
    Op op;
    op.name = "init_conv.layer1";
    op.pack = [=](const HostWeights& w, float* A) {
        const float* wt = w.get("init_conv.layer1.conv.weight");  // [16][cin][5][5]
        for (int o = 0; o < 16; ++o)
            for (int k = 0; k < cin * 25; ++k) A[ow + (size_t)k * 16 + o] = wt[(size_t)o * cin * 25 + k];
        affine_unit(w, "init_conv.layer1.conv", "init_conv.layer1.bn", 16, A + oa, A + oa + 16);
        copy_vec(w, "init_conv.layer1.act.weight", 16, A + oa + 32);
    };
    const size_t smem = ((size_t)cin * 25 * 16 + S1_CC * S1_TR * S1_TC) * sizeof(float);
    cudaFuncSetAttribute(stem1_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const size_t off_s1 = td(t_s1).off;
    op.launch = [=](int n, const float* x, float*, cudaStream_t st) {
        dim3 grid(cdiv(W / 2, S1_TX), cdiv(H / 2, S1_TY), n);
        stem1_kernel<T><<<grid, S1_TX * S1_TY, smem, st>>>(x, cin, H, W, reinterpret_cast<T*>(P->ws + off_s1),
                                                          P->arena + ow, P->arena + oa, P->arena + oa + 16,
                                                          P->arena + oa + 32);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
    // max-pool 4x4 of the raw input -> channels [0,cin) of init_down (segment.py:29)
    Op pool;
    pool.name = "init_conv.maxpool4";
    pool.pack = [](const HostWeights&, float*) {};
    const size_t off_init = td(t_init).off;
    const int cs_init = td(t_init).Cs;
    pool.launch = [=](int n, const float* x, float*, cudaStream_t st) {
        int total = n * cin * (H / 4) * (W / 4);
        pool4_kernel<T><<<cdiv(total, 256), 256, 0, st>>>(x, cin, H, W, reinterpret_cast<T*>(P->ws + off_init),
                                                         cs_init, total);
        return cudaGetLastError();
    };
    P->ops.push_back(pool);
    // layer2 -> channels [cin, cin+16) (segment.py:25-26, :31)
    emit_convk("init_conv.layer2", "init_conv.layer2", t_s1, t_init, cin, 16, 5, 5, 2, 2, ACT_PRELU);
}

template <typename T>
void Builder<T>::emit_convk(const std::string& name, const std::string& unit, int tin, int tout, int co_off,
                            int Cout, int KH, int KW, int stride, int pad, int act) {
    isg_plan* P = this->P;
    const TensorD ti = td(tin), to = td(tout);
    const int K = KH * KW * ti.Cs;
    size_t ow = arena((size_t)K * Cout), oa = arena(3 * (size_t)Cout);
    macs(Cout, ti.C, KH, KW, to.H, to.W);
    Op op;
    op.name = name;
    const int Cin = ti.C, Cs = ti.Cs;
    op.pack = [=](const HostWeights& w, float* A) {
        pack_convk(w.get(unit + ".conv.weight"), Cout, Cin, Cs, KH, KW, A + ow);
        affine_unit(w, unit + ".conv", unit + ".bn", Cout, A + oa, A + oa + Cout);
        if (act == ACT_PRELU) copy_vec(w, unit + ".act.weight", Cout, A + oa + 2 * Cout);
    };
    const size_t smem = (size_t)K * Cout * sizeof(float);
    op.launch = [=](int n, const float*, float*, cudaStream_t st) {
        ConvKParams p;
        p.in = P->ws + ti.off; p.Hi = ti.H; p.Wi = ti.W; p.Cs_in = ti.Cs;
        p.out = P->ws + to.off; p.Ho = to.H; p.Wo = to.W; p.Cs_out = to.Cs; p.co_off = co_off;
        p.KH = KH; p.KW = KW; p.stride = stride; p.pad = pad;
        p.w = P->arena + ow; p.scale = P->arena + oa; p.shift = P->arena + oa + Cout;
        p.slope = act == ACT_PRELU ? P->arena + oa + 2 * Cout : nullptr;
        p.act = act; p.total_pix = n * to.H * to.W;
        if (Cout == 16) convk_kernel<T, 16><<<cdiv(p.total_pix, 128), 128, smem, st>>>(p);
        else convk_kernel<T, 4><<<cdiv(p.total_pix, 128), 128, smem, st>>>(p);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
}

template <typename T>
void Builder<T>::emit_dw(const std::string& name, const std::string& conv, const std::string& bn,
                         const std::string& slope, int tin, int tout, int KH, int KW, int dil, int act) {
    isg_plan* P = this->P;
    const TensorD ti = td(tin), to = td(tout);
    const int C = ti.C, Cs = ti.Cs, taps = KH * KW;
    size_t ow = arena((size_t)taps * Cs), oa = arena(3 * (size_t)Cs);
    macs(C, 1, KH, KW, ti.H, ti.W);
    Op op;
    op.name = name;
    op.pack = [=](const HostWeights& w, float* A) {
        const float* wt = w.get(conv + ".weight");  // [C][1][KH][KW]
        std::memset(A + ow, 0, sizeof(float) * taps * Cs);
        std::memset(A + oa, 0, sizeof(float) * 3 * Cs);
        for (int c = 0; c < C; ++c)
            for (int t = 0; t < taps; ++t) A[ow + (size_t)t * Cs + c] = wt[(size_t)c * taps + t];
        if (bn.empty()) affine_bias(w, conv + ".bias", C, A + oa, A + oa + Cs);
        else affine_unit(w, conv, bn, C, A + oa, A + oa + Cs);
        if (!slope.empty()) copy_vec(w, slope, C, A + oa + 2 * Cs);
    };
    op.launch = [=](int n, const float*, float*, cudaStream_t st) {
        DwParams p;
        p.in = P->ws + ti.off; p.out = P->ws + to.off; p.H = ti.H; p.W = ti.W; p.Cs = Cs;
        p.total = n * ti.H * ti.W * (Cs / 4);
        p.KH = KH; p.KW = KW; p.dil_h = dil; p.dil_w = dil;
        p.pad_h = dil * (KH / 2); p.pad_w = dil * (KW / 2);
        p.w = P->arena + ow; p.scale = P->arena + oa; p.shift = P->arena + oa + Cs;
        p.slope = slope.empty() ? nullptr : P->arena + oa + 2 * Cs; p.act = act;
        dw_kernel<T><<<cdiv(p.total, 256), 256, 0, st>>>(p);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
}

template <typename T>
void Builder<T>::emit_convt4(const std::string& name, const std::string& prefix, int tin, int tout) {
    isg_plan* P = this->P;
    const TensorD ti = td(tin), to = td(tout);
    const int C = ti.C, Cs = ti.Cs;
    size_t ow = arena((size_t)16 * Cs * C), oa = arena(2 * (size_t)C);
    macs(C, C, 4, 4, ti.H, ti.W);  // each input pixel meets all 16 taps once
    Op op;
    op.name = name;
    op.pack = [=](const HostWeights& w, float* A) {
        pack_convt(w.get(prefix + ".convs.1.weight"), C, C, Cs, 4, A + ow);
        affine_unit(w, prefix + ".convs.1", prefix + ".convs.2", C, A + oa, A + oa + C);
    };
    const size_t smem = (size_t)16 * Cs * C * sizeof(float);
    op.launch = [=](int n, const float*, float*, cudaStream_t st) {
        ConvT4Params p;
        p.in = P->ws + ti.off; p.out = P->ws + to.off; p.Hi = ti.H; p.Wi = ti.W;
        p.Cs_in = Cs; p.Cs_out = to.Cs; p.total_pix = n * to.H * to.W;
        p.w = P->arena + ow; p.scale = P->arena + oa; p.shift = P->arena + oa + C;
        if (C == 16) convt4_kernel<T, 16><<<cdiv(p.total_pix, 128), 128, smem, st>>>(p);
        else convt4_kernel<T, 4><<<cdiv(p.total_pix, 128), 128, smem, st>>>(p);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
}

template <typename T>
void Builder<T>::emit_tail(int tin) {
    isg_plan* P = this->P;
    const TensorD ti = td(tin);
    size_t owt = arena(64 * 16 * 4), ow3 = arena(36), ob = arena(8);
    macs(4, 16, 8, 8, ti.H, ti.W);
    macs(1, 4, 3, 3, 4 * ti.H, 4 * ti.W);
    Op op;
    op.name = "bottle6_1+bottle6_2";
    op.pack = [=](const HostWeights& w, float* A) {
        pack_convt(w.get("bottle6_1.weight"), 16, 4, 16, 8, A + owt);
        const float* w3 = w.get("bottle6_2.weight");  // [1][4][3][3]
        for (int c = 0; c < 4; ++c)
            for (int t = 0; t < 9; ++t) A[ow3 + t * 4 + c] = w3[c * 9 + t];
        std::memcpy(A + ob, w.get("bottle6_1.bias"), 4 * sizeof(float));
        A[ob + 4] = w.get("bottle6_2.bias")[0];
    };
    op.launch = [=](int n, const float*, float* logits, cudaStream_t st) {
        TailParams p;
        p.in = P->ws + ti.off; p.Hq = ti.H; p.Wq = ti.W; p.Cs_in = ti.Cs; p.logits = logits;
        p.wt = P->arena + owt; p.bt = P->arena + ob; p.w3 = P->arena + ow3;
        p.b3 = P->arena_host[ob + 4];
        dim3 grid(cdiv(4 * ti.W, TL), cdiv(4 * ti.H, TL), n);
        tail_kernel<T><<<grid, 256, 0, st>>>(p);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
}

template <typename T>
void Builder<T>::emit_pw(const PwSpec& s) {
    isg_plan* P = this->P;
    const TensorD to = td(s.out);
    const int Cout = to.C, CO = pw_chunk(Cout), CoutPad = round_up(Cout, CO);
    auto dims = [&](const Src& r, int& C0, int& Cs0, int& C1, int& Cs1) {
        C0 = td(r.t0).C; Cs0 = td(r.t0).Cs;
        C1 = r.t1 >= 0 ? td(r.t1).C : 0; Cs1 = r.t1 >= 0 ? td(r.t1).Cs : 0;
    };
    int a0, as0, a1, as1, b0 = 0, bs0 = 0, b1 = 0, bs1 = 0;
    dims(s.a, a0, as0, a1, as1);
    if (s.b_kind == BR_GEMM) dims(s.b, b0, bs0, b1, bs1);
    const int Ka = as0 + as1, Kb = bs0 + bs1;
    size_t owa = arena((size_t)Ka * CoutPad), oaa = arena(3 * (size_t)CoutPad);
    size_t owb = arena((size_t)Kb * CoutPad), oab = arena(2 * (size_t)CoutPad), oso = arena(CoutPad);
    macs(Cout, a0 + a1, 1, 1, to.H, to.W);
    if (s.b_kind == BR_GEMM) macs(Cout, b0 + b1, 1, 1, to.H, to.W);
    Op op;
    op.name = s.name;
    const PwSpec sp = s;
    op.pack = [=](const HostWeights& w, float* A) {
        std::memset(A + oaa, 0, sizeof(float) * 3 * CoutPad);
        std::memset(A + oab, 0, sizeof(float) * 2 * CoutPad);
        std::memset(A + oso, 0, sizeof(float) * CoutPad);
        pack_pw(w.get(sp.wa + ".weight"), Cout, CoutPad, a0, as0, a1, as1, A + owa);
        if (sp.a_bias_only) affine_bias(w, sp.wa + ".bias", Cout, A + oaa, A + oaa + CoutPad);
        else affine_unit(w, sp.wa, sp.bna, Cout, A + oaa, A + oaa + CoutPad);
        if (!sp.slope_a.empty()) copy_vec(w, sp.slope_a, Cout, A + oaa + 2 * CoutPad);
        if (sp.b_kind == BR_GEMM) {
            pack_pw(w.get(sp.wb + ".weight"), Cout, CoutPad, b0, bs0, b1, bs1, A + owb);
            if (sp.b_bias_only) affine_bias(w, sp.wb + ".bias", Cout, A + oab, A + oab + CoutPad);
            else affine_unit(w, sp.wb, sp.bnb, Cout, A + oab, A + oab + CoutPad);
        }
        if (!sp.slope_out.empty()) copy_vec(w, sp.slope_out, Cout, A + oso);
    };
    auto mk = [=](const Src& r) {
        PwSrc q;
        q.p0 = P->ws + P->tensors[r.t0].off; q.Cs0 = P->tensors[r.t0].Cs;
        q.p1 = r.t1 >= 0 ? P->ws + P->tensors[r.t1].off : nullptr;
        q.Cs1 = r.t1 >= 0 ? P->tensors[r.t1].Cs : 0;
        q.mode = r.mode;
        return q;
    };
    const size_t smem = (size_t)(Ka + Kb) * CO * sizeof(float);
    op.launch = [=](int n, const float*, float*, cudaStream_t st) {
        PwParams p;
        p.total_pix = n * to.H * to.W; p.Ho = to.H; p.Wo = to.W;
        p.a = mk(sp.a); p.wa = P->arena + owa; p.scale_a = P->arena + oaa; p.shift_a = P->arena + oaa + CoutPad;
        p.slope_a = sp.slope_a.empty() ? nullptr : P->arena + oaa + 2 * CoutPad; p.act_a = sp.act_a;
        p.b_kind = sp.b_kind;
        if (sp.b_kind == BR_GEMM) p.b = mk(sp.b); else { p.b = PwSrc{nullptr, 0, nullptr, 0, 0}; }
        p.wb = P->arena + owb; p.scale_b = P->arena + oab; p.shift_b = P->arena + oab + CoutPad;
        p.res = sp.res >= 0 ? P->ws + P->tensors[sp.res].off : nullptr;
        p.res_Cs = sp.res >= 0 ? P->tensors[sp.res].Cs : 0;
        p.act_out = sp.act_out; p.slope_out = sp.slope_out.empty() ? nullptr : P->arena + oso;
        p.out = P->ws + to.off; p.out_Cs = to.Cs; p.Cout = Cout; p.CoutPad = CoutPad;
        p.pool_out = sp.pool_out >= 0 ? P->ws + P->tensors[sp.pool_out].off : nullptr;
        p.pool_Cs = sp.pool_out >= 0 ? P->tensors[sp.pool_out].Cs : 0;
        dim3 grid(cdiv(p.total_pix, 128), CoutPad / CO);
        if (CO == 32) pw_kernel<T, 32><<<grid, 128, smem, st>>>(p);
        else if (CO == 16) pw_kernel<T, 16><<<grid, 128, smem, st>>>(p);
        else pw_kernel<T, 4><<<grid, 128, smem, st>>>(p);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
}
