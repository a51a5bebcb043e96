// Builder<T>::emit_tc -- lowers one GEMM-shaped conv to tc_conv_kernel (bf16 mode only).
// Included inside namespace isg::{anon}.

struct TcMaps {
    CUtensorMap a[3], out;
};

template <typename T>
bool Builder<T>::emit_tc(const TcSpec& s) {
    isg_plan* P = this->P;
    if (sizeof(T) != 2 || !P->use_tc) return false;
    const TensorD to = td(s.out);
    const int N = to.Cs;                       // Cs is a multiple of 16 in bf16 mode
    const int h = s.kind == TC_C3X3 ? 1 : 0;
    const int W = to.W, H = to.H, Wp = W + 2 * h;
    if (Wp > 256 || N > 256 || N % 16) return false;
    // K slices
    int nsrc = (int)s.parts.size(), planes[3] = {0, 0, 0}, nk16 = 0;
    if (nsrc < 1 || nsrc > 3) return false;
    for (int i = 0; i < nsrc; ++i) {
        const TensorD ti = td(s.parts[i].tensor);
        planes[i] = (s.kind == TC_S2D ? 4 : 1) * ti.Cs / 8;
        nk16 += planes[i] / 2;
    }
    if (nk16 > 24) return false;
    const int ntaps = s.kind == TC_C3X3 ? 9 : 1;
    const int w_bytes = ntaps * nk16 * N * 32;
    // tile height: TMEM columns n_mt*N <= 256 (two CTAs per SM), shared memory <= 100 KB
    int TH = 0, n_mt = 0, a_bytes = 0, src_off[3] = {0, 0, 0};
    size_t smem = 0;
    const int mt_cap = std::max(1, 256 / N);
    for (int th = std::min(H, 255 - 2 * h); th >= 1; --th) {
        int mt = cdiv(th * Wp, 128);
        if (mt > mt_cap) continue;
        int R = th + 2 * h, plane = R * Wp * 16, off = 0;
        for (int i = 0; i < nsrc; ++i) { src_off[i] = off; off = round_up(off + planes[i] * plane, 128); }
        int overrun = std::max(0, mt * 128 + 2 * h * Wp + 2 * h - R * Wp) * 16;
        int ab = round_up(off + overrun, 128);
        size_t need = (size_t)w_bytes + ab + (size_t)(N / 8) * th * Wp * 16 + 64;
        if (need > 100 * 1024 && th > 1) continue;
        TH = th; n_mt = mt; a_bytes = ab; smem = need;
        break;
    }
    if (!TH || smem > 200 * 1024) return false;
    const int R = TH + 2 * h, plane_bytes = R * Wp * 16;

    // weights + epilogue vectors in the arena
    size_t ow = arena((size_t)w_bytes / 4), osh = arena(N), osl = arena(N);
    for (auto& pt : s.parts) {
        const TensorD ti = td(pt.tensor);
        macs(to.C, ti.C, s.kind == TC_S2D ? 2 : (h ? 3 : 1), s.kind == TC_S2D ? 2 : (h ? 3 : 1), H, W);
    }
    std::vector<TcPart> parts = s.parts;
    std::vector<TensorD> tin;
    for (auto& pt : parts) tin.push_back(td(pt.tensor));
    const int Cout = to.C, kind = s.kind;
    const std::string slope = s.slope;
    Op op;
    op.name = s.name + " [tcgen05]";
    op.pack = [=](const HostWeights& w, float* A) {
        uint16_t* wb = reinterpret_cast<uint16_t*>(A + ow);
        std::memset(wb, 0, w_bytes);
        std::vector<float> shift(N, 0.f), scale(N), sh(N);
        int kbase = 0;  // K index (storage channels) where this part starts
        for (size_t i = 0; i < parts.size(); ++i) {
            const TcPart& pt = parts[i];
            const TensorD& ti = tin[i];
            const float* wt = w.get(pt.conv + ".weight");
            if (pt.bias_only) affine_bias(w, pt.conv + ".bias", Cout, scale.data(), sh.data());
            else affine_unit(w, pt.conv, pt.bn, Cout, scale.data(), sh.data());
            if (pt.cin_off == 0) for (int o = 0; o < Cout; ++o) shift[o] += sh[o];
            const int taps_w = kind == TC_S2D ? 4 : (kind == TC_C3X3 ? 9 : 1);
            const int cin_total = pt.cin_total;
            for (int o = 0; o < Cout; ++o)
                for (int c = 0; c < ti.C; ++c)
                    for (int t = 0; t < taps_w; ++t) {
                        float v = wt[((size_t)o * cin_total + pt.cin_off + c) * taps_w + t] * scale[o];
                        // K position and MMA tap of this weight
                        int k, tap;
                        if (kind == TC_S2D) { k = kbase + (t >> 1) * 2 * ti.Cs + (t & 1) * ti.Cs + c; tap = 0; }
                        else { k = kbase + c; tap = kind == TC_C3X3 ? t : 0; }
                        size_t blk = (size_t)tap * nk16 + k / 16;
                        size_t idx = blk * (size_t)N * 16 + (size_t)((k % 16) / 8) * N * 8 + (size_t)o * 8 + (k % 8);
                        wb[idx] = tc::f2bf(v);
                    }
            kbase += (kind == TC_S2D ? 4 : 1) * ti.Cs;
        }
        std::memcpy(A + osh, shift.data(), sizeof(float) * N);
        std::memset(A + osl, 0, sizeof(float) * N);
        if (!slope.empty()) copy_vec(w, slope, Cout, A + osl);
    };

    auto maps = std::make_shared<TcMaps>();
    const int maxB = P->max_batch;
    P->finalizers.push_back([=]() -> int {
        for (size_t i = 0; i < tin.size(); ++i) {
            const TensorD& ti = tin[i];
            int rc = kind == TC_S2D
                         ? tc::make_s2d_map(&maps->a[i], P->ws + ti.off, maxB, ti.H, ti.W, ti.Cs, Wp, R, s.name.c_str())
                         : tc::make_act_map(&maps->a[i], P->ws + ti.off, maxB, ti.H, ti.W, ti.Cs, Wp, R, s.name.c_str());
            if (rc) return rc;
        }
        for (size_t i = tin.size(); i < 3; ++i) maps->a[i] = maps->a[0];
        return tc::make_act_map(&maps->out, P->ws + to.off, maxB, H, W, to.Cs, Wp, TH, s.name.c_str());
    });

    tc::TcParams q;
    std::memset(&q, 0, sizeof q);
    q.n_src = nsrc;
    int kk = 0;
    for (int i = 0; i < nsrc; ++i) {
        q.planes[i] = planes[i]; q.src_off[i] = src_off[i];
        for (int j = 0; j < planes[i] / 2; ++j) q.a_off[kk++] = src_off[i] + 2 * j * plane_bytes;
    }
    q.Wp = Wp; q.R = R; q.hx = h; q.hy = h;
    q.batch_in_y = kind == TC_S2D ? 1 : 0; q.Hin = H;   // the s2d map's y dimension is b*Ho + oy
    q.ntaps = ntaps;
    for (int t = 0; t < ntaps; ++t) q.tap_off[t] = h ? (t / 3) * Wp + (t % 3) : 0;
    q.nk16 = nk16; q.TH = TH; q.n_mt = n_mt; q.N = N; q.H = H; q.W = W;
    q.tiles_per_img = cdiv(H, TH);
    q.act = s.act; q.w_bytes = w_bytes; q.stage_pix = TH * Wp; q.a_bytes = a_bytes;
    q.res_mode = s.res >= 0 ? s.res_mode : tc::RES_NONE;
    const int res_t = s.res;
    cudaFuncSetAttribute(tc::tc_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    op.launch = [=](int n, const float*, float*, cudaStream_t st) {
        tc::TcParams p = q;
        p.shift = P->arena + osh; p.slope = P->arena + osl;
        p.wpack = reinterpret_cast<const bf16*>(P->arena + ow);
        if (res_t >= 0) {
            const TensorD& tr = P->tensors[res_t];
            p.res = reinterpret_cast<const bf16*>(P->ws + tr.off);
            p.res_Cs = tr.Cs; p.res_H = tr.H; p.res_W = tr.W;
        }
        tc::tc_conv_kernel<<<n * p.tiles_per_img, tc::TC_THREADS, smem, st>>>(maps->a[0], maps->a[1], maps->a[2],
                                                                            maps->out, p);
        return cudaGetLastError();
    };
    P->ops.push_back(op);
    return true;
}
