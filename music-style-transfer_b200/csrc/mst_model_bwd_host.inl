// mst_model_bwd_host.inl — host-side orchestration of losses + backward pass (included by mst_model.cu inside
// its anonymous namespace, after the forward pass).  Hand-derived reverse of forward_extract / forward_song_info /
// forward_apply; no autograd.  The same function runs twice: once "dry" (Launcher::dry, Arena::dry) to size the
// workspace, once for real — so the plan and the run can never disagree.
//
// Parameter gradients are ACCUMULATED into the flat gradient buffer (+=), which reproduces train-model.py's
// iter_size accumulation (:126,151-154) and makes a batch of N samples the sum of N batch-1 backward passes.

struct TrainIn {
    const float *instr_target;   // (N, 41)
    const float *bpm_target;     // (N)
    int normalize;
    float *leaves;               // (N, kLossLeaves) device output
};

struct Bwd {
    Launcher &L; Ctx &c; Arena &ar;
    Arena &zar;                               // region of the buffers that must start at zero (see Z)
    float *tmpA = nullptr, *tmpB = nullptr;   // ping-pong scratch of the staged reductions

    // Buffers that must start at zero come from a region of their own (zar), cleared by ONE launch before the pass starts
    // instead of one fill per buffer on the critical chain.
    float *A(long long n) { return ar.take(n); }
    float *Z(long long n) { return zar.take(n); }
    // weight gradients / bias column sums straight into the flat gradient buffer: leaf work, off the dX chain (Launcher::leaf).
    // leaf_dw = false where another kernel of the caller's stream accumulates into the same parameter.
    bool leaf_dw = true;
    bool in_leaf_term = false;    // inside branch_bwd's leaf re-entry (see there)
    void dw(const float *dY, long long ldy, const float *X, long long ldx, float *gW, long long ldw, int M, int Nout, int K) {
        if (leaf_dw) L.leaf([&] { L.linear_dw(dY, ldy, X, ldx, gW, ldw, M, Nout, K); });
        else L.linear_dw(dY, ldy, X, ldx, gW, ldw, M, Nout, K);
    }
    void bias_grad(const float *dY, float *gb, long long rows, int cols, float *gb2 = nullptr) {
        if (leaf_dw) L.leaf([&] { L.colsum(dY, gb, rows, cols, gb2); });
        else L.colsum(dY, gb, rows, cols, gb2);
    }

    // out[kept rows, cols] += sum over the dropped axes of in[(n,c,b,t,f), cols]; runs of dropped axes are reduced
    // innermost first, each as one split "sum over the middle axis".  `out` must hold valid data (zeros or a gradient).
    void reduce_to(const float *in, Dims5 d, int cols, int mask, float *out) {
        int dims[5] = {d.n, d.c, d.b, d.t, d.f};
        bool keep[5] = {(mask & 16) != 0, (mask & 8) != 0, (mask & 4) != 0, (mask & 2) != 0, (mask & 1) != 0};
        for (int i = 0; i < 5; i++) if (dims[i] == 1) keep[i] = true;   // nothing to sum
        // count runs
        int n_runs = 0;
        for (int i = 0; i < 5; i++) if (!keep[i] && (i == 0 || keep[i - 1])) n_runs++;
        if (n_runs == 0) { long long n = cols; for (int i = 0; i < 5; i++) n *= dims[i]; L.axpy(out, in, n); return; }
        const float *cur = in;
        int run_i = 0;
        for (int hi = 4; hi >= 0; hi--) {
            if (keep[hi]) continue;
            int lo = hi;
            while (lo > 0 && !keep[lo - 1]) lo--;
            long long Aouter = 1, Lr = 1, inner = cols;
            for (int i = 0; i < lo; i++) Aouter *= dims[i];
            for (int i = lo; i <= hi; i++) Lr *= dims[i];
            for (int i = hi + 1; i < 5; i++) inner *= dims[i];
            run_i++;
            float *dst;
            if (run_i == n_runs) dst = out;
            else { dst = (cur == tmpA) ? tmpB : tmpA; L.fill(dst, Aouter * inner, 0.f); }
            reduce_split(cur, dst, Aouter, (int)Lr, inner);
            for (int i = lo; i <= hi; i++) { dims[i] = 1; keep[i] = true; }
            cur = dst;
            hi = lo;
        }
    }
    void reduce_split(const float *in, float *out, long long Aouter, int Lr, long long inner) {
        if (L.dry) return;
        const long long total = Aouter * inner;
        long long chunks = std::max<long long>(1, std::min<long long>((300000 + total - 1) / total, (Lr + 3) / 4));
        const int l_per = (int)((Lr + chunks - 1) / chunks);
        chunks = (Lr + l_per - 1) / l_per;
        dim3 grid((unsigned)grid_for(total, 256, 148 * 8), (unsigned)chunks);
        k_reduce_mid_split<<<grid, 256, 0, L.st>>>(in, out, Aouter, Lr, inner, l_per); L.check();
    }

    // Y = act(X W^T + b), rows M: given dY (gradient w.r.t. Y, overwritten with the pre-activation gradient):
    //   gb += colsum, gW += dpre^T X, dX (+)= dpre W.   Yact == nullptr: no activation.
    void lin_bwd(float *dY, const float *Yact, int M, int Nout, const float *X, long long ldx, int K,
                 ParamId w, int w_col_off, int ldw, int bias /* ParamId or -1 */, float *dX, long long lddx, int acc_dx) {
        if (Yact && bias >= 0) L.act_grad_colsum(dY, Yact, dY, GRD(c, (ParamId)bias), M, Nout);
        else {
            if (Yact) L.act_grad(dY, Yact, dY, (long long)M * Nout);
            if (bias >= 0) bias_grad(dY, GRD(c, (ParamId)bias), M, Nout);
        }
        dw(dY, Nout, X, ldx, GRD(c, w) + w_col_off, ldw, M, Nout, K);
        if (dX) L.linear_dx(dY, Nout, PRM(c, w) + w_col_off, ldw, dX, lddx, M, Nout, K, acc_dx);
    }
    // one term of a decomposed "Linear over a broadcast concat": z_s = W[:, off:off+K] a_s, a_s = leaky(X W2^T + b2).
    // dZ (rows x Nout) is the gradient w.r.t. z_s at the source's own row space.
    void branch_bwd(const float *dZ, int rows, int Nout, ParamId W, int col_off, int ldw, const float *a_s, int K,
                    const float *X, long long ldx, int Kin, ParamId w2, ParamId b2, float *dX, long long lddx, int acc_dx) {
        if (!dX && leaf_dw && !in_leaf_term) {
            // a term whose input needs no gradient (instrument features, mode, bpm, the grids): nothing downstream reads what
            // it computes, so all of it is leaf work (Launcher::leaf), off the chain of the caller's branch
            in_leaf_term = true;
            L.leaf([&] { branch_bwd(dZ, rows, Nout, W, col_off, ldw, a_s, K, X, ldx, Kin, w2, b2, nullptr, 0, 0); });
            in_leaf_term = false;
            return;
        }
        float *d_a = A((long long)rows * K);
        dw(dZ, Nout, a_s, K, GRD(c, W) + col_off, ldw, rows, Nout, K);
        L.linear_dx(dZ, Nout, PRM(c, W) + col_off, ldw, d_a, K, rows, Nout, K, 0);
        lin_bwd(d_a, a_s, rows, K, X, ldx, Kin, w2, 0, Kin, b2, dX, lddx, acc_dx);
    }
    // one LSTM direction backward: dout -> dgates (recurrence), then the gradients of W_ih, W_hh, both biases and
    // dX (+)= dgates W_ih.  The two halves are separate so that independent recurrences can share one launch.
    LstmArgs lstm_bwd_args(LstmBuf &lb, const float *dout, long long out_ld, int out_off, int reverse, ParamId whh) {
        LstmArgs a{};
        a.w = PRM(c, whh); a.dout = dout; a.out_ld = out_ld; a.out_off = out_off;
        a.gates = lb.gates; a.cells = lb.cells; a.hprev = lb.hprev; a.dgates = lb.dgates;
        a.S = lb.S; a.T = lb.T; a.H = lb.H; a.reverse = reverse;
        return a;
    }
    void lstm_bwd_grads(LstmBuf &lb, const float *X, ParamId wih, ParamId whh, ParamId bih, ParamId bhh, float *dX, int acc_dx) {
        const int rows = lb.S * lb.T, H4 = 4 * lb.H;
        L.gemm_cls = 16;
        dw(lb.dgates, H4, X, lb.I, GRD(c, wih), lb.I, rows, H4, lb.I);
        dw(lb.dgates, H4, lb.hprev, lb.H, GRD(c, whh), lb.H, rows, H4, lb.H);
        bias_grad(lb.dgates, GRD(c, bih), rows, H4, GRD(c, bhh));      // b_ih and b_hh see the same gradient
        if (dX) L.linear_dx(lb.dgates, H4, PRM(c, wih), lb.I, dX, lb.I, rows, H4, lb.I, acc_dx);
        L.gemm_cls = 1;
    }
    void lstm_bwd(LstmBuf &lb, const float *dout, long long out_ld, int out_off, int reverse, const float *X,
                  ParamId wih, ParamId whh, ParamId bih, ParamId bhh, float *dX, int acc_dx) {
        L.lstm(lstm_bwd_args(lb, dout, out_ld, out_off, reverse, whh), true);
        lstm_bwd_grads(lb, X, wih, whh, bih, bhh, dX, acc_dx);
    }
    void combine_bwd(const float *x, long long n_stride, long long k_stride, int K, long long M, const float *ss,
                     const float *out, const float *dout, float *dx, int accumulate) {
        CombineArgs a{};
        a.x = x; a.n_stride = n_stride; a.k_stride = k_stride; a.N = c.N; a.K = K; a.M = M; a.ss = (float *)ss;
        a.out = (float *)out; a.dout = dout; a.stats = A((long long)c.N * (K + 1)); a.dx = dx; a.accumulate = accumulate;
        L.combine_bwd(a);
    }
};

// ---------------------------------------------------------------------------------------------- losses
struct LossBufs { double *sums_p, *sums_u; float *coef_p, *coef_u, *d_instr, *d_mode, *d_bpm_pre; };

static LossBufs run_losses(Launcher &L, Arena &ar, int N, int U, long long cells_p, long long cells_u,
                           const float *pred_p, const float *tgt_p, const float *pred_u, const float *tgt_u,
                           const float *instr, const float *instr_t, const float *mode, const float *mode_t,
                           const float *bpm, const float *bpm_t, int normalize, float *leaves, bool want_grad) {
    LossBufs b{};
    b.sums_p = (double *)ar.take((long long)N * kLossSums * 2);
    b.sums_u = U ? (double *)ar.take((long long)N * kLossSums * 2) : nullptr;
    if (want_grad) {
        b.coef_p = ar.take((long long)N * kLossCoefs); b.coef_u = U ? ar.take((long long)N * kLossCoefs) : nullptr;
        b.d_instr = ar.take((long long)N * 41); b.d_mode = ar.take((long long)N * 2); b.d_bpm_pre = ar.take(N);
    }
    if (L.dry) return b;
    L.fill((float *)b.sums_p, (long long)N * kLossSums * 2, 0.f);
    if (U) L.fill((float *)b.sums_u, (long long)N * kLossSums * 2, 0.f);
    if (U) {                                         // the unpitched sums on branch 0, beside the pitched ones
        L.wait(0, -1); L.on(0);
        LossSumsArgs a{pred_u, tgt_u, cells_u, b.sums_u};
        dim3 g((unsigned)std::max<long long>(1, std::min<long long>((cells_u + 8191) / 8192, 148 * 4)), N);
        k_loss_sums<2><<<g, 256, 0, L.st>>>(a); L.check();
        L.on(-1);
    }
    {
        LossSumsArgs a{pred_p, tgt_p, cells_p, b.sums_p};
        dim3 g((unsigned)std::max<long long>(1, std::min<long long>((cells_p + 8191) / 8192, 148 * 4)), N);
        k_loss_sums<5><<<g, 256, 0, L.st>>>(a); L.check();
    }
    if (U) L.wait(-1, 0);
    LossTreeArgs t{};
    t.sums_p = b.sums_p; t.sums_u = b.sums_u; t.instr = instr; t.instr_target = instr_t; t.mode = mode; t.mode_target = mode_t;
    t.bpm = bpm; t.bpm_target = bpm_t; t.leaves = leaves; t.coef_p = b.coef_p; t.coef_u = b.coef_u;
    t.d_instr = b.d_instr; t.d_mode = b.d_mode; t.d_bpm_pre = b.d_bpm_pre; t.N = N; t.normalize = normalize; t.seed = 1.f;
    k_loss_tree<<<(N + 63) / 64, 64, 0, L.st>>>(t); L.check();
    return b;
}

// ---------------------------------------------------------------------------------------------- backward
static void backward_impl(Launcher &L, Ctx &c, Arena &ar, Arena &zar, const TrainIn &ti) {
    const int N = c.N, C = c.C, B = c.B, T = c.T, NC = N * C, U = c.U;
    const long long Rb = c.Rb, Rf = c.Rf, Rbf = c.Rbf, Nb = c.Nb;
    Bwd w{L, c, ar, zar};
    w.tmpA = ar.take(Rf * 32 + Rb * 120); w.tmpB = ar.take(Rf * 32 + Rb * 120);
    // the tail kernels read the small weights from the constant banks: reload them from THESE parameters (another
    // forward / predict_apply with other parameters may have run since the forward pass of this step)
    load_me_constants(L, c);
    load_psa_constants(L, c);

    // ---------------- get_total_loss (model.py:935-997) + its derivative
    // NOTE the positional quirk of train-model.py:115-122 is resolved by the Python caller: here "mode" is the
    // cross-entropy pair and "bpm" the squared-error pair.
    LossBufs lb = run_losses(L, ar, N, U, (long long)C * B * T * 560, (long long)B * T * 470, c.psa_out, c.P, c.usa_out, c.Ug,
                             c.sim_instr, ti.instr_target, c.sim_mode, c.mode, c.sim_bpm, ti.bpm_target, ti.normalize,
                             ti.leaves, true);

    float *d_style = w.Z((long long)N * 256), *d_rhythm = w.Z(Rbf * 32);
    float *d_pce_beats = w.Z(Rb * 64), *d_pce_bars = w.Z(Nb * 128);
    float *d_uce_beats = U ? w.Z(Nb * T * 64) : nullptr, *d_uce_bars = U ? w.Z(Nb * 128) : nullptr;

    // Branch 0: the unpitched applier's backward and the SongInfoModel heads, beside the pitched applier's tail kernels on the
    // caller's stream.  Both sides add into d_style / d_rhythm: branch 0 first, the pitched applier after the join below.
    L.wait(0, -1); L.on(0);
    // ---------------- UnpitchedStyleApplier (model.py:703-724)
    if (U) {
        float *dH = w.A(Rbf * 376), *dZs = w.Z((long long)N * 10 * 376);
        if (!L.dry) {
            UsaBwdArgs a{};
            a.f.bias = PRM(c, usa_nl_b); a.f.Wl = PRM(c, usa_lin_w); a.f.bl = PRM(c, usa_lin_b); a.f.Zr = c.usa_zr; a.f.Zs = c.usa_zs;
            a.f.H = c.usa_h; a.f.out = nullptr; a.f.N = N; a.f.B = B; a.f.T = T;
            a.target = c.Ug; a.coef = lb.coef_u; a.dHpre = dH; a.gWl = GRD(c, usa_lin_w); a.gbl = GRD(c, usa_lin_b);
            k_usa_bwd<<<(int)std::min<long long>((Rbf + kUsaWarps - 1) / kUsaWarps, 148 * 4), kUsaWarps * 32, 0, L.st>>>(a); L.check();
        }
        w.bias_grad(dH, GRD(c, usa_nl_b), Rbf, 376);
        w.reduce_to(dH, Dims5{N, 1, B, T, 10}, 376, M_N | M_F, dZs);
        // style branch: a_s = usa_s650 viewed (N*10, 65); its producer is style_linear 256 -> 650 on N rows
        {
            float *d_a = w.A((long long)N * 650);
            w.dw(dZs, 376, c.usa_s650, 65, GRD(c, usa_nl_w), 82, N * 10, 376, 65);
            L.linear_dx(dZs, 376, PRM(c, usa_nl_w), 82, d_a, 65, N * 10, 376, 65, 0);
            w.lin_bwd(d_a, c.usa_s650, N, 650, c.style, 256, 256, usa_sl_w, 0, 256, usa_sl_b, d_style, 256, 1);
        }
        w.branch_bwd(dH, (int)Rbf, 376, usa_nl_w, 65, 82, c.usa_r17, 17, c.rhythm, 32, 32, usa_rl_w, usa_rl_b, d_rhythm, 32, 1);
    }
    // ---------------- SongInfoModel (model.py:513-562)
    float *d_rf = w.Z((long long)N * 8);
    {
        auto head = [&](float *dOut, int nout, ParamId sw, ParamId sb, int ks, const float *sa, ParamId rw, ParamId rb, int kr,
                        const float *ra, ParamId fw, ParamId fb) {
            w.bias_grad(dOut, GRD(c, fb), N, nout);
            w.branch_bwd(dOut, N, nout, fw, 0, ks + kr, sa, ks, c.style, 256, 256, sw, sb, d_style, 256, 1);
            w.branch_bwd(dOut, N, nout, fw, ks, ks + kr, ra, kr, c.sim_rf, 8, 8, rw, rb, d_rf, 8, 1);
        };
        head(lb.d_instr, 41, sim_si_w, sim_si_b, 8, c.sim_si, sim_ri_w, sim_ri_b, 10, c.sim_ri, sim_i_w, sim_i_b);
        head(lb.d_mode, 2, sim_sm_w, sim_sm_b, 2, c.sim_sm, sim_rm_w, sim_rm_b, 2, c.sim_rm, sim_m_w, sim_m_b);
        head(lb.d_bpm_pre, 1, sim_sb_w, sim_sb_b, 2, c.sim_sb, sim_rb_w, sim_rb_b, 2, c.sim_rb, sim_b_w, sim_b_b);
    }
    // the two SongInfoModel recurrences continue from the heads on a branch of their own (two serial kernels of n_bars steps);
    // their gradient w.r.t. the rhythm goes to a buffer of its own and is added where the rhythm combine needs it
    float *d_rhythm_sim = w.A(Rbf * 32);
    L.wait(kBranchSim, 0); L.on(kBranchSim);
    {
        float *d_h2 = w.Z(Nb * 8), *d_last1 = w.A(Nb * 9), *d_h1 = w.Z(Nb * T * 9);
        L.put_last(d_rf, d_h2, N, B, 8);
        w.lstm_bwd(c.sim_l2, d_h2, 8, 0, 0, c.sim_last1, sim_br_wih, sim_br_whh, sim_br_bih, sim_br_bhh, d_last1, 0);
        L.put_last(d_last1, d_h1, Nb, T, 9);
        w.lstm_bwd(c.sim_l1, d_h1, 9, 0, 0, c.rhythm, sim_bl_wih, sim_bl_whh, sim_bl_bih, sim_bl_bhh, d_rhythm_sim, 0);
    }
    L.on(-1);
    // ---------------- PitchedStyleApplier (model.py:624-675)
    float *d_melody = w.A(Rbf * 448);
    float *dZr = w.A(Rbf * 450), *dZi = w.Z((long long)NC * 450), *dZs = w.A((long long)N * 450);
    {
        // the tcgen05 formulation of this backward (k_psa_bwd_tc, mst_model_psa_tc.cuh): on request only, see Launcher::psa_bwd_tc
        const bool tc_tail = L.psa_bwd_tc() && ptc_bwd_smem(C) <= 72 * 1024;
        if (!L.dry && tc_tail) {
            PsaBwdTcArgs a{};
            a.Zr = c.psa_zr; a.Zci = c.psa_zci; a.pred = c.psa_out; a.target = c.P; a.coef = lb.coef_p;
            a.dZr = dZr; a.dZi = dZi; a.dps = d_melody; a.dps_stride = 448; a.gWl = GRD(c, psa_lin_w);
            a.N = N; a.C = C; a.B = B; a.T = T;
            a.ctas_per_n = (int)std::max<long long>(1, std::min<long long>((148 * 3 + N - 1) / N, (long long)B * T * 10));
            const size_t smem = ptc_bwd_smem(C);
            cudaFuncSetAttribute(k_psa_bwd_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k_psa_bwd_tc<<<N * a.ctas_per_n, kPtcThreads, smem, L.st>>>(a); L.check();
        }
        if (!L.dry && !tc_tail) {
            // the per-note channel sums of the pre-sigmoid gradient travel through the (not yet written) dmelody buffer
            PsaBwdArgs3 a{};
            a.Zr = c.psa_zr; a.Zci = c.psa_zci; a.melody = c.melody; a.target = c.P; a.coef = lb.coef_p;
            a.Wl = PRM(c, psa_lin_w);
            a.dZr = dZr; a.dZi = dZi; a.dps = d_melody; a.dps_stride = 448;
            a.gWl = GRD(c, psa_lin_w);
            a.N = N; a.C = C; a.B = B; a.T = T;
            a.ctas_per_n = (int)std::max<long long>(1, std::min<long long>(148 / N, ((long long)B * T * 10 + kPsaBwdWarps - 1) / kPsaBwdWarps));
            // fewer warps per CTA when the dZi accumulators of a many-channel song would not fit beside the staging
            int warps = kPsaBwdWarps;
            while (warps > 1 && sizeof(PsaBwdWarpSmem) * warps + (size_t)C * 1800 > 227 * 1024) warps--;
            const size_t smem = sizeof(PsaBwdWarpSmem) * warps + (size_t)C * 1800;
            a.warps = warps;
            cudaFuncSetAttribute(k_psa_bwd3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k_psa_bwd3<<<N * a.ctas_per_n, warps * 32, smem, L.st>>>(a); L.check();
        }
        if (!L.dry) {
            PsaMelBwdArgs m{};
            m.melody = c.melody; m.dps = d_melody; m.dps_stride = 448; m.dmelody = d_melody;
            m.gWm = GRD(c, psa_ml_w); m.gbm = GRD(c, psa_ml_b); m.gWl = GRD(c, psa_lin_w); m.gbl = GRD(c, psa_lin_b);
            m.units = Rbf;
            const size_t msmem = sizeof(PsaMelWarpSmem) * kPsaMelWarps;
            cudaFuncSetAttribute(k_psa_mel_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msmem);
            k_psa_mel_bwd<<<(int)std::min<long long>((Rbf + kPsaMelWarps - 1) / kPsaMelWarps, 148), kPsaMelWarps * 32, msmem, L.st>>>(m); L.check();
        }
        L.reduce_mid(dZi, dZs, N, C, 450, 0);
        w.bias_grad(dZs, GRD(c, psa_oct_b), N, 450);
    }
    // ---------------- MelodyEncoder (model.py:252-297) on branch 1: it needs d_melody only.  What it adds to d_pce_beats / d_pce_bars
    // lands before the PitchedRhythmEncoder's contributions (the caller's stream joins branch 1 in front of that block); its staged
    // reductions use a scratch pair of their own.
    L.wait(1, -1); L.on(1);
    {
        float *keepA = w.tmpA, *keepB = w.tmpB;
        w.tmpA = w.A(Rb * 120); w.tmpB = w.A(Rb * 120);
        float *stats = w.Z((long long)N * (C + 1)), *dod = w.A(Rb * 120), *dsum = w.A((long long)N * B * T * 448);
        if (!L.dry) {
            MeArgs2 a = me_args(c);
            a.dmelody = d_melody; a.dsum = dsum; a.stats = stats; a.dod = dod;
            a.gWc = GRD(c, me_ch_w); a.gbc = GRD(c, me_ch_b); a.gWl = GRD(c, me_lin_w); a.gbl = GRD(c, me_lin_b);
            k_me_bwd_stats2<<<(int)std::min<long long>(((long long)N * B * T + 7) / 8, 148 * 8), 256, 0, L.st>>>(a); L.check();
            // one CTA per SM at most (207 registers: one resident CTA anyway): every warp flushes its 80 register accumulators once
            k_me_bwd2<<<(int)std::min<long long>((Rb + 7) / 8, 148), 256, 0, L.st>>>(a); L.check();
            k_me_bwd_cells<<<(int)std::min<long long>((Rb + 7) / 8, 148 * 2), 256, 0, L.st>>>(a); L.check();
        }
        const Dims5 d{N, C, B, T, 1};
        float *dz1 = w.Z(Nb * 120), *dz2 = w.Z((long long)NC * 120);
        w.bias_grad(dod, GRD(c, me_oct_b), Rb, 120);
        w.reduce_to(dod, d, 120, M_N | M_B, dz1);
        w.reduce_to(dod, d, 120, M_N | M_C, dz2);
        w.branch_bwd(dod, (int)Rb, 120, me_oct_w, 0, 112, c.me_a1, 36, c.pce_beats, 64, 64, me_bl_w, me_bl_b, d_pce_beats, 64, 1);
        w.branch_bwd(dz1, (int)Nb, 120, me_oct_w, 36, 112, c.me_a2, 68, c.pce_bars, 128, 128, me_br_w, me_br_b, d_pce_bars, 128, 1);
        w.branch_bwd(dz2, NC, 120, me_oct_w, 104, 112, c.me_a3, 8, c.F, 51, 51, me_il_w, me_il_b, nullptr, 0, 0);
        w.tmpA = keepA; w.tmpB = keepB;
    }
    L.on(-1);
    // the style / rhythm gradients of branch 0 (unpitched applier, SongInfoModel heads) are in: the pitched applier adds its own
    L.wait(-1, 0);
    {
        w.branch_bwd(dZs, N, 450, psa_oct_w, 0, 88, c.psa_s66, 66, c.style, 256, 256, psa_sl_w, psa_sl_b, d_style, 256, 1);
        w.branch_bwd(dZr, (int)Rbf, 450, psa_oct_w, 66, 88, c.psa_r10, 10, c.rhythm, 32, 32, psa_rl_w, psa_rl_b, d_rhythm, 32, 1);
        w.branch_bwd(dZi, NC, 450, psa_oct_w, 76, 88, c.psa_i12, 12, c.F, 51, 51, psa_il_w, psa_il_b, nullptr, 0, 0);
    }
    // ---------------- StyleEncoder (model.py:179-200) on branch 0: d_style is complete once the heads above are through, and
    // the backward of its 192-wide recurrence is the longest serial kernel of the step; the caller's stream meanwhile runs the
    // SongInfoModel recurrences, the MelodyEncoder and both rhythm encoders.  Branch 0 only WRITES buffers of its own (d_bars,
    // d_bp, its parameter gradients): what it adds to d_pce_bars / d_uce_bars is applied after the join below.
    float *d_bars = w.A(Nb * 128), *d_bp = U ? w.A(2 * Nb * 128) : nullptr;
    L.wait(0, -1); L.on(0);
    {
        float *d_x = w.A((long long)NC * 256);
        w.combine_bwd(c.se_x, (long long)C * 256, 256, C, 256, c.se_ss, c.style, d_style, d_x, 0);
        L.act_grad(d_x, c.se_x, d_x, (long long)NC * 256);
        w.bias_grad(d_x, GRD(c, se_lin_b), NC, 256);
        float *dzn = w.Z((long long)N * 256);
        w.reduce_to(d_x, Dims5{N, C, 1, 1, 1}, 256, M_N, dzn);
        float *d_last = w.A((long long)N * 192), *d_h = w.Z(Nb * 192);
        w.dw(dzn, 256, c.se_last, 192, GRD(c, se_lin_w), 251, N, 256, 192);
        L.linear_dx(dzn, 256, PRM(c, se_lin_w), 251, d_last, 192, N, 256, 192, 0);
        w.branch_bwd(d_x, NC, 256, se_lin_w, 192, 251, c.se_a2, 39, c.F, 51, 51, se_il_w, se_il_b, nullptr, 0, 0);
        w.branch_bwd(dzn, N, 256, se_lin_w, 231, 251, c.se_a3, 13, c.mode, 2, 2, se_ml_w, se_ml_b, nullptr, 0, 0);
        w.branch_bwd(dzn, N, 256, se_lin_w, 244, 251, c.se_a4, 7, c.bpm, 1, 1, se_bp_w, se_bp_b, nullptr, 0, 0);
        L.put_last(d_last, d_h, N, B, 192);
        w.lstm_bwd(c.se_l, d_h, 192, 0, 0, c.bars, se_lstm_wih, se_lstm_whh, se_lstm_bih, se_lstm_bhh, d_bars, 0);
    }
    if (U) {
        const long long Mb = (long long)B * 128;
        w.combine_bwd(c.bars_pair, 2 * Mb, Mb, 2, Mb, c.bars_ss, c.bars_c, d_bars, d_bp, 0);
    }
    L.on(-1);
    // ---------------- combine(pitched, unpitched) of the rhythm (model.py:767)
    L.wait(-1, kBranchSim);
    L.axpy(d_rhythm, d_rhythm_sim, Rbf * 32);
    float *d_pre_out, *d_ure_out = nullptr;
    if (U) {
        const long long Mr = (long long)B * T * 320;
        float *d_rp = w.A(2 * Rbf * 32);
        d_pre_out = w.Z(Rbf * 32); d_ure_out = w.Z(Rbf * 32);
        w.combine_bwd(c.rh_pair, 2 * Mr, Mr, 2, Mr, c.rh_ss, c.rh_c, d_rhythm, d_rp, 0);
        if (!L.dry) { k_pair_unpack_add<<<grid_for(Rbf * 32), 256, 0, L.st>>>(d_rp, d_pre_out, d_ure_out, N, Mr); L.check(); }
    } else {
        d_pre_out = d_rhythm;
    }
    L.wait(-1, 1);                                   // MelodyEncoder's contributions to d_pce_beats / d_pce_bars are in
    if (U) {
        // ---------------- UnpitchedRhythmEncoder (model.py:418-443), on branch 1 (free again) beside the pitched one: it adds
        // into d_uce_beats / d_uce_bars only, which nothing else touches before the join in front of the bars' unpack below
        L.wait(1, -1); L.on(1);
        {
            const Dims5 d{N, 1, B, T, 10};
            float *d_x = w.A(Rbf * 32);
            w.combine_bwd(c.ure_x, (long long)B * T * 320, (long long)B * T * 320, 1, (long long)B * T * 320, c.ure_ss, c.ure_out, d_ure_out, d_x, 0);
            L.act_grad(d_x, c.ure_x, d_x, Rbf * 32);
            w.bias_grad(d_x, GRD(c, ure_lin_b), Rbf, 32);
            float *dz0 = w.Z(Nb * T * 32), *dz1 = w.Z(Nb * 32), *dz3 = w.Z((long long)N * 32);
            w.reduce_to(d_x, d, 32, M_N | M_B | M_T, dz0);
            w.reduce_to(d_x, d, 32, M_N | M_B, dz1);
            w.reduce_to(d_x, d, 32, M_N, dz3);
            w.branch_bwd(dz0, (int)(Nb * T), 32, ure_lin_w, 0, 109, c.ure_a1, 48, c.uce_beats, 64, 64, ure_bl_w, ure_bl_b, d_uce_beats, 64, 1);
            w.branch_bwd(dz1, (int)Nb, 32, ure_lin_w, 48, 109, c.ure_a2, 40, c.uce_bars, 128, 128, ure_br_w, ure_br_b, d_uce_bars, 128, 1);
            w.branch_bwd(d_x, (int)Rbf, 32, ure_lin_w, 88, 109, c.ure_a3, 16, c.Ug, 94, 94, ure_ch_w, ure_ch_b, nullptr, 0, 0);
            w.branch_bwd(dz3, N, 32, ure_lin_w, 104, 109, c.ure_a4, 5, c.bpm, 1, 1, ure_bp_w, ure_bp_b, nullptr, 0, 0);
        }
        L.on(-1);
    }
    // ---------------- PitchedRhythmEncoder (model.py:346-381)
    {
        const Dims5 d{N, C, B, T, 10};
        const long long M = (long long)B * T * 320;
        float *d_x = w.A(Rf * 32);
        w.combine_bwd(c.pre_x, C * M, M, C, M, c.pre_ss, c.pre_out, d_pre_out, d_x, 0);
        L.act_grad(d_x, c.pre_x, d_x, Rf * 32);
        w.bias_grad(d_x, GRD(c, pre_lin_b), Rf, 32);
        float *dz0 = w.Z(Rb * 32), *dz1 = w.Z(Nb * 32), *dz3 = w.Z((long long)NC * 32), *dz4 = w.Z((long long)N * 32);
        w.reduce_to(d_x, d, 32, M_N | M_C | M_B | M_T, dz0);
        w.reduce_to(dz0, Dims5{N, C, B, T, 1}, 32, M_N | M_B, dz1);
        w.reduce_to(dz0, Dims5{N, C, B, T, 1}, 32, M_N | M_C, dz3);
        w.reduce_to(dz3, Dims5{N, C, 1, 1, 1}, 32, M_N, dz4);
        w.branch_bwd(dz0, (int)Rb, 32, pre_lin_w, 0, 135, c.pre_a1, 48, c.pce_beats, 64, 64, pre_bl_w, pre_bl_b, d_pce_beats, 64, 1);
        w.branch_bwd(dz1, (int)Nb, 32, pre_lin_w, 48, 135, c.pre_a2, 40, c.pce_bars, 128, 128, pre_br_w, pre_br_b, d_pce_bars, 128, 1);
        {   // channels branch: the weight gradient of channels_linear runs over the occupied cells only
            float *d_a = w.A(Rf * 16);
            w.dw(d_x, 32, c.pre_a3, 16, GRD(c, pre_lin_w) + 88, 135, (int)Rf, 32, 16);
            L.linear_dx(d_x, 32, PRM(c, pre_lin_w) + 88, 135, d_a, 16, (int)Rf, 32, 16, 0);
            L.act_grad(d_a, c.pre_a3, d_a, Rf * 16);
            w.bias_grad(d_a, GRD(c, pre_ch_b), Rf, 16);
            if (!L.dry) {
                k_chlin_sparse_wgrad<<<(int)std::min<long long>((Rb + 7) / 8, 148 * 8), 256, 0, L.st>>>(c.cells_off, c.cells_rec, d_a, GRD(c, pre_ch_w), Rb);
                L.check();
            }
        }
        w.branch_bwd(dz3, NC, 32, pre_lin_w, 104, 135, c.pre_a4, 21, c.F, 51, 51, pre_il_w, pre_il_b, nullptr, 0, 0);
        w.branch_bwd(dz4, N, 32, pre_lin_w, 125, 135, c.pre_a5, 5, c.mode, 2, 2, pre_ml_w, pre_ml_b, nullptr, 0, 0);
        w.branch_bwd(dz4, N, 32, pre_lin_w, 130, 135, c.pre_a6, 5, c.bpm, 1, 1, pre_bp_w, pre_bp_b, nullptr, 0, 0);
    }
    // ---------------- join branch 0: the StyleEncoder's gradient w.r.t. the bars reaches the channel encoders (model.py:766)
    L.wait(-1, 0);
    if (U) L.wait(-1, 1);                            // the unpitched rhythm encoder's backward
    if (U) {
        if (!L.dry) { k_pair_unpack_add<<<grid_for(Nb * 128), 256, 0, L.st>>>(d_bp, d_pce_bars, d_uce_bars, N, (long long)B * 128); L.check(); }
    } else {
        L.axpy(d_pce_bars, d_bars, Nb * 128);
    }
    // ---------------- bar-axis BiLSTMs of both channel encoders: their four recurrences are independent -> one launch
    {
        LstmArgs jobs[4];
        int nj = 0;
        jobs[nj++] = w.lstm_bwd_args(c.pce_brf, d_pce_bars, 128, 0, 0, pce_br_whh);
        jobs[nj++] = w.lstm_bwd_args(c.pce_brr, d_pce_bars, 128, 64, 1, pce_br_whh_r);
        if (U) {
            jobs[nj++] = w.lstm_bwd_args(c.uce_brf, d_uce_bars, 128, 0, 0, uce_br_whh);
            jobs[nj++] = w.lstm_bwd_args(c.uce_brr, d_uce_bars, 128, 64, 1, uce_br_whh_r);
        }
        L.lstm_multi(jobs, nj, true);
    }
    if (U) {
        // ---------------- UnpitchedChannelsEncoder (model.py:128-141), on branch 0 beside the pitched encoder's backward
        L.wait(0, -1); L.on(0);
        {
            float *d_in = w.A(Nb * 64), *d_last = w.A(Nb * 64), *d_lin = w.A(Nb * T * 64), *gWp = w.Z(64 * 940);
            w.lstm_bwd_grads(c.uce_brf, c.uce_bars_in, uce_br_wih, uce_br_whh, uce_br_bih, uce_br_bhh, d_in, 0);
            w.lstm_bwd_grads(c.uce_brr, c.uce_bars_in, uce_br_wih_r, uce_br_whh_r, uce_br_bih_r, uce_br_bhh_r, d_in, 1);
            w.combine_bwd(c.uce_last, (long long)B * 64, (long long)B * 64, 1, (long long)B * 64, c.uce_ss, c.uce_bars_in, d_in, d_last, 0);
            L.put_last(d_last, d_uce_beats, Nb, T, 64);
            w.lstm_bwd(c.uce_bl, d_uce_beats, 64, 0, 0, c.uce_lin, uce_bl_wih, uce_bl_whh, uce_bl_bih, uce_bl_bhh, d_lin, 0);
            L.act_grad(d_lin, c.uce_lin, d_lin, Nb * T * 64);
            w.bias_grad(d_lin, GRD(c, uce_lin_b), Nb * T, 64);
            L.linear_dw(d_lin, 64, c.Ug, 940, gWp, 940, (int)(Nb * T), 64, 940);
            if (!L.dry) { k_permute_cols<<<grid_for(64 * 940), 256, 0, L.st>>>(gWp, GRD(c, uce_lin_w), 64, 940, 2, 1, 1); L.check(); }
        }
        L.on(-1);
    }
    // ---------------- PitchedChannelsEncoder (model.py:77-99)
    {
        float *d_in = w.A(Nb * 64), *d_last = w.A((long long)NC * B * 64), *d_x = w.A(Rb * 64);
        w.lstm_bwd_grads(c.pce_brf, c.pce_bars_in, pce_br_wih, pce_br_whh, pce_br_bih, pce_br_bhh, d_in, 0);
        w.lstm_bwd_grads(c.pce_brr, c.pce_bars_in, pce_br_wih_r, pce_br_whh_r, pce_br_bih_r, pce_br_bhh_r, d_in, 1);
        w.combine_bwd(c.pce_last, (long long)C * B * 64, (long long)B * 64, C, (long long)B * 64, c.pce_ss, c.pce_bars_in, d_in, d_last, 0);
        L.put_last(d_last, d_pce_beats, (long long)NC * B, T, 64);
        w.lstm_bwd(c.pce_bl, d_pce_beats, 64, 0, 0, c.pce_x, pce_bl_wih, pce_bl_whh, pce_bl_bih, pce_bl_bhh, d_x, 0);
        L.act_grad(d_x, c.pce_x, d_x, Rb * 64);
        w.bias_grad(d_x, GRD(c, pce_lin_b), Rb, 64);
        // instruments branch (columns 456.. of linear.weight are not permuted)
        float *dzi = w.Z((long long)NC * 64);
        w.reduce_to(d_x, Dims5{N, C, B, T, 1}, 64, M_N | M_C, dzi);
        w.leaf_dw = false;            // k_permute_cols below adds into every column of pce_lin_w's gradient on this stream
        w.branch_bwd(dzi, NC, 64, pce_lin_w, 456, 514, c.pce_il, 58, c.F, 51, 51, pce_il_w, pce_il_b, nullptr, 0, 0);
        w.leaf_dw = true;
        // conv branch: z = conv_out (rows, 456) . Wp[:, :456]^T in the permuted column order
        float *gWp = w.Z(64 * 514), *d_conv = w.A(Rb * 456), *gT = w.Z(700 * 57);
        L.linear_dw(d_x, 64, c.pce_conv, 456, gWp, 514, (int)Rb, 64, 456);
        if (!L.dry) { k_permute_cols<<<grid_for(64 * 514), 256, 0, L.st>>>(gWp, GRD(c, pce_lin_w), 64, 514, 1, 1, 1); L.check(); }
        L.gemm_cls = 32;
        L.linear_dx(d_x, 64, c.wpcelin_p, 514, d_conv, 456, (int)Rb, 64, 456, 0);
        L.gemm_cls = 1;
        L.act_grad(d_conv, c.pce_conv, d_conv, Rb * 456);
        w.bias_grad(d_conv, GRD(c, pce_conv_b), Rb * 8, 57);
        // gT[k][o] = sum over occupied cells of p[feat] * d_conv[row][oct][o]  (k = f * 70 + tap * 5 + feat)
        if (!L.dry) {
            k_conv_sparse_wgrad<<<(int)std::min<long long>((Rb + 7) / 8, 148 * 8), 256, 0, L.st>>>(c.cells_off, c.cells_rec, d_conv, gT, Rb); L.check();
            k_conv_wgrad_unpack<<<grid_for(57 * 700), 256, 0, L.st>>>(gT, GRD(c, pce_conv_w)); L.check();
        }
    }
    if (U) L.wait(-1, 0);
}

static void backward(Launcher &L, Ctx &c, Arena &ar, const TrainIn &ti) {
    // sizing pass: what the pass carves from the work region and from the zero region, which follows it
    Launcher P; P.st = nullptr; P.dry = true; P.bf16 = L.bf16;
    Arena a1{nullptr, ar.off, true}, z1{nullptr, 0, true};
    TrainIn none{};
    backward_impl(P, c, a1, z1, none);
    const size_t z_base = mst_align_up(a1.off, 256);
    Arena zar{ar.base + z_base, 0, ar.dry};
    if (!L.dry) L.fill((float *)(ar.base + z_base), (long long)(z1.off / sizeof(float)), 0.f);
    backward_impl(L, c, ar, zar, ti);
    ar.off = z_base + z1.off;
}

static size_t plan_backward(Ctx &c, Arena &ar) {
    Launcher L; L.st = nullptr; L.dry = true;
    TrainIn ti{};
    backward(L, c, ar, ti);
    return ar.off;
}
