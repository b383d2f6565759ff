// mst_model.cu — training step of the style-transfer network (style/model.py) on sm_100a: host-side runtime.
//
// One flat fp32 parameter buffer (980 325 values, table below) + one flat gradient buffer; activations live in
// a caller-provided arena.  A batch is N *independent* batch-1 samples of identical shape (C, B, T, drums):
// every reduction of the reference that spans the batch dimension (combine, the loss sums) is segmented per
// sample, and gradients are summed over samples — exactly train-model.py's iter_size accumulation.
#include <math.h>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "mst_model_sparse.cuh"
#include "mst_model_gemm_tc.cuh"
#include "mst_model_gemm_t5.cuh"
#include "mst_model_gemm_b16.cuh"
#include "mst_model_psa_tc.cuh"
#include "mst_model_gemm_tma.cuh"
#include "mst_model_lstm_cluster.cuh"
#include "mst_model_tails_bwd.cuh"
#include "mst_model_me.cuh"

using namespace mstm;

// ------------------------------------------------------------------------------------------------ parameters
// X(id, state-dict name, numel).  Names are the reference's (style/model.py); Python maps by name.
// octave_linear / scale_degree_linear weights (and biases) are adjacent on purpose: they act as one matrix.
#define MST_PARAMS(X) \
    X(pce_conv_w, "pitched_channels_encoder.beats_conv.module.weight", 57 * 50 * 14) \
    X(pce_conv_b, "pitched_channels_encoder.beats_conv.module.bias", 57) \
    X(pce_il_w, "pitched_channels_encoder.instruments_linear.weight", 58 * 51) \
    X(pce_il_b, "pitched_channels_encoder.instruments_linear.bias", 58) \
    X(pce_lin_w, "pitched_channels_encoder.linear.weight", 64 * 514) \
    X(pce_lin_b, "pitched_channels_encoder.linear.bias", 64) \
    X(pce_bl_wih, "pitched_channels_encoder.beats_lstm.module.weight_ih_l0", 256 * 64) \
    X(pce_bl_whh, "pitched_channels_encoder.beats_lstm.module.weight_hh_l0", 256 * 64) \
    X(pce_bl_bih, "pitched_channels_encoder.beats_lstm.module.bias_ih_l0", 256) \
    X(pce_bl_bhh, "pitched_channels_encoder.beats_lstm.module.bias_hh_l0", 256) \
    X(pce_br_wih, "pitched_channels_encoder.bars_lstm.weight_ih_l0", 256 * 64) \
    X(pce_br_whh, "pitched_channels_encoder.bars_lstm.weight_hh_l0", 256 * 64) \
    X(pce_br_bih, "pitched_channels_encoder.bars_lstm.bias_ih_l0", 256) \
    X(pce_br_bhh, "pitched_channels_encoder.bars_lstm.bias_hh_l0", 256) \
    X(pce_br_wih_r, "pitched_channels_encoder.bars_lstm.weight_ih_l0_reverse", 256 * 64) \
    X(pce_br_whh_r, "pitched_channels_encoder.bars_lstm.weight_hh_l0_reverse", 256 * 64) \
    X(pce_br_bih_r, "pitched_channels_encoder.bars_lstm.bias_ih_l0_reverse", 256) \
    X(pce_br_bhh_r, "pitched_channels_encoder.bars_lstm.bias_hh_l0_reverse", 256) \
    X(uce_lin_w, "unpitched_channels_encoder.linear.weight", 64 * 940) \
    X(uce_lin_b, "unpitched_channels_encoder.linear.bias", 64) \
    X(uce_bl_wih, "unpitched_channels_encoder.beats_lstm.module.weight_ih_l0", 256 * 64) \
    X(uce_bl_whh, "unpitched_channels_encoder.beats_lstm.module.weight_hh_l0", 256 * 64) \
    X(uce_bl_bih, "unpitched_channels_encoder.beats_lstm.module.bias_ih_l0", 256) \
    X(uce_bl_bhh, "unpitched_channels_encoder.beats_lstm.module.bias_hh_l0", 256) \
    X(uce_br_wih, "unpitched_channels_encoder.bars_lstm.weight_ih_l0", 256 * 64) \
    X(uce_br_whh, "unpitched_channels_encoder.bars_lstm.weight_hh_l0", 256 * 64) \
    X(uce_br_bih, "unpitched_channels_encoder.bars_lstm.bias_ih_l0", 256) \
    X(uce_br_bhh, "unpitched_channels_encoder.bars_lstm.bias_hh_l0", 256) \
    X(uce_br_wih_r, "unpitched_channels_encoder.bars_lstm.weight_ih_l0_reverse", 256 * 64) \
    X(uce_br_whh_r, "unpitched_channels_encoder.bars_lstm.weight_hh_l0_reverse", 256 * 64) \
    X(uce_br_bih_r, "unpitched_channels_encoder.bars_lstm.bias_ih_l0_reverse", 256) \
    X(uce_br_bhh_r, "unpitched_channels_encoder.bars_lstm.bias_hh_l0_reverse", 256) \
    X(se_lstm_wih, "style_encoder.bars_lstm.weight_ih_l0", 768 * 128) \
    X(se_lstm_whh, "style_encoder.bars_lstm.weight_hh_l0", 768 * 192) \
    X(se_lstm_bih, "style_encoder.bars_lstm.bias_ih_l0", 768) \
    X(se_lstm_bhh, "style_encoder.bars_lstm.bias_hh_l0", 768) \
    X(se_il_w, "style_encoder.instruments_linear.weight", 39 * 51) \
    X(se_il_b, "style_encoder.instruments_linear.bias", 39) \
    X(se_ml_w, "style_encoder.mode_linear.weight", 13 * 2) \
    X(se_ml_b, "style_encoder.mode_linear.bias", 13) \
    X(se_bp_w, "style_encoder.bpm_linear.weight", 7) \
    X(se_bp_b, "style_encoder.bpm_linear.bias", 7) \
    X(se_lin_w, "style_encoder.linear.weight", 256 * 251) \
    X(se_lin_b, "style_encoder.linear.bias", 256) \
    X(me_bl_w, "melody_encoder.beats_linear.weight", 36 * 64) \
    X(me_bl_b, "melody_encoder.beats_linear.bias", 36) \
    X(me_br_w, "melody_encoder.bars_linear.weight", 68 * 128) \
    X(me_br_b, "melody_encoder.bars_linear.bias", 68) \
    X(me_il_w, "melody_encoder.instruments_linear.weight", 8 * 51) \
    X(me_il_b, "melody_encoder.instruments_linear.bias", 8) \
    X(me_oct_w, "melody_encoder.octave_linear.weight", 64 * 112) \
    X(me_deg_w, "melody_encoder.scale_degree_linear.weight", 56 * 112) \
    X(me_oct_b, "melody_encoder.octave_linear.bias", 64) \
    X(me_deg_b, "melody_encoder.scale_degree_linear.bias", 56) \
    X(me_ch_w, "melody_encoder.channels_linear.weight", 7 * 5) \
    X(me_ch_b, "melody_encoder.channels_linear.bias", 7) \
    X(me_lin_w, "melody_encoder.linear.weight", 8 * 15) \
    X(me_lin_b, "melody_encoder.linear.bias", 8) \
    X(pre_bl_w, "pitched_rhythm_encoder.beats_linear.weight", 48 * 64) \
    X(pre_bl_b, "pitched_rhythm_encoder.beats_linear.bias", 48) \
    X(pre_br_w, "pitched_rhythm_encoder.bars_linear.weight", 40 * 128) \
    X(pre_br_b, "pitched_rhythm_encoder.bars_linear.bias", 40) \
    X(pre_ch_w, "pitched_rhythm_encoder.channels_linear.weight", 16 * 280) \
    X(pre_ch_b, "pitched_rhythm_encoder.channels_linear.bias", 16) \
    X(pre_il_w, "pitched_rhythm_encoder.instruments_linear.weight", 21 * 51) \
    X(pre_il_b, "pitched_rhythm_encoder.instruments_linear.bias", 21) \
    X(pre_ml_w, "pitched_rhythm_encoder.mode_linear.weight", 5 * 2) \
    X(pre_ml_b, "pitched_rhythm_encoder.mode_linear.bias", 5) \
    X(pre_bp_w, "pitched_rhythm_encoder.bpm_linear.weight", 5) \
    X(pre_bp_b, "pitched_rhythm_encoder.bpm_linear.bias", 5) \
    X(pre_lin_w, "pitched_rhythm_encoder.linear.weight", 32 * 135) \
    X(pre_lin_b, "pitched_rhythm_encoder.linear.bias", 32) \
    X(ure_bl_w, "unpitched_rhythm_encoder.beats_linear.weight", 48 * 64) \
    X(ure_bl_b, "unpitched_rhythm_encoder.beats_linear.bias", 48) \
    X(ure_br_w, "unpitched_rhythm_encoder.bars_linear.weight", 40 * 128) \
    X(ure_br_b, "unpitched_rhythm_encoder.bars_linear.bias", 40) \
    X(ure_ch_w, "unpitched_rhythm_encoder.channels_linear.weight", 16 * 94) \
    X(ure_ch_b, "unpitched_rhythm_encoder.channels_linear.bias", 16) \
    X(ure_bp_w, "unpitched_rhythm_encoder.bpm_linear.weight", 5) \
    X(ure_bp_b, "unpitched_rhythm_encoder.bpm_linear.bias", 5) \
    X(ure_lin_w, "unpitched_rhythm_encoder.linear.weight", 32 * 109) \
    X(ure_lin_b, "unpitched_rhythm_encoder.linear.bias", 32) \
    X(sim_bl_wih, "song_info_model.beats_lstm.module.weight_ih_l0", 36 * 320) \
    X(sim_bl_whh, "song_info_model.beats_lstm.module.weight_hh_l0", 36 * 9) \
    X(sim_bl_bih, "song_info_model.beats_lstm.module.bias_ih_l0", 36) \
    X(sim_bl_bhh, "song_info_model.beats_lstm.module.bias_hh_l0", 36) \
    X(sim_br_wih, "song_info_model.bars_lstm.weight_ih_l0", 32 * 9) \
    X(sim_br_whh, "song_info_model.bars_lstm.weight_hh_l0", 32 * 8) \
    X(sim_br_bih, "song_info_model.bars_lstm.bias_ih_l0", 32) \
    X(sim_br_bhh, "song_info_model.bars_lstm.bias_hh_l0", 32) \
    X(sim_si_w, "song_info_model.style_instruments_linear.weight", 8 * 256) \
    X(sim_si_b, "song_info_model.style_instruments_linear.bias", 8) \
    X(sim_ri_w, "song_info_model.rhythm_instruments_linear.weight", 10 * 8) \
    X(sim_ri_b, "song_info_model.rhythm_instruments_linear.bias", 10) \
    X(sim_i_w, "song_info_model.instruments_linear.weight", 41 * 18) \
    X(sim_i_b, "song_info_model.instruments_linear.bias", 41) \
    X(sim_sm_w, "song_info_model.style_mode_linear.weight", 2 * 256) \
    X(sim_sm_b, "song_info_model.style_mode_linear.bias", 2) \
    X(sim_rm_w, "song_info_model.rhythm_mode_linear.weight", 2 * 8) \
    X(sim_rm_b, "song_info_model.rhythm_mode_linear.bias", 2) \
    X(sim_m_w, "song_info_model.mode_linear.weight", 2 * 4) \
    X(sim_m_b, "song_info_model.mode_linear.bias", 2) \
    X(sim_sb_w, "song_info_model.style_bpm_linear.weight", 2 * 256) \
    X(sim_sb_b, "song_info_model.style_bpm_linear.bias", 2) \
    X(sim_rb_w, "song_info_model.rhythm_bpm_linear.weight", 2 * 8) \
    X(sim_rb_b, "song_info_model.rhythm_bpm_linear.bias", 2) \
    X(sim_b_w, "song_info_model.bpm_linear.weight", 4) \
    X(sim_b_b, "song_info_model.bpm_linear.bias", 1) \
    X(psa_sl_w, "pitched_style_applier.style_linear.weight", 66 * 256) \
    X(psa_sl_b, "pitched_style_applier.style_linear.bias", 66) \
    X(psa_rl_w, "pitched_style_applier.rhythm_linear.weight", 10 * 32) \
    X(psa_rl_b, "pitched_style_applier.rhythm_linear.bias", 10) \
    X(psa_il_w, "pitched_style_applier.instruments_linear.weight", 12 * 51) \
    X(psa_il_b, "pitched_style_applier.instruments_linear.bias", 12) \
    X(psa_oct_w, "pitched_style_applier.octave_linear.weight", 240 * 88) \
    X(psa_deg_w, "pitched_style_applier.scale_degree_linear.weight", 210 * 88) \
    X(psa_oct_b, "pitched_style_applier.octave_linear.bias", 240) \
    X(psa_deg_b, "pitched_style_applier.scale_degree_linear.bias", 210) \
    X(psa_ml_w, "pitched_style_applier.melody_linear.weight", 20 * 8) \
    X(psa_ml_b, "pitched_style_applier.melody_linear.bias", 20) \
    X(psa_lin_w, "pitched_style_applier.linear.weight", 5 * 50) \
    X(psa_lin_b, "pitched_style_applier.linear.bias", 5) \
    X(usa_sl_w, "unpitched_style_applier.style_linear.weight", 650 * 256) \
    X(usa_sl_b, "unpitched_style_applier.style_linear.bias", 650) \
    X(usa_rl_w, "unpitched_style_applier.rhythm_linear.weight", 17 * 32) \
    X(usa_rl_b, "unpitched_style_applier.rhythm_linear.bias", 17) \
    X(usa_nl_w, "unpitched_style_applier.notes_linear.weight", 376 * 82) \
    X(usa_nl_b, "unpitched_style_applier.notes_linear.bias", 376) \
    X(usa_lin_w, "unpitched_style_applier.linear.weight", 2 * 8) \
    X(usa_lin_b, "unpitched_style_applier.linear.bias", 2)

enum ParamId {
#define X(id, name, n) id,
    MST_PARAMS(X)
#undef X
    kNumParams
};
static const char *const kParamNames[kNumParams] = {
#define X(id, name, n) name,
    MST_PARAMS(X)
#undef X
};
static const int kParamNumel[kNumParams] = {
#define X(id, name, n) n,
    MST_PARAMS(X)
#undef X
};
static long long g_param_off[kNumParams + 1];
static bool g_param_off_ready = false;
static void init_param_offsets() {
    if (g_param_off_ready) return;
    long long off = 0;
    for (int i = 0; i < kNumParams; i++) { g_param_off[i] = off; off += kParamNumel[i]; }
    g_param_off[kNumParams] = off;
    g_param_off_ready = true;
}

extern "C" int mst_model_param_count(void) { return kNumParams; }
extern "C" const char *mst_model_param_name(int i) { return (i >= 0 && i < kNumParams) ? kParamNames[i] : nullptr; }
extern "C" long long mst_model_param_numel(int i) { return (i >= 0 && i < kNumParams) ? kParamNumel[i] : -1; }
extern "C" long long mst_model_param_offset(int i) { init_param_offsets(); return (i >= 0 && i <= kNumParams) ? g_param_off[i] : -1; }

// ------------------------------------------------------------------------------------------------ small glue kernels
namespace {

// out[s, k] = in[((s * T) + T - 1) * H + k]  (x[:, -1] of an LSTM output); backward adds into the same place
__global__ void __launch_bounds__(256) k_take_last(const float *__restrict__ in, float *__restrict__ out, long long S, int T, int H) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < S * H; i += (long long)gridDim.x * blockDim.x) {
        const long long s = i / H; const int k = (int)(i - s * H);
        out[i] = in[(s * T + T - 1) * H + k];
    }
}
__global__ void __launch_bounds__(256) k_put_last(const float *__restrict__ src, float *__restrict__ dst, long long S, int T, int H) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < S * H; i += (long long)gridDim.x * blockDim.x) {
        const long long s = i / H; const int k = (int)(i - s * H);
        dst[(s * T + T - 1) * H + k] += src[i];
    }
}
__global__ void __launch_bounds__(256) k_add2(const float *a, const float *b, float *out, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = a[i] + b[i];
}
// pair buffer [N, 2, M] <- a[N, M], b[N, M]   (and back, adding into da / db)
__global__ void __launch_bounds__(256) k_pair_pack(const float *a, const float *b, float *pair, int N, long long M) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)N * M; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / M, m = i - n * M;
        pair[(n * 2) * M + m] = a[i];
        pair[(n * 2 + 1) * M + m] = b[i];
    }
}
__global__ void __launch_bounds__(256) k_pair_unpack_add(const float *dpair, float *da, float *db, int N, long long M) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)N * M; i += (long long)gridDim.x * blockDim.x) {
        const long long n = i / M, m = i - n * M;
        da[i] += dpair[(n * 2) * M + m];
        db[i] += dpair[(n * 2 + 1) * M + m];
    }
}
// bpm head (model.py:552-554): out = sigmoid(x) * 150 + 50
__global__ void k_bpm_head(const float *x, float *out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = sigmoidf(x[i]) * 150.f + 50.f;
}

struct Arena {
    char *base; size_t off; bool dry;
    float *take(long long n_floats) {
        const size_t o = off;
        off = mst_align_up(off + (size_t)(n_floats > 0 ? n_floats : 1) * sizeof(float), 256);
        return dry ? nullptr : (float *)(base + o);
    }
};

inline int grid_for(long long n, int threads = 256, int cap = 148 * 16) {
    long long b = (n + threads - 1) / threads;
    return (int)std::max<long long>(1, std::min<long long>(b, cap));
}

// Side streams of one caller stream.  Independent branches of the model (the two channel encoders, the rhythm / melody / style
// encoders, the two style appliers and their backward passes) run on them concurrently: the sequential recurrences keep a
// handful of SMs busy for hundreds of microseconds, and at N = 1 every kernel is that small.  Branches are forked from and
// joined back to the caller's stream with events, so at return all work is ordered on the caller's stream as before, and a
// CUDA graph capture of the call records the branches as parallel graph nodes.  One set per (device, caller stream): two
// callers on different streams never share side streams (a capture on one must not swallow the other's launches).
// MST_MODEL_STREAMS=0 keeps everything on the caller's stream.
constexpr int kSideStreams = 4, kSideEvents = 32;
constexpr int kBranchSim = 3;        // backward of the SongInfoModel recurrences (beside the pitched applier and the StyleEncoder)
constexpr int kBranchLeaf = 2;       // weight-gradient GEMMs and bias column sums of the backward pass: nothing downstream reads them
struct StreamSet {
    cudaStream_t s[kSideStreams];
    cudaEvent_t ev[kSideEvents];
    int next_ev = 0;
};
static StreamSet *stream_set_for(cudaStream_t main_st, cudaError_t *err) {
    static std::mutex mu;
    static std::vector<std::pair<std::pair<int, cudaStream_t>, StreamSet *>> sets;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { *err = e; return nullptr; }
    std::lock_guard<std::mutex> lock(mu);
    for (auto &kv : sets) if (kv.first.first == dev && kv.first.second == main_st) return kv.second;
    StreamSet *ss = new StreamSet();
    for (int i = 0; i < kSideStreams && e == cudaSuccess; i++) e = cudaStreamCreateWithFlags(&ss->s[i], cudaStreamNonBlocking);
    for (int i = 0; i < kSideEvents && e == cudaSuccess; i++) e = cudaEventCreateWithFlags(&ss->ev[i], cudaEventDisableTiming);
    if (e != cudaSuccess) { *err = e; delete ss; return nullptr; }
    sets.push_back({{dev, main_st}, ss});
    return ss;
}

struct Launcher {
    cudaStream_t st;              // the stream launches go to: the caller's, or a side stream between on(k) and on(-1)
    cudaStream_t main_st = nullptr;
    StreamSet *ss = nullptr;      // null: no concurrency, every branch runs on the caller's stream in program order
    unsigned forked = 0;          // side streams that hold work not yet joined into the caller's stream
    int launches = 0;
    bool dry = false;             // planning pass: carve the arena, launch nothing
    bool bf16 = false;            // bf16 dense mode (MstModelShape.precision = MST_PRECISION_BF16)
    // development knobs: attribute the mode's error / time to its two halves
    // MST_BF16_MASK: which GEMM classes take the bf16 kernel (1 linear forward, 2 input gradient, 4 weight gradient,
    // 8 LSTM input projection, 16 LSTM backward products, 32 the input gradient towards the conv features: the conv weight
    // gradient behind it is a near-cancelling sum over a few thousand occupied cells (|g| 1e-5 of the step's largest), where
    // bf16 rounding of dY shows as 2-8 % of its norm; that GEMM is bound by its 456-wide output write in either mode).  Default 23: everything but the LSTM input projections - their
    // inputs carry the raw bpm / un-normalised features, and a 2^-9 rounding of those moves the gates of a 200-800 step
    // recurrence enough to push gradients 8 % off the fp32 oracle (tools/bf16_diag.py), against <= 1 % for all other classes
    int gemm_cls = 1;
    bool bf16_gemm() const {
        static const bool off = getenv("MST_BF16_NO_GEMM") != nullptr;
        static const int mask = getenv("MST_BF16_MASK") ? atoi(getenv("MST_BF16_MASK")) : 23;
        return bf16 && !off && (mask & gemm_cls) != 0;
    }
    // The tcgen05 formulations of the PitchedStyleApplier tail (mst_model_psa_tc.cuh) are correct (tests/test_gpu_bf16.py) but
    // measured SLOWER than the FP32 SIMT kernels (profiles/r2_psa_tc_*.txt: building the A tile costs more instructions per
    // row than the 75 packed FMAs it removes), so they run only on request: MST_PSA_FWD_TC=1 / MST_PSA_BWD_TC=1 (read per call).
    bool psa_fwd_tc() const { return bf16 && getenv("MST_PSA_FWD_TC") != nullptr; }
    bool psa_bwd_tc() const { return bf16 && getenv("MST_PSA_BWD_TC") != nullptr; }
    cudaError_t err = cudaSuccess;
    // ---- branches.  begin(stream): call once per entry point.  wait(k, on): branch k continues only after everything queued
    // so far on branch `on` (-1 = the caller's stream).  on(k): the launches that follow go to branch k.  join_all(): the
    // caller's stream waits for every branch (every entry point ends with it; a capture would fail on unjoined work).
    void begin(void *stream) {
        st = main_st = (cudaStream_t)stream;
        static const bool off = getenv("MST_MODEL_STREAMS") && atoi(getenv("MST_MODEL_STREAMS")) == 0;
        if (!dry && !off) ss = stream_set_for(main_st, &err);
    }
    cudaStream_t stream_of(int k) const { return k < 0 ? main_st : ss->s[k]; }
    void wait(int k, int on_k) {
        if (!ss || dry || k == on_k) return;
        if (on_k >= 0 && !(forked & (1u << on_k))) return;      // nothing pending there (and, under capture, not part of it)
        cudaEvent_t ev = ss->ev[ss->next_ev];
        ss->next_ev = (ss->next_ev + 1) % kSideEvents;
        cudaError_t e = cudaEventRecord(ev, stream_of(on_k));
        if (e == cudaSuccess) e = cudaStreamWaitEvent(stream_of(k), ev, 0);
        if (e != cudaSuccess && err == cudaSuccess) err = e;
        if (k >= 0) forked |= 1u << k;
        if (k < 0 && on_k >= 0) forked &= ~(1u << on_k);
    }
    int cur = -1;                 // the branch launches currently go to
    void on(int k) { if (ss && !dry) { st = stream_of(k); cur = k; if (k >= 0) forked |= 1u << k; } }
    // run f's launches on the leaf branch, after what is queued on the current one (the gradient they read is ready there)
    template <class F> void leaf(F f) {
        if (!ss || dry) { f(); return; }
        const int back = cur;
        wait(kBranchLeaf, back); on(kBranchLeaf);
        f();
        on(back);
    }
    void join_all() {
        if (!ss || dry) return;
        on(-1);
        for (int k = 0; k < kSideStreams; k++) if (forked & (1u << k)) wait(-1, k);
    }
    void check() { cudaError_t e = cudaGetLastError(); if (e != cudaSuccess && err == cudaSuccess) err = e; launches++; }

    // C[M,N] (op)= act(A·B + bias) with explicit strides
    // tensor-core path (mst_model_gemm_tc.cuh) for everything but the tiny problems
    static bool use_tc(int M, int N, int K) { return (long long)M * N * K >= (1ll << 21); }
    template <int BN>
    void gemm_tc(const GemmArgs &g, int split_k) {
        dim3 grid((g.N + BN - 1) / BN, (g.M + kTcBM - 1) / kTcBM, split_k);
        const bool ak = g.a_sk == 1, bk = g.b_sk == 1;
        if (ak && bk) k_gemm_tc<BN, true, true><<<grid, kTcThreads, 0, st>>>(g);
        else if (ak) k_gemm_tc<BN, true, false><<<grid, kTcThreads, 0, st>>>(g);
        else if (bk) k_gemm_tc<BN, false, true><<<grid, kTcThreads, 0, st>>>(g);
        else k_gemm_tc<BN, false, false><<<grid, kTcThreads, 0, st>>>(g);
    }
    // operand access mode of the tcgen05 GEMM (mst_model_gemm_t5.cuh): 0 = k-contiguous rows on 16-byte boundaries,
    // 1 = on 8-byte boundaries, 2 = any strides
    static int t5_mode(const float *p, long long s_row, long long s_k) {
        if (s_k != 1) return 2;
        if ((s_row & 3) == 0 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) return 0;
        if ((s_row & 1) == 0 && (reinterpret_cast<uintptr_t>(p) & 7) == 0) return 1;
        return 2;
    }
    template <int BN, int AM, int BMD, int NS>
    void gemm_t5_launch(const GemmArgs &g, dim3 grid) {
        static bool attr_set[64] = {};            // per device: the attribute belongs to the device's copy of the function
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            cudaFuncSetAttribute(k_gemm_t5<BN, AM, BMD, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t5_smem_bytes<BN, NS>());
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
        k_gemm_t5<BN, AM, BMD, NS><<<grid, kT5Block, t5_smem_bytes<BN, NS>(), st>>>(g);
    }
    template <int BN, int NS>
    void gemm_t5_ns(const GemmArgs &g, dim3 grid) {
        const int am = t5_mode(g.A, g.a_si, g.a_sk);
        const int bm = t5_mode(g.B, g.b_sj, g.b_sk) == 0 ? 0 : 2;          // the small operand: aligned or generic
        if (am == 0) { if (bm == 0) gemm_t5_launch<BN, 0, 0, NS>(g, grid); else gemm_t5_launch<BN, 0, 2, NS>(g, grid); }
        else if (am == 1) { if (bm == 0) gemm_t5_launch<BN, 1, 0, NS>(g, grid); else gemm_t5_launch<BN, 1, 2, NS>(g, grid); }
        else { if (bm == 0) gemm_t5_launch<BN, 2, 0, NS>(g, grid); else gemm_t5_launch<BN, 2, 2, NS>(g, grid); }
    }
    // bf16 dense mode (mst_model_gemm_b16.cuh): same tiling, bf16 operands, one MMA per k-step
    template <int BN, int AM, int BMD>
    void gemm_b16_launch(const GemmArgs &g, dim3 grid) {
        static bool attr_set[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            cudaFuncSetAttribute(k_gemm_b16<BN, AM, BMD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b16_smem_bytes<BN>());
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
        k_gemm_b16<BN, AM, BMD><<<grid, kT5Block, b16_smem_bytes<BN>(), st>>>(g);
    }
    template <int BN>
    void gemm_b16(const GemmArgs &g, int split_k) {
        dim3 grid((g.N + BN - 1) / BN, (g.M + kT5BM - 1) / kT5BM, split_k);
        const int am = t5_mode(g.A, g.a_si, g.a_sk);
        const int bm = t5_mode(g.B, g.b_sj, g.b_sk) == 0 ? 0 : 2;
        if (am == 0) { if (bm == 0) gemm_b16_launch<BN, 0, 0>(g, grid); else gemm_b16_launch<BN, 0, 2>(g, grid); }
        else if (am == 1) { if (bm == 0) gemm_b16_launch<BN, 1, 0>(g, grid); else gemm_b16_launch<BN, 1, 2>(g, grid); }
        else { if (bm == 0) gemm_b16_launch<BN, 2, 0>(g, grid); else gemm_b16_launch<BN, 2, 2>(g, grid); }
    }
    // tensor-map TMA path (bf16 dense mode; MST_BF16_NO_TMA=1 or a driver without cuTensorMapEncodeTiled falls back to k_gemm_b16)
    bool use_tma() const {
        static const bool off = getenv("MST_BF16_NO_TMA") != nullptr;
        return bf16_gemm_any() && !off && !dry && tensor_map_encoder() != nullptr;
    }
    bool bf16_gemm_any() const { static const bool off = getenv("MST_BF16_NO_GEMM") != nullptr; return bf16 && !off; }
    // C[M, N<=64] = act(A16[M, K] . B16[N, K]^T + bias): both operands bf16, K-contiguous, 16-byte aligned rows
    void gemm_tma64(const __nv_bfloat16 *A, long long lda, long long M, const __nv_bfloat16 *B, long long ldb, int N, int K, float *C,
                    long long ldc, const float *bias, int act, int accumulate) {
        if (dry || M <= 0) return;
        CUtensorMap ma, mb;
        if (!make_bf16_map(&ma, A, M, K, lda, kTmaBM) || !make_bf16_map(&mb, B, N, K, ldb, 64)) {
            if (err == cudaSuccess) err = cudaErrorInvalidValue;
            return;
        }
        static bool attr_set[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !attr_set[dev]) {
            cudaFuncSetAttribute(k_gemm_b16_tma<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tma_gemm_smem_bytes<64>());
            if (dev >= 0 && dev < 64) attr_set[dev] = true;
        }
        GemmTmaArgs g{C, bias, ldc, (int)M, N, K, act, accumulate};
        dim3 grid((N + 63) / 64, (unsigned)((M + kTmaBM - 1) / kTmaBM));
        k_gemm_b16_tma<64><<<grid, kTmaThreads, tma_gemm_smem_bytes<64>(), st>>>(ma, mb, g);
        check();
    }
    template <int BN>
    void gemm_t5(const GemmArgs &g, int split_k) {
        dim3 grid((g.N + BN - 1) / BN, (g.M + kT5BM - 1) / kT5BM, split_k);
        // two stages, two CTAs per SM.  (A 4-stage ring at one CTA per SM was measured 5-30 % SLOWER on every training
        // shape, tools/gemm_probe.py: the loaders are latency-bound, so resident loader warps matter more than ring depth.)
        gemm_t5_ns<BN, 2>(g, grid);
    }
    // C[M,N] (op)= act(A·B + bias) with explicit strides
    void gemm(const float *A, long long a_si, long long a_sk, const float *B, long long b_sk, long long b_sj, float *C, long long c_si,
              int M, int N, int K, const float *bias, int act, int accumulate, int split_k = 1, int kernel = 0) {
        static const int log_shapes = getenv("MST_GEMM_LOG") ? atoi(getenv("MST_GEMM_LOG")) : 0;   // development aid: list the step's GEMMs (1: planning pass, 2: launches)
        if ((log_shapes == 1 && dry) || (log_shapes == 2 && !dry && M > 0 && N > 0 && K > 0)) fprintf(stderr, "gemm M=%d N=%d K=%d a=(%lld,%lld) b=(%lld,%lld) c=%lld act=%d acc=%d split=%d\n", M, N, K, a_si, a_sk, b_sk, b_sj, c_si, act, accumulate, split_k);
        if (dry || M <= 0 || N <= 0 || K <= 0) return;
        GemmArgs g{A, B, bias, C, M, N, K, a_si, a_sk, b_sk, b_sj, c_si, act, accumulate};
        if (split_k > 1) g.accumulate = 2;
        if (kernel == 0 && K <= kSkMaxK && N <= 512 && N >= 128 && M >= 4096 && a_sk == 1 && split_k == 1 && accumulate != 2) {
            // short contraction, wide output, many rows: the row-streaming kernel (mst_model_gemm_tc.cuh)
            const int threads = (N + 31) / 32 * 32;
            const int blocks = std::min((M + kSkRows - 1) / kSkRows, 148 * (threads <= 512 ? 2 : 1));
            switch ((K + 3) / 4) {
                case 1: k_gemm_smallk<4><<<blocks, threads, 0, st>>>(g); break;
                case 2: k_gemm_smallk<8><<<blocks, threads, 0, st>>>(g); break;
                case 3: k_gemm_smallk<12><<<blocks, threads, 0, st>>>(g); break;
                case 4: k_gemm_smallk<16><<<blocks, threads, 0, st>>>(g); break;
                case 5: k_gemm_smallk<20><<<blocks, threads, 0, st>>>(g); break;
                default: k_gemm_smallk<24><<<blocks, threads, 0, st>>>(g); break;
            }
        } else if (kernel == 4 || (kernel == 0 && bf16_gemm() && use_tc(M, N, K) && K >= 24)) {
            if (N <= 16) gemm_b16<16>(g, split_k); else if (N <= 32) gemm_b16<32>(g, split_k); else gemm_b16<64>(g, split_k);
        } else if (kernel == 3 || (kernel == 0 && use_tc(M, N, K) && K >= 24)) {
            if (N <= 16) gemm_t5<16>(g, split_k); else if (N <= 32) gemm_t5<32>(g, split_k); else gemm_t5<64>(g, split_k);
        } else if (kernel == 2 || (kernel == 0 && use_tc(M, N, K))) {
            if (N <= 32) gemm_tc<32>(g, split_k); else gemm_tc<64>(g, split_k);
        } else if (kernel == 0 && ((N + kGemmTile - 1) / kGemmTile) * ((M + kGemmTile - 1) / kGemmTile) * split_k <= 16 &&
                   (long long)M * N <= 8192) {
            // a handful of 64 x 64 tiles would leave the machine idle: one warp per output element
            dim3 grid((unsigned)(((long long)M * N + kGemmWarpThreads / 32 - 1) / (kGemmWarpThreads / 32)), split_k);
            k_gemm_warp<<<grid, kGemmWarpThreads, 0, st>>>(g);
        } else {
            dim3 grid((N + kGemmTile - 1) / kGemmTile, (M + kGemmTile - 1) / kGemmTile, split_k);
            k_gemm<0><<<grid, kGemmThreads, 0, st>>>(g);
        }
        check();
    }
    // Y[M,N] = act(X[M,K]·W[N,K]^T + bias)      (nn.Linear); W row stride ldw (column block of a wider weight)
    void linear(const float *X, long long ldx, const float *W, long long ldw, const float *bias, float *Y, long long ldy,
                int M, int N, int K, int act, int accumulate = 0) {
        const int keep = gemm_cls;
        if (gemm_cls != 8) gemm_cls = 1;
        gemm(X, ldx, 1, W, 1, ldw, Y, ldy, M, N, K, bias, act, accumulate);
        gemm_cls = keep;
    }
    // dX[M,K] (+)= dY[M,N]·W[N,K]
    void linear_dx(const float *dY, long long ldy, const float *W, long long ldw, float *dX, long long ldx, int M, int N, int K, int accumulate) {
        const int keep = gemm_cls;
        if (gemm_cls != 16 && gemm_cls != 32) gemm_cls = 2;
        gemm(dY, ldy, 1, W, ldw, 1, dX, ldx, M, K, N, nullptr, 0, accumulate);
        gemm_cls = keep;
    }
    // dW[N,K] += dY[M,N]^T·X[M,K]   (atomic accumulate, K-dimension of the GEMM = M rows, split over CTAs)
    void linear_dw(const float *dY, long long ldy, const float *X, long long ldx, float *dW, long long ldw, int M, int N, int K) {
        // few output tiles, long reduction: split the rows over enough CTAs to fill the 148 SMs twice
        const bool tc = use_tc(N, K, M);
        const int tm = tc ? kTcBM : kGemmTile, tn = tc ? (K <= 16 ? 16 : (K <= 32 ? 32 : 64)) : kGemmTile;
        const int tiles = ((N + tm - 1) / tm) * ((K + tn - 1) / tn);
        const int split = std::max(1, std::min(std::min(512, M / 128), (2 * 148 + tiles - 1) / tiles));
        const int keep = gemm_cls;
        if (gemm_cls != 16) gemm_cls = 4;
        gemm(dY, 1, ldy, X, ldx, 1, dW, ldw, N, K, M, nullptr, 0, 2, split);
        gemm_cls = keep;
    }
    void colsum(const float *in, float *out, long long rows, int cols, float *out2 = nullptr) {
        if (dry || rows <= 0) return;
        dim3 grid((unsigned)std::max<long long>(1, std::min<long long>((rows + 63) / 64, 256)), (cols + 31) / 32);
        k_colsum<<<grid, 256, 0, st>>>(in, out, rows, cols, out2); check();
    }
    void act_grad_colsum(const float *dY, const float *Y, float *dPre, float *out, long long rows, int cols) {
        if (dry || rows <= 0) return;
        dim3 grid((unsigned)std::max<long long>(1, std::min<long long>((rows + 63) / 64, 148 * 4 / ((cols + 31) / 32) + 1)), (cols + 31) / 32);
        k_act_grad_colsum<<<grid, 256, 0, st>>>(dY, Y, dPre, out, rows, cols); check();
    }
    void fill(float *p, long long n, float v) { if (!dry && n > 0) { k_fill<<<grid_for(n), 256, 0, st>>>(p, n, v); check(); } }
    void axpy(float *y, const float *x, long long n) { if (!dry && n > 0) { k_axpy<<<grid_for(n), 256, 0, st>>>(y, x, n); check(); } }
    void act_grad(const float *dY, const float *Y, float *dPre, long long n) { if (dry || n <= 0) return; k_act_grad<<<grid_for(n), 256, 0, st>>>(dY, Y, dPre, n); check(); }
    void reduce_mid(const float *in, float *out, long long A, int L, long long inner, int accumulate) {
        if (dry) return;
        k_reduce_mid<<<grid_for(A * inner), 256, 0, st>>>(in, out, A, L, inner, accumulate); check();
    }
    void bcast(BcastArgs &a) {
        if (dry) return;
        const long long prefixes = (long long)a.d.n * a.d.c * a.d.b * a.d.t;
        bool vec = a.cols % 4 == 0 && (reinterpret_cast<uintptr_t>(a.Y) & 15) == 0;
        for (int s = 0; s < a.n_src; s++) vec = vec && (reinterpret_cast<uintptr_t>(a.src[s]) & 15) == 0;
        const int cv = vec ? a.cols / 4 : a.cols;
        int lpr = 32;
        while (lpr > 1 && lpr / 2 >= cv) lpr /= 2;
        if (vec) k_bcast_add<4><<<grid_for(prefixes * lpr), 256, 0, st>>>(a, lpr);
        else k_bcast_add<1><<<grid_for(prefixes * lpr), 256, 0, st>>>(a, lpr);
        check();
    }
    void take_last(const float *in, float *out, long long S, int T, int H) { if (dry) return; k_take_last<<<grid_for(S * H), 256, 0, st>>>(in, out, S, T, H); check(); }
    void put_last(const float *src, float *dst, long long S, int T, int H) { if (dry) return; k_put_last<<<grid_for(S * H), 256, 0, st>>>(src, dst, S, T, H); check(); }

    static size_t lstm_smem(int H, int G, bool w_smem) { return sizeof(float) * ((size_t)2 * G * H + (size_t)G * 4 * H + (size_t)4 * G * H + (w_smem ? (size_t)H * 4 * H : 0)); }
    // up to kLstmMaxJobs independent recurrences of equal H in one launch
    void lstm_multi(LstmArgs *jobs, int n_jobs, bool backward) {
        if (dry || n_jobs <= 0) return;
        LstmJobs pack{};
        int max_blocks = 0, G = kLstmG;
        for (int i = 0; i < n_jobs; i++) if (jobs[i].S < 148 * 2) G = 1;
        const int H = jobs[0].H;
        const int w_smem = ((size_t)H * 4 * H * sizeof(float) <= 96 * 1024) ? 1 : 0;
        // wide state, few sequences: 4-CTA cluster with W_hh resident in (distributed) shared memory
        if (!w_smem && n_jobs == 1 && H % kClSize == 0 && H * kClFwdParts <= 1024 && jobs[0].S <= 148 &&
            std::max(lstm_cl_fwd_smem(H), lstm_cl_bwd_smem(H)) <= 227 * 1024) {
            const int grid = (jobs[0].S + kClG - 1) / kClG * kClSize;
            const bool regs = H == 192;       // StyleEncoder.bars_lstm: 48 / 96 weights per thread fit the register file
            if (backward) {
                if (regs) {
                    cudaFuncSetAttribute(k_lstm_bwd_cl<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lstm_cl_bwd_smem(H));
                    k_lstm_bwd_cl<96><<<grid, H / kClSize * kClBwdParts, lstm_cl_bwd_smem(H), st>>>(jobs[0]);
                } else {
                    cudaFuncSetAttribute(k_lstm_bwd_cl<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lstm_cl_bwd_smem(H));
                    k_lstm_bwd_cl<0><<<grid, H / kClSize * kClBwdParts, lstm_cl_bwd_smem(H), st>>>(jobs[0]);
                }
            } else {
                if (regs) {
                    cudaFuncSetAttribute(k_lstm_fwd_cl<48>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lstm_cl_fwd_smem(H));
                    k_lstm_fwd_cl<48><<<grid, H * kClFwdParts, lstm_cl_fwd_smem(H), st>>>(jobs[0]);
                } else {
                    cudaFuncSetAttribute(k_lstm_fwd_cl<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lstm_cl_fwd_smem(H));
                    k_lstm_fwd_cl<0><<<grid, H * kClFwdParts, lstm_cl_fwd_smem(H), st>>>(jobs[0]);
                }
            }
            check();
            return;
        }
        for (int i = 0; i < n_jobs; i++) {
            jobs[i].G = G; jobs[i].w_smem = w_smem;
            pack.j[i] = jobs[i];
            max_blocks = std::max(max_blocks, (jobs[i].S + G - 1) / G);
        }
        if (G == 1 && H == 8 && n_jobs == 1) {
            // one narrow sequence per warp: shuffles instead of shared memory and barriers (mst_model_kernels2.cuh)
            const int blocks = std::min(jobs[0].S, 148 * 16);
            if (backward) k_lstm_bwd_w8<<<blocks, 32, 0, st>>>(jobs[0]); else k_lstm_fwd_w8<<<blocks, 32, 0, st>>>(jobs[0]);
            check();
            return;
        }
        if (G == 1 && H == 64) {
            // one long sequence per CTA: W_hh rows in registers (mst_model_kernels2.cuh)
            dim3 grid(std::min(max_blocks, std::max(1, 148 * 4 / n_jobs)), n_jobs);
            if (backward) k_lstm_bwd_r64<<<grid, 256, 0, st>>>(pack); else k_lstm_fwd_r64<<<grid, 256, 0, st>>>(pack);
            check();
            return;
        }
        const size_t smem = lstm_smem(H, G, w_smem);
        const int threads = std::max(32, (4 * H + 31) / 32 * 32);
        // the kernels loop over the sequence groups: enough CTAs to fill every SM, W_hh staged once per CTA
        const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(2048 / threads, (227 * 1024) / (smem + 1024)));
        dim3 grid(std::min(max_blocks, std::max(1, 148 * per_sm / n_jobs)), n_jobs);
        if (backward) {
            cudaFuncSetAttribute(k_lstm_bwd, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
            k_lstm_bwd<<<grid, threads, smem, st>>>(pack);
        } else {
            cudaFuncSetAttribute(k_lstm_fwd, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
            k_lstm_fwd<<<grid, threads, smem, st>>>(pack);
        }
        check();
    }
    void lstm(LstmArgs a, bool backward) { lstm_multi(&a, 1, backward); }
    void combine_fwd(CombineArgs a) {
        if (dry) return;
        fill(a.ss, (long long)a.N * a.K, 0.f);
        dim3 g1((unsigned)std::max<long long>(1, std::min<long long>((a.M + 2047) / 2048, 64)), a.N * a.K);
        k_combine_sumsq<<<g1, 256, 0, st>>>(a); check();
        dim3 g2((unsigned)std::max<long long>(1, std::min<long long>((a.M + 255) / 256, 256)), a.N);
        k_combine_fwd<<<g2, 256, 0, st>>>(a); check();
    }
    void combine_bwd(CombineArgs a) {
        if (dry) return;
        fill(a.stats, (long long)a.N * (a.K + 1), 0.f);
        dim3 g1((unsigned)std::max<long long>(1, std::min<long long>((a.M + 2047) / 2048, 64)), a.N);
        k_combine_bwd_stats<<<g1, 256, 0, st>>>(a); check();
        dim3 g2((unsigned)std::max<long long>(1, std::min<long long>((a.M + 255) / 256, 256)), a.N);
        k_combine_bwd<<<g2, 256, 0, st>>>(a); check();
    }
};

}  // namespace

// ------------------------------------------------------------------------------------------------ context
namespace {

__global__ void __launch_bounds__(256) k_transpose(const float *__restrict__ in, float *__restrict__ out, int rows, int cols) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows * cols; i += gridDim.x * blockDim.x) {
        const int r = i / cols, c = i - r * cols;
        out[c * rows + r] = in[i];
    }
}

struct LstmBuf {
    float *xproj, *gates, *cells, *hprev, *dgates, *bias_sum, *wt;
    int S, T, H, I;
    void carve(Arena &ar, int S_, int T_, int H_, int I_) {
        S = S_; T = T_; H = H_; I = I_;
        const long long st = (long long)S * T;
        xproj = ar.take(st * 4 * H); gates = ar.take(st * 4 * H); cells = ar.take(st * H); hprev = ar.take(st * H);
        dgates = ar.take(st * 4 * H); bias_sum = ar.take(4 * H); wt = ar.take((long long)H * 4 * H);
    }
};

struct Ctx {
    int N, C, B, T, U;
    bool fused;                     // forward_extract + _song_info + _apply run in one call: pieces move between them (see forward_extract)
    long long Rb, Rf, Rbf, Nb;      // beat rows (n,c,b,t), fraction rows, (n,b,t,f) rows, bar rows (n,b)
    const float *P, *Ug, *F, *mode, *bpm;   // inputs
    const float *prm;               // flat parameters
    float *grad;                    // flat gradients (may be null for inference)
    // permuted weights
    float *wconv_p, *wpcelin_p, *wucelin_p, *wconv_t, *wprech_t;
    __nv_bfloat16 *pce_conv_h, *wpce_h;   // bf16 dense mode: bf16 copies of the conv features and of linear.weight[:, :456] for the TMA-fed GEMM
    // occupied cells of the pitched grid, CSR by beat row
    uint32_t *cells_cnt, *cells_off; CellRec *cells_rec;
    // PCE
    float *pce_conv, *pce_il, *pce_zi, *pce_zc, *pce_x, *pce_beats, *pce_last, *pce_ss, *pce_bars_in, *pce_bars;
    LstmBuf pce_bl, pce_brf, pce_brr;
    // PRE
    float *pre_a1, *pre_a2, *pre_a3, *pre_a4, *pre_a5, *pre_a6, *pre_z[6], *pre_x, *pre_ss, *pre_out;
    // UCE / URE
    float *uce_lin, *uce_beats, *uce_last, *uce_ss, *uce_bars_in, *uce_bars;
    LstmBuf uce_bl, uce_brf, uce_brr;
    float *ure_a1, *ure_a2, *ure_a3, *ure_a4, *ure_z[4], *ure_x, *ure_ss, *ure_out;
    // pair combines
    float *bars_pair, *bars_ss, *bars_c, *rh_pair, *rh_ss, *rh_c;
    const float *bars, *rhythm;     // what the rest of the model consumes
    // SE
    LstmBuf se_l;
    float *se_h, *se_last, *se_a2, *se_a3, *se_a4, *se_z[4], *se_x, *se_ss, *style;
    // ME
    float *me_a1, *me_a2, *me_a3, *me_z[3], *me_od, *me_ss, *melody;
    // SIM
    LstmBuf sim_l1, sim_l2;
    float *sim_h1, *sim_last1, *sim_h2, *sim_rf, *sim_si, *sim_ri, *sim_instr, *sim_sm, *sim_rm, *sim_mode, *sim_sb, *sim_rb, *sim_bpm_pre, *sim_bpm;
    // PSA / USA
    float *psa_s66, *psa_r10, *psa_i12, *psa_zs, *psa_zr, *psa_zi, *psa_zci, *psa_out;
    float *usa_s650, *usa_r17, *usa_zs, *usa_zr, *usa_h, *usa_out;
    // backward scratch (carved by plan_backward)
    float *scratch[8]; long long scratch_floats[8];
};

const float *PRM(const Ctx &c, ParamId id) { return c.prm + g_param_off[id]; }
float *GRD(const Ctx &c, ParamId id) { return c.grad + g_param_off[id]; }

size_t plan_forward(Ctx &c, Arena &ar) {
    const long long N = c.N, C = c.C, B = c.B, T = c.T;
    c.Rb = N * C * B * T; c.Rf = c.Rb * 10; c.Rbf = N * B * T * 10; c.Nb = N * B;
    const long long Rb = c.Rb, Rf = c.Rf, Rbf = c.Rbf, Nb = c.Nb, NC = N * C;
    c.wconv_p = ar.take(57 * 700); c.wpcelin_p = ar.take(64 * 514); c.wucelin_p = ar.take(64 * 940);
    c.wconv_t = ar.take(700 * 57); c.wprech_t = ar.take(280 * 16);
    c.cells_cnt = (uint32_t *)ar.take(Rb); c.cells_off = (uint32_t *)ar.take(Rb + 1);
    c.cells_rec = (CellRec *)ar.take(Rb * 560 * (long long)(sizeof(CellRec) / sizeof(float)));   // worst case: every cell occupied
    c.pce_conv_h = (__nv_bfloat16 *)ar.take(Rb * 228); c.wpce_h = (__nv_bfloat16 *)ar.take(64 * 256);
    c.pce_conv = ar.take(Rb * 456); c.pce_il = ar.take(NC * 58); c.pce_zi = ar.take(NC * 64); c.pce_zc = ar.take(Rb * 64);
    c.pce_x = ar.take(Rb * 64); c.pce_beats = ar.take(Rb * 64); c.pce_last = ar.take(NC * B * 64); c.pce_ss = ar.take(NC);
    c.pce_bars_in = ar.take(Nb * 64); c.pce_bars = ar.take(Nb * 128);
    c.pce_bl.carve(ar, (int)(NC * B), (int)T, 64, 64); c.pce_brf.carve(ar, (int)N, (int)B, 64, 64); c.pce_brr.carve(ar, (int)N, (int)B, 64, 64);
    c.pre_a1 = ar.take(Rb * 48); c.pre_a2 = ar.take(Nb * 40); c.pre_a3 = ar.take(Rf * 16); c.pre_a4 = ar.take(NC * 21);
    c.pre_a5 = ar.take(N * 5); c.pre_a6 = ar.take(N * 5);
    const long long pre_rows[6] = {Rb, Nb, Rf, NC, N, N};
    for (int i = 0; i < 6; i++) c.pre_z[i] = ar.take(pre_rows[i] * 32);
    c.pre_x = ar.take(Rf * 32); c.pre_ss = ar.take(NC); c.pre_out = ar.take(Rbf * 32);
    if (c.U) {
        c.uce_lin = ar.take(Nb * T * 64); c.uce_beats = ar.take(Nb * T * 64); c.uce_last = ar.take(Nb * 64); c.uce_ss = ar.take(N);
        c.uce_bars_in = ar.take(Nb * 64); c.uce_bars = ar.take(Nb * 128);
        c.uce_bl.carve(ar, (int)Nb, (int)T, 64, 64); c.uce_brf.carve(ar, (int)N, (int)B, 64, 64); c.uce_brr.carve(ar, (int)N, (int)B, 64, 64);
        c.ure_a1 = ar.take(Nb * T * 48); c.ure_a2 = ar.take(Nb * 40); c.ure_a3 = ar.take(Rbf * 16); c.ure_a4 = ar.take(N * 5);
        const long long ure_rows[4] = {Nb * T, Nb, Rbf, N};
        for (int i = 0; i < 4; i++) c.ure_z[i] = ar.take(ure_rows[i] * 32);
        c.ure_x = ar.take(Rbf * 32); c.ure_ss = ar.take(N); c.ure_out = ar.take(Rbf * 32);
        c.bars_pair = ar.take(N * 2 * B * 128); c.bars_ss = ar.take(N * 2); c.bars_c = ar.take(Nb * 128);
        c.rh_pair = ar.take(2 * Rbf * 32); c.rh_ss = ar.take(N * 2); c.rh_c = ar.take(Rbf * 32);
    }
    c.se_l.carve(ar, (int)N, (int)B, 192, 128);
    c.se_h = ar.take(Nb * 192); c.se_last = ar.take(N * 192); c.se_a2 = ar.take(NC * 39); c.se_a3 = ar.take(N * 13); c.se_a4 = ar.take(N * 7);
    const long long se_rows[4] = {N, NC, N, N};
    for (int i = 0; i < 4; i++) c.se_z[i] = ar.take(se_rows[i] * 256);
    c.se_x = ar.take(NC * 256); c.se_ss = ar.take(NC); c.style = ar.take(N * 256);
    c.me_a1 = ar.take(Rb * 36); c.me_a2 = ar.take(Nb * 68); c.me_a3 = ar.take(NC * 8);
    const long long me_rows[3] = {Rb, Nb, NC};
    for (int i = 0; i < 3; i++) c.me_z[i] = ar.take(me_rows[i] * 120);
    c.me_od = ar.take(Rb * 120); c.me_ss = ar.take(NC); c.melody = ar.take(Rbf * 56 * 8);
    c.sim_l1.carve(ar, (int)Nb, (int)T, 9, 320); c.sim_l2.carve(ar, (int)N, (int)B, 8, 9);
    c.sim_h1 = ar.take(Nb * T * 9); c.sim_last1 = ar.take(Nb * 9); c.sim_h2 = ar.take(Nb * 8); c.sim_rf = ar.take(N * 8);
    c.sim_si = ar.take(N * 8); c.sim_ri = ar.take(N * 10); c.sim_instr = ar.take(N * 41);
    c.sim_sm = ar.take(N * 2); c.sim_rm = ar.take(N * 2); c.sim_mode = ar.take(N * 2);
    c.sim_sb = ar.take(N * 2); c.sim_rb = ar.take(N * 2); c.sim_bpm_pre = ar.take(N); c.sim_bpm = ar.take(N);
    c.psa_s66 = ar.take(N * 66); c.psa_r10 = ar.take(Rbf * 10); c.psa_i12 = ar.take(NC * 12);
    c.psa_zs = ar.take(N * 450); c.psa_zr = ar.take(Rbf * 450); c.psa_zi = ar.take(NC * 450); c.psa_zci = ar.take(NC * 450); c.psa_out = ar.take(Rf * 56 * 5);
    if (c.U) {
        c.usa_s650 = ar.take(N * 650); c.usa_r17 = ar.take(Rbf * 17); c.usa_zs = ar.take(N * 10 * 376); c.usa_zr = ar.take(Rbf * 376);
        c.usa_h = ar.take(Rbf * 376); c.usa_out = ar.take(Rbf * 47 * 2);
    }
    return ar.off;
}

// LSTM weights: bias_ih + bias_hh and W_hh^T, refreshed every step because the weights change
void prep_lstm(Launcher &L, const Ctx &c, LstmBuf &lb, ParamId whh, ParamId bih, ParamId bhh) {
    k_add2<<<grid_for(4 * lb.H), 256, 0, L.st>>>(PRM(c, bih), PRM(c, bhh), lb.bias_sum, 4 * lb.H); L.check();
    k_transpose<<<grid_for(4 * lb.H * lb.H), 256, 0, L.st>>>(PRM(c, whh), lb.wt, 4 * lb.H, lb.H); L.check();
}

// weight preparation, split by the branch that needs it first (forward_extract)
void prep_weights_main(Launcher &L, Ctx &c) {
    k_permute_cols<<<grid_for(57 * 700), 256, 0, L.st>>>(PRM(c, pce_conv_w), c.wconv_p, 57, 700, 0, 0, 0); L.check();
    k_permute_cols<<<grid_for(64 * 514), 256, 0, L.st>>>(PRM(c, pce_lin_w), c.wpcelin_p, 64, 514, 1, 0, 0); L.check();
    k_transpose<<<grid_for(57 * 700), 256, 0, L.st>>>(c.wconv_p, c.wconv_t, 57, 700); L.check();
    k_transpose<<<grid_for(16 * 280), 256, 0, L.st>>>(PRM(c, pre_ch_w), c.wprech_t, 16, 280); L.check();
}
void prep_weights_uce(Launcher &L, Ctx &c) {
    k_permute_cols<<<grid_for(64 * 940), 256, 0, L.st>>>(PRM(c, uce_lin_w), c.wucelin_p, 64, 940, 2, 0, 0); L.check();
    prep_lstm(L, c, c.uce_bl, uce_bl_whh, uce_bl_bih, uce_bl_bhh);
    prep_lstm(L, c, c.uce_brf, uce_br_whh, uce_br_bih, uce_br_bhh);
    prep_lstm(L, c, c.uce_brr, uce_br_whh_r, uce_br_bih_r, uce_br_bhh_r);
}
void prep_weights_lstm_first(Launcher &L, Ctx &c) {     // what the pitched chain needs at its first recurrences
    prep_lstm(L, c, c.pce_bl, pce_bl_whh, pce_bl_bih, pce_bl_bhh);
    prep_lstm(L, c, c.pce_brf, pce_br_whh, pce_br_bih, pce_br_bhh);
    prep_lstm(L, c, c.pce_brr, pce_br_whh_r, pce_br_bih_r, pce_br_bhh_r);
}
void prep_weights_lstm_late(Launcher &L, Ctx &c) {      // StyleEncoder and SongInfoModel recurrences: hundreds of microseconds later
    prep_lstm(L, c, c.se_l, se_lstm_whh, se_lstm_bih, se_lstm_bhh);
    prep_lstm(L, c, c.sim_l1, sim_bl_whh, sim_bl_bih, sim_bl_bhh);
    prep_lstm(L, c, c.sim_l2, sim_br_whh, sim_br_bih, sim_br_bhh);
}

// input projection of one LSTM direction for all time steps + the recurrence's arguments
LstmArgs lstm_project(Launcher &L, const Ctx &c, LstmBuf &lb, const float *X, ParamId wih, float *out, long long out_ld, int out_off, int reverse) {
    L.gemm_cls = 8;
    L.linear(X, lb.I, PRM(c, wih), lb.I, lb.bias_sum, lb.xproj, 4 * lb.H, lb.S * lb.T, 4 * lb.H, lb.I, 0);
    L.gemm_cls = 1;
    LstmArgs a{};
    a.xproj = lb.xproj; a.w = lb.wt; a.out = out; a.out_ld = out_ld; a.out_off = out_off;
    a.gates = lb.gates; a.cells = lb.cells; a.hprev = lb.hprev; a.S = lb.S; a.T = lb.T; a.H = lb.H; a.reverse = reverse;
    return a;
}

// run one LSTM direction: xproj = X·W_ih^T + (b_ih + b_hh), then the recurrence
void lstm_forward(Launcher &L, const Ctx &c, LstmBuf &lb, const float *X, ParamId wih, float *out, long long out_ld, int out_off, int reverse) {
    L.gemm_cls = 8;
    L.linear(X, lb.I, PRM(c, wih), lb.I, lb.bias_sum, lb.xproj, 4 * lb.H, lb.S * lb.T, 4 * lb.H, lb.I, 0);
    L.gemm_cls = 1;
    LstmArgs a{};
    a.xproj = lb.xproj; a.w = lb.wt; a.out = out; a.out_ld = out_ld; a.out_off = out_off;
    a.gates = lb.gates; a.cells = lb.cells; a.hprev = lb.hprev; a.S = lb.S; a.T = lb.T; a.H = lb.H; a.reverse = reverse;
    L.lstm(a, false);
}

}  // namespace

// ------------------------------------------------------------------------------------------------ forward
namespace {

// masks over (n, c, b, t, f)
constexpr int M_N = 16, M_C = 8, M_B = 4, M_T = 2, M_F = 1;

void bcast_fwd(Launcher &L, Dims5 d, int cols, const float *bias, float *Y, int act, std::initializer_list<std::pair<const float *, int>> srcs) {
    BcastArgs a{};
    int i = 0;
    for (auto &s : srcs) { a.src[i] = s.first; a.mask[i] = s.second; i++; }
    a.n_src = i; a.bias = bias; a.Y = Y; a.d = d; a.cols = cols; a.act = act;
    L.bcast(a);
}

// channels_linear / linear of the MelodyEncoder into the constant bank (device -> device, stream ordered)
void load_me_constants(Launcher &L, const Ctx &c) {
    if (L.dry) return;
    const struct { ParamId id; int off, n; } parts[4] = {{me_ch_w, kMeWc, 35}, {me_ch_b, kMeBc, 7}, {me_lin_w, kMeWl, 120}, {me_lin_b, kMeBl, 8}};
    for (auto &p : parts) {
        cudaError_t e = cudaMemcpyToSymbolAsync(c_me, PRM(c, p.id), p.n * sizeof(float), p.off * sizeof(float), cudaMemcpyDeviceToDevice, L.st);
        if (e != cudaSuccess && L.err == cudaSuccess) L.err = e;
    }
}
MeArgs2 me_args(const Ctx &c) {
    MeArgs2 a{};
    a.od = c.me_od; a.off = c.cells_off; a.recs = c.cells_rec; a.ss = c.me_ss; a.melody = c.melody;
    a.N = c.N; a.C = c.C; a.B = c.B; a.T = c.T;
    return a;
}

void forward_extract(Launcher &L, Ctx &c) {
    const int N = c.N, C = c.C, B = c.B, T = c.T, NC = N * C;
    const long long Rb = c.Rb, Rf = c.Rf, Rbf = c.Rbf, Nb = c.Nb;
    // Branches (Launcher::wait / on): 1 = the LSTM weight preparations, needed by the caller's stream at its first recurrence;
    // 0 = the UnpitchedChannelsEncoder up to its bars input (model.py:128-138), independent of the pitched chain.  Both start
    // after whatever the caller queued before this call (the previous optimiser step).
    LstmArgs bar_jobs[4];
    int n_bar_jobs = 2;
    L.wait(1, -1); L.on(1);
    prep_weights_lstm_first(L, c);
    if (c.U) {
        L.wait(0, -1); L.on(0);
        prep_weights_uce(L, c);
        L.linear(c.Ug, 940, c.wucelin_p, 940, PRM(c, uce_lin_b), c.uce_lin, 64, (int)(Nb * T), 64, 940, 1);
        lstm_forward(L, c, c.uce_bl, c.uce_lin, uce_bl_wih, c.uce_beats, 64, 0, 0);
        L.take_last(c.uce_beats, c.uce_last, Nb, T, 64);
        { CombineArgs a{}; a.x = c.uce_last; a.M = (long long)B * 64; a.n_stride = a.M; a.k_stride = a.M; a.N = N; a.K = 1;
          a.ss = c.uce_ss; a.out = c.uce_bars_in; L.combine_fwd(a); }
        bar_jobs[2] = lstm_project(L, c, c.uce_brf, c.uce_bars_in, uce_br_wih, c.uce_bars, 128, 0, 0);
        bar_jobs[3] = lstm_project(L, c, c.uce_brr, c.uce_bars_in, uce_br_wih_r, c.uce_bars, 128, 64, 1);
        n_bar_jobs = 4;
    }
    L.on(-1);
    prep_weights_main(L, c);
    // ---------------- occupied cells of the piano roll, CSR by beat row (mst_model_sparse.cuh)
    {
        const int g = (int)std::min<long long>((Rb + 7) / 8, 148 * 8);
        k_cells_scan<0><<<g, 256, 0, L.st>>>(c.P, Rb, c.cells_cnt, nullptr, nullptr); L.check();
        k_cells_offsets<<<1, 1024, 0, L.st>>>(c.cells_cnt, c.cells_off, Rb); L.check();
        k_cells_scan<1><<<g, 256, 0, L.st>>>(c.P, Rb, nullptr, c.cells_off, c.cells_rec); L.check();
        // ---------------- PitchedChannelsEncoder (model.py:77-99): beats_conv + leaky over the occupied cells
        k_conv_sparse_fwd<<<g, 256, 0, L.st>>>(c.cells_off, c.cells_rec, c.wconv_t, PRM(c, pce_conv_b), c.pce_conv, Rb,
                                               L.use_tma() ? c.pce_conv_h : nullptr); L.check();
    }
    L.linear(c.F, 51, PRM(c, pce_il_w), 51, PRM(c, pce_il_b), c.pce_il, 58, NC, 58, 51, 1);
    L.linear(c.pce_il, 58, c.wpcelin_p + 456, 514, nullptr, c.pce_zi, 64, NC, 64, 58, 0);
    if (L.use_tma()) {
        // the largest contraction of the model, fed by tensor-map TMA from the bf16 copies (mst_model_gemm_tma.cuh)
        k_cast_bf16<<<grid_for(64 * 512), 256, 0, L.st>>>(c.wpcelin_p, 514, c.wpce_h, 512, 64, 456, 512); L.check();
        L.gemm_tma64(c.pce_conv_h, 456, Rb, c.wpce_h, 512, 64, 456, c.pce_zc, 64, nullptr, 0, 0);
    } else {
        L.linear(c.pce_conv, 456, c.wpcelin_p, 514, nullptr, c.pce_zc, 64, (int)Rb, 64, 456, 0);
    }
    bcast_fwd(L, Dims5{N, C, B, T, 1}, 64, PRM(c, pce_lin_b), c.pce_x, 1, {{c.pce_zc, M_N | M_C | M_B | M_T}, {c.pce_zi, M_N | M_C}});
    L.wait(-1, 1);                                   // LSTM weights of the channel encoder ready
    // queued on branch 1 only now (host order), so that the wait above does not cover it: the weights of the later recurrences
    // and the linears that read only instrument features / mode / bpm; the caller's stream takes them in front of the StyleEncoder
    L.on(1);
    prep_weights_lstm_late(L, c);
    {   // linears that read only the instrument features / mode / bpm: off the chains that end in the style vector
        const float *slw = PRM(c, se_lin_w);
        L.linear(c.F, 51, PRM(c, se_il_w), 51, PRM(c, se_il_b), c.se_a2, 39, NC, 39, 51, 1);
        L.linear(c.mode, 2, PRM(c, se_ml_w), 2, PRM(c, se_ml_b), c.se_a3, 13, N, 13, 2, 1);
        L.linear(c.bpm, 1, PRM(c, se_bp_w), 1, PRM(c, se_bp_b), c.se_a4, 7, N, 7, 1, 1);
        L.linear(c.se_a2, 39, slw + 192, 251, nullptr, c.se_z[1], 256, NC, 256, 39, 0);
        L.linear(c.se_a3, 13, slw + 231, 251, nullptr, c.se_z[2], 256, N, 256, 13, 0);
        L.linear(c.se_a4, 7, slw + 244, 251, nullptr, c.se_z[3], 256, N, 256, 7, 0);
        if (c.fused) {
            L.linear(c.F, 51, PRM(c, psa_il_w), 51, PRM(c, psa_il_b), c.psa_i12, 12, NC, 12, 51, 1);
            L.linear(c.psa_i12, 12, PRM(c, psa_oct_w) + 76, 88, nullptr, c.psa_zi, 450, NC, 450, 12, 0);
        }
    }
    L.on(-1);
    lstm_forward(L, c, c.pce_bl, c.pce_x, pce_bl_wih, c.pce_beats, 64, 0, 0);
    L.take_last(c.pce_beats, c.pce_last, (long long)NC * B, T, 64);
    { CombineArgs a{}; a.x = c.pce_last; a.n_stride = (long long)C * B * 64; a.k_stride = (long long)B * 64; a.N = N; a.K = C; a.M = (long long)B * 64;
      a.ss = c.pce_ss; a.out = c.pce_bars_in; L.combine_fwd(a); }
    // the bar-axis BiLSTMs of both channel encoders are independent serial recurrences: one joint launch
    bar_jobs[0] = lstm_project(L, c, c.pce_brf, c.pce_bars_in, pce_br_wih, c.pce_bars, 128, 0, 0);
    bar_jobs[1] = lstm_project(L, c, c.pce_brr, c.pce_bars_in, pce_br_wih_r, c.pce_bars, 128, 64, 1);
    L.wait(-1, 0);                                   // the unpitched encoder's bars input
    L.lstm_multi(bar_jobs, n_bar_jobs, false);
    // From here three independent branches: 0 = both rhythm encoders and their combine (then the SongInfoModel recurrences,
    // forward_song_info), 1 = MelodyEncoder, caller's stream = combine of the bars + StyleEncoder (its 192-wide recurrence
    // over the bars is the longest serial kernel of the forward pass).
    L.wait(-1, 1);                                   // the late weights and static linears (queued long ago)
    L.wait(0, -1); L.wait(1, -1);
    L.on(0);
    // ---------------- PitchedRhythmEncoder (model.py:346-381)
    const float *plw = PRM(c, pre_lin_w);
    L.linear(c.pce_beats, 64, PRM(c, pre_bl_w), 64, PRM(c, pre_bl_b), c.pre_a1, 48, (int)Rb, 48, 64, 1);
    L.linear(c.pce_bars, 128, PRM(c, pre_br_w), 128, PRM(c, pre_br_b), c.pre_a2, 40, (int)Nb, 40, 128, 1);
    k_chlin_sparse_fwd<<<(int)std::min<long long>((Rb + 7) / 8, 148 * 8), 256, 0, L.st>>>(c.cells_off, c.cells_rec, c.wprech_t, PRM(c, pre_ch_b), c.pre_a3, Rb); L.check();
    L.linear(c.F, 51, PRM(c, pre_il_w), 51, PRM(c, pre_il_b), c.pre_a4, 21, NC, 21, 51, 1);
    L.linear(c.mode, 2, PRM(c, pre_ml_w), 2, PRM(c, pre_ml_b), c.pre_a5, 5, N, 5, 2, 1);
    L.linear(c.bpm, 1, PRM(c, pre_bp_w), 1, PRM(c, pre_bp_b), c.pre_a6, 5, N, 5, 1, 1);
    L.linear(c.pre_a1, 48, plw + 0, 135, nullptr, c.pre_z[0], 32, (int)Rb, 32, 48, 0);
    L.linear(c.pre_a2, 40, plw + 48, 135, nullptr, c.pre_z[1], 32, (int)Nb, 32, 40, 0);
    L.linear(c.pre_a3, 16, plw + 88, 135, nullptr, c.pre_z[2], 32, (int)Rf, 32, 16, 0);
    L.linear(c.pre_a4, 21, plw + 104, 135, nullptr, c.pre_z[3], 32, NC, 32, 21, 0);
    L.linear(c.pre_a5, 5, plw + 125, 135, nullptr, c.pre_z[4], 32, N, 32, 5, 0);
    L.linear(c.pre_a6, 5, plw + 130, 135, nullptr, c.pre_z[5], 32, N, 32, 5, 0);
    bcast_fwd(L, Dims5{N, C, B, T, 10}, 32, PRM(c, pre_lin_b), c.pre_x, 1,
              {{c.pre_z[0], M_N | M_C | M_B | M_T}, {c.pre_z[1], M_N | M_B}, {c.pre_z[2], M_N | M_C | M_B | M_T | M_F},
               {c.pre_z[3], M_N | M_C}, {c.pre_z[4], M_N}, {c.pre_z[5], M_N}});
    { CombineArgs a{}; a.x = c.pre_x; a.M = (long long)B * T * 320; a.n_stride = C * a.M; a.k_stride = a.M; a.N = N; a.K = C;
      a.ss = c.pre_ss; a.out = c.pre_out; L.combine_fwd(a); }
    c.bars = c.pce_bars; c.rhythm = c.pre_out;
    if (c.U) {
        // ---------------- UnpitchedRhythmEncoder (model.py:418-443)
        const float *ulw = PRM(c, ure_lin_w);
        L.linear(c.uce_beats, 64, PRM(c, ure_bl_w), 64, PRM(c, ure_bl_b), c.ure_a1, 48, (int)(Nb * T), 48, 64, 1);
        L.linear(c.uce_bars, 128, PRM(c, ure_br_w), 128, PRM(c, ure_br_b), c.ure_a2, 40, (int)Nb, 40, 128, 1);
        L.linear(c.Ug, 94, PRM(c, ure_ch_w), 94, PRM(c, ure_ch_b), c.ure_a3, 16, (int)Rbf, 16, 94, 1);
        L.linear(c.bpm, 1, PRM(c, ure_bp_w), 1, PRM(c, ure_bp_b), c.ure_a4, 5, N, 5, 1, 1);
        L.linear(c.ure_a1, 48, ulw + 0, 109, nullptr, c.ure_z[0], 32, (int)(Nb * T), 32, 48, 0);
        L.linear(c.ure_a2, 40, ulw + 48, 109, nullptr, c.ure_z[1], 32, (int)Nb, 32, 40, 0);
        L.linear(c.ure_a3, 16, ulw + 88, 109, nullptr, c.ure_z[2], 32, (int)Rbf, 32, 16, 0);
        L.linear(c.ure_a4, 5, ulw + 104, 109, nullptr, c.ure_z[3], 32, N, 32, 5, 0);
        bcast_fwd(L, Dims5{N, 1, B, T, 10}, 32, PRM(c, ure_lin_b), c.ure_x, 1,
                  {{c.ure_z[0], M_N | M_B | M_T}, {c.ure_z[1], M_N | M_B}, {c.ure_z[2], M_N | M_B | M_T | M_F}, {c.ure_z[3], M_N}});
        { CombineArgs a{}; a.x = c.ure_x; a.M = (long long)B * T * 320; a.n_stride = a.M; a.k_stride = a.M; a.N = N; a.K = 1;
          a.ss = c.ure_ss; a.out = c.ure_out; L.combine_fwd(a); }
        // ---------------- combine(pitched, unpitched) for the rhythm (model.py:767)
        k_pair_pack<<<grid_for(Rbf * 32), 256, 0, L.st>>>(c.pre_out, c.ure_out, c.rh_pair, N, (long long)B * T * 320); L.check();
        { CombineArgs a{}; a.x = c.rh_pair; a.M = (long long)B * T * 320; a.n_stride = 2 * a.M; a.k_stride = a.M; a.N = N; a.K = 2;
          a.ss = c.rh_ss; a.out = c.rh_c; L.combine_fwd(a); }
        c.rhythm = c.rh_c;
    }
    if (c.fused) {
        // the appliers' linears that read only the rhythm follow it on this branch (forward_apply skips them)
        L.linear(c.rhythm, 32, PRM(c, psa_rl_w), 32, PRM(c, psa_rl_b), c.psa_r10, 10, (int)Rbf, 10, 32, 1);
        L.linear(c.psa_r10, 10, PRM(c, psa_oct_w) + 66, 88, nullptr, c.psa_zr, 450, (int)Rbf, 450, 10, 0);
        if (c.U) {
            L.linear(c.rhythm, 32, PRM(c, usa_rl_w), 32, PRM(c, usa_rl_b), c.usa_r17, 17, (int)Rbf, 17, 32, 1);
            L.linear(c.usa_r17, 17, PRM(c, usa_nl_w) + 65, 82, nullptr, c.usa_zr, 376, (int)Rbf, 376, 17, 0);
        }
    }
    L.on(-1);
    if (c.U) {
        // ---------------- combine(pitched, unpitched) for the bars (model.py:766)
        k_pair_pack<<<grid_for(Nb * 128), 256, 0, L.st>>>(c.pce_bars, c.uce_bars, c.bars_pair, N, (long long)B * 128); L.check();
        { CombineArgs a{}; a.x = c.bars_pair; a.M = (long long)B * 128; a.n_stride = 2 * a.M; a.k_stride = a.M; a.N = N; a.K = 2;
          a.ss = c.bars_ss; a.out = c.bars_c; L.combine_fwd(a); }
        c.bars = c.bars_c;
    }
    // ---------------- StyleEncoder (model.py:179-200)
    lstm_forward(L, c, c.se_l, c.bars, se_lstm_wih, c.se_h, 192, 0, 0);
    L.take_last(c.se_h, c.se_last, N, B, 192);
    const float *slw = PRM(c, se_lin_w);
    L.linear(c.se_last, 192, slw + 0, 251, nullptr, c.se_z[0], 256, N, 256, 192, 0);     // se_z[1..3]: branch 1, top of this function
    bcast_fwd(L, Dims5{N, C, 1, 1, 1}, 256, PRM(c, se_lin_b), c.se_x, 1,
              {{c.se_z[0], M_N}, {c.se_z[1], M_N | M_C}, {c.se_z[2], M_N}, {c.se_z[3], M_N}});
    { CombineArgs a{}; a.x = c.se_x; a.M = 256; a.n_stride = (long long)C * 256; a.k_stride = 256; a.N = N; a.K = C;
      a.ss = c.se_ss; a.out = c.style; L.combine_fwd(a); }
    // fused forward: the caller's stream takes what branch 0 has produced so far (rhythm, psa_zr, usa_zr) - recorded NOW, before
    // forward_song_info queues the SongInfoModel on that branch, so the PitchedStyleApplier does not wait for it
    if (c.fused) L.wait(-1, 0);
    // ---------------- MelodyEncoder (model.py:252-297); octave+degree weights act as one (120, 112) matrix
    L.on(1);
    const float *mow = PRM(c, me_oct_w);
    L.linear(c.pce_beats, 64, PRM(c, me_bl_w), 64, PRM(c, me_bl_b), c.me_a1, 36, (int)Rb, 36, 64, 1);
    L.linear(c.pce_bars, 128, PRM(c, me_br_w), 128, PRM(c, me_br_b), c.me_a2, 68, (int)Nb, 68, 128, 1);
    L.linear(c.F, 51, PRM(c, me_il_w), 51, PRM(c, me_il_b), c.me_a3, 8, NC, 8, 51, 1);
    L.linear(c.me_a1, 36, mow + 0, 112, nullptr, c.me_z[0], 120, (int)Rb, 120, 36, 0);
    L.linear(c.me_a2, 68, mow + 36, 112, nullptr, c.me_z[1], 120, (int)Nb, 120, 68, 0);
    L.linear(c.me_a3, 8, mow + 104, 112, nullptr, c.me_z[2], 120, NC, 120, 8, 0);
    bcast_fwd(L, Dims5{N, C, B, T, 1}, 120, PRM(c, me_oct_b), c.me_od, 1,
              {{c.me_z[0], M_N | M_C | M_B | M_T}, {c.me_z[1], M_N | M_B}, {c.me_z[2], M_N | M_C}});
    {
        load_me_constants(L, c);
        MeArgs2 a = me_args(c);
        L.fill(c.me_ss, NC, 0.f);
        k_me_sumsq2<<<(int)std::min<long long>((Rb + 7) / 8, 148 * 8), 256, 0, L.st>>>(a); L.check();
        k_me_fwd2<<<(int)std::min<long long>(((long long)N * B * T + 7) / 8, 148 * 8), 256, 0, L.st>>>(a); L.check();
    }
    L.on(-1);
    // returns with branch 0 (rhythm) and branch 1 (melody) still running: forward_song_info / forward_apply wait where they
    // consume them, the entry points end with join_all()
}

void forward_song_info(Launcher &L, Ctx &c) {   // SongInfoModel (model.py:513-562)
    const int N = c.N, B = c.B, T = c.T;
    // the two recurrences read the rhythm only: they follow the rhythm encoders on branch 0 while the caller's stream is still
    // in the StyleEncoder (standalone use, mst_model_song_info / _predict_apply: the branch starts from the caller's stream)
    if (!c.fused && !(L.forked & 1u)) L.wait(0, -1);   // fused: branch 0 carries the rhythm itself
    L.on(0);
    lstm_forward(L, c, c.sim_l1, c.rhythm, sim_bl_wih, c.sim_h1, 9, 0, 0);
    L.take_last(c.sim_h1, c.sim_last1, c.Nb, T, 9);
    lstm_forward(L, c, c.sim_l2, c.sim_last1, sim_br_wih, c.sim_h2, 8, 0, 0);
    L.take_last(c.sim_h2, c.sim_rf, N, B, 8);
    if (c.fused) {
        L.wait(0, -1);                               // fused forward: the heads stay on branch 0, after the style (caller's stream)
    } else {
        L.on(-1);
        L.wait(-1, 0);                               // the heads read the style (caller's stream) and sim_rf (branch 0)
    }
    auto head = [&](ParamId sw, ParamId sb, int ks, float *sa, ParamId rw, ParamId rb, int kr, float *ra, ParamId fw, ParamId fb, int nout, float *out) {
        L.linear(c.style, 256, PRM(c, sw), 256, PRM(c, sb), sa, ks, N, ks, 256, 1);
        L.linear(c.sim_rf, 8, PRM(c, rw), 8, PRM(c, rb), ra, kr, N, kr, 8, 1);
        L.linear(sa, ks, PRM(c, fw), ks + kr, PRM(c, fb), out, nout, N, nout, ks, 0);
        L.linear(ra, kr, PRM(c, fw) + ks, ks + kr, nullptr, out, nout, N, nout, kr, 0, 1);
    };
    head(sim_si_w, sim_si_b, 8, c.sim_si, sim_ri_w, sim_ri_b, 10, c.sim_ri, sim_i_w, sim_i_b, 41, c.sim_instr);
    head(sim_sm_w, sim_sm_b, 2, c.sim_sm, sim_rm_w, sim_rm_b, 2, c.sim_rm, sim_m_w, sim_m_b, 2, c.sim_mode);
    head(sim_sb_w, sim_sb_b, 2, c.sim_sb, sim_rb_w, sim_rb_b, 2, c.sim_rb, sim_b_w, sim_b_b, 1, c.sim_bpm_pre);
    k_bpm_head<<<(N + 63) / 64, 64, 0, L.st>>>(c.sim_bpm_pre, c.sim_bpm, N); L.check();
    L.on(-1);
}

// melody_linear / linear of the PitchedStyleApplier into the constant bank (device -> device, stream ordered)
void load_psa_constants(Launcher &L, const Ctx &c) {
    if (L.dry) return;
    const struct { ParamId id; int off, n; } parts[4] = {{psa_ml_w, kPsaWm, 160}, {psa_ml_b, kPsaBm, 20}, {psa_lin_w, kPsaWl, 250}, {psa_lin_b, kPsaBl, 5}};
    for (auto &p : parts) {
        cudaError_t e = cudaMemcpyToSymbolAsync(c_psa, PRM(c, p.id), p.n * sizeof(float), p.off * sizeof(float), cudaMemcpyDeviceToDevice, L.st);
        if (e != cudaSuccess && L.err == cudaSuccess) L.err = e;
    }
}

void forward_apply(Launcher &L, Ctx &c, bool unpitched) {
    const int N = c.N, C = c.C, B = c.B, T = c.T, NC = N * C;
    const long long Rbf = c.Rbf;
    // style, rhythm (branch 0 unless forward_song_info joined it) and melody (branch 1) meet here; the two appliers are
    // independent of each other: the unpitched one runs on branch 0
    // fused forward: the rhythm-only linears were computed on branch 0 (forward_extract) and the caller's stream has already
    // taken them; branch 0 now holds the SongInfoModel, which nothing here reads
    if (!c.fused && (L.forked & 1u)) L.wait(-1, 0);
    if (unpitched) {
        // ---------------- UnpitchedStyleApplier (model.py:703-724)
        if (!c.fused) L.wait(0, -1);                 // fused: branch 0 waited for the style before the SongInfoModel heads
        L.on(0);
        const float *nw = PRM(c, usa_nl_w);
        L.linear(c.style, 256, PRM(c, usa_sl_w), 256, PRM(c, usa_sl_b), c.usa_s650, 650, N, 650, 256, 1);
        L.linear(c.usa_s650, 65, nw + 0, 82, nullptr, c.usa_zs, 376, N * 10, 376, 65, 0);
        if (!c.fused) {
            L.linear(c.rhythm, 32, PRM(c, usa_rl_w), 32, PRM(c, usa_rl_b), c.usa_r17, 17, (int)Rbf, 17, 32, 1);
            L.linear(c.usa_r17, 17, nw + 65, 82, nullptr, c.usa_zr, 376, (int)Rbf, 376, 17, 0);
        }
        UsaArgs a{};
        a.bias = PRM(c, usa_nl_b); a.Wl = PRM(c, usa_lin_w); a.bl = PRM(c, usa_lin_b); a.Zr = c.usa_zr; a.Zs = c.usa_zs;
        a.H = c.usa_h; a.out = c.usa_out; a.N = N; a.B = B; a.T = T;
        k_usa_fwd<<<(int)std::min<long long>((Rbf + kUsaWarps - 1) / kUsaWarps, 148 * 4), kUsaWarps * 32, 0, L.st>>>(a); L.check();
        L.on(-1);
    }
    // ---------------- PitchedStyleApplier (model.py:624-675); octave+degree weights = one (450, 88) matrix
    const float *pow_ = PRM(c, psa_oct_w);
    L.linear(c.style, 256, PRM(c, psa_sl_w), 256, PRM(c, psa_sl_b), c.psa_s66, 66, N, 66, 256, 1);
    L.linear(c.psa_s66, 66, pow_ + 0, 88, nullptr, c.psa_zs, 450, N, 450, 66, 0);
    if (!c.fused) {
        L.linear(c.rhythm, 32, PRM(c, psa_rl_w), 32, PRM(c, psa_rl_b), c.psa_r10, 10, (int)Rbf, 10, 32, 1);
        L.linear(c.F, 51, PRM(c, psa_il_w), 51, PRM(c, psa_il_b), c.psa_i12, 12, NC, 12, 51, 1);
        L.linear(c.psa_r10, 10, pow_ + 66, 88, nullptr, c.psa_zr, 450, (int)Rbf, 450, 10, 0);
        L.linear(c.psa_i12, 12, pow_ + 76, 88, nullptr, c.psa_zi, 450, NC, 450, 12, 0);
    }
    {
        load_psa_constants(L, c);
        k_psa_zci<<<grid_for((long long)NC * 450), 256, 0, L.st>>>(PRM(c, psa_oct_b), c.psa_zi, c.psa_zs, c.psa_zci, N, C); L.check();
        if (L.forked & 2u) L.wait(-1, 1);            // the melody
        PsaArgs2 a{c.psa_zr, c.psa_zci, c.melody, c.psa_out, N, C, B, T};
        if (L.psa_fwd_tc() && !L.dry) {
            // bf16 dense mode: the 5 x 50 contraction per note row as tcgen05 tiles (mst_model_psa_tc.cuh)
            cudaFuncSetAttribute(k_psa_fwd_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPtcSmem);
            k_psa_fwd_tc<<<(int)std::min<long long>(Rbf, 148 * 4), kPtcThreads, kPtcSmem, L.st>>>(a); L.check();
        } else if (!L.dry) {
            const size_t smem = sizeof(PsaWarpSmem) * kPsaWarps;
            cudaFuncSetAttribute(k_psa_fwd3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            k_psa_fwd3<<<(int)std::min<long long>((Rbf + kPsaWarps - 1) / kPsaWarps, 148), kPsaWarps * 32, smem, L.st>>>(a); L.check();
        }
    }
    if (unpitched || c.fused) L.wait(-1, 0);         // the unpitched applier (and, fused, the SongInfoModel heads)
}

}  // namespace

// ------------------------------------------------------------------------------------------------ losses + backward
namespace {
#include "mst_model_bwd_host.inl"
}  // namespace

// ------------------------------------------------------------------------------------------------ C ABI

static int setup_ctx(Ctx &c, const MstModelShape *s, void *ws, size_t ws_bytes, bool training, bool dry) {
    init_param_offsets();
    memset(&c, 0, sizeof(c));
    c.N = s->n_samples; c.C = s->n_channels; c.B = s->n_bars; c.T = s->n_beats; c.U = s->has_unpitched ? 1 : 0;
    if (c.N < 1 || c.C < 1 || c.B < 1 || c.T < 1 || c.C > 64) { mst_set_error("mst_model: bad shape"); return MST_E_INVALID; }
    Arena ar{(char *)ws, 0, dry};
    plan_forward(c, ar);
    if (training) plan_backward(c, ar);
    if (!dry && ar.off > ws_bytes) { mst_set_error("mst_model: workspace too small (%zu < %zu)", ws_bytes, ar.off); return MST_E_CAPACITY; }
    return (int)0;
}

extern "C" size_t mst_model_workspace_bytes(const MstModelShape *s, int training) {
    Ctx c;
    init_param_offsets();
    memset(&c, 0, sizeof(c));
    c.N = s->n_samples; c.C = s->n_channels; c.B = s->n_bars; c.T = s->n_beats; c.U = s->has_unpitched ? 1 : 0;
    Arena ar{nullptr, 0, true};
    plan_forward(c, ar);
    if (training) plan_backward(c, ar);
    return ar.off;
}

static void fill_outputs(const Ctx &c, MstModelOutputs *o) {
    o->instruments = c.sim_instr; o->mode = c.sim_mode; o->bpm = c.sim_bpm;
    o->pitched = c.psa_out; o->unpitched = c.U ? c.usa_out : nullptr;
    o->style = c.style; o->melody = c.melody; o->rhythm = (float *)c.rhythm;
}

extern "C" int mst_model_forward(const float *d_params, const MstModelShape *shape, const float *d_pitched,
                                 const float *d_unpitched, const float *d_features, const float *d_mode, const float *d_bpm,
                                 MstModelOutputs *outputs, void *d_workspace, size_t workspace_bytes, int training,
                                 void *stream, int *n_launches) {
    if (!d_params || !shape || !d_pitched || !d_features || !d_mode || !d_bpm || !outputs || !d_workspace ||
        (shape->has_unpitched && !d_unpitched)) {
        mst_set_error("mst_model_forward: null pointer argument");
        return MST_E_INVALID;
    }
    Ctx c;
    int rc = setup_ctx(c, shape, d_workspace, workspace_bytes, training != 0, false);
    if (rc) return rc;
    c.P = d_pitched; c.Ug = d_unpitched; c.F = d_features; c.mode = d_mode; c.bpm = d_bpm; c.prm = d_params;
    Launcher L; L.bf16 = shape->precision == MST_PRECISION_BF16; L.begin(stream);
    c.fused = true;
    forward_extract(L, c);
    forward_song_info(L, c);
    forward_apply(L, c, c.U != 0);
    L.join_all();
    if (L.err != cudaSuccess) { mst_set_error("mst_model_forward: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    fill_outputs(c, outputs);
    if (n_launches) *n_launches = L.launches;
    return MST_OK;
}

extern "C" int mst_model_train_step(const float *d_params, float *d_grads, const MstModelShape *shape, const float *d_pitched,
                                    const float *d_unpitched, const float *d_features, const float *d_mode, const float *d_bpm,
                                    const float *d_instruments_target, const float *d_bpm_target, int normalize,
                                    float *d_losses, MstModelOutputs *outputs, void *d_workspace, size_t workspace_bytes,
                                    void *stream, int *n_launches) {
    if (!d_params || !d_grads || !shape || !d_pitched || !d_features || !d_mode || !d_bpm || !d_instruments_target ||
        !d_bpm_target || !d_losses || !d_workspace || (shape->has_unpitched && !d_unpitched)) {
        mst_set_error("mst_model_train_step: null pointer argument");
        return MST_E_INVALID;
    }
    init_param_offsets();
    Ctx c;
    memset(&c, 0, sizeof(c));
    c.N = shape->n_samples; c.C = shape->n_channels; c.B = shape->n_bars; c.T = shape->n_beats; c.U = shape->has_unpitched ? 1 : 0;
    if (c.N < 1 || c.C < 1 || c.B < 1 || c.T < 1 || c.C > 64) { mst_set_error("mst_model: bad shape"); return MST_E_INVALID; }
    if (mst_model_workspace_bytes(shape, 1) > workspace_bytes) {
        mst_set_error("mst_model_train_step: workspace too small (%zu < %zu)", workspace_bytes, mst_model_workspace_bytes(shape, 1));
        return MST_E_CAPACITY;
    }
    Arena ar{(char *)d_workspace, 0, false};
    plan_forward(c, ar);
    c.P = d_pitched; c.Ug = d_unpitched; c.F = d_features; c.mode = d_mode; c.bpm = d_bpm; c.prm = d_params; c.grad = d_grads;
    Launcher L; L.bf16 = shape->precision == MST_PRECISION_BF16; L.begin(stream);
    c.fused = true;
    forward_extract(L, c);
    forward_song_info(L, c);
    forward_apply(L, c, c.U != 0);
    TrainIn ti{d_instruments_target, d_bpm_target, normalize, d_losses};
    backward(L, c, ar, ti);
    L.join_all();
    if (L.err != cudaSuccess) { mst_set_error("mst_model_train_step: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    if (outputs) fill_outputs(c, outputs);
    if (n_launches) *n_launches = L.launches;
    return MST_OK;
}

extern "C" int mst_model_predict_apply(const float *d_params, const MstModelShape *shape, const float *d_style, const float *d_melody,
                                       const float *d_rhythm, const float *d_features, MstModelOutputs *outputs,
                                       void *d_workspace, size_t workspace_bytes, void *stream, int *n_launches) {
    if (!d_params || !shape || !d_style || !d_melody || !d_rhythm || !d_features || !outputs || !d_workspace) {
        mst_set_error("mst_model_predict_apply: null pointer argument");
        return MST_E_INVALID;
    }
    Ctx c;
    int rc = setup_ctx(c, shape, d_workspace, workspace_bytes, false, false);
    if (rc) return rc;
    c.F = d_features; c.prm = d_params;
    c.style = (float *)d_style; c.melody = (float *)d_melody; c.rhythm = d_rhythm;   // read only on this path
    Launcher L; L.bf16 = shape->precision == MST_PRECISION_BF16; L.begin(stream);
    prep_lstm(L, c, c.sim_l1, sim_bl_whh, sim_bl_bih, sim_bl_bhh);
    prep_lstm(L, c, c.sim_l2, sim_br_whh, sim_br_bih, sim_br_bhh);
    forward_song_info(L, c);
    forward_apply(L, c, c.U != 0);
    L.join_all();
    if (L.err != cudaSuccess) { mst_set_error("mst_model_predict_apply: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    fill_outputs(c, outputs);
    if (n_launches) *n_launches = L.launches;
    return MST_OK;
}

extern "C" int mst_model_song_info(const float *d_params, const MstModelShape *shape, const float *d_style, const float *d_rhythm,
                                   MstModelOutputs *outputs, void *d_workspace, size_t workspace_bytes, void *stream, int *n_launches) {
    if (!d_params || !shape || !d_style || !d_rhythm || !outputs || !d_workspace) {
        mst_set_error("mst_model_song_info: null pointer argument");
        return MST_E_INVALID;
    }
    Ctx c;
    int rc = setup_ctx(c, shape, d_workspace, workspace_bytes, false, false);
    if (rc) return rc;
    c.prm = d_params;
    c.style = (float *)d_style; c.rhythm = d_rhythm;                                   // read only on this path
    Launcher L; L.bf16 = shape->precision == MST_PRECISION_BF16; L.begin(stream);
    prep_lstm(L, c, c.sim_l1, sim_bl_whh, sim_bl_bih, sim_bl_bhh);
    prep_lstm(L, c, c.sim_l2, sim_br_whh, sim_br_bih, sim_br_bhh);
    forward_song_info(L, c);
    L.join_all();
    if (L.err != cudaSuccess) { mst_set_error("mst_model_song_info: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    memset(outputs, 0, sizeof(*outputs));
    outputs->instruments = c.sim_instr; outputs->mode = c.sim_mode; outputs->bpm = c.sim_bpm;
    outputs->style = c.style; outputs->rhythm = (float *)c.rhythm;
    if (n_launches) *n_launches = L.launches;
    return MST_OK;
}

extern "C" int mst_model_backward(const float *d_params, float *d_grads, const MstModelShape *shape, const float *d_pitched,
                                  const float *d_unpitched, const float *d_features, const float *d_mode, const float *d_bpm,
                                  const float *d_instruments_target, const float *d_bpm_target, int normalize,
                                  float *d_losses, void *d_workspace, size_t workspace_bytes, void *stream, int *n_launches) {
    if (!d_params || !d_grads || !shape || !d_pitched || !d_features || !d_mode || !d_bpm || !d_instruments_target ||
        !d_bpm_target || !d_losses || !d_workspace || (shape->has_unpitched && !d_unpitched)) {
        mst_set_error("mst_model_backward: null pointer argument");
        return MST_E_INVALID;
    }
    init_param_offsets();
    Ctx c;
    memset(&c, 0, sizeof(c));
    c.N = shape->n_samples; c.C = shape->n_channels; c.B = shape->n_bars; c.T = shape->n_beats; c.U = shape->has_unpitched ? 1 : 0;
    if (c.N < 1 || c.C < 1 || c.B < 1 || c.T < 1 || c.C > 64) { mst_set_error("mst_model: bad shape"); return MST_E_INVALID; }
    if (mst_model_workspace_bytes(shape, 1) > workspace_bytes) { mst_set_error("mst_model_backward: workspace too small"); return MST_E_CAPACITY; }
    Arena ar{(char *)d_workspace, 0, false};
    plan_forward(c, ar);
    c.P = d_pitched; c.Ug = d_unpitched; c.F = d_features; c.mode = d_mode; c.bpm = d_bpm; c.prm = d_params; c.grad = d_grads;
    c.bars = c.U ? c.bars_c : c.pce_bars; c.rhythm = c.U ? c.rh_c : c.pre_out;     // as forward_extract left them
    Launcher L; L.bf16 = shape->precision == MST_PRECISION_BF16; L.begin(stream);
    TrainIn ti{d_instruments_target, d_bpm_target, normalize, d_losses};
    backward(L, c, ar, ti);
    L.join_all();
    if (L.err != cudaSuccess) { mst_set_error("mst_model_backward: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    if (n_launches) *n_launches = L.launches;
    return MST_OK;
}

extern "C" size_t mst_model_loss_workspace_bytes(const MstModelShape *shape) {
    return (size_t)(shape ? shape->n_samples : 1) * 1024 + 4096;
}

extern "C" int mst_model_loss(const MstModelShape *shape, const MstLossIn *in, float *d_losses, void *d_workspace,
                              size_t workspace_bytes, void *stream) {
    if (!shape || !in || !d_losses || !d_workspace || !in->pitched_pred || !in->pitched_target || !in->instruments_logits ||
        !in->instruments_target || !in->mode_logits || !in->mode_target || !in->bpm_pred || !in->bpm_target) {
        mst_set_error("mst_model_loss: null pointer argument");
        return MST_E_INVALID;
    }
    if (workspace_bytes < mst_model_loss_workspace_bytes(shape)) { mst_set_error("mst_model_loss: workspace too small"); return MST_E_CAPACITY; }
    const int N = shape->n_samples, U = (in->unpitched_pred && in->unpitched_target) ? 1 : 0;
    Arena ar{(char *)d_workspace, 0, false};
    Launcher L; L.st = (cudaStream_t)stream;
    run_losses(L, ar, N, U, (long long)shape->n_channels * shape->n_bars * shape->n_beats * 560, (long long)shape->n_bars * shape->n_beats * 470,
               in->pitched_pred, in->pitched_target, in->unpitched_pred, in->unpitched_target, in->instruments_logits,
               in->instruments_target, in->mode_logits, in->mode_target, in->bpm_pred, in->bpm_target, in->normalize, d_losses, false);
    if (L.err != cudaSuccess) { mst_set_error("mst_model_loss: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    return MST_OK;
}

// ---- mst_model_loss without any allocation by anybody: the scratch of the loss kernels (per-sample partial sums) lives in a
// file-scope device array that every GPU loading the library holds until exit.  No counterpart in the reference; added
// for callers that cannot allocate (graph capture, serving loops).
constexpr size_t kLossNoAllocBytes = (size_t)MST_LOSS_NOALLOC_MAX_SAMPLES * 1024 + 4096;
static __device__ __align__(256) unsigned char g_loss_noalloc_scratch[kLossNoAllocBytes];

// The array's name is a host-side stand-in in host code: the device address comes from cudaGetSymbolAddress, asked once per
// device with that device current.
static int loss_noalloc_scratch(void **out) {
    static void *ptr[64] = {};
    static std::mutex mu;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { mst_set_error("mst_model_loss_noalloc: cudaGetDevice: %s", cudaGetErrorString(e)); return MST_E_CUDA; }
    if (dev < 0 || dev >= 64) { mst_set_error("mst_model_loss_noalloc: device ordinal %d out of range", dev); return MST_E_INVALID; }
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_loss_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            mst_set_error("mst_model_loss_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return MST_E_CUDA;
        }
        ptr[dev] = p;
    }
    *out = ptr[dev];
    return MST_OK;
}

extern "C" int mst_model_loss_noalloc(const MstModelShape *shape, const MstLossIn *in, float *d_losses, void *stream) {
    if (!shape || !in || !d_losses || !in->pitched_pred || !in->pitched_target || !in->instruments_logits ||
        !in->instruments_target || !in->mode_logits || !in->mode_target || !in->bpm_pred || !in->bpm_target) {
        mst_set_error("mst_model_loss_noalloc: null pointer argument");
        return MST_E_INVALID;
    }
    if (shape->n_samples < 1 || shape->n_samples > MST_LOSS_NOALLOC_MAX_SAMPLES) {
        mst_set_error("mst_model_loss_noalloc: n_samples = %d is outside 1..%d (MST_LOSS_NOALLOC_MAX_SAMPLES: the static scratch holds "
                      "%zu bytes); use mst_model_loss with a caller workspace for larger batches", shape->n_samples,
                      MST_LOSS_NOALLOC_MAX_SAMPLES, kLossNoAllocBytes);
        return MST_E_CAPACITY;
    }
    if (mst_device_count() <= 0) { mst_set_error("mst_model_loss_noalloc: no CUDA device"); return MST_E_NO_DEVICE; }
    void *scratch = nullptr;
    const int rc = loss_noalloc_scratch(&scratch);
    if (rc != MST_OK) return rc;
    const int N = shape->n_samples, U = (in->unpitched_pred && in->unpitched_target) ? 1 : 0;
    const long long cells_p = (long long)shape->n_channels * shape->n_bars * shape->n_beats * 560, cells_u = (long long)shape->n_bars * shape->n_beats * 470;
    {   // planning pass: what run_losses carves must fit the static array
        Arena plan{nullptr, 0, true};
        Launcher P; P.dry = true;
        run_losses(P, plan, N, U, cells_p, cells_u, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, false);
        if (plan.off > kLossNoAllocBytes) {
            mst_set_error("mst_model_loss_noalloc: internal: scratch carve %zu > %zu", plan.off, kLossNoAllocBytes);
            return MST_E_CAPACITY;
        }
    }
    Arena ar{(char *)scratch, 0, false};
    Launcher L; L.st = (cudaStream_t)stream;
    // run_losses clears the partial sums it accumulates into (the array keeps the previous call's contents)
    run_losses(L, ar, N, U, cells_p, cells_u,
               in->pitched_pred, in->pitched_target, in->unpitched_pred, in->unpitched_target, in->instruments_logits,
               in->instruments_target, in->mode_logits, in->mode_target, in->bpm_pred, in->bpm_target, in->normalize, d_losses, false);
    if (L.err != cudaSuccess) { mst_set_error("mst_model_loss_noalloc: %s", cudaGetErrorString(L.err)); return MST_E_CUDA; }
    return MST_OK;
}

extern "C" int mst_adam_step(float *d_params, float *d_grads, float *d_m, float *d_v, long long n, int step, float lr,
                             float beta1, float beta2, float eps, float grad_scale, int zero_grad, void *stream) {
    if (!d_params || !d_grads || !d_m || !d_v || n <= 0 || step < 1) { mst_set_error("mst_adam_step: bad argument"); return MST_E_INVALID; }
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    k_adam<<<grid_for(n, 256, 148 * 8), 256, 0, (cudaStream_t)stream>>>(d_params, d_grads, d_m, d_v, n, beta1, beta2, eps,
                                                                     (float)((double)lr / bc1), (float)(1.0 / sqrt(bc2)), grad_scale, zero_grad);
    MST_LAUNCH_CHECK("k_adam");
    return MST_OK;
}

extern "C" int mst_model_gemm(const float *d_A, long long a_si, long long a_sk, const float *d_B, long long b_sk, long long b_sj,
                              float *d_C, long long c_si, int M, int N, int K, const float *d_bias, int act, int accumulate,
                              int split_k, int kernel, void *stream) {
    if (!d_A || !d_B || !d_C || M <= 0 || N <= 0 || K <= 0 || split_k < 1 || kernel < 0 || kernel > 4) {
        mst_set_error("mst_model_gemm: bad argument");
        return MST_E_INVALID;
    }
    if (mst_device_count() <= 0) { mst_set_error("mst_model_gemm: no CUDA device"); return MST_E_NO_DEVICE; }
    Launcher L;
    L.st = (cudaStream_t)stream;
    L.gemm(d_A, a_si, a_sk, d_B, b_sk, b_sj, d_C, c_si, M, N, K, d_bias, act, accumulate, split_k, kernel);
    if (L.err != cudaSuccess) { mst_set_error(cudaGetErrorString(L.err)); return MST_E_CUDA; }
    return MST_OK;
}
