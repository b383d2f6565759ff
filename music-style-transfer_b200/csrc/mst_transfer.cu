// mst_transfer.cu — the host decisions of style_transfer.apply_style (style/style_transfer.py:101-122) made on the device,
// for a batch of styles: which instruments to render (top-k of the 41 predicted logits, with the percussion special case),
// the mode (arg-max of the two logits), the tempo (bpm2tempo(round(bpm))), the instrument one-hot features apply_style
// feeds the appliers (style/data.py:122-127) and the decode table entry of each output song.  With this kernel between
// mst_model_predict_apply (song-info heads) and mst_model_predict_apply / mst_decode_batch_device, a whole
// "composition x S styles" transfer runs without a host round trip until the finished messages are read back.
#include <math.h>

#include "mst_device.cuh"

namespace {

constexpr int kLogits = 41;        // 40 included programs + percussion (style/data.py:21,29)
constexpr int kPercussion = 40;    // percussion_id (style/data.py:31)
constexpr int kFeat = 51;          // instrument_size: 40 program one-hot + 11 group one-hot (style/data.py:28)

struct SelectTables {
    int16_t program[40];           // sorted included programs (OneHotEncoder categories, style/data.py:23-25)
    int16_t group[40];             // index of the program's group among the 11 sorted group names (:26-27)
};

// One warp per style.  Rank of logit i in np.argsort(-logits): number of logits that sort before it (ties by index).
__global__ void __launch_bounds__(32)
k_transfer_select(const float *__restrict__ instr, const float *__restrict__ mode_logits, const float *__restrict__ bpm,
                  const int32_t *__restrict__ n_instruments, SelectTables tab, int max_channels,
                  float *__restrict__ features, MstDecodeSong *__restrict__ songs, MstTransferChoice *__restrict__ choice) {
    const int s = blockIdx.x, lane = threadIdx.x;
    __shared__ float key[kLogits];
    __shared__ int order[kLogits];
    const float *x = instr + (size_t)s * kLogits;
    for (int i = lane; i < kLogits; i += 32) key[i] = -x[i];
    __syncwarp();
    for (int i = lane; i < kLogits; i += 32) {
        int rank = 0;
        const float ki = key[i];
        for (int j = 0; j < kLogits; j++) rank += (key[j] < ki || (key[j] == ki && j < i)) ? 1 : 0;
        order[rank] = i;
    }
    for (int i = lane; i < max_channels * kFeat; i += 32) features[(size_t)s * max_channels * kFeat + i] = 0.f;
    __syncwarp();
    if (lane == 0) {
        int n = n_instruments[s];
        n = max(0, min(n, kLogits));
        if (n == 1 && order[0] == kPercussion) n = 2;                 // :107-108: a lone percussion pick takes one more
        MstTransferChoice c;
        c.n_pitched = 0; c.has_unpitched = 0;
        for (int k = 0; k < MST_MAX_DECODE_CHANNELS; k++) c.instruments[k] = -1;
        for (int k = 0; k < n; k++) {
            const int id = order[k];
            if (id == kPercussion) { c.has_unpitched = 1; continue; }  // :109-110
            if (c.n_pitched < MST_MAX_DECODE_CHANNELS) c.instruments[c.n_pitched] = tab.program[id];
            if (c.n_pitched < max_channels) {                          // encode_instruments (style/data.py:122-127)
                float *f = features + ((size_t)s * max_channels + c.n_pitched) * kFeat;
                f[id] = 1.f;
                f[40 + tab.group[id]] = 1.f;
            }
            c.n_pitched++;
        }
        c.mode = mode_logits[2 * s] >= mode_logits[2 * s + 1] ? 0 : 1;    // mode[0].argmax() == 0 -> major (:118-121; first max wins)
        // info['tempo'] = bpm2tempo(round(float(bpm)))  (:103): Python round() = half to even, bpm2tempo = int(round(6e7 / bpm))
        const double b = rint((double)bpm[s]);
        c.tempo = (int32_t)rint(6e7 / b);
        choice[s] = c;
        MstDecodeSong d = songs[s];
        // save() (:78-86): channels 0..15 without 9, zipped with the instruments and the grid's channels
        const int rendered = min(c.n_pitched, max_channels);
        d.n_pitched = min(rendered, 15);
        d.has_unpitched = c.has_unpitched;
        d.mode = c.mode;
        d.tempo = c.tempo;
        int ch = 0;
        for (int k = 0; k < MST_MAX_DECODE_CHANNELS; k++) {
            if (ch == 9) ch++;
            d.channel_id[k] = (uint8_t)min(ch, 15);
            d.instrument_id[k] = k < c.n_pitched ? c.instruments[k] : 0;
            ch++;
        }
        songs[s] = d;
    }
}

}  // namespace

extern "C" int mst_transfer_select(const float *d_instruments_logits, const float *d_mode_logits, const float *d_bpm,
                                   const int32_t *d_n_instruments, uint32_t n_styles, const int16_t *h_programs,
                                   const int16_t *h_groups, int32_t max_channels, float *d_features, MstDecodeSong *d_songs,
                                   MstTransferChoice *d_choice, void *stream) {
    if (!d_instruments_logits || !d_mode_logits || !d_bpm || !d_n_instruments || !h_programs || !h_groups || !d_features ||
        !d_songs || !d_choice || max_channels < 1) {
        mst_set_error("mst_transfer_select: bad argument");
        return MST_E_INVALID;
    }
    if (mst_device_count() == 0) { mst_set_error("no CUDA device: libmst_b200 has no CPU fallback"); return MST_E_NO_DEVICE; }
    if (n_styles == 0) return MST_OK;
    SelectTables tab;
    for (int i = 0; i < 40; i++) { tab.program[i] = h_programs[i]; tab.group[i] = h_groups[i]; }
    k_transfer_select<<<n_styles, 32, 0, (cudaStream_t)stream>>>(d_instruments_logits, d_mode_logits, d_bpm, d_n_instruments, tab,
                                                               max_channels, d_features, d_songs, d_choice);
    MST_LAUNCH_CHECK("k_transfer_select");
    return MST_OK;
}
