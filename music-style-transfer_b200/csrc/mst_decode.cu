// mst_decode.cu — piano-roll grids -> ordered MIDI note messages on sm_100a (B200).
//
// Replaces, for a whole batch of songs per call, the reference's
//   hard_output        style/model.py:818-832          (velocity gate > f32(.01), accidental arg-max gate > f32(.1))
//   vchannel2qchannel  style/midi_conversion.py:518-562 (nonzero-velocity cells in (bar, beat, fraction, row) order)
//   qchannel2channel   style/midi_conversion.py:458-488 (scale location -> pitch, location -> tick, on/off pairs)
//   decode_midi        style/style_transfer.py:145-158
//   create_midi        style/midi.py:120-168            (int(velocity*127), stable sort by time, delta + 1 s clip)
//
// Pipeline (the grid is read from HBM exactly once):
//   k_dec_plan     one CTA: chunk prefix over the songs (a chunk = one bar of one channel)
//   k_dec_scan     warp per chunk: streaming read, velocity gate, ballot/popc ordered compaction of the hits
//                  into a 32-record slot per chunk (+ optional write-back of the gate like the in-place `*=`)
//   k_dec_offsets  CTA per song: scan of the chunk counts -> note offsets, song totals
//   k_dec_songs    one CTA: message offsets over the songs, capacity check
//   k_dec_emit     warp per chunk: pitch / tick / duration / velocity of every hit -> (time key, packed message)
//   k_dec_order    CTA per song: stable radix sort by time, delta + clip, end_of_track
#include <string.h>

#include <algorithm>

#include "mst_device.cuh"
#include "mst_sort.cuh"

namespace {

constexpr int kSlotCap = 32;                 // hit records kept per chunk; denser chunks are re-read by k_dec_emit
constexpr uint32_t kTimeBias = 1u << 30;     // sort keys = time + bias (times may be negative: negative durations)
constexpr int kScanWarps = 8;

struct __align__(16) Hit {
    uint32_t cell;       // (beat * 10 + fraction) * rows + row inside the chunk
    float dur;
    float vel;
    uint32_t acc;        // 0 flat, 1 none, 2 sharp
};

struct DecWorkspace {
    uint32_t *chunk_prefix;   // [S+1] first chunk of each song
    uint32_t *chunk_count;    // [n_chunks] hits
    uint32_t *chunk_off;      // [n_chunks] song-local note offset
    Hit *slots;               // [n_chunks][kSlotCap]
    uint32_t *song_notes;     // [S]
    int32_t *song_status;     // [S]
    uint32_t *keys_a, *vals_a, *keys_b, *vals_b;   // [msg_capacity]
    uint32_t *packed;         // [msg_capacity] is_on | channel << 8 | note << 16 | velocity << 24
};

__constant__ int8_t d_abs_intervals[2][7] = {{0, 2, 4, 5, 7, 9, 11}, {0, 2, 3, 5, 7, 8, 10}};   // scales.py:49-51
__constant__ int8_t d_frac_num[10] = {0, 1, 1, 1, 3, 1, 5, 2, 3, 7};    // sorted fractions, midi_conversion.py:358-362
__constant__ int8_t d_frac_den[10] = {1, 8, 4, 3, 8, 2, 8, 3, 4, 8};

// a song table entry the kernels can index safely (the Python wrapper guards this too; a C caller may not)
__device__ __forceinline__ bool song_valid(const MstDecodeSong &d) {
    return d.n_pitched >= 0 && d.n_pitched <= MST_MAX_DECODE_CHANNELS && d.n_bars > 0 && d.n_beats > 0 &&
           d.ticks_per_beat > 0 && d.tempo > 0 && d.mode >= 0 && d.mode <= 1;
}
__device__ __forceinline__ uint32_t song_chunks(const MstDecodeSong &d) {
    if (!song_valid(d)) return 0u;          // flagged MST_DECODE_BAD_VALUE by k_dec_plan, decodes nothing
    return (uint32_t)(d.n_pitched + (d.has_unpitched ? 1 : 0)) * (uint32_t)d.n_bars;
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1)
k_dec_plan(const MstDecodeSong *__restrict__ songs, uint32_t n_songs, DecWorkspace ws, uint32_t n_chunks) {
    __shared__ uint32_t tmp[33];
    uint32_t running = 0;
    for (uint32_t c0 = 0; c0 < n_songs; c0 += blockDim.x) {
        const uint32_t s = c0 + threadIdx.x;
        const uint32_t v = s < n_songs ? song_chunks(songs[s]) : 0u;
        uint32_t tot;
        const uint32_t ex = block_exclusive_scan(v, tmp, &tot);
        if (s < n_songs) {
            ws.chunk_prefix[s] = running + ex;
            ws.song_status[s] = song_valid(songs[s]) ? MST_DECODE_OK : MST_DECODE_BAD_VALUE;
        }
        running += tot;
    }
    __syncthreads();
    if (running > n_chunks) {
        // the song table needs more chunks than the caller's count (which sized the workspace): decode nothing and say so
        for (uint32_t s = threadIdx.x; s <= n_songs; s += blockDim.x) {
            ws.chunk_prefix[s] = 0;
            if (s < n_songs) ws.song_status[s] = MST_DECODE_CAPACITY;
        }
    } else if (threadIdx.x == 0) ws.chunk_prefix[n_songs] = running;
}

// chunk -> (song, channel, bar) and the chunk's float base / geometry
struct ChunkGeom {
    uint32_t song, channel, bar;
    bool pitched;
    uint32_t rows, feats, cells;   // cells = n_beats * 10 * rows
    uint64_t base;                 // float offset of the chunk in the grid buffer
};

__device__ __forceinline__ ChunkGeom chunk_geom(const MstDecodeSong *songs, uint32_t n_songs, const uint32_t *prefix,
                                                uint32_t chunk) {
    uint32_t lo = 0, hi = n_songs;
    while (hi - lo > 1) { uint32_t mid = (lo + hi) >> 1; if (prefix[mid] <= chunk) lo = mid; else hi = mid; }
    ChunkGeom g;
    g.song = lo;
    const MstDecodeSong &d = songs[lo];
    const uint32_t local = chunk - prefix[lo];
    g.channel = local / (uint32_t)d.n_bars;
    g.bar = local % (uint32_t)d.n_bars;
    g.pitched = g.channel < (uint32_t)d.n_pitched;
    g.rows = g.pitched ? MST_N_PITCHED_ROWS : MST_N_UNPITCHED_ROWS;
    g.feats = g.pitched ? MST_N_PITCHED_FEATS : MST_N_UNPITCHED_FEATS;
    g.cells = (uint32_t)d.n_beats * MST_N_FRACTIONS * g.rows;
    const uint64_t bar_floats = (uint64_t)g.cells * g.feats;
    g.base = g.pitched ? d.pitched_offset + ((uint64_t)g.channel * d.n_bars + g.bar) * bar_floats
                       : d.unpitched_offset + (uint64_t)g.bar * bar_floats;
    return g;
}

// hard_output's accidental gate (model.py:825-828) + vchannel2qchannel's precedence (:532-539)
__device__ __forceinline__ uint32_t pick_accidental(float a0, float a1, float a2) {
    float mx = fmaxf(a0, fmaxf(a1, a2));
    if (a0 != a0 || a1 != a1 || a2 != a2) mx = __int_as_float(0x7fc00000);   // torch.max propagates NaN
    const float thr = 0.1f;
    if (a0 == mx && a0 > thr) return 0;
    if (a1 == mx && a1 > thr) return 1;
    if (a2 == mx && a2 > thr) return 2;
    return 1;
}

// ------------------------------------------------------------------------------------------------
// k_dec_scan: the only pass over the grid.  One warp per chunk walks the cells in memory order
// (= bar, beat, fraction, row order), 4 strips of 32 cells per iteration for memory-level parallelism.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kScanWarps * 32)
k_dec_scan(float *__restrict__ grids, const MstDecodeSong *__restrict__ songs, uint32_t n_songs, DecWorkspace ws,
           uint32_t n_chunks) {
    const unsigned lane = lane_id();
    const float thr_v = 0.01f;
    n_chunks = min(n_chunks, ws.chunk_prefix[n_songs]);       // 0 when k_dec_plan rejected the table
    for (uint32_t chunk = blockIdx.x * kScanWarps + (threadIdx.x >> 5); chunk < n_chunks; chunk += gridDim.x * kScanWarps) {
        const ChunkGeom g = chunk_geom(songs, n_songs, ws.chunk_prefix, chunk);
        const bool write_back = songs[g.song].write_back != 0;
        float *p = grids + g.base;
        Hit *slot = ws.slots + (size_t)chunk * kSlotCap;
        uint32_t count = 0;
        for (uint32_t b = 0; b < g.cells; b += 128) {
            float vel[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t i = b + u * 32 + lane;
                vel[u] = i < g.cells ? p[(size_t)i * g.feats + 1] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const uint32_t i = b + u * 32 + lane;
                const float v = vel[u];
                const float gated = v * ((v > thr_v) ? 1.0f : 0.0f);    // velocity *= (velocity > .01).float()
                const bool hit = (i < g.cells) && !(gated == 0.0f);     // np.nonzero: NaN counts
                if (write_back && i < g.cells && !(v == gated) ) p[(size_t)i * g.feats + 1] = gated;
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (hit) {
                    const uint32_t k = count + __popc(m & lanemask_lt());
                    if (k < kSlotCap) {
                        Hit h;
                        h.cell = i; h.vel = gated; h.dur = p[(size_t)i * g.feats];
                        h.acc = g.pitched ? pick_accidental(p[(size_t)i * 5 + 2], p[(size_t)i * 5 + 3], p[(size_t)i * 5 + 4]) : 1u;
                        slot[k] = h;
                    }
                }
                count += __popc(m);
            }
        }
        if (lane == 0) ws.chunk_count[chunk] = count;
    }
}

// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_dec_offsets(DecWorkspace ws) {
    __shared__ uint32_t tmp[33];
    const uint32_t song = blockIdx.x;
    const uint32_t c0 = ws.chunk_prefix[song], c1 = ws.chunk_prefix[song + 1];
    uint32_t running = 0;
    for (uint32_t b = c0; b < c1; b += blockDim.x) {
        const uint32_t i = b + threadIdx.x;
        const uint32_t v = i < c1 ? ws.chunk_count[i] : 0u;
        uint32_t tot;
        const uint32_t ex = block_exclusive_scan(v, tmp, &tot);
        if (i < c1) ws.chunk_off[i] = running + ex;
        running += tot;
    }
    if (threadIdx.x == 0) ws.song_notes[song] = running;
}

__global__ void __launch_bounds__(1024, 1)
k_dec_songs(const MstDecodeSong *__restrict__ songs, uint32_t n_songs, DecWorkspace ws, MstDecodeResult *results,
            uint64_t msg_capacity) {
    __shared__ unsigned long long w64[33];
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
    unsigned long long running = 0;
    for (uint32_t c0 = 0; c0 < n_songs; c0 += blockDim.x) {
        const uint32_t s = c0 + threadIdx.x;
        const unsigned long long v = s < n_songs ? 2ull * ws.song_notes[s] : 0ull;
        unsigned long long inc = v;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= (unsigned)o) inc += t;
        }
        if (lane == 31) w64[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            unsigned long long x = w64[lane], xi = x;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long t = __shfl_up_sync(0xffffffffu, xi, o);
                if (lane >= (unsigned)o) xi += t;
            }
            w64[lane] = xi - x;
            if (lane == 31) w64[32] = xi;
        }
        __syncthreads();
        const unsigned long long off = running + w64[warp] + inc - v;
        const unsigned long long tot = w64[32];
        __syncthreads();
        if (s < n_songs) {
            MstDecodeResult r;
            r.n_messages = (uint32_t)v;
            r.msg_offset = off;
            r.eot_delta = 0;
            // mido.second2tick(1, tpb, tempo) = int(round(1 / (tempo * 1e-6 / tpb)))   (midi.py:121-123)
            const double scale = __ddiv_rn(__dmul_rn((double)songs[s].tempo, 1e-6), (double)songs[s].ticks_per_beat);
            r.max_delta = (int32_t)rint(__ddiv_rn(1.0, scale));
            r.status = v == 0 ? MST_DECODE_NO_MESSAGES : (off + v > msg_capacity ? MST_DECODE_CAPACITY : MST_DECODE_OK);
            if (ws.song_status[s] != MST_DECODE_OK) r.status = ws.song_status[s];     // rejected by k_dec_plan
            results[s] = r;
        }
        running += tot;
    }
}

// one hit -> note_on + note_off (qchannel2channel :458-488, create_midi's int(velocity*127) midi.py:147)
__device__ __forceinline__ void emit_hit(const MstDecodeSong &d, const ChunkGeom &g, const Hit &h, uint32_t local_idx,
                                         uint64_t msg_base, DecWorkspace &ws) {
    const uint32_t row = h.cell % g.rows;
    const uint32_t bf = h.cell / g.rows;
    const uint32_t frac = bf % MST_N_FRACTIONS, beat = bf / MST_N_FRACTIONS;
    int note, channel;
    if (g.pitched) {
        const int degree = row % 7 + 1;
        int octave = row / 7;
        int interval = d_abs_intervals[d.mode][degree - 1] + d.tonic;          // scale_loc2key_octave :269-283
        if (h.acc == 2) interval += 1; else if (h.acc == 0) interval -= 1;
        if (interval < 0) { octave -= 1; interval += 12; }
        else if (interval >= 12) { octave += 1; interval -= 12; }
        note = 12 * (octave + 1) + interval;                                   // note2note_id :320-324
        channel = d.channel_id[g.channel];
    } else {
        note = (int)row + 35;                                                  // :549
        channel = d.unpitched_channel_id;
    }
    int err = 0;
    const float prod = __fmul_rn(h.dur, (float)d.ticks_per_beat);             // np.float32 * int stays f32 (:558)
    if (!(prod == prod) || prod > 1.0e9f || prod < -1.0e9f) err = MST_DECODE_BAD_VALUE;
    const int qdur = err ? 0 : (int)prod;                                      // int(): truncation
    const long long t = (long long)g.bar * d.ticks_per_bar + (long long)beat * d.ticks_per_beat +
                        ((long long)d_frac_num[frac] * d.ticks_per_beat) / d_frac_den[frac];   // loc2ticks :459-464
    if (t > (1ll << 30)) err = MST_DECODE_BAD_VALUE;
    const float vf = __fmul_rn(h.vel, 127.0f);
    int vi = 0;
    if (!(vf == vf)) err = MST_DECODE_BAD_VALUE; else vi = (int)vf;
    if (!err && vi > 127) err = MST_DECODE_VELOCITY;                           // the assert at midi.py:148
    if (err) { atomicMax(&ws.song_status[g.song], err); return; }
    const uint64_t j = msg_base + 2ull * local_idx;
    ws.keys_a[j] = (uint32_t)((int)t + (int)kTimeBias);
    ws.vals_a[j] = 2u * local_idx;
    ws.packed[j] = 1u | ((uint32_t)channel << 8) | ((uint32_t)note << 16) | ((uint32_t)vi << 24);
    ws.keys_a[j + 1] = (uint32_t)((int)t + qdur + (int)kTimeBias);
    ws.vals_a[j + 1] = 2u * local_idx + 1u;
    ws.packed[j + 1] = 0u | ((uint32_t)channel << 8) | ((uint32_t)note << 16);
}

__global__ void __launch_bounds__(kScanWarps * 32)
k_dec_emit(const float *__restrict__ grids, const MstDecodeSong *__restrict__ songs, uint32_t n_songs, DecWorkspace ws,
           const MstDecodeResult *__restrict__ results, uint32_t n_chunks) {
    const unsigned lane = lane_id();
    const float thr_v = 0.01f;
    n_chunks = min(n_chunks, ws.chunk_prefix[n_songs]);       // 0 when k_dec_plan rejected the table
    for (uint32_t chunk = blockIdx.x * kScanWarps + (threadIdx.x >> 5); chunk < n_chunks; chunk += gridDim.x * kScanWarps) {
        const uint32_t count = ws.chunk_count[chunk];
        if (count == 0) continue;
        const ChunkGeom g = chunk_geom(songs, n_songs, ws.chunk_prefix, chunk);
        if (results[g.song].status != MST_DECODE_OK) continue;
        const MstDecodeSong &d = songs[g.song];
        const uint64_t msg_base = results[g.song].msg_offset;
        const uint32_t first = ws.chunk_off[chunk];
        if (count <= kSlotCap) {
            if (lane < count) emit_hit(d, g, ws.slots[(size_t)chunk * kSlotCap + lane], first + lane, msg_base, ws);
            continue;
        }
        // dense chunk: walk it again (re-gating an already written-back velocity is idempotent)
        const float *p = grids + g.base;
        uint32_t k0 = 0;
        for (uint32_t b = 0; b < g.cells; b += 32) {
            const uint32_t i = b + lane;
            const float v = i < g.cells ? p[(size_t)i * g.feats + 1] : 0.f;
            const float gated = v * ((v > thr_v) ? 1.0f : 0.0f);
            const bool hit = (i < g.cells) && !(gated == 0.0f);
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (hit) {
                Hit h;
                h.cell = i; h.vel = gated; h.dur = p[(size_t)i * g.feats];
                h.acc = g.pitched ? pick_accidental(p[(size_t)i * 5 + 2], p[(size_t)i * 5 + 3], p[(size_t)i * 5 + 4]) : 1u;
                emit_hit(d, g, h, first + k0 + __popc(m & lanemask_lt()), msg_base, ws);
            }
            k0 += __popc(m);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_dec_order: sorted(messages, key=time) per channel (:487) then over all channels (midi.py:157) == one
// stable sort by time of the channel-major emission order; then the delta / clip loop of midi.py:161-167.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads, 1)
k_dec_order(const MstDecodeSong *__restrict__ songs, DecWorkspace ws, MstDecodeResult *results, MstOutMsg *out) {
    __shared__ uint32_t whist[kSortWarps][256];
    __shared__ uint32_t scan_tmp[33];
    __shared__ int skip_flag;
    const uint32_t song = blockIdx.x;
    MstDecodeResult r = results[song];
    if (r.status != MST_DECODE_OK) return;
    const int err = ws.song_status[song];
    if (err != MST_DECODE_OK) {
        if (threadIdx.x == 0) results[song].status = err;
        return;
    }
    const uint32_t n = r.n_messages;
    uint32_t *ka = ws.keys_a + r.msg_offset, *va = ws.vals_a + r.msg_offset;
    uint32_t *kb = ws.keys_b + r.msg_offset, *vb = ws.vals_b + r.msg_offset;
    radix_pass(ka, va, kb, vb, n, 0, whist, scan_tmp, &skip_flag);
    radix_pass(kb, vb, ka, va, n, 8, whist, scan_tmp, &skip_flag);
    radix_pass(ka, va, kb, vb, n, 16, whist, scan_tmp, &skip_flag);
    radix_pass(kb, vb, ka, va, n, 24, whist, scan_tmp, &skip_flag);
    const uint32_t *packed = ws.packed + r.msg_offset;
    MstOutMsg *o = out + r.msg_offset;
    for (uint32_t i = threadIdx.x; i < n; i += kSortThreads) {
        const int t = (int)(ka[i] - kTimeBias);
        const int prev = i ? (int)(ka[i - 1] - kTimeBias) : 0;
        int delta = min(t - prev, r.max_delta);
        delta = max(0, delta);
        const uint32_t pk = packed[va[i]];
        MstOutMsg m;
        m.time = t; m.delta = delta; m.is_on = pk & 1u; m.channel = (pk >> 8) & 255u;
        m.note = (pk >> 16) & 255u; m.velocity = pk >> 24;
        o[i] = m;
    }
    if (threadIdx.x == 0) {
        const int last = (int)(ka[n - 1] - kTimeBias);
        const int end = songs[song].duration >= 0 ? songs[song].duration : last + songs[song].ticks_per_bar;   // midi.py:158
        results[song].eot_delta = max(0, min(end - last, r.max_delta));
    }
}

}  // namespace

// ================================================================================================
// host side
// ================================================================================================
static size_t dec_carve(DecWorkspace *ws, void *base, uint32_t S, uint64_t n_chunks, uint64_t cap) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = mst_align_up(off + bytes, 256); return base ? (char *)base + o : (char *)nullptr; };
    const uint64_t nc = n_chunks ? n_chunks : 1, mc = cap ? cap : 1;
    ws->chunk_prefix = (uint32_t *)take(((size_t)S + 1) * 4);
    ws->chunk_count = (uint32_t *)take(nc * 4);
    ws->chunk_off = (uint32_t *)take(nc * 4);
    ws->slots = (Hit *)take(nc * kSlotCap * sizeof(Hit));
    ws->song_notes = (uint32_t *)take((size_t)S * 4);
    ws->song_status = (int32_t *)take((size_t)S * 4);
    ws->keys_a = (uint32_t *)take(mc * 4); ws->vals_a = (uint32_t *)take(mc * 4);
    ws->keys_b = (uint32_t *)take(mc * 4); ws->vals_b = (uint32_t *)take(mc * 4);
    ws->packed = (uint32_t *)take(mc * 4);
    return off;
}

extern "C" size_t mst_decode_workspace_bytes(uint32_t n_songs, uint64_t n_chunks, uint64_t msg_capacity) {
    DecWorkspace ws;
    return dec_carve(&ws, nullptr, n_songs ? n_songs : 1, n_chunks, msg_capacity);
}

extern "C" int mst_decode_batch_device(float *d_grids, const MstDecodeSong *d_songs, uint32_t n_songs, uint64_t n_chunks,
                                       MstOutMsg *d_msgs, uint64_t msg_capacity, MstDecodeResult *d_results,
                                       void *d_workspace, size_t workspace_bytes, void *stream_) {
    if (!d_grids || !d_songs || !d_results || !d_workspace || (!d_msgs && msg_capacity)) {
        mst_set_error("mst_decode_batch_device: null pointer argument");
        return MST_E_INVALID;
    }
    if (n_songs == 0) return MST_OK;
    if (n_chunks >= 0xFFFFFFFFull || msg_capacity >= 0xFFFFFFFFull) {
        mst_set_error("mst_decode_batch_device: at most 2^32-1 chunks / messages per batch");
        return MST_E_INVALID;
    }
    DecWorkspace ws;
    const size_t need = dec_carve(&ws, d_workspace, n_songs, n_chunks, msg_capacity);
    if (need > workspace_bytes) {
        mst_set_error("mst_decode_batch_device: workspace too small (%zu < %zu)", workspace_bytes, need);
        return MST_E_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const uint32_t nch = (uint32_t)n_chunks;
    const uint32_t chunk_blocks = (uint32_t)std::max<uint64_t>(1, std::min<uint64_t>((n_chunks + kScanWarps - 1) / kScanWarps, (uint64_t)sms * 8));
    k_dec_plan<<<1, 1024, 0, stream>>>(d_songs, n_songs, ws, nch);
    MST_LAUNCH_CHECK("k_dec_plan");
    k_dec_scan<<<chunk_blocks, kScanWarps * 32, 0, stream>>>(d_grids, d_songs, n_songs, ws, nch);
    MST_LAUNCH_CHECK("k_dec_scan");
    k_dec_offsets<<<n_songs, 256, 0, stream>>>(ws);
    MST_LAUNCH_CHECK("k_dec_offsets");
    k_dec_songs<<<1, 1024, 0, stream>>>(d_songs, n_songs, ws, d_results, msg_capacity);
    MST_LAUNCH_CHECK("k_dec_songs");
    k_dec_emit<<<chunk_blocks, kScanWarps * 32, 0, stream>>>(d_grids, d_songs, n_songs, ws, d_results, nch);
    MST_LAUNCH_CHECK("k_dec_emit");
    k_dec_order<<<n_songs, kSortThreads, 0, stream>>>(d_songs, ws, d_results, d_msgs);
    MST_LAUNCH_CHECK("k_dec_order");
    return MST_OK;
}
