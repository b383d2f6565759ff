// mst_encode.cu — MIDI events -> piano-roll grids on sm_100a (B200).
//
// Replaces, for a whole batch of songs per call, the reference's
//   read_midi            style/midi_conversion.py:216-232  (merge_tracks :37, split_channels :55,
//                                                            get_midi_info :117, group_channel_messages :182)
//   iter_inputs' filter  style/data.py:53-58
//   get_input            style/data.py:66-100  (channel2nchannel :371, merge_nchannels data.py:103,
//                                               get_keys_dist :340, get_scale scales.py:214,
//                                               nchannel2vchannel :564 = :408 + :425 + :490)
//   prepare_input        style/data.py:130-143 (bar crop + f32 cast)
//
// Pipeline (one launch each, everything stays in HBM):
//   k_sort      per-song CTA: stable partition by class + sortedness check (full stable LSD radix sort of
//               (class, tick) only for songs whose tracks share a channel) -> merged order, grouped per channel
//   k_walk<0>   warp per (song, channel): program / CC7 state scan, per-instrument message counts
//   k_info      warp per song: song info (tempo histogram, signatures, bars), channel-list table, filter,
//               output slots and sizes
//   k_offsets   one CTA: prefix sums of grid floats over the batch, arena check, end of the written range
//   k_fill_flat zero fill of the batch's arena range at the HBM write bandwidth (caller's stream)   } concurrent
//   k_walk<1>   same scan again, now writing NoteMessage records into per-list storage              } (high-priority
//   k_pair      warp per list: note pairing (match within a 32-message strip, open-note table across) } side stream)
//   k_scale     per-song CTA: key histogram in reference summation order + 24-candidate scale choice  }
//   k_cells<1>  per-song CTA: f64 dual-grid quantisation, scale-aware rows, note scatter (global atomicMax)
// Comparison path (MST_GRID_TILES=1): k_cells<0> bins the cells per (slot, bar) and k_grid (persistent CTAs) builds
// (slot, bar) tiles in shared memory and writes them once with TMA bulk stores (cp.async.bulk shared -> global).
#include <math_constants.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>

#include "mst_device.cuh"
#include "mst_sort.cuh"

namespace {

constexpr uint32_t kClassGlobal = 16;     // messages without a channel attribute that matter
constexpr uint32_t kClassPrefixHi = 17;   // channel_prefix with channel > 15 (only counts for `duration`)
constexpr uint32_t kClassEnd = 18;        // first class that is not walked
constexpr uint32_t kClassIgnored = 30;
constexpr uint32_t kClassDropped = 31;    // tick > MST_MAX_TICK (midi_conversion.py:52,59)
constexpr int kSegStride = 20;
constexpr int kMaxLists = 16 * MST_MAX_LISTS_PER_CHANNEL;
constexpr int kTileBeats = 4;
constexpr uint32_t kNoOpen = 0xFFFFFFFFu;

struct __align__(16) NoteMsg {   // NoteMessage, midi_conversion.py:101-106
    uint32_t time;
    uint8_t is_on;
    uint8_t note;
    uint16_t list;
    double velocity;
};
static_assert(sizeof(NoteMsg) == 16, "NoteMsg layout");

struct __align__(8) Note {       // Note, midi_conversion.py:286-306 (fields the encoder needs)
    int32_t note_id;
    int32_t time;
    int32_t end_time;
    int32_t pad;
    double velocity;
};
static_assert(sizeof(Note) == 24, "Note layout");

struct __align__(16) Cell {      // one quantised note, addressed inside its (slot, bar) bin
    uint32_t bin;                // global bin id
    uint32_t off_acc;            // float offset inside the bar slab | accidental << 28
    float dur;
    float vel;
};

struct ListRec {                 // one read_midi channel entry (midi_conversion.py:225-230)
    int16_t channel;
    int16_t instrument;
    int16_t out_slot;            // pitched: 0..C-1, unpitched: C..C+U-1, dropped by the filter: -1
    int16_t pitched;
    uint32_t msg_offset;         // song-local
    uint32_t n_messages;
    uint32_t n_notes;
    uint32_t note_offset;        // song-local
};

struct ChanRec {
    int32_t n_slots;
    int32_t error;
    int32_t first_on;
    int32_t last_on;
    int16_t slot_instr[MST_MAX_LISTS_PER_CHANNEL];
    uint32_t slot_msgs[MST_MAX_LISTS_PER_CHANNEL];
    uint32_t slot_ons[MST_MAX_LISTS_PER_CHANNEL];
};

struct __align__(16) TileDesc {   // one k_grid work item: up to kTileBeats beats of one (slot, bar)
    uint64_t dst;            // float offset in the arena
    uint32_t floats;
    uint32_t cell_lo, cell_hi;     // range in cells_sorted (absolute)
    uint32_t beat_lo_off;          // float offset of the tile's first beat inside the bar slab
    uint32_t pad[2];
};
static_assert(sizeof(TileDesc) == 32, "TileDesc layout");

struct SongPlan {
    uint64_t out_floats;   // requested grid floats (pitched + unpitched)
    uint32_t n_bins;       // (C+U) * n_bars_out
    uint32_t n_tiles;      // n_bins * ceil(n_beats / kTileBeats)
    uint32_t bin_base;
    uint32_t tile_base;
    uint32_t n_lists;
    uint32_t fits;
};

struct Workspace {
    uint32_t *keys_a, *vals_a, *keys_b, *vals_b, *tmp_c;
    MstEvent *sorted_ev;
    NoteMsg *msgs;
    Note *notes;
    Cell *cells, *cells_sorted;
    uint32_t *seg_start;      // [S][kSegStride]
    ChanRec *chans;           // [S][16]
    ListRec *lists;           // [S][kMaxLists]
    int16_t *chan_slot_list;  // [S][16][16]
    SongPlan *plans;          // [S]
    uint32_t *tile_prefix;    // [S+1]
    uint32_t *bin_count, *bin_start, *bin_fill;  // [max_bins + 1]
    uint32_t *totals;         // [4]: total tiles, total bins
    TileDesc *tiles;          // [max_bins + 1] (a tile never covers more than one bin)
    uint64_t max_bins;
};

// songs whose info is complete: OK, or planned but not written because the arena is full
__device__ __forceinline__ bool song_active(int status) { return status == MST_SONG_OK || status == MST_SONG_ARENA_FULL; }

__device__ __forceinline__ uint32_t event_class(const MstEvent &e) {
    if (e.tick > MST_MAX_TICK) return kClassDropped;
    if (e.kind <= MST_EV_PITCHWHEEL) return e.channel & 15u;
    if (e.kind == MST_EV_CHANNEL_PREFIX) return e.channel <= 15 ? e.channel : kClassPrefixHi;
    if (e.kind == MST_EV_META_IGNORED) return kClassIgnored;
    return kClassGlobal;   // set_tempo, time_signature, key_signature, unknown
}

// ------------------------------------------------------------------------------------------------
// k_sort: merge_tracks' stable sort by absolute tick (:37-49) and split_channels' routing (:55-66)
// in one stable LSD radix sort on key = class << 24 | tick.  Stability keeps (track, in-track index)
// order for equal ticks because the input is track-major.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kSortThreads, 1)
k_sort(const MstEvent *__restrict__ ev, const MstSongIn *__restrict__ songs, Workspace ws) {
    __shared__ uint32_t whist[kSortWarps][256];
    __shared__ uint32_t scan_tmp[33];
    __shared__ int skip_flag;
    const uint32_t song = blockIdx.x;
    const uint64_t base = songs[song].event_offset;
    const uint32_t n = songs[song].n_events;
    const MstEvent *e = ev + base;
    uint32_t *ka = ws.keys_a + base, *va = ws.vals_a + base, *kb = ws.keys_b + base, *vb = ws.vals_b + base;
    // the sort key is packed to the bits this song needs: class << tb | tick with tb = bit length of the song's largest
    // tick, so a 128-bar song at 480 ticks per beat (18 tick bits + 5 class bits) takes three 8-bit passes, not four
    __shared__ uint32_t max_tick_s;
    if (threadIdx.x == 0) max_tick_s = 1;
    __syncthreads();
    // events are read ONCE, four per thread and step so that four 16-byte loads are in flight per lane (a CTA streams
    // ~0.5 MB of events: with one load per lane in flight the loop runs at the latency, not the bandwidth)
    uint32_t mt = 0;
    for (uint32_t i0 = threadIdx.x; i0 < n; i0 += 4 * kSortThreads) {
        MstEvent x[4];
#pragma unroll
        for (int u = 0; u < 4; u++) if (i0 + u * kSortThreads < n) x[u] = e[i0 + u * kSortThreads];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t i = i0 + u * kSortThreads;
            if (i >= n) break;
            const uint32_t cls = event_class(x[u]);
            const uint32_t tick = cls == kClassDropped ? 0u : x[u].tick;
            mt = max(mt, tick);
            ka[i] = (cls << 24) | tick;
        }
    }
    mt = (uint32_t)warp_reduce_max((int)mt);            // ticks are <= MST_MAX_TICK < 2^31
    if (lane_id() == 0) atomicMax(&max_tick_s, mt);
    for (uint32_t i = threadIdx.x; i < n; i += kSortThreads) va[i] = i;
    __syncthreads();
    // Fast path.  Every track is already in tick order (delta times are non-negative), so when no two tracks share a
    // class (the usual format-1 layout: one channel per track; also any format-0 file) the stable partition by class IS
    // the (class, tick) order.  One pass on the class byte (long runs of equal digits: near-coalesced scatter), then a
    // sortedness check; only songs that fail it take the full radix sort below.
    __shared__ int unsorted_s;
    if (threadIdx.x == 0) unsorted_s = 0;
    radix_pass(ka, va, kb, vb, n, 24, whist, scan_tmp, &skip_flag);       // ends with a CTA barrier
    {
        int bad = 0;
        for (uint32_t i = threadIdx.x; i + 1 < n; i += kSortThreads) bad |= kb[i] > kb[i + 1];
        if (bad) unsorted_s = 1;
    }
    __syncthreads();
    int tb = 24;
    uint32_t tick_mask = 0xFFFFFFu;
    uint32_t *kin = kb, *vin = vb, *kout = ka, *vout = va;
    if (unsorted_s) {
        tb = 32 - __clz(max_tick_s);
        tick_mask = (1u << tb) - 1u;
        for (uint32_t i = threadIdx.x; i < n; i += kSortThreads) {  // own elements again: no barrier needed in between
            const uint32_t k = ka[i];
            ka[i] = ((k >> 24) << tb) | (k & 0xFFFFFFu);
            va[i] = i;
        }
        __syncthreads();
        kin = ka; vin = va; kout = kb; vout = vb;
        for (int shift = 0; shift < tb + 5; shift += 8) {
            radix_pass(kin, vin, kout, vout, n, shift, whist, scan_tmp, &skip_flag);
            uint32_t *t = kin; kin = kout; kout = t;
            t = vin; vin = vout; vout = t;
        }
    }
    // gather the events into merged order, leave (class << 24 | tick, source index) in keys_a / vals_a for k_info, and
    // find the class segments
    MstEvent *se = ws.sorted_ev + base;
    for (uint32_t i0 = threadIdx.x; i0 < n; i0 += 4 * kSortThreads) {
        uint32_t k[4], v[4];
        MstEvent x[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t i = i0 + u * kSortThreads;
            k[u] = i < n ? kin[i] : 0u;
            v[u] = i < n ? vin[i] : 0u;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) if (i0 + u * kSortThreads < n) x[u] = e[v[u]];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t i = i0 + u * kSortThreads;
            if (i >= n) break;
            se[i] = x[u];
            ka[i] = ((k[u] >> tb) << 24) | (k[u] & tick_mask);
            va[i] = v[u];
        }
    }
    __syncthreads();
    if (threadIdx.x <= kClassEnd) {
        uint32_t target = threadIdx.x << 24;
        uint32_t lo = 0, hi = n;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (ka[mid] < target) lo = mid + 1; else hi = mid;
        }
        ws.seg_start[(size_t)song * kSegStride + threadIdx.x] = lo;
    }
}

// ------------------------------------------------------------------------------------------------
// k_walk: group_channel_messages (:182-210) as a warp-per-channel scan.  Program / CC7 state is carried
// across 32-message strips; inside a strip the "last program_change before me" comes from a ballot.
// WRITE=0 counts messages per (channel, instrument); WRITE=1 emits NoteMessage records into the
// per-list storage laid out by k_info.
// ------------------------------------------------------------------------------------------------
template <int WRITE>
__global__ void __launch_bounds__(128)
k_walk(const MstSongIn *__restrict__ songs, uint32_t n_songs, Workspace ws, const MstSongInfo *__restrict__ info) {
    const unsigned lane = lane_id();
    const uint32_t warps_per_block = blockDim.x >> 5;
    const uint32_t n_units = n_songs * 16u;
    for (uint32_t unit = blockIdx.x * warps_per_block + (threadIdx.x >> 5); unit < n_units;
         unit += gridDim.x * warps_per_block) {
        const uint32_t song = unit >> 4, c = unit & 15u;
        if (WRITE && !song_active(info[song].status)) continue;
        const uint64_t base = songs[song].event_offset;
        const uint32_t *seg = ws.seg_start + (size_t)song * kSegStride;
        const uint32_t s0 = seg[c], s1 = seg[c + 1];
        const MstEvent *se = ws.sorted_ev + base;
        ChanRec *cr = ws.chans + (size_t)song * 16 + c;

        int instr = (c == 9) ? -1 : 0;   // get_instrument_id(0, channel_id), midi.py:90-93
        int vol = 96;                    // default_volume
        int n_slots = 0, error = 0;
        int slot_instr = 0;              // lane s holds slot s
        uint32_t slot_msgs = 0, slot_ons = 0;
        int first_on = INT_MAX, last_on = -1;
        // WRITE: lane s holds the list index and the write cursor of slot s
        int slot_list = -1;
        uint32_t slot_cursor = 0;
        if (WRITE) {
            n_slots = cr->n_slots;
            if (lane < (unsigned)n_slots) {
                slot_instr = cr->slot_instr[lane];
                slot_list = ws.chan_slot_list[((size_t)song * 16 + c) * 16 + lane];
                if (slot_list >= 0) slot_cursor = ws.lists[(size_t)song * kMaxLists + slot_list].msg_offset;
            }
        }
        for (uint32_t b = s0; b < s1; b += 32) {
            const uint32_t i = b + lane;
            const bool valid = i < s1;
            MstEvent x;
            x.kind = 255; x.d1 = 0; x.d2 = 0; x.tick = 0;
            if (valid) x = se[i];
            const bool is_pc = valid && x.kind == MST_EV_PROGRAM_CHANGE;
            const bool is_cc7 = valid && x.kind == MST_EV_CONTROL_CHANGE && x.d1 == 7;
            const unsigned pc_mask = __ballot_sync(0xffffffffu, is_pc);
            const unsigned cc_mask = __ballot_sync(0xffffffffu, is_cc7);
            unsigned m = pc_mask & lanemask_lt();
            int v = __shfl_sync(0xffffffffu, (int)x.d1, m ? 31 - __clz(m) : 0);
            int my_instr = m ? ((c == 9) ? -1 : v) : instr;
            m = cc_mask & lanemask_lt();
            v = __shfl_sync(0xffffffffu, (int)x.d2, m ? 31 - __clz(m) : 0);
            int my_vol = m ? v : vol;
            if (pc_mask) {
                v = __shfl_sync(0xffffffffu, (int)x.d1, 31 - __clz(pc_mask));
                instr = (c == 9) ? -1 : v;
            }
            if (cc_mask) vol = __shfl_sync(0xffffffffu, (int)x.d2, 31 - __clz(cc_mask));

            const bool is_note = valid && (x.kind == MST_EV_NOTE_ON || x.kind == MST_EV_NOTE_OFF);
            const int prod = (int)x.d2 * my_vol;     // velocity * volume (:199); 0 <=> velocity == 0.0
            const bool is_on = is_note && x.kind == MST_EV_NOTE_ON && prod != 0;
            if (valid && x.kind == MST_EV_NOTE_ON && x.d2 > 0) {   // get_midi_info's notes_on filter (:123)
                first_on = min(first_on, (int)x.tick);
                last_on = max(last_on, (int)x.tick);
            }
            // instrument -> slot (dict insertion order of instrument_id2messages, :184,203)
            int my_slot = -1;
            for (int s = 0; s < n_slots; s++) {
                int si = __shfl_sync(0xffffffffu, slot_instr, s);
                if (is_note && my_instr == si) my_slot = s;
            }
            if (!WRITE) {
                while (true) {
                    unsigned need = __ballot_sync(0xffffffffu, is_note && my_slot < 0);
                    if (!need) break;
                    if (n_slots == MST_MAX_LISTS_PER_CHANNEL) { error = 1; break; }
                    int ni = __shfl_sync(0xffffffffu, my_instr, __ffs(need) - 1);
                    if (lane == (unsigned)n_slots) slot_instr = ni;
                    if (is_note && my_slot < 0 && my_instr == ni) my_slot = n_slots;
                    n_slots++;
                }
                for (int s = 0; s < n_slots; s++) {
                    unsigned cm = __ballot_sync(0xffffffffu, is_note && my_slot == s);
                    unsigned om = __ballot_sync(0xffffffffu, is_on && my_slot == s);
                    if (lane == (unsigned)s) { slot_msgs += __popc(cm); slot_ons += __popc(om); }
                }
                if (error) break;
            } else {
                for (int s = 0; s < n_slots; s++) {
                    unsigned cm = __ballot_sync(0xffffffffu, is_note && my_slot == s);
                    int list = __shfl_sync(0xffffffffu, slot_list, s);
                    uint32_t cur = __shfl_sync(0xffffffffu, slot_cursor, s);
                    if (is_note && my_slot == s && list >= 0) {
                        NoteMsg nm;
                        nm.time = x.tick;
                        nm.is_on = is_on ? 1 : 0;
                        nm.note = x.d1 & 127;
                        nm.list = (uint16_t)list;
                        nm.velocity = (double)prod / (double)(127 * 127);   // :199
                        ws.msgs[base + cur + __popc(cm & lanemask_lt())] = nm;
                    }
                    if (lane == (unsigned)s) slot_cursor += __popc(cm);
                }
            }
        }
        if (!WRITE) {
            first_on = warp_reduce_min(first_on);
            last_on = warp_reduce_max(last_on);
            if (lane == 0) { cr->n_slots = n_slots; cr->error = error; cr->first_on = first_on; cr->last_on = last_on; }
            if (lane < MST_MAX_LISTS_PER_CHANNEL) {
                cr->slot_instr[lane] = (int16_t)slot_instr;
                cr->slot_msgs[lane] = slot_msgs;
                cr->slot_ons[lane] = slot_ons;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_info: get_midi_info (:117-179), read_midi's channel entries (:216-232), iter_inputs' filter
// (data.py:53-58), get_input's instrument grouping (data.py:69-77,97) and the grid sizes.
// One warp per song; the order-dependent walks run on lane 0.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
k_info(const MstSongIn *__restrict__ songs, uint32_t n_songs, Workspace ws, MstFilter filter, MstSongInfo *info_out) {
    const unsigned lane = lane_id();
    const uint32_t warps_per_block = blockDim.x >> 5;
    for (uint32_t song = blockIdx.x * warps_per_block + (threadIdx.x >> 5); song < n_songs;
         song += gridDim.x * warps_per_block) {
        const uint64_t base = songs[song].event_offset;
        const uint32_t n_events = songs[song].n_events;
        const int tpb = songs[song].ticks_per_beat;
        const uint32_t *seg = ws.seg_start + (size_t)song * kSegStride;
        const uint32_t *keys = ws.keys_a + base, *vals = ws.vals_a + base;
        const MstEvent *se = ws.sorted_ev + base;
        const ChanRec *chans = ws.chans + (size_t)song * 16;
        MstSongInfo *info = info_out + song;
        SongPlan *plan = ws.plans + song;

        // zero the info record (488 B = 122 words) and the plan
        for (unsigned i = lane; i < sizeof(MstSongInfo) / 4; i += 32) ((uint32_t *)info)[i] = 0;
        // tempo hash table (keys_b = tempo, vals_b = accumulated ticks, tmp_c = insertion order)
        uint32_t *h_tempo = ws.keys_b + base, *h_total = ws.vals_b + base, *h_order = ws.tmp_c + base;
        for (uint32_t i = lane; i < n_events; i += 32) h_tempo[i] = 0xFFFFFFFFu;

        int first_on = INT_MAX, last_on = -1, duration = -1, too_many = 0;
        if (lane < 16) { first_on = chans[lane].first_on; last_on = chans[lane].last_on; too_many = chans[lane].error; }
        // duration = max tick over messages that carry a channel (:126): classes 0..15 and 17, not the globals
        if (lane < kClassEnd && lane != kClassGlobal && seg[lane + 1] > seg[lane]) duration = (int)(keys[seg[lane + 1] - 1] & 0xFFFFFFu);
        first_on = warp_reduce_min(first_on);
        last_on = warp_reduce_max(last_on);
        duration = warp_reduce_max(duration);
        too_many = warp_reduce_max(too_many);
        __syncwarp();
        if (lane != 0) continue;

        plan->out_floats = 0; plan->n_bins = 0; plan->n_tiles = 0; plan->n_lists = 0; plan->fits = 0;
        info->ticks_per_beat = tpb;
        info->numerator = 4; info->denominator_log2 = 2; info->key_signature = -1;
        if (last_on < 0) { info->status = MST_SONG_NO_NOTES; continue; }
        info->duration = duration;

        // ---- global messages in merged order (:142-165)
        int tempo = 500000, tempo_change_time = 0;
        int numerator = 4, den_log2 = 2, key = -1, status = MST_SONG_OK;
        uint32_t n_entries = 0;
        auto tempo_add = [&](int t, int span) {
            uint32_t h = ((uint32_t)t * 2654435761u) % n_events;
            while (h_tempo[h] != 0xFFFFFFFFu && h_tempo[h] != (uint32_t)t) h = (h + 1 == n_events) ? 0 : h + 1;
            if (h_tempo[h] == 0xFFFFFFFFu) { h_tempo[h] = (uint32_t)t; h_total[h] = 0; h_order[n_entries++] = h; }
            h_total[h] = (uint32_t)((int)h_total[h] + span);
        };
        for (uint32_t r = seg[kClassGlobal]; r < seg[kClassGlobal + 1]; r++) {
            const MstEvent x = se[r];
            const int t = (int)x.tick;
            const bool during = first_on <= t && t <= last_on;
            if (x.kind == MST_EV_TIME_SIGNATURE) {
                if (x.d1 != numerator || x.d2 != den_log2) {
                    if (during) { status = MST_SONG_TIME_SIG_CHANGED; break; }
                    numerator = x.d1; den_log2 = x.d2;
                }
            } else if (x.kind == MST_EV_KEY_SIGNATURE) {
                if ((int)x.aux != key) {
                    if (during) { status = MST_SONG_KEY_SIG_CHANGED; break; }
                    key = (int)x.aux;
                }
            } else if (x.kind == MST_EV_SET_TEMPO) {
                if ((int)x.aux != tempo) {
                    tempo_add(tempo, t - tempo_change_time);
                    tempo = (int)x.aux;
                    tempo_change_time = t;
                }
            } else {
                status = MST_SONG_UNKNOWN_MESSAGE; break;
            }
        }
        info->numerator = numerator; info->denominator_log2 = den_log2; info->key_signature = key;
        if (status == MST_SONG_OK && (tpb <= 0 || numerator == 0)) status = MST_SONG_BAD_TIMING;
        if (status == MST_SONG_OK && too_many) status = MST_SONG_TOO_MANY_INSTRUMENTS;
        if (status != MST_SONG_OK) { info->status = status; continue; }
        const int tpbar = tpb * numerator;                     // :167
        const int n_bars = (duration + tpbar - 1) / tpbar;     // ceil of :169 (math.ceil at :585)
        info->ticks_per_bar = tpbar; info->n_bars = n_bars; info->n_beats = numerator;
        tempo_add(tempo, duration - tempo_change_time);        // :172
        int best = -1, best_total = 0;
        for (uint32_t k = 0; k < n_entries; k++) {             // :173,176 first maximum among non-zero totals
            int tot = (int)h_total[h_order[k]];
            if (tot == 0) continue;
            if (best < 0 || tot > best_total) { best = (int)h_order[k]; best_total = tot; }
        }
        if (best < 0) { info->status = MST_SONG_BAD_TIMING; continue; }
        info->tempo = (int)h_tempo[best];
        info->bpm = (int)rint(60 * 1e6 / (double)info->tempo * 4 / 4.);   // round(mido.tempo2bpm(tempo)), :177

        // ---- channels in discovery order (first message in merged order, :57-66)
        int order[16]; unsigned long long okey[16]; int n_ch = 0;
        for (int c = 0; c < 16; c++) {
            if (seg[c + 1] == seg[c]) continue;
            unsigned long long k = ((unsigned long long)(keys[seg[c]] & 0xFFFFFFu) << 32) | vals[seg[c]];
            int p = n_ch++;
            while (p > 0 && okey[p - 1] > k) { okey[p] = okey[p - 1]; order[p] = order[p - 1]; p--; }
            okey[p] = k; order[p] = c;
        }
        // ---- channel entries (:220-230), filter (data.py:53-56), grouping by instrument (data.py:69-77)
        ListRec *lists = ws.lists + (size_t)song * kMaxLists;
        int16_t *csl = ws.chan_slot_list + (size_t)song * 256;
        int n_lists = 0, C = 0, U = 0;
        uint32_t msg_off = 0, note_off = 0;
        for (int oc = 0; oc < n_ch; oc++) {
            const int c = order[oc];
            const ChanRec &cr = chans[c];
            for (int s = 0; s < cr.n_slots; s++) {
                if (cr.slot_ons[s] == 0) { csl[c * 16 + s] = -1; continue; }   // no note_on: not a channel entry (:224)
                ListRec &L = lists[n_lists];
                const int id = cr.slot_instr[s];
                L.channel = (int16_t)c; L.instrument = (int16_t)id;
                L.msg_offset = msg_off; L.n_messages = cr.slot_msgs[s];
                L.n_notes = cr.slot_ons[s]; L.note_offset = note_off;
                L.pitched = (id >= 0 && id <= 119) ? 1 : 0;                    // is_pitched, midi.py:96-101
                msg_off += L.n_messages; note_off += L.n_notes;
                bool keep = true;
                if (filter.apply) {
                    bool ok = (id == -1) || (id >= 0 && id < 128 && filter.included[id]);
                    keep = ok && (int)L.n_messages >= filter.min_n_messages;
                }
                int slot = -1;
                if (keep) {
                    if (L.pitched) {
                        for (int k = 0; k < C; k++) if (info->pitched_instruments[k] == id) slot = k;
                        if (slot < 0) { slot = C; info->pitched_instruments[C++] = (int16_t)id; }
                    } else {
                        for (int k = 0; k < U; k++) if (info->unpitched_instruments[k] == id) slot = k;
                        if (slot < 0) { slot = U; info->unpitched_instruments[U++] = (int16_t)id; }
                        slot += 1000;   // rebased below once C is final
                    }
                }
                L.out_slot = (int16_t)slot;
                csl[c * 16 + s] = (int16_t)n_lists;
                n_lists++;
            }
        }
        for (int l = 0; l < n_lists; l++) if (lists[l].out_slot >= 1000) lists[l].out_slot = (int16_t)(lists[l].out_slot - 1000 + C);
        info->n_lists = n_lists; info->n_pitched = C; info->n_unpitched = U;
        plan->n_lists = (uint32_t)n_lists;
        if (C == 0) { info->status = MST_SONG_NO_PITCHED; continue; }   // data.py:57
        int n_out = n_bars + 1;                                         // :493
        if (filter.crop_channel_bars > 0) n_out = min(n_out, filter.crop_channel_bars / C);   // train-model.py:101
        info->n_bars_out = n_out;
        plan->out_floats = (uint64_t)n_out * numerator * ((uint64_t)C * MST_PITCHED_BEAT_FLOATS + (uint64_t)U * MST_UNPITCHED_BEAT_FLOATS);
        plan->n_bins = (uint32_t)((C + U) * n_out);
        plan->n_tiles = plan->n_bins * (uint32_t)((numerator + kTileBeats - 1) / kTileBeats);
    }
}

// ------------------------------------------------------------------------------------------------
// k_offsets: exclusive prefix sums over the songs of the batch (grid floats, bins, tiles).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1)
k_offsets(uint32_t n_songs, Workspace ws, MstSongInfo *info, uint64_t arena_floats) {
    __shared__ unsigned long long w64[33];
    __shared__ uint32_t w32[33];
    __shared__ unsigned long long fill_end;
    if (threadIdx.x == 0) fill_end = 0;
    __syncthreads();
    unsigned long long run_floats = 0;
    uint32_t run_bins = 0, run_tiles = 0;
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
    for (uint32_t c0 = 0; c0 < n_songs; c0 += blockDim.x) {
        const uint32_t song = c0 + threadIdx.x;
        const bool valid = song < n_songs;
        unsigned long long f = valid ? ws.plans[song].out_floats : 0ull;
        // 64-bit block exclusive scan
        unsigned long long inc = f;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= (unsigned)o) inc += t;
        }
        if (lane == 31) w64[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            unsigned long long s = w64[lane], si = s;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long t = __shfl_up_sync(0xffffffffu, si, o);
                if (lane >= (unsigned)o) si += t;
            }
            w64[lane] = si - s;
            if (lane == 31) w64[32] = si;
        }
        __syncthreads();
        const unsigned long long off = run_floats + w64[warp] + inc - f;
        const unsigned long long chunk_total = w64[32];
        __syncthreads();
        const bool ok = valid && info[song].status == MST_SONG_OK;
        bool fits = ok && off + f <= arena_floats;
        uint32_t bins = fits ? ws.plans[song].n_bins : 0u, tiles = fits ? ws.plans[song].n_tiles : 0u;
        uint32_t tb, tt;
        uint32_t bin_off = block_exclusive_scan(bins, w32, &tb);
        uint32_t tile_off = block_exclusive_scan(tiles, w32, &tt);
        if (valid) {
            SongPlan &p = ws.plans[song];
            p.fits = fits ? 1u : 0u;
            p.bin_base = run_bins + bin_off;
            p.tile_base = run_tiles + tile_off;
            ws.tile_prefix[song] = run_tiles + tile_off;
            if (!fits) { p.n_bins = 0; p.n_tiles = 0; }
            if (ok) {
                if (!fits) info[song].status = MST_SONG_ARENA_FULL;
                const uint64_t pf = (uint64_t)info[song].n_pitched * info[song].n_bars_out * info[song].n_beats * MST_PITCHED_BEAT_FLOATS;
                info[song].pitched_offset = off;
                info[song].unpitched_offset = off + pf;
            }
        }
        if (fits) atomicMax(&fill_end, off + f);     // songs that fit are a prefix of the offset order
        run_floats += chunk_total; run_bins += tb; run_tiles += tt;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        ws.tile_prefix[n_songs] = run_tiles;
        ws.totals[0] = run_tiles;
        ws.totals[1] = run_bins;
        *reinterpret_cast<unsigned long long *>(ws.totals + 2) = fill_end;   // floats [0, fill_end) hold the batch's grids
    }
}

// ------------------------------------------------------------------------------------------------
// k_pair: channel2nchannel (:371-406).  One warp per channel entry; a strip of 32 NoteMessages is
// paired with match_any on the pitch (next message on the same pitch closes the note), an open-note
// table in shared memory carries notes across strips.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_pair(const MstSongIn *__restrict__ songs, Workspace ws, const MstSongInfo *__restrict__ info) {
    __shared__ uint32_t open_tab[8][128];
    const uint32_t song = blockIdx.x;
    if (!song_active(info[song].status)) return;
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5;
    const uint64_t base = songs[song].event_offset;
    const ListRec *lists = ws.lists + (size_t)song * kMaxLists;
    const uint32_t n_lists = ws.plans[song].n_lists;
    uint32_t *tab = open_tab[warp];
    for (uint32_t l = warp; l < n_lists; l += 8) {
        const ListRec L = lists[l];
        if (L.out_slot < 0) continue;
        for (int i = lane; i < 128; i += 32) tab[i] = kNoOpen;
        __syncwarp();
        const NoteMsg *m = ws.msgs + base + L.msg_offset;
        Note *notes = ws.notes + base + L.note_offset;
        uint32_t n_on = 0;
        for (uint32_t b = 0; b < L.n_messages; b += 32) {
            const uint32_t i = b + lane;
            const bool valid = i < L.n_messages;
            const unsigned act = __ballot_sync(0xffffffffu, valid);
            NoteMsg x;
            x.time = 0; x.is_on = 0; x.note = 0; x.velocity = 0;
            if (valid) x = m[i];
            const unsigned on_mask = __ballot_sync(0xffffffffu, valid && x.is_on);
            const uint32_t my_idx = n_on + __popc(on_mask & lanemask_lt());
            n_on += __popc(on_mask);
            unsigned peers = valid ? __match_any_sync(act, (unsigned)x.note) : 0u;
            const unsigned after = peers & lanemask_gt(), before = peers & lanemask_lt();
            // the next message on my pitch inside the strip closes me (:381-384)
            const int next_lane = after ? __ffs(after) - 1 : 0;
            const uint32_t next_time = __shfl_sync(0xffffffffu, x.time, next_lane);
            if (valid) {
                if (!before) {   // first message on this pitch in the strip: closes a note left open earlier
                    uint32_t o = tab[x.note];
                    if (o != kNoOpen) notes[o].end_time = (int32_t)x.time;
                }
                if (x.is_on) {
                    Note nt;
                    nt.note_id = x.note; nt.time = (int32_t)x.time;
                    nt.end_time = after ? (int32_t)next_time : (int32_t)x.time;   // unclosed: duration 0 (:393)
                    nt.pad = 0; nt.velocity = x.velocity;
                    notes[my_idx] = nt;
                }
            }
            __syncwarp();
            if (valid && !after) tab[x.note] = x.is_on ? my_idx : kNoOpen;
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_scale: get_keys_dist (:340-346) + data.py:79-83 + get_scale (scales.py:160-221).
// ------------------------------------------------------------------------------------------------
struct PySum {   // Python >= 3.12 builtin sum(): Neumaier compensation (bltinmodule.c)
    double s, c;
    bool started;
    __device__ void add(double x) {
        if (!started) { s = 0.0; c = 0.0; started = true; }
        double t = s + x;
        if (fabs(s) >= fabs(x)) c += (s - t) + x; else c += (x - t) + s;
        s = t;
    }
    __device__ double result() const {
        if (!started) return 0.0;
        double r = s;
        if (c != 0.0 && isfinite(c)) r += c;
        return r;
    }
};

__device__ double np_sum(const double *a, int n) {   // numpy pairwise_sum, n < 128
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; i++) res = __dadd_rn(res, a[i]);
        return res;
    }
    double r[8];
    for (int j = 0; j < 8; j++) r[j] = a[j];
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
        for (int j = 0; j < 8; j++) r[j] = __dadd_rn(r[j], a[i + j]);
    double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                           __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
    for (; i < n; i++) res = __dadd_rn(res, a[i]);
    return res;
}

__constant__ double c_profiles[2][12] = {
    {6.35, 2.23, 3.48, 2.33, 4.38, 4.09, 2.52, 5.19, 2.39, 3.66, 2.29, 2.88},   // scales.py:111
    {6.33, 2.68, 3.52, 5.38, 2.60, 3.53, 2.54, 4.75, 3.98, 2.69, 3.34, 3.17},   // scales.py:114
};
__constant__ int8_t c_abs_intervals[2][7] = {{0, 2, 4, 5, 7, 9, 11}, {0, 2, 3, 5, 7, 8, 10}};
__constant__ int8_t c_typical[2][10] = {{0, 2, 4, 5, 6, 7, 9, 10, 11, 0}, {0, 1, 2, 3, 5, 7, 8, 9, 10, 11}};
__constant__ int8_t c_degree[2][12] = {{1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 7, 7}, {1, 2, 2, 3, 4, 4, 5, 5, 6, 6, 7, 7}};
__constant__ int8_t c_accidental[2][12] = {{1, 0, 1, 0, 1, 1, 2, 1, 2, 1, 0, 1}, {1, 0, 1, 1, 0, 1, 0, 1, 1, 2, 1, 2}};
__constant__ int8_t c_col8[8] = {0, 1, 2, 4, 5, 6, 8, 9};
__constant__ int8_t c_col3[3] = {0, 3, 7};

struct __align__(16) ScaleTerm {   // duration * velocity of one note, placed at its merged-order rank
    double prod;
    int32_t key;
    int32_t pad;
};
static_assert(sizeof(ScaleTerm) == sizeof(Cell), "ScaleTerm aliases the cell scratch");

constexpr int kScaleThreads = 512;     // 16 warps: one per pitched instrument of a 15 + 1 track song in a single round

__global__ void __launch_bounds__(kScaleThreads)
k_scale(const MstSongIn *__restrict__ songs, Workspace ws, MstSongInfo *info_out) {
    __shared__ double sec[MST_MAX_PITCHED * 12];
    __shared__ uint8_t started[MST_MAX_PITCHED * 12];
    __shared__ double hist[12];
    __shared__ double losses[24], covs[24], looses[24], ces[24];
    __shared__ uint32_t inst_base[MST_MAX_PITCHED + 1];
    __shared__ uint16_t inst_lists[MST_MAX_PITCHED];
    const uint32_t song = blockIdx.x;
    MstSongInfo *info = info_out + song;
    if (!song_active(info->status)) return;
    const uint64_t base = songs[song].event_offset;
    const ListRec *lists = ws.lists + (size_t)song * kMaxLists;
    const int n_lists = (int)ws.plans[song].n_lists;
    const int C = info->n_pitched;
    const Note *notes = ws.notes + base;
    ScaleTerm *terms = reinterpret_cast<ScaleTerm *>(ws.cells + base);   // free until k_cells runs
    const unsigned lane = lane_id(), warp = threadIdx.x >> 5, nwarps = kScaleThreads / 32;

    // notes per pitched instrument (merge_nchannels, data.py:103-114) and their term ranges
    for (int p = threadIdx.x; p < C; p += kScaleThreads) {
        uint32_t cnt = 0; int nl = 0;
        for (int l = 0; l < n_lists; l++) if (lists[l].pitched && lists[l].out_slot == p) { cnt += lists[l].n_notes; nl++; }
        inst_base[p + 1] = cnt;
        inst_lists[p] = (uint16_t)nl;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        inst_base[0] = 0;
        for (int p = 0; p < C; p++) inst_base[p + 1] += inst_base[p];
    }
    __syncthreads();
    // each note goes to its rank in sorted(notes, key=time) of the concatenated lists (stable):
    // own position + notes of earlier lists with time <= t + notes of later lists with time < t
    for (int l = 0; l < n_lists; l++) {
        const ListRec L = lists[l];
        if (!L.pitched || L.out_slot < 0) continue;
        const int p = L.out_slot;
        const Note *nt = notes + L.note_offset;
        const bool multi = inst_lists[p] > 1;
#pragma unroll 4
        for (uint32_t j = threadIdx.x; j < L.n_notes; j += kScaleThreads) {
            const Note x = nt[j];
            uint32_t rank = j;
            if (multi) {
                for (int l2 = 0; l2 < n_lists; l2++) {
                    if (l2 == l || !lists[l2].pitched || lists[l2].out_slot != p) continue;
                    const Note *o = notes + lists[l2].note_offset;
                    const int thr = (l2 < l) ? x.time + 1 : x.time;   // count notes with time < thr
                    uint32_t lo = 0, hi = lists[l2].n_notes;
                    while (lo < hi) { uint32_t mid = (lo + hi) >> 1; if (o[mid].time < thr) lo = mid + 1; else hi = mid; }
                    rank += lo;
                }
            }
            ScaleTerm t;
            t.prod = __dmul_rn((double)(x.end_time - x.time), x.velocity);   // duration * velocity (:342)
            t.key = x.note_id % 12;
            t.pad = 0;
            terms[inst_base[p] + rank] = t;
        }
    }
    __syncthreads();
    // Python sum() per (instrument, pitch class) in merged order: lane k owns class k.  Inside a strip of 32
    // terms every lane walks the set bits of its own class mask, so the adds of different classes overlap.
    for (int p = warp; p < C; p += nwarps) {
        PySum acc; acc.started = false; acc.s = 0; acc.c = 0;
        const uint32_t t0 = inst_base[p], t1 = inst_base[p + 1];
        // the next strip's terms are fetched while the current strip's dependent adds run
        ScaleTerm nxt; nxt.prod = 0.0; nxt.key = -1;
        if (t0 + lane < t1) nxt = terms[t0 + lane];
        for (uint32_t b = t0; b < t1; b += 32) {
            const double prod = nxt.prod; const int key = nxt.key;
            nxt.prod = 0.0; nxt.key = -1;
            if (b + 32 + lane < t1) nxt = terms[b + 32 + lane];
            unsigned mine = 0;
#pragma unroll
            for (int k = 0; k < 12; k++) {
                const unsigned mk = __ballot_sync(0xffffffffu, key == k);
                if (lane == (unsigned)k) mine = mk;
            }
            int iters = __popc(mine);
            iters = warp_reduce_max(iters);
            for (int it = 0; it < iters; it++) {
                const int src = mine ? __ffs(mine) - 1 : 0;
                const double x = __shfl_sync(0xffffffffu, prod, src);
                if (mine) { acc.add(x); mine &= mine - 1; }
            }
        }
        if (lane < 12) { sec[p * 12 + lane] = acc.result(); started[p * 12 + lane] = acc.started ? 1 : 0; }
    }
    __syncthreads();
    if (threadIdx.x < 12) {
        // mido.tick2second, then the pandas column sum over instruments (NaN = key absent, skipped)
        const double scale = __ddiv_rn(__dmul_rn((double)info->tempo, 1e-6), (double)info->ticks_per_beat);
        double h = 0.0; bool any = false;
        for (int p = 0; p < C; p++) {
            if (!started[p * 12 + threadIdx.x]) continue;
            double v = __dmul_rn(sec[p * 12 + threadIdx.x], scale);
            h = any ? __dadd_rn(h, v) : v;
            any = true;
        }
        hist[threadIdx.x] = h;
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        __shared__ double kd[12];
        if (threadIdx.x == 0) {
            double total = np_sum(hist, 12);
            for (int i = 0; i < 12; i++) { kd[i] = __ddiv_rn(hist[i], total); info->keys_dist[i] = kd[i]; }
        }
        __syncwarp();
        if (threadIdx.x < 24) {
            const int m = threadIdx.x / 12, t = threadIdx.x % 12;
            double prof[12], g[12];
            for (int i = 0; i < 12; i++) prof[i] = c_profiles[m][i];
            const double ptot = np_sum(prof, 12);
            for (int i = 0; i < 7; i++) g[i] = kd[(c_abs_intervals[m][i] + t) % 12];
            const double coverage = np_sum(g, 7);
            const int nt = m ? 10 : 9;
            for (int i = 0; i < nt; i++) g[i] = kd[(c_typical[m][i] + t) % 12];
            const double loose = np_sum(g, nt);
            for (int i = 0; i < 12; i++) {
                double d = kd[(i + t) % 12];
                if (d < 1e-12) d = 1e-12; else if (d > 1.0) d = 1.0;   // np.clip; NaN propagates
                g[i] = __dmul_rn(__ddiv_rn(prof[i], ptot), log(d));
            }
            const double ce = __ddiv_rn(-np_sum(g, 12), 12.0);
            losses[threadIdx.x] = __dmul_rn(__dmul_rn(ce, __dsub_rn(1.5, coverage)), __dsub_rn(2.0, loose));
            covs[threadIdx.x] = coverage; looses[threadIdx.x] = loose; ces[threadIdx.x] = ce;
        }
        __syncwarp();
        if (threadIdx.x == 0) {   // min(scales, key=loss): first minimum, NaN never smaller (scales.py:216)
            int best = 0; double second = CUDART_INF;
            for (int i = 1; i < 24; i++) {
                if (losses[i] < losses[best]) { second = losses[best]; best = i; }
                else if (losses[i] < second) second = losses[i];
            }
            info->tonic = best % 12; info->mode = best / 12;
            info->scale_coverage = covs[best]; info->scale_loose_coverage = looses[best];
            info->scale_cross_entropy = ces[best]; info->scale_loss = losses[best];
            info->scale_margin = second - losses[best];
        }
    }
}

// ------------------------------------------------------------------------------------------------
// k_cells: nchannel2kchannel (:408-423), kchannel2qchannel (:425-456, round_number utils/math.py:14-19),
// note2idx (:597-609) and the feature values of qchannel2vchannel (:501-511), then a counting sort of
// the resulting cells by (slot, bar) bin.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double py_floordiv(double vx, double wx) {   // CPython float_divmod, positive wx
    double mod = fmod(vx, wx);
    double div = __ddiv_rn(__dsub_rn(vx, mod), wx);
    if (mod != 0.0 && ((wx < 0) != (mod < 0))) div = __dsub_rn(div, 1.0);
    if (div != 0.0) {
        double fl = floor(div);
        if (__dsub_rn(div, fl) > 0.5) fl = __dadd_rn(fl, 1.0);
        return fl;
    }
    return 0.0;
}

__device__ __forceinline__ int quantise(int time, int tpb, int tpbar, int *qtime, int *bar, int *beat) {
    double best_q = 0, best_err = 0; int best_d = 8;
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const int d = k ? 3 : 8;
        const double ticks = __ddiv_rn((double)tpb, (double)d);
        double rpos = fmod((double)time, ticks);          // time >= 0, ticks > 0: Python % == fmod
        const double rneg = fabs(__dsub_rn(rpos, ticks));
        double q, err;
        if (rpos < rneg) { q = __dsub_rn((double)time, rpos); err = rpos; }
        else { q = __dadd_rn((double)time, rneg); err = rneg; }
        if (k == 0 || err < best_err) { best_q = q; best_err = err; best_d = d; }   // min(): 8 wins ties
    }
    const int qt = (int)best_q;                            // int(): truncation (:447)
    *qtime = qt;
    const int b = qt / tpbar; int rem = qt - b * tpbar;
    const int bt = rem / tpb; rem -= bt * tpb;
    *bar = b; *beat = bt;
    const int quants = (int)py_floordiv((double)rem, __ddiv_rn((double)tpb, (double)best_d));
    if (best_d == 8) return (quants >= 0 && quants < 8) ? c_col8[quants] : -1;
    return (quants >= 0 && quants < 3) ? c_col3[quants] : -1;
}

// DIRECT = 1 (the default path): the arena range of the batch has been zero-filled by k_fill, so every quantised note
// goes straight to its cell with the same max / flag semantics (integer atomicMax on the non-negative f32 bit patterns is
// np.maximum; order-independent, hence deterministic) and nothing is binned.  DIRECT = 0 bins the cells per (slot, bar)
// for the shared-memory tile writer k_grid (kept for comparison: MST_GRID_TILES=1).
template <int DIRECT>
__global__ void __launch_bounds__(256)
k_cells(const MstSongIn *__restrict__ songs, Workspace ws, MstSongInfo *info_out, float *__restrict__ arena) {
    __shared__ uint32_t scan_tmp[33];
    __shared__ uint32_t s_count;
    const uint32_t song = blockIdx.x;
    MstSongInfo *info = info_out + song;
    if (info->status != MST_SONG_OK) return;   // ARENA_FULL songs are planned but not written
    const SongPlan plan = ws.plans[song];
    const uint64_t base = songs[song].event_offset;
    const ListRec *lists = ws.lists + (size_t)song * kMaxLists;
    const int tpb = info->ticks_per_beat, tpbar = info->ticks_per_bar, nb = info->n_beats;
    const int n_out = info->n_bars_out, n_grid = info->n_bars + 1;
    const int tonic = info->tonic, mode = info->mode;
    const uint64_t p_off = info->pitched_offset, u_off = info->unpitched_offset;
    const uint32_t Cp = (uint32_t)info->n_pitched;
    uint32_t *bin_count = ws.bin_count + plan.bin_base, *bin_start = ws.bin_start + plan.bin_base;
    uint32_t *bin_fill = ws.bin_fill + plan.bin_base;
    Cell *cells = ws.cells + base;
    if (!DIRECT) {
        for (uint32_t i = threadIdx.x; i < plan.n_bins; i += blockDim.x) { bin_count[i] = 0; bin_fill[i] = 0; }
        if (threadIdx.x == 0) s_count = 0;
        __syncthreads();
    }
    uint32_t scattered = 0;
    for (uint32_t l = 0; l < plan.n_lists; l++) {
        const ListRec L = lists[l];
        if (L.out_slot < 0) continue;
        const Note *notes = ws.notes + base + L.note_offset;
        Note nxt;                                                    // the next note is fetched while this one is quantised
        if (threadIdx.x < L.n_notes) nxt = notes[threadIdx.x];
        for (uint32_t j = threadIdx.x; j < L.n_notes; j += blockDim.x) {
            const Note nt = nxt;
            if (j + blockDim.x < L.n_notes) nxt = notes[j + blockDim.x];
            int row, acc = 1;
            if (L.pitched) {
                int octave = nt.note_id / 12 - 1;                       // note_id2key_octave (:309-317)
                int interval = nt.note_id % 12 - tonic;                 // note2scale_loc (:244-266)
                if (interval < 0) { octave -= 1; interval += 12; }
                row = octave * 7 + c_degree[mode][interval] - 1;        // note2idx (:597-604)
                acc = c_accidental[mode][interval];
                if (row < 0 || row >= MST_N_PITCHED_ROWS) continue;
            } else {
                if (nt.note_id < 35 || nt.note_id > 81) continue;       // :605-609
                row = nt.note_id - 35;
                acc = 7;                                                // no accidental features
            }
            int qtime, bar, beat;
            const int col = quantise(nt.time, tpb, tpbar, &qtime, &bar, &beat);
            if (col < 0 || bar >= n_grid || beat >= nb) continue;
            scattered++;
            if (bar >= n_out) continue;                                  // cropped by prepare_input
            const int qdur = nt.end_time - qtime;                       // :448, may be negative
            const double duration = __ddiv_rn((double)qdur, (double)tpb);   // :501
            const uint32_t rows = L.pitched ? MST_N_PITCHED_ROWS : MST_N_UNPITCHED_ROWS;
            const uint32_t feats = L.pitched ? MST_N_PITCHED_FEATS : MST_N_UNPITCHED_FEATS;
            if (DIRECT) {
                const uint64_t slab = L.pitched
                    ? p_off + ((uint64_t)L.out_slot * n_out + bar) * nb * MST_PITCHED_BEAT_FLOATS
                    : u_off + ((uint64_t)(L.out_slot - Cp) * n_out + bar) * nb * MST_UNPITCHED_BEAT_FLOATS;
                float *cell = arena + slab + (((uint32_t)beat * MST_N_FRACTIONS + col) * rows + row) * feats;
                const float dur = (float)duration, vel = (float)nt.velocity;       // f64 -> f32 of prepare_input (data.py:136)
                if (dur > 0.f) atomicMax((int *)cell, __float_as_int(dur));        // np.maximum with 0 (:514)
                atomicMax((int *)cell + 1, __float_as_int(vel));
                if (acc < 3) cell[2 + acc] = 1.0f;
                continue;
            }
            Cell c;
            c.bin = plan.bin_base + (uint32_t)L.out_slot * n_out + bar;
            c.off_acc = (((uint32_t)beat * MST_N_FRACTIONS + col) * rows + row) * feats | ((uint32_t)acc << 28);
            c.dur = (float)duration;                                     // f64 -> f32 of prepare_input (data.py:136)
            c.vel = (float)nt.velocity;
            const uint32_t pos = atomicAdd(&s_count, 1u);
            cells[pos] = c;
            atomicAdd(&bin_count[c.bin - plan.bin_base], 1u);
        }
    }
    // n_notes (informative)
    for (int o = 16; o; o >>= 1) scattered += __shfl_xor_sync(0xffffffffu, scattered, o);
    if (lane_id() == 0 && scattered) atomicAdd((unsigned *)&info->n_notes, scattered);
    if (DIRECT) return;
    __syncthreads();
    // exclusive scan of the song's bins; bin_start is an absolute index into cells_sorted
    uint32_t running = 0;
    for (uint32_t c0 = 0; c0 < plan.n_bins; c0 += blockDim.x) {
        const uint32_t i = c0 + threadIdx.x;
        const uint32_t v = i < plan.n_bins ? bin_count[i] : 0u;
        uint32_t tot;
        const uint32_t ex = block_exclusive_scan(v, scan_tmp, &tot);
        if (i < plan.n_bins) bin_start[i] = (uint32_t)base + running + ex;
        running += tot;
    }
    __syncthreads();
    Cell *sorted = ws.cells_sorted;
    const uint32_t n_cells = s_count;
    for (uint32_t i = threadIdx.x; i < n_cells; i += blockDim.x) {
        const Cell c = cells[i];
        const uint32_t b = c.bin - plan.bin_base;
        const uint32_t pos = bin_start[b] + atomicAdd(&bin_fill[b], 1u);
        sorted[pos] = c;
    }
    // tile descriptors for k_grid: (slot, bar, beat chunk) -> destination + cell range
    const uint32_t C = (uint32_t)info->n_pitched;
    const uint32_t chunks = ((uint32_t)nb + kTileBeats - 1) / kTileBeats;
    TileDesc *tiles = ws.tiles + plan.tile_base;
    for (uint32_t local = threadIdx.x; local < plan.n_tiles; local += blockDim.x) {
        const uint32_t bin = local / chunks, chunk = local % chunks;
        const uint32_t slot = bin / (uint32_t)n_out, bar = bin % (uint32_t)n_out;
        const bool pitched = slot < C;
        const uint32_t beat_floats = pitched ? MST_PITCHED_BEAT_FLOATS : MST_UNPITCHED_BEAT_FLOATS;
        const uint32_t beat_lo = chunk * kTileBeats, beats = min((uint32_t)kTileBeats, (uint32_t)nb - beat_lo);
        const uint64_t slab = pitched
            ? info->pitched_offset + ((uint64_t)slot * n_out + bar) * nb * beat_floats
            : info->unpitched_offset + ((uint64_t)(slot - C) * n_out + bar) * nb * beat_floats;
        TileDesc t;
        t.dst = slab + (uint64_t)beat_lo * beat_floats;
        t.floats = beats * beat_floats;
        t.beat_lo_off = beat_lo * beat_floats;
        const uint32_t cnt = bin_count[bin];
        t.cell_lo = cnt ? bin_start[bin] : 0u;
        t.cell_hi = t.cell_lo + cnt;
        t.pad[0] = 0; t.pad[1] = 0;
        tiles[local] = t;
    }
}

// ------------------------------------------------------------------------------------------------
// k_grid: qchannel2vchannel's scatter-max (:490-516) + np.stack (data.py:91-95) + f32 cast.
// Persistent CTAs over the tile descriptors k_cells wrote.  A tile = up to kTileBeats beats of one
// (slot, bar) = 44.8 KB (pitched, 4/4).  Tile buffers in shared memory are kept all-zero between uses:
// notes are scattered with atomicMax on the (non-negative) f32 bit patterns, the tile leaves with one
// TMA bulk store (cp.async.bulk shared -> global), and once that store has read the buffer only the
// touched cells are reset.  Empty tiles are stored straight from a read-only zero buffer.
// ------------------------------------------------------------------------------------------------
constexpr int kGridThreads = 256;
constexpr int kGridBufs = 4;
constexpr uint32_t kTileFloatsMax = kTileBeats * MST_PITCHED_BEAT_FLOATS;   // 11200 floats = 44800 B

__device__ __forceinline__ TileDesc load_tile(const TileDesc *tiles, uint32_t tile, uint32_t n_tiles) {
    TileDesc t;
    if (tile < n_tiles) {
        const uint4 *p = reinterpret_cast<const uint4 *>(tiles + tile);
        const uint4 a = __ldg(p), b = __ldg(p + 1);
        t.dst = ((uint64_t)a.y << 32) | a.x; t.floats = a.z; t.cell_lo = a.w;
        t.cell_hi = b.x; t.beat_lo_off = b.y; t.pad[0] = 0; t.pad[1] = 0;
    } else {
        t.dst = 0; t.floats = 0; t.cell_lo = 0; t.cell_hi = 0; t.beat_lo_off = 0; t.pad[0] = 0; t.pad[1] = 0;
    }
    return t;
}

constexpr uint32_t kNoCell = 0xFFFFFFFFu;

// scatter (or reset) one cell into a tile buffer; returns the tile-relative offset|acc word, kNoCell when
// the cell belongs to another beat chunk of the bar
__device__ __forceinline__ uint32_t apply_cell(float *buf, const Cell &c, uint32_t beat_lo_off, uint32_t floats) {
    const uint32_t off = c.off_acc & 0x0FFFFFFFu;
    if (off < beat_lo_off || off >= beat_lo_off + floats) return kNoCell;
    float *cell = buf + (off - beat_lo_off);
    const uint32_t acc = c.off_acc >> 28;
    if (c.dur > 0.f) atomicMax((int *)cell, __float_as_int(c.dur));       // np.maximum with 0 (:514)
    atomicMax((int *)cell + 1, __float_as_int(c.vel));
    if (acc < 3) cell[2 + acc] = 1.0f;
    return (off - beat_lo_off) | (acc << 28);
}
__device__ __forceinline__ void reset_cell(float *buf, uint32_t word) {
    if (word == kNoCell) return;
    float *cell = buf + (word & 0x0FFFFFFFu);
    const uint32_t acc = word >> 28;
    cell[0] = 0.f; cell[1] = 0.f;
    if (acc < 3) cell[2 + acc] = 0.f;
}

__global__ void __launch_bounds__(kGridThreads, 1)
k_grid(Workspace ws, float *__restrict__ arena) {
    extern __shared__ __align__(128) float smem[];
    float *zero_buf = smem;                                  // never written after init
    float *bufs = smem + kTileFloatsMax;                     // kGridBufs tile buffers
    __shared__ uint32_t touched[kGridBufs][kGridThreads];    // cells each buffer's last tile wrote
    __shared__ uint32_t over_lo[kGridBufs], over_hi[kGridBufs], over_beat[kGridBufs], over_floats[kGridBufs];
    const uint32_t n_tiles = ws.totals[0];
    const TileDesc *tiles = ws.tiles;
    const Cell *cells = ws.cells_sorted;
    for (uint32_t i = threadIdx.x; i < (1 + kGridBufs) * kTileFloatsMax; i += kGridThreads) smem[i] = 0.f;
    for (int b = 0; b < kGridBufs; b++) touched[b][threadIdx.x] = kNoCell;
    if (threadIdx.x < kGridBufs) { over_lo[threadIdx.x] = 0; over_hi[threadIdx.x] = 0; }
    fence_proxy_async_smem();
    __syncthreads();
    uint32_t used = 0;   // tiles that went through a scatter buffer
    // software pipeline: descriptors two tiles ahead, this thread's cell one tile ahead
    const uint32_t stride = gridDim.x;
    TileDesc d0 = load_tile(tiles, blockIdx.x, n_tiles);
    TileDesc d1 = load_tile(tiles, blockIdx.x + stride, n_tiles);
    Cell c0;
    c0.bin = 0; c0.off_acc = 0; c0.dur = 0.f; c0.vel = 0.f;
    if (d0.cell_lo + threadIdx.x < d0.cell_hi) c0 = cells[d0.cell_lo + threadIdx.x];
    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += stride) {
        const TileDesc t = d0;
        const Cell mine = c0;
        const TileDesc d2 = load_tile(tiles, tile + 2 * stride, n_tiles);
        if (d1.cell_lo + threadIdx.x < d1.cell_hi) c0 = cells[d1.cell_lo + threadIdx.x];
        d0 = d1; d1 = d2;
        float *dst = arena + t.dst;
        if (t.cell_hi == t.cell_lo) {
            if (threadIdx.x == 0) { bulk_store_s2g(dst, zero_buf, t.floats * 4u); bulk_commit(); }
            continue;
        }
        const uint32_t b = used % kGridBufs;
        float *buf = bufs + (size_t)b * kTileFloatsMax;
        if (used >= kGridBufs) {
            // the bulk store that last used this buffer must have finished reading it.  Bulk groups are
            // committed in tile order by thread 0; zero-buffer stores in between only make this stricter.
            if (threadIdx.x == 0) bulk_wait_read<kGridBufs - 1>();
            __syncthreads();
            reset_cell(buf, touched[b][threadIdx.x]);
            for (uint32_t i = over_lo[b] + threadIdx.x; i < over_hi[b]; i += kGridThreads) {   // rare: > 256 cells
                const Cell c = cells[i];
                const uint32_t off = c.off_acc & 0x0FFFFFFFu;
                if (off >= over_beat[b] && off < over_beat[b] + over_floats[b])
                    reset_cell(buf, (off - over_beat[b]) | (c.off_acc & 0xF0000000u));
            }
            __syncthreads();
        }
        uint32_t word = kNoCell;
        if (t.cell_lo + threadIdx.x < t.cell_hi) word = apply_cell(buf, mine, t.beat_lo_off, t.floats);
        touched[b][threadIdx.x] = word;
        const uint32_t olo = min(t.cell_lo + (uint32_t)kGridThreads, t.cell_hi);
        for (uint32_t i = olo + threadIdx.x; i < t.cell_hi; i += kGridThreads) apply_cell(buf, cells[i], t.beat_lo_off, t.floats);
        if (threadIdx.x == 0) { over_lo[b] = olo; over_hi[b] = t.cell_hi; over_beat[b] = t.beat_lo_off; over_floats[b] = t.floats; }
        fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) { bulk_store_s2g(dst, buf, t.floats * 4u); bulk_commit(); }
        used++;
    }
    if (threadIdx.x == 0) bulk_wait<0>();
}

// ------------------------------------------------------------------------------------------------
// k_fill: the dense part of qchannel2vchannel / np.stack — 99.7 % of a grid is zeros.  Persistent CTAs stream the batch's
// arena range [0, fill_end) from a read-only zero tile in shared memory with TMA bulk stores (cp.async.bulk shared ->
// global, 32 KB each, one issuing thread per CTA, no barrier in the loop); k_cells<1> then drops the notes in.
// A write-only stream is bounded by the HBM write bandwidth (tools/write_peak_probe.py), not by the copy figure.
// ------------------------------------------------------------------------------------------------
constexpr int kFillThreads = 128;
constexpr uint32_t kFillChunkFloats = 8192;      // 32 KB per bulk store

__global__ void __launch_bounds__(kFillThreads)
k_fill(Workspace ws, float *__restrict__ arena) {
    extern __shared__ __align__(128) float zero_tile[];
    for (uint32_t i = threadIdx.x; i < kFillChunkFloats; i += kFillThreads) zero_tile[i] = 0.f;
    fence_proxy_async_smem();
    __syncthreads();
    if (threadIdx.x != 0) return;
    const unsigned long long total = *reinterpret_cast<const unsigned long long *>(ws.totals + 2);
    const unsigned long long n_chunks = (total + kFillChunkFloats - 1) / kFillChunkFloats;
    for (unsigned long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        const unsigned long long lo = c * kFillChunkFloats;
        const uint32_t fl = (uint32_t)min((unsigned long long)kFillChunkFloats, total - lo);
        bulk_store_s2g(arena + lo, zero_tile, fl * 4u);
        bulk_commit();
    }
    bulk_wait<0>();
}

// k_fill_flat (the default): plain 16-byte stores, one CTA per 16 KB, no loop; the grid covers the arena and CTAs past
// fill_end exit.  Measured on the 512-song batch (39 GB): 5.2 ms = 7.5 TB/s alone, against 6.15 ms for the persistent bulk
// store variant above (MST_FILL_MODE=0) and 5.6-5.9 ms for persistent grid-stride stores at 32-512 CTAs per SM.
__global__ void __launch_bounds__(256)
k_fill_flat(Workspace ws, float *__restrict__ arena) {
    const unsigned long long total4 = *reinterpret_cast<const unsigned long long *>(ws.totals + 2) / 4;
    float4 *dst = reinterpret_cast<float4 *>(arena);
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    const unsigned long long i = (unsigned long long)blockIdx.x * 1024 + threadIdx.x;
#pragma unroll
    for (int u = 0; u < 4; u++) if (i + u * 256 < total4) dst[i + u * 256] = z;
}

}  // namespace

// ================================================================================================
// host side
// ================================================================================================
static size_t carve(Workspace *ws, void *base, uint64_t E, uint32_t S, uint64_t max_bins) {
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = mst_align_up(off + bytes, 256); return base ? (char *)base + o : (char *)nullptr; };
    const uint64_t En = E ? E : 1;
    ws->keys_a = (uint32_t *)take(En * 4); ws->vals_a = (uint32_t *)take(En * 4);
    ws->keys_b = (uint32_t *)take(En * 4); ws->vals_b = (uint32_t *)take(En * 4);
    ws->tmp_c = (uint32_t *)take(En * 4);
    ws->sorted_ev = (MstEvent *)take(En * sizeof(MstEvent));
    ws->msgs = (NoteMsg *)take(En * sizeof(NoteMsg));
    ws->notes = (Note *)take(En * sizeof(Note));
    ws->cells = (Cell *)take(En * sizeof(Cell));
    ws->cells_sorted = (Cell *)take(En * sizeof(Cell));
    ws->seg_start = (uint32_t *)take((size_t)S * kSegStride * 4);
    ws->chans = (ChanRec *)take((size_t)S * 16 * sizeof(ChanRec));
    ws->lists = (ListRec *)take((size_t)S * kMaxLists * sizeof(ListRec));
    ws->chan_slot_list = (int16_t *)take((size_t)S * 256 * 2);
    ws->plans = (SongPlan *)take((size_t)S * sizeof(SongPlan));
    ws->tile_prefix = (uint32_t *)take(((size_t)S + 1) * 4);
    ws->bin_count = (uint32_t *)take((max_bins + 1) * 4);
    ws->bin_start = (uint32_t *)take((max_bins + 1) * 4);
    ws->bin_fill = (uint32_t *)take((max_bins + 1) * 4);
    ws->totals = (uint32_t *)take(16);
    ws->tiles = (TileDesc *)take((max_bins + 1) * sizeof(TileDesc));
    ws->max_bins = max_bins;
    return off;
}

static uint64_t max_bins_for(uint64_t arena_floats) {
    // a bin is one bar of one slot: at least one beat of an unpitched slab
    return arena_floats / MST_UNPITCHED_BEAT_FLOATS + 1;
}

extern "C" size_t mst_encode_workspace_bytes(uint64_t n_events_total, uint32_t n_songs, uint64_t arena_floats) {
    Workspace ws;
    return carve(&ws, nullptr, n_events_total, n_songs ? n_songs : 1, max_bins_for(arena_floats));
}

// optional device-side timing of the two halves of the pipeline (bench.py's roofline numbers)
static bool g_profile = false;
static cudaEvent_t g_ev[4] = {nullptr, nullptr, nullptr, nullptr};   // pipeline start, fill start, fill end, pipeline end
extern "C" int mst_profile_enable(int on) {
    if (on && !g_ev[0]) for (int i = 0; i < 4; i++) MST_CUDA_CHECK(cudaEventCreate(&g_ev[i]));
    g_profile = on != 0;
    return MST_OK;
}
// which: 0 = the whole pipeline (k_sort .. k_cells), 1 = the grid-writing kernel (k_fill_flat, or k_grid on the tile path);
// valid after the stream has been synchronised.  The two overlap (side stream), so they do not add up.
extern "C" float mst_profile_encode_ms(int which) {
    if (!g_ev[0]) return -1.f;
    float ms = -1.f;
    if (cudaEventElapsedTime(&ms, g_ev[which ? 1 : 0], g_ev[which ? 2 : 3]) != cudaSuccess) { cudaGetLastError(); return -1.f; }
    return ms;
}

// high-priority side stream for the latency-bound front-end kernels that run next to the zero fill; fork / join events.  One set per
// process (one process per GPU); the enqueue sequence is serialised so that callers on several host threads
// (utils/parallel.py:34-39 consumes iter_inputs off the main thread) cannot interleave the event records.
static cudaStream_t g_side = nullptr;
static cudaEvent_t g_fork = nullptr, g_offs = nullptr, g_join = nullptr;
static int g_side_dev = -1;
static std::mutex g_enqueue_mutex;
static int side_stream_ready() {
    int dev = 0;
    MST_CUDA_CHECK(cudaGetDevice(&dev));
    if (g_side && dev == g_side_dev) return MST_OK;
    if (g_side) { cudaStreamDestroy(g_side); cudaEventDestroy(g_fork); cudaEventDestroy(g_offs); cudaEventDestroy(g_join); g_side = nullptr; }
    int prio_lo = 0, prio_hi = 0;
    MST_CUDA_CHECK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    MST_CUDA_CHECK(cudaStreamCreateWithPriority(&g_side, cudaStreamNonBlocking, prio_hi));
    MST_CUDA_CHECK(cudaEventCreateWithFlags(&g_fork, cudaEventDisableTiming));
    MST_CUDA_CHECK(cudaEventCreateWithFlags(&g_offs, cudaEventDisableTiming));
    MST_CUDA_CHECK(cudaEventCreateWithFlags(&g_join, cudaEventDisableTiming));
    g_side_dev = dev;
    return MST_OK;
}

static int g_num_sms = 0;
static int num_sms() {
    if (!g_num_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_num_sms <= 0) g_num_sms = 148;
    }
    return g_num_sms;
}

extern "C" int mst_encode_batch_device(const MstEvent *d_events, uint64_t n_events_total,
                                       const MstSongIn *d_songs, uint32_t n_songs,
                                       const MstFilter *h_filter, MstSongInfo *d_info,
                                       float *d_arena, uint64_t arena_floats,
                                       void *d_workspace, size_t workspace_bytes, void *stream_) {
    return mst_encode_batch_device_front(d_events, n_events_total, d_songs, n_songs, h_filter, d_info, d_arena, arena_floats,
                                         d_workspace, workspace_bytes, stream_, nullptr, nullptr);
}

// The same with a hook that enqueues producer kernels of d_events / d_songs (the device-side SMF parser) on the stream the
// front end runs on - the library's high-priority side stream when the fill overlaps - right in front of k_sort.
int mst_encode_batch_device_front(const MstEvent *d_events, uint64_t n_events_total,
                                  const MstSongIn *d_songs, uint32_t n_songs,
                                  const MstFilter *h_filter, MstSongInfo *d_info,
                                  float *d_arena, uint64_t arena_floats,
                                  void *d_workspace, size_t workspace_bytes, void *stream_,
                                  int (*front_hook)(void *ctx, cudaStream_t front), void *hook_ctx) {
    if (!d_songs || !d_info || !d_workspace || (!d_events && n_events_total) || (!d_arena && arena_floats)) {
        mst_set_error("mst_encode_batch_device: null pointer argument");
        return MST_E_INVALID;
    }
    if (n_songs == 0) return MST_OK;
    if (reinterpret_cast<uintptr_t>(d_arena) & 15u) {
        mst_set_error("mst_encode_batch_device: the arena must be 16-byte aligned");
        return MST_E_INVALID;
    }
    if (n_events_total >= 0xFFFFFFFFull) {
        mst_set_error("mst_encode_batch_device: at most 2^32-1 events per batch");
        return MST_E_INVALID;
    }
    Workspace ws;
    const size_t need = carve(&ws, d_workspace, n_events_total, n_songs, max_bins_for(arena_floats));
    if (need > workspace_bytes) {
        mst_set_error("mst_encode_batch_device: workspace too small (%zu < %zu)", workspace_bytes, need);
        return MST_E_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    MstFilter filter;
    if (h_filter) filter = *h_filter; else { memset(&filter, 0, sizeof(filter)); }
    const int sms = num_sms();

    static const bool tiles_path = getenv("MST_GRID_TILES") != nullptr;      // comparison path: shared-memory tile writer
    static const int fill_mode = getenv("MST_FILL_MODE") ? atoi(getenv("MST_FILL_MODE")) : 2;
    static const bool overlap = getenv("MST_ENCODE_NO_OVERLAP") == nullptr;
    std::lock_guard<std::mutex> lock(g_enqueue_mutex);
    if (g_profile) MST_CUDA_CHECK(cudaEventRecord(g_ev[0], stream));
    // The zero fill needs only k_offsets' result and streams at the HBM write bandwidth; every other kernel but the final
    // note scatter is latency-bound.  Those run on a HIGH-PRIORITY side stream, so that their CTAs are picked first whenever
    // a fill's short-lived CTAs retire: k_walk<1> / k_pair / k_scale hide behind this batch's fill, and when the caller
    // alternates two streams (bench.py, iter_inputs' prefetch) the next batch's k_sort .. k_offsets hide behind it too.
    // The fill and k_cells stay on the caller's stream; k_cells waits for both.
    const bool fill = !tiles_path && arena_floats != 0;
    const bool side = fill && overlap;
    cudaStream_t front = stream;
    if (side) {
        if (side_stream_ready() != MST_OK) return MST_E_CUDA;
        MST_CUDA_CHECK(cudaEventRecord(g_fork, stream));
        MST_CUDA_CHECK(cudaStreamWaitEvent(g_side, g_fork, 0));
        front = g_side;
    }
    if (front_hook) {
        const int hrc = front_hook(hook_ctx, front);
        if (hrc != MST_OK) return hrc;
    }
    k_sort<<<n_songs, kSortThreads, 0, front>>>(d_events, d_songs, ws);
    MST_LAUNCH_CHECK("k_sort");
    const uint32_t walk_blocks = (uint32_t)std::min<uint64_t>(((uint64_t)n_songs * 16u + 3) / 4, (uint64_t)sms * 16);
    k_walk<0><<<walk_blocks, 128, 0, front>>>(d_songs, n_songs, ws, d_info);
    MST_LAUNCH_CHECK("k_walk<0>");
    const uint32_t info_blocks = (uint32_t)std::min<uint64_t>(((uint64_t)n_songs + 3) / 4, (uint64_t)sms * 16);
    k_info<<<info_blocks, 128, 0, front>>>(d_songs, n_songs, ws, filter, d_info);
    MST_LAUNCH_CHECK("k_info");
    k_offsets<<<1, 1024, 0, front>>>(n_songs, ws, d_info, arena_floats);
    MST_LAUNCH_CHECK("k_offsets");
    if (fill) {
        if (side) {
            MST_CUDA_CHECK(cudaEventRecord(g_offs, g_side));
            MST_CUDA_CHECK(cudaStreamWaitEvent(stream, g_offs, 0));
        }
        if (g_profile) MST_CUDA_CHECK(cudaEventRecord(g_ev[1], stream));
        if (fill_mode == 0) k_fill<<<sms * 2, kFillThreads, kFillChunkFloats * sizeof(float), stream>>>(ws, d_arena);
        else k_fill_flat<<<(unsigned)((arena_floats / 4 + 1023) / 1024), 256, 0, stream>>>(ws, d_arena);
        MST_LAUNCH_CHECK("k_fill");
        if (g_profile) MST_CUDA_CHECK(cudaEventRecord(g_ev[2], stream));
    }
    k_walk<1><<<walk_blocks, 128, 0, front>>>(d_songs, n_songs, ws, d_info);
    MST_LAUNCH_CHECK("k_walk<1>");
    k_pair<<<n_songs, 256, 0, front>>>(d_songs, ws, d_info);
    MST_LAUNCH_CHECK("k_pair");
    k_scale<<<n_songs, kScaleThreads, 0, front>>>(d_songs, ws, d_info);
    MST_LAUNCH_CHECK("k_scale");
    if (side) MST_CUDA_CHECK(cudaEventRecord(g_join, g_side));
    if (tiles_path) {
        k_cells<0><<<n_songs, 256, 0, stream>>>(d_songs, ws, d_info, d_arena);
        MST_LAUNCH_CHECK("k_cells");
        const size_t grid_smem = (size_t)(1 + kGridBufs) * kTileFloatsMax * sizeof(float);
        static bool attr_set = false;
        if (!attr_set) {
            MST_CUDA_CHECK(cudaFuncSetAttribute(k_grid, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)grid_smem));
            attr_set = true;
        }
        if (g_profile) MST_CUDA_CHECK(cudaEventRecord(g_ev[1], stream));
        k_grid<<<sms, kGridThreads, grid_smem, stream>>>(ws, d_arena);
        MST_LAUNCH_CHECK("k_grid");
        if (g_profile) MST_CUDA_CHECK(cudaEventRecord(g_ev[2], stream));
    } else {
        if (side) MST_CUDA_CHECK(cudaStreamWaitEvent(stream, g_join, 0));
        k_cells<1><<<n_songs, 256, 0, stream>>>(d_songs, ws, d_info, d_arena);
        MST_LAUNCH_CHECK("k_cells");
    }
    if (g_profile) {
        if (!fill && !tiles_path) { MST_CUDA_CHECK(cudaEventRecord(g_ev[1], stream)); MST_CUDA_CHECK(cudaEventRecord(g_ev[2], stream)); }
        MST_CUDA_CHECK(cudaEventRecord(g_ev[3], stream));
    }
    return MST_OK;
}
