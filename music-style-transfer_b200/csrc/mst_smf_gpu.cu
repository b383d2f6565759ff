// mst_smf_gpu.cu — Standard MIDI File bytes -> MstEvent records ON THE DEVICE, and the encode entry point that takes raw
// file bytes as its wire format (mst_encode_batch_smf_async).
//
// Replaces mido.MidiFile(path) of load_midi_from_file (style/midi.py:104-108) for a whole batch of files, with the parsing
// rules of the host parser mst_smf_parse (csrc/mst_smf.cpp, which states them against mido >= 1.3): the same events in the
// same order for every file that loads, and "load failed" for exactly the files mst_smf_parse rejects.
//
// Why: the host-buffer path is bound by the host -> device copy once several ranks share one host (8 GPUs: 13.6 - 18.3 GB/s
// per rank while the kernels run, profiles/r2_h2d_8gpu.txt).  A 16-byte MstEvent is 4 - 5x the bytes the file spends on the
// event (delta-time varint + running status), so the file itself is the compact wire format, and the parse - which bounded
// the file-fed path on the host (71 us per file per core) - becomes about a millisecond of latency-bound device work
// per 512-file batch, most of which overlaps the zero fill of another slot (DESIGN.md section 4 has the measurements).
//
// Work decomposition.  A file is a header chunk and n_tracks track chunks; a track chunk can only be parsed sequentially
// (variable-length delta times, running status), but tracks are independent once their byte ranges are known, and those
// follow from the chunk headers alone when the file is well-formed (a track whose events do not end exactly at its chunk
// end makes mst_smf_parse fail: it reads on until the end of the file).  So:
//   k_smf_tracks   one thread per file: header, walk of the chunk headers -> (file, start, size) per track
//   k_smf_parse    one thread per track: 16-byte MstEvent records (one 16-byte store each) into a staging area indexed by the
//                  track's byte position, the record count, and a flag on the file when the track is malformed
//   k_smf_offsets  exclusive scan of the counts (tracks of failed files count 0) -> record offsets, the MstSongIn table
//   k_smf_compact  one warp per track: staging -> the dense, track-major record array the encoder reads
//   k_smf_status   after the encoder: MST_SONG_LOAD_FAILED for the files that did not load
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <utility>
#include <vector>

#include "mst_device.cuh"

namespace {

struct TrackRec {
    uint32_t file;
    uint32_t index;      // track number inside the file (MstEvent.track)
    uint32_t start;      // byte offset of the first event, relative to the file
    uint32_t size;       // chunk size from the MTrk header
};

struct SmfWorkspace {
    uint8_t *bytes;            // the batch's files, concatenated
    MstFileIn *files;          // [F]  (device copy; k_smf_tracks voids descriptors that point outside the buffer)
    uint32_t *track_base;      // [F + 1]
    int32_t *division;         // [F]
    int32_t *file_fail;        // [F]  != 0: load failed
    TrackRec *tracks;          // [max_tracks]
    uint32_t *track_count;     // [max_tracks]
    uint32_t *track_offset;    // [max_tracks + 1]
    MstSongIn *songs;          // [F]
    MstEvent *events;          // [max_events]  the batch's records, track-major per file, dense
    MstEvent *staging;         // [max_events + 8]  where the parser writes: track at record index (its first byte's offset) / 2
    uint32_t *totals;          // [0] tracks, [1] events
    uint64_t max_tracks, max_events;
    uint64_t n_bytes;          // size of the batch buffer: a file must lie inside it
};

__device__ __forceinline__ uint32_t be32(const uint8_t *b) {
    return ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3];
}

// ------------------------------------------------------------------------------------------------
// k_smf_tracks: one CTA for the batch.  Pass 1 reads the headers (mst_smf.cpp:69-82), a block scan places each file's
// tracks in the track table, pass 2 walks the chunk headers (:96-101).  A file that is cut short of its last chunk headers
// keeps the tracks found so far and is flagged (the sequential parser fails on it at that point).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1)
k_smf_tracks(SmfWorkspace ws, uint32_t n_files) {
    __shared__ unsigned scan_tmp[33];
    __shared__ unsigned s_running;
    if (threadIdx.x == 0) s_running = 0;
    __syncthreads();
    for (uint32_t f0 = 0; f0 < n_files; f0 += blockDim.x) {
        const uint32_t f = f0 + threadIdx.x;
        uint32_t ntr = 0, first = 0;
        int fail = 0, division = 0;
        const uint8_t *p = nullptr;
        uint32_t n = 0;
        if (f < n_files) {
            const MstFileIn fi = ws.files[f];
            const bool inside = fi.byte_offset <= ws.n_bytes && fi.n_bytes <= ws.n_bytes - fi.byte_offset;
            p = ws.bytes + (inside ? fi.byte_offset : 0);
            n = inside ? fi.n_bytes : 0;                              // a descriptor that points outside the buffer loads nothing
            if (!inside) ws.files[f].n_bytes = 0;                     // (the parser kernels read the descriptor again)
            if (n < 8) fail = 1;
            else if (p[0] != 'M' || p[1] != 'T' || p[2] != 'h' || p[3] != 'd') fail = 2;
            else {
                const uint32_t hsize = be32(p + 4);
                const uint32_t avail = n - 8 < hsize ? n - 8 : hsize;
                if (avail < 6) fail = 1;
                else {
                    const int nt = (int16_t)((p[10] << 8) | p[11]);
                    division = (int16_t)((p[12] << 8) | p[13]);
                    first = 8 + avail;
                    // a chunk header takes 8 bytes: more tracks than that cannot be found (the walk below flags the file)
                    ntr = nt > 0 ? min((uint32_t)nt, (n - first) / 8u + 1u) : 0u;
                }
            }
        }
        unsigned total;
        const unsigned ex = block_exclusive_scan(ntr, scan_tmp, &total);
        const uint32_t base = s_running + ex;
        if (f < n_files) {
            ws.track_base[f] = base;
            ws.division[f] = division;
            uint64_t pos = first;
            for (uint32_t t = 0; t < ntr; t++) {
                TrackRec r{f, t, 0u, 0u};
                if (!fail) {
                    if (pos + 8 > n) fail = 1;
                    else if (p[pos] != 'M' || p[pos + 1] != 'T' || p[pos + 2] != 'r' || p[pos + 3] != 'k') fail = 3;
                    else {
                        r.size = be32(p + pos + 4);
                        r.start = (uint32_t)(pos + 8);
                        pos += 8ull + r.size;
                    }
                }
                if (fail) r.start = 0xFFFFFFFFu;                     // not a track: parsed as nothing
                if (base + t < ws.max_tracks) ws.tracks[base + t] = r;
            }
            // the header promised more tracks than fit the table bound: they cannot exist in n bytes either
            if (!fail && ntr > 0) {
                const int nt = (int16_t)((p[10] << 8) | p[11]);
                if ((uint32_t)nt > ntr) fail = 1;
            }
            ws.file_fail[f] = fail;
        }
        __syncthreads();
        if (threadIdx.x == 0) s_running += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        ws.track_base[n_files] = s_running;
        ws.totals[0] = s_running;
    }
}

// ------------------------------------------------------------------------------------------------
// The track parser (mst_smf.cpp:102-196), one thread per track.  EMIT = 1 writes the records; EMIT = 0 only counts
// them (not instantiated since records go to a staging area in ONE pass, see k_smf_parse).
// ------------------------------------------------------------------------------------------------
struct Rd {
    const uint8_t *p;
    uint32_t n, pos;
    bool eof;
    __device__ __forceinline__ int byte() {
        if (pos >= n) { eof = true; return 0; }
        return p[pos++];
    }
    __device__ __forceinline__ uint64_t varint() {
        uint64_t v = 0;
        while (true) {
            const int b = byte();
            if (eof) return 0;
            v = v >= (1ull << 56) ? (1ull << 62) : ((v << 7) | (uint64_t)(b & 0x7F));
            if (b < 0x80) return v;
        }
    }
    __device__ __forceinline__ bool skip(uint64_t k) {
        if (k > (uint64_t)(n - pos)) { eof = true; pos = n; return false; }
        pos += (uint32_t)k;
        return true;
    }
};

template <int EMIT>
__device__ __forceinline__ void put(MstEvent *out, uint32_t &count, uint64_t tick, uint32_t kind, uint32_t ch, uint32_t d1,
                                    uint32_t d2, uint32_t aux, uint32_t track) {
    if (tick > MST_MAX_TICK) return;                                 // mst_smf.cpp:86
    if (EMIT) {
        uint4 w;                                                     // MstEvent layout, one 16-byte store
        w.x = (uint32_t)tick;
        w.y = aux;
        w.z = kind | (ch << 8) | (d1 << 16) | (d2 << 24);
        w.w = track > 65535u ? 65535u : track;
        reinterpret_cast<uint4 *>(out)[count] = w;
    }
    count++;
}

// returns 0 when the track ends exactly at its chunk end, else a LoadFail code of mst_smf.cpp
constexpr int kSmfBuf = 256;     // bytes of a track held in shared memory per lane

// `buf`: this lane's column of the warp's staging area, 4-byte word j of the lane at buf[j * 32] (lane-interleaved)
template <int EMIT>
__device__ int parse_track(const uint8_t *file, uint32_t n, const TrackRec tr, MstEvent *out, uint32_t *n_out, uint32_t *buf) {
    uint32_t buf_rel = tr.start + 0x40000000u;   // file-relative position of the first buffered byte, modulo 2^32 (nothing buffered yet)
    Rd r{file, n, tr.start, false};
    const uint32_t start = tr.start, size = tr.size, t = tr.index;
    int last_status = -1;
    uint64_t tick = 0;
    uint32_t count = 0;
    while (r.pos - start != size) {
        if (r.pos - start > size) return 1;                          // ran past the chunk end: the sequential parser reads on to EOF
        // ---- fast path: a channel voice message (99 % of a file) whose delta time takes at most four bytes lies inside one
        // 8-byte window: three 4-byte words from shared memory, then funnel shifts and selects only - no per-byte load latency, and the 32 tracks
        // of a warp run the same instructions whatever their message types.  Everything else (meta, sysex, system messages,
        // long delta times, the last 7 bytes of a file, malformed input) takes the byte-wise path below, which starts over at
        // the same position; the two paths emit the same records (the byte-wise one is mst_smf.cpp's loop).
        if (r.pos + 8 <= n) {
            // the window comes from the lane's 256-byte buffer in shared memory (4-byte words, lane-interleaved: conflict-free),
            // refilled by sixteen independent 16-byte loads: one global-memory latency per ~60 messages instead of one per
            // message (next to a zero fill that saturates the memory system, a dependent global load costs microseconds)
            uint32_t off = r.pos - buf_rel;                          // (wraps to a huge value when pos lies in front of the buffer)
            if (off > (uint32_t)(kSmfBuf - 8)) {
                const uintptr_t addr = reinterpret_cast<uintptr_t>(file + r.pos);
                const uint4 *g = reinterpret_cast<const uint4 *>(addr & ~(uintptr_t)15);
                off = (uint32_t)(addr & 15);
                buf_rel = r.pos - off;
                uint4 v[kSmfBuf / 16];
#pragma unroll
                for (int k = 0; k < kSmfBuf / 16; k++) v[k] = g[k];
#pragma unroll
                for (int k = 0; k < kSmfBuf / 16; k++) {
                    buf[(4 * k + 0) * 32] = v[k].x; buf[(4 * k + 1) * 32] = v[k].y;
                    buf[(4 * k + 2) * 32] = v[k].z; buf[(4 * k + 3) * 32] = v[k].w;
                }
            }
            const uint32_t wj = off >> 2, sh = (off & 3) * 8;
            const uint32_t w0 = buf[wj * 32], w1 = buf[(wj + 1) * 32], w2 = buf[(wj + 2) * 32];
            const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);   // bytes 0..3 and 4..7 of the window
            const uint32_t term = ~lo & 0x80808080u;                 // bit 7 of the first delta byte that ends the varint
            if (term != 0) {
                const uint32_t nd = (uint32_t)__ffs((int)term) >> 3;    // bytes of the delta time: 1..4
                const uint32_t b0 = lo & 0x7F, b1 = (lo >> 8) & 0x7F, b2 = (lo >> 16) & 0x7F, b3 = (lo >> 24) & 0x7F;
                const uint32_t v1 = b0, v2 = (v1 << 7) | b1, v3 = (v2 << 7) | b2, v4 = (v3 << 7) | b3;
                const uint32_t delta = nd == 1 ? v1 : (nd == 2 ? v2 : (nd == 3 ? v3 : v4));
                uint32_t w = nd == 4 ? hi : __funnelshift_r(lo, hi, 8 * nd);     // the four bytes behind the delta time
                const uint32_t sb = w & 0xFF;
                const bool running = sb < 0x80;
                const int status = running ? last_status : (int)sb;
                if (status >= 0x80 && status < 0xF0) {               // (running status without a status, or >= 0xF0: byte-wise path)
                    if (!running) w >>= 8;
                    const uint32_t need = (status & 0xE0) == 0xC0 ? 1u : 2u;
                    const uint32_t d0 = w & 0xFF, d1 = need == 2 ? ((w >> 8) & 0xFF) : 0u;
                    if (d0 > 127 || d1 > 127) return 5;
                    tick = ((uint64_t)delta > (1ull << 62) - tick) ? (1ull << 62) : tick + delta;
                    last_status = status;
                    const uint32_t kind = (uint32_t)((status >> 4) - 8);
                    const bool payload = kind == MST_EV_NOTE_OFF || kind == MST_EV_NOTE_ON || kind == MST_EV_CONTROL_CHANGE ||
                                         kind == MST_EV_PROGRAM_CHANGE;
                    put<EMIT>(out, count, tick, kind, (uint32_t)(status & 0x0F), payload ? d0 : 0u, payload ? d1 : 0u, 0, t);
                    r.pos += nd + (running ? 0u : 1u) + need;
                    continue;
                }
            }
        }
        const uint64_t delta = r.varint();
        if (r.eof) return 1;
        tick = (delta > (1ull << 62) - tick) ? (1ull << 62) : tick + delta;
        int status = r.byte();
        if (r.eof) return 1;
        int peek = -1;
        if (status < 0x80) {
            if (last_status < 0) return 4;
            peek = status;
            status = last_status;
        } else if (status != 0xFF) {
            last_status = status;
        }
        if (status == 0xFF) {
            const int type = r.byte();
            const uint64_t len = r.varint();
            if (r.eof) return 1;
            const uint32_t dpos = r.pos;
            if (!r.skip(len)) return 1;
            const uint8_t *d = file + dpos;
            switch (type) {
                case 0x51:
                    if (len < 3) return 8;
                    put<EMIT>(out, count, tick, MST_EV_SET_TEMPO, 0, 0, 0, ((uint32_t)d[0] << 16) | ((uint32_t)d[1] << 8) | d[2], t);
                    break;
                case 0x58:
                    if (len < 4) return 8;
                    put<EMIT>(out, count, tick, MST_EV_TIME_SIGNATURE, 0, d[0], d[1], 0, t);
                    break;
                case 0x59: {
                    if (len < 2) return 8;
                    const int sf = (int8_t)d[0], mi = d[1];
                    if (sf < -7 || sf > 7 || mi > 1) return 7;
                    put<EMIT>(out, count, tick, MST_EV_KEY_SIGNATURE, 0, 0, 0, (uint32_t)((sf + 7) * 2 + mi), t);
                    break;
                }
                case 0x20:
                    if (len < 1) return 8;
                    put<EMIT>(out, count, tick, MST_EV_CHANNEL_PREFIX, d[0], 0, 0, 0, t);
                    break;
                case 0x09:
                    put<EMIT>(out, count, tick, MST_EV_UNKNOWN, 0, 0, 0, 0x0900u, t);
                    break;
                default:
                    break;
            }
        } else if (status == 0xF0 || status == 0xF7) {
            const uint64_t len = r.varint();
            if (r.eof) return 1;
            const uint32_t dpos = r.pos;
            if (!r.skip(len)) return 1;
            uint32_t a = dpos, b = dpos + (uint32_t)len;
            if (b > a && file[a] == 0xF0) a++;
            if (b > a && file[b - 1] == 0xF7) b--;
            for (uint32_t i = a; i < b; i++) if (file[i] > 127) return 9;
        } else if (status < 0xF0) {
            const int hi = status & 0xF0;
            const int need = (hi == 0xC0 || hi == 0xD0) ? 1 : 2;
            int d0 = 0, d1 = 0, have = 0;
            if (peek >= 0) { d0 = peek; have = 1; }
            while (have < need) {
                const int b = r.byte();
                if (r.eof) return 1;
                if (have == 0) d0 = b; else d1 = b;
                have++;
            }
            if (d0 > 127 || (need == 2 && d1 > 127)) return 5;
            const uint32_t kind = (uint32_t)((hi >> 4) - 8);
            const bool payload = kind == MST_EV_NOTE_OFF || kind == MST_EV_NOTE_ON || kind == MST_EV_CONTROL_CHANGE ||
                                 kind == MST_EV_PROGRAM_CHANGE;
            put<EMIT>(out, count, tick, kind, (uint32_t)(status & 0x0F), payload ? (uint32_t)d0 : 0u,
                      payload && need == 2 ? (uint32_t)d1 : 0u, 0, t);
        } else {
            int ndata;
            switch (status) {
                case 0xF1: case 0xF3: ndata = 1; break;
                case 0xF2: ndata = 2; break;
                case 0xF6: case 0xF8: case 0xFA: case 0xFB: case 0xFC: case 0xFE: ndata = 0; break;
                default: return 6;
            }
            if (peek >= 0) return 6;
            for (int i = 0; i < ndata; i++) {
                const int b = r.byte();
                if (r.eof) return 1;
                if (b > 127) return 5;
            }
            put<EMIT>(out, count, tick, MST_EV_UNKNOWN, 0, 0, 0, (uint32_t)status, t);
        }
    }
    *n_out = count;
    return 0;
}

// A record takes at least two bytes of its track, and the byte ranges of tracks are disjoint, so the track that starts at byte
// B of the batch buffer can write its records from staging index (B + 1) / 2 on without reaching the next track's: ONE parse
// pass (no counting pass in front of it); k_smf_compact then moves the records to their dense places.
__device__ __forceinline__ uint64_t staging_index(const MstFileIn &fi, const TrackRec &tr) { return (fi.byte_offset + tr.start + 1) / 2; }

__global__ void __launch_bounds__(32)
k_smf_parse(SmfWorkspace ws) {
    __shared__ uint32_t s_buf[(kSmfBuf / 4 + 2) * 32];              // (+2 words: the window reads one word past its last byte)
    const uint32_t n_tracks = min((uint64_t)ws.totals[0], ws.max_tracks);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_tracks; i += gridDim.x * blockDim.x) {
        const TrackRec tr = ws.tracks[i];
        uint32_t count = 0;
        int rc = 0;
        if (tr.start != 0xFFFFFFFFu) {
            const MstFileIn fi = ws.files[tr.file];
            rc = parse_track<1>(ws.bytes + fi.byte_offset, fi.n_bytes, tr, ws.staging + staging_index(fi, tr), &count, s_buf + threadIdx.x);
        }
        ws.track_count[i] = rc ? 0u : count;
        if (rc) atomicExch(&ws.file_fail[tr.file], rc);
    }
}

// one warp per track: staging -> dense (16-byte records, coalesced)
__global__ void __launch_bounds__(256)
k_smf_compact(SmfWorkspace ws) {
    const uint32_t n_tracks = min((uint64_t)ws.totals[0], ws.max_tracks);
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, n_warps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i = warp; i < n_tracks; i += n_warps) {
        const TrackRec tr = ws.tracks[i];
        if (tr.start == 0xFFFFFFFFu || ws.file_fail[tr.file]) continue;
        const uint32_t count = ws.track_count[i];
        const uint4 *src = reinterpret_cast<const uint4 *>(ws.staging + staging_index(ws.files[tr.file], tr));
        uint4 *dst = reinterpret_cast<uint4 *>(ws.events + ws.track_offset[i]);
        for (uint32_t k = lane_id(); k < count; k += 32) dst[k] = src[k];
    }
}

// ------------------------------------------------------------------------------------------------
// k_smf_offsets: one CTA.  Exclusive scan of the per-track record counts (0 for the tracks of failed files) and the
// MstSongIn table (event_offset, n_events, ticks_per_beat = the header's division, mst_smf.cpp:79-81).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(1024, 1)
k_smf_offsets(SmfWorkspace ws, uint32_t n_files) {
    __shared__ unsigned scan_tmp[33];
    __shared__ unsigned s_running;
    const uint32_t n_tracks = min((uint64_t)ws.totals[0], ws.max_tracks);
    if (threadIdx.x == 0) s_running = 0;
    __syncthreads();
    for (uint32_t i0 = 0; i0 < n_tracks; i0 += blockDim.x) {
        const uint32_t i = i0 + threadIdx.x;
        unsigned v = 0;
        if (i < n_tracks && !ws.file_fail[ws.tracks[i].file]) v = ws.track_count[i];
        unsigned total;
        const unsigned ex = block_exclusive_scan(v, scan_tmp, &total);
        if (i < n_tracks) ws.track_offset[i] = s_running + ex;
        __syncthreads();
        if (threadIdx.x == 0) s_running += total;
        __syncthreads();
    }
    if (threadIdx.x == 0) { ws.track_offset[n_tracks] = s_running; ws.totals[1] = s_running; }
    __syncthreads();
    for (uint32_t f = threadIdx.x; f < n_files; f += blockDim.x) {
        const uint32_t lo = min(ws.track_base[f], n_tracks), hi = min(ws.track_base[f + 1], n_tracks);
        MstSongIn s;
        s.event_offset = ws.track_offset[lo];
        s.n_events = ws.file_fail[f] ? 0u : ws.track_offset[hi] - ws.track_offset[lo];
        s.ticks_per_beat = ws.division[f];
        ws.songs[f] = s;
    }
}

__global__ void __launch_bounds__(256)
k_smf_status(const int32_t *__restrict__ file_fail, MstSongInfo *info, uint32_t n_files) {
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < n_files && file_fail[f]) info[f].status = MST_SONG_LOAD_FAILED;
}

// max records a batch of n_bytes can hold: an event takes at least two bytes (delta time + one data byte under running status)
inline uint64_t smf_max_events(uint64_t n_bytes) { return n_bytes / 2 + 1; }
inline uint64_t smf_max_tracks(uint64_t n_bytes, uint32_t n_files) { return n_bytes / 8 + n_files + 1; }

size_t smf_carve(SmfWorkspace *ws, void *base, uint64_t n_bytes, uint32_t F) {
    const uint64_t E = smf_max_events(n_bytes), T = smf_max_tracks(n_bytes, F);
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off = mst_align_up(off + bytes, 256); return (char *)base + o; };
    ws->bytes = (uint8_t *)take(n_bytes ? n_bytes : 1);
    ws->files = (MstFileIn *)take((size_t)F * sizeof(MstFileIn));
    ws->track_base = (uint32_t *)take((size_t)(F + 1) * 4);
    ws->division = (int32_t *)take((size_t)F * 4);
    ws->file_fail = (int32_t *)take((size_t)F * 4);
    ws->tracks = (TrackRec *)take(T * sizeof(TrackRec));
    ws->track_count = (uint32_t *)take(T * 4);
    ws->track_offset = (uint32_t *)take((T + 1) * 4);
    ws->songs = (MstSongIn *)take((size_t)F * sizeof(MstSongIn));
    ws->events = (MstEvent *)take(E * sizeof(MstEvent));
    ws->staging = (MstEvent *)take((E + 8) * sizeof(MstEvent));
    ws->totals = (uint32_t *)take(64);
    ws->max_tracks = T; ws->max_events = E; ws->n_bytes = n_bytes;
    return off;
}

}  // namespace

// one high-priority stream (+ fork / join events) per device and caller stream
struct ParseStream { cudaStream_t s; cudaEvent_t fork, join; };
static ParseStream *parse_stream_for(cudaStream_t caller) {
    static std::mutex mu;
    static std::vector<std::pair<std::pair<int, cudaStream_t>, ParseStream *>> table;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    for (auto &kv : table) if (kv.first.first == dev && kv.first.second == caller) return kv.second;
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);                      // hi = greatest priority (numerically lowest)
    ParseStream *ps = new ParseStream();
    if (cudaStreamCreateWithPriority(&ps->s, cudaStreamNonBlocking, hi) != cudaSuccess ||
        cudaEventCreateWithFlags(&ps->fork, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&ps->join, cudaEventDisableTiming) != cudaSuccess) { delete ps; return nullptr; }
    table.push_back({{dev, caller}, ps});
    return ps;
}

// the copies of the batch on `stream`
static int smf_upload(const SmfWorkspace &ws, const uint8_t *h_bytes, uint64_t n_bytes_total, const MstFileIn *h_files,
                      uint32_t n_files, cudaStream_t stream) {
    if (n_bytes_total) MST_CUDA_CHECK(cudaMemcpyAsync(ws.bytes, h_bytes, n_bytes_total, cudaMemcpyHostToDevice, stream));
    MST_CUDA_CHECK(cudaMemcpyAsync(ws.files, h_files, (size_t)n_files * sizeof(MstFileIn), cudaMemcpyHostToDevice, stream));
    return MST_OK;
}

// the four parser kernels on `stream`
static int smf_parse_enqueue(const SmfWorkspace &ws, uint32_t n_files, cudaStream_t stream) {
    k_smf_tracks<<<1, 1024, 0, stream>>>(ws, n_files);
    MST_LAUNCH_CHECK("k_smf_tracks");
    // thread per track: the track count lives on the device, the kernels loop over it
    // (one warp per CTA: the few hundred warps of a batch spread over all SMs)
    const unsigned blocks = (unsigned)std::min<uint64_t>((ws.max_tracks + 31) / 32, 148 * 32);
    k_smf_parse<<<blocks, 32, 0, stream>>>(ws);
    MST_LAUNCH_CHECK("k_smf_parse");
    k_smf_offsets<<<1, 1024, 0, stream>>>(ws, n_files);
    MST_LAUNCH_CHECK("k_smf_offsets");
    k_smf_compact<<<148 * 8, 256, 0, stream>>>(ws);
    MST_LAUNCH_CHECK("k_smf_compact");
    return MST_OK;
}

extern "C" size_t mst_smf_parse_device_workspace_bytes(uint64_t n_bytes_total, uint32_t n_files) {
    SmfWorkspace ws;
    return mst_align_up(smf_carve(&ws, nullptr, n_bytes_total, n_files ? n_files : 1), 256);
}

extern "C" int mst_smf_parse_batch_device(const uint8_t *h_bytes, uint64_t n_bytes_total, const MstFileIn *h_files, uint32_t n_files,
                                          MstEvent *h_events, uint64_t capacity, MstSongIn *h_songs, int32_t *h_status,
                                          uint64_t *n_events_total, void *d_workspace, size_t workspace_bytes, void *stream_) {
    if (mst_device_count() == 0) { mst_set_error("no CUDA device: libmst_b200 has no CPU fallback"); return MST_E_NO_DEVICE; }
    if (!h_files || !h_songs || !h_status || !n_events_total || !d_workspace || (!h_bytes && n_bytes_total) || (!h_events && capacity)) {
        mst_set_error("mst_smf_parse_batch_device: null pointer argument");
        return MST_E_INVALID;
    }
    *n_events_total = 0;
    if (n_files == 0) return MST_OK;
    if (n_bytes_total >= 0xFFFFFFFFull) { mst_set_error("mst_smf_parse_batch_device: at most 2^32-1 bytes of files per batch"); return MST_E_INVALID; }
    if (mst_smf_parse_device_workspace_bytes(n_bytes_total, n_files) > workspace_bytes) {
        mst_set_error("mst_smf_parse_batch_device: workspace too small");
        return MST_E_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    SmfWorkspace ws;
    smf_carve(&ws, d_workspace, n_bytes_total, n_files);
    int rc = smf_upload(ws, h_bytes, n_bytes_total, h_files, n_files, stream);
    if (rc == MST_OK) rc = smf_parse_enqueue(ws, n_files, stream);
    if (rc != MST_OK) return rc;
    uint32_t totals[2] = {0, 0};
    MST_CUDA_CHECK(cudaMemcpyAsync(totals, ws.totals, sizeof(totals), cudaMemcpyDeviceToHost, stream));
    MST_CUDA_CHECK(cudaMemcpyAsync(h_songs, ws.songs, (size_t)n_files * sizeof(MstSongIn), cudaMemcpyDeviceToHost, stream));
    MST_CUDA_CHECK(cudaMemcpyAsync(h_status, ws.file_fail, (size_t)n_files * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
    MST_CUDA_CHECK(cudaStreamSynchronize(stream));
    *n_events_total = totals[1];
    if (totals[1] > capacity) return MST_E_CAPACITY;
    if (totals[1]) MST_CUDA_CHECK(cudaMemcpyAsync(h_events, ws.events, (size_t)totals[1] * sizeof(MstEvent), cudaMemcpyDeviceToHost, stream));
    MST_CUDA_CHECK(cudaStreamSynchronize(stream));
    return MST_OK;
}

extern "C" size_t mst_encode_smf_workspace_bytes(uint64_t n_bytes_total, uint32_t n_files, uint64_t arena_floats) {
    SmfWorkspace ws;
    const uint32_t F = n_files ? n_files : 1;
    return mst_align_up(mst_encode_workspace_bytes(smf_max_events(n_bytes_total), F, arena_floats), 256) +
           mst_align_up(smf_carve(&ws, nullptr, n_bytes_total, F), 256) + mst_align_up((size_t)F * sizeof(MstSongInfo), 256);
}

extern "C" int mst_encode_batch_smf_async(const uint8_t *h_bytes, uint64_t n_bytes_total, const MstFileIn *h_files, uint32_t n_files,
                                          const MstFilter *h_filter, MstSongInfo *h_info, float *d_arena, uint64_t arena_floats,
                                          void *d_workspace, size_t workspace_bytes, void *stream_) {
    if (mst_device_count() == 0) { mst_set_error("no CUDA device: libmst_b200 has no CPU fallback"); return MST_E_NO_DEVICE; }
    if (!h_files || !h_info || !d_workspace || (!h_bytes && n_bytes_total)) {
        mst_set_error("mst_encode_batch_smf_async: null pointer argument");
        return MST_E_INVALID;
    }
    if (n_files == 0) return MST_OK;
    if (n_bytes_total >= 0xFFFFFFFFull) {
        mst_set_error("mst_encode_batch_smf_async: at most 2^32-1 bytes of files per batch");
        return MST_E_INVALID;
    }
    const size_t need = mst_encode_smf_workspace_bytes(n_bytes_total, n_files, arena_floats);
    if (need > workspace_bytes) {
        mst_set_error("mst_encode_batch_smf_async: workspace too small (%zu < %zu)", workspace_bytes, need);
        return MST_E_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const uint64_t E = smf_max_events(n_bytes_total);
    const size_t inner = mst_align_up(mst_encode_workspace_bytes(E, n_files, arena_floats), 256);
    SmfWorkspace ws;
    const size_t smf_bytes = mst_align_up(smf_carve(&ws, (char *)d_workspace + inner, n_bytes_total, n_files), 256);
    MstSongInfo *d_info = (MstSongInfo *)((char *)d_workspace + inner + smf_bytes);
    // The copies go to the caller's stream; the parser kernels run on a HIGH-PRIORITY stream of this caller stream's own
    // (forked and joined by events): they are a few hundred long-lived single-warp CTAs of latency-bound work, and at equal
    // priority the block scheduler would not start them before another slot's zero fill has dispatched its last CTA.
    int prc = smf_upload(ws, h_bytes, n_bytes_total, h_files, n_files, stream);
    if (prc != MST_OK) return prc;
    ParseStream *ps = parse_stream_for(stream);
    if (!ps) { mst_set_error("mst_encode_batch_smf_async: cannot create the parser stream"); return MST_E_CUDA; }
    MST_CUDA_CHECK(cudaEventRecord(ps->fork, stream));
    MST_CUDA_CHECK(cudaStreamWaitEvent(ps->s, ps->fork, 0));
    prc = smf_parse_enqueue(ws, n_files, ps->s);
    if (prc != MST_OK) return prc;
    MST_CUDA_CHECK(cudaEventRecord(ps->join, ps->s));
    MST_CUDA_CHECK(cudaStreamWaitEvent(stream, ps->join, 0));
    int rc = mst_encode_batch_device(ws.events, E, ws.songs, n_files, h_filter, d_info, d_arena, arena_floats, d_workspace, inner, stream_);
    if (rc != MST_OK) return rc;
    k_smf_status<<<(n_files + 255) / 256, 256, 0, stream>>>(ws.file_fail, d_info, n_files);
    MST_LAUNCH_CHECK("k_smf_status");
    MST_CUDA_CHECK(cudaMemcpyAsync(h_info, d_info, (size_t)n_files * sizeof(MstSongInfo), cudaMemcpyDeviceToHost, stream));
    return MST_OK;
}
