// mst_smf.cpp — Standard MIDI File bytes <-> MstEvent records (host side of libmst_b200.so).
//
// Replaces mido.MidiFile(path) as used by load_midi_from_file (style/midi.py:104-108) and, for the
// decode side, create_midi's track serialisation + mid.save (style/midi.py:125-168,
// style/style_transfer.py:98).  mido is an un-vendored, un-pinned dependency of the reference
// (requirements.txt:2), so this follows mido >= 1.3's published parsing rules:
//   - running status; meta events do not touch it; every other status byte (sysex included) sets it
//   - data bytes > 127, undefined status bytes, a missing MThd/MTrk tag, a truncated file and an
//     undecodable key signature all make the load fail (the reference then gets None)
//   - events keep their delta times per track; here they are accumulated to absolute ticks
// Only what the encoder consumes is emitted: channel voice messages, set_tempo, time_signature,
// key_signature, channel_prefix (it carries a `.channel`, midi_conversion.py:61) and one
// MST_EV_UNKNOWN per message whose type the reference does not know (:109,164).
#include <stdint.h>
#include <string.h>

#include <vector>

#include "../../include/mst_b200.h"

namespace {

enum LoadFail {
    kFailEof = 1,         // EOFError
    kFailNoHeader = 2,    // IOError: MThd not found
    kFailNoTrack = 3,     // IOError: no MTrk header at start of track
    kFailRunning = 4,     // IOError: running status without last_status
    kFailDataByte = 5,    // IOError: data byte must be in range 0..127
    kFailStatus = 6,      // IOError: undefined status byte
    kFailKeySig = 7,      // KeySignatureError
    kFailMeta = 8,        // truncated meta payload (mido raises IndexError; treated as a load failure)
    kFailSysex = 9,       // ValueError: sysex data byte out of range
};

struct Reader {
    const uint8_t *p;
    size_t n, pos;
    bool eof;
    int byte() {
        if (pos >= n) { eof = true; return 0; }
        return p[pos++];
    }
    // mido's read_variable_int takes any number of 7-bit groups; values beyond 2^56 saturate here (they can only
    // describe a payload longer than any file or a tick beyond MST_MAX_TICK, both handled by the callers)
    uint64_t varint() {
        uint64_t v = 0;
        while (true) {
            int b = byte();
            if (eof) return 0;
            v = v >= (1ull << 56) ? (1ull << 62) : ((v << 7) | (uint64_t)(b & 0x7F));
            if (b < 0x80) return v;
        }
    }
    bool skip(uint64_t k) {
        if (k > (uint64_t)(n - pos)) { eof = true; pos = n; return false; }   // no wrap-around for huge k
        pos += (size_t)k;
        return true;
    }
};

inline uint32_t be32(const uint8_t *b) { return ((uint32_t)b[0] << 24) | ((uint32_t)b[1] << 16) | ((uint32_t)b[2] << 8) | b[3]; }

}  // namespace

extern "C" int64_t mst_smf_parse(const uint8_t *bytes, size_t n_bytes, MstEvent *out, size_t capacity,
                                 int32_t *ticks_per_beat, int32_t *n_tracks_out, size_t *n_needed) {
    Reader r{bytes, n_bytes, 0, false};
    if (n_needed) *n_needed = 0;
    if (n_bytes < 8) return -(100 + kFailEof);
    if (memcmp(bytes, "MThd", 4) != 0) return -(100 + kFailNoHeader);
    const uint32_t hsize = be32(bytes + 4);
    r.pos = 8;
    const size_t avail = n_bytes - 8 < hsize ? n_bytes - 8 : hsize;
    if (avail < 6) return -(100 + kFailEof);
    const int16_t n_tracks = (int16_t)((bytes[10] << 8) | bytes[11]);
    const int16_t division = (int16_t)((bytes[12] << 8) | bytes[13]);
    r.pos = 8 + avail;
    if (ticks_per_beat) *ticks_per_beat = division;
    if (n_tracks_out) *n_tracks_out = n_tracks;

    size_t count = 0;
    auto emit = [&](uint64_t tick, uint8_t kind, uint8_t ch, uint8_t d1, uint8_t d2, uint32_t aux, int track) {
        if (tick > MST_MAX_TICK) return;   // split_channels drops these (midi_conversion.py:52,59)
        if (out && count < capacity) {
            MstEvent e;
            e.tick = (uint32_t)tick; e.aux = aux; e.kind = kind; e.channel = ch; e.d1 = d1; e.d2 = d2;
            e.track = (uint16_t)(track > 65535 ? 65535 : track); e.reserved = 0;
            out[count] = e;
        }
        count++;
    };

    for (int t = 0; t < n_tracks; t++) {
        if (r.pos + 8 > r.n) return -(100 + kFailEof);
        if (memcmp(r.p + r.pos, "MTrk", 4) != 0) return -(100 + kFailNoTrack);
        const uint32_t size = be32(r.p + r.pos + 4);
        r.pos += 8;
        const size_t start = r.pos;
        int last_status = -1;
        uint64_t tick = 0;
        while (r.pos - start != size) {
            const uint64_t delta = r.varint();
            if (r.eof) return -(100 + kFailEof);
            tick = (delta > (1ull << 62) - tick) ? (1ull << 62) : tick + delta;   // saturate: beyond MST_MAX_TICK anyway
            int status = r.byte();
            if (r.eof) return -(100 + kFailEof);
            int peek = -1;
            if (status < 0x80) {
                if (last_status < 0) return -(100 + kFailRunning);
                peek = status;
                status = last_status;
            } else if (status != 0xFF) {
                last_status = status;
            }
            if (status == 0xFF) {
                const int type = r.byte();
                const uint64_t len = r.varint();
                if (r.eof) return -(100 + kFailEof);
                const size_t dpos = r.pos;
                if (!r.skip(len)) return -(100 + kFailEof);
                const uint8_t *d = r.p + dpos;
                switch (type) {
                    case 0x51:
                        if (len < 3) return -(100 + kFailMeta);
                        emit(tick, MST_EV_SET_TEMPO, 0, 0, 0, ((uint32_t)d[0] << 16) | ((uint32_t)d[1] << 8) | d[2], t);
                        break;
                    case 0x58:
                        if (len < 4) return -(100 + kFailMeta);
                        emit(tick, MST_EV_TIME_SIGNATURE, 0, d[0], d[1], 0, t);
                        break;
                    case 0x59: {
                        if (len < 2) return -(100 + kFailMeta);
                        const int sf = (int8_t)d[0], mi = d[1];
                        if (sf < -7 || sf > 7 || mi > 1) return -(100 + kFailKeySig);
                        emit(tick, MST_EV_KEY_SIGNATURE, 0, 0, 0, (uint32_t)((sf + 7) * 2 + mi), t);
                        break;
                    }
                    case 0x20:
                        if (len < 1) return -(100 + kFailMeta);
                        emit(tick, MST_EV_CHANNEL_PREFIX, d[0], 0, 0, 0, t);
                        break;
                    case 0x09:   // device_name: a meta type outside known_messages
                        emit(tick, MST_EV_UNKNOWN, 0, 0, 0, 0x0900u, t);
                        break;
                    default:     // messages_to_ignore (midi_conversion.py:80-98), incl. unknown_meta
                        break;
                }
            } else if (status == 0xF0 || status == 0xF7) {
                const uint64_t len = r.varint();
                if (r.eof) return -(100 + kFailEof);
                const size_t dpos = r.pos;
                if (!r.skip(len)) return -(100 + kFailEof);
                size_t a = dpos, b = dpos + (size_t)len;
                if (b > a && r.p[a] == 0xF0) a++;
                if (b > a && r.p[b - 1] == 0xF7) b--;
                for (size_t i = a; i < b; i++) if (r.p[i] > 127) return -(100 + kFailSysex);
            } else if (status < 0xF0) {
                const int hi = status & 0xF0;
                const int need = (hi == 0xC0 || hi == 0xD0) ? 1 : 2;
                int data[2] = {0, 0};
                int have = 0;
                if (peek >= 0) data[have++] = peek;
                while (have < need) {
                    data[have++] = r.byte();
                    if (r.eof) return -(100 + kFailEof);
                }
                for (int i = 0; i < need; i++) if (data[i] > 127) return -(100 + kFailDataByte);
                const uint8_t kind = (uint8_t)((hi >> 4) - 8);
                // polytouch / aftertouch / pitchwheel only matter through their tick and channel
                const bool payload = kind == MST_EV_NOTE_OFF || kind == MST_EV_NOTE_ON ||
                                     kind == MST_EV_CONTROL_CHANGE || kind == MST_EV_PROGRAM_CHANGE;
                emit(tick, kind, (uint8_t)(status & 0x0F), payload ? (uint8_t)data[0] : 0,
                     payload ? (uint8_t)data[1] : 0, 0, t);
            } else {
                int ndata;
                switch (status) {
                    case 0xF1: case 0xF3: ndata = 1; break;
                    case 0xF2: ndata = 2; break;
                    case 0xF6: case 0xF8: case 0xFA: case 0xFB: case 0xFC: case 0xFE: ndata = 0; break;
                    default: return -(100 + kFailStatus);
                }
                if (peek >= 0) return -(100 + kFailStatus);
                for (int i = 0; i < ndata; i++) {
                    const int b = r.byte();
                    if (r.eof) return -(100 + kFailEof);
                    if (b > 127) return -(100 + kFailDataByte);
                }
                emit(tick, MST_EV_UNKNOWN, 0, 0, 0, (uint32_t)status, t);
            }
        }
    }
    if (n_needed) *n_needed = count;
    if (out && count > capacity) return MST_E_CAPACITY;
    return (int64_t)count;
}

// ------------------------------------------------------------------------------------------------
// SMF writer: create_midi's track layout (style/midi.py:125-168) serialised the way mido's
// MidiFile.save does (running status for channel messages, meta events reset it).
// ------------------------------------------------------------------------------------------------
namespace {
struct Writer {
    uint8_t *out;
    size_t cap, n;
    void put(uint8_t b) { if (out && n < cap) out[n] = b; n++; }
    void varint(uint32_t v) {
        uint8_t tmp[5]; int k = 0;
        tmp[k++] = v & 0x7F; v >>= 7;
        while (v) { tmp[k++] = (v & 0x7F) | 0x80; v >>= 7; }
        while (k) put(tmp[--k]);
    }
};
}  // namespace

extern "C" int64_t mst_smf_write(const MstDecodeSong *song, const MstOutMsg *msgs, uint32_t n_messages,
                                 int32_t eot_delta, uint8_t *out, size_t capacity, size_t *n_needed) {
    Writer w{out, capacity, 0};
    const char hdr[8] = {'M', 'T', 'h', 'd', 0, 0, 0, 6};
    for (int i = 0; i < 8; i++) w.put((uint8_t)hdr[i]);
    w.put(0); w.put(1);                       // type 1 (MidiFile() default)
    w.put(0); w.put(1);                       // one track
    w.put((uint8_t)((song->ticks_per_beat >> 8) & 0xFF)); w.put((uint8_t)(song->ticks_per_beat & 0xFF));
    const char trk[4] = {'M', 'T', 'r', 'k'};
    for (int i = 0; i < 4; i++) w.put((uint8_t)trk[i]);
    const size_t len_pos = w.n;
    for (int i = 0; i < 4; i++) w.put(0);
    const size_t start = w.n;
    // MetaMessage('time_signature', numerator, denominator): clocks_per_click 24, notated_32nd 8 (mido defaults)
    w.put(0); w.put(0xFF); w.put(0x58); w.put(4);
    w.put((uint8_t)song->numerator); w.put((uint8_t)song->denominator_log2); w.put(24); w.put(8);
    w.put(0); w.put(0xFF); w.put(0x51); w.put(3);
    w.put((uint8_t)((song->tempo >> 16) & 0xFF)); w.put((uint8_t)((song->tempo >> 8) & 0xFF)); w.put((uint8_t)(song->tempo & 0xFF));
    int running = -1;
    for (int c = 0; c < song->n_pitched; c++) {           // one program_change per non-9 channel (:139-145)
        const int ch = song->channel_id[c];
        if (ch == 9) continue;
        const int prog = song->instrument_id[c];
        if (prog < 0 || prog > 127 || ch > 15) return MST_E_INVALID;   // mido rejects the Message
        const int status = 0xC0 | ch;
        w.put(0);
        if (status != running) w.put((uint8_t)status);
        w.put((uint8_t)prog);
        running = status;
    }
    for (uint32_t i = 0; i < n_messages; i++) {
        const MstOutMsg &m = msgs[i];
        const int status = (m.is_on ? 0x90 : 0x80) | (m.channel & 0x0F);
        w.varint((uint32_t)m.delta);
        if (status != running) w.put((uint8_t)status);
        w.put(m.note); w.put(m.velocity);
        running = status;
    }
    w.varint((uint32_t)eot_delta);
    w.put(0xFF); w.put(0x2F); w.put(0);
    const size_t len = w.n - start;
    if (out && w.n <= capacity) {
        out[len_pos] = (uint8_t)(len >> 24); out[len_pos + 1] = (uint8_t)(len >> 16);
        out[len_pos + 2] = (uint8_t)(len >> 8); out[len_pos + 3] = (uint8_t)len;
    }
    if (n_needed) *n_needed = w.n;
    if (!out || w.n > capacity) return MST_E_CAPACITY;
    return (int64_t)w.n;
}

// ------------------------------------------------------------------------------------------------
// Batch loader: replaces the per-file `load_midi_from_file` (style/midi.py:104-108) + the one-thread prefetcher
// (style/utils/parallel.py:34-51) in front of the encoder.  A pool of C++ threads reads and parses the files straight
// into the caller's (page-locked) MstEvent / MstSongIn buffers, so that mst_encode_batch_host_async can start its H2D
// copy from them without another pass over the data.
// ------------------------------------------------------------------------------------------------
#include <stdio.h>

#include <atomic>
#include <thread>

extern "C" int mst_smf_load_batch(const char *const *paths, uint32_t n_files, int32_t n_threads, MstEvent *h_events,
                                  uint64_t capacity, MstSongIn *h_songs, int32_t *h_status, uint64_t *n_events_total) {
    if (!paths || !h_songs || !h_status || !n_events_total || (!h_events && capacity)) return MST_E_INVALID;
    *n_events_total = 0;
    if (n_files == 0) return MST_OK;
    unsigned nt = n_threads > 0 ? (unsigned)n_threads : std::thread::hardware_concurrency();
    if (nt == 0) nt = 1;
    if (nt > n_files) nt = n_files;
    std::vector<std::vector<MstEvent>> parsed(n_files);
    std::vector<int32_t> tpb(n_files, 0);
    std::atomic<uint32_t> next(0);
    auto parse_worker = [&]() {
        std::vector<uint8_t> bytes;
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= n_files) return;
            h_status[i] = -1;                                   // unreadable file (OSError in the reference)
            FILE *f = paths[i] ? fopen(paths[i], "rb") : nullptr;
            if (!f) continue;
            fseek(f, 0, SEEK_END);
            const long size = ftell(f);
            fseek(f, 0, SEEK_SET);
            if (size < 0) { fclose(f); continue; }
            bytes.resize((size_t)size);
            const size_t got = size ? fread(bytes.data(), 1, (size_t)size, f) : 0;
            fclose(f);
            if (got != (size_t)size) continue;
            std::vector<MstEvent> &ev = parsed[i];
            ev.resize((size_t)size / 2 + 1);                    // an event takes at least two bytes (delta + data byte)
            int32_t division = 0, n_tracks = 0;
            size_t need = 0;
            const int64_t n = mst_smf_parse(bytes.data(), bytes.size(), ev.data(), ev.size(), &division, &n_tracks, &need);
            if (n < 0) { ev.clear(); h_status[i] = (int32_t)n; continue; }   // -(100 + reason): mido would raise, the reference gets None
            ev.resize((size_t)n);
            tpb[i] = division;
            h_status[i] = 0;
        }
    };
    {
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nt; t++) pool.emplace_back(parse_worker);
        parse_worker();
        for (auto &t : pool) t.join();
    }
    uint64_t total = 0;
    for (uint32_t i = 0; i < n_files; i++) {
        h_songs[i].event_offset = total;
        h_songs[i].n_events = (uint32_t)parsed[i].size();
        h_songs[i].ticks_per_beat = tpb[i];
        total += parsed[i].size();
    }
    *n_events_total = total;
    if (total > capacity) return MST_E_CAPACITY;
    next = 0;
    auto copy_worker = [&]() {
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= n_files) return;
            if (!parsed[i].empty()) memcpy(h_events + h_songs[i].event_offset, parsed[i].data(), parsed[i].size() * sizeof(MstEvent));
        }
    };
    {
        std::vector<std::thread> pool;
        for (unsigned t = 1; t < nt; t++) pool.emplace_back(copy_worker);
        copy_worker();
        for (auto &t : pool) t.join();
    }
    return MST_OK;
}

// ------------------------------------------------------------------------------------------------
// General SMF serialiser of an MstEvent array (track-major, absolute ticks): the inverse of mst_smf_parse for the event
// kinds the encoder consumes.  Used to put synthetic corpora on disk as real .mid files (bench.py's file-fed legs, tests).
// ------------------------------------------------------------------------------------------------
extern "C" int64_t mst_smf_write_events(const MstEvent *events, uint64_t n_events, int32_t ticks_per_beat, uint8_t *out,
                                        size_t capacity, size_t *n_needed) {
    if (!events && n_events) return MST_E_INVALID;
    Writer w{out, capacity, 0};
    uint32_t n_tracks = 0;
    for (uint64_t i = 0; i < n_events; i++) if (events[i].track + 1u > n_tracks) n_tracks = events[i].track + 1u;
    if (n_tracks == 0) n_tracks = 1;
    const char hdr[8] = {'M', 'T', 'h', 'd', 0, 0, 0, 6};
    for (int i = 0; i < 8; i++) w.put((uint8_t)hdr[i]);
    w.put(0); w.put(1);
    w.put((uint8_t)(n_tracks >> 8)); w.put((uint8_t)n_tracks);
    w.put((uint8_t)((ticks_per_beat >> 8) & 0xFF)); w.put((uint8_t)(ticks_per_beat & 0xFF));
    uint64_t i = 0;
    for (uint32_t t = 0; t < n_tracks; t++) {
        const char trk[4] = {'M', 'T', 'r', 'k'};
        for (int k = 0; k < 4; k++) w.put((uint8_t)trk[k]);
        const size_t len_pos = w.n;
        for (int k = 0; k < 4; k++) w.put(0);
        const size_t start = w.n;
        uint32_t prev = 0;
        int running = -1;
        for (; i < n_events && events[i].track == t; i++) {
            const MstEvent &e = events[i];
            w.varint(e.tick >= prev ? e.tick - prev : 0);
            prev = e.tick >= prev ? e.tick : prev;
            if (e.kind <= MST_EV_PITCHWHEEL) {
                const int status = ((e.kind + 8) << 4) | (e.channel & 0x0F);
                if (status != running) w.put((uint8_t)status);
                running = status;
                w.put(e.d1 & 0x7F);
                if (e.kind != MST_EV_PROGRAM_CHANGE && e.kind != MST_EV_AFTERTOUCH) w.put(e.d2 & 0x7F);
            } else {
                running = -1;                                    // a meta event cancels running status in a file
                w.put(0xFF);
                switch (e.kind) {
                    case MST_EV_SET_TEMPO: w.put(0x51); w.put(3); w.put((uint8_t)(e.aux >> 16)); w.put((uint8_t)(e.aux >> 8)); w.put((uint8_t)e.aux); break;
                    case MST_EV_TIME_SIGNATURE: w.put(0x58); w.put(4); w.put(e.d1); w.put(e.d2); w.put(24); w.put(8); break;
                    case MST_EV_KEY_SIGNATURE: w.put(0x59); w.put(2); w.put((uint8_t)(int8_t)((int)(e.aux / 2) - 7)); w.put((uint8_t)(e.aux & 1)); break;
                    case MST_EV_CHANNEL_PREFIX: w.put(0x20); w.put(1); w.put(e.channel); break;
                    case MST_EV_UNKNOWN: w.put(0x09); w.put(0); break;   // device_name: a type outside known_messages
                    default: w.put(0x01); w.put(0); break;               // an ignored meta (empty text)
                }
            }
        }
        w.put(0); w.put(0xFF); w.put(0x2F); w.put(0);
        const size_t len = w.n - start;
        if (out && w.n <= capacity) {
            out[len_pos] = (uint8_t)(len >> 24); out[len_pos + 1] = (uint8_t)(len >> 16);
            out[len_pos + 2] = (uint8_t)(len >> 8); out[len_pos + 3] = (uint8_t)len;
        }
    }
    if (n_needed) *n_needed = w.n;
    if (!out || w.n > capacity) return MST_E_CAPACITY;
    return (int64_t)w.n;
}

// ------------------------------------------------------------------------------------------------
// mst_files_read_batch: the files of a batch, read as they are into one (page-locked) buffer for the device-side parser
// (mst_encode_batch_smf_async).  Sizes first (stat), then the reads on a thread pool; an unreadable file gets n_bytes = 0
// and status -1 (OSError in the reference's load_midi_from_file -> the file is skipped; here the device parser reports
// MST_SONG_LOAD_FAILED for a 0-byte file).  Files start on 16-byte boundaries.
// ------------------------------------------------------------------------------------------------
extern "C" int mst_files_read_batch(const char *const *paths, uint32_t n_files, int32_t n_threads, uint8_t *h_bytes,
                                    uint64_t capacity, MstFileIn *h_files, int32_t *h_status, uint64_t *n_bytes_total) {
    if (!paths || !h_files || !h_status || !n_bytes_total || (!h_bytes && capacity)) return MST_E_INVALID;
    *n_bytes_total = 0;
    if (n_files == 0) return MST_OK;
    uint64_t total = 0;
    for (uint32_t i = 0; i < n_files; i++) {
        h_status[i] = -1;
        h_files[i].byte_offset = total; h_files[i].n_bytes = 0; h_files[i].reserved = 0;
        FILE *f = paths[i] ? fopen(paths[i], "rb") : nullptr;
        if (!f) continue;
        long size = -1;
        if (fseek(f, 0, SEEK_END) == 0) size = ftell(f);
        fclose(f);
        if (size < 0 || (uint64_t)size >= 0xFFFFFFFFull) continue;
        h_files[i].n_bytes = (uint32_t)size;
        h_status[i] = 0;
        total += ((uint64_t)size + 15) & ~15ull;
    }
    *n_bytes_total = total;
    if (total > capacity) return MST_E_CAPACITY;
    unsigned nt = n_threads > 0 ? (unsigned)n_threads : std::thread::hardware_concurrency();
    if (nt == 0) nt = 1;
    if (nt > n_files) nt = n_files;
    std::atomic<uint32_t> next(0);
    auto worker = [&]() {
        for (;;) {
            const uint32_t i = next.fetch_add(1);
            if (i >= n_files) return;
            if (h_status[i] != 0 || h_files[i].n_bytes == 0) continue;
            FILE *f = fopen(paths[i], "rb");
            size_t got = 0;
            if (f) { got = fread(h_bytes + h_files[i].byte_offset, 1, h_files[i].n_bytes, f); fclose(f); }
            if (got != h_files[i].n_bytes) { h_files[i].n_bytes = 0; h_status[i] = -1; }   // changed under us: treated as unreadable
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < nt; t++) pool.emplace_back(worker);
    worker();
    for (auto &t : pool) t.join();
    return MST_OK;
}
