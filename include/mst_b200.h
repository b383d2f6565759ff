/*
 * mst_b200.h — C ABI of the B200-native hot path of marcinp7/music-style-transfer.
 *
 * Plain C: fixed-width scalars, raw pointers and sizes only.  No torch / C++ types cross
 * this boundary.  The reference is pure Python and has no FFI of its own,
 * so each entry point names the reference function(s) it replaces (file:line relative to the
 * reference tree) and INTEGRATION.md shows the ctypes binding a maintainer would add.
 *
 * Conventions
 *   - every function returns 0 on success, a negative MST_E_* code on an API error
 *     (mst_last_error() gives the text); per-song problems never fail the call: they are
 *     reported in MstSongInfo.status (mirrors the reference raising MidiFormatError per file,
 *     style/data.py:47);
 *   - "d_" pointers are device pointers, "h_" pointers are host pointers;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); device entry points
 *     are asynchronous on that stream, host ("_host") entry points return after completion;
 *   - the caller owns every buffer; scratch sizes come from the *_workspace_bytes calls.
 */
#ifndef MST_B200_H
#define MST_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MST_ABI_VERSION 2

/* ---------------------------------------------------------------- errors */
#define MST_OK                0
#define MST_E_INVALID        -1   /* bad argument */
#define MST_E_CUDA           -2   /* CUDA runtime error (text in mst_last_error) */
#define MST_E_CAPACITY       -3   /* a caller-provided buffer is too small */
#define MST_E_NO_DEVICE      -4   /* no CUDA device: the product path has no CPU fallback */

/* per-song status (MstSongInfo.status) */
#define MST_SONG_OK                   0
#define MST_SONG_TIME_SIG_CHANGED     1  /* "Time signature changed"  midi_conversion.py:153 */
#define MST_SONG_KEY_SIG_CHANGED      2  /* "Key signature changed"   midi_conversion.py:157 */
#define MST_SONG_UNKNOWN_MESSAGE      3  /* "Unknown message type"    midi_conversion.py:165,191 */
#define MST_SONG_NO_NOTES             4  /* no note_on with velocity>0: min() of empty list, :125 */
#define MST_SONG_NO_PITCHED           5  /* filtered song has no pitched channel, data.py:57 */
#define MST_SONG_BAD_TIMING           6  /* ticks_per_beat<=0 or numerator==0 (reference crashes) */
#define MST_SONG_TOO_MANY_INSTRUMENTS 7  /* >MST_MAX_LISTS_PER_CHANNEL programs with notes on one channel */
#define MST_SONG_ARENA_FULL           8  /* grids did not fit the caller's arena; info is valid */
#define MST_SONG_LOAD_FAILED          9  /* SMF parse error: load_midi_from_file returns None, midi.py:104-108 */

/* ---------------------------------------------------------------- events */
/* One MIDI event, 16 bytes.  Replaces mido message objects as consumed by merge_tracks /
 * split_channels (style/midi_conversion.py:37-66).  A song's events are laid out track after
 * track, each track in file order, `tick` already accumulated per track (absolute ticks);
 * events beyond MST_MAX_TICK are dropped by the encoder exactly like :52,59. */
typedef struct MstEvent {
    uint32_t tick;     /* absolute tick inside its track */
    uint32_t aux;      /* set_tempo: microseconds per beat; key_signature: (sf+7)*2+minor */
    uint8_t  kind;     /* MST_EV_* */
    uint8_t  channel;  /* channel voice messages: 0..15; channel_prefix: its data byte */
    uint8_t  d1;       /* note | control | program | time_signature numerator */
    uint8_t  d2;       /* velocity | value | time_signature log2(denominator) */
    uint16_t track;    /* informative */
    uint16_t reserved;
} MstEvent;

#define MST_MAX_TICK 10000000u

#define MST_EV_NOTE_OFF        0   /* 0x8n */
#define MST_EV_NOTE_ON         1   /* 0x9n */
#define MST_EV_POLYTOUCH       2   /* 0xAn */
#define MST_EV_CONTROL_CHANGE  3   /* 0xBn */
#define MST_EV_PROGRAM_CHANGE  4   /* 0xCn */
#define MST_EV_AFTERTOUCH      5   /* 0xDn */
#define MST_EV_PITCHWHEEL      6   /* 0xEn */
#define MST_EV_SET_TEMPO      16
#define MST_EV_TIME_SIGNATURE 17
#define MST_EV_KEY_SIGNATURE  18
#define MST_EV_CHANNEL_PREFIX 19   /* meta 0x20: carries a .channel attribute in mido, so the
                                      reference routes it into the channel lists (:61) */
#define MST_EV_META_IGNORED   20   /* any type in messages_to_ignore (:80-98) without channel */
#define MST_EV_UNKNOWN        21   /* type outside known_messages (:109): rejects the song */

/* One song inside a batch. */
typedef struct MstSongIn {
    uint64_t event_offset;   /* index of the song's first event in the batch event array */
    uint32_t n_events;
    int32_t  ticks_per_beat; /* SMF header division (mid.ticks_per_beat) */
} MstSongIn;

/* Channel filter of iter_inputs (style/data.py:53-58) / get_model_input
 * (style/style_transfer.py:62). */
typedef struct MstFilter {
    int32_t apply;              /* 0: keep every (channel, instrument) list */
    int32_t min_n_messages;     /* keep lists with at least this many note messages */
    uint8_t included[128];      /* included[p] != 0: program p passes; percussion (-1) always */
    int32_t crop_channel_bars;  /* >0: write only the first crop_channel_bars // C bars
                                   (prepare_input, data.py:133-137 with train-model.py:101) */
    int32_t reserved;
} MstFilter;

#define MST_MAX_PITCHED   120   /* programs 0..119 are pitched (midi.py:96-101) */
#define MST_MAX_UNPITCHED 12    /* -1 (channel 9) and sound effects 120..127 */
#define MST_MAX_LISTS_PER_CHANNEL 16

#define MST_N_FRACTIONS       10  /* sorted {i/8} U {i/3}, midi_conversion.py:358-362 */
#define MST_N_PITCHED_ROWS    56  /* 8 octaves x 7 degrees, :366 */
#define MST_N_PITCHED_FEATS   5   /* duration, velocity, flat, natural, sharp (:502-510) */
#define MST_N_UNPITCHED_ROWS  47  /* percussion keys 35..81, :367 */
#define MST_N_UNPITCHED_FEATS 2
#define MST_PITCHED_BEAT_FLOATS   (MST_N_FRACTIONS * MST_N_PITCHED_ROWS * MST_N_PITCHED_FEATS)     /* 2800 */
#define MST_UNPITCHED_BEAT_FLOATS (MST_N_FRACTIONS * MST_N_UNPITCHED_ROWS * MST_N_UNPITCHED_FEATS) /* 940 */

/* Song-level result: the reference's `info` dict (get_midi_info, :117-179) + info['scale']
 * (style/scales.py:160-221) + the shapes/instrument lists get_input returns (data.py:66-100). */
typedef struct MstSongInfo {
    int32_t status;            /* MST_SONG_* */
    int32_t ticks_per_beat;
    int32_t numerator;         /* info['time_signature'] */
    int32_t denominator_log2;
    int32_t key_signature;     /* -1 = None, else (sf+7)*2+minor */
    int32_t duration;          /* max tick over channel messages */
    int32_t ticks_per_bar;
    int32_t n_bars;            /* ceil(duration / ticks_per_bar); grids hold n_bars+1 bars */
    int32_t n_beats;
    int32_t tempo;
    int32_t bpm;
    int32_t tonic;             /* 0..11 = C..B */
    int32_t mode;              /* 0 major (Ionian), 1 minor (Aeolian) */
    int32_t n_pitched;         /* C */
    int32_t n_unpitched;       /* U */
    int32_t n_bars_out;        /* bars actually written per channel (n_bars+1, or the crop) */
    double  scale_coverage;
    double  scale_loose_coverage;
    double  scale_cross_entropy;
    double  scale_loss;
    double  scale_margin;      /* runner-up loss - best loss (near-tie audit) */
    double  keys_dist[12];
    uint64_t pitched_offset;   /* float offset of (C, n_bars_out, n_beats, 10, 56, 5) in the arena */
    uint64_t unpitched_offset; /* float offset of (U, n_bars_out, n_beats, 10, 47, 2) */
    int32_t n_notes;           /* notes scattered (after dropping out-of-range rows) */
    int32_t n_lists;           /* (channel, instrument) lists before filtering (read_midi channels) */
    int16_t pitched_instruments[MST_MAX_PITCHED];
    int16_t unpitched_instruments[MST_MAX_UNPITCHED];
} MstSongInfo;

/* ---------------------------------------------------------------- library */
int         mst_abi_version(void);
const char *mst_last_error(void);
int         mst_device_count(void);   /* 0 when no GPU is visible */

/* Optional device-side timing (CUDA events) of the encode pipeline: which = 0 the whole pipeline
 * (sort .. note scatter), 1 = the grid-writing kernel (the zero fill), which overlaps part of it.
 * Read after a stream synchronise.  Used by bench.py for the roofline of the dominant kernel. */
int   mst_profile_enable(int on);
float mst_profile_encode_ms(int which);

/* ---------------------------------------------------------------- SMF I/O (host, C++) */
/* Replaces mido.MidiFile(path) as used by load_midi_from_file (style/midi.py:104-108):
 * parses Standard MIDI File bytes into MstEvent records (absolute ticks, track-major order).
 * Returns the number of events (>=0) or MST_E_CAPACITY (needed count in *n_needed) or
 * -(100+reason) for the load failures on which the reference returns None. */
int64_t mst_smf_parse(const uint8_t *bytes, size_t n_bytes, MstEvent *h_events, size_t capacity,
                      int32_t *ticks_per_beat, int32_t *n_tracks, size_t *n_needed);

/* Batch loader: `load_midi_from_file` (style/midi.py:104-108) for n_files paths on a pool of n_threads C++ threads
 * (<= 0: one per hardware thread), replacing the reference's single prefetch thread (style/utils/parallel.py:34-51).
 * Events land in h_events (page-locked memory makes the following mst_encode_batch_host_async copy asynchronous),
 * h_songs[i] = {offset, count, ticks_per_beat} of file i, h_status[i] = 0, -1 (unreadable: the reference's OSError ->
 * None) or the negative mst_smf_parse failure code (mido raises -> None); failed files get zero events.
 * Returns MST_OK, or MST_E_CAPACITY with the needed record count in *n_events_total. */
int mst_smf_load_batch(const char *const *paths, uint32_t n_files, int32_t n_threads, MstEvent *h_events,
                       uint64_t capacity, MstSongIn *h_songs, int32_t *h_status, uint64_t *n_events_total);

/* Serialise an MstEvent array (track-major, absolute ticks) as a type-1 Standard MIDI File: the inverse of mst_smf_parse
 * for the event kinds the encoder consumes (synthetic corpora on disk for bench.py's file-fed legs and for tests).
 * Returns bytes written or MST_E_CAPACITY (needed size in *n_needed). */
int64_t mst_smf_write_events(const MstEvent *events, uint64_t n_events, int32_t ticks_per_beat, uint8_t *out,
                             size_t capacity, size_t *n_needed);

/* ---------------------------------------------------------------- encode (MIDI -> grids) */
/* Replaces read_midi (midi_conversion.py:216-232), iter_inputs' filter (data.py:53-58),
 * get_input (data.py:66-100) incl. ChannelConverter.channel2nchannel / nchannel2vchannel
 * (:371-456,490-516), get_keys_dist (:340) + get_scale (scales.py:214), and prepare_input's
 * crop + f32 cast (data.py:130-143) for a whole batch of songs.
 * d_arena must be 16-byte aligned (MST_E_INVALID otherwise); the grids of the songs that fit are
 * packed from its start, song after song, and that range is the only part of it that is written. */
size_t mst_encode_workspace_bytes(uint64_t n_events_total, uint32_t n_songs, uint64_t arena_floats);

int mst_encode_batch_device(const MstEvent *d_events, uint64_t n_events_total,
                            const MstSongIn *d_songs, uint32_t n_songs,
                            const MstFilter *h_filter,
                            MstSongInfo *d_info,
                            float *d_arena, uint64_t arena_floats,
                            void *d_workspace, size_t workspace_bytes,
                            void *stream);

/* Same through host buffers: H2D of events/songs, kernels, D2H of the song infos (and of the
 * grids when h_arena != NULL; otherwise they stay in d_arena for the device consumer).
 * The workspace additionally stages the events/songs/infos: size it with the _host_ query. */
size_t mst_encode_host_workspace_bytes(uint64_t n_events_total, uint32_t n_songs, uint64_t arena_floats);

int mst_encode_batch_host(const MstEvent *h_events, uint64_t n_events_total,
                          const MstSongIn *h_songs, uint32_t n_songs,
                          const MstFilter *h_filter,
                          MstSongInfo *h_info,
                          float *d_arena, uint64_t arena_floats, float *h_arena,
                          void *d_workspace, size_t workspace_bytes,
                          void *stream);

/* The same without the final synchronisation (grids stay in d_arena): copies and kernels are only ENQUEUED on `stream`,
 * so a caller that alternates two streams / workspaces / arenas overlaps the H2D of batch i + 1 with the kernels of
 * batch i (what the prefetch thread of the reference, utils/parallel.py:34-39, does for its Python encoder).
 * h_events / h_songs / h_info must be page-locked and stay valid until the stream has drained; h_info is only
 * readable after that. */
int mst_encode_batch_host_async(const MstEvent *h_events, uint64_t n_events_total,
                                const MstSongIn *h_songs, uint32_t n_songs,
                                const MstFilter *h_filter,
                                MstSongInfo *h_info,
                                float *d_arena, uint64_t arena_floats,
                                void *d_workspace, size_t workspace_bytes,
                                void *stream);


/* 8-byte wire format of an event for the host path (halves the H2D volume, which is what bounds the host-buffer path when
 * several ranks share a host socket): tick:32 | kind:4 (channel kinds as they are, meta kinds 16..21 as 8..13) | channel:4 |
 * d1:8 | d2:8 | track:8; set_tempo carries its 24-bit aux in the (d1, d2, track) bytes, key_signature its code in d1.
 * mst_events_pack converts MstEvent records on the host; mst_encode_batch_packed_async is mst_encode_batch_host_async on
 * packed records (widened on the device); the workspace is sized by mst_encode_packed_workspace_bytes. */
typedef uint64_t MstPackedEvent;
void   mst_events_pack(const MstEvent *events, uint64_t n, MstPackedEvent *out);
size_t mst_encode_packed_workspace_bytes(uint64_t n_events_total, uint32_t n_songs, uint64_t arena_floats);
int    mst_encode_batch_packed_async(const MstPackedEvent *h_events, uint64_t n_events_total, const MstSongIn *h_songs,
                                     uint32_t n_songs, const MstFilter *h_filter, MstSongInfo *h_info, float *d_arena,
                                     uint64_t arena_floats, void *d_workspace, size_t workspace_bytes, void *stream);

/* Raw Standard MIDI File bytes as the wire format of the host path: the files of a batch lie concatenated in one
 * (page-locked) host buffer, are copied as they are and PARSED ON THE DEVICE (one thread per track chunk), then encoded.
 * Replaces mido.MidiFile(path) in load_midi_from_file (style/midi.py:104-108) + everything mst_encode_batch_host_async
 * replaces, for a batch of files: same events and same "load failed" verdicts as mst_smf_parse, reported as
 * MST_SONG_LOAD_FAILED in the file's MstSongInfo.  A file spends 3 - 4 bytes on an event where MstEvent spends 16 and
 * MstPackedEvent 8: this is the form that keeps the host -> device copy off the critical path when several GPUs share
 * one host.  The workspace is sized from the BYTE count (an event takes at least 2 bytes, a track chunk at least 8). */
typedef struct MstFileIn {
    uint64_t byte_offset;   /* of the file's first byte in the batch buffer */
    uint32_t n_bytes;
    uint32_t reserved;
} MstFileIn;
/* Reads the files as they are into one buffer (C++ thread pool, no Python in the loop; stat first, then the reads): file i
 * starts at h_files[i].byte_offset (16-byte aligned); an unreadable file gets n_bytes = 0 and h_status[i] = -1 (the
 * device parser then reports MST_SONG_LOAD_FAILED for it).  MST_E_CAPACITY with the needed size in *n_bytes_total when
 * the buffer is too small. */
int    mst_files_read_batch(const char *const *paths, uint32_t n_files, int32_t n_threads, uint8_t *h_bytes, uint64_t capacity,
                            MstFileIn *h_files, int32_t *h_status, uint64_t *n_bytes_total);
/* The device-side parser alone (synchronous; for callers that want the events): same results as mst_smf_parse per file -
 * h_songs[i] = (event offset, count, division), h_status[i] != 0 where the load fails, *n_events_total records in h_events
 * (MST_E_CAPACITY with the needed count when `capacity` is too small). */
size_t mst_smf_parse_device_workspace_bytes(uint64_t n_bytes_total, uint32_t n_files);
int    mst_smf_parse_batch_device(const uint8_t *h_bytes, uint64_t n_bytes_total, const MstFileIn *h_files, uint32_t n_files,
                                  MstEvent *h_events, uint64_t capacity, MstSongIn *h_songs, int32_t *h_status,
                                  uint64_t *n_events_total, void *d_workspace, size_t workspace_bytes, void *stream);
size_t mst_encode_smf_workspace_bytes(uint64_t n_bytes_total, uint32_t n_files, uint64_t arena_floats);
int    mst_encode_batch_smf_async(const uint8_t *h_bytes, uint64_t n_bytes_total, const MstFileIn *h_files, uint32_t n_files,
                                  const MstFilter *h_filter, MstSongInfo *h_info, float *d_arena, uint64_t arena_floats,
                                  void *d_workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------- decode (grids -> MIDI events) */
/* Replaces hard_output (style/model.py:818-832), ChannelConverter.vchannel2channel
 * (midi_conversion.py:518-573 = vchannel2qchannel + qchannel2channel), decode_midi
 * (style/style_transfer.py:145-158) and create_midi's ordering / delta / 1-second clip
 * (style/midi.py:120-168) for a batch of songs. */
#define MST_MAX_DECODE_CHANNELS 32

#define MST_DECODE_OK             0
#define MST_DECODE_NO_MESSAGES    1   /* nothing above the velocity gate: messages[-1] raises (midi.py:158) */
#define MST_DECODE_BAD_VALUE      2   /* non-finite duration / time out of range (int() raises) */
#define MST_DECODE_VELOCITY       3   /* int(velocity*127) > 127: the assert at midi.py:148 */
#define MST_DECODE_CAPACITY       4   /* message buffer too small; n_messages is the needed count */

typedef struct MstDecodeSong {
    uint64_t pitched_offset;     /* float offset of (C, n_bars, n_beats, 10, 56, 5) in the grid buffer */
    uint64_t unpitched_offset;   /* float offset of (1, n_bars, n_beats, 10, 47, 2) */
    int32_t  n_pitched;          /* channels to decode (len(channels_info) zipped with the grid) */
    int32_t  has_unpitched;
    int32_t  n_bars;
    int32_t  n_beats;
    int32_t  ticks_per_beat;
    int32_t  ticks_per_bar;
    int32_t  tempo;
    int32_t  tonic;              /* info['scale']['key'] as 0..11 */
    int32_t  mode;               /* 0 major, 1 minor */
    int32_t  duration;           /* info.get('duration'); < 0 = absent (last time + ticks_per_bar) */
    int32_t  numerator;
    int32_t  denominator_log2;
    uint8_t  channel_id[MST_MAX_DECODE_CHANNELS];      /* MIDI channel of each pitched grid channel */
    int16_t  instrument_id[MST_MAX_DECODE_CHANNELS];   /* program written by create_midi */
    int32_t  unpitched_channel_id;                     /* 9 */
    int32_t  write_back;         /* != 0: zero gated velocities in the input like hard_output's in-place `*=` */
} MstDecodeSong;

typedef struct MstOutMsg {       /* one note message of the output track, in final order */
    int32_t time;                /* absolute tick */
    int32_t delta;               /* max(0, min(time - previous, second2tick(1 s)))  (midi.py:161-167) */
    uint8_t is_on;               /* 1: 0x9n, 0: 0x8n */
    uint8_t channel;
    uint8_t note;
    uint8_t velocity;            /* int(velocity * 127), 0 for note_off */
} MstOutMsg;

typedef struct MstDecodeResult {
    int32_t  status;             /* MST_DECODE_* */
    uint32_t n_messages;
    uint64_t msg_offset;         /* index of the song's first message in the output array */
    int32_t  eot_delta;          /* delta time of end_of_track */
    int32_t  max_delta;          /* second2tick(1, ticks_per_beat, tempo) */
} MstDecodeResult;

/* n_chunks >= sum over songs of (n_pitched + has_unpitched) * n_bars: an upper bound is enough (it sizes the workspace;
 * mst_transfer_select decides n_pitched / has_unpitched on the device).  A song table that needs MORE chunks than
 * n_chunks, or an entry with n_pitched > MST_MAX_DECODE_CHANNELS / non-positive bars, beats, ticks or tempo, is
 * reported per song (MST_DECODE_CAPACITY / MST_DECODE_BAD_VALUE), never indexed.
 * d_grids is written only when a song asks for write_back. */
size_t mst_decode_workspace_bytes(uint32_t n_songs, uint64_t n_chunks, uint64_t msg_capacity);

int mst_decode_batch_device(float *d_grids,
                            const MstDecodeSong *d_songs, uint32_t n_songs, uint64_t n_chunks,
                            MstOutMsg *d_msgs, uint64_t msg_capacity,
                            MstDecodeResult *d_results,
                            void *d_workspace, size_t workspace_bytes, void *stream);

/* SMF serialisation of one decoded song (create_midi's track + mid.save, style/midi.py:125-168 and
 * style_transfer.py:98): type-1 file, one track, time_signature + set_tempo + one program_change per
 * non-9 channel + the messages with running status + end_of_track.  Returns bytes written or
 * MST_E_CAPACITY (needed size in *n_needed). */
int64_t mst_smf_write(const MstDecodeSong *song, const MstOutMsg *h_msgs, uint32_t n_messages,
                      int32_t eot_delta, uint8_t *out, size_t capacity, size_t *n_needed);

/* ---------------------------------------------------------------- style transfer glue (style/style_transfer.py) */
/* apply_style's host decisions (style/style_transfer.py:101-122), made on the device for a batch of S styles so that a
 * "composition x S styles" transfer needs no host round trip between the song-info heads and the finished messages:
 *   instruments = argsort(-logits)[:n] (+1 when the single pick is percussion, :106-108), percussion removed (:109-110),
 *   mapped through the one-hot encoder's sorted categories (:112-116); mode = argmax of the two logits (:118-121);
 *   tempo = bpm2tempo(round(bpm)) with Python's half-to-even round (:103).
 * Outputs: d_features (S, max_channels, 51) = encode_instruments of the chosen programs (zero rows beyond n_pitched),
 * d_choice[s], and d_songs[s] (in/out): n_pitched, has_unpitched, mode, tempo, channel_id (0..15 without 9, :78-80) and
 * instrument_id are filled in, every other field (offsets, bars, ticks, tonic, ...) is the caller's. */
typedef struct MstTransferChoice {
    int32_t n_pitched;         /* instruments chosen, percussion removed */
    int32_t has_unpitched;     /* percussion was among the picks */
    int32_t mode;              /* 0 major, 1 minor */
    int32_t tempo;             /* microseconds per beat */
    int16_t instruments[MST_MAX_DECODE_CHANNELS];   /* programs, descending logit; -1 beyond n_pitched */
} MstTransferChoice;

int mst_transfer_select(const float *d_instruments_logits /* (S,41) */, const float *d_mode_logits /* (S,2) */,
                        const float *d_bpm /* (S) */, const int32_t *d_n_instruments /* (S) */, uint32_t n_styles,
                        const int16_t *h_programs /* [40] sorted included programs */,
                        const int16_t *h_groups /* [40] group one-hot index of each */, int32_t max_channels,
                        float *d_features, MstDecodeSong *d_songs, MstTransferChoice *d_choice, void *stream);

/* ---------------------------------------------------------------- model (style/model.py) */
/* Replaces StyleTransferModel.forward / extract_style / predict_song_info / apply_style
 * (style/model.py:751-793) with its nine sub-modules, get_total_loss (:935-997), loss.backward() and
 * torch.optim.Adam.step (train-model.py:113-154) for a batch of N independent batch-1 samples of one shape.
 * Parameters live in ONE flat fp32 buffer; mst_model_param_* describe its layout by state-dict name. */
/* arithmetic of the dense contractions (conv / linear / LSTM input projections of style/model.py:36-724) */
#define MST_PRECISION_FP32 0   /* fp32-accurate: tcgen05 kind::tf32 with the 3-term split (1e-5 / 1e-4 parity with torch fp32) */
#define MST_PRECISION_BF16 1   /* BASELINE.json configs[3] "bf16 dense layers": bf16 operands, fp32 accumulation in TMEM
                                  (tcgen05 kind::f16), for the row-space GEMMs and the per-note PitchedStyleApplier tail;
                                  tolerance 1e-3 on losses / 2e-2 on gradients */

typedef struct MstModelShape {
    int32_t n_samples;      /* N independent samples (the reference runs N = 1) */
    int32_t n_channels;     /* C pitched channels */
    int32_t n_bars;         /* B */
    int32_t n_beats;        /* beats per bar (time-signature numerator) */
    int32_t has_unpitched;  /* drums grid present */
    int32_t precision;      /* MST_PRECISION_* */
} MstModelShape;

typedef struct MstModelOutputs {   /* device pointers into the workspace, valid until it is reused */
    float *instruments;  /* (N, 41) logits */
    float *mode;         /* (N, 2) logits */
    float *bpm;          /* (N) */
    float *pitched;      /* (N, C, B, nb, 10, 56, 5) */
    float *unpitched;    /* (N, B, nb, 10, 47, 2) or NULL */
    float *style;        /* (N, 256) */
    float *melody;       /* (N, B, nb, 10, 56, 8) */
    float *rhythm;       /* (N, B, nb, 10, 32) */
} MstModelOutputs;

int         mst_model_param_count(void);
const char *mst_model_param_name(int i);
long long   mst_model_param_numel(int i);
long long   mst_model_param_offset(int i);   /* i == count: total number of parameters */

size_t mst_model_workspace_bytes(const MstModelShape *shape, int training);

/* d_pitched (N,C,B,nb,10,56,5), d_unpitched (N,B,nb,10,47,2) or NULL, d_features (N,C,51), d_mode (N,2), d_bpm (N).
 * training != 0 keeps what the backward pass needs (workspace sized with training = 1). */
int mst_model_forward(const float *d_params, const MstModelShape *shape, const float *d_pitched,
                      const float *d_unpitched, const float *d_features, const float *d_mode, const float *d_bpm,
                      MstModelOutputs *outputs, void *d_workspace, size_t workspace_bytes, int training,
                      void *stream, int *n_launches);

/* One training iteration of train-model.py:111-126 for N independent samples: forward, get_total_loss
 * (style/model.py:935-997, the positional quirk of train-model.py:115-122 resolved by the caller: "mode" is the
 * cross-entropy pair (mode logits vs the one-hot input mode), "bpm" the squared-error pair (predicted bpm vs
 * d_bpm_target)), and loss.backward().  Parameter gradients are ADDED into d_grads (flat, same layout as d_params):
 * the sum over the N samples and over successive calls, i.e. the reference's iter_size accumulation (:151-154).
 * d_losses (N, MST_N_LOSS_LEAVES) receives every leaf of the nested loss dict per sample (MST_LOSS_* indices).
 * The workspace is sized by mst_model_workspace_bytes(shape, 1).  outputs may be NULL. */
#define MST_N_LOSS_LEAVES 16
#define MST_LOSS_TOTAL            0   /* ['total'] */
#define MST_LOSS_CHANNELS         1   /* ['channels_loss']['total'] */
#define MST_LOSS_P_TOTAL          2   /* ['channels_loss']['pitched']['total'] */
#define MST_LOSS_P_NOTES          3
#define MST_LOSS_P_VELOCITY       4
#define MST_LOSS_P_DURATION       5
#define MST_LOSS_P_ACCIDENTALS    6
#define MST_LOSS_U_TOTAL          7   /* ['channels_loss']['unpitched'][...]; 0 when there is no unpitched grid */
#define MST_LOSS_U_NOTES          8
#define MST_LOSS_U_VELOCITY       9
#define MST_LOSS_U_DURATION      10
#define MST_LOSS_SONG_TOTAL      11   /* ['song_info_loss']['total'] */
#define MST_LOSS_INSTRUMENTS     12
#define MST_LOSS_MODE            13
#define MST_LOSS_BPM             14

int mst_model_train_step(const float *d_params, float *d_grads, const MstModelShape *shape, const float *d_pitched,
                         const float *d_unpitched, const float *d_features, const float *d_mode, const float *d_bpm,
                         const float *d_instruments_target /* (N,41) get_used_instruments, data.py:159 */,
                         const float *d_bpm_target /* (N) info['bpm'] */, int normalize,
                         float *d_losses, MstModelOutputs *outputs, void *d_workspace, size_t workspace_bytes,
                         void *stream, int *n_launches);

/* predict_song_info + apply_style (style/model.py:775-782, as style_transfer.py:101-131 uses them) from caller-provided
 * style (N,256), melody (N,B,nb,10,56,8) and rhythm (N,B,nb,10,32): the style vector may come from another song, and
 * n_channels / d_features / has_unpitched describe the instruments to render.  Fills outputs->instruments, mode, bpm,
 * pitched and unpitched (style / melody / rhythm echo the inputs). */
int mst_model_predict_apply(const float *d_params, const MstModelShape *shape, const float *d_style, const float *d_melody,
                            const float *d_rhythm, const float *d_features, MstModelOutputs *outputs,
                            void *d_workspace, size_t workspace_bytes, void *stream, int *n_launches);

/* predict_song_info alone (SongInfoModel, style/model.py:513-562,775-777): the three heads from style (N,256) and rhythm
 * (N,B,nb,10,32); shape->n_channels is ignored by the computation (the workspace is the forward workspace of the shape).
 * Fills outputs->instruments, mode, bpm.  With mst_transfer_select and mst_model_predict_apply this is the device-side
 * apply_style of style_transfer.py:101-131. */
int mst_model_song_info(const float *d_params, const MstModelShape *shape, const float *d_style, const float *d_rhythm,
                        MstModelOutputs *outputs, void *d_workspace, size_t workspace_bytes, void *stream, int *n_launches);

/* get_total_loss + loss.backward() on the workspace a preceding mst_model_forward(..., training = 1) call filled for the
 * SAME shape, inputs and parameters (the reference's call pattern: model(...), get_total_loss(...), loss.backward(),
 * train-model.py:113-126).  Nothing else may have used the workspace or run another forward on the device in between. */
int mst_model_backward(const float *d_params, float *d_grads, const MstModelShape *shape, const float *d_pitched,
                       const float *d_unpitched, const float *d_features, const float *d_mode, const float *d_bpm,
                       const float *d_instruments_target, const float *d_bpm_target, int normalize,
                       float *d_losses, void *d_workspace, size_t workspace_bytes, void *stream, int *n_launches);

/* get_total_loss alone (style/model.py:935-997) on caller-provided predictions / targets, N samples segmented. */
typedef struct MstLossIn {
    const float *pitched_pred, *pitched_target;         /* (N,C,B,nb,10,56,5) */
    const float *unpitched_pred, *unpitched_target;     /* (N,1,B,nb,10,47,2) or both NULL */
    const float *instruments_logits, *instruments_target; /* (N,41) */
    const float *mode_logits, *mode_target;             /* (N,2): cross-entropy against argmax(mode_target) */
    const float *bpm_pred, *bpm_target;                 /* (N): ((pred - target) / 150)^2 */
    int32_t normalize;
    int32_t reserved;
} MstLossIn;
size_t mst_model_loss_workspace_bytes(const MstModelShape *shape);
int mst_model_loss(const MstModelShape *shape, const MstLossIn *in, float *d_losses, void *d_workspace,
                   size_t workspace_bytes, void *stream);

/* mst_model_loss for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments, results and
 * error codes as mst_model_loss minus the workspace: the kernels' scratch (per-sample partial sums) is a file-scope
 * `static __device__` array of MST_LOSS_NOALLOC_MAX_SAMPLES * 1024 + 4096 bytes (266 240 B) that every GPU loading the library
 * holds until exit; its device address is resolved once per device (cudaGetSymbolAddress with that device current).
 * Neither this call nor anything below it calls cudaMalloc / cudaFree or an allocating library routine.
 * shape->n_samples outside 1..MST_LOSS_NOALLOC_MAX_SAMPLES returns MST_E_CAPACITY (use mst_model_loss then).
 * All calls on a device share the one array (it keeps the previous call's contents; each call clears the sums it
 * accumulates into, on `stream`): calls whose kernels may overlap on the GPU (different streams without an ordering
 * between them) race on it and must be serialised by the caller. */
#define MST_LOSS_NOALLOC_MAX_SAMPLES 256
int mst_model_loss_noalloc(const MstModelShape *shape, const MstLossIn *in, float *d_losses, void *stream);

/* torch.optim.Adam.step (train-model.py:89,152; defaults betas (.9,.999), eps 1e-8, no weight decay) on the flat
 * buffers, `step` = 1-based optimizer step count (bias correction), lr already scheduled by the caller
 * (StepLR(200, .9), :90,154).  grad_scale multiplies the gradient first (1 = the reference's summed gradients);
 * zero_grad != 0 clears d_grads in the same pass (optimizer.zero_grad(), :153). */
int mst_adam_step(float *d_params, float *d_grads, float *d_m, float *d_v, long long n, int step, float lr,
                  float beta1, float beta2, float eps, float grad_scale, int zero_grad, void *stream);

/* Dense building block of the model path, exposed for parity tests and profiling: C[M,N] (op)= act(A.B + bias) with
 * explicit element strides (A(i,k) = A[i*a_si + k*a_sk], B(k,j) = B[k*b_sk + j*b_sj], C row stride c_si).  This is the
 * nn.Linear / Distributed(Linear) forward (model.py:23-33), input-gradient and weight-gradient product.
 * act: 0 none, 1 LeakyReLU(.01).  accumulate: 0 store, 1 +=, 2 atomic +=.  split_k > 1 splits K over CTAs (atomic).
 * kernel: 0 = the library's own choice, 1 = fp32 SIMT tiles, 2 = mma.sync TF32 (3-term split: fp32-accurate),
 * 3 = tcgen05.mma kind::tf32 with the accumulator in TMEM (same 3-term split), 4 = tcgen05.mma kind::f16 with bf16
 * operands (the MST_PRECISION_BF16 mode: NOT fp32-accurate). */
int mst_model_gemm(const float *d_A, long long a_si, long long a_sk, const float *d_B, long long b_sk, long long b_sj,
                   float *d_C, long long c_si, int M, int N, int K, const float *d_bias, int act, int accumulate,
                   int split_k, int kernel, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* MST_B200_H */
