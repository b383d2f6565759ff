/*
 * mst_oracle.c — CPU restatement of the reference's MIDI <-> piano-roll codec.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library; the product path (music-style-transfer_b200/)
 * never does and has no CPU fallback.
 *
 * Parity status: PINNED against outputs of the unmodified reference executed in the build
 * container (oracle/gen_golden.py -> tests/golden/*.npz, all 34 bundled MIDI files plus seeded
 * synthetic quirk cases).  The reference ships no tests or golden vectors of its own, and its SMF
 * parsing lives in the un-vendored, un-pinned third-party `mido`; see DESIGN.md "Oracle".
 *
 * Every function cites the reference lines it follows (paths relative to the reference tree).
 * Plain sequential C on purpose: it is the checker, not the thing measured.
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/mst_b200.h"

/* ------------------------------------------------------------------ small helpers */

/* CPython float % float for positive divisor (Objects/floatobject.c float_rem), as used by
 * round_number (style/utils/math.py:15). */
static double py_fmod_pos(double v, double w) {
    double mod = fmod(v, w);
    if (mod != 0.0) {
        if ((w < 0) != (mod < 0)) mod += w;
    } else {
        mod = copysign(0.0, w);
    }
    return mod;
}

/* CPython float // float (float_divmod -> floor part), as used at midi_conversion.py:451. */
static double py_floordiv(double vx, double wx) {
    double mod = fmod(vx, wx);
    double div = (vx - mod) / wx;
    if (mod != 0.0) {
        if ((wx < 0) != (mod < 0)) {
            mod += wx;
            div -= 1.0;
        }
    }
    double floordiv;
    if (div != 0.0) {
        floordiv = floor(div);
        if (div - floordiv > 0.5) floordiv += 1.0;
    } else {
        floordiv = copysign(0.0, vx / wx);
    }
    return floordiv;
}

/* numpy pairwise_sum for n < 128 contiguous doubles (numpy/_core/src/umath/loops_utils.h.src):
 * n < 8 plain loop, else 8 accumulators combined as a tree, remainder added in order. */
static double np_sum(const double *a, int n) {
    if (n < 8) {
        double res = 0.0;  /* numpy starts from the first element; 0.+a0 is exact */
        for (int i = 0; i < n; i++) res += a[i];
        return res;
    }
    double r[8];
    for (int j = 0; j < 8; j++) r[j] = a[j];
    int i;
    for (i = 8; i < n - (n % 8); i += 8)
        for (int j = 0; j < 8; j++) r[j] += a[i + j];
    double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
    for (; i < n; i++) res += a[i];
    return res;
}

/* Python >= 3.12 builtin sum() over floats: Neumaier compensated summation
 * (Python/bltinmodule.c builtin_sum_impl).  Used by get_keys_dist (midi_conversion.py:341-342). */
typedef struct { double s, c; int started; } PySum;
static void pysum_add(PySum *p, double x) {
    if (!p->started) { p->s = 0.0; p->c = 0.0; p->started = 1; }
    double t = p->s + x;
    if (fabs(p->s) >= fabs(x)) p->c += (p->s - t) + x;
    else p->c += (x - t) + p->s;
    p->s = t;
}
static double pysum_result(const PySum *p) {
    if (!p->started) return 0.0;
    double r = p->s;
    if (p->c != 0.0 && isfinite(p->c)) r += p->c;
    return r;
}

/* ------------------------------------------------------------------ scale tables */
/* Mode.interval2degree (style/scales.py:54-63) + degree2accidental / note2scale_loc
 * (midi_conversion.py:235-266), folded into LUTs indexed by (pitch class - tonic) mod 12.
 * accidental: 0 flat, 1 none, 2 sharp (the feature order of :504-509). */
static const int8_t k_degree[2][12] = {
    /* major  */ {1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 7, 7},
    /* minor  */ {1, 2, 2, 3, 4, 4, 5, 5, 6, 6, 7, 7},
};
static const int8_t k_accidental[2][12] = {
    /* major  */ {1, 0, 1, 0, 1, 1, 2, 1, 2, 1, 0, 1},
    /* minor  */ {1, 0, 1, 1, 0, 1, 0, 1, 1, 2, 1, 2},
};
/* Mode.absolute_intervals (scales.py:49-51) for major and create_mode(major, -2) */
static const int8_t k_abs_intervals[2][7] = {
    {0, 2, 4, 5, 7, 9, 11},
    {0, 2, 3, 5, 7, 8, 10},
};

/* style/scales.py:111-124 */
static const double k_major_profile[12] = {6.35, 2.23, 3.48, 2.33, 4.38, 4.09, 2.52, 5.19, 2.39, 3.66, 2.29, 2.88};
static const double k_minor_profile[12] = {6.33, 2.68, 3.52, 5.38, 2.60, 3.53, 2.54, 4.75, 3.98, 2.69, 3.34, 3.17};
static const int k_typical_major[9] = {0, 2, 4, 5, 6, 7, 9, 10, 11};
static const int k_typical_minor[10] = {0, 1, 2, 3, 5, 7, 8, 9, 10, 11};

/* fraction column tables: sorted {0,1/8,1/4,1/3,3/8,1/2,5/8,2/3,3/4,7/8} (midi_conversion.py:358-364) */
static const int8_t k_col8[8] = {0, 1, 2, 4, 5, 6, 8, 9};
static const int8_t k_col3[3] = {0, 3, 7};
static const int8_t k_frac_num[10] = {0, 1, 1, 1, 3, 1, 5, 2, 3, 7};
static const int8_t k_frac_den[10] = {1, 8, 4, 3, 8, 2, 8, 3, 4, 8};

static int is_pitched(int instrument_id) { return instrument_id >= 0 && instrument_id <= 119; }

/* ------------------------------------------------------------------ intermediate records */
typedef struct MstoNoteMsg {   /* NoteMessage, midi_conversion.py:101-106 */
    uint32_t time;
    uint8_t  is_on;
    uint8_t  note;
    uint16_t list;
    double   velocity;
} MstoNoteMsg;

typedef struct MstoList {      /* one read_midi channel entry, midi_conversion.py:225-230 */
    int32_t channel_id;
    int32_t instrument_id;
    uint32_t msg_offset;
    uint32_t n_messages;
} MstoList;

typedef struct MstoNote {      /* Note, midi_conversion.py:286-306 */
    int32_t note_id;
    int32_t time;
    int32_t end_time;
    uint32_t seq;              /* position in the concatenation: makes qsort stable */
    double  velocity;
} MstoNote;

/* ------------------------------------------------------------------ stage R: read_midi */
typedef struct { uint32_t tick; uint32_t idx; } SortKey;
static int cmp_sortkey(const void *a, const void *b) {
    const SortKey *x = (const SortKey *)a, *y = (const SortKey *)b;
    if (x->tick != y->tick) return x->tick < y->tick ? -1 : 1;
    return x->idx < y->idx ? -1 : (x->idx > y->idx);
}

static int has_channel(int kind) { return kind <= MST_EV_PITCHWHEEL || kind == MST_EV_CHANNEL_PREFIX; }

/* tempo -> accumulated ticks, insertion ordered (get_midi_info's tempo2total_time, :120,159-176) */
typedef struct { int32_t tempo; int64_t total; } TempoSlot;

/*
 * read_midi (midi_conversion.py:216-232) = merge_tracks (:37-49) + split_channels (:55-66) +
 * get_midi_info (:117-179) + group_channel_messages (:182-210).
 * lists/msgs must hold up to 16*MST_MAX_LISTS_PER_CHANNEL lists and n_events messages.
 */
int msto_read_midi(const MstEvent *ev, uint32_t n_events, int32_t tpb, MstSongInfo *info,
                   MstoList *lists, MstoNoteMsg *msgs, uint32_t *n_msgs_out) {
    memset(info, 0, sizeof(*info));
    info->ticks_per_beat = tpb;
    info->numerator = 4;
    info->denominator_log2 = 2;
    info->key_signature = -1;
    *n_msgs_out = 0;

    /* merge_tracks: stable sort by absolute tick; split_channels drops time > 1e7 */
    SortKey *order = (SortKey *)malloc(sizeof(SortKey) * (n_events ? n_events : 1));
    uint32_t n = 0;
    for (uint32_t i = 0; i < n_events; i++)
        if (ev[i].tick <= MST_MAX_TICK) { order[n].tick = ev[i].tick; order[n].idx = i; n++; }
    qsort(order, n, sizeof(SortKey), cmp_sortkey);

    /* split_channels: discovery-ordered channel keys (0..255 because channel_prefix carries one) */
    int discovered[256]; int n_disc = 0; uint8_t seen[256]; memset(seen, 0, sizeof(seen));
    int64_t first_on = -1, last_on = -1, duration = -1;
    for (uint32_t r = 0; r < n; r++) {
        const MstEvent *e = &ev[order[r].idx];
        if (!has_channel(e->kind)) continue;
        int ch = e->channel;
        if (!seen[ch]) { seen[ch] = 1; discovered[n_disc++] = ch; }
        if ((int64_t)e->tick > duration) duration = e->tick;
        if (e->kind == MST_EV_NOTE_ON && e->d2 > 0) {
            if (first_on < 0 || (int64_t)e->tick < first_on) first_on = e->tick;
            if ((int64_t)e->tick > last_on) last_on = e->tick;
        }
    }
    if (first_on < 0) { info->status = MST_SONG_NO_NOTES; free(order); return 0; }
    info->duration = (int32_t)duration;

    /* get_midi_info: walk the global messages in merged order */
    int32_t tempo = 500000; int64_t tempo_change_time = 0;
    TempoSlot *slots = (TempoSlot *)malloc(sizeof(TempoSlot) * (n + 2));
    int n_slots = 0;
    int status = MST_SONG_OK;
    int numerator = 4, den_log2 = 2, key = -1;
    for (uint32_t r = 0; r < n && status == MST_SONG_OK; r++) {
        const MstEvent *e = &ev[order[r].idx];
        if (has_channel(e->kind) || e->kind == MST_EV_META_IGNORED) continue;
        int during = (first_on <= (int64_t)e->tick && (int64_t)e->tick <= last_on);
        if (e->kind == MST_EV_TIME_SIGNATURE) {
            if (e->d1 != numerator || e->d2 != den_log2) {
                if (during) { status = MST_SONG_TIME_SIG_CHANGED; break; }
                numerator = e->d1; den_log2 = e->d2;
            }
        } else if (e->kind == MST_EV_KEY_SIGNATURE) {
            if ((int)e->aux != key) {
                if (during) { status = MST_SONG_KEY_SIG_CHANGED; break; }
                key = (int)e->aux;
            }
        } else if (e->kind == MST_EV_SET_TEMPO) {
            if ((int32_t)e->aux != tempo) {
                int s; for (s = 0; s < n_slots; s++) if (slots[s].tempo == tempo) break;
                if (s == n_slots) { slots[n_slots].tempo = tempo; slots[n_slots].total = 0; n_slots++; }
                slots[s].total += (int64_t)e->tick - tempo_change_time;
                tempo = (int32_t)e->aux;
                tempo_change_time = e->tick;
            }
        } else {
            status = MST_SONG_UNKNOWN_MESSAGE;
        }
    }
    info->numerator = numerator; info->denominator_log2 = den_log2; info->key_signature = key;
    if (status != MST_SONG_OK) { info->status = status; free(order); free(slots); return 0; }
    if (tpb <= 0 || numerator == 0) { info->status = MST_SONG_BAD_TIMING; free(order); free(slots); return 0; }

    info->ticks_per_bar = tpb * numerator;                       /* :167 */
    info->n_bars = (int32_t)((duration + info->ticks_per_bar - 1) / info->ticks_per_bar); /* ceil of :169 */
    info->n_beats = numerator;                                   /* :170 */
    {
        int s; for (s = 0; s < n_slots; s++) if (slots[s].tempo == tempo) break;
        if (s == n_slots) { slots[n_slots].tempo = tempo; slots[n_slots].total = 0; n_slots++; }
        slots[s].total += duration - tempo_change_time;         /* :172 */
        int best = -1;
        for (s = 0; s < n_slots; s++) {                          /* :173,176 first max among non-zero */
            if (slots[s].total == 0) continue;
            if (best < 0 || slots[s].total > slots[best].total) best = s;
        }
        if (best < 0) { info->status = MST_SONG_BAD_TIMING; free(order); free(slots); return 0; }
        info->tempo = slots[best].tempo;
        /* round(mido.tempo2bpm(tempo)): 60*1e6/tempo*4/4., banker's rounding (:177) */
        info->bpm = (int32_t)nearbyint(60 * 1e6 / (double)info->tempo * 4 / 4.);
    }
    free(slots);

    /* group_channel_messages per channel in discovery order */
    int n_lists = 0; uint32_t n_msgs = 0;
    MstoNoteMsg *tmp = (MstoNoteMsg *)malloc(sizeof(MstoNoteMsg) * (n ? n : 1));
    for (int d = 0; d < n_disc; d++) {
        int ch = discovered[d];
        if (ch > 15) continue;   /* only channel_prefix metas live here: never a note_on */
        int instrument_id = (ch == 9) ? -1 : 0;                   /* get_instrument_id, midi.py:90-93 */
        int volume = 96;
        int slot_instr[MST_MAX_LISTS_PER_CHANNEL]; int n_slots_ch = 0;
        uint32_t n_tmp = 0;
        for (uint32_t r = 0; r < n; r++) {
            const MstEvent *e = &ev[order[r].idx];
            if (!has_channel(e->kind) || e->channel != ch) continue;
            if (e->kind == MST_EV_PROGRAM_CHANGE) {
                instrument_id = (ch == 9) ? -1 : e->d1;
            } else if (e->kind == MST_EV_CONTROL_CHANGE && e->d1 == 7) {
                volume = e->d2;
            } else if (e->kind == MST_EV_NOTE_ON || e->kind == MST_EV_NOTE_OFF) {
                double velocity = (double)(e->d2 * volume) / (double)(127 * 127);   /* :199 */
                int is_on = (velocity == 0.0) ? 0 : (e->kind == MST_EV_NOTE_ON);
                int s; for (s = 0; s < n_slots_ch; s++) if (slot_instr[s] == instrument_id) break;
                if (s == n_slots_ch) {
                    if (n_slots_ch == MST_MAX_LISTS_PER_CHANNEL) {
                        info->status = MST_SONG_TOO_MANY_INSTRUMENTS; free(order); free(tmp); return 0;
                    }
                    slot_instr[n_slots_ch++] = instrument_id;
                }
                tmp[n_tmp].time = e->tick; tmp[n_tmp].is_on = (uint8_t)is_on; tmp[n_tmp].note = e->d1;
                tmp[n_tmp].list = (uint16_t)s; tmp[n_tmp].velocity = velocity; n_tmp++;
            }
        }
        for (int s = 0; s < n_slots_ch; s++) {
            uint32_t cnt = 0; int any_on = 0;
            for (uint32_t i = 0; i < n_tmp; i++) if (tmp[i].list == s) { cnt++; any_on |= tmp[i].is_on; }
            if (!any_on) continue;                                 /* :224 */
            lists[n_lists].channel_id = ch; lists[n_lists].instrument_id = slot_instr[s];
            lists[n_lists].msg_offset = n_msgs; lists[n_lists].n_messages = cnt;
            for (uint32_t i = 0; i < n_tmp; i++) if (tmp[i].list == s) {
                msgs[n_msgs] = tmp[i]; msgs[n_msgs].list = (uint16_t)n_lists; n_msgs++;
            }
            n_lists++;
        }
    }
    free(tmp); free(order);
    info->n_lists = n_lists;
    *n_msgs_out = n_msgs;
    return n_lists;
}

/* ------------------------------------------------------------------ stage G: get_input */

/* channel2nchannel (midi_conversion.py:371-406): note pairing inside one list. */
static uint32_t pair_notes(const MstoNoteMsg *m, uint32_t n, MstoNote *out) {
    int32_t open[128]; for (int i = 0; i < 128; i++) open[i] = -1;
    uint32_t k = 0;
    for (uint32_t i = 0; i < n; i++) {
        int p = m[i].note;
        if (open[p] >= 0) { out[open[p]].end_time = (int32_t)m[i].time; open[p] = -1; }
        if (m[i].is_on) {
            out[k].note_id = p; out[k].time = (int32_t)m[i].time; out[k].end_time = (int32_t)m[i].time;
            out[k].velocity = m[i].velocity; open[p] = (int32_t)k; k++;
        }
    }
    return k;
}

static int cmp_note_time(const void *a, const void *b) {   /* (time, seq) = stable sort by time */
    const MstoNote *x = (const MstoNote *)a, *y = (const MstoNote *)b;
    if (x->time != y->time) return x->time < y->time ? -1 : 1;
    return x->seq < y->seq ? -1 : (x->seq > y->seq);
}

/* get_scales + get_key_scores + get_scale (style/scales.py:160-221), cross_entropy
 * (utils/metrics.py:4-8), normalize_dist (utils/math.py:4-11). */
static void choose_scale(const double *keys_dist_in, MstSongInfo *info) {
    double profile[2][12];
    for (int m = 0; m < 2; m++) {
        const double *p = m ? k_minor_profile : k_major_profile;
        double tot = np_sum(p, 12);
        for (int i = 0; i < 12; i++) profile[m][i] = p[i] / tot;
    }
    double best = 0, second = INFINITY; int best_idx = -1;
    double best_cov = 0, best_loose = 0, best_ce = 0;
    for (int m = 0; m < 2; m++) {
        double kd[12]; memcpy(kd, keys_dist_in, sizeof(kd));
        for (int t = 0; t < 12; t++) {
            double g[12];
            for (int i = 0; i < 7; i++) g[i] = kd[k_abs_intervals[m][i]];
            double coverage = np_sum(g, 7);
            int nt = m ? 10 : 9; const int *typ = m ? k_typical_minor : k_typical_major;
            for (int i = 0; i < nt; i++) g[i] = kd[typ[i]];
            double loose = np_sum(g, nt);
            for (int i = 0; i < 12; i++) {
                double d = kd[i];
                /* np.clip(dist, 1e-12, 1.): NaN propagates */
                if (d < 1e-12) d = 1e-12; else if (d > 1.0) d = 1.0;
                g[i] = profile[m][i] * log(d);
            }
            double ce = -np_sum(g, 12) / 12;
            double loss = ce * (1.5 - coverage) * (2 - loose);
            /* min(scales, key=loss): first minimum; NaN never compares smaller */
            if (best_idx < 0) { best = loss; best_idx = m * 12 + t; best_cov = coverage; best_loose = loose; best_ce = ce; }
            else if (loss < best) { second = best; best = loss; best_idx = m * 12 + t; best_cov = coverage; best_loose = loose; best_ce = ce; }
            else if (loss < second) second = loss;
            double first = kd[0]; for (int i = 0; i < 11; i++) kd[i] = kd[i + 1]; kd[11] = first;  /* :211 */
        }
    }
    info->tonic = best_idx % 12; info->mode = best_idx / 12;
    info->scale_coverage = best_cov; info->scale_loose_coverage = best_loose;
    info->scale_cross_entropy = best_ce; info->scale_loss = best; info->scale_margin = second - best;
}

/* kchannel2qchannel (midi_conversion.py:425-456) + round_number (utils/math.py:14-19) for one note.
 * Returns the fraction column; fills qtime/bar/beat. */
static int quantise(int32_t time, int32_t tpb, int32_t tpbar, int32_t *qtime, int32_t *bar, int32_t *beat) {
    double best_q = 0, best_err = 0; int best_d = 0;
    const int divisors[2] = {8, 3};
    for (int k = 0; k < 2; k++) {
        double ticks = (double)tpb / (double)divisors[k];
        double rpos = py_fmod_pos((double)time, ticks);
        double rneg = fabs(rpos - ticks);
        double q, err;
        if (rpos < rneg) { q = (double)time - rpos; err = rpos; }
        else { q = (double)time + rneg; err = -rneg; }
        double tot = fabs(err);
        if (k == 0 || tot < best_err) { best_q = q; best_err = tot; best_d = divisors[k]; }
    }
    int32_t qt = (int32_t)best_q;            /* int(): truncation, :447 */
    *qtime = qt;
    int32_t b = qt / tpbar, rem = qt % tpbar;    /* qt >= 0 */
    int32_t bt = rem / tpb; rem = rem % tpb;
    *bar = b; *beat = bt;
    int quants = (int)py_floordiv((double)rem, (double)tpb / (double)best_d);
    if (best_d == 8) return (quants >= 0 && quants < 8) ? k_col8[quants] : -1;
    return (quants >= 0 && quants < 3) ? k_col3[quants] : -1;
}

/*
 * get_input (style/data.py:66-100) on the lists selected by `keep` (the caller's filter, data.py:53-56),
 * followed by prepare_input's crop and f32 cast (data.py:133-143).
 * Pass pitched/unpitched == NULL to only plan (info is filled, sizes known).
 */
typedef struct MstoScratch { float *p; uint64_t cap_p; float *u; uint64_t cap_u; } MstoScratch;

int msto_get_input(const MstoList *lists, int n_lists, const uint8_t *keep, const MstoNoteMsg *msgs,
                   MstSongInfo *info, int crop_channel_bars,
                   float *pitched, uint64_t pitched_cap, float *unpitched, uint64_t unpitched_cap,
                   MstoScratch *grow) {
    int32_t tpb = info->ticks_per_beat, tpbar = info->ticks_per_bar, nb = info->n_beats;
    /* group_by instrument_id in first-seen order (data.py:69) */
    int instr[MST_MAX_PITCHED + MST_MAX_UNPITCHED]; int n_instr = 0;
    uint32_t total_msgs = 0;
    for (int l = 0; l < n_lists; l++) {
        if (keep && !keep[l]) continue;
        total_msgs += lists[l].n_messages;
        int i; for (i = 0; i < n_instr; i++) if (instr[i] == lists[l].instrument_id) break;
        if (i == n_instr) instr[n_instr++] = lists[l].instrument_id;
    }
    int C = 0, U = 0;
    for (int i = 0; i < n_instr; i++) {
        if (is_pitched(instr[i])) info->pitched_instruments[C++] = (int16_t)instr[i];
        else info->unpitched_instruments[U++] = (int16_t)instr[i];
    }
    info->n_pitched = C; info->n_unpitched = U;
    if (C == 0) { info->status = MST_SONG_NO_PITCHED; return 0; }

    /* notes per instrument: concat lists in order, stable sort by time (merge_nchannels, data.py:103-114) */
    MstoNote *notes = (MstoNote *)malloc(sizeof(MstoNote) * (total_msgs ? total_msgs : 1));
    uint32_t *inst_off = (uint32_t *)calloc(n_instr + 1, sizeof(uint32_t));
    uint32_t k = 0;
    for (int i = 0; i < n_instr; i++) {
        inst_off[i] = k;
        for (int l = 0; l < n_lists; l++) {
            if ((keep && !keep[l]) || lists[l].instrument_id != instr[i]) continue;
            k += pair_notes(msgs + lists[l].msg_offset, lists[l].n_messages, notes + k);
        }
        for (uint32_t j = inst_off[i]; j < k; j++) notes[j].seq = j;
        qsort(notes + inst_off[i], k - inst_off[i], sizeof(MstoNote), cmp_note_time);
    }
    inst_off[n_instr] = k;

    /* get_keys_dist (midi_conversion.py:340-346) per pitched instrument, summed (data.py:79-83) */
    double scale_t2s = (double)info->tempo * 1e-6 / (double)tpb;   /* mido.tick2second */
    double hist[12]; int hist_set[12];
    for (int i = 0; i < 12; i++) { hist[i] = 0.0; hist_set[i] = 0; }
    for (int i = 0; i < n_instr; i++) {
        if (!is_pitched(instr[i])) continue;
        PySum ps[12]; memset(ps, 0, sizeof(ps));
        for (uint32_t j = inst_off[i]; j < inst_off[i + 1]; j++) {
            int key = notes[j].note_id % 12;
            double dur = (double)(notes[j].end_time - notes[j].time);
            pysum_add(&ps[key], dur * notes[j].velocity);
        }
        for (int key = 0; key < 12; key++) if (ps[key].started) {
            double sec = pysum_result(&ps[key]) * scale_t2s;
            hist[key] = hist_set[key] ? hist[key] + sec : sec;   /* column sum over instruments, NaN skipped */
            hist_set[key] = 1;
        }
    }
    double total = np_sum(hist, 12);
    for (int i = 0; i < 12; i++) { hist[i] = hist[i] / total; info->keys_dist[i] = hist[i]; }
    choose_scale(hist, info);

    /* grids */
    int n_bars_grid = info->n_bars + 1;                                     /* :493 */
    int n_out = n_bars_grid;
    if (crop_channel_bars > 0) { int mx = crop_channel_bars / C; if (mx < n_out) n_out = mx; }
    info->n_bars_out = n_out;
    uint64_t p_floats = (uint64_t)C * n_out * nb * MST_PITCHED_BEAT_FLOATS;
    uint64_t u_floats = (uint64_t)U * n_out * nb * MST_UNPITCHED_BEAT_FLOATS;
    info->pitched_offset = 0; info->unpitched_offset = p_floats;
    int n_scattered = 0;
    if (grow) {   /* output allocated here, like the reference's np.zeros (midi_conversion.py:590-595) */
        if (grow->cap_p < p_floats) { free(grow->p); grow->p = (float *)malloc(p_floats * sizeof(float)); grow->cap_p = p_floats; }
        if (grow->cap_u < u_floats) { free(grow->u); grow->u = (float *)malloc(u_floats * sizeof(float)); grow->cap_u = u_floats; }
        pitched = grow->p; pitched_cap = grow->cap_p; unpitched = grow->u; unpitched_cap = grow->cap_u;
    }
    int write = (pitched != NULL);
    if (write) {
        if (p_floats > pitched_cap || (U && (unpitched == NULL || u_floats > unpitched_cap))) {
            info->status = MST_SONG_ARENA_FULL; free(notes); free(inst_off); return 0;
        }
        memset(pitched, 0, p_floats * sizeof(float));
        if (U) memset(unpitched, 0, u_floats * sizeof(float));
    }
    int ci = 0, ui = 0;
    for (int i = 0; i < n_instr; i++) {
        int pitched_i = is_pitched(instr[i]);
        int slot = pitched_i ? ci++ : ui++;
        for (uint32_t j = inst_off[i]; j < inst_off[i + 1]; j++) {
            const MstoNote *nt = &notes[j];
            int row, acc = 1;
            if (pitched_i) {
                /* note_id2key_octave (:309-317) + note2scale_loc (:244-266) + note2idx (:597-604) */
                int octave = nt->note_id / 12 - 1, pc = nt->note_id % 12;
                int interval = pc - info->tonic;
                if (interval < 0) { octave -= 1; interval += 12; }
                int degree = k_degree[info->mode][interval];
                acc = k_accidental[info->mode][interval];
                row = octave * 7 + degree - 1;
                if (row < 0 || row >= MST_N_PITCHED_ROWS) continue;
            } else {
                if (nt->note_id < 35 || nt->note_id > 81) continue;   /* :605-609 */
                row = nt->note_id - 35;
            }
            int32_t qtime, bar, beat;
            int col = quantise(nt->time, tpb, tpbar, &qtime, &bar, &beat);
            if (col < 0 || bar >= n_bars_grid || beat >= nb) { continue; }  /* reference would raise */
            n_scattered++;
            if (!write || bar >= n_out) continue;
            int32_t qdur = nt->end_time - qtime;                      /* :448 */
            double duration = (double)qdur / (double)tpb;             /* :501 */
            if (pitched_i) {
                float *cell = pitched + ((((uint64_t)slot * n_out + bar) * nb + beat) * MST_N_FRACTIONS + col)
                                            * (MST_N_PITCHED_ROWS * MST_N_PITCHED_FEATS) + row * MST_N_PITCHED_FEATS;
                float d = (float)duration, v = (float)nt->velocity;  /* np.maximum (:514) then f32 (data.py:136) */
                if (d > cell[0]) cell[0] = d;
                if (v > cell[1]) cell[1] = v;
                cell[2 + acc] = 1.0f;
            } else {
                float *cell = unpitched + ((((uint64_t)slot * n_out + bar) * nb + beat) * MST_N_FRACTIONS + col)
                                              * (MST_N_UNPITCHED_ROWS * MST_N_UNPITCHED_FEATS) + row * MST_N_UNPITCHED_FEATS;
                float d = (float)duration, v = (float)nt->velocity;
                if (d > cell[0]) cell[0] = d;
                if (v > cell[1]) cell[1] = v;
            }
        }
    }
    info->n_notes = n_scattered;
    free(notes); free(inst_off);
    return 0;
}

/* iter_inputs' channel filter (style/data.py:53-56). */
static void apply_filter(const MstoList *lists, int n_lists, const MstFilter *f, uint8_t *keep) {
    for (int l = 0; l < n_lists; l++) {
        keep[l] = 1;
        if (!f || !f->apply) continue;
        int id = lists[l].instrument_id;
        int ok = (id == -1) || (id >= 0 && id < 128 && f->included[id]);
        if (!ok || (int32_t)lists[l].n_messages < f->min_n_messages) keep[l] = 0;
    }
}

/*
 * Whole encode of one song: events -> info + f32 grids (what iter_inputs + prepare_input yield).
 * pitched == NULL: plan only.  Returns 0; per-song problems are in info->status.
 */
static int encode_song_impl(const MstEvent *ev, uint32_t n_events, int32_t tpb, const MstFilter *filter,
                     MstSongInfo *info, float *pitched, uint64_t pitched_cap,
                     float *unpitched, uint64_t unpitched_cap, MstoScratch *grow) {
    MstoList lists[16 * MST_MAX_LISTS_PER_CHANNEL];
    MstoNoteMsg *msgs = (MstoNoteMsg *)malloc(sizeof(MstoNoteMsg) * (n_events ? n_events : 1));
    uint32_t n_msgs = 0;
    int n_lists = msto_read_midi(ev, n_events, tpb, info, lists, msgs, &n_msgs);
    if (info->status == MST_SONG_OK) {
        uint8_t keep[16 * MST_MAX_LISTS_PER_CHANNEL];
        apply_filter(lists, n_lists, filter, keep);
        msto_get_input(lists, n_lists, keep, msgs, info, filter ? filter->crop_channel_bars : 0,
                       pitched, pitched_cap, unpitched, unpitched_cap, grow);
    }
    free(msgs);
    return 0;
}

int msto_encode_song(const MstEvent *ev, uint32_t n_events, int32_t tpb, const MstFilter *filter,
                     MstSongInfo *info, float *pitched, uint64_t pitched_cap,
                     float *unpitched, uint64_t unpitched_cap) {
    return encode_song_impl(ev, n_events, tpb, filter, info, pitched, pitched_cap, unpitched, unpitched_cap, NULL);
}

/* Timed CPU baseline (bench.py cpu_baseline / --impl reference): a batch of songs on n_threads host
 * threads (pthreads, dynamic work queue), each writing its grids into a thread-private buffer.
 * Returns a checksum of a strided sample of all grids. */
typedef struct {
    const MstEvent *ev; const MstSongIn *songs; uint32_t n_songs; const MstFilter *filter; MstSongInfo *infos;
    volatile uint32_t *next; double checksum;
} BatchJob;

static void *batch_worker(void *arg) {
    BatchJob *job = (BatchJob *)arg;
    MstoScratch sc = {NULL, 0, NULL, 0};
    double checksum = 0.0;
    for (;;) {
        uint32_t s = __atomic_fetch_add(job->next, 1u, __ATOMIC_RELAXED);
        if (s >= job->n_songs) break;
        encode_song_impl(job->ev + job->songs[s].event_offset, job->songs[s].n_events, job->songs[s].ticks_per_beat,
                         job->filter, &job->infos[s], NULL, 0, NULL, 0, &sc);
        if (job->infos[s].status == MST_SONG_OK) {
            uint64_t n = job->infos[s].unpitched_offset;
            for (uint64_t i = 0; i < n; i += 4099) checksum += sc.p[i];
        }
    }
    free(sc.p); free(sc.u);
    job->checksum = checksum;
    return NULL;
}

double msto_encode_batch(const MstEvent *ev, const MstSongIn *songs, uint32_t n_songs, const MstFilter *filter,
                         MstSongInfo *infos, int n_threads) {
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 256) n_threads = 256;
    pthread_t th[256]; BatchJob jobs[256];
    volatile uint32_t next = 0;
    for (int t = 0; t < n_threads; t++) {
        jobs[t].ev = ev; jobs[t].songs = songs; jobs[t].n_songs = n_songs; jobs[t].filter = filter;
        jobs[t].infos = infos; jobs[t].next = &next; jobs[t].checksum = 0.0;
        if (t > 0) pthread_create(&th[t], NULL, batch_worker, &jobs[t]);
    }
    batch_worker(&jobs[0]);
    double checksum = jobs[0].checksum;
    for (int t = 1; t < n_threads; t++) { pthread_join(th[t], NULL); checksum += jobs[t].checksum; }
    return checksum;
}

/* Export of the stage-R result for tests of read_midi parity (lists + message counts). */
int msto_read_midi_lists(const MstEvent *ev, uint32_t n_events, int32_t tpb, MstSongInfo *info,
                         int32_t *channel_ids, int32_t *instrument_ids, int32_t *n_messages, int cap) {
    MstoList lists[16 * MST_MAX_LISTS_PER_CHANNEL];
    MstoNoteMsg *msgs = (MstoNoteMsg *)malloc(sizeof(MstoNoteMsg) * (n_events ? n_events : 1));
    uint32_t n_msgs = 0;
    int n_lists = msto_read_midi(ev, n_events, tpb, info, lists, msgs, &n_msgs);
    for (int l = 0; l < n_lists && l < cap; l++) {
        channel_ids[l] = lists[l].channel_id; instrument_ids[l] = lists[l].instrument_id;
        n_messages[l] = (int32_t)lists[l].n_messages;
    }
    free(msgs);
    return n_lists;
}

size_t msto_sizeof_song_info(void) { return sizeof(MstSongInfo); }
size_t msto_sizeof_event(void) { return sizeof(MstEvent); }

/* ================================================================== decode (grids -> messages) */

typedef struct { int32_t time; uint32_t seq; uint8_t is_on, channel, note, velocity; } DecMsg;
static int cmp_decmsg(const void *a, const void *b) {   /* (time, seq): stable sort by time */
    const DecMsg *x = (const DecMsg *)a, *y = (const DecMsg *)b;
    if (x->time != y->time) return x->time < y->time ? -1 : 1;
    return x->seq < y->seq ? -1 : (x->seq > y->seq);
}

/*
 * decode_midi (style/style_transfer.py:145-158): hard_output (style/model.py:818-832) ->
 * vchannel2qchannel (midi_conversion.py:518-562) -> qchannel2channel (:458-488, scale_loc2key_octave
 * :269-283) -> create_midi (style/midi.py:120-168).  `grids` is the float buffer the song's offsets
 * point into; with song->write_back the gated velocities are zeroed in place like `velocity *= mask`.
 * Environment pin: numpy >= 2 keeps np.float32 * int in f32.
 */
int msto_decode_song(float *grids, const MstDecodeSong *d, MstOutMsg *out, uint32_t capacity,
                     MstDecodeResult *res) {
    const float thr_v = 0.01f, thr_a = 0.1f;   /* torch compares f32 with the scalar cast to f32 */
    const int nb = d->n_beats, tpb = d->ticks_per_beat, tpbar = d->ticks_per_bar;
    memset(res, 0, sizeof(*res));
    /* mido.second2tick(1, tpb, tempo) = int(round(1 / (tempo * 1e-6 / tpb))), then int() (midi.py:121-123) */
    {
        double scale = (double)d->tempo * 1e-6 / (double)tpb;
        res->max_delta = (int32_t)nearbyint(1.0 / scale);
    }
    uint64_t cells = (uint64_t)d->n_bars * nb * MST_N_FRACTIONS;
    uint64_t cap_msgs = 2 * ((uint64_t)d->n_pitched * cells * MST_N_PITCHED_ROWS +
                             (d->has_unpitched ? cells * MST_N_UNPITCHED_ROWS : 0));
    if (cap_msgs > (1u << 26)) cap_msgs = 1u << 26;
    DecMsg *m = (DecMsg *)malloc(sizeof(DecMsg) * (cap_msgs ? cap_msgs : 1));
    uint32_t n = 0; int status = MST_DECODE_OK;
    int n_ch = d->n_pitched + (d->has_unpitched ? 1 : 0);
    for (int c = 0; c < n_ch && status == MST_DECODE_OK; c++) {
        const int pitched = c < d->n_pitched;
        const int rows = pitched ? MST_N_PITCHED_ROWS : MST_N_UNPITCHED_ROWS;
        const int feats = pitched ? MST_N_PITCHED_FEATS : MST_N_UNPITCHED_FEATS;
        float *g = pitched ? grids + d->pitched_offset + (uint64_t)c * cells * rows * feats
                           : grids + d->unpitched_offset;
        const int channel = pitched ? d->channel_id[c] : d->unpitched_channel_id;
        for (uint64_t cell = 0; cell < cells && status == MST_DECODE_OK; cell++) {
            const int frac = (int)(cell % MST_N_FRACTIONS);
            const int beat = (int)((cell / MST_N_FRACTIONS) % nb);
            const int64_t bar = (int64_t)(cell / MST_N_FRACTIONS / nb);
            for (int row = 0; row < rows; row++) {
                float *v = g + (cell * rows + row) * feats;
                float vel = v[1];
                float gated = vel * ((vel > thr_v) ? 1.0f : 0.0f);       /* velocity *= (velocity > .01).float() */
                if (d->write_back) v[1] = gated;
                if (gated == 0.0f) continue;                            /* np.nonzero (:526); NaN counts as non-zero */
                int note;
                if (pitched) {
                    float a0 = v[2], a1 = v[3], a2 = v[4];
                    float mx = a0; if (a1 > mx) mx = a1; if (a2 > mx) mx = a2;
                    if (a0 != a0 || a1 != a1 || a2 != a2) mx = NAN;      /* torch.max propagates NaN */
                    int f0 = (a0 == mx) && (a0 > thr_a), f1 = (a1 == mx) && (a1 > thr_a), f2 = (a2 == mx) && (a2 > thr_a);
                    int acc = f0 ? 0 : (f1 ? 1 : (f2 ? 2 : 1));          /* flat > natural > sharp, else none (:532-539) */
                    int degree = row % 7 + 1, octave = row / 7;
                    int interval = k_abs_intervals[d->mode][degree - 1] + d->tonic;   /* :270-271 */
                    if (acc == 2) interval += 1; else if (acc == 0) interval -= 1;
                    if (interval < 0) { octave -= 1; interval += 12; }
                    else if (interval >= 12) { octave += 1; interval -= 12; }
                    note = 12 * (octave + 1) + interval;                 /* note2note_id (:320-324) */
                } else {
                    note = row + 35;                                     /* :549 */
                }
                float dur = v[0];
                float prod = dur * (float)tpb;                           /* np.float32 * int -> f32 (:558) */
                if (!(prod == prod) || prod > 1.0e9f || prod < -1.0e9f) { status = MST_DECODE_BAD_VALUE; break; }
                int32_t qdur = (int32_t)prod;                            /* int(): truncation */
                int64_t t = bar * tpbar + (int64_t)beat * tpb + ((int64_t)k_frac_num[frac] * tpb) / k_frac_den[frac];
                if (t > (1 << 30)) { status = MST_DECODE_BAD_VALUE; break; }
                float vf = gated * 127.0f;                               /* int(msg.velocity * max_velocity) (midi.py:147) */
                if (!(vf == vf)) { status = MST_DECODE_BAD_VALUE; break; }
                int vi = (int)vf;
                if (vi > 127) { status = MST_DECODE_VELOCITY; break; }
                if (n + 2 > cap_msgs) { status = MST_DECODE_BAD_VALUE; break; }
                m[n].time = (int32_t)t; m[n].seq = n; m[n].is_on = 1; m[n].channel = (uint8_t)channel;
                m[n].note = (uint8_t)note; m[n].velocity = (uint8_t)vi; n++;
                m[n].time = (int32_t)(t + qdur); m[n].seq = n; m[n].is_on = 0; m[n].channel = (uint8_t)channel;
                m[n].note = (uint8_t)note; m[n].velocity = 0; n++;
            }
        }
    }
    if (status == MST_DECODE_OK && n == 0) status = MST_DECODE_NO_MESSAGES;
    res->status = status;
    res->n_messages = n;
    if (status != MST_DECODE_OK) { free(m); return 0; }
    /* per-channel stable sort by time (:487) followed by the global stable sort (midi.py:157) == one
     * stable sort of the channel-major emission order */
    qsort(m, n, sizeof(DecMsg), cmp_decmsg);
    if (n > capacity) { res->status = MST_DECODE_CAPACITY; free(m); return 0; }
    int32_t prev = 0;
    for (uint32_t i = 0; i < n; i++) {
        int32_t delta = m[i].time - prev; if (delta > res->max_delta) delta = res->max_delta;
        prev = m[i].time;
        if (delta < 0) delta = 0;
        out[i].time = m[i].time; out[i].delta = delta; out[i].is_on = m[i].is_on; out[i].channel = m[i].channel;
        out[i].note = m[i].note; out[i].velocity = m[i].velocity;
    }
    {
        int32_t last = m[n - 1].time;
        int32_t end = d->duration >= 0 ? d->duration : last + tpbar;    /* midi.py:158 */
        int32_t delta = end - last; if (delta > res->max_delta) delta = res->max_delta;
        if (delta < 0) delta = 0;
        res->eot_delta = delta;
    }
    free(m);
    return 0;
}

size_t msto_sizeof_decode_song(void) { return sizeof(MstDecodeSong); }
size_t msto_sizeof_out_msg(void) { return sizeof(MstOutMsg); }
size_t msto_sizeof_decode_result(void) { return sizeof(MstDecodeResult); }
