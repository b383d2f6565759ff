// mst_lib.cu — library-level entry points of libmst_b200.so and the host-buffer ("_host") variants
// that wrap the device pipelines with the H2D / D2H copies (the e2e path of bench.py).
#include <stdarg.h>
#include <string.h>

#include <algorithm>

#include "mst_device.cuh"

static thread_local char g_error[512] = "";

void mst_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

extern "C" int mst_abi_version(void) { return MST_ABI_VERSION; }
extern "C" const char *mst_last_error(void) { return g_error; }

extern "C" int mst_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" size_t mst_encode_host_workspace_bytes(uint64_t n_events_total, uint32_t n_songs, uint64_t arena_floats) {
    size_t inner = mst_encode_workspace_bytes(n_events_total, n_songs, arena_floats);
    return mst_align_up(inner, 256) + mst_align_up((n_events_total ? n_events_total : 1) * sizeof(MstEvent), 256) +
           mst_align_up((size_t)(n_songs ? n_songs : 1) * sizeof(MstSongIn), 256) +
           mst_align_up((size_t)(n_songs ? n_songs : 1) * sizeof(MstSongInfo), 256);
}

extern "C" int mst_encode_batch_host_async(const MstEvent *h_events, uint64_t n_events_total,
                                           const MstSongIn *h_songs, uint32_t n_songs,
                                           const MstFilter *h_filter, MstSongInfo *h_info,
                                           float *d_arena, uint64_t arena_floats,
                                           void *d_workspace, size_t workspace_bytes, void *stream_) {
    if (mst_device_count() == 0) { mst_set_error("no CUDA device: libmst_b200 has no CPU fallback"); return MST_E_NO_DEVICE; }
    if (!h_songs || !h_info || !d_workspace || (!h_events && n_events_total)) {
        mst_set_error("mst_encode_batch_host: null pointer argument");
        return MST_E_INVALID;
    }
    if (n_songs == 0) return MST_OK;
    const size_t need = mst_encode_host_workspace_bytes(n_events_total, n_songs, arena_floats);
    if (need > workspace_bytes) {
        mst_set_error("mst_encode_batch_host: workspace too small (%zu < %zu)", workspace_bytes, need);
        return MST_E_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    const size_t inner = mst_align_up(mst_encode_workspace_bytes(n_events_total, n_songs, arena_floats), 256);
    char *p = (char *)d_workspace + inner;
    MstEvent *d_events = (MstEvent *)p; p += mst_align_up((n_events_total ? n_events_total : 1) * sizeof(MstEvent), 256);
    MstSongIn *d_songs = (MstSongIn *)p; p += mst_align_up((size_t)n_songs * sizeof(MstSongIn), 256);
    MstSongInfo *d_info = (MstSongInfo *)p;
    if (n_events_total)
        MST_CUDA_CHECK(cudaMemcpyAsync(d_events, h_events, n_events_total * sizeof(MstEvent), cudaMemcpyHostToDevice, stream));
    MST_CUDA_CHECK(cudaMemcpyAsync(d_songs, h_songs, (size_t)n_songs * sizeof(MstSongIn), cudaMemcpyHostToDevice, stream));
    int rc = mst_encode_batch_device(d_events, n_events_total, d_songs, n_songs, h_filter, d_info, d_arena, arena_floats,
                                     d_workspace, inner, stream_);
    if (rc != MST_OK) return rc;
    MST_CUDA_CHECK(cudaMemcpyAsync(h_info, d_info, (size_t)n_songs * sizeof(MstSongInfo), cudaMemcpyDeviceToHost, stream));
    return MST_OK;
}

extern "C" int mst_encode_batch_host(const MstEvent *h_events, uint64_t n_events_total,
                                     const MstSongIn *h_songs, uint32_t n_songs,
                                     const MstFilter *h_filter, MstSongInfo *h_info,
                                     float *d_arena, uint64_t arena_floats, float *h_arena,
                                     void *d_workspace, size_t workspace_bytes, void *stream_) {
    int rc = mst_encode_batch_host_async(h_events, n_events_total, h_songs, n_songs, h_filter, h_info, d_arena, arena_floats,
                                         d_workspace, workspace_bytes, stream_);
    if (rc != MST_OK || n_songs == 0) return rc;
    cudaStream_t stream = (cudaStream_t)stream_;
    MST_CUDA_CHECK(cudaStreamSynchronize(stream));
    if (h_arena) {
        // grids are contiguous from offset 0 in song order: copy the used prefix
        uint64_t used = 0;
        for (uint32_t s = 0; s < n_songs; s++) {
            if (h_info[s].status != MST_SONG_OK) continue;
            const uint64_t end = h_info[s].unpitched_offset +
                (uint64_t)h_info[s].n_unpitched * h_info[s].n_bars_out * h_info[s].n_beats * MST_UNPITCHED_BEAT_FLOATS;
            if (end > used) used = end;
        }
        if (used) {
            MST_CUDA_CHECK(cudaMemcpyAsync(h_arena, d_arena, used * sizeof(float), cudaMemcpyDeviceToHost, stream));
            MST_CUDA_CHECK(cudaStreamSynchronize(stream));
        }
    }
    return MST_OK;
}


// ------------------------------------------------------------------------------------------------
// 8-byte wire format of an event for the host path.  The H2D copy of 16-byte MstEvent records (274 MB per 512-song batch)
// is what bounds the host-buffer path once several ranks share one host socket (profiles/r1: e2e scaling efficiency 0.57
// at 8 GPUs); every field of an event fits 64 bits:
//   bits  0..31 tick | 32..35 kind (channel kinds 0..6 as they are, meta kinds 16..21 as 8..13) | 36..39 channel |
//   40..47 d1 | 48..55 d2 | 56..63 track (saturated at 255; informative only)
// and the 24-bit aux of set_tempo travels in the (d1, d2, track) bytes, the key-signature code in d1 - those kinds use
// neither.  The records are widened on the device (k_events_unpack: 8 B read + 16 B written per event, < 1 % of the step).
// ------------------------------------------------------------------------------------------------
static inline uint64_t mst_pack_one(const MstEvent &e) {
    const uint32_t kind = e.kind < 16 ? e.kind : (uint32_t)(e.kind - 8);
    uint32_t d1 = e.d1, d2 = e.d2, tr = e.track > 255 ? 255u : e.track;
    if (e.kind == MST_EV_SET_TEMPO) { d1 = e.aux & 0xFF; d2 = (e.aux >> 8) & 0xFF; tr = (e.aux >> 16) & 0xFF; }
    else if (e.kind == MST_EV_KEY_SIGNATURE) d1 = e.aux & 0xFF;
    return (uint64_t)e.tick | ((uint64_t)(kind & 15) << 32) | ((uint64_t)(e.channel & 15) << 36) | ((uint64_t)d1 << 40) |
           ((uint64_t)d2 << 48) | ((uint64_t)tr << 56);
}

extern "C" void mst_events_pack(const MstEvent *events, uint64_t n, MstPackedEvent *out) {
    for (uint64_t i = 0; i < n; i++) out[i] = mst_pack_one(events[i]);
}

__global__ void __launch_bounds__(256) k_events_unpack(const MstPackedEvent *__restrict__ in, MstEvent *__restrict__ out, uint64_t n) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t p = in[i];
        const uint32_t k4 = (uint32_t)(p >> 32) & 15u, kind = k4 < 8 ? k4 : k4 + 8;
        uint32_t d1 = (uint32_t)(p >> 40) & 0xFF, d2 = (uint32_t)(p >> 48) & 0xFF, tr = (uint32_t)(p >> 56) & 0xFF, aux = 0;
        if (kind == MST_EV_SET_TEMPO) { aux = d1 | (d2 << 8) | (tr << 16); d1 = d2 = tr = 0; }
        else if (kind == MST_EV_KEY_SIGNATURE) { aux = d1; d1 = 0; }
        uint4 w;                                                 // one 16-byte store per event (MstEvent layout)
        w.x = (uint32_t)p;                                       // tick
        w.y = aux;
        w.z = kind | (((uint32_t)(p >> 36) & 15u) << 8) | (d1 << 16) | (d2 << 24);
        w.w = tr;                                                // track | reserved << 16
        reinterpret_cast<uint4 *>(out)[i] = w;
    }
}

extern "C" size_t mst_encode_packed_workspace_bytes(uint64_t n_events_total, uint32_t n_songs, uint64_t arena_floats) {
    return mst_align_up(mst_encode_host_workspace_bytes(n_events_total, n_songs, arena_floats), 256) +
           mst_align_up((n_events_total ? n_events_total : 1) * sizeof(MstPackedEvent), 256);
}

extern "C" int mst_encode_batch_packed_async(const MstPackedEvent *h_events, uint64_t n_events_total, const MstSongIn *h_songs,
                                             uint32_t n_songs, const MstFilter *h_filter, MstSongInfo *h_info, float *d_arena,
                                             uint64_t arena_floats, void *d_workspace, size_t workspace_bytes, void *stream_) {
    if (mst_device_count() == 0) { mst_set_error("no CUDA device: libmst_b200 has no CPU fallback"); return MST_E_NO_DEVICE; }
    if (!h_songs || !h_info || !d_workspace || (!h_events && n_events_total)) {
        mst_set_error("mst_encode_batch_packed_async: null pointer argument");
        return MST_E_INVALID;
    }
    if (n_songs == 0) return MST_OK;
    const size_t need = mst_encode_packed_workspace_bytes(n_events_total, n_songs, arena_floats);
    if (need > workspace_bytes) {
        mst_set_error("mst_encode_batch_packed_async: workspace too small (%zu < %zu)", workspace_bytes, need);
        return MST_E_CAPACITY;
    }
    cudaStream_t stream = (cudaStream_t)stream_;
    // layout: [device pipeline workspace][MstEvent staging][songs][infos]  (as the 16-byte host path)  [packed staging]
    const size_t host_ws = mst_align_up(mst_encode_host_workspace_bytes(n_events_total, n_songs, arena_floats), 256);
    const size_t inner = mst_align_up(mst_encode_workspace_bytes(n_events_total, n_songs, arena_floats), 256);
    char *p = (char *)d_workspace + inner;
    MstEvent *d_events = (MstEvent *)p; p += mst_align_up((n_events_total ? n_events_total : 1) * sizeof(MstEvent), 256);
    MstSongIn *d_songs = (MstSongIn *)p; p += mst_align_up((size_t)n_songs * sizeof(MstSongIn), 256);
    MstSongInfo *d_info = (MstSongInfo *)p;
    MstPackedEvent *d_packed = (MstPackedEvent *)((char *)d_workspace + host_ws);
    if (n_events_total) {
        MST_CUDA_CHECK(cudaMemcpyAsync(d_packed, h_events, n_events_total * sizeof(MstPackedEvent), cudaMemcpyHostToDevice, stream));
        const unsigned blocks = (unsigned)std::min<uint64_t>((n_events_total + 255) / 256, 148 * 16);
        k_events_unpack<<<blocks, 256, 0, stream>>>(d_packed, d_events, n_events_total);
        MST_LAUNCH_CHECK("k_events_unpack");
    }
    MST_CUDA_CHECK(cudaMemcpyAsync(d_songs, h_songs, (size_t)n_songs * sizeof(MstSongIn), cudaMemcpyHostToDevice, stream));
    int rc = mst_encode_batch_device(d_events, n_events_total, d_songs, n_songs, h_filter, d_info, d_arena, arena_floats,
                                     d_workspace, inner, stream_);
    if (rc != MST_OK) return rc;
    MST_CUDA_CHECK(cudaMemcpyAsync(h_info, d_info, (size_t)n_songs * sizeof(MstSongInfo), cudaMemcpyDeviceToHost, stream));
    return MST_OK;
}
