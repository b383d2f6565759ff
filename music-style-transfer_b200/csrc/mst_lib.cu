// mst_lib.cu — library-level entry points of libmst_b200.so and the host-buffer ("_host") variants
// that wrap the device pipelines with the H2D / D2H copies (the e2e path of bench.py).
#include <stdarg.h>
#include <string.h>

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
