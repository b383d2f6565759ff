"""CPU check of the device-side SMF parser's LOGIC (csrc/mst_smf_gpu.cu): the per-track parser (`parse_track`, with its
shared-memory window fast path) is plain C++ apart from a handful of CUDA spellings, so this test compiles that part of the
.cu file for the host behind a few shims (`__device__`, `__funnelshift_r`, `__ffs`, `uint4`), mirrors k_smf_tracks' chunk
walk sequentially, and compares records and load verdicts with the host parser mst_smf_parse on the files of
tests/test_gpu_smf.py plus several thousand random mutations.  The GPU tests check the real kernels; this one catches a
logic slip on the build machine, where no GPU exists."""
import ctypes
import os
import shutil
import subprocess

import numpy as np
import pytest

import test_gpu_smf as cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, 'music-style-transfer_b200', 'csrc', 'mst_smf_gpu.cu')

SHIMS = r'''
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include <algorithm>
#include <vector>
#include "%(header)s"
#define __device__
#define __forceinline__ inline
#define __ffs __builtin_ffs
struct uint4 { uint32_t x, y, z, w; };
static inline uint32_t __funnelshift_r(uint32_t lo, uint32_t hi, uint32_t sh) { sh &= 31; return sh ? (lo >> sh) | (hi << (32 - sh)) : lo; }
static uint32_t g_buf[80 * 32];
'''

DRIVER = r'''
// sequential mirror of k_smf_tracks (header + chunk walk) followed by parse_track per track
extern "C" int mirror_parse(const uint8_t *p, uint32_t n, MstEvent *out, int32_t *division) {
    if (n < 8) return -1;
    if (memcmp(p, "MThd", 4)) return -2;
    const uint32_t hsize = ((uint32_t)p[4] << 24) | ((uint32_t)p[5] << 16) | ((uint32_t)p[6] << 8) | p[7];
    const uint32_t avail = n - 8 < hsize ? n - 8 : hsize;
    if (avail < 6) return -1;
    const int nt = (int16_t)((p[10] << 8) | p[11]);
    *division = (int16_t)((p[12] << 8) | p[13]);
    const uint32_t first = 8 + avail;
    const uint32_t ntr = nt > 0 ? std::min((uint32_t)nt, (n - first) / 8u + 1u) : 0u;
    std::vector<TrackRec> tracks;
    int fail = 0;
    uint64_t pos = first;
    for (uint32_t t = 0; t < ntr; t++) {
        TrackRec r{0, t, 0u, 0u};
        if (!fail) {
            if (pos + 8 > n) fail = 1;
            else if (memcmp(p + pos, "MTrk", 4)) fail = 3;
            else {
                r.size = ((uint32_t)p[pos + 4] << 24) | ((uint32_t)p[pos + 5] << 16) | ((uint32_t)p[pos + 6] << 8) | p[pos + 7];
                r.start = (uint32_t)(pos + 8);
                pos += 8ull + r.size;
            }
        }
        tracks.push_back(r);
    }
    if (!fail && ntr > 0 && (uint32_t)nt > ntr) fail = 1;
    if (fail) return -fail;
    uint32_t total = 0;
    for (auto &tr : tracks) {
        uint32_t c = 0;
        const int rc = parse_track<1>(p, n, tr, out + total, &c, g_buf + 7);
        if (rc) return -100 - rc;
        total += c;
    }
    return (int)total;
}
'''


@pytest.fixture(scope='module')
def mirror(tmp_path_factory):
    if shutil.which('g++') is None:
        pytest.skip('g++ not available')
    src = open(SRC).read()
    try:
        track_rec = src[src.index('struct TrackRec {'):src.index('struct SmfWorkspace {')]
        parser = src[src.index('struct Rd {'):src.index('// A record takes at least two bytes of its track')]
    except ValueError:
        pytest.fail('mst_smf_gpu.cu no longer has the markers this test cuts the parser out by')
    parser = parser.replace('reinterpret_cast<uint4 *>(out)[count] = w;', 'memcpy(&out[count], &w, 16);')
    d = tmp_path_factory.mktemp('smf_mirror')
    cpp = d / 'mirror.cpp'
    cpp.write_text(SHIMS % {'header': os.path.join(ROOT, 'include', 'mst_b200.h')} + track_rec + parser + DRIVER)
    so = d / 'mirror.so'
    r = subprocess.run(['g++', '-O1', '-shared', '-fPIC', '-o', str(so), str(cpp)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    lib = ctypes.CDLL(str(so))
    lib.mirror_parse.restype = ctypes.c_int
    return lib


def mirror_parse(mirror, blob, shift):
    from mst_native import abi
    # the file at an arbitrary alignment inside a padded buffer: the window loads read whole 16-byte lines around it
    buf = np.full(len(blob) + 640, 0xAB, np.uint8)
    buf[32 + shift:32 + shift + len(blob)] = np.frombuffer(blob, np.uint8)
    out = np.zeros(len(blob) // 2 + 2, dtype=abi.EVENT_DTYPE)
    div = ctypes.c_int32(0)
    n = mirror.mirror_parse(ctypes.c_void_p(buf.ctypes.data + 32 + shift), len(blob), out.ctypes.data_as(ctypes.c_void_p), ctypes.byref(div))
    return None if n < 0 else (out[:n].copy(), div.value)


def compare(mirror, blobs, tag):
    from mst_native import lib
    host = [lib.smf_parse(b) for b in blobs]
    dev = [mirror_parse(mirror, b, i % 16) for i, b in enumerate(blobs)]
    cases.assert_same_parse(host, dev, tag)
    return sum(h is not None for h in host)


def test_parser_logic_on_real_and_hostile_files(mirror):
    blobs = cases.bundled_blobs() + cases.golden_event_blobs()
    assert compare(mirror, blobs, 'real') >= 60
    hostile = cases.hostile_files(cases.bundled_blobs()[0])
    n_ok = compare(mirror, list(hostile.values()), 'hostile')
    assert 10 < n_ok < len(hostile) - 10


def test_parser_logic_on_random_mutations(mirror):
    rng = np.random.default_rng(99)
    sources = [b for b in cases.bundled_blobs() + cases.golden_event_blobs() if len(b) < 60000][:10]
    blobs = []
    for src in sources:
        a = np.frombuffer(src, dtype=np.uint8)
        for k in range(500):
            m = a.copy()
            kind = k % 3
            if kind == 0:
                for p in rng.integers(14, len(m), size=int(rng.integers(1, 6))):
                    m[p] = rng.integers(0, 256)
            elif kind == 1:
                pos = int(rng.integers(0, len(m)))
                m = np.delete(m, slice(pos, pos + int(rng.integers(1, 4))))
            else:
                pos = int(rng.integers(0, len(m)))
                m = np.insert(m, pos, rng.integers(0, 256, size=int(rng.integers(1, 4)), dtype=np.uint8))
            blobs.append(m.tobytes())
    n_ok = compare(mirror, blobs, 'mutations')
    assert 200 < n_ok < len(blobs) - 200, n_ok
