"""Builds libmst_b200.so in-tree with nvcc for sm_100a (B200).

    python music-style-transfer_b200/build.py [--force] [--verbose]

No torch dependency: the library is plain CUDA runtime + C ABI (include/mst_b200.h).  The codec
translation units are compiled with -fmad=false: their f64 arithmetic must match CPython / numpy
operation by operation (no FMA contraction).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(HERE, 'libmst_b200.so')
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')

ARCH = ['-gencode', 'arch=compute_100a,code=sm_100a']
COMMON = ['-O3', '-lineinfo', '-std=c++17', '-Xcompiler', '-fPIC', '-Xcompiler', '-Wall',
          '-Xcompiler', '-fno-strict-aliasing']

# (source, extra flags)
UNITS = [
    ('mst_lib.cu', []),
    ('mst_encode.cu', ['-fmad=false']),
    ('mst_decode.cu', ['-fmad=false']),
    ('mst_model.cu', []),
    ('mst_transfer.cu', []),
    ('mst_smf.cpp', []),
    ('mst_smf_gpu.cu', []),
]


def _sources():
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(HERE, '..', 'include', 'mst_b200.h'))
    return deps


def up_to_date():
    if not os.path.exists(OUT):
        return False
    t = os.path.getmtime(OUT)
    return all(os.path.getmtime(d) <= t for d in _sources())


def build(force=False, verbose=False):
    if not force and up_to_date():
        return OUT
    objs = []
    build_dir = os.path.join(HERE, 'build')
    os.makedirs(build_dir, exist_ok=True)
    procs = []
    for src, extra in UNITS:
        obj = os.path.join(build_dir, os.path.splitext(src)[0] + '.o')
        cmd = [NVCC, *ARCH, *COMMON, *extra, '-c', os.path.join(CSRC, src), '-o', obj]
        if verbose:
            cmd.insert(1, '-Xptxas=-v')
            print(' '.join(cmd))
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out = p.communicate()[0].decode()
        if p.returncode != 0 or verbose:
            sys.stdout.write(out)
        if p.returncode != 0:
            failed = True
            print(f'nvcc failed on {src}')
    if failed:
        raise RuntimeError('build of libmst_b200.so failed')
    cmd = [NVCC, *ARCH, '-shared', '-o', OUT, *objs, '-lcudart']
    subprocess.check_call(cmd)
    return OUT


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, verbose='--verbose' in sys.argv))
