"""Joins the GEMM shapes a training step logs (MST_GEMM_LOG=2, stderr) with the ncu launch list of the same run:
per GEMM launch the algorithmic bytes (operands read once + result written), the duration and the achieved GB/s.
Usage: python tools/gemm_report.py launches.csv gemm_log.txt [steps]"""
import csv
import re
import sys


def main(csv_path, log_path, steps=3):
    rows = list(csv.DictReader(l for l in open(csv_path) if l.startswith('"')))
    gem = [r for r in rows if 'k_gemm' in r['Kernel Name']]
    pat = re.compile(r'gemm M=(\d+) N=(\d+) K=(\d+) a=\((-?\d+),(-?\d+)\) b=\((-?\d+),(-?\d+)\) c=(-?\d+) act=(\d+) acc=(\d+) split=(\d+)')
    shapes = [tuple(int(x) for x in m.groups()) for m in map(pat.search, open(log_path)) if m]
    n = len(gem) // steps
    per = len(shapes) // steps
    assert n == per, (len(gem), len(shapes))
    gem, shapes = gem[-n:], shapes[-n:]
    out = []
    for r, s in zip(gem, shapes):
        M, N, K = s[:3]
        us = float(r['Metric Value']) / 1e3
        byts = 4.0 * (M * K + K * N + M * N * (2 if s[9] else 1))
        name = r['Kernel Name'].split('(')[0].replace('void ', '').replace('mstm::', '')
        out.append((us, name, M, N, K, s[10], byts / us / 1e3, 2.0 * M * N * K / us / 1e6))
    tot = sum(o[0] for o in out)
    print(f'{len(out)} GEMM launches per step, {tot / 1e3:.3f} ms')
    for us, name, M, N, K, split, gbs, tf in sorted(out, reverse=True)[:60]:
        print(f'{us:9.1f} us  {name:24s} M={M:8d} N={N:5d} K={K:8d} split={split:4d}  {gbs:7.0f} GB/s  {tf:6.1f} TFLOP/s')


if __name__ == '__main__':
    main(sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 3)
