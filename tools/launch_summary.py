"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, mean, share."""
import collections
import csv
import sys


def main(path, skip=0):
    rows = list(csv.DictReader(l for l in open(path) if l.startswith('"')))
    rows = rows[skip:]
    agg = collections.OrderedDict()
    for r in rows:
        k = r['Kernel Name'].split('(')[0].replace('<unnamed>::', '').replace('mstm::', '').replace('void ', '')
        agg.setdefault(k, []).append(float(r['Metric Value']) / 1e3)
    tot = sum(sum(v) for v in agg.values())
    print(f'{len(rows)} launches, {tot / 1e3:.3f} ms total')
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f'{k:28s} n={len(v):4d} mean={sum(v) / len(v):10.1f} us  max={max(v):10.1f} us  share={sum(v) / tot:.3f}')


if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0)
