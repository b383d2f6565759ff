"""Per-kernel table of one training step from an .ncu-rep captured over every launch of the step: launches, total time,
share, and the time-weighted DRAM throughput / SM throughput / FMA-pipe / tensor-pipe utilisation and achieved occupancy.
Usage: python tools/ncu_step_table.py report.ncu-rep > profiles/table.txt"""
import collections
import csv
import subprocess
import sys

COLS = [('gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed', 'mem%'),
        ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm%'),
        ('sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'fma%'),
        ('sm__inst_executed_pipe_tensor.sum', 'tensor-inst'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%')]


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return 0.0


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    it = hdr.index('gpu__time_duration.sum')
    scale = {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(units[it], 1e-3)
    rd, wr = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
    bscale = lambda i: {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(units[i], 1.0)
    agg = collections.OrderedDict()
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')].split('(')[0].replace('void ', '').replace('mstm::', '').replace('<unnamed>::', '')
        us = num(r[it]) * scale
        a = agg.setdefault(name, {'n': 0, 'us': 0.0, 'bytes': 0.0, **{c: 0.0 for _, c in COLS}})
        a['n'] += 1
        a['us'] += us
        a['bytes'] += num(r[rd]) * bscale(rd) + num(r[wr]) * bscale(wr)
        for k, c in COLS:
            if k in hdr:
                v = num(r[hdr.index(k)])
                a[c] += v if c == 'tensor-inst' else v * us
    tot = sum(a['us'] for a in agg.values())
    print(f'{sum(a["n"] for a in agg.values())} launches, {tot / 1e3:.3f} ms under ncu (cold caches, serialised)')
    print(f'{"kernel":30s} {"n":>4s} {"ms":>8s} {"share":>6s} {"DRAM GB/s":>10s} ' + ' '.join(f'{c:>7s}' for _, c in COLS if c != 'tensor-inst') + '  tensor-inst')
    for name, a in sorted(agg.items(), key=lambda kv: -kv[1]['us']):
        w = max(a['us'], 1e-9)
        gbs = a['bytes'] / (a['us'] * 1e-6) / 1e9 if a['us'] else 0.0
        print(f'{name:30s} {a["n"]:4d} {a["us"] / 1e3:8.3f} {a["us"] / tot:6.3f} {gbs:10.0f} ' +
              ' '.join(f'{a[c] / w:7.1f}' for _, c in COLS if c != 'tensor-inst') + f'  {a["tensor-inst"]:.0f}')


if __name__ == '__main__':
    main(sys.argv[1])
