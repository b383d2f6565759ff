"""Summarise an .ncu-rep (one line block per kernel): duration, occupancy, issue rate, DRAM traffic, pipe utilisation,
top stall reasons.  Usage: python tools/ncu_summary.py report.ncu-rep > profiles/summary.txt"""
import csv
import subprocess
import sys

KEYS = [
    ('gpu__time_duration.sum', 'duration'),
    ('launch__registers_per_thread', 'registers/thread'),
    ('launch__occupancy_limit_registers', 'occupancy limit (registers, blocks)'),
    ('launch__occupancy_limit_shared_mem', 'occupancy limit (shared memory, blocks)'),
    ('sm__warps_active.avg.pct_of_peak_sustained_active', 'achieved occupancy %'),
    ('smsp__issue_active.avg.pct', 'issue slots busy %'),
    ('smsp__inst_executed.sum', 'warp instructions'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'SM throughput % of peak'),
    ('gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed', 'memory throughput % of peak'),
    ('dram__bytes_read.sum', 'DRAM read'),
    ('dram__bytes_write.sum', 'DRAM write'),
    ('dram__throughput.avg.pct_of_peak_sustained_elapsed', 'DRAM throughput % of peak'),
    ('lts__t_bytes.sum', 'L2 traffic'),
    ('sm__pipe_tensor_op_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'tensor pipe (HMMA) active %'),
    ('sm__inst_executed_pipe_tensor.sum', 'tensor pipe instructions'),
    ('sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'FMA pipe active %'),
    ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'shared-memory bank conflicts'),
]


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    stall = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')]
        print(f'== {name}   grid {r[hdr.index("Grid Size")]} block {r[hdr.index("Block Size")]}')
        for k, label in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f'   {label:42s} {r[i]} {units[i]}')
        st = sorted(((float(r[hdr.index(h)] or 0), h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')])
                     for h in stall), reverse=True)[:5]
        print('   top stalls (warps per issue):              ' + ', '.join(f'{n} {v:.2f}' for v, n in st))


if __name__ == '__main__':
    main(sys.argv[1])
