#!/bin/bash
mkdir -p gpurun_out
timeout 240 python -m pytest tests/test_gpu_smf.py -m gpu -x -q > gpurun_out/n_smf.log 2>&1
rc=$?; echo "smf rc=$rc"; tail -5 gpurun_out/n_smf.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 600 python bench.py --no-train --no-long > gpurun_out/n_bench.json 2> gpurun_out/n_bench.err
echo "bench rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/n_bench.json').read().strip().splitlines()[-1])
e=d['e2e']; print('device', d['value'], 'smf', e['value'], e['ms_per_step_per_rank'], 'events8', e['packed_events']['value'], e['packed_events']['ms_per_step_per_rank'])
f=d['files']; print('files smf', f['value'], f['files_per_s'], 'host', f['host_parse_path']['value'])
PY
tail -3 gpurun_out/n_bench.err
