"""Per-kernel timeline (start, duration, stream) of one N = 1 training iteration under the CUDA graph, via torch.profiler
(development tool: shows what the critical path of the branch schedule is)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
import torch
from torch.profiler import ProfilerActivity, profile
from mst_native import synth, trainer as tr

graph = (sys.argv[1] if len(sys.argv) > 1 else '1') == '1'
sd = synth.make_state_dict(108)
s = synth.synthetic_sample(seed=7, C=4, B=200, nb=4, drums=True)
dev = {k: v.cuda().contiguous() for k, v in s.items() if torch.is_tensor(v)}
bt = torch.full((1,), 120., device='cuda')
t = tr.Trainer(sd, graph=graph, precision='fp32')
losses = torch.empty(1, 16, device='cuda')


def step():
    t.accumulate(dev['pitched'], dev['unpitched'], dev['feats'], dev['mode'], dev['bpm'], dev['used'], bt, losses)
    t.optimizer_step()


for _ in range(5):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        step()
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
# last iteration: kernels after the last k_adam but one
names = [e.name for e in ev]
adam = [i for i, n in enumerate(names) if 'k_adam' in n]
lo = adam[-2] + 1 if len(adam) >= 2 else 0
sel = ev[lo:adam[-1] + 1]
t0 = sel[0].time_range.start
streams = {}
print(f'{len(sel)} kernels, span {(sel[-1].time_range.end - t0):.1f} us')
for e in sel:
    st = getattr(e, 'stream', None)
    if st is None:
        st = -1
    sid = streams.setdefault(st, len(streams))
    name = e.name.split('(')[0].replace('void ', '').replace('(anonymous namespace)::', '').replace('mstm::', '')
    print(f'{e.time_range.start - t0:9.1f} {e.time_range.end - e.time_range.start:8.1f}  s{sid}  {name[:60]}')
