"""Times the training iteration through the C-ABI Trainer at N = 1 (graph replay and eager) and at larger N (development probe,
not the bench).   python tools/n1_probe.py [Ns=1,16] [steps=20] [precision=fp32]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
import torch
from mst_native import synth, trainer as tr


def main():
    Ns = [int(x) for x in (sys.argv[1].split(',') if len(sys.argv) > 1 else ['1', '16'])]
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    precision = sys.argv[3] if len(sys.argv) > 3 else 'fp32'
    sd = synth.make_state_dict(108)
    samples = [synth.synthetic_sample(seed=7 + i, C=4, B=200, nb=4, drums=True) for i in range(2)]
    for N in Ns:
        cat = lambda k: torch.cat([samples[i % 2][k] for i in range(N)]).cuda().contiguous()
        p, u, f, m, b, it = cat('pitched'), cat('unpitched'), cat('feats'), cat('mode'), cat('bpm'), cat('used')
        bt = torch.full((N,), 120., device='cuda')
        modes = (True, False) if N <= 4 else (False,)
        if os.environ.get('MST_PROBE_GRAPH'):
            modes = (os.environ['MST_PROBE_GRAPH'] == '1',)
        for graph in modes:
            t = tr.Trainer(sd, graph=graph, precision=precision)
            losses = torch.empty(N, 16, device='cuda')
            for _ in range(int(os.environ.get('MST_PROBE_WARMUP', '4'))):
                t.accumulate(p, u, f, m, b, it, bt, losses); t.optimizer_step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                t.accumulate(p, u, f, m, b, it, bt, losses); t.optimizer_step()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / steps
            print(f'N={N} graph={graph} {precision} streams={os.environ.get("MST_MODEL_STREAMS", "on")}: {ms:.3f} ms/step  '
                  f'{N / ms * 1e3:.1f} samples/s  launches {t.launches}  loss {float(losses[0, 0]):.6f}', flush=True)
            del t
            torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
