"""Times mst_model_train_step at SURVEY.md config 4 shapes (development probe, not the bench)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import torch
import model_oracle as mo
from mst_native import trainer as tr


def main():
    Ns = [int(x) for x in (sys.argv[1].split(',') if len(sys.argv) > 1 else ['1', '4', '16'])]
    C, B, nb = (int(x) for x in (sys.argv[2].split(',') if len(sys.argv) > 2 else ['4', '200', '4']))
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    sd = mo.make_state_dict(108)
    t = tr.Trainer(sd)
    for N in Ns:
        samples = [mo.synthetic_sample(seed=7 + i, C=C, B=B, nb=nb, drums=True) for i in range(min(N, 4))]
        cat = lambda k: torch.cat([samples[i % len(samples)][k] for i in range(N)]).cuda().contiguous()
        p, u, f, m, b = cat('pitched'), cat('unpitched'), cat('feats'), cat('mode'), cat('bpm')
        it = torch.cat([mo.used_instruments(samples[i % len(samples)]['feats'], samples[0]['unpitched']) for i in range(N)]).cuda()
        bt = torch.full((N,), 120., device='cuda')
        for _ in range(2):
            t.accumulate(p, u, f, m, b, it, bt); t.optimizer_step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            losses = t.accumulate(p, u, f, m, b, it, bt); t.optimizer_step()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        ws = t.workspace(N, C, B, nb, True)
        print(f'N={N} C={C} B={B} nb={nb}: {ms:.2f} ms/step  {N / ms * 1e3:.1f} samples/s  launches {t.launches}  '
              f'workspace {ws.bytes / 2**30:.2f} GiB  loss {float(losses[0, 0]):.4f}', flush=True)
        t._ws.clear()


if __name__ == '__main__':
    main()
