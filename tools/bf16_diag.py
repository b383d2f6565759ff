"""Per-parameter gradient error of the bf16 dense mode against the fp32 torch oracle, with the mode's two halves
(tcgen05 GEMMs / tcgen05 PitchedStyleApplier tail) switched on separately (development probe).

    python tools/bf16_diag.py C B nb density     (runs itself three times with the MST_BF16_NO_* knobs)"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ('music-style-transfer_b200', 'oracle', 'tests'):
    sys.path.insert(0, os.path.join(ROOT, p))


def child(C, B, nb, density):
    import torch
    import model_oracle as mo
    from mst_native import model as mm
    sd = mo.make_state_dict(108)
    flat = mm.flatten_state_dict(sd)
    samples = [mo.synthetic_sample(seed=5 + i, C=C, B=B, nb=nb, drums=True, density=density) for i in range(2)]
    cat = lambda k: torch.cat([s[k] for s in samples]).cuda().contiguous()
    it = torch.cat([mo.used_instruments(s['feats'], s['unpitched']) for s in samples]).cuda()
    bt = torch.tensor([120., 120.], device='cuda')
    ws = mm.Workspace(2, C, B, nb, True, training=True, precision='bf16')
    grads = torch.zeros_like(flat)
    losses, _ = ws.train_step(flat, grads, cat('pitched'), cat('unpitched'), cat('feats'), cat('mode'), cat('bpm'), it, bt)
    torch.cuda.synchronize()
    sdg = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    tot = 0.0
    for s in samples:
        l, _ = mo.training_losses(sdg, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'], s['bpm_target'])
        l['total'].sum().backward()
        tot = float(l['total'])
    rows = []
    for name, off, numel in mm.param_table():
        g = grads[off:off + numel].cpu()
        r = sdg[name].grad.reshape(-1) if sdg[name].grad is not None else torch.zeros(numel)
        rows.append((float((g - r).norm() / (r.norm() + 1e-12)), name, float(r.norm())))
    rows.sort(reverse=True)
    print(f'  loss {float(losses[1, 0]):.6f} vs oracle {tot:.6f}')
    for e, n, rn in rows[:12]:
        print(f'  {e:9.3e}  {n}  (|g| {rn:.2e})')


if __name__ == '__main__':
    if os.environ.get('MST_DIAG_CHILD'):
        child(*(int(x) for x in sys.argv[1:4]), float(sys.argv[4]))
    else:
        args = sys.argv[1:5] if len(sys.argv) >= 5 else ['4', '12', '4', '0.003']
        variants = [('bf16 GEMMs + tcgen05 tail', {}), ('bf16 GEMMs only', {'MST_BF16_NO_TAILS': '1'}),
                    ('tcgen05 tail only', {'MST_BF16_NO_GEMM': '1'})]
        if os.environ.get('MST_DIAG_MASKS'):
            names = {1: 'linear forward', 2: 'input gradient', 4: 'weight gradient', 8: 'LSTM projection', 16: 'LSTM backward products'}
            variants = []
            for m in (int(x) for x in os.environ['MST_DIAG_MASKS'].split(',')):
                variants.append(('bf16 GEMM classes: ' + ' + '.join(v for k, v in names.items() if m & k), {'MST_BF16_NO_TAILS': '1', 'MST_BF16_MASK': str(m)}))
        for label, env in variants:
            print(f'== {label}  (C, B, nb, density) = {args}', flush=True)
            subprocess.run([sys.executable, os.path.abspath(__file__)] + args, env=dict(os.environ, MST_DIAG_CHILD='1', **env))
