"""CPU tests (gloo, world_size 2) of the data-parallel host logic in mst_native/trainer.py: sample sharding,
ragged-batch balancing, the StepLR schedule and the flat-gradient all-reduce.  The gradient itself comes from the
torch oracle here (the CUDA step needs a GPU): the property checked is the one the design rests on — the SUM over
ranks of per-shard gradients equals the gradient accumulated over the whole global batch (train-model.py's
iter_size semantics), through the same allreduce_sum_ the Trainer calls."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mst_native import trainer as tr


def test_shard_samples_partitions_the_batch():
    for n in (1, 2, 7, 8, 64, 65):
        for world in (1, 2, 3, 8):
            spans = [tr.shard_samples(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def test_balanced_assignment_is_deterministic_and_balanced():
    costs = [800, 100, 790, 400, 410, 50, 60, 300]
    a = tr.balanced_assignment(costs, 2)
    assert a == tr.balanced_assignment(costs, 2)
    assert sorted(i for part in a for i in part) == list(range(len(costs)))
    loads = [sum(costs[i] for i in part) for part in a]
    assert abs(loads[0] - loads[1]) <= 100
    assert tr.balanced_assignment([5, 5, 5], 4)[3] == []


def test_step_lr_matches_torch_scheduler():
    p = torch.zeros(1, requires_grad=True)
    opt = torch.optim.Adam([p], lr=0.01)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=200, gamma=0.9)      # train-model.py:89-90
    for done in range(0, 650):
        assert abs(opt.param_groups[0]['lr'] - tr.step_lr(0.01, done)) < 1e-12
        opt.step(); sched.step()


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    import model_oracle as mo
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    torch.set_num_threads(2)
    case = dict(C=2, B=3, nb=4, drums=True, density=0.02)
    samples = [mo.synthetic_sample(seed=40 + i, **case) for i in range(3)]      # global batch of 3 over 2 ranks
    lo, hi = tr.shard_samples(len(samples), world, rank)
    sd = {k: v.clone().requires_grad_(True) for k, v in mo.make_state_dict(108).items()}
    names = sorted(sd)
    for s in samples[lo:hi]:
        losses, _ = mo.training_losses(sd, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'], s['bpm_target'])
        losses['total'].sum().backward()
    flat = torch.cat([(sd[k].grad if sd[k].grad is not None else torch.zeros_like(sd[k])).reshape(-1) for k in names])
    tr.allreduce_sum_(flat)
    if rank == 0:
        torch.save(flat, out)
    dist.barrier()
    dist.destroy_process_group()


def test_allreduce_of_shard_gradients_equals_accumulated_gradient(tmp_path):
    import model_oracle as mo
    out = str(tmp_path / 'flat.pt')
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    case = dict(C=2, B=3, nb=4, drums=True, density=0.02)
    sd = {k: v.clone().requires_grad_(True) for k, v in mo.make_state_dict(108).items()}
    for i in range(3):
        s = mo.synthetic_sample(seed=40 + i, **case)
        losses, _ = mo.training_losses(sd, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'], s['bpm_target'])
        losses['total'].sum().backward()
    want = torch.cat([(sd[k].grad if sd[k].grad is not None else torch.zeros_like(sd[k])).reshape(-1) for k in sorted(sd)])
    assert got.numel() == 980325
    assert float((got - want).norm() / want.norm()) < 1e-5


def test_allreduce_is_a_noop_without_a_process_group():
    x = torch.arange(5.)
    assert torch.equal(tr.allreduce_sum_(x.clone()), x)


@pytest.mark.gpu
def test_trainer_matches_reference_optimizer_trajectory():
    """Two optimizer steps (iter_size = 2 songs each) of the CUDA trainer vs torch autograd + torch.optim.Adam on the
    oracle (train-model.py:97-154)."""
    import model_oracle as mo
    case = dict(C=2, B=6, nb=4, drums=True, density=0.01)
    samples = [mo.synthetic_sample(seed=70 + i, **case) for i in range(4)]
    sd0 = mo.make_state_dict(108)
    t = tr.Trainer(sd0)
    ref = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    opt = torch.optim.Adam(list(ref.values()), lr=0.01)
    for step in range(2):
        for s in samples[2 * step:2 * step + 2]:
            losses, _ = mo.training_losses(ref, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'], s['bpm_target'])
            losses['total'].sum().backward()
            it = mo.used_instruments(s['feats'], s['unpitched']).cuda()
            got = t.accumulate(s['pitched'].cuda(), s['unpitched'].cuda(), s['feats'].cuda(), s['mode'].cuda(), s['bpm'].cuda(),
                               it, torch.tensor([float(s['bpm_target'])], device='cuda'))
            assert abs(float(got[0, 0]) - float(losses['total'])) < 2e-5 * float(losses['total'])
        opt.step(); opt.zero_grad()
        t.optimizer_step()
    ours = t.state_dict(sd0)
    worst = max(float((ours[k].cpu() - ref[k].detach()).abs().max()) for k in ref)
    # Adam's first steps move every weight by ~lr regardless of gradient scale: a sign flip of a ~0 gradient is the
    # only way to differ by more than float noise
    assert worst < 2e-3, worst
    close = sum(int(((ours[k].cpu() - ref[k].detach()).abs() < 1e-5).sum()) for k in ref)
    assert close > 0.995 * 980325, close


def test_ragged_plan_covers_the_batch_once_and_groups_by_shape():
    shapes = [(4, 200, 4, True), (3, 266, 4, True), (4, 200, 4, True), (8, 100, 4, False), (4, 200, 4, True), (2, 150, 4, True)]
    for world in (1, 2, 3):
        plans = [tr.ragged_plan(shapes, world, r) for r in range(world)]
        seen = sorted(i for p in plans for idx in p.values() for i in idx)
        assert seen == list(range(len(shapes)))                                   # every sample trained exactly once
        for p in plans:
            for shape, idx in p.items():
                assert all(tuple(shapes[i]) == shape for i in idx)
        loads = [sum(shapes[i][0] * shapes[i][1] for idx in p.values() for i in idx) for p in plans]
        assert max(loads) - min(loads) <= 800


@pytest.mark.gpu
def test_accumulate_ragged_equals_per_sample_accumulation():
    """A ragged list of songs (three shapes) through accumulate_ragged == the same songs one at a time (iter_size semantics),
    and the two halves ragged_plan gives to two ranks sum to the same gradient."""
    from mst_native import synth
    cases = [dict(C=2, B=5, nb=4, drums=True), dict(C=3, B=4, nb=4, drums=False), dict(C=2, B=5, nb=4, drums=True), dict(C=1, B=7, nb=3, drums=True)]
    samples = [synth.synthetic_sample(seed=90 + i, density=0.02, **c) for i, c in enumerate(cases)]
    sd = synth.make_state_dict(108)
    one = tr.Trainer(sd)
    for s in samples:
        cu = lambda t: t.cuda().contiguous()
        one.accumulate(cu(s['pitched']), None if s['unpitched'] is None else cu(s['unpitched']), cu(s['feats']), cu(s['mode']), cu(s['bpm']),
                       cu(s['used']), torch.tensor([120.], device='cuda'))
    rag = tr.Trainer(sd)
    leaves = rag.accumulate_ragged(samples, world=1, rank=0)
    assert sorted(leaves) == [0, 1, 2, 3]
    assert float((rag.grads - one.grads).norm() / one.grads.norm()) < 1e-5
    halves = torch.zeros_like(one.grads)
    for r in range(2):
        t = tr.Trainer(sd)
        t.accumulate_ragged(samples, world=2, rank=r)
        halves += t.grads
    assert float((halves - one.grads).norm() / one.grads.norm()) < 1e-5
