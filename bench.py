#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native music-style-transfer hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[2]): bulk encode of a synthetic MIDI
corpus — 16 tracks (15 pitched + drums), 128 bars, tpb 480, ~33 k events per file.  One "step" =
one pass of the whole encode pipeline (libmst_b200.so, mst_encode_batch_device) over one batch of
`--batch` songs whose events are already resident in HBM; batches cycle through a pool of distinct
songs that is larger than L2, and every step rewrites a grid arena that is far larger than L2.

Prints ONE JSON line (see the task contract): value = song-bars/s over all ranks, `roofline` for
the dominant kernel (k_fill_flat, HBM-write-bound), `e2e` through the host-buffer C-ABI entry point
(mst_encode_batch_smf_async: the batch's raw .mid bytes from page-locked memory, parsed on the device;
`e2e.packed_events` = the same loop on 8-byte event records parsed on the host beforehand), and
`cpu_baseline` (the C oracle port on the host cores, rank 0, N=1 only).
Further keys on the same line: `decode`, `files` (files on disk through the product EventPipeline, both wire
formats), `training` (configs[3], bf16 dense mode) / `training_fp32` / `training_long` / `training_n1` (the
reference's own batch size, CUDA graph + drop-in call pattern), each with its own `e2e`, and LAST `summary`.
`--impl reference` times the CPU implementation on the same config: the C oracle port on all host threads for
the codec, the unmodified reference model from baseline/_ref for the training step.
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))

import numpy as np  # noqa: E402

METRIC = 'piano-roll bars/sec encoded'
UNIT = 'song-bars/s'
KERNELS_PER_STEP = 9   # k_sort, k_walk<0>, k_info, k_offsets, k_fill_flat, k_walk<1>, k_pair, k_scale, k_cells


def workload_config(args):
    return {
        'workload': 'configs[2]: synthetic MIDI corpus bulk encode (16 tracks = 15 pitched + drums, 128 bars, '
                    'tpb 480, 1/16 + triplet onsets, ~33k events/file)',
        'songs_per_step': args.batch,
        'distinct_songs_in_pool': args.pool,
        'corpus_files_equivalent': args.batch * args.steps,
        'l2_policy': 'inputs larger than L2: >=270 MB of events and a multi-GB grid arena per step',
        'streams': 'two alternating output rings (arena + stream); the library adds one high-priority side stream',
        'filter': 'iter_inputs filter: popular_instruments, min_n_messages=100',
    }


def measured_peak_hbm():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        with open(path) as f:
            return float(json.load(f)['hbm_gbs']), 'measured (MEASURED_PEAKS.json)'
    except Exception:
        return 6650.0, 'fallback (B200_PROFILING.md)'


class ClockSampler:
    """nvidia-smi clocks / throttle reasons while the GPU is under load."""
    FIELDS = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
              'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
              'clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.samples = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', f'--query-gpu={self.FIELDS}', '--format=csv,noheader,nounits', '-lms', '100',
                 '-i', str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        sm, smax, reasons = [], [], set()
        for t, line in self.samples:
            if not (t0 <= t <= t1 + 0.2):
                continue
            parts = [p.strip() for p in line.split(',')]
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'),
                                 parts[3:7]):
                if val == 'Active':
                    reasons.add(name)
        return {'sm_mhz': statistics.median(sm) if sm else None, 'sm_max_mhz': max(smax) if smax else None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def run_decode(args, world, rank, torch, dist, L, lib, abi, batch, filt, arena):
    """Decode leg (the inverse path of style_transfer.py: hard_output gate + compaction + ordering) over the grids the
    encoder just wrote: one step = mst_decode_batch_device over one batch of songs, grids resident in HBM."""
    batch.encode(filt, arena)
    infos = batch.infos()
    dsongs = []
    for info in infos:
        C = min(int(info['n_pitched']), 15)
        dsongs.append(abi.make_decode_song(info, C, int(info['n_unpitched']) > 0, int(info['n_bars_out']),
                                           instrument_ids=list(info['pitched_instruments'][:C]), duration=int(info['duration']),
                                           write_back=False, pitched_offset=int(info['pitched_offset']),
                                           unpitched_offset=int(info['unpitched_offset'])))
    dsongs = np.array(dsongs, dtype=abi.DECODE_SONG_DTYPE)
    S, n_chunks = len(dsongs), lib.decode_chunks(dsongs)
    grid_bytes = float(sum(4 * int(d['n_bars']) * int(d['n_beats']) * (int(d['n_pitched']) * abi.PITCHED_BEAT_FLOATS +
                                                                      int(d['has_unpitched']) * abi.UNPITCHED_BEAT_FLOATS)
                           for d in dsongs))
    res, msgs = lib.decode_batch(arena, dsongs)           # sizing pass (+ correctness: every song decodes)
    assert (res['status'] == 0).all(), 'decode leg: non-OK song'
    n_msgs = int(res['n_messages'].astype(np.int64).sum())
    cap = n_msgs + 1024
    d_songs = torch.from_numpy(dsongs.view(np.uint8).reshape(-1)).cuda()
    d_res = torch.empty(S * abi.DECODE_RESULT_DTYPE.itemsize, dtype=torch.uint8, device='cuda')
    d_msgs = torch.empty(cap * abi.OUT_MSG_DTYPE.itemsize, dtype=torch.uint8, device='cuda')
    wsb = L.mst_decode_workspace_bytes(S, n_chunks, cap)
    ws = torch.empty(wsb, dtype=torch.uint8, device='cuda')
    stream = torch.cuda.current_stream().cuda_stream
    h_res = torch.empty(S * abi.DECODE_RESULT_DTYPE.itemsize, dtype=torch.uint8).pin_memory()
    h_msgs = torch.empty(cap * abi.OUT_MSG_DTYPE.itemsize, dtype=torch.uint8).pin_memory()

    def step():
        lib.check(L.mst_decode_batch_device(arena.data_ptr(), d_songs.data_ptr(), S, n_chunks, d_msgs.data_ptr(), cap,
                                            d_res.data_ptr(), ws.data_ptr(), wsb, stream))

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device='cuda')
    tot = torch.tensor([grid_bytes, float(n_msgs)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    s_per_step = float(ms.item()) / 1e3 / args.steps
    # e2e: results + messages read back to pinned host memory every step (the grids come from the model on device).  Two
    # slots (stream + device / host result buffers + workspace): the D2H of step i overlaps the kernels of step i + 1, the
    # host consumes step i - 1's results while step i runs
    slots = [dict(stream=torch.cuda.Stream(), d_res=d_res, d_msgs=d_msgs, ws=ws, h_res=h_res, h_msgs=h_msgs),
             dict(stream=torch.cuda.Stream(), d_res=torch.empty_like(d_res), d_msgs=torch.empty_like(d_msgs), ws=torch.empty_like(ws),
                  h_res=torch.empty_like(h_res).pin_memory(), h_msgs=torch.empty_like(h_msgs).pin_memory())]
    seen = [0]

    def e2e_step(i):
        sl = slots[i & 1]
        with torch.cuda.stream(sl['stream']):
            lib.check(L.mst_decode_batch_device(arena.data_ptr(), d_songs.data_ptr(), S, n_chunks, sl['d_msgs'].data_ptr(), cap,
                                                sl['d_res'].data_ptr(), sl['ws'].data_ptr(), wsb, sl['stream'].cuda_stream))
            sl['h_res'].copy_(sl['d_res'], non_blocking=True)
            sl['h_msgs'].copy_(sl['d_msgs'], non_blocking=True)
        if i > 0:
            prev = slots[(i - 1) & 1]
            prev['stream'].synchronize()
            seen[0] += int(prev['h_res'].numpy().view(abi.DECODE_RESULT_DTYPE)['n_messages'].astype(np.int64).sum())

    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(args.steps):
        e2e_step(i)
    slots[(args.steps - 1) & 1]['stream'].synchronize()
    e2e_s = (time.perf_counter() - t0) / args.steps
    assert seen[0] == n_msgs * (args.steps - 1), 'decode e2e: message counts changed between steps'
    peak, peak_src = measured_peak_hbm()
    alg = grid_bytes + 12.0 * n_msgs
    return {
        'metric': 'grid cells/sec decoded', 'value': float(tot[0].item()) / 20 / s_per_step, 'unit': 'cells/s',
        'messages_per_s': float(tot[1].item()) / s_per_step, 'ms_per_step': s_per_step * 1e3, 'songs_per_step': S,
        'messages_per_step': n_msgs, 'gpu_launches_per_step': 6,
        'roofline': {'kernel': 'decode pipeline (k_dec_scan dominant: reads the grid once)', 'bound': 'hbm',
                     'achieved': alg / s_per_step / 1e9, 'peak': peak, 'unit': 'GB/s', 'frac': alg / s_per_step / 1e9 / peak,
                     'traffic': None, 'peak_source': peak_src, 'algorithmic_bytes_per_step': alg},
        'e2e': {'value': float(tot[0].item()) / 20 / e2e_s, 'unit': 'cells/s', 'h2d_bytes_per_step': 0,
                'd2h_bytes_per_step': int(h_res.numel() + n_msgs * abi.OUT_MSG_DTYPE.itemsize)},
    }


TRAIN_SHAPE = dict(C=4, B=200, nb=4, drums=True)     # the 800 channel-bar crop of train-model.py:101
LONG_SHAPE = dict(C=8, B=800, nb=4, drums=True)      # 8x the default sequence length (C.B = 6400)


def training_config(per_gpu, world, sh=None):
    sh = sh or TRAIN_SHAPE
    name = 'configs[4]: long-sequence training' if sh['C'] * sh['B'] > 800 else 'configs[3]: synthetic-batch training'
    return {'workload': f'{name} of the style model (C={sh["C"]} pitched channels x B={sh["B"]} bars, 4/4, '
                        'drums present, Bernoulli(0.003) grids, seed-108 weights), forward + get_total_loss + backward + '
                        'all-reduce + Adam per step', 'per_gpu_batch': per_gpu, 'global_batch': per_gpu * world,
            'gradient': 'sum over samples (iter_size semantics)', 'parallelism': f'dp{world}',
            'l2_policy': 'inputs larger than L2: >=155 MB of grids + >=0.8 GB of activations per sample'}


def k_grid_traffic(songs_per_step, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the grid-writing kernel from the committed ncu
    capture (profiles/r1_k_grid_traffic.json); only valid for the kernel and batch size it was captured at."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'r1_k_grid_traffic.json')) as f:
            t = json.load(f)
        same = int(t['songs_per_step']) == int(songs_per_step) and t.get('kernel', 'k_grid') == kernel
        return int(t['dram_bytes_per_launch']) if same else None
    except (OSError, ValueError, KeyError):
        return None


def algorithmic_bytes_per_sample(C, B, nb, drums):
    """read the input grids twice (fwd, bwd) + write and read the prediction grids once each."""
    return 4 * 4 * nb * 10 * (280 * C + (94 if drums else 0)) * B


def run_training(args, world, rank, torch, dist, sh=None, N=None, precision='bf16', e2e=True):
    """Training leg (BASELINE.json configs[3], or configs[4] with sh=LONG_SHAPE): samples/s over all ranks,
    device-timed, max over ranks.  precision: 'bf16' = the mode configs[3] names (bf16 dense layers: tcgen05 kind::f16
    GEMMs + PitchedStyleApplier tail), 'fp32' = the fp32-accurate 3xTF32 mode."""
    from mst_native import synth, trainer as tr   # workload generators live in the product package (no oracle/ import here)
    N = N or args.train_batch
    sh = sh or TRAIN_SHAPE
    distinct = [synth.synthetic_sample(seed=7 + 100 * rank + i, **sh) for i in range(min(N, 4))]
    pick = lambda k: torch.cat([distinct[i % len(distinct)][k] for i in range(N)]).contiguous()
    host = {k: pick(k).pin_memory() for k in ('pitched', 'unpitched', 'feats', 'mode', 'bpm')}
    host['it'] = pick('used').pin_memory()
    host['bt'] = torch.full((N,), 120.).pin_memory()
    dev = {k: v.cuda() for k, v in host.items()}
    t = tr.Trainer(synth.make_state_dict(108), precision=precision)
    losses = torch.empty(N, 16, device='cuda')
    h_losses = torch.empty(N, 16).pin_memory()

    def step():
        t.accumulate(dev['pitched'], dev['unpitched'], dev['feats'], dev['mode'], dev['bpm'], dev['it'], dev['bt'], losses)
        t.optimizer_step()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.train_steps):
        step()
    ev1.record()
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    s_per_step = float(ms.item()) / 1e3 / args.train_steps
    value = N * world / s_per_step
    launches = t.launches
    loss0 = float(losses[0, 0])

    peak, peak_src = measured_peak_hbm()
    alg = algorithmic_bytes_per_sample(**sh)
    base = {
        'metric': 'training samples/sec', 'value': value, 'unit': 'samples/s', 'ms_per_step': s_per_step * 1e3,
        'steps': args.train_steps, 'config': training_config(N, world, sh), 'dtype': 'bf16 operands, f32 accumulate' if precision == 'bf16' else 'f32 (3xTF32)',
        'loss_total_sample0': loss0, 'gpu_launches_per_step': launches,
        'roofline': {'kernel': 'whole training step (all kernels)', 'bound': 'hbm',
                     'achieved': alg * N / s_per_step / 1e9, 'peak': peak, 'unit': 'GB/s',
                     'frac': alg * N / s_per_step / 1e9 / peak, 'traffic': None, 'peak_source': peak_src,
                     'algorithmic_bytes_per_sample': alg}}
    if not e2e:
        del t, dev
        torch.cuda.empty_cache()
        return base
    # e2e: pinned host inputs -> H2D -> step -> D2H of the loss leaves, every step.  Two device slots: the H2D of batch
    # i + 1 runs on a copy stream while batch i trains (what a pin_memory DataLoader with one batch of prefetch does);
    # the loss leaves of every step are read back and consumed on the host one step later.
    slots = [dev, {k: torch.empty_like(v) for k, v in dev.items()}]
    h_loss = [h_losses, torch.empty(N, 16).pin_memory()]
    copy_stream = torch.cuda.Stream()
    ready = [torch.cuda.Event() for _ in range(2)]
    free = [torch.cuda.Event() for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]
    seen = [0.0]

    def upload(i):
        sl = i & 1
        copy_stream.wait_event(free[sl])                 # the step that read this slot has finished
        with torch.cuda.stream(copy_stream):
            for k in host:
                slots[sl][k].copy_(host[k], non_blocking=True)
            ready[sl].record(copy_stream)

    def e2e_step(i):
        sl = i & 1
        upload(i + 1)
        cur = torch.cuda.current_stream()
        cur.wait_event(ready[sl])
        d = slots[sl]
        t.accumulate(d['pitched'], d['unpitched'], d['feats'], d['mode'], d['bpm'], d['it'], d['bt'], losses)
        t.optimizer_step()
        free[sl].record(cur)
        h_loss[sl].copy_(losses, non_blocking=True)
        done[sl].record(cur)
        if i > 0:
            done[1 - sl].synchronize()
            seen[0] += float(h_loss[1 - sl][:, 0].sum())  # the host consumes the previous step's losses

    for ev in free:
        ev.record()
    upload(0)
    for i in range(2):
        e2e_step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(2, 2 + args.train_steps):
        e2e_step(i)
    done[(1 + args.train_steps) & 1].synchronize()
    torch.cuda.current_stream().synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device='cuda')
    torch.cuda.synchronize()
    barrier()
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    h2d = int(sum(v.numel() * v.element_size() for v in host.values()))
    del slots, dev
    torch.cuda.empty_cache()
    e2e_events = training_e2e_from_events(args, world, rank, torch, dist, t, sh, N, losses, barrier)
    out = dict(base)
    out.update({
        'e2e': e2e_events,
        'e2e_dense_grids': {'value': N * world * args.train_steps / float(e2e_s.item()), 'unit': 'samples/s',
                            'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': int(h_losses.numel() * 4),
                            'note': 'the same step fed with dense f32 grids from pinned host memory (what the reference '
                                    'uploads, data.py:144-149): two device slots, H2D of batch i+1 overlaps the step of '
                                    'batch i; PCIe-bound'},
    })
    del t
    torch.cuda.empty_cache()
    return out


def run_files(args, torch, lib, abi, synth, filt):
    """File-fed encode through the product pipeline (mst_native.pipeline.EventPipeline, what style.data.iter_inputs runs on):
    .mid FILES on disk -> C++ thread-pool parse into page-locked buffers -> mst_encode_batch_host_async on rotating slots.
    The same files go through the CPU arm (same native parser, then the C oracle port on all host threads)."""
    import shutil
    import tempfile
    from mst_native.pipeline import EventPipeline
    n_files = args.file_songs
    events, songs = synth.synth_corpus(n_files, seed=777)
    per = synth.split_songs(events, songs)
    tmp = tempfile.mkdtemp(prefix='mst_bench_')
    try:
        paths, n_bytes = [], 0
        for i, (e, sg) in enumerate(zip(per, songs)):
            data = lib.smf_write_events(e, int(sg['ticks_per_beat']))
            paths.append(os.path.join(tmp, f'song{i:05d}.mid'))
            with open(paths[-1], 'wb') as f:
                f.write(data)
            n_bytes += len(data)
        batch = min(128, n_files)
        probe = lib.DeviceBatch(*lib.make_songs(per[:batch], [int(t) for t in songs['ticks_per_beat'][:batch]]))
        probe.encode(filt, None)
        arena_floats = int(1.1 * sum(lib.grid_floats(r) for r in probe.infos()))
        del probe
        def timed(wire):
            pipe = EventPipeline(filt, batch_files=batch, slots=3, arena_floats=arena_floats, wire=wire)

            def one_pass():
                bars = ok = 0
                for _, r, _ in pipe.run(paths):
                    if int(r['status']) == 0:
                        ok += 1
                        bars += int(r['n_bars_out'])
                torch.cuda.synchronize()
                return ok, bars
            one_pass()                                            # warm-up: page cache, pinned buffers, workspaces
            t0 = time.perf_counter()
            ok, bars = one_pass()
            dt = time.perf_counter() - t0
            pipe.close()
            del pipe
            torch.cuda.empty_cache()
            return ok, bars, dt
        ok, bars, dt = timed('smf')
        ok_e, bars_e, dt_e = timed('events')
        assert (ok, bars) == (ok_e, bars_e), ((ok, bars), (ok_e, bars_e))
        out = {'metric': 'piano-roll bars/sec encoded from .mid files on disk', 'value': bars / dt, 'unit': UNIT, 'files_per_s': ok / dt,
               'files': n_files, 'files_ok': ok, 'file_bytes': n_bytes, 'batch_files': batch, 'ms_per_file': dt / n_files * 1e3,
               'note': 'EventPipeline(wire="smf"): mst_files_read_batch (C++ threads) -> pinned file bytes -> mst_encode_batch_smf_async '
                       '(device-side parse), three slots',
               'host_parse_path': {'value': bars_e / dt_e, 'unit': UNIT, 'files_per_s': ok_e / dt_e,
                                   'note': 'EventPipeline(wire="events"): mst_smf_load_batch parses on the host cores, '
                                           'mst_encode_batch_host_async uploads 16-byte event records'}}
        # CPU arm on the same files: same native parser on all cores + the C oracle port of the encoder
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import oracle_lib
        cores = os.cpu_count() or 1
        sub = paths[:min(n_files, max(2 * cores, 32))]
        buf = np.empty(sum(len(per[i]) for i in range(len(sub))) + 16, dtype=abi.EVENT_DTYPE)
        t0 = time.perf_counter()
        total, sg2, status = lib.smf_load_batch(sub, buf, n_threads=cores)
        t_parse = time.perf_counter() - t0
        infos, _ = oracle_lib.encode_batch(buf[:total], sg2, filt, cores)
        dt_cpu = time.perf_counter() - t0
        out['cpu_baseline'] = {'value': float(infos['n_bars_out'][infos['status'] == 0].sum()) / dt_cpu, 'unit': UNIT, 'cores': cores,
                               'kind': 'port', 'parse_share': t_parse / dt_cpu,
                               'sample': f'{len(sub)} of the same files: native SMF parser + oracle/mst_oracle.c, {cores} threads'}
        return out
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def run_training_n1(args, torch):
    """N = 1: the batch the reference and the shipped train-model.py actually run (one song per iteration, an optimizer
    step every iter_size = 2 iterations, train-model.py:97-154), through the drop-in call pattern
        model(...) -> get_total_loss(...) -> loss.backward() -> optimizer.step()
    with one CUDA graph per entry point and shape, and through the C-ABI Trainer (one graph per iteration)."""
    from mst_native import synth, trainer as tr
    from style import model as m
    sh = TRAIN_SHAPE
    samples = [synth.synthetic_sample(seed=7 + i, **sh) for i in range(2)]
    host = [{k: v.pin_memory() for k, v in s.items() if torch.is_tensor(v)} for s in samples]
    dev = [{k: v.cuda() for k, v in h.items()} for h in host]
    model = synth.build_model(108)
    opt = torch.optim.Adam(model.parameters(), lr=.01)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=200, gamma=.9)
    opt.zero_grad()

    def iteration(d, i):
        (ip, mp, bp), pp, up = model(d['mode'], d['bpm'], d['pitched'], d['feats'], d['unpitched'])
        losses = m.get_total_loss(ip, d['used'], bp, 120, mp, d['mode'], pp, d['pitched'], up, d['unpitched'], normalize=True)
        losses['total'].backward()
        vals = [float(x) for x in (losses['total'], losses['channels_loss']['total'], losses['song_info_loss']['total'])]
        if i % 2 == 1:
            opt.step(); opt.zero_grad(); sched.step()
        return vals[0]

    K = max(4, 2 * args.train_steps)
    for i in range(6):
        loss0 = iteration(dev[i & 1], i)
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(K):
        iteration(dev[i & 1], i)
    ev1.record()
    torch.cuda.synchronize()
    ms_dropin = ev0.elapsed_time(ev1) / K
    # e2e: the sample's tensors come from pinned host memory every iteration (what prepare_input's .to(device) does)
    slots = [{k: torch.empty_like(v) for k, v in dev[0].items()} for _ in range(2)]
    t0 = time.perf_counter()
    for i in range(K):
        for k, v in host[i & 1].items():
            slots[i & 1][k].copy_(v, non_blocking=True)
        iteration(slots[i & 1], i)
    torch.cuda.synchronize()
    e2e_s = (time.perf_counter() - t0) / K
    h2d = int(sum(v.numel() * v.element_size() for v in host[0].values()))
    del model, opt, slots
    torch.cuda.empty_cache()
    # the same iteration through the C-ABI Trainer: ONE graph replay (forward + loss + backward) + fused Adam every 2nd
    out = {}
    for graph in (True, False):
        t = tr.Trainer(synth.make_state_dict(108), graph=graph, precision='fp32')
        losses = torch.empty(1, 16, device='cuda')

        def tstep(i):
            d = dev[i & 1]
            t.accumulate(d['pitched'], d['unpitched'], d['feats'], d['mode'], d['bpm'], d['used'], torch.full((1,), 120., device='cuda'), losses)
            if i % 2 == 1:
                t.optimizer_step()
        for i in range(6):
            tstep(i)
        torch.cuda.synchronize()
        ev0.record()
        for i in range(K):
            tstep(i)
        ev1.record()
        torch.cuda.synchronize()
        out['graph' if graph else 'launches'] = ev0.elapsed_time(ev1) / K
        n_launches = t.launches
        del t
        torch.cuda.empty_cache()
    return {
        'metric': 'training iterations/sec at N = 1', 'value': 1e3 / out['graph'], 'unit': 'samples/s',
        'ms_per_iteration': out['graph'], 'ms_per_iteration_without_graph': out['launches'], 'kernels_per_iteration': n_launches,
        'dropin': {'ms_per_iteration': ms_dropin, 'value': 1e3 / ms_dropin, 'unit': 'samples/s',
                   'note': 'style.model drop-in: model() + get_total_loss() + loss.backward() + torch.optim.Adam.step(); graphs per '
                           'entry point; ONE device->host copy of the loss leaves per iteration', 'loss_total': loss0},
        'e2e': {'value': 1.0 / e2e_s, 'unit': 'samples/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 64,
                'note': 'drop-in iteration with the sample uploaded from pinned host memory every iteration'},
        'config': training_config(1, 1), 'dtype': 'f32 (3xTF32)',
    }


def training_e2e_from_events(args, world, rank, torch, dist, t, sh, N, losses, barrier):
    """Training throughput through the drop-in pipeline with HOST event buffers (train-model.py:97-154 on the CUDA path):
    every step uploads the songs' MIDI events (pinned host memory), encodes them on the device with prepare_input's crop
    (mst_encode_batch_host_async, crop_channel_bars = C.B), gathers the batch tensors from the encoder's arena, runs
    forward + loss + backward + all-reduce + Adam, and reads the loss leaves back.  The grids never cross PCIe.
    Two slots: the upload + encode of batch i + 1 (side stream) overlaps the training step of batch i."""
    sys.path.insert(0, os.path.join(ROOT, 'music-style-transfer_b200'))
    import numpy as np
    from mst_native import abi, lib, synth
    from style import data as sdata
    L = lib.load()
    C, B, nb = sh['C'], sh['B'], sh['nb']
    filt = abi.make_filter(synth.POPULAR, 100, crop_channel_bars=C * B)
    # candidate songs: C pitched tracks + drums over B bars; keep those the encoder turns into exactly (C, B) + drums
    # (tracks that drew the same program are merged by merge_nchannels, data.py:103-114)
    ev, sg = synth.synth_corpus(N + 8, seed=4321 + 17 * rank, n_bars=B, n_tracks=C + 1)
    ev = ev.copy()
    pc = ev['kind'] == abi.EV_PROGRAM_CHANGE              # one distinct program per pitched track: C channels after merging
    ev['d1'][pc] = np.array(sorted(synth.POPULAR), dtype=np.uint8)[(ev['track'][pc].astype(np.int64) - 1) % len(synth.POPULAR)]
    per_song = synth.split_songs(ev, sg)
    cand = lib.DeviceBatch(ev, sg)
    cand.encode(filt, None)
    plan = cand.infos()
    keep = [i for i, r in enumerate(plan) if int(r['status']) in (0, 8) and int(r['n_pitched']) == C
            and int(r['n_unpitched']) == 1 and int(r['n_bars_out']) == B and int(r['n_beats']) == nb][:N]
    assert len(keep) == N, f'only {len(keep)} of {len(plan)} synthetic songs have the ({C}, {B}) shape'
    del cand
    events, songs = lib.make_songs([per_song[i] for i in keep], [int(sg['ticks_per_beat'][i]) for i in keep])
    recs = [plan[i] for i in keep]
    P, U = C * B * nb * abi.PITCHED_BEAT_FLOATS, B * nb * abi.UNPITCHED_BEAT_FLOATS
    h_events = torch.from_numpy(events.view(np.uint8).reshape(-1)).pin_memory()
    h_songs = torch.from_numpy(songs.view(np.uint8).reshape(-1)).pin_memory()
    # the small per-song tensors prepare_input / get_used_instruments derive from the song infos (data.py:122-169)
    feats = torch.stack([torch.as_tensor(sdata.encode_instruments([int(x) for x in r['pitched_instruments'][:C]]),
                                         dtype=torch.float32) for r in recs])
    small = {'feats': feats,
             'mode': torch.tensor([[1., 0.] if int(r['mode']) == 0 else [0., 1.] for r in recs]),
             'bpm': torch.tensor([float(r['bpm']) for r in recs]),
             'it': torch.cat([sdata.get_used_instruments(feats[i:i + 1], True) for i in range(N)]),
             'bt': torch.tensor([float(r['bpm']) for r in recs])}
    small = {k: v.contiguous().pin_memory() for k, v in small.items()}
    arena_floats = N * (P + U)
    wsb = L.mst_encode_host_workspace_bytes(len(events), N, arena_floats)
    slots = []
    for _ in range(2):
        slots.append({'stream': torch.cuda.Stream(), 'ws': torch.empty(wsb, dtype=torch.uint8, device='cuda'),
                      'arena': torch.empty(arena_floats, dtype=torch.float32, device='cuda'),
                      'info': torch.empty(N * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8).pin_memory(),
                      'pitched': torch.empty(N, C, B, nb, 10, 56, 5, device='cuda'),
                      'unpitched': torch.empty(N, 1, B, nb, 10, 47, 2, device='cuda'),
                      'small': {k: torch.empty_like(v, device='cuda') for k, v in small.items()},
                      'h_loss': torch.empty(N, 16).pin_memory(),
                      'ready': torch.cuda.Event(), 'free': torch.cuda.Event(), 'done': torch.cuda.Event()})
    seen = [0.0, 0]

    def produce(i):
        sl = slots[i & 1]
        st = sl['stream']
        st.wait_event(sl['free'])                        # the training step that read this slot has finished
        lib.check(L.mst_encode_batch_host_async(
            h_events.data_ptr(), len(events), h_songs.data_ptr(), N, filt.reshape(1).ctypes.data, sl['info'].data_ptr(),
            sl['arena'].data_ptr(), arena_floats, sl['ws'].data_ptr(), wsb, st.cuda_stream))
        with torch.cuda.stream(st):
            # songs are laid out one after the other, each (pitched, unpitched): gather the two batch tensors
            per = sl['arena'].view(N, P + U)
            sl['pitched'].view(N, P).copy_(per[:, :P])
            sl['unpitched'].view(N, U).copy_(per[:, P:])
            for k, v in small.items():
                sl['small'][k].copy_(v, non_blocking=True)
            sl['ready'].record(st)

    def e2e_step(i):
        sl = slots[i & 1]
        produce(i + 1)
        cur = torch.cuda.current_stream()
        cur.wait_event(sl['ready'])
        d = sl['small']
        t.accumulate(sl['pitched'], sl['unpitched'], d['feats'], d['mode'], d['bpm'], d['it'], d['bt'], losses)
        t.optimizer_step()
        sl['free'].record(cur)
        sl['h_loss'].copy_(losses, non_blocking=True)
        sl['done'].record(cur)
        if i > 0:
            prev = slots[1 - (i & 1)]
            prev['done'].synchronize()
            seen[0] += float(prev['h_loss'][:, 0].sum())          # the host consumes the previous step's losses
            info = prev['info'].numpy().view(abi.SONG_INFO_DTYPE)  # ... and its song infos
            seen[1] += int((info['status'] == 0).sum())

    for sl in slots:
        sl['free'].record()
    produce(0)
    for i in range(2):
        e2e_step(i)
    barrier()
    chk = slots[0]['info'].numpy().view(abi.SONG_INFO_DTYPE)
    assert all(int(r['status']) == 0 and int(r['pitched_offset']) == i * (P + U) and int(r['unpitched_offset']) == i * (P + U) + P
               for i, r in enumerate(chk)), 'arena layout differs from the batch gather'
    seen[1] = 0
    t0 = time.perf_counter()
    for i in range(2, 2 + args.train_steps):
        e2e_step(i)
    slots[(1 + args.train_steps) & 1]['done'].synchronize()
    torch.cuda.current_stream().synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device='cuda')
    torch.cuda.synchronize()
    barrier()
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    assert seen[1] >= N * (args.train_steps - 1), 'the encoder rejected songs inside the timed region'
    h2d = int(h_events.numel() + h_songs.numel() + sum(v.numel() * v.element_size() for v in small.values()))
    out = {'value': N * world * args.train_steps / float(dt.item()), 'unit': 'samples/s', 'h2d_bytes_per_step': h2d,
           'd2h_bytes_per_step': int(N * 16 * 4 + N * abi.SONG_INFO_DTYPE.itemsize),
           'events_per_step': int(len(events)), 'loss_total_sample0': float(slots[0]['h_loss'][0, 0]),
           'note': 'host MIDI events -> H2D -> mst_encode_batch_host_async (crop = C.B channel-bars) -> batch gather -> '
                   'training step -> D2H of loss leaves and song infos, every step; upload + encode of batch i+1 on a '
                   'side stream overlap the training step of batch i; the grids never cross PCIe'}
    del slots
    torch.cuda.empty_cache()
    return out


def cpu_training_baseline_here(threads, warmup=3, steps=10):
    """Runs INSIDE the --cpu-train-baseline child process: the reference's own training iteration on the host cores.
    baseline/_ref holds the unmodified reference (oracle/install_reference.py); its `style.model` + `style.utils.pytorch`
    need torch alone, so StyleTransferModel / get_total_loss / backward / torch.optim.Adam are the reference's code
    (`kind: "reference"`).  When baseline/_ref is absent the plain-torch oracle restatement is timed instead (`"port"`)."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import model_oracle as mo            # sample generator (and the fallback model)
    torch.set_num_threads(threads)
    s = mo.synthetic_sample(seed=7, **TRAIN_SHAPE)
    used = mo.used_instruments(s['feats'], s['unpitched'])
    ref_root = os.path.join(ROOT, 'baseline', '_ref')
    kind = 'port'
    if os.path.exists(os.path.join(ref_root, 'style', 'model.py')):
        sys.path.insert(0, ref_root)
        import style.model as m
        assert os.path.abspath(m.__file__).startswith(os.path.abspath(ref_root)), m.__file__
        kind = 'reference'
        torch.manual_seed(108)       # train-model.py:52-85: construction order fixes the weights
        isz, nin = 51, 41            # style.data.instrument_size / n_instruments (style.data itself needs mido)
        pce = m.PitchedChannelsEncoder(64, 128, isz); uce = m.UnpitchedChannelsEncoder(64, 128)
        pre = m.PitchedRhythmEncoder(32, 64, 128, isz); ure = m.UnpitchedRhythmEncoder(32, 64, 128)
        se = m.StyleEncoder(256, 128, isz); me = m.MelodyEncoder(8, 64, 128, isz)
        sim = m.SongInfoModel(8, 256, 32, nin); psa = m.PitchedStyleApplier(256, 8, 32, isz); usa = m.UnpitchedStyleApplier(256, 32)
        model = m.StyleTransferModel(pce, uce, se, me, pre, ure, sim, psa, usa)
        opt = torch.optim.Adam(model.parameters(), lr=0.01)

        def it():
            (ip, mp, bp), pp, up = model(s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'])
            losses = m.get_total_loss(ip, used, bp, 120, mp, s['mode'], pp, s['pitched'], up, s['unpitched'], normalize=True)
            losses['total'].backward()
            opt.step(); opt.zero_grad()
            return float(losses['total'])
    else:
        sd = {k: v.clone().requires_grad_(True) for k, v in mo.make_state_dict(108).items()}
        opt = torch.optim.Adam(list(sd.values()), lr=0.01)

        def it():
            losses, _ = mo.training_losses(sd, s['mode'], s['bpm'], s['pitched'], s['feats'], s['unpitched'], s['bpm_target'])
            losses['total'].sum().backward()
            opt.step(); opt.zero_grad()
            return float(losses['total'])
    for _ in range(warmup):
        loss0 = it()
    t0 = time.perf_counter()
    for _ in range(steps):
        it()
    dt = (time.perf_counter() - t0) / steps
    return {'value': 1.0 / dt, 'unit': 'samples/s', 'cores': threads, 'kind': kind, 'ms_per_iteration': dt * 1e3,
            'loss_first_iteration': loss0,
            'sample': f'{warmup} warm-up + {steps} timed iterations of one C=4 x B=200 sample (fwd + get_total_loss + bwd + Adam), '
                      f'torch {torch.__version__} CPU, {threads} threads'
                      + ('; the unmodified reference style/model.py from baseline/_ref' if kind == 'reference' else '; oracle/model_oracle.py')}


def cpu_training_baseline(threads, warmup=3, steps=10):
    """The CPU arm of the training metric, in a child process (the reference package and the drop-in are both `style`)."""
    r = subprocess.run([sys.executable, os.path.abspath(__file__), '--cpu-train-baseline', str(threads), str(warmup), str(steps)],
                       capture_output=True, text=True, timeout=900)
    for line in reversed(r.stdout.strip().splitlines()):
        if line.startswith('{'):
            return json.loads(line)
    return {'value': None, 'unit': 'samples/s', 'cores': threads, 'kind': 'unavailable', 'sample': (r.stderr or r.stdout)[-300:]}


def run_reference(args):
    """CPU arm.  Codec: the oracle port (oracle/mst_oracle.c, pthreads over all host cores) - the reference's codec is
    pure Python on top of mido, which is neither in the image nor in the wheelhouse, and the C port is a strictly faster
    restatement of the same algorithm.  Training: the unmodified reference model from baseline/_ref."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import oracle_lib
    from mst_native import abi, synth
    cores = os.cpu_count() or 1
    # a step = a bounded sample of the configs[2] workload: `sample` songs (same generator, same filter), a DIFFERENT
    # slice of the corpus every step
    sample = max(cores * 2, 32)
    n_steps = args.warmup + args.steps
    events, songs = synth.synth_corpus(sample * min(n_steps, 4), seed=1234)
    per = synth.split_songs(events, songs)
    from mst_native import lib as _l
    slices = []
    for k in range(min(n_steps, 4)):
        slices.append(_l.make_songs(per[k * sample:(k + 1) * sample], [int(t) for t in songs['ticks_per_beat'][k * sample:(k + 1) * sample]]))
    filt = abi.make_filter(synth.POPULAR, 100)
    for i in range(args.warmup):
        oracle_lib.encode_batch(*slices[i % len(slices)], filt, cores)
    t0 = time.perf_counter()
    bars = 0
    for i in range(args.steps):
        infos, _ = oracle_lib.encode_batch(*slices[(args.warmup + i) % len(slices)], filt, cores)
        bars += int(infos['n_bars_out'][infos['status'] == 0].sum())
    dt = time.perf_counter() - t0
    value = bars / dt
    cfg = workload_config(args)
    cfg.update({'songs_per_step': sample, 'distinct_songs_in_pool': sample * len(slices), 'corpus_files_equivalent': sample * args.steps,
                'streams': f'{cores} host threads', 'l2_policy': 'n/a (host)',
                'note': 'the CPU arm encodes a bounded sample per step (songs_per_step above), not the 512 songs of the GPU arm'})
    training = cpu_training_baseline(cores)
    training['config'] = training_config(1, 1)
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
        'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64 quantise -> f32 grid',
        'data': 'synthetic', 'config': cfg,
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port',
                         'sample': f'{sample} songs of the corpus per step, {args.steps} steps, {len(slices)} distinct slices'},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'training': training,
        'summary': {'encode_song_bars_per_s': value, 'train_sps': training.get('value'), 'train_kind': training.get('kind'),
                    'cores': cores},
    }))


def main():
    if len(sys.argv) >= 2 and sys.argv[1] == '--cpu-train-baseline':
        threads, warmup, steps = (int(x) for x in (sys.argv[2:5] + ['3', '10'])[:3]) if len(sys.argv) > 2 else (os.cpu_count() or 1, 3, 10)
        print(json.dumps(cpu_training_baseline_here(threads, warmup, steps)))
        return
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--batch', type=int, default=512, help='songs per step per GPU')
    ap.add_argument('--pool', type=int, default=1024, help='distinct songs generated per GPU')
    ap.add_argument('--train-batch', type=int, default=64, help='training samples per step per GPU')
    ap.add_argument('--train-steps', type=int, default=0, help='timed training steps (default: min(steps, 10))')
    ap.add_argument('--no-train', action='store_true', help='skip the training leg')
    ap.add_argument('--e2e-slots', type=int, default=3, help='host-path slots (stream + workspace + arena) in rotation')
    ap.add_argument('--long-batch', type=int, default=8, help='long-sequence (C=8, B=800) samples per step per GPU')
    ap.add_argument('--no-long', action='store_true', help='skip the long-sequence training leg')
    ap.add_argument('--file-songs', type=int, default=1024, help='songs written as .mid files for the file-fed encode leg (0: skip)')
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == 'ours' else args.warmup
    args.pool = max(args.pool, args.batch)
    args.train_steps = args.train_steps or max(1, min(args.steps, 10))
    if args.impl == 'reference':
        run_reference(args)
        return

    import torch
    import torch.distributed as dist
    from mst_native import abi, lib, synth

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: the B200 path has no CPU fallback')
    torch.cuda.set_device(local_rank)
    if world > 1:
        # stdout carries ONE JSON line; NCCL prints its banner / INFO lines to stdout unless NCCL_DEBUG_FILE names a file
        if os.environ.get('MST_NCCL_DEBUG'):
            os.environ['NCCL_DEBUG'] = os.environ['MST_NCCL_DEBUG']
        if os.environ.get('NCCL_DEBUG'):
            # keep what the environment asked for, but route the log (banner included) away from stdout
            os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    L = lib.require_gpu()

    # ---- synthetic corpus shard of this rank, resident in HBM (events are NOT part of the timed copies)
    n_batches = max(1, args.pool // args.batch)
    filt = abi.make_filter(synth.POPULAR, 100)
    batches, host_batches = [], []
    for b in range(n_batches):
        events, songs = synth.synth_corpus(args.batch, seed=1234 + 1000 * rank + b)
        batches.append(lib.DeviceBatch(events, songs))
        host_batches.append((events, songs))
    plans = []
    for batch in batches:
        batch.encode(filt, None)
        plans.append(batch.infos())
    floats = [sum(lib.grid_floats(r) for r in p if r['status'] in (0, 8)) for p in plans]
    bars = [int(sum(int(r['n_bars_out']) for r in p if r['status'] in (0, 8))) for p in plans]
    chan_bars = [int(sum(int(r['n_bars_out']) * (int(r['n_pitched']) + int(r['n_unpitched'])) for r in p
                         if r['status'] in (0, 8))) for p in plans]
    arena = torch.empty(max(floats), dtype=torch.float32, device='cuda')
    # two output rings (arena + stream), alternated step by step: the library runs the latency-bound front end of a batch
    # on its own high-priority stream, so the next batch's sort / walk / info hide behind the current batch's zero fill
    arenas = [arena, torch.empty_like(arena)]
    rings = [torch.cuda.Stream(), torch.cuda.Stream()]

    def step(i):
        batches[i % n_batches].encode(filt, arenas[i & 1], stream=rings[i & 1])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank) if rank == 0 else None
    for i in range(args.warmup):
        step(i)
    # sanity: the run really produced grids
    torch.cuda.synchronize()
    check = batches[(args.warmup - 1) % n_batches].infos()
    assert all(int(r['status']) == 0 for r in check), 'warm-up step reported non-OK songs'

    # ---- timed region: K steps, device timed, max over ranks
    barrier()
    t_load0 = time.time()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for r in rings:
        r.wait_event(ev0)
    for i in range(args.steps):
        step(i)
    for r in rings:
        done = torch.cuda.Event()
        done.record(r)
        torch.cuda.current_stream().wait_event(done)
    ev1.record()
    barrier()
    elapsed_ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device='cuda')
    total_bars = torch.tensor([sum(bars[i % n_batches] for i in range(args.steps))], dtype=torch.float64, device='cuda')
    total_chan_bars = torch.tensor([sum(chan_bars[i % n_batches] for i in range(args.steps))], dtype=torch.float64,
                                   device='cuda')
    if world > 1:
        dist.all_reduce(elapsed_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(total_bars, op=dist.ReduceOp.SUM)
        dist.all_reduce(total_chan_bars, op=dist.ReduceOp.SUM)
    elapsed_s = float(elapsed_ms.item()) / 1e3
    value = float(total_bars.item()) / elapsed_s

    # ---- roofline of the dominant kernel (the grid-writing fill), CUDA events recorded by the library on its stream
    lib.check(L.mst_profile_enable(1))
    grid_ms, front_ms, grid_bytes = [], [], []
    for i in range(args.steps):
        step(i)
        torch.cuda.synchronize()
        grid_ms.append(L.mst_profile_encode_ms(1))
        front_ms.append(L.mst_profile_encode_ms(0))
        p = plans[i % n_batches]
        n_cells = int(sum(int(r['n_notes']) for r in batches[i % n_batches].infos())) if i < n_batches else None
        grid_bytes.append(floats[i % n_batches] * 4)
    lib.check(L.mst_profile_enable(0))
    peak, peak_src = measured_peak_hbm()
    g_ms = statistics.mean(grid_ms)
    g_bytes = statistics.mean(grid_bytes)
    achieved = g_bytes / (g_ms * 1e-3) / 1e9
    ev_bytes = statistics.mean(len(h[0]) * 16 for h in host_batches)
    whole_ms = statistics.mean(front_ms)      # which = 0: k_sort .. k_cells on the caller's stream
    tiles = bool(os.environ.get('MST_GRID_TILES'))
    roofline = {
        'kernel': 'k_grid' if tiles else 'k_fill_flat', 'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
        'frac': achieved / peak, 'traffic': k_grid_traffic(args.batch, 'k_grid' if tiles else 'k_fill_flat'),
        'peak_source': peak_src + '; copy bandwidth (read + write bytes) - a write-only stream can exceed it: torch '
                                  'fill_ of 8 GiB measures 7535 GB/s on this part (tools/write_peak_probe.py)',
        'algorithmic_bytes_per_launch': g_bytes, 'ms_per_launch': g_ms,
        'frac_of_write_only_peak': achieved / 7535.0, 'write_only_peak_source': 'profiles/r1_write_peak_probe.txt',
        'share_of_step': g_ms / whole_ms,
        'pipeline': {'whole_ms': whole_ms, 'not_hidden_behind_the_fill_ms': whole_ms - g_ms,
                     'note': 'the zero fill runs on a side stream next to k_walk<1> / k_pair / k_scale; k_sort, k_walk<0>, '
                             'k_info, k_offsets run before it and k_cells (note scatter) after it',
                     'whole_step_achieved_GBps': (g_bytes + ev_bytes) / (elapsed_s / args.steps) / 1e9},
    }

    # ---- e2e: host buffers through the C ABI (H2D of events + song table, kernels, D2H of the infos), every step.
    # Three slots (stream + workspace + arena + info buffer) rotate, so the H2D of batch i + 2 is already queued while batch i
    # drains (with two, a late k_cells of batch i delays the upload of batch i + 2 into a bubble: 7.3M vs 9.4M song-bars/s)
    # and overlaps the kernels of
    # batch i: the same prefetch the reference does with a thread (utils/parallel.py:34-39).  All copies are timed.
    # the host buffers hold the 8-byte wire format of the events (mst_events_pack, done once outside the timed region: it is
    # the form a loader writes); mst_encode_batch_packed_async widens them on the device.  Half the H2D volume of MstEvent
    # records: what matters once several ranks share one host socket
    # Round 2b: the headline e2e leg uploads the FILES THEMSELVES (Standard MIDI File bytes, 3-4 bytes per event) and parses
    # them on the device (mst_encode_batch_smf_async); the packed-event leg is measured beside it for comparison.
    def song_bytes(events, songs):
        blobs = []
        for srec in songs:
            o, k = int(srec['event_offset']), int(srec['n_events'])
            blobs.append(lib.smf_write_events(events[o:o + k], int(srec['ticks_per_beat'])))
        return lib.concat_files(blobs)
    smf = [song_bytes(*h) for h in host_batches]
    h_bytes = [torch.from_numpy(b).pin_memory() for b, _ in smf]
    h_files = [torch.from_numpy(f.view(np.uint8).reshape(-1)).pin_memory() for _, f in smf]
    n_by = [len(b) for b, _ in smf]
    h_events = [torch.from_numpy(lib.events_pack(h[0]).view(np.uint8)).pin_memory() for h in host_batches]
    h_songs = [torch.from_numpy(h[1].view(np.uint8).reshape(-1)).pin_memory() for h in host_batches]
    n_ev = [len(h[0]) for h in host_batches]
    wsb = max(max(L.mst_encode_packed_workspace_bytes(n, args.batch, arena.numel()) for n in n_ev),
              max(L.mst_encode_smf_workspace_bytes(n, args.batch, arena.numel()) for n in n_by))
    for b in batches:
        b._ws = None   # release the device-path workspaces before allocating the host-path ones
    n_slots = args.e2e_slots
    slots = []
    for k in range(n_slots):
        slots.append({'stream': torch.cuda.Stream(),
                      'ws': torch.empty(wsb, dtype=torch.uint8, device='cuda'),
                      'arena': arenas[k] if k < len(arenas) else torch.empty_like(arena),
                      'info': torch.empty(args.batch * abi.SONG_INFO_DTYPE.itemsize, dtype=torch.uint8).pin_memory()})

    def run_e2e(wire):
        e2e_bars = [0]

        def e2e_drain(sl):
            """Wait for the slot's batch and consume its infos (what a caller does before handing the grids on)."""
            sl['stream'].synchronize()
            if sl.get('busy'):
                info = sl['info'].numpy().view(abi.SONG_INFO_DTYPE)
                ok = info['status'] == 0
                e2e_bars[0] += int(info['n_bars_out'][ok].sum())
                sl['busy'] = False

        def e2e_step(i):
            j = i % n_batches
            sl = slots[i % n_slots]
            e2e_drain(sl)
            if wire == 'smf':
                lib.check(L.mst_encode_batch_smf_async(
                    h_bytes[j].data_ptr(), n_by[j], h_files[j].data_ptr(), args.batch, filt.reshape(1).ctypes.data,
                    sl['info'].data_ptr(), sl['arena'].data_ptr(), sl['arena'].numel(), sl['ws'].data_ptr(), wsb,
                    sl['stream'].cuda_stream))
            else:
                lib.check(L.mst_encode_batch_packed_async(
                    h_events[j].data_ptr(), n_ev[j], h_songs[j].data_ptr(), args.batch, filt.reshape(1).ctypes.data,
                    sl['info'].data_ptr(), sl['arena'].data_ptr(), sl['arena'].numel(), sl['ws'].data_ptr(), wsb,
                    sl['stream'].cuda_stream))
            sl['busy'] = True

        for i in range(3):
            e2e_step(i)
        for sl in slots:
            e2e_drain(sl)
        warm_bars = e2e_bars[0]
        e2e_bars[0] = 0
        barrier()
        t0 = time.perf_counter()
        for i in range(args.steps):
            e2e_step(i)
        for sl in slots:
            e2e_drain(sl)
        torch.cuda.synchronize()
        e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device='cuda')
        barrier()
        e2e_per_rank = [float(e2e_s.item())]
        if world > 1:
            each = [torch.zeros_like(e2e_s) for _ in range(world)]
            dist.all_gather(each, e2e_s)
            e2e_per_rank = [float(t.item()) for t in each]
            dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
        e2e_total = torch.tensor([float(e2e_bars[0])], dtype=torch.float64, device='cuda')
        if world > 1:
            dist.all_reduce(e2e_total, op=dist.ReduceOp.SUM)
        h2d = (statistics.mean(n_by) + args.batch * 16) if wire == 'smf' else (statistics.mean(n_ev) * 8 + args.batch * 16)
        assert warm_bars == sum(bars[i % n_batches] for i in range(3)), (wire, warm_bars)      # same songs encoded as on the device path
        return {'value': float(e2e_total.item()) / float(e2e_s.item()), 'unit': UNIT,
                'h2d_bytes_per_step': int(h2d),
                'd2h_bytes_per_step': int(args.batch * abi.SONG_INFO_DTYPE.itemsize),
                'ms_per_step_per_rank': [round(t / args.steps * 1e3, 3) for t in e2e_per_rank],
                'h2d_GBps_per_rank': [round(h2d * args.steps / t / 1e9, 2) for t in e2e_per_rank]}

    e2e = run_e2e('smf')
    e2e['bytes_per_event'] = round(statistics.mean(n_by) / statistics.mean(n_ev), 2)
    e2e['note'] = ('mst_encode_batch_smf_async: the batch\'s Standard MIDI File bytes are uploaded as they are and parsed on the '
                   'device (thread per track), three rotating slots (H2D of batch i+1, i+2 overlaps the kernels of batch i); grids '
                   'stay in HBM for the training consumer (SURVEY.md 8(b)); infos are read back and consumed every step')
    e2e['packed_events'] = run_e2e('events')
    e2e['packed_events']['note'] = 'the same loop on mst_encode_batch_packed_async (8-byte event records parsed on the host beforehand)'
    t_load1 = time.time()
    # ---- decode leg over the grids of batch 0 (round trip of the encoder's output)
    decode = run_decode(args, world, rank, torch, dist, L, lib, abi, batches[0], filt, arena)
    files_leg = None
    if world == 1 and args.file_songs > 0:
        del slots
        for b in batches[1:]:
            b._ws = None
        torch.cuda.empty_cache()
        files_leg = run_files(args, torch, lib, abi, synth, filt)
        slots = None
    # ---- training leg (configs[3]); frees the encode buffers first
    training = training_long = training_fp32 = training_n1 = None
    if not args.no_train:
        del arena, arenas, slots, batches, h_events, h_songs
        torch.cuda.empty_cache()
        training = run_training(args, world, rank, torch, dist, precision='bf16')           # configs[3] as named: bf16 dense layers
        training_fp32 = run_training(args, world, rank, torch, dist, precision='fp32', e2e=False)
        if not args.no_long:
            training_long = run_training(args, world, rank, torch, dist, sh=LONG_SHAPE, N=args.long_batch, precision='bf16')
        if world == 1:
            training_n1 = run_training_n1(args, torch)
        t_load1 = time.time()
    clocks = sampler.stop(t_load0, t_load1) if sampler else None

    # ---- CPU baseline (rank 0, N=1): the oracle port, one thread, bounded sample
    cpu_baseline = None
    if rank == 0 and world == 1:
        sys.path.insert(0, os.path.join(ROOT, 'oracle'))
        import oracle_lib
        sample = 256
        ev_s, sg_s = host_batches[0][0], host_batches[0][1][:sample]
        t0 = time.perf_counter()
        infos, _ = oracle_lib.encode_batch(ev_s, sg_s, filt, 1)
        dt = time.perf_counter() - t0
        cpu_baseline = {'value': float(infos['n_bars_out'][infos['status'] == 0].sum()) / dt, 'unit': UNIT,
                        'cores': 1, 'kind': 'port',
                        'sample': f'first {sample} songs of batch 0 (same generator, same filter), one pass'}
        if training is not None:
            training['cpu_baseline'] = cpu_training_baseline(min(os.cpu_count() or 1, 32))

    if rank == 0:
        print(json.dumps({
            'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': elapsed_s / args.steps * 1e3, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64 quantise -> f32 grid', 'data': 'synthetic',
            'config': workload_config(args),
            'channel_bars_per_s': float(total_chan_bars.item()) / elapsed_s,
            'roofline': roofline, 'e2e': e2e, 'cpu_baseline': cpu_baseline, 'clocks': clocks,
            'gpu_launches': KERNELS_PER_STEP * args.steps +
            (training['gpu_launches_per_step'] * args.train_steps if training else 0) +
            (training_long['gpu_launches_per_step'] * args.train_steps if training_long else 0),
            'decode': decode, 'files': files_leg, 'training': training, 'training_fp32': training_fp32, 'training_long': training_long,
            'training_n1': training_n1,
            # compact digest, LAST so that a truncated stdout tail still carries the training half of the metric
            'summary': {
                'encode_song_bars_per_s': value, 'encode_e2e': e2e['value'], 'encode_e2e_events8': e2e['packed_events']['value'], 'encode_frac': roofline['frac'],
                'decode_frac': decode['roofline']['frac'],
                'files_song_bars_per_s': files_leg['value'] if files_leg else None,
                'files_cpu_song_bars_per_s': files_leg['cpu_baseline']['value'] if files_leg else None,
                'train_sps': training['value'] if training else None,
                'train_e2e_sps': training['e2e']['value'] if training else None,
                'train_ms': training['ms_per_step'] if training else None,
                'train_frac': training['roofline']['frac'] if training else None,
                'train_dtype': 'bf16' if training else None,
                'train_fp32_sps': training_fp32['value'] if training_fp32 else None,
                'long_sps': training_long['value'] if training_long else None,
                'n1_ms': training_n1['ms_per_iteration'] if training_n1 else None,
                'n1_dropin_ms': training_n1['dropin']['ms_per_iteration'] if training_n1 else None,
                'train_cpu_sps': (training.get('cpu_baseline') or {}).get('value') if training else None,
                'n_gpus': world},
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
