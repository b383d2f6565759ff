"""Host -> device bandwidth of every rank while ALL ranks copy at once (development probe for the multi-GPU host-buffer path).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29519 tools/h2d_probe.py
Prints per rank: GPU PCI address, its NUMA node as sysfs reports it, the CPUs the process may run on, and GB/s for 137 MB and
274 MB pinned copies, first with the process where the launcher put it, then after binding it to the GPU's NUMA node and
re-allocating the pinned buffer there (first touch)."""
import os
import time

import torch
import torch.distributed as dist


def node_cpus(node):
    try:
        txt = open(f'/sys/devices/system/node/node{node}/cpulist').read().strip()
    except OSError:
        return None
    cpus = set()
    for part in txt.split(','):
        a, _, b = part.partition('-')
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def measure(buf_host, buf_dev, reps, stream):
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    with torch.cuda.stream(stream):
        for _ in range(reps):
            buf_dev.copy_(buf_host, non_blocking=True)
    stream.synchronize()
    dt = time.perf_counter() - t0
    return buf_host.numel() * reps / dt / 1e9


def main():
    rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    props = torch.cuda.get_device_properties(local)
    bdf = f'{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0'
    try:
        node = int(open(f'/sys/bus/pci/devices/{bdf}/numa_node').read())
    except OSError:
        node = -2
    stream = torch.cuda.Stream()
    out = [f'rank {rank} gpu {bdf} numa_node {node} affinity {sorted(os.sched_getaffinity(0))[:4]}..({len(os.sched_getaffinity(0))})']
    for phase in ('as launched', 'bound to the GPU node'):
        if phase != 'as launched':
            cpus = node_cpus(node) if node >= 0 else None
            if not cpus:
                out.append('  no NUMA node to bind to')
                break
            os.sched_setaffinity(0, cpus & os.sched_getaffinity(0) or cpus)
        for mb in (137, 274):
            h = torch.empty(mb << 20, dtype=torch.uint8)
            h.fill_(1)                                  # first touch on the current node
            h = h.pin_memory()
            d = torch.empty_like(h, device='cuda')
            measure(h, d, 3, stream)
            out.append(f'  {phase}: {mb} MB x 20 copies, all ranks at once: {measure(h, d, 20, stream):.1f} GB/s')
            del h, d
    gathered = [None] * world
    dist.all_gather_object(gathered, '\n'.join(out))
    if rank == 0:
        print('\n'.join(gathered), flush=True)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
