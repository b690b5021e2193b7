"""Host-fabric ceiling for the posterior export: aggregate device -> pinned-host copy bandwidth with 1, 2, 4 and 8 GPUs
of one box copying at the same time (the export's end-to-end number is D2H of fp32 posteriors, 7.8 KB per frame).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 \
      tools/d2h_ceiling.py

Every rank owns one GPU and one pinned buffer; in phase n only ranks < n copy (256 MiB x 12, timed with CUDA events
between two barriers), the others wait.  Rank 0 prints one JSON line."""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
n_bytes, reps = 256 << 20, 12
src = torch.empty(n_bytes, dtype=torch.uint8, device='cuda').random_(0, 255)
dst = torch.empty(n_bytes, dtype=torch.uint8).pin_memory()
out = {'bytes_per_copy': n_bytes, 'copies': reps, 'cpus': os.cpu_count()}


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()


for direction in ('d2h', 'h2d'):
    n = 1
    while n <= world:
        active = rank < n
        for _ in range(2):
            if active:
                (dst.copy_(src, non_blocking=True) if direction == 'd2h' else src.copy_(dst, non_blocking=True))
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if active:
            for _ in range(reps):
                (dst.copy_(src, non_blocking=True) if direction == 'd2h' else src.copy_(dst, non_blocking=True))
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) if active else 0.0], device='cuda')
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        barrier()
        out['%s_%dgpu_GBps' % (direction, n)] = round(n * reps * n_bytes / (ms.item() * 1e-3) / 1e9, 1)
        n *= 2
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
