"""Multi-rank host logic on CPU: two gloo processes exercise the data-parallel plumbing of b200/dist.py
(compressed-gradient all-reduce over the kept weights, utterance sharding for the posterior export)."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for _p in (os.path.join(ROOT, 'pytorch-kaldi-cgs_b200'), ROOT, HERE):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from b200 import dist as bdist
    from helpers import build_product_nets
    r, w, _ = bdist.init_from_env('gloo')
    assert (r, w) == (rank, world) and dist.is_initialized()
    nets, _ = build_product_nets('train', D=12, seed=2234, device='cpu')   # same seed: identical masks on all ranks
    bucket = bdist.GradBucket(bdist.lstm_mlp_entries(list(nets.values())), torch.device('cpu'))
    sig = torch.tensor([bucket.mask_signature()], dtype=torch.int64)
    sigs = [torch.zeros_like(sig) for _ in range(world)]
    dist.all_gather(sigs, sig)
    assert all(int(s) == int(sig) for s in sigs), 'mask index differs across ranks'
    # rank-dependent gradients; masked-out entries carry garbage that must neither travel nor change
    g = torch.Generator().manual_seed(100 + rank)
    params = [p for n in nets.values() for name, p in n.named_parameters() if not name.endswith('mask')]
    for p in params:
        p.grad = torch.randn(p.shape, generator=g)
    before = [p.grad.clone() for p in params]
    bucket.all_reduce_mean()
    torch.save({'after': [p.grad.clone() for p in params], 'before': before, 'numel': bucket.numel},
               os.path.join(out_dir, 'rank%d.pt' % rank))
    # overlapped per-group buckets: hooks fire during a real backward; same result as the single bucket
    for p in params:
        p.grad = None
    ob = bdist.OverlappedGradBuckets(bdist.lstm_mlp_groups(list(nets.values())), torch.device('cpu'))
    assert ob.numel == bucket.numel
    trainable = [p for p in params if p.requires_grad]
    fired_early = []
    for step in range(2):
        for p in trainable:
            p.grad = None
        loss = sum((p * float(rank + 1 + step)).sum() for p in trainable[::2])     # every other parameter gets a gradient
        loss.backward()
        fired_early.append(sum(ob.fired))
        ob.all_reduce_mean()
    torch.save({'grads': [None if p.grad is None else p.grad.clone() for p in trainable], 'fired_early': fired_early},
               os.path.join(out_dir, 'overlap%d.pt' % rank))
    ob.remove_hooks()
    # utterance sharding: every rank computes the same partition
    lengths = [int(v) for v in np.random.RandomState(3).randint(50, 900, size=37)]
    shards = bdist.shard_utterances(lengths, world)
    torch.save(shards, os.path.join(out_dir, 'shards%d.pt' % rank))
    dist.barrier()
    dist.destroy_process_group()


def _mismatch_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from b200 import dist as bdist
    from helpers import build_product_nets
    bdist.init_from_env('gloo')
    nets, _ = build_product_nets('train', D=12, seed=2234 + rank, device='cpu')     # per-rank seed: masks differ
    raised = []
    for make in (lambda: bdist.GradBucket(bdist.lstm_mlp_entries(list(nets.values())), torch.device('cpu')),
                 lambda: bdist.OverlappedGradBuckets(bdist.lstm_mlp_groups(list(nets.values())), torch.device('cpu'))):
        try:
            make()
            raised.append(False)
        except RuntimeError as e:
            raised.append('masks differ across ranks' in str(e))
    torch.save(raised, os.path.join(out_dir, 'mismatch%d.pt' % rank))
    dist.barrier()
    dist.destroy_process_group()


def test_mask_mismatch_across_ranks_is_refused_world2(tmp_path):
    """the compressed all-reduce must not start when the kept-entry index differs between ranks."""
    world = 2
    mp.spawn(_mismatch_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert torch.load(os.path.join(str(tmp_path), 'mismatch%d.pt' % r)) == [True, True]


def test_grad_bucket_allreduce_and_sharding_world2(tmp_path):
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    res = [torch.load(os.path.join(str(tmp_path), 'rank%d.pt' % r)) for r in range(world)]
    from b200 import dist as bdist
    from helpers import build_product_nets
    nets, _ = build_product_nets('train', D=12, seed=2234, device='cpu')
    entries = [(p, m) for p, m in bdist.lstm_mlp_entries(list(nets.values()))]
    total_kept = 0
    dense_total = 0
    for i, (p, m) in enumerate(entries):
        mean = (res[0]['before'][i] + res[1]['before'][i]) / 2
        for r in range(world):
            got = res[r]['after'][i]
            if m is None:
                assert torch.allclose(got, mean, atol=1e-6)
            else:
                keep = m.detach() != 0
                assert torch.allclose(got[keep], mean[keep], atol=1e-6)
                assert torch.equal(got[~keep], res[r]['before'][i][~keep]), 'masked-out entries must not be touched'
        total_kept += int((m.detach() != 0).sum()) if m is not None else p.numel()
        dense_total += p.numel()
    assert res[0]['numel'] == total_kept < dense_total      # only the kept weights travel
    # overlapped buckets, second step: gradient (rank + 2) on each rank -> mean 2.5 on kept entries, own value elsewhere
    ov = [torch.load(os.path.join(str(tmp_path), 'overlap%d.pt' % r)) for r in range(world)]
    trainable = [(p, m) for p, m in entries if p.requires_grad]
    for r in range(world):
        assert ov[r]['fired_early'][0] <= ov[r]['fired_early'][1] and ov[r]['fired_early'][1] >= 1, ov[r]['fired_early']
        for i, (p, m) in enumerate(trainable):
            got = ov[r]['grads'][i]
            if i % 2:
                assert got is None
                continue
            if m is None:
                assert torch.allclose(got, torch.full_like(got, 2.5))
            else:
                keep = m.detach() != 0
                assert torch.allclose(got[keep], torch.full_like(got[keep], 2.5))
                assert torch.allclose(got[~keep], torch.full_like(got[~keep], float(r + 2)))
    s0 = torch.load(os.path.join(str(tmp_path), 'shards0.pt'))
    s1 = torch.load(os.path.join(str(tmp_path), 'shards1.pt'))
    assert s0 == s1
    flat = sorted(i for s in s0 for i in s)
    assert flat == list(range(37))
    lengths = [int(v) for v in np.random.RandomState(3).randint(50, 900, size=37)]
    loads = [sum(lengths[i] for i in s) for s in s0]
    assert max(loads) - min(loads) <= max(lengths)          # length-balanced
