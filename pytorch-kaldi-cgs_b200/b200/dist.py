"""Data-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink on the GPU box,
gloo in the CPU tests).

The reference has no working multi-GPU path (utils.py:1783-1784 wraps the nets in nn.DataParallel,
which core.py:123 then breaks).  The B200 design:
  * training  -- rank r runs its own minibatches; after backward ONE all-reduce averages a single
    contiguous buffer that holds the gradients of every trainable parameter (weights only where
    the HCGS mask keeps them: masked-out entries are structurally zero on every rank and are not
    sent); BN statistics stay per rank like in the reference (no sync-BN).
  * export    -- utterances are sharded by length-balanced round robin, no collective.
"""
import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """torchrun-style rendezvous (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT)."""
    import os
    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = 'nccl' if torch.cuda.is_available() else 'gloo'
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


class GradBucket:
    """Flat gradient buffer over the kept entries of all trainable parameters.

    ``entries`` = list of (param, mask_or_None).  For a masked weight only the indices where the mask
    is non-zero travel; everything else is copied whole.  The index lists are built once.
    """

    def __init__(self, entries, device, check=True):
        self.items = []
        off = 0
        for p, m in entries:
            if not p.requires_grad:
                continue
            if m is not None:
                idx = torch.nonzero(m.detach().reshape(-1) != 0, as_tuple=False).reshape(-1).to(device)
                n = idx.numel()
            else:
                idx, n = None, p.numel()
            self.items.append((p, idx, off, n))
            off += n
        self.numel = off
        self.flat = torch.zeros(off, device=device, dtype=torch.float32)
        if check:
            assert_same_masks(self.mask_signature(), device)

    def mask_signature(self):
        """Hash of the whole kept-entry index (every index, position weighted).  The compressed all-reduce sums
        buffer slot k of every rank: if the HCGS masks differed across ranks (per-rank numpy seed, another
        checkpoint) gradients of DIFFERENT weights would be added silently -- hence assert_same_masks()."""
        s = 0
        for _, idx, off, n in self.items:
            h = 0
            if idx is not None and n:
                wgt = (torch.arange(n, device=idx.device, dtype=torch.int64) % 8191) + 1
                h = int((idx.to(torch.int64) * wgt).sum().item()) % (2 ** 61 - 1)
            s = (s * 1000003 + n * 31 + h) % (2 ** 61 - 1)
        return s

    def gather(self):
        for p, idx, off, n in self.items:
            if p.grad is None:
                self.flat[off:off + n].zero_()
            elif idx is None:
                self.flat[off:off + n].copy_(p.grad.reshape(-1))
            else:
                torch.index_select(p.grad.reshape(-1), 0, idx, out=self.flat[off:off + n])

    def scatter(self):
        for p, idx, off, n in self.items:
            if p.grad is None:
                continue
            if idx is None:
                p.grad.reshape(-1).copy_(self.flat[off:off + n])
            else:
                p.grad.reshape(-1).index_copy_(0, idx, self.flat[off:off + n])

    def all_reduce_mean(self):
        """gradient <- mean over ranks (what one process would compute on the concatenated batch
        when every rank's loss is a mean over its own frames)."""
        if not dist.is_initialized() or dist.get_world_size() == 1:
            return
        self.gather()
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        self.flat.div_(dist.get_world_size())
        self.scatter()


def assert_same_masks(signature, device):
    """All ranks must hold the same kept-entry index: all-reduce MIN and MAX of the signature, raise on a mismatch
    (on every rank, so that no rank is left waiting in a collective)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    t = torch.tensor([signature, -signature], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    lo, hi = int(t[0].item()), -int(t[1].item())
    if lo != hi:
        raise RuntimeError('HCGS masks differ across ranks (kept-entry index signature %d on this rank, min %d / max %d '
                           'over ranks): seed numpy identically on every rank before constructing the nets, or load the '
                           'same checkpoint' % (signature, lo, hi))


def lstm_mlp_entries(nets):
    """(param, mask) pairs for LSTM / MLP plugin instances."""
    out = []
    for net in nets:
        masks = {}
        if hasattr(net, 'hcgsx') and getattr(net, 'lstm_hcgs', False):
            for i in range(net.N_lstm_lay):
                for g in 'fioc':
                    masks['w%sx.%d.weight' % (g, i)] = net.hcgsx[i].mask
                    masks['u%sh.%d.weight' % (g, i)] = net.hcgsh[i].mask
        if hasattr(net, 'hcgs') and getattr(net, 'mlp_hcgs', False):
            for i in range(net.N_dnn_lay):
                masks['wx.%d.weight' % i] = net.hcgs[i].mask
        for name, p in net.named_parameters():
            if name.endswith('mask'):
                continue            # masks never receive a gradient
            out.append((p, masks.get(name)))
    return out


def lstm_mlp_groups(nets):
    """The entries of lstm_mlp_entries cut into groups whose gradients become ready together during backward: one
    group per LSTM layer (input-norm parameters ride with layer 0) and one per MLP net, in the order in which
    backward completes them (heads first, then the LSTM layers from the top)."""
    groups = []
    for net in nets:
        entries = lstm_mlp_entries([net])
        names = [n for n, _ in net.named_parameters() if not n.endswith('mask')]
        assert len(names) == len(entries)
        if hasattr(net, 'N_lstm_lay'):
            per = [[] for _ in range(net.N_lstm_lay)]
            for n, e in zip(names, entries):
                parts = n.split('.')
                i = int(parts[1]) if len(parts) > 2 and parts[1].isdigit() and int(parts[1]) < net.N_lstm_lay else 0
                per[i].append(e)
            groups = [g for g in reversed(per) if g] + groups        # LSTM layers after whatever sits above them
        else:
            groups = [entries] + groups
    return [g for g in groups if g]


class OverlappedGradBuckets:
    """Per-group GradBuckets whose all-reduce overlaps the rest of backward: an autograd hook counts the gradients
    of a group as they are accumulated and launches the group's gather / all-reduce / scatter on a side stream when
    the last one has arrived; ``all_reduce_mean()`` (called where the single-bucket version does its whole work,
    after backward) only launches what has not fired and makes the main stream wait.  Same result as one GradBucket."""

    def __init__(self, groups, device):
        self.buckets = [GradBucket(g, device, check=False) for g in groups]
        self.groups = [[p for p, _ in g if p.requires_grad] for g in groups]
        self.calibrated = False
        self.total = [len(g) for g in self.groups]
        self.left = list(self.total)
        self.fired = [False] * len(groups)
        self.numel = sum(b.numel for b in self.buckets)
        self.stream = torch.cuda.Stream(device=device) if device.type == 'cuda' else None
        self.handles = []
        assert_same_masks(self.mask_signature(), device)
        for bi, g in enumerate(groups):
            for p, _ in g:
                if p.requires_grad:
                    self.handles.append(p.register_post_accumulate_grad_hook(lambda _p, bi=bi: self._ready(bi)))

    def mask_signature(self):
        s = 0
        for b in self.buckets:
            s = (s * 1000003 + b.mask_signature()) % (2 ** 61 - 1)
        return s

    def _ready(self, bi):
        self.left[bi] -= 1
        if self.left[bi] == 0 and not self.fired[bi]:
            self._launch(bi)

    def _launch(self, bi):
        self.fired[bi] = True
        if not dist.is_initialized() or dist.get_world_size() == 1:
            return
        if self.stream is None:
            self.buckets[bi].all_reduce_mean()
            return
        self.stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self.stream):
            self.buckets[bi].all_reduce_mean()

    def all_reduce_mean(self):
        for bi in range(len(self.buckets)):
            if not self.fired[bi]:
                self._launch(bi)
        if self.stream is not None:
            torch.cuda.current_stream().wait_stream(self.stream)
        if not self.calibrated:
            # parameters of unused options (e.g. the LayerNorm / BatchNorm modules the classes always create) never
            # receive a gradient: count only those that did in the first step, or no group would ever fire early
            self.total = [sum(1 for p in g if p.grad is not None) for g in self.groups]
            self.calibrated = True
        self.left = list(self.total)
        self.fired = [False] * len(self.buckets)

    def remove_hooks(self):
        for h in self.handles:
            h.remove()
        self.handles = []


def shard_utterances(lengths, world):
    """Length-balanced round robin: sort by length (descending), deal to the least-loaded rank."""
    order = sorted(range(len(lengths)), key=lambda i: -lengths[i])
    load = [0] * world
    shards = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: load[k])
        shards[r].append(i)
        load[r] += lengths[i]
    return [sorted(s) for s in shards]
