"""Chunk pipeline around the hot path: the reference's minibatch loop for one chunk
(core.py:156-273) with the chunk resident in HBM.

The reference keeps the chunk on the host, and per minibatch zero-fills a [max_len, batch, dim] tensor, copies every
sentence into it slice by slice behind a random number of leading zero rows (core.py:183-200), uploads it, and reads
the running loss back for its progress bar.  Here the chunk's [rows, dim] matrix is uploaded once, the offsets of ALL
minibatches are drawn up front -- same `random.randint` calls in the same order, nothing else in the loop consumes
Python's `random` -- and uploaded as three small tables; each minibatch is then one `hcgs_assemble_batch` launch on the
device followed by the graph step.  Loss / error sums stay on the device and are read once per chunk.

The forward (`to_do = forward`) half of the chunk loop is b200/export.py.
"""
import random

import numpy as np
import torch

from . import native


class ChunkPlan:
    """Host-side minibatch tables of one chunk: beg[i][k] (first chunk row), length[i][k], left[i][k] (leading zero
    rows) and max_len[i]; N_batches = int(N_snt / batch_size) like core.py:157-159 (trailing sentences are dropped)."""

    def __init__(self, data_end_index, batch_size, rng=random):
        end = np.asarray(data_end_index, dtype=np.int64).reshape(-1)
        if batch_size < 1:
            raise ValueError('batch_size must be >= 1')
        if end.size and (end[0] < 0 or np.any(np.diff(end) < 0)):
            raise ValueError('data_end_index must be non-decreasing')
        lens = np.diff(np.concatenate([[0], end]))
        self.batch_size = int(batch_size)
        self.n_batches = int(end.size / batch_size)
        n = self.n_batches * self.batch_size
        self.beg = (end[:n] - lens[:n]).reshape(self.n_batches, self.batch_size)
        self.length = lens[:n].astype(np.int32).reshape(self.n_batches, self.batch_size)
        self.max_len = self.length.max(axis=1).astype(np.int64) if n else np.zeros(0, np.int64)
        left = np.empty((self.n_batches, self.batch_size), dtype=np.int32)
        for i in range(self.n_batches):            # one randint per sentence, in sentence order (core.py:193)
            m = int(self.max_len[i])
            for k in range(self.batch_size):
                left[i, k] = rng.randint(0, m - int(self.length[i, k]))
        self.left = left
        self.rows = int(end[-1]) if end.size else 0

    def frames(self):
        return int(self.length.sum())


class ResidentChunk:
    """The chunk matrix and the minibatch tables in HBM; `minibatch(i)` assembles minibatch i on the device."""

    def __init__(self, data_set, plan, device):
        if isinstance(data_set, np.ndarray):
            data_set = torch.from_numpy(np.ascontiguousarray(data_set, dtype=np.float32))
        data_set = data_set.float().contiguous()
        if data_set.dim() != 2 or data_set.shape[0] < plan.rows:
            raise ValueError('data_set must be [rows >= %d, dim]' % plan.rows)
        self.plan = plan
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise native.NativeError('the chunk pipeline assembles minibatches on the GPU; there is no CPU path')
        if not data_set.is_cuda:
            data_set = data_set.pin_memory()
        self.data = data_set.to(self.device, non_blocking=True)
        self.beg = torch.from_numpy(plan.beg.copy()).to(self.device)
        self.length = torch.from_numpy(plan.length.copy()).to(self.device)
        self.left = torch.from_numpy(plan.left.copy()).to(self.device)

    def minibatch(self, i, out=None):
        return native.assemble_batch(self.data, self.beg[i], self.length[i], self.left[i], int(self.plan.max_len[i]), out)


def run_chunk(graph, optimizers, data_set, data_end_index, batch_size, to_do='train', device=None, frozen=(),
              rng=random):
    """core.py:156-273 for a sequential model, `to_do` in ('train', 'valid'): returns (loss_tot, err_tot) as Python
    floats = the means over the chunk's minibatches of outs['loss_final'] / outs['err_final'].
    graph: b200.graph.Graph; optimizers: name -> optimizer (train only); frozen: names with arch_freeze = True."""
    if to_do not in ('train', 'valid'):
        raise ValueError("run_chunk handles to_do = 'train' / 'valid'; posterior export is b200/export.py")
    if device is None:
        device = next(p for n in graph.nets.values() for p in n.parameters()).device
    plan = ChunkPlan(data_end_index, batch_size, rng)
    chunk = ResidentChunk(data_set, plan, device)
    loss_sum = torch.zeros((), device=chunk.device)
    err_sum = torch.zeros((), device=chunk.device)
    for i in range(plan.n_batches):
        inp = chunk.minibatch(i)
        max_len = int(plan.max_len[i])
        if to_do == 'train':
            outs = graph.forward(inp, max_len, plan.batch_size, 'train')
            for opt in optimizers.values():
                opt.zero_grad()
            outs['loss_final'].backward()
            for name, opt in optimizers.items():
                if name not in frozen:
                    opt.step()
        else:
            with torch.no_grad():
                outs = graph.forward(inp, max_len, plan.batch_size, 'valid')
        loss_sum += outs['loss_final'].detach()
        err_sum += outs['err_final'].detach()
    if plan.n_batches == 0:
        return float('nan'), float('nan')
    both = (torch.stack([loss_sum, err_sum]) / plan.n_batches).cpu()
    return float(both[0]), float(both[1])
