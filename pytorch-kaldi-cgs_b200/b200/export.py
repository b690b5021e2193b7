"""Posterior export (forward pass that feeds the Kaldi decoder).

Reference: core.run_nn with to_do='forward' (core.py:183-249) forwards ONE utterance per step
(batch_size = 1, :186), subtracts ``log(counts/sum(counts))`` on the host (:241-245) and appends
a Kaldi matrix record per utterance to ``*_to_decode.ark`` (data_io.write_mat, :770-806).

Here many utterances share one launch: they are sorted by length, cut into groups of ``batch``
and RIGHT-padded with zero frames to the group's longest utterance.  In forward mode the nets
are causal and stateless across sequences (eval batch norm = running statistics, dropout = the
scalar (1-p)), so frames ``t < len`` of every sequence are exactly what the batch_size=1 pass
computes; the padded tail is dropped on the device (``hcgs_unpad_rows`` packs every utterance's rows into one
contiguous block), so the host copy moves only what is written.  The prior is subtracted inside the fused
output kernel (``hcgs_out_logsoftmax_prior_fwd``); device->host copies go through two pinned
buffers on a side stream so that the copy of group i overlaps the compute of group i+1.
Multi-GPU: ``dist.shard_utterances`` deals the utterances to the ranks, no collective; every rank
writes its own ark (the decoder globs ``*_to_decode.ark``, run_exp.py:459-460).
"""
import numpy as np
import torch

import posterior_io

from . import dist as bdist
from . import native


def plan_groups(lengths, batch, max_frames=None):
    """Indices grouped for one launch each: ascending length (like data_io.py:78 sorts the chunk), ``batch``
    utterances per group, optionally capped at ``max_frames`` padded frames per group."""
    order = sorted(range(len(lengths)), key=lambda i: (lengths[i], i))
    groups, cur = [], []
    for i in order:
        t_max = lengths[i]            # ascending: the newest utterance is the longest of the group
        if cur and (len(cur) >= batch or (max_frames and (len(cur) + 1) * t_max > max_frames)):
            groups.append(cur)
            cur = []
        cur.append(i)
    if cur:
        groups.append(cur)
    return groups


def pack_group(feats, idx, out=None):
    """[T_max, B, D] right-padded minibatch of the utterances ``idx`` (host tensor).  ``out``: a flat (pinned)
    float32 buffer to build it in."""
    t_max = max(int(feats[i].shape[0]) for i in idx)
    d = int(feats[idx[0]].shape[1])
    if out is None:
        x = torch.zeros(t_max, len(idx), d)
    else:
        x = out[:t_max * len(idx) * d].view(t_max, len(idx), d)
        x.zero_()
    for b, i in enumerate(idx):
        x[:feats[i].shape[0], b] = torch.as_tensor(feats[i])
    return x


def _pinned(buf, numel):
    """grow-only pinned float32 buffer (cudaHostAlloc is slow: allocate once, reuse across groups and runs)"""
    if buf is None or buf.numel() < numel:
        buf = torch.empty(numel, dtype=torch.float32, pin_memory=True)
    return buf


class PosteriorExporter:
    """forward_fn(x[T,B,D] cuda) -> log-posteriors [T*B, P] (prior already subtracted when ``log_prior`` was
    handed to the output net).  ``run`` yields (utterance index, float32 ndarray [T_i, P]) in group order
    (ascending length).  The arrays are contiguous VIEWS of one of two pinned buffers: consume (write) each one
    before asking for the matrices of the group after the next."""

    def __init__(self, forward_fn, device, batch=512, max_frames=None):
        self.forward_fn, self.device, self.batch, self.max_frames = forward_fn, device, batch, max_frames
        self.copy_stream = torch.cuda.Stream(device=device)
        self.frames = self.h2d_bytes = self.d2h_bytes = 0
        self.host_in = [None, None]
        self.host_out = [None, None]
        self.in_free = [None, None]

    def run(self, feats):
        lengths = [int(f.shape[0]) for f in feats]
        groups = plan_groups(lengths, self.batch, self.max_frames)
        main = torch.cuda.current_stream(self.device)
        host = self.host_out
        done = [torch.cuda.Event(), torch.cuda.Event()]
        # per-group tables for hcgs_unpad_rows (first packed row and length of every utterance), one upload for the run
        offs, lens_g = [], []
        for idx in groups:
            ln = np.asarray([lengths[i] for i in idx], dtype=np.int64)
            offs.append(np.concatenate([[0], np.cumsum(ln)[:-1]]) if len(idx) else ln)
            lens_g.append(ln.astype(np.int32))
        if groups:
            d_off = torch.from_numpy(np.concatenate(offs).astype(np.int64)).to(self.device)
            d_len = torch.from_numpy(np.concatenate(lens_g)).to(self.device)
        first = 0
        pending = None
        for g, idx in enumerate(groups):
            slot = g % 2
            t_g = max(lengths[i] for i in idx)
            self.host_in[slot] = _pinned(self.host_in[slot], t_g * len(idx) * int(feats[idx[0]].shape[1]))
            if self.in_free[slot] is not None:
                self.in_free[slot].synchronize()        # the upload that last used this staging buffer has finished
            x = pack_group(feats, idx, self.host_in[slot]).to(self.device, non_blocking=True)
            self.in_free[slot] = torch.cuda.Event()
            self.in_free[slot].record(main)
            with torch.no_grad():
                post = self.forward_fn(x)
            t_max, b = x.shape[0], x.shape[1]
            rows = int(lens_g[g].sum())
            # every utterance one contiguous [len, P] block, padding rows dropped, in ONE pass on the device: the host copy
            # then moves exactly the matrices that get written
            post2d = post.view(t_max * b, -1)
            packed = torch.empty(rows, post2d.shape[1], device=self.device, dtype=torch.float32)
            native.unpad_rows(post2d, d_off[first:first + b], d_len[first:first + b], t_max, packed)
            first += b
            host[slot] = _pinned(host[slot], packed.numel())
            dst = host[slot][:packed.numel()].view(packed.shape)
            self.copy_stream.wait_stream(main)
            with torch.cuda.stream(self.copy_stream):
                dst.copy_(packed, non_blocking=True)
                done[slot].record(self.copy_stream)
            packed.record_stream(self.copy_stream)
            self.d2h_bytes += packed.numel() * 4
            self.h2d_bytes += x.numel() * 4
            if pending is not None:
                yield from self._drain(*pending)
            pending = (dst, done[slot], idx, lengths, offs[g])
        if pending is not None:
            yield from self._drain(*pending)

    def _drain(self, buf, ev, idx, lengths, off):
        ev.synchronize()
        arr = buf.numpy()
        for b, i in enumerate(idx):
            self.frames += lengths[i]
            yield i, arr[int(off[b]):int(off[b]) + lengths[i]]


def export_ark(forward_fn, feats, keys, ark_path, device, batch=512, rank=0, world=1):
    """Write this rank's share of the utterances as Kaldi float-matrix records (posterior_io.mat_record);
    returns the number of frames written.  Records appear in ascending utterance length, the order in which
    the reference's chunk is sorted (data_io.py:78); an ark is keyed, the decoder does not depend on the order."""
    lengths = [int(f.shape[0]) for f in feats]
    mine = bdist.shard_utterances(lengths, world)[rank] if world > 1 else list(range(len(feats)))
    ex = PosteriorExporter(forward_fn, device, batch)
    with open(ark_path, 'wb') as fd:
        for j, m in ex.run([feats[i] for i in mine]):
            fd.write(posterior_io.mat_record(m, keys[mine[j]]))
    return ex.frames
