"""CPU-drawn dropout masks one forward ahead, bit-identical to drawing them on demand.

The reference draws one mask per recurrent layer and forward on the CPU generator
(``torch.bernoulli(torch.Tensor(B, H).fill_(1 - drop))``, neural_networks.py:843-851): 1.8 ms of host time per layer at
B = 528, H = 550, i.e. 7 of the 19 ms the host needs to enqueue a train step.  The masks only depend on the state of the
default CPU generator, so a worker thread can draw the masks of the NEXT forward from a *copy* of that state while the
host enqueues the rest of the step.  At the next forward the copy is valid if and only if the default generator is still
in the state the copy was taken from (nobody seeded it or drew from it in between); then the masks are used and the
default generator is advanced to the state the worker ended in -- exactly what drawing on demand would have done.
Otherwise the prefetched masks are dropped and the masks are drawn on demand.
"""
import threading
from concurrent.futures import ThreadPoolExecutor

import torch

_pool = None
_lock = threading.Lock()
_BERNOULLI = torch.bernoulli   # tests replay recorded masks by patching torch.bernoulli: then nothing is prefetched


def _executor():
    global _pool
    with _lock:
        if _pool is None:
            _pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix='hcgs-dropmask')
    return _pool


def _draw(state, spec):
    gen = torch.Generator(device='cpu')
    gen.set_state(state)
    masks = [_BERNOULLI(torch.Tensor(b, h).fill_(p), generator=gen) for (b, h, p) in spec]
    return masks, gen.get_state()


class DropMaskPrefetcher:
    """One per module.  ``take(spec)`` returns the masks for ``spec = [(B, H, keep_prob), ...]`` (one per layer, in
    drawing order) and leaves the default generator as if they had been drawn one by one; ``prefetch(spec)`` starts
    drawing the masks of the next call."""

    def __init__(self):
        self._pending = None   # (spec, state the copy was taken from, future)
        self.hits = self.misses = 0

    def take(self, spec):
        pend, self._pending = self._pending, None
        if pend is not None:
            pspec, state0, fut = pend
            try:
                masks, state1 = fut.result()
            except Exception:   # the worker is an optimisation only
                masks = None
            if (masks is not None and torch.bernoulli is _BERNOULLI and pspec == spec and
                    torch.equal(torch.get_rng_state(), state0)):
                torch.set_rng_state(state1)
                self.hits += 1
                return masks
        self.misses += 1
        return [torch.bernoulli(torch.Tensor(b, h).fill_(p)) for (b, h, p) in spec]

    def prefetch(self, spec):
        if torch.bernoulli is not _BERNOULLI:
            return
        state0 = torch.get_rng_state()
        self._pending = (spec, state0, _executor().submit(_draw, state0.clone(), list(spec)))
