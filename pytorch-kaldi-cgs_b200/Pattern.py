"""8x8 "pattern" masks (reference: Pattern.py:25-133 and sparsity/sparsity.py:999-1049, 1112-1146).

Three things live here because the reference's LSTM/MLP reach them through two modules:
  * ``Pattern`` -- the mask module with the ``pattern`` / ``coo`` / ``pattern_coo`` modes;
  * ``find_top_k_by_kmeans`` -- mine ``pattern_num`` kernels from a weight matrix (k-means over the
    per-block top-``nnz`` indicator vectors);
  * ``apply_patterns`` -- per block pick the kernel with the largest sum |w|*pattern.
This is mask *generation* (host-side, run once and cached across chunks, core.py:129-131); the
masks it returns are then compiled into a block index like any other mask.

Parity notes: one ``np.random.choice(pattern_num, 1)[0]`` per block in row-major block
order from the global legacy RNG; freshly drawn patterns use ``choice(n, nnz)`` WITH replacement;
ties in ``apply_patterns`` (``out >= max``) add overlapping patterns, so mask values may exceed 1.
"""
import math
import os

import numpy as np
import torch
import torch.nn.functional as F
from torch.nn.modules.module import Module
from torch.nn.parameter import Parameter


def _blocks(shape, pshape):
    if shape[0] % pshape[0] != 0:
        raise AssertionError(f'Error:{shape[0]} can not be divisible by {pshape[0]}')
    if shape[1] % pshape[1] != 0:
        raise AssertionError(f'Error:{shape[1]} can not be divisible by {pshape[1]}')
    return shape[0] // pshape[0], shape[1] // pshape[1]


class Pattern(Module):
    def __init__(self, dense_features, pattern_mode, pattern_shape, pattern_nnz, pattern_num, save_dir,
                 pattern_from_file=False):
        super(Pattern, self).__init__()
        self.dense_features = dense_features
        self.pattern_from_file = pattern_from_file
        self.pattern_mode = pattern_mode
        self.pattern_shape = np.array(pattern_shape)
        self.pattern_nnz = int(pattern_nnz)
        self.pattern_num = int(pattern_num)
        self.mask = self.get_mask()

    def get_mask(self, dense=True):
        if self.pattern_mode == 'pattern':
            return self.pattern_mask()
        if self.pattern_mode == 'coo':
            return self.coo_mask()
        if self.pattern_mode == 'pattern_coo':
            return self.pattern_coo_mask()
        return Parameter(torch.ones(self.dense_features.shape))

    def get_pattern(self, pattern_nnz):
        if self.pattern_from_file:
            # the reference tests an undefined name here (Pattern.py:63-67); the intent is the file path
            path = self.pattern_from_file
            if not os.path.exists(path):
                raise FileNotFoundError('Pattern file "%s" does not exist' % path)
            return np.split(np.load(path), self.pattern_num, 0)
        n = int(np.prod(self.pattern_shape))
        out = []
        for _ in range(self.pattern_num):
            m = np.zeros(n)
            m[np.random.choice(m.shape[0], pattern_nnz)] = 1     # with replacement, like the reference
            out.append(m.reshape(self.pattern_shape))
        return out

    def _dense_np(self):
        d = self.dense_features
        return d.detach().cpu().numpy() if torch.is_tensor(d) else np.asarray(d)

    def _fill(self, block_fn):
        shape = tuple(self.dense_features.shape)
        ps = self.pattern_shape
        nbr, nbc = _blocks(shape, ps)
        mask = np.zeros(shape)
        for i in range(nbr):
            for j in range(nbc):
                sl = (slice(i * ps[0], (i + 1) * ps[0]), slice(j * ps[1], (j + 1) * ps[1]))
                mask[sl] = block_fn(sl)
        self.mask = mask
        return Parameter(torch.from_numpy(mask))

    def pattern_mask(self):
        self.pattern = self.get_pattern(self.pattern_nnz)
        return self._fill(lambda sl: self.pattern[np.random.choice(self.pattern_num, 1)[0]])

    def coo_mask(self):
        dense = self._dense_np()

        def top(sl):
            blk = dense[sl].flatten()
            m = np.zeros(blk.shape)
            m[np.argsort(np.abs(blk))[-self.pattern_nnz:]] = 1
            return m.reshape(self.pattern_shape)
        return self._fill(top)

    def pattern_coo_mask(self):
        dense = self._dense_np()
        self.pat_nnz = math.ceil(self.pattern_nnz / 2)
        self.coo_nnz = self.pattern_nnz - self.pat_nnz
        self.pattern = self.get_pattern(self.pat_nnz)

        def mix(sl):
            # NB: like the reference, flatten() of a contiguous pattern is a copy, the stored
            # pattern is not modified
            m = self.pattern[np.random.choice(self.pattern_num, 1)[0]].flatten()
            blk = dense[sl].flatten()
            m[np.argsort(np.abs(np.array(blk) * (np.ones_like(m) - m)))[-self.coo_nnz:]] = 1
            return m.reshape(self.pattern_shape)
        return self._fill(mix)


def comb_num(n, m):
    return math.factorial(n) // (math.factorial(m) * math.factorial(n - m))


def _kernel_device():
    return torch.device('cuda') if torch.cuda.is_available() else torch.device('cpu')


def find_top_k_by_kmeans(raw_w, pattern_num, pattern_shape, pattern_nnz, stride):
    """sparsity/sparsity.py:999-1049: returns a [pattern_num, 1, ph, pw] 0/1 kernel stack."""
    from sklearn.cluster import KMeans
    ph, pw = int(pattern_shape[0]), int(pattern_shape[1])
    p_num_x = (raw_w.size(0) - ph) // stride[0] + 1
    p_num_y = (raw_w.size(1) - pw) // stride[1] + 1
    pattern_num = min(pattern_num, comb_num(ph * pw, pattern_nnz))
    w = torch.abs(raw_w.detach()).cpu()
    if w.dim() == 2:
        w = w.unsqueeze(2)
    cands = []
    for k in range(w.size(2)):
        for i in range(p_num_x):
            for j in range(p_num_y):
                sub = w[i * stride[0]: i * stride[0] + ph, j * stride[1]: j * stride[1] + pw, k]
                value, _ = torch.topk(sub.abs().flatten(), pattern_nnz)
                cand = torch.where(abs(sub) < value[-1], torch.zeros_like(sub), torch.ones_like(sub))
                cands.append(cand.numpy().flatten())
    clf = KMeans(n_clusters=pattern_num)
    clf.fit(cands)
    kernel = torch.zeros((len(clf.cluster_centers_), 1, ph, pw))
    for p_i, center in enumerate(clf.cluster_centers_):
        c = torch.from_numpy(center)
        index = c.sort()[1][-pattern_nnz:]
        pat = torch.zeros_like(c)
        pat[index] = 1
        kernel[p_i, 0] = pat.reshape(ph, pw)
    return kernel.to(_kernel_device())


def apply_patterns(raw_w, kernel):
    """sparsity/sparsity.py:1112-1146: per block keep the pattern(s) maximising sum |w|*pattern."""
    w = torch.abs(raw_w.detach()).to(kernel.device)
    ph, pw = kernel.size(2), kernel.size(3)
    stride = (ph, pw)
    unsq = w.dim() == 2
    if unsq:
        w = w.unsqueeze(2)
    w = w.permute(2, 0, 1)
    # fp32 scores: a TF32 convolution (the cuDNN default on the device) could move the per-block arg-max
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        out = F.conv2d(w.unsqueeze(1), kernel, stride=stride, padding=0)
    out_max = torch.max(out, dim=1)[0].unsqueeze(1).repeat(1, kernel.size(0), 1, 1)
    idx = torch.where(out >= out_max, torch.ones_like(out), torch.zeros_like(out))
    with torch.backends.cudnn.flags(enabled=True, allow_tf32=False):
        mask = F.conv_transpose2d(idx, kernel, bias=None, stride=stride, padding=0, output_padding=0, groups=1)
    mask = mask.squeeze(1).permute(1, 2, 0)
    if unsq:
        mask = mask.squeeze(2)
    return mask
