"""Guided (magnitude-driven) HCGS masks (reference: guided_hcgs.py:9-76, guided_cgs_base.py,
guided_choices.py).  Same block structure as hcgs.conn_mat, but each block-row keeps the
column blocks with the largest mean |W| instead of a random draw; all levels use the
integer block count (no float quirk at the top level, guided_hcgs.py:19-27)."""
import os

import numpy as np
import torch

import cgs_base
from hcgs import mask_device


def conn_mat_numpy(n_in, n_out, block_sizes, drop_ratios, w_mat):
    levels = cgs_base.levels_from_lists(list(block_sizes), list(drop_ratios))
    w_abs = torch.as_tensor(w_mat).detach().abs().to('cpu', torch.float32)
    return cgs_base.fill_levels(int(n_in), int(n_out), levels, top_level_float_count=False, w_abs=w_abs)


def conn_mat_device(n_in, n_out, block_sizes, drop_ratios, w_mat):
    """The same mask with the weights left on the GPU.  The reference copies every block-row strip of |W| to
    the host and pools it there, strip by strip and level by level (guided_choices.py:4-30: ~740 device->host copies and
    AvgPool2d calls for one 550 x 550 matrix).  Here a level is ONE launch: every pooling window of every kept block gets a
    thread that adds |w| in the CPU kernel's order (hcgs_window_abs_mean, bit-identical means), the few hundred means come
    back in one copy, the ranking is the reference's own np.argsort, and the kept blocks of the last level are written as
    ones on the device.  Returns a float32 CUDA tensor [n_in, n_out]."""
    from b200 import native
    levels = cgs_base.levels_from_lists(list(block_sizes), list(drop_ratios))
    w = w_mat.detach()
    if w.dtype != torch.float32 or w.stride(-1) != 1:
        w = w.float().contiguous()
    n_in, n_out = int(n_in), int(n_out)
    rects = np.array([[0, n_in, 0, n_out]], dtype=np.int32)
    for bs, drop in levels:
        keep = 1 - float(drop) / 100
        wins, shape = [], []
        for r0, r1, c0, c1 in rects:
            if c1 - c0 < bs:
                # the reference's AvgPool2d raises for a window wider than the strip: same error from the same call
                return torch.from_numpy(conn_mat_numpy(n_in, n_out, block_sizes, drop_ratios, w.cpu())).to(w.device)
            rs = np.arange(r0, r1, bs, dtype=np.int32)
            cs = np.arange(c0, c1, bs, dtype=np.int32)
            g = np.empty((len(rs), len(cs), 4), dtype=np.int32)
            g[:, :, 0] = rs[:, None]
            g[:, :, 1] = np.minimum(rs + bs, r1)[:, None]
            g[:, :, 2] = cs[None, :]
            g[:, :, 3] = np.minimum(cs + bs, c1)[None, :]
            wins.append(g.reshape(-1, 4))
            shape.append((len(rs), len(cs)))
        wins = np.concatenate(wins, 0)
        means = native.window_abs_mean(w, torch.from_numpy(wins).to(w.device)).cpu().numpy()
        nxt, pos = [], 0
        for nr, nc in shape:
            n_sel = int(round(nc * keep))
            k = means[pos:pos + nr * nc].reshape(nr, nc)
            for i in range(nr):
                for c in np.argsort(k[i])[-n_sel:]:          # guided_choices.py:28 (n_sel == 0 keeps every block, as there)
                    nxt.append(wins[pos + i * nc + int(c)])
            pos += nr * nc
        rects = np.asarray(nxt, dtype=np.int32).reshape(-1, 4)
    mask = torch.zeros(n_in, n_out, device=w.device, dtype=torch.float32)
    if len(rects):
        native.fill_rects(mask, torch.from_numpy(np.ascontiguousarray(rects)).to(w.device), 1.0)
    return mask


def conn_mat(n_in, n_out, block_sizes, drop_ratios, w_mat, mat_num='1', dir='/home/dkadetot/saved_mat',
             equal_blks_for_input=True, for_test=False):
    if not len(block_sizes) == len(drop_ratios):
        print('block size and drop ratio should have the same length!')
        raise SystemExit(0)
    if not equal_blks_for_input:
        raise NotImplementedError('equal_blks_for_input=False is not runnable in the reference either')
    if not for_test and torch.is_tensor(w_mat) and w_mat.is_cuda:
        t = conn_mat_device(n_in, n_out, block_sizes, drop_ratios, w_mat)
        t.requires_grad_(False)
        return t
    mask = conn_mat_numpy(n_in, n_out, block_sizes, drop_ratios, w_mat)
    if for_test:
        import scipy.io as sio
        sio.savemat(os.path.join(dir, 'conn_mat%s.mat' % mat_num), {'CM%s' % mat_num: mask})
        return mask
    t = torch.from_numpy(mask).to(mask_device())
    t.requires_grad_(False)
    return t
