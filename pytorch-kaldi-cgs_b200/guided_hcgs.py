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


def conn_mat(n_in, n_out, block_sizes, drop_ratios, w_mat, mat_num='1', dir='/home/dkadetot/saved_mat',
             equal_blks_for_input=True, for_test=False):
    if not len(block_sizes) == len(drop_ratios):
        print('block size and drop ratio should have the same length!')
        raise SystemExit(0)
    if not equal_blks_for_input:
        raise NotImplementedError('equal_blks_for_input=False is not runnable in the reference either')
    mask = conn_mat_numpy(n_in, n_out, block_sizes, drop_ratios, w_mat)
    if for_test:
        import scipy.io as sio
        sio.savemat(os.path.join(dir, 'conn_mat%s.mat' % mat_num), {'CM%s' % mat_num: mask})
        return mask
    t = torch.from_numpy(mask).to(mask_device())
    t.requires_grad_(False)
    return t
