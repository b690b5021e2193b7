"""Random HCGS connection masks -- top level (reference: hcgs.py:8-74, duplicated at :77-143).

``conn_mat(n_in, n_out, ...)`` keeps the reference's (confusing) argument names:
``n_in`` is the number of mask *rows* (= out_features of the linear) and ``n_out`` the
number of columns, because HCGS.py calls it as ``conn_mat(out_features, in_features, ..)``.

Bit-exactness contract: global legacy NumPy RNG, depth-first draw order,
float block count at this level only, Python-3 banker's ``round``.  See cgs_base.fill_levels.
"""
import os

import numpy as np
import torch

import cgs_base


def mask_device():
    """The reference forces ``cuda`` (hcgs.py:71-72).  On a GPU-less host (unit tests,
    mask-only tooling) we return a CPU tensor instead of crashing; every compute entry
    point still refuses to run without the CUDA library."""
    return torch.device('cuda') if torch.cuda.is_available() else torch.device('cpu')


def conn_mat_numpy(n_in, n_out, block_sizes, drop_ratios):
    levels = cgs_base.levels_from_lists(list(block_sizes), list(drop_ratios))
    return cgs_base.fill_levels(int(n_in), int(n_out), levels, top_level_float_count=True)


def conn_mat(n_in, n_out, block_sizes, drop_ratios, mat_num='1', dir='/home/dkadetot/saved_mat',
             equal_blks_for_input=True, for_test=False):
    if not len(block_sizes) == len(drop_ratios):
        print('block size and drop ratio should have the same length!')
        raise SystemExit(0)  # the reference calls exit() here (hcgs.py:9-11)
    if not equal_blks_for_input:
        raise NotImplementedError(
            'equal_blks_for_input=False cannot run in the reference either (hcgs.py:48 range(float))')
    mask = conn_mat_numpy(n_in, n_out, block_sizes, drop_ratios)
    if for_test:
        import scipy.io as sio
        sio.savemat(os.path.join(dir, 'conn_mat%s.mat' % mat_num), {'CM%s' % mat_num: mask})
        return mask
    t = torch.from_numpy(mask).to(mask_device())
    t.requires_grad_(False)
    return t
