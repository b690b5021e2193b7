"""Inner levels of guided HCGS masks (reference: guided_cgs_base.py:5-59)."""
import torch

import cgs_base


def conn_mat(n_in, n_out, block_sizes, drop_ratios, w_mat, equal_blks_for_input=False, recursive_call=1):
    if len(block_sizes) and not equal_blks_for_input:
        raise NotImplementedError('only equal_blks_for_input=True is supported')
    levels = cgs_base.levels_from_lists(list(reversed(list(block_sizes))), list(reversed(list(drop_ratios))))
    w_abs = torch.as_tensor(w_mat).detach().to('cpu', torch.float32)
    return cgs_base.fill_levels(int(n_in), int(n_out), levels, False, w_abs=w_abs)
