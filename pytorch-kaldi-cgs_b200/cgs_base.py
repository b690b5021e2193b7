"""Hierarchical coarse-grain-sparsity (HCGS) connection masks -- recursion core.

Drop-in for the reference module of the same name (reference: cgs_base.py:4-58,
the inner levels of hcgs.conn_mat; guided_cgs_base.py:5-59 for the guided family).
The mask is data, but its *bits* are part of the contract: the reference draws one
``np.random.choice(n_cols, n_sel, False)`` per block-row from the process-global
legacy NumPy RNG and recurses into every kept block before drawing for the next
block-row, so the draw order is a depth-first walk.  This file reproduces exactly
that walk with one generic routine instead of the reference's hand-unrolled
branches (ragged edge blocks are simply clipped slices).

Only ``equal_blks_for_input=True`` is meaningful: the reference's other branch
cannot run at the top level (``range(float)``, hcgs.py:48) and HCGS.py never
selects it, so it is rejected loudly here.
"""
import numpy as np


def _n_blocks(n, bs):
    """ceil(n / bs) the way the reference counts (floor division, +1 if ragged)."""
    return n // bs + (1 if n % bs != 0 else 0)


def guided_block_choice(w_abs_strip, n_blk, n_sel, bs):
    """Column blocks of one block-row strip ranked by mean |W| (guided_choices.py:4-30).

    Same operator sequence as the reference (torch AvgPool2d on a [1,1,r,c] CPU
    tensor, ragged tail pooled separately, ``np.argsort`` of the block means,
    last ``n_sel`` taken) so that near-ties resolve identically.  Note the
    reference's ``[-0:]`` quirk: ``n_sel == 0`` selects *every* block.
    """
    import torch
    w = torch.as_tensor(w_abs_strip).detach().to('cpu', torch.float32)
    r, c = w.shape
    t = w.reshape(1, 1, r, c)
    pool = torch.nn.AvgPool2d(bs, bs) if r == bs else torch.nn.AvgPool2d((r, bs), bs)
    out = pool(t)
    if c % bs != 0:
        x = (n_blk - 1) * bs
        c1 = c - x
        t2 = w[:, x:x + c1].reshape(1, 1, r, c1)
        pool2 = torch.nn.AvgPool2d(bs, c1) if (r == bs and c1 >= bs) else torch.nn.AvgPool2d((r, c1), c1)
        out = torch.cat((out, pool2(t2)), 3)
    means = out[0, 0, :, :].numpy()
    return np.argsort(means[0])[-n_sel:]


def fill_levels(n_rows, n_cols, levels, top_level_float_count=False, w_abs=None):
    """Return the float32 0/1 mask ``[n_rows, n_cols]`` for ``levels``.

    levels : sequence of ``(block_size, drop_percent)`` from the coarsest level in.
    top_level_float_count : the reference's outermost random level (hcgs.py:18-28)
        counts column blocks with true division before rounding the number kept,
        which differs from the inner levels (cgs_base.py:11-20) for ragged widths
        (e.g. 550/32 -> 18.1875*0.25 -> 5 kept, where the integer count gives 4).
    w_abs : ``None`` for random masks (global legacy RNG, like the reference); a
        ``[n_rows, n_cols]`` |W| matrix for guided masks (block choice = top-k mean |W|).
    """
    if len(levels) == 0:
        # leaf: fully connected block (cgs_base.py:6-7)
        return np.full((n_rows, n_cols), 1, dtype='float32')
    bs, drop = levels[0]
    bs = int(bs)
    keep = 1 - float(drop) / 100
    n_blk_rows = _n_blocks(n_rows, bs)
    n_blk_cols = _n_blocks(n_cols, bs)
    if top_level_float_count:
        cols_f = n_cols / bs + (1 if n_cols % bs != 0 else 0)
        n_sel = int(round(cols_f * keep))
    else:
        n_sel = int(round(n_blk_cols * keep))
    mask = np.full((n_rows, n_cols), 0, dtype='float32')
    for i in range(n_blk_rows):
        r0, r1 = i * bs, min((i + 1) * bs, n_rows)
        if w_abs is None:
            picks = np.random.choice(n_blk_cols, n_sel, False)
        else:
            picks = guided_block_choice(w_abs[r0:r1, :], n_blk_cols, n_sel, bs)
        for c in picks:
            c0, c1 = int(c) * bs, min((int(c) + 1) * bs, n_cols)
            sub_w = None if w_abs is None else w_abs[r0:r1, c0:c1]
            mask[r0:r1, c0:c1] = fill_levels(r1 - r0, c1 - c0, levels[1:], False, sub_w)
    return mask


def levels_from_lists(block_sizes, drop_ratios):
    if len(block_sizes) != len(drop_ratios):
        raise ValueError('block size and drop ratio should have the same length!')
    return [(int(b), float(d)) for b, d in zip(block_sizes, drop_ratios)]


def conn_mat(n_in, n_out, block_sizes, drop_ratios, equal_blks_for_input=False, recursive_call=1):
    """Reference-compatible entry (cgs_base.py:4).  ``block_sizes``/``drop_ratios`` hold the
    *remaining* levels innermost-first (the reference pops from the end of the list)."""
    if len(block_sizes) and not equal_blks_for_input:
        raise NotImplementedError(
            'only equal_blks_for_input=True is supported (the reference cannot run the other branch, hcgs.py:48)')
    levels = levels_from_lists(list(reversed(list(block_sizes))), list(reversed(list(drop_ratios))))
    return fill_levels(int(n_in), int(n_out), levels, top_level_float_count=False)
