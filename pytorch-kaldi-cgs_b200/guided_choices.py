"""Top-k block choice by mean |W| (reference: guided_choices.py:4-30)."""
from cgs_base import guided_block_choice


def guided_array_rows(w_mat, n_blk, n_blk_sels, blk_size):
    return guided_block_choice(w_mat, n_blk, n_blk_sels, blk_size)
