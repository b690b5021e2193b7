"""CPU restatement of the reference's minibatch assembly (test infrastructure; never imported by the product path).

Follows core.py:156-200 of the reference for sequential models: N_batches = int(N_snt / batch_size) (trailing
sentences of the chunk are dropped), per minibatch max_len = the longest of its sentences, a zeroed
[max_len, batch_size, inp_dim] tensor, and every sentence copied behind `random.randint(0, max_len - snt_len)` leading
zero rows -- one draw per sentence, in sentence order, from Python's global `random`.

Pinned by tests/golden/chunk_batches.npz = the `inp` tensors the reference's own core.run_nn handed to forward_model
(tests/golden/make_golden.py gen_chunk_batches), and the value of the next random.random() after the loop.
"""
import random

import numpy as np


def sentence_lengths(data_end_index):
    """core.py:171-173: arr_snt_len."""
    end = np.asarray(data_end_index, dtype=np.int64)
    lens = np.empty_like(end)
    lens[0] = end[0]
    lens[1:] = end[1:] - end[:-1]
    return lens


def assemble_minibatches(data_set, data_end_index, batch_size, rng=random):
    """Returns [(inp [max_len, batch_size, inp_dim] float32, max_len, left offsets)] for the chunk (core.py:180-200)."""
    data_set = np.asarray(data_set, dtype=np.float32)
    end = np.asarray(data_end_index, dtype=np.int64)
    lens = sentence_lengths(end)
    n_batches = int(len(end) / batch_size)
    out = []
    snt_index, beg_snt = 0, 0
    for _ in range(n_batches):
        max_len = int(max(lens[snt_index:snt_index + batch_size]))
        inp = np.zeros((max_len, batch_size, data_set.shape[1]), dtype=np.float32)
        lefts = []
        for k in range(batch_size):
            snt_len = int(end[snt_index] - beg_snt)
            n_left = rng.randint(0, max_len - snt_len)
            inp[n_left:n_left + snt_len, k, :] = data_set[beg_snt:beg_snt + snt_len, :]
            lefts.append(n_left)
            beg_snt = int(end[snt_index])
            snt_index += 1
        out.append((inp, max_len, lefts))
    return out
