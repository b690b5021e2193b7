"""ORACLE (test infrastructure only) -- CPU restatement of the reference's integer/host logic.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package; the product (pytorch-kaldi-cgs_b200/) never does.

Pinned: every function here is checked against fixtures produced by running the reference
itself (tests/golden/make_golden.py -> tests/golden/*.npz) in tests/test_oracle_pinned.py.

Functions
  conn_mat          hcgs.py:8-74 + cgs_base.py:4-58   random hierarchical block mask
  quantize_w        quantized_modules.py:77-97        Quantize(balanced=False)
  quantize_inp      quantized_modules.py:99-119       Quantize_inp
  prune_mask        quantized_modules.py:15-28        magnitude mask at np.percentile
  load_counts       data_io.py:148-152
  kaldi_mat_bytes   data_io.py:770-806                write_mat record
"""
import struct

import numpy as np


def _inner(n_in, n_out, block_sizes, drop_ratios):
    """cgs_base.conn_mat (cgs_base.py:4-58), equal_blks_for_input=True branch.
    block_sizes/drop_ratios are innermost-first and are consumed with pop() like the reference."""
    block_sizes, drop_ratios = list(block_sizes), list(drop_ratios)
    if len(block_sizes) == 0:
        return np.full((n_in, n_out), 1, dtype='float32')                     # :6-7
    bs = block_sizes.pop()                                                    # :8
    keep = 1 - float(drop_ratios.pop()) / 100                                 # :9-10
    nbr = n_in // bs + (1 if n_in % bs != 0 else 0)                           # :12-15
    nbc = n_out // bs + (1 if n_out % bs != 0 else 0)                         # :16-18
    n_sel = int(round(nbc * keep))                                            # :20
    m = np.full((n_in, n_out), 0, dtype='float32')
    for i in range(nbr):                                                      # :22 and :41 (last row drawn the same way)
        choices = np.random.choice(nbc, n_sel, False)                         # :24 / :41
        for j in range(n_sel):
            c = choices[j]
            view = m[i * bs:(i + 1) * bs, c * bs:(c + 1) * bs]                # numpy clips ragged edges (:27-40)
            view[...] = _inner(view.shape[0], view.shape[1], block_sizes, drop_ratios)
    return m


def conn_mat(n_in, n_out, block_sizes, drop_ratios):
    """hcgs.conn_mat (hcgs.py:8-74): rows = n_in, cols = n_out (the reference's naming)."""
    assert len(block_sizes) == len(drop_ratios)
    bsz, drs = list(block_sizes), list(drop_ratios)
    bsz.reverse()                                                             # :12-13
    drs.reverse()
    bs = bsz.pop()                                                            # :15
    keep = 1 - float(drs.pop()) / 100                                         # :16-17
    nbr = n_in / bs                                                           # :18 true division
    if n_in % bs != 0:
        nbr += 1
    nbc = n_out / bs                                                          # :22
    if n_out % bs != 0:
        nbc += 1
    n_sel = int(round(nbc * keep))                                            # :26 on the FLOAT count
    nbr, nbc = int(nbr), int(nbc)                                             # :27-28
    m = np.full((n_in, n_out), 0, dtype='float32')
    for i in range(nbr):
        choices = np.random.choice(nbc, n_sel, False)                         # :29 / :41
        for j in range(n_sel):
            c = choices[j]
            view = m[i * bs:(i + 1) * bs, c * bs:(c + 1) * bs]
            view[...] = _inner(view.shape[0], view.shape[1], bsz, drs)        # :36 / :40 / :45
    return m


def quantize_w(w, bits):
    """Returns (clamped master, quantised copy); float32 arithmetic like torch."""
    w = np.clip(np.asarray(w, dtype=np.float32), -1, 1)
    s = np.float32(2 ** (bits - 1))
    q = (np.ceil(np.abs(w) * s) / s).astype(np.float32) * np.sign(w).astype(np.float32)
    return w, q.astype(np.float32)


def quantize_inp(x, bits, repeat=1):
    x = np.asarray(x, dtype=np.float32)
    for _ in range(repeat):
        mx, mn = np.abs(x.max()), np.abs(x.min())
        var = mx if mx > mn else mn
        if var == 0:
            return x
        s = np.float32(2 ** (bits - 1))
        t = (x / var).astype(np.float32)
        t = (np.ceil(np.abs(t) * s) / s).astype(np.float32)
        x = (t * var).astype(np.float32) * np.sign(x).astype(np.float32)
    return x


def prune_mask(w, perc):
    a = np.abs(np.asarray(w, dtype=np.float32))
    thr = np.percentile(a.flatten(), perc)
    return (a > thr).astype(np.float32)


def load_counts(path):
    with open(path) as f:
        row = next(f).strip().strip('[]').strip()
    return np.array([np.float32(v) for v in row.split()])


def log_prior(counts):
    """core.py:245: np.log(counts / np.sum(counts)) in float32."""
    counts = np.asarray(counts, dtype=np.float32)
    return np.log(counts / np.sum(counts))


def kaldi_mat_bytes(m, key=''):
    """data_io.write_mat record: [key ' '] '\\0B' 'FM ' '\\4' rows '\\4' cols data (float32)."""
    m = np.ascontiguousarray(m)
    assert m.dtype == np.float32
    out = b''
    if key != '':
        out += (key + ' ').encode('latin1')
    out += b'\0B' + b'FM ' + b'\x04' + struct.pack('<I', m.shape[0]) + b'\x04' + struct.pack('<I', m.shape[1])
    return out + m.tobytes()
