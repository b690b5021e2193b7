"""Posterior wire format (reference: data_io.py:148-152 load_counts, :770-806 write_mat; the
normalisation itself is core.py:241-249).

The Kaldi decoder consumes ``*_to_decode.ark`` records ``<utt-id> \\0B FM \\4<rows>\\4<cols><float32 data>``;
these bytes are the contract with kaldi_decoding_scripts/decode_dnn.sh and must not change.
"""
import struct

import numpy as np
import torch


def load_counts(class_counts_file):
    """First line of a Kaldi count file ``[ c0 c1 ... ]`` parsed value by value as float32."""
    with open(class_counts_file) as f:
        row = next(f).strip().strip('[]').strip()
    return np.array([np.float32(v) for v in row.split()])


def log_prior_from_counts(counts):
    """``np.log(counts / np.sum(counts))`` in float32, exactly as core.py:245 evaluates it."""
    counts = np.asarray(counts, dtype=np.float32)
    return np.log(counts / np.sum(counts))


def log_prior_tensor(counts, device):
    """Device copy handed to the fused log-softmax kernel (computed on the host so that the
    float32 rounding of the prior is bit-identical to the reference's NumPy expression)."""
    return torch.from_numpy(log_prior_from_counts(counts).astype(np.float32)).to(device)


def mat_record(m, key=''):
    """Bytes of one binary Kaldi matrix record (float32 or float64)."""
    m = np.ascontiguousarray(m)
    if m.dtype == np.float32:
        tag = b'FM '
    elif m.dtype == np.float64:
        tag = b'DM '
    else:
        raise TypeError("'%s', please use 'float32' or 'float64'" % m.dtype)
    head = (key + ' ').encode('latin1') if key != '' else b''
    return b''.join([head, b'\0B', tag, b'\x04', struct.pack('<I', m.shape[0]), b'\x04',
                     struct.pack('<I', m.shape[1]), m.tobytes()])


def write_mat(output_folder, file_or_fd, m, key=''):
    """Reference signature (data_io.write_mat); ``file_or_fd`` is an open binary file or a path."""
    if hasattr(file_or_fd, 'write'):
        file_or_fd.write(mat_record(m, key))
    else:
        with open(file_or_fd, 'wb') as fd:
            fd.write(mat_record(m, key))


def read_ark(path):
    """Minimal reader of the records written above (tests and tooling)."""
    out = []
    with open(path, 'rb') as f:
        data = f.read()
    pos = 0
    while pos < len(data):
        sp = data.index(b' ', pos)
        key = data[pos:sp].decode('latin1')
        pos = sp + 1
        assert data[pos:pos + 2] == b'\0B'
        tag = data[pos + 2:pos + 5]
        dt = np.float32 if tag == b'FM ' else np.float64
        rows = struct.unpack('<I', data[pos + 6:pos + 10])[0]
        cols = struct.unpack('<I', data[pos + 11:pos + 15])[0]
        pos += 15
        n = rows * cols * np.dtype(dt).itemsize
        out.append((key, np.frombuffer(data[pos:pos + n], dtype=dt).reshape(rows, cols)))
        pos += n
    return out
