"""Guided HCGS masks built with the weights on the GPU (guided_hcgs.py:9-76, guided_choices.py:4-30).

The host implementation (guided_hcgs.conn_mat_numpy) is pinned to the reference's own output (tests/golden/masks.npz,
tests/test_oracle_pinned.py::test_guided_mask_matches_reference); the device path must produce the same BITS: its block
means have to equal the CPU AvgPool2d means exactly, otherwise near-ties would rank differently."""
import numpy as np
import pytest
import torch

import guided_hcgs
from helpers import load_golden


def _seq_mean(a, r0, r1, c0, c1):
    s = np.float32(0)
    for r in range(r0, r1):
        for c in range(c0, c1):
            s = np.float32(s + abs(a[r, c]))
    return np.float32(s / np.float32((r1 - r0) * (c1 - c0)))


def test_row_major_fp32_sum_is_the_cpu_avgpool_arithmetic():
    """The premise of the device kernel, checked on the CPU: AvgPool2d on a CPU float tensor == fp32 row-major
    accumulation / window area, bit for bit (full and ragged windows)."""
    torch.manual_seed(3)
    for r, c, bs in [(8, 40, 8), (5, 37, 8), (4, 8, 4), (32, 70, 32)]:
        w = torch.randn(r, c).abs()
        pool = torch.nn.AvgPool2d(bs, bs) if r == bs else torch.nn.AvgPool2d((r, bs), bs)
        out = pool(w.reshape(1, 1, r, c))[0, 0, 0].numpy()
        for j in range(c // bs):
            assert out[j] == _seq_mean(w.numpy(), 0, r, j * bs, (j + 1) * bs)
        if c % bs:
            c1 = c % bs
            tail = torch.nn.AvgPool2d((r, c1), c1)(w[:, c - c1:].reshape(1, 1, r, c1))[0, 0, 0, 0].item()
            assert np.float32(tail) == _seq_mean(w.numpy(), 0, r, c - c1, c)


@pytest.mark.gpu
def test_window_means_are_bit_identical_to_the_host_pooling():
    from b200 import native
    torch.manual_seed(5)
    w = torch.randn(70, 101)
    wins = np.array([[0, 32, 0, 32], [32, 64, 96, 101], [64, 70, 32, 64], [3, 7, 8, 12], [0, 70, 0, 101]], dtype=np.int32)
    got = native.window_abs_mean(w.cuda(), torch.from_numpy(wins).cuda()).cpu().numpy()
    for g, (r0, r1, c0, c1) in zip(got, wins):
        assert g == _seq_mean(w.numpy(), r0, r1, c0, c1)


@pytest.mark.gpu
@pytest.mark.parametrize('n_in,n_out,blocks,drops', [(48, 40, [8, 4], [50, 50]), (550, 550, [32, 4], [75, 62.5]),
                                                     (550, 40, [32, 4], [75, 62.5]), (1944, 550, [64, 4], [50, 25]),
                                                     (100, 70, [16], [50]), (36, 70, [8, 2], [50, 50]),
                                                     (64, 64, [16, 4, 2], [50, 50, 50])])
def test_device_guided_mask_equals_host_mask(n_in, n_out, blocks, drops):
    torch.manual_seed(n_in + n_out)
    w = torch.randn(n_in, n_out)
    if n_in == 550:
        # the state guided HCGS starts from: weights trained under a random HCGS mask (whole blocks exactly zero)
        from oracle import hcgs_oracle
        np.random.seed(1)
        w = w * torch.from_numpy(hcgs_oracle.conn_mat(n_in, n_out, list(blocks), list(drops)).astype(np.float32))
    host = guided_hcgs.conn_mat_numpy(n_in, n_out, list(blocks), list(drops), w)
    dev = guided_hcgs.conn_mat(n_in, n_out, list(blocks), list(drops), w.cuda())
    assert dev.is_cuda and dev.dtype == torch.float32 and not dev.requires_grad
    assert np.array_equal(dev.cpu().numpy(), host)


@pytest.mark.gpu
def test_device_guided_mask_matches_reference_fixture():
    g = load_golden('masks')
    w = torch.from_numpy(g['guided_w'])
    m = guided_hcgs.conn_mat(48, 40, [8, 4], [50, 50], w.cuda())
    ref = np.unpackbits(g['guided_bits'])[:48 * 40].reshape(48, 40)
    assert np.array_equal(m.cpu().numpy().astype(np.uint8), ref)


@pytest.mark.gpu
def test_window_wider_than_the_strip_fails_like_the_reference():
    """A ragged block narrower than the next level's block size makes the reference's AvgPool2d raise; both paths do."""
    w = torch.randn(33, 65)
    with pytest.raises(RuntimeError):
        guided_hcgs.conn_mat_numpy(33, 65, [8, 2], [50, 50], w)
    with pytest.raises(RuntimeError):
        guided_hcgs.conn_mat(33, 65, [8, 2], [50, 50], w.cuda())
