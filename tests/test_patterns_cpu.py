"""Pattern masks on the CPU: the oracle (oracle/pattern_oracle.py) and the product's HOST code
(pytorch-kaldi-cgs_b200/Pattern.py) against fixtures produced by running the reference itself
(tests/golden/patterns.npz, train_pattern*.npz; make_golden.py gen_patterns / gen_pattern_model).
Bit-exact, including the position of the global numpy RNG stream afterwards."""
import ast
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN
from oracle import pattern_oracle


def _load(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)


def _file_patterns(g, name):
    return np.unpackbits(g['file/' + name])[:128 * 8].reshape(128, 8).astype(np.float64)


def _cases(g):
    return [ast.literal_eval(str(c)) for c in g['cases']]


def test_shipped_pattern_files_are_16_patterns_of_4():
    g = _load('patterns')
    for f in ('b08b08_k04_n16_pattern.npy', 'b08b08_k04_n16_pattern_v1.npy', 'b08b08_k04_n16_pattern_v2.npy'):
        pats = pattern_oracle.file_patterns(_file_patterns(g, f), 16)
        assert len(pats) == 16 and all(p.shape == (8, 8) and p.sum() == 4 for p in pats)


def test_oracle_pattern_masks_bit_exact():
    g = _load('patterns')
    for k, (r, c, mode, from_file, seed) in enumerate(_cases(g)):
        w = g['w_%d' % k]
        np.random.seed(seed)
        pats = pattern_oracle.file_patterns(_file_patterns(g, from_file), 16) if from_file else None
        if mode == 'pattern':
            m = pattern_oracle.pattern_mask((r, c), (8, 8), pats or pattern_oracle.draw_patterns(16, (8, 8), 4))
        elif mode == 'coo':
            m = pattern_oracle.coo_mask(w, (8, 8), 4)
        else:
            m = pattern_oracle.pattern_coo_mask(w, (8, 8), 4, 16, pats)
        ref = np.unpackbits(g['bits_%d' % k])[:r * c].reshape(r, c)
        assert np.array_equal(m.astype(np.uint8), ref), (r, c, mode, from_file)
        assert np.random.randint(0, 2 ** 31 - 1) == int(g['next_rand_%d' % k][0]), (mode, from_file)


def test_product_pattern_masks_bit_exact(tmp_path):
    import Pattern
    g = _load('patterns')
    for k, (r, c, mode, from_file, seed) in enumerate(_cases(g)):
        w = g['w_%d' % k]
        path = False
        if from_file:
            path = str(tmp_path / from_file)
            np.save(path, _file_patterns(g, from_file))
        np.random.seed(seed)
        m = Pattern.Pattern(w, mode, [8, 8], 4, 16, '', pattern_from_file=path)
        mask = m.mask.detach().numpy()
        assert mask.dtype == np.float64 and tuple(mask.shape) == (r, c)
        ref = np.unpackbits(g['bits_%d' % k])[:r * c].reshape(r, c)
        assert np.array_equal(mask.astype(np.uint8), ref), (r, c, mode, from_file)
        assert np.random.randint(0, 2 ** 31 - 1) == int(g['next_rand_%d' % k][0]), (mode, from_file)
    with pytest.raises(AssertionError):
        Pattern.Pattern(np.zeros((550, 40)), 'pattern', [8, 8], 4, 16, '')      # 550 % 8 != 0 (Pattern.py:79-82)
    with pytest.raises(FileNotFoundError):
        Pattern.Pattern(np.zeros((8, 8)), 'pattern', [8, 8], 4, 16, '', pattern_from_file=str(tmp_path / 'nope.npy'))


def test_apply_patterns_incl_ties():
    import Pattern
    g = _load('patterns')
    w = g['apply_w']
    for name in ('a', 'b'):
        kern, ref = g['apply_%s_kernel' % name], g['apply_%s_mask' % name]
        got_o = pattern_oracle.apply_patterns(w, kern)
        got_p = Pattern.apply_patterns(torch.from_numpy(w), torch.from_numpy(kern)).cpu().numpy()
        assert got_o.dtype == np.float32 and np.array_equal(got_o, ref), name
        assert got_p.dtype == np.float32 and np.array_equal(got_p, ref), name
    assert g['apply_b_mask'].max() >= 2          # ties add overlapping patterns (sparsity.py:1136-1137)
    assert g['apply_a_mask'][8:16, 0:8].max() >= 1 and g['apply_a_mask'][8:16, 0:8].sum() > 4   # the constant block


def test_find_top_k_by_kmeans_matches_reference():
    import Pattern
    g = _load('patterns')
    w = g['kmeans_w']
    for impl in ('oracle', 'product'):
        np.random.seed(32)
        if impl == 'oracle':
            kern = pattern_oracle.find_top_k_by_kmeans(w, 16, (8, 8), 4)
        else:
            kern = Pattern.find_top_k_by_kmeans(torch.from_numpy(w), 16, [8, 8], 4, [8, 8]).cpu().numpy()
        assert kern.shape == (16, 1, 8, 8) and np.array_equal(kern, g['kmeans_kernel']), impl
        assert np.random.randint(0, 2 ** 31 - 1) == int(g['kmeans_next_rand'][0]), impl


def test_model_level_pattern_masks_from_reference_weights():
    """The masks the reference's LSTM / MLP built in train_pattern.npz follow from its (initial) weights and mined
    kernels through apply_patterns: oracle and product reproduce them bit for bit."""
    import Pattern
    g = _load('train_pattern')
    n = 0
    for key in g.files:
        if not key.startswith('pattern_mask/LSTM_layers/'):
            continue
        _, _, name, i = key.split('/')
        wname = name[len('pattern_mask_'):]
        w = g['init/LSTM_layers/%s.%s.weight' % (wname, i)]
        kern = g['pattern/LSTM_layers/pattern_%s/%s' % (wname, i)]
        assert np.array_equal(pattern_oracle.apply_patterns(w, kern), g[key]), key
        got = Pattern.apply_patterns(torch.from_numpy(w), torch.from_numpy(kern)).cpu().numpy()
        assert np.array_equal(got, g[key]), key
        n += 1
    assert n == 16
    w, kern = g['init/MLP_layers/wx.0.weight'], g['pattern/MLP_layers/0']
    assert np.array_equal(pattern_oracle.apply_patterns(w, kern), g['pattern_mask/MLP_layers/0'])
