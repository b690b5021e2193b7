"""CPU-only checks of the product's host side: drop-in construction parity, state_dict surface,
option handling, graph interpreter parsing, the C-ABI symbol table, and the no-fallback rule."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT
from helpers import NETS, build_product_nets, golden_state, load_golden


@pytest.mark.parametrize('name,to_do,kw', [
    ('train_small', 'train', {}),
    ('train_htanh', 'train', dict(seed=11, lstm_over=dict(lstm_act='htanh,htanh', lstm_lay='20,28', lstm_drop='0.0,0.1'))),
    ('eval_forward', 'forward', dict(seed=77)),
    ('train_quant', 'train', dict(seed=5, lstm_over=dict(lstm_quant='True'), mlp_over=dict(mlp_quant='True'))),
])
def test_construction_is_bit_identical_to_reference(name, to_do, kw):
    """Same seed -> same initial weights, masks, BN buffers, under the same state_dict keys."""
    g = load_golden(name)
    nets, _ = build_product_nets(to_do, **kw)
    for net_name, net in nets.items():
        ref = golden_state(g, 'init', net_name)
        sd = net.state_dict()
        assert list(sd.keys()) == list(ref.keys()), net_name
        for k, v in sd.items():
            assert torch.equal(v.cpu(), ref[k]), (net_name, k)


def test_state_dict_roundtrip_and_plugin_surface():
    nets, _ = build_product_nets('train')
    lstm, cd = nets['LSTM_layers'], nets['MLP_layers']
    assert lstm.out_dim == 24 and cd.out_dim == 10
    for n in (lstm, cd):
        for attr in ('prune', 'guided_hcgs', 'apply_guided_hcgs', 'if_pattern', 'skip_regularization'):
            assert hasattr(n, attr)
    sd = {k: v.clone() for k, v in lstm.state_dict().items()}
    nets2, _ = build_product_nets('train', seed=1)
    nets2['LSTM_layers'].load_state_dict(sd)
    for k, v in nets2['LSTM_layers'].state_dict().items():
        assert torch.equal(v, sd[k])
    # masks are parameters (travel in checkpoints) but the optimizer never sees a gradient for them
    assert 'hcgsx.0.mask' in dict(lstm.named_parameters())


def test_missing_cgs_keys_default_off_and_alias_spelling():
    import neural_networks as nb
    base = dict(dnn_lay='16,8', dnn_drop='0.1,0.0', dnn_use_batchnorm='True,False', dnn_use_laynorm='False,False',
                dnn_use_laynorm_inp='False', dnn_use_batchnorm_inp='False', dnn_act='relu,softmax', to_do='train')
    m = nb.MLP(base, 12)           # TIMIT_MLP_mfcc_basic style section: no CGS keys at all
    assert not m.mlp_hcgs and not m.mlp_quant and not m.prune and not m.if_pattern
    np.random.seed(0)
    m2 = nb.MLP(dict(base, mlp_hcgs='True', hcgs_block='8,4', hcgs_drop='50,25'), 12)   # LibriSpeech_CGS spelling
    assert m2.hcgs_sparse == [50.0, 25.0] and tuple(m2.hcgs[0].mask.shape) == (16, 12)
    with pytest.raises(ValueError):
        nb.LSTM(dict(lstm_lay='8', lstm_drop='0', lstm_use_batchnorm='False', lstm_use_laynorm='False',
                     lstm_use_laynorm_inp='False', lstm_use_batchnorm_inp='False', lstm_act='tanh',
                     lstm_orthinit='True', lstm_bidir='False'), 4)


def test_no_cpu_fallback():
    """Without a CUDA device the compute entry points must raise, never compute on the host."""
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from b200 import native
    nets, _ = build_product_nets('train')
    with pytest.raises(native.NativeError):
        nets['LSTM_layers'](torch.zeros(3, 2, 12))
    with pytest.raises(native.NativeError):
        nets['MLP_layers'](torch.zeros(6, 24))


def test_graph_parser():
    from b200.graph import parse_model
    ops = parse_model('out_dnn1=compute(LSTM_layers,fbank)\n\tloss_final=sum(loss_cd,loss_mono_w)\n'
                      '\tloss_mono_w=mult_constant(loss_mono,1.0)')
    assert ops[0] == ('out_dnn1', 'compute', 'LSTM_layers', 'fbank')
    assert ops[2] == ('loss_mono_w', 'mult_constant', 'loss_mono', '1.0')


def test_c_abi_exports_every_declared_symbol():
    from b200 import native
    header = open(os.path.join(ROOT, 'include', 'hcgs_b200.h')).read()
    declared = set(re.findall(r'^\s*(?:int|const char\*)\s+(hcgs_\w+)\s*\(', header, flags=re.M))
    assert len(declared) >= 20
    lib = ctypes.CDLL(native.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), 'library does not export %s' % name
    assert declared == set(native.SIGNATURES), declared ^ set(native.SIGNATURES)
    assert native.load().hcgs_abi_version() == 1


def test_percentile_threshold_matches_numpy():
    """quantized_modules.percentile_threshold (device-side np.percentile of the pruning path, reference
    quantized_modules.py:15-28) against numpy on the same float32 data, including the float32 cast the reference's
    `tensor > threshold` comparison applies, ties, zeros (HCGS-masked weights) and both interpolation branches."""
    import numpy as np
    import torch
    from quantized_modules import percentile_threshold
    rng = np.random.RandomState(0)
    for n, perc, zeros in [(1000, 70, 0.0), (577, 70, 0.9), (24 * 24, 70, 0.75), (10, 33.3, 0.0), (7, 50, 0.0),
                           (302500, 70, 0.89), (5, 99.9, 0.0), (64, 0, 0.0), (64, 100, 0.0)]:
        w = rng.randn(n).astype(np.float32)
        w[rng.rand(n) < zeros] = 0.0
        a = np.abs(w)
        want = np.percentile(a, perc)
        got = percentile_threshold(torch.from_numpy(a), perc)
        assert got.dtype == torch.float32 and got.dim() == 0
        assert got.item() == float(np.float32(want)), (n, perc, got.item(), want)
        assert torch.equal(torch.from_numpy(a) > got, torch.from_numpy(a) > want)


def test_drop_mask_prefetch_is_bit_identical_to_drawing_on_demand():
    """b200/rng.py: masks drawn one forward ahead on a copy of the generator state == the reference's on-demand draws
    (neural_networks.py:843-851), also when somebody seeds or uses the default generator in between."""
    import torch
    from b200 import rng
    spec = [(7, 5, 0.8), (7, 3, 0.6)]

    def on_demand():
        return [torch.bernoulli(torch.Tensor(b, h).fill_(p)) for (b, h, p) in spec]

    # reference stream: forward, forward, foreign draw, forward, re-seed, forward
    torch.manual_seed(11)
    ref = [on_demand(), on_demand()]
    foreign_ref = torch.rand(3)
    ref.append(on_demand())
    torch.manual_seed(5)
    ref.append(on_demand())
    tail_ref = torch.rand(2)

    torch.manual_seed(11)
    pf = rng.DropMaskPrefetcher()
    got = []
    got.append(pf.take(spec)); pf.prefetch(spec)
    got.append(pf.take(spec)); pf.prefetch(spec)          # served from the prefetch
    foreign = torch.rand(3)                                  # somebody else draws: the prefetch must be dropped
    got.append(pf.take(spec)); pf.prefetch(spec)
    torch.manual_seed(5)                                     # re-seeded: dropped again
    got.append(pf.take(spec)); pf.prefetch(spec)
    tail = torch.rand(2)                                     # a pending prefetch never advances the default generator
    assert torch.equal(foreign, foreign_ref) and torch.equal(tail, tail_ref)
    for a, b in zip(ref, got):
        for x, y in zip(a, b):
            assert torch.equal(x, y)
    # a different shape request drops the prefetch as well
    torch.manual_seed(3)
    want = [torch.bernoulli(torch.Tensor(2, 2).fill_(0.5))]
    torch.manual_seed(3)
    pf.prefetch(spec)
    assert torch.equal(pf.take([(2, 2, 0.5)])[0], want[0])
