"""GRU / liGRU / minimalGRU / RNN on the CPU: the oracle's restatement (oracle/model_oracle.gated_rnn_forward)
against fixtures produced by running the reference's own dense classes (tests/golden/train_{gru,...}.npz, made by
make_golden.py gen_gru_family), and the plugin classes' construction (state_dict keys, order and initial values)
against the reference's."""
import numpy as np
import pytest
import torch

from helpers import MLP_OPTS, load_golden
from oracle import model_oracle
from test_oracle_pinned import NETS, _run_oracle_train

FAMILY = [   # (fixture, option prefix, class name, overrides, spec)
    ('train_gru', 'gru', 'GRU', {}, dict(cell='gru', act=['tanh', 'tanh'], use_bn=[True, True], drop=[0.2, 0.2])),
    ('train_gru_bidir_nobn', 'gru', 'GRU',
     {'gru_bidir': 'True', 'gru_use_batchnorm': 'False,False', 'gru_drop': '0.0,0.1'},
     dict(cell='gru', act=['tanh', 'tanh'], use_bn=[False, False], drop=[0.0, 0.1], bidir=True)),
    ('train_ligru', 'ligru', 'liGRU', {'ligru_act': 'relu,relu'},
     dict(cell='ligru', act=['relu', 'relu'], use_bn=[True, True], drop=[0.2, 0.2])),
    ('train_minimalgru', 'minimalgru', 'minimalGRU', {},
     dict(cell='minimalgru', act=['tanh', 'tanh'], use_bn=[True, True], drop=[0.2, 0.2])),
    ('train_rnn', 'rnn', 'RNN', {'rnn_act': 'relu,relu'},
     dict(cell='rnn', act=['relu', 'relu'], use_bn=[True, True], drop=[0.2, 0.2])),
]


def family_opts(prefix, cls, **over):
    p = prefix
    o = {'arch_name': 'LSTM_layers', 'arch_library': 'neural_networks', 'arch_class': cls, 'arch_seq_model': 'True',
         p + '_lay': '24,16', p + '_drop': '0.2,0.2', p + '_use_laynorm_inp': 'False', p + '_use_batchnorm_inp': 'False',
         p + '_use_laynorm': 'False,False', p + '_use_batchnorm': 'True,True', p + '_bidir': 'False',
         p + '_act': 'tanh,tanh', p + '_orthinit': 'True', 'arch_lr': '0.0016', 'arch_opt': 'rmsprop',
         'opt_momentum': '0.0', 'opt_alpha': '0.95', 'opt_eps': '1e-8', 'opt_centered': 'False',
         'opt_weight_decay': '0.0', 'out_folder': ''}
    o.update(over)
    return o


def family_specs(spec):
    lstm = dict(lay=[24, 16], hcgs=False, **spec)
    cd = dict(lay=[10], act=['softmax'], use_bn=[False], drop=[0.0], hcgs=True)
    mono = dict(lay=[5], act=['softmax'], use_bn=[False], drop=[0.0], hcgs=False)
    return {'lstm': lstm, 'cd': cd, 'mono': mono}


def build_family_nets(prefix, cls, over, to_do='train', D=12, seed=0, device='cpu', extra=None):
    """the recurrent net + the two MLP heads in the cfg's model order with run_nn's seeding"""
    import neural_networks as nn_b200
    torch.manual_seed(seed)
    np.random.seed(seed)
    a1 = dict(family_opts(prefix, cls, **over), to_do=to_do, use_cuda=str(device != 'cpu'))
    a1.update(extra or {})
    a2 = dict(MLP_OPTS, to_do=to_do, use_cuda=str(device != 'cpu'))
    a3 = dict(a2, arch_name='MLP_layers2', dnn_lay='5', mlp_hcgs='False', arch_lr='0.0004')
    rec = getattr(nn_b200, cls)(a1, D)
    cd = nn_b200.MLP(a2, rec.out_dim)
    mono = nn_b200.MLP(a3, rec.out_dim)
    nets = {'LSTM_layers': rec, 'MLP_layers': cd, 'MLP_layers2': mono}
    for n in nets.values():
        n.to(device)
        n.train() if to_do == 'train' else n.eval()
    return nets, (a1, a2, a3)


@pytest.mark.parametrize('name,prefix,cls,over,spec', FAMILY)
def test_oracle_gated_rnn_matches_reference(name, prefix, cls, over, spec):
    g = load_golden(name)
    steps = int(g['dims'][3])
    params, outs = _run_oracle_train(g, family_specs(spec), steps)
    for s, (out, grads) in enumerate(outs):
        ref_loss = g['loss/%d' % s]
        assert abs(out['loss_final'].item() - ref_loss[0]) <= 1e-6 * abs(ref_loss[0])
        np.testing.assert_allclose(out['out_dnn1'].detach().numpy().reshape(g['out_dnn1/%d' % s].shape),
                                   g['out_dnn1/%d' % s], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(out['out_dnn2'].detach().numpy(), g['out_dnn2/%d' % s], rtol=1e-5, atol=1e-6)
    out, grads = outs[-1]
    n_checked = 0
    for n, gr in grads['lstm'].items():
        key = 'grad/LSTM_layers/' + n
        if key not in g.files:
            assert gr is None
            continue
        np.testing.assert_allclose(gr.numpy(), g[key], rtol=2e-4, atol=1e-7, err_msg=key)
        n_checked += 1
    assert n_checked >= 2 * 2 * len(model_oracle.CELL_GATES[spec['cell']])
    for n, v in params['lstm'].items():
        if n.endswith('num_batches_tracked'):
            continue
        np.testing.assert_allclose(v.detach().numpy(), g['final/LSTM_layers/' + n], rtol=1e-4, atol=2e-6, err_msg=n)


@pytest.mark.parametrize('name,prefix,cls,over,spec', FAMILY)
def test_plugin_construction_is_bit_identical_to_reference(name, prefix, cls, over, spec):
    g = load_golden(name)
    seed = 30 + [f[0] for f in FAMILY].index(name)
    nets, _ = build_family_nets(prefix, cls, over, seed=seed)
    sd = nets['LSTM_layers'].state_dict()
    ref_keys = [k[len('init/LSTM_layers/'):] for k in g.files if k.startswith('init/LSTM_layers/')]
    assert list(sd.keys()) == ref_keys                     # same modules, same registration order
    for k, v in sd.items():
        assert np.array_equal(v.numpy(), g['init/LSTM_layers/' + k]), k
    assert nets['LSTM_layers'].out_dim == g['init/MLP_layers/wx.0.weight'].shape[1]
    # the heads built after it see the same RNG stream as in the reference
    assert np.array_equal(nets['MLP_layers'].state_dict()['wx.0.weight'].numpy(), g['init/MLP_layers/wx.0.weight'])


def test_hcgs_keys_add_masks_without_disturbing_the_reference_parameters():
    """<prefix>_hcgs = True adds hcgsx / hcgsh mask parameters (numpy RNG); every reference parameter keeps its value."""
    g = load_golden('train_gru')
    nets, _ = build_family_nets('gru', 'GRU', {}, seed=30, extra=dict(gru_hcgs='True', hcgsx_block='8,4',
                                                                      hcgsx_sparse='50,50', hcgsh_block='8,4',
                                                                      hcgsh_sparse='50,50'))
    sd = nets['LSTM_layers'].state_dict()
    for k in g.files:
        if k.startswith('init/LSTM_layers/'):
            assert np.array_equal(sd[k[len('init/LSTM_layers/'):]].numpy(), g[k]), k
    m = sd['hcgsh.0.mask']
    assert tuple(m.shape) == (24, 24) and 0 < m.mean().item() < 0.5 and tuple(sd['hcgsx.1.mask'].shape) == (16, 24)
