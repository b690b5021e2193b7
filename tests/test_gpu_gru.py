"""GRU / liGRU / minimalGRU / RNN on the B200: the plugin classes through the C ABI (csrc/gru.cu) against
 (a) tensors dumped from the reference's own dense classes (all-ones masks) and
 (b) the CPU oracle with HCGS masks on the recurrent and input matrices, more sequences than one cluster holds.
fp32: loss / outputs rel <= 1e-5, gradients 1e-4 of the tensor's max."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from helpers import NETS, golden_state, load_golden, make_graph  # noqa: E402
from oracle import model_oracle  # noqa: E402
from test_gpu_model import _ReplayBernoulli, _state_to_cpu, close  # noqa: E402
from test_gru_family_cpu import FAMILY, build_family_nets, family_specs  # noqa: E402


@pytest.mark.parametrize('name,prefix,cls,over,spec', FAMILY)
def test_dense_cells_match_reference_fixture(name, prefix, cls, over, spec):
    from b200 import native
    from b200.graph import make_optimizer
    g = load_golden(name)
    T, B, D, steps = [int(v) for v in g['dims']]
    nets, sections = build_family_nets(prefix, cls, over, device='cuda')
    for n, net in nets.items():
        net.load_state_dict(golden_state(g, 'init', n))
    graph = make_graph(nets, D)
    opts = {n: make_optimizer(nets[n], sec) for n, sec in zip(nets, sections)}
    nl = nets['LSTM_layers'].N_lay
    drops = [torch.from_numpy(g['drop/%d' % i]) for i in range(steps * nl)]
    native.reset_launch_count()
    with _ReplayBernoulli(drops):
        for s in range(steps):
            out = graph.forward(torch.from_numpy(g['inp/%d' % s]).cuda(), T, B, 'train', ['out_dnn2'])
            for o in opts.values():
                o.zero_grad()
            out['loss_final'].backward()
            ref = g['loss/%d' % s]
            assert abs(out['loss_final'].item() - ref[0]) <= 1e-5 * abs(ref[0]), (s, out['loss_final'].item(), ref[0])
            close(out['out_dnn1'].reshape(T * B, -1), g['out_dnn1/%d' % s], 1e-5, 'out_dnn1')
            close(out['out_dnn2'], g['out_dnn2/%d' % s], 1e-5, 'out_dnn2')
            if s == steps - 1:
                seen = 0
                for k, p in nets['LSTM_layers'].named_parameters():
                    key = 'grad/LSTM_layers/' + k
                    if key not in g.files:
                        assert p.grad is None, k
                        continue
                    close(p.grad, g[key], 1e-4, key)
                    seen += 1
                assert seen >= 4 * len(model_oracle.CELL_GATES[spec['cell']])
            for o in opts.values():
                o.step()
    assert native.LAUNCHES.get('hcgs_gru_fwd', 0) == steps * nl and native.LAUNCHES.get('hcgs_gru_bwd', 0) == steps * nl
    sd = nets['LSTM_layers'].state_dict()
    for k, v in sd.items():
        ref = g['final/LSTM_layers/' + k]
        if k.endswith('num_batches_tracked'):
            assert int(v.item()) == int(ref)
            continue
        close(v, ref, 1e-4, 'final ' + k)


@pytest.mark.parametrize('prefix,cls,cell,act', [('gru', 'GRU', 'gru', 'tanh'), ('ligru', 'liGRU', 'ligru', 'relu'),
                                                 ('minimalgru', 'minimalGRU', 'minimalgru', 'tanh'),
                                                 ('rnn', 'RNN', 'rnn', 'relu')])
@pytest.mark.parametrize('H,T,B,blk,sp', [(72, 9, 11, '8,4', '50,50'), (550, 20, 8, '32,4', '75,62.5')])
def test_hcgs_masked_cells_match_oracle(prefix, cls, cell, act, H, T, B, blk, sp):
    """block-sparse recurrent and input matrices (not in the reference: parity target is the oracle's masked-dense
    restatement), ragged H = 550 and a batch that spans two clusters"""
    D = 40
    over = {prefix + '_lay': '%d,%d' % (H, H), prefix + '_act': '%s,%s' % (act, act), prefix + '_hcgs': 'True',
            'hcgsx_block': blk, 'hcgsx_sparse': sp, 'hcgsh_block': blk, 'hcgsh_sparse': sp}
    nets, _ = build_family_nets(prefix, cls, over, D=D, seed=7, device='cuda')
    params = _state_to_cpu(nets)
    gen = torch.Generator().manual_seed(2)
    inp = torch.zeros(T, B, D + 2)
    inp[:, :, :D] = torch.randn(T, B, D, generator=gen)
    inp[:, :, D] = torch.randint(0, 10, (T, B), generator=gen).float()
    inp[:, :, D + 1] = torch.randint(0, 5, (T, B), generator=gen).float()
    inp[T - 3:, B - 1, :] = 0
    drops = [torch.bernoulli(torch.full((B, H), 0.8), generator=gen) for _ in range(2)]
    for d in params.values():
        for k, v in d.items():
            if v.dtype == torch.float32 and not k.endswith(('running_mean', 'running_var', 'mask')):
                v.requires_grad_(True)
    specs = family_specs(dict(cell=cell, act=[act, act], use_bn=[True, True], drop=[0.2, 0.2]))
    specs['lstm'].update(lay=[H, H], hcgs=True)
    ref = model_oracle.model_forward(params, specs, inp, D, True, drops)
    ref['loss_final'].backward()
    with _ReplayBernoulli(drops):
        out = make_graph(nets, D).forward(inp.cuda(), T, B, 'train', ['out_dnn2'])
    out['loss_final'].backward()
    torch.cuda.synchronize()
    assert abs(out['loss_final'].item() - ref['loss_final'].item()) <= 1e-5 * abs(ref['loss_final'].item())
    close(out['out_dnn1'].reshape(T, B, -1), ref['out_dnn1'], 1e-5, 'h')
    close(out['out_dnn2'], ref['out_dnn2'], 1e-5, 'log-posteriors')
    rec = nets['LSTM_layers']
    for name, p in rec.named_parameters():
        rg = params['lstm'][name].grad
        if rg is None:
            continue
        if name[0] in 'wu' and name.endswith('.weight'):
            i = int(name.split('.')[1])
            keep = ((rec.hcgsx if name[0] == 'w' else rec.hcgsh)[i].mask.detach().cpu() != 0).float()
            close(p.grad.cpu() * keep, rg * keep, 1e-4, name)       # kept entries
        else:
            close(p.grad, rg, 1e-4, name)


def test_gru_eval_forward_uses_running_statistics_and_scalar_drop():
    over = {'gru_lay': '40,40', 'gru_hcgs': 'True', 'hcgsx_block': '8,4', 'hcgsx_sparse': '50,50', 'hcgsh_block': '8,4',
            'hcgsh_sparse': '50,50'}
    nets, _ = build_family_nets('gru', 'GRU', over, to_do='forward', D=16, seed=3, device='cuda')
    rec = nets['LSTM_layers']
    with torch.no_grad():
        for name, buf in rec.named_buffers():
            if name.endswith('running_var'):
                buf.uniform_(0.5, 2.0)
            elif name.endswith('running_mean'):
                buf.normal_(0, 0.3)
    params = _state_to_cpu(nets)
    x = torch.randn(13, 5, 16, generator=torch.Generator().manual_seed(1))
    spec = dict(cell='gru', lay=[40, 40], act=['tanh', 'tanh'], use_bn=[True, True], drop=[0.2, 0.2], hcgs=True)
    with torch.no_grad():
        want = model_oracle.gated_rnn_forward(params['lstm'], spec, x, None, False)
        got = rec(x.cuda())
    close(got, want, 1e-5, 'eval h')
