"""Pin the oracle (oracle/*.py) and the product's HOST logic against fixtures produced by running
the reference itself (tests/golden/make_golden.py).  CPU only -- no compute call into the native
library happens here."""
import ast
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN

from oracle import hcgs_oracle, model_oracle


def _load(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)


def _mask_cases():
    g = _load('masks')
    return g, [ast.literal_eval(str(c)) for c in g['cases']]


def test_oracle_masks_bit_exact():
    g, cases = _mask_cases()
    for k, (r, c, bs, dr, seed) in enumerate(cases):
        np.random.seed(seed)
        m = hcgs_oracle.conn_mat(r, c, bs, dr)
        ref = np.unpackbits(g['bits_%d' % k])[:r * c].reshape(r, c)
        assert np.array_equal(m.astype(np.uint8), ref), (r, c, bs, dr, seed)
        # the RNG stream must be at the same position afterwards (draw count and order identical)
        assert np.random.randint(0, 2 ** 31 - 1) == int(g['next_rand_%d' % k][0])


def test_product_masks_bit_exact():
    import hcgs
    g, cases = _mask_cases()
    for k, (r, c, bs, dr, seed) in enumerate(cases):
        np.random.seed(seed)
        m = hcgs.conn_mat_numpy(r, c, bs, dr)
        ref = np.unpackbits(g['bits_%d' % k])[:r * c].reshape(r, c)
        assert m.dtype == np.float32 and np.array_equal(m.astype(np.uint8), ref), (r, c, bs, dr, seed)
        assert np.random.randint(0, 2 ** 31 - 1) == int(g['next_rand_%d' % k][0])


def test_product_mask_entry_points():
    import cgs_base
    import hcgs
    from HCGS import HCGS
    np.random.seed(3)
    a = hcgs.conn_mat(48, 40, [8, 4], [50, 50])
    np.random.seed(3)
    b = HCGS(40, 48, [8, 4], [50, 50]).mask
    assert torch.equal(a.cpu(), b.detach().cpu()) and tuple(b.shape) == (48, 40)
    # lists are not mutated (the reference reverses/pops the caller's lists)
    bs, dr = [8, 4], [50, 50]
    hcgs.conn_mat(16, 16, bs, dr)
    assert bs == [8, 4] and dr == [50, 50]
    with pytest.raises(NotImplementedError):
        hcgs.conn_mat(16, 16, [8], [50], equal_blks_for_input=False)
    np.random.seed(1)
    inner = cgs_base.conn_mat(8, 8, [4], [50], equal_blks_for_input=True)
    assert inner.shape == (8, 8) and inner.sum() == 32
    # degenerate: in=40, blk=64 keeps round(1.625*0.5) ... the float-count quirk gives an all-zero mask
    assert hcgs.conn_mat_numpy(40, 64, [64, 4], [50, 25]).sum() == 0


def test_guided_mask_matches_reference():
    import guided_hcgs
    g = _load('masks')
    w = torch.from_numpy(g['guided_w'])
    m = guided_hcgs.conn_mat_numpy(48, 40, [8, 4], [50, 50], w)
    ref = np.unpackbits(g['guided_bits'])[:48 * 40].reshape(48, 40)
    assert np.array_equal(m.astype(np.uint8), ref)


def test_quantisation_restatement():
    q = _load('quant')
    master, qq = hcgs_oracle.quantize_w(q['w'], 8)
    assert np.array_equal(master, q['master']) and np.array_equal(qq, q['q'])
    assert np.array_equal(hcgs_oracle.quantize_inp(q['x'], 16, 1), q['xq1'])
    assert np.array_equal(hcgs_oracle.quantize_inp(q['x'], 16, 4), q['xq4'])
    assert np.array_equal(hcgs_oracle.prune_mask(q['w'], 70), q['prune70'])
    # balanced=True + if_forward (the in-place save path) differs from balanced=False only by a clamp
    assert np.array_equal(q['qf'], q['q'])
    t = model_oracle.quantize_w_t(torch.from_numpy(q['w']), 8)
    assert np.array_equal(t.numpy(), q['q'])
    t = model_oracle.quantize_inp_t(torch.from_numpy(q['x']), 16, 4)
    assert np.array_equal(t.numpy(), q['xq4'])


def test_posterior_wire_format(tmp_path):
    import posterior_io
    g = _load('posterior')
    cfile = tmp_path / 'c.count'
    cfile.write_text(str(g['counts_text']))
    for mod in (hcgs_oracle, posterior_io):
        counts = mod.load_counts(str(cfile))
        assert counts.dtype == np.float32 and np.array_equal(counts, g['counts'])
    lp_o = hcgs_oracle.log_prior(g['counts'])
    lp_p = posterior_io.log_prior_from_counts(g['counts'])
    assert np.array_equal(lp_o, lp_p)
    post = g['logp'] - lp_o
    assert post.dtype == np.float32 and np.array_equal(post, g['post'])
    assert np.array_equal(model_oracle.posterior(g['logp'], g['counts']), g['post'])
    blob = bytes(g['ark'])
    assert hcgs_oracle.kaldi_mat_bytes(g['post'], 'utt_0001') == blob
    assert posterior_io.mat_record(g['post'], 'utt_0001') == blob
    p = tmp_path / 'o.ark'
    with open(p, 'wb') as fd:
        posterior_io.write_mat('.', fd, g['post'], 'utt_0001')
        posterior_io.write_mat('.', fd, g['post'][:3], 'utt_0002')
    recs = posterior_io.read_ark(str(p))
    assert [k for k, _ in recs] == ['utt_0001', 'utt_0002'] and np.array_equal(recs[1][1], g['post'][:3])
    with pytest.raises(TypeError):
        posterior_io.mat_record(np.zeros((2, 2), dtype=np.int32))


# ---- floating-point path: oracle vs tensors dumped from the reference -------------------------------

def _params(g, stage, net):
    pre = '%s/%s/' % (stage, net)
    return {k[len(pre):]: torch.from_numpy(g[k].copy()) for k in g.files if k.startswith(pre)}


def golden_specs(lstm_over=None, quant=False, quant_inp=0, prune=0):
    q = dict(quant=8 if quant else 0, quant_inp=quant_inp, prune=prune)
    lstm = dict(lay=[24, 24], act=['tanh', 'tanh'], use_bn=[True, True], drop=[0.2, 0.2], hcgs=True, **q)
    lstm.update(lstm_over or {})
    cd = dict(lay=[10], act=['softmax'], use_bn=[False], drop=[0.0], hcgs=True, **q)
    mono = dict(lay=[5], act=['softmax'], use_bn=[False], drop=[0.0], hcgs=False, **q)
    return {'lstm': lstm, 'cd': cd, 'mono': mono}


NETS = {'lstm': 'LSTM_layers', 'cd': 'MLP_layers', 'mono': 'MLP_layers2'}


def _run_oracle_train(g, specs, steps):
    params = {k: _params(g, 'init', n) for k, n in NETS.items()}
    for d in params.values():
        for k, v in d.items():
            if v.dtype == torch.float32 and not k.endswith(('running_mean', 'running_var', 'mask')):
                v.requires_grad_(True)
    T, B, D, _ = [int(v) for v in g['dims']]
    opts = []
    for k in NETS:
        lr = 0.0004 if k == 'mono' else 0.0016
        opts.append(torch.optim.RMSprop([v for v in params[k].values() if v.requires_grad], lr=lr, alpha=0.95,
                                        eps=1e-8, momentum=0.0, centered=False, weight_decay=0.0))
    outs = []
    nl = len(specs['lstm']['lay'])
    for s in range(steps):
        inp = torch.from_numpy(g['inp/%d' % s])
        drops = [torch.from_numpy(g['drop/%d' % (s * nl + i)]) for i in range(nl)]
        for o in opts:
            o.zero_grad()
        out = model_oracle.model_forward(params, specs, inp, D, True, drops)
        out['loss_final'].backward()
        grads = {k: {n: (v.grad.clone() if v.grad is not None else None) for n, v in params[k].items()
                     if v.requires_grad} for k in NETS}
        for o in opts:
            o.step()
        outs.append((out, grads))
    return params, outs


@pytest.mark.parametrize('name,over', [('train_small', None),
                                       ('train_htanh', dict(lay=[20, 28], act=['htanh', 'htanh'], drop=[0.0, 0.1])),
                                       # reference with the `self.bidir = 0` override (:835) lifted, SURVEY f4
                                       ('train_bidir', dict(lay=[24, 16], bidir=True)),
                                       # LayerNorm of h_t inside the time loop (:1092-1095)
                                       ('train_laynorm', dict(lay=[24, 20], use_ln=[True, True])),
                                       ('train_laynorm_nobn', dict(use_ln=[True, False], use_bn=[False, True],
                                                                   act=['tanh', 'relu'], drop=[0.0, 0.2]))])
def test_oracle_train_matches_reference(name, over):
    g = _load(name)
    steps = int(g['dims'][3])
    specs = golden_specs(over)
    params, outs = _run_oracle_train(g, specs, steps)
    for s, (out, grads) in enumerate(outs):
        ref_loss = g['loss/%d' % s]
        assert abs(out['loss_final'].item() - ref_loss[0]) <= 1e-6 * abs(ref_loss[0])
        assert abs(out['err_final'].item() - ref_loss[1]) < 1e-7
        np.testing.assert_allclose(out['out_dnn1'].detach().numpy().reshape(g['out_dnn1/%d' % s].shape), g['out_dnn1/%d' % s], rtol=1e-5, atol=1e-6)
        np.testing.assert_allclose(out['out_dnn2'].detach().numpy(), g['out_dnn2/%d' % s], rtol=1e-5, atol=1e-6)
    # gradients of the last step: identical on every kept weight; the reference's dense gradient
    # on masked-out entries is the documented difference
    out, grads = outs[-1]
    for k, net in NETS.items():
        for n, gr in grads[k].items():
            key = 'grad/%s/%s' % (net, n)
            if key not in g.files:
                assert gr is None or n.endswith('mask')
                continue
            ref = g[key]
            mine = gr.numpy()
            mask = None
            if k == 'lstm' and n.endswith('weight') and n[0] in 'wu' and not n.startswith('bn'):
                mask = params['lstm'][('hcgsx.%s.mask' if n[0] == 'w' else 'hcgsh.%s.mask') % n.split('.')[1]]
            if k == 'cd' and n == 'wx.0.weight':
                mask = params['cd']['hcgs.0.mask']
            if mask is not None:
                ref = ref * mask.detach().numpy()
            np.testing.assert_allclose(mine, ref, rtol=2e-4, atol=1e-7, err_msg=key)
    # parameters after the optimizer steps (kept entries)
    for k, net in NETS.items():
        for n, v in params[k].items():
            ref = g['final/%s/%s' % (net, n)]
            if n.endswith('num_batches_tracked'):
                continue
            mine = v.detach().numpy()
            if n.endswith('weight') and (n[0] in 'wu') and not n.startswith('bn') and k != 'mono':
                mk = {'lstm': ('hcgsx.%s.mask' if n[0] == 'w' else 'hcgsh.%s.mask') % n.split('.')[1], 'cd': 'hcgs.0.mask'}[k]
                m = params[k][mk].detach().numpy()
                mine, ref = mine * m, ref * m
            np.testing.assert_allclose(mine, ref, rtol=1e-4, atol=2e-6, err_msg='final ' + n)


def test_oracle_quantised_train_matches_reference():
    g = _load('train_quant')
    specs = golden_specs(quant=True)
    params, outs = _run_oracle_train(g, specs, 1)
    out, grads = outs[0]
    ref_loss = g['loss/0']
    assert abs(out['loss_final'].item() - ref_loss[0]) <= 2e-6 * abs(ref_loss[0])
    np.testing.assert_allclose(out['out_dnn2'].detach().numpy(), g['out_dnn2/0'], rtol=1e-5, atol=2e-6)
    for n in ('ufh.0.weight', 'wcx.1.weight'):
        mask = params['lstm'][('hcgsx.%s.mask' if n[0] == 'w' else 'hcgsh.%s.mask') % n.split('.')[1]].detach().numpy()
        np.testing.assert_allclose(grads['lstm'][n].numpy(), g['grad/LSTM_layers/' + n] * mask, rtol=3e-4, atol=1e-7)


def test_oracle_quantised_activations_and_pruning_match_reference():
    """The shipped QuantExp configuration: 8-bit weights, 16-bit Quantize_inp of every layer input and of h_t at
    every time step (the stacked output holds the quantised h for all steps but the last), magnitude pruning 70."""
    g = _load('train_quant_inp')
    specs = golden_specs(quant=True, quant_inp=16, prune=70)
    params, outs = _run_oracle_train(g, specs, 1)
    out, grads = outs[0]
    ref_loss = g['loss/0']
    assert abs(out['loss_final'].item() - ref_loss[0]) <= 2e-6 * abs(ref_loss[0])
    np.testing.assert_allclose(out['out_dnn2'].detach().numpy(), g['out_dnn2/0'], rtol=1e-5, atol=2e-6)
    # the reference's gradients are dense; the oracle (like the product) carries the masks in the graph, so the
    # comparison is on the kept entries: HCGS mask x pruning mask of the initial weights
    def kept(net, n, hmask):
        w0 = torch.from_numpy(g['init/%s/%s' % (net, n)]) * hmask
        return (hmask * model_oracle.prune_t(w0, 70)[1]).numpy()
    for n in ('ufh.0.weight', 'wcx.1.weight'):
        hm = params['lstm'][('hcgsx.%s.mask' if n[0] == 'w' else 'hcgsh.%s.mask') % n.split('.')[1]].detach()
        m = kept('LSTM_layers', n, hm)
        assert 0 < m.sum() < m.size
        np.testing.assert_allclose(grads['lstm'][n].numpy() * m, g['grad/LSTM_layers/' + n] * m, rtol=3e-4, atol=2e-6)
    m = kept('MLP_layers', 'wx.0.weight', params['cd']['hcgs.0.mask'].detach())
    np.testing.assert_allclose(grads['cd']['wx.0.weight'].numpy() * m, g['grad/MLP_layers/wx.0.weight'] * m,
                               rtol=3e-4, atol=2e-4 * np.abs(g['grad/MLP_layers/wx.0.weight']).max())


@pytest.mark.parametrize('name', ['valid_small', 'eval_forward'])
def test_oracle_eval_matches_reference(name):
    g = _load(name)
    params = {k: _params(g, 'init', n) for k, n in NETS.items()}
    T, B, D, _ = [int(v) for v in g['dims']]
    inp = torch.from_numpy(g['inp/0'])
    with torch.no_grad():
        out = model_oracle.model_forward(params, golden_specs(), inp, D, False, None)
    np.testing.assert_allclose(out['out_dnn2'].numpy(), g['out_dnn2/0'], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(out['out_dnn1'].numpy().reshape(g['out_dnn1/0'].shape), g['out_dnn1/0'], rtol=1e-5, atol=1e-6)
    if 'loss/0' in g.files:
        assert abs(out['loss_final'].item() - g['loss/0'][0]) <= 1e-6 * abs(g['loss/0'][0])
