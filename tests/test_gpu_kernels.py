"""Kernel-level parity (B200): every C-ABI entry point against a plain fp32 restatement on the CPU.

fp32 tolerance: relative 1e-5 against the magnitude of the result (summation order differs
between a tiled GPU kernel and torch's CPU GEMM; inputs are O(1))."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import hcgs_oracle, model_oracle  # noqa: E402


def dev():
    return torch.device('cuda:0')


def close(a, b, rtol=1e-5, what=''):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    scale = max(b.abs().max().item(), 1e-30)
    err = (a - b).abs().max().item()
    assert err <= rtol * scale + 1e-7, '%s: max err %.3e vs scale %.3e' % (what, err, scale)


def random_mask(rows, cols, blocks, drops, seed):
    np.random.seed(seed)
    return torch.from_numpy(hcgs_oracle.conn_mat(rows, cols, blocks, drops))


CASES = [(550, 40, [32, 4], [75, 62.5]), (550, 550, [32, 4], [75, 62.5]), (70, 45, [16, 4], [50, 50]),
         (48, 550, [64, 4], [50, 25]), (33, 65, [8, 2], [50, 50])]


@pytest.mark.parametrize('n_out,n_in,blocks,drops', CASES)
@pytest.mark.parametrize('n_w', [1, 4])
def test_block_sparse_gemm_family(n_out, n_in, blocks, drops, n_w):
    from b200 import native
    torch.manual_seed(n_out + n_in + n_w)
    mask = random_mask(n_out, n_in, blocks, drops, 7)
    plan = native.Plan(mask)
    assert plan.nnz == int(mask.sum().item())
    M = 203
    ws = [(torch.randn(n_out, n_in) * mask) for _ in range(n_w)]
    bias = [torch.randn(n_out) for _ in range(n_w)]
    # x is a column slice of a wider tensor (the reference slices features out of `inp`)
    big = torch.randn(M, n_in + 3)
    x = big[:, 1:1 + n_in]
    ref = torch.cat([x @ w.t() + b for w, b in zip(ws, bias)], 1)
    d = dev()
    wd, bd = [w.to(d) for w in ws], [b.to(d) for b in bias]
    xd = big.to(d)[:, 1:1 + n_in]
    y = torch.empty(M, n_w * n_out, device=d)
    native.bsgemm_nt(plan, wd, xd, y, bd)
    close(y, ref, what='nt')
    dy = torch.randn(M, n_w * n_out)
    ref_dx = sum(dy[:, g * n_out:(g + 1) * n_out] @ ws[g] for g in range(n_w))
    dx = torch.empty(M, n_in, device=d)
    native.bsgemm_nn(plan, wd, dy.to(d), dx)
    close(dx, ref_dx, what='nn')
    dws = [torch.zeros(n_out, n_in, device=d) for _ in range(n_w)]
    native.bs_sddmm(plan, dy.to(d), xd, dws, mask.to(d))
    for g in range(n_w):
        close(dws[g], (dy[:, g * n_out:(g + 1) * n_out].t() @ x) * mask, what='sddmm')
    outs = [torch.empty(n_out, device=d) for _ in range(n_w)]
    native.colsum(dy.to(d), n_w, n_out, outs)
    for g in range(n_w):
        close(outs[g], dy[:, g * n_out:(g + 1) * n_out].sum(0), what='colsum')
    plan.destroy()


def test_gemm_large_m_and_dense_plan():
    from b200 import native
    torch.manual_seed(1)
    n_out, n_in, M = 96, 429, 3000
    plan = native.Plan(torch.ones(n_out, n_in))
    assert plan.nz_tiles == plan.total_tiles
    w, x, dy = torch.randn(n_out, n_in) * 0.1, torch.randn(M, n_in), torch.randn(M, n_out)
    d = dev()
    y = torch.empty(M, n_out, device=d)
    native.bsgemm_nt(plan, [w.to(d)], x.to(d), y)
    close(y, x @ w.t(), what='nt dense')
    dw = [torch.zeros(n_out, n_in, device=d)]
    native.bs_sddmm(plan, dy.to(d), x.to(d), dw, None)
    close(dw[0], dy.t() @ x, what='sddmm dense split-M')


@pytest.mark.parametrize('M,G,C', [(200, 4, 24), (4097, 4, 550), (130, 1, 1024), (64, 1, 5)])
def test_batch_norm_fwd_bwd(M, G, C):
    from b200 import native
    torch.manual_seed(M + C)
    d = dev()
    x = (torch.randn(M, G * C) * 2 + 0.5)
    gam = [torch.rand(C) + 0.5 for _ in range(G)]
    bet = [torch.randn(C) for _ in range(G)]
    rm = [torch.randn(C) * 0.1 for _ in range(G)]
    rv = [torch.rand(C) + 0.5 for _ in range(G)]
    dy = torch.randn(M, G * C)
    # reference per group with autograd
    ys, dxs, dgs, dbs, rms, rvs = [], [], [], [], [], []
    for g in range(G):
        xg = x[:, g * C:(g + 1) * C].clone().requires_grad_(True)
        ga, be = gam[g].clone().requires_grad_(True), bet[g].clone().requires_grad_(True)
        r1, r2 = rm[g].clone(), rv[g].clone()
        yg = F.batch_norm(xg, r1, r2, ga, be, True, 0.05, 1e-5)
        yg.backward(dy[:, g * C:(g + 1) * C])
        ys.append(yg.detach()); dxs.append(xg.grad); dgs.append(ga.grad); dbs.append(be.grad)
        rms.append(r1); rvs.append(r2)
    xd, y = x.to(d), torch.empty(M, G * C, device=d)
    gd, bd = [t.to(d) for t in gam], [t.to(d) for t in bet]
    rmd, rvd = [t.to(d) for t in rm], [t.to(d) for t in rv]
    sm, si = torch.empty(G * C, device=d), torch.empty(G * C, device=d)
    ws = torch.empty(2 * G * C * 64, device=d)
    native.bn_fwd(xd, y, G, C, gd, bd, rmd, rvd, 1e-5, 0.05, True, sm, si, ws)
    close(y, torch.cat(ys, 1), what='bn y')
    for g in range(G):
        close(rmd[g], rms[g], what='running_mean')
        close(rvd[g], rvs[g], what='running_var')
    dx = torch.empty(M, G * C, device=d)
    dg, db = [torch.empty(C, device=d) for _ in range(G)], [torch.empty(C, device=d) for _ in range(G)]
    native.bn_bwd(dy.to(d), xd, G, C, gd, sm, si, dx, dg, db, ws)
    close(dx, torch.cat(dxs, 1), rtol=2e-5, what='bn dx')
    for g in range(G):
        close(dg[g], dgs[g], rtol=2e-5, what='dgamma')
        close(db[g], dbs[g], rtol=2e-5, what='dbeta')
    # eval mode uses the running statistics
    y2 = torch.empty(M, G * C, device=d)
    native.bn_fwd(xd, y2, G, C, gd, bd, rmd, rvd, 1e-5, 0.05, False, None, None, None)
    ref2 = torch.cat([F.batch_norm(x[:, g * C:(g + 1) * C], rms[g], rvs[g], gam[g], bet[g], False, 0.05, 1e-5)
                      for g in range(G)], 1)
    close(y2, ref2, what='bn eval')


def _lstm_layer_oracle(x, us, pre, drop, act, need_grad=True):
    """One layer of the recurrence only (pre-activations given), with autograd."""
    T, B, H4 = pre.shape
    H = H4 // 4
    h = torch.zeros(B, H)
    c = h
    hs = []
    for k in range(T):
        zf = pre[k, :, 0:H] + h @ us[0].t()
        zi = pre[k, :, H:2 * H] + h @ us[1].t()
        zo = pre[k, :, 2 * H:3 * H] + h @ us[2].t()
        zc = pre[k, :, 3 * H:] + h @ us[3].t()
        c = torch.sigmoid(zi) * model_oracle.act(act, zc) * drop + torch.sigmoid(zf) * c
        h = torch.sigmoid(zo) * model_oracle.act(act, c)
        hs.append(h)
    return torch.stack(hs)


@pytest.mark.parametrize('H,B,T,blocks,drops,act,cl', [
    (24, 3, 7, [8, 4], [50, 50], 'tanh', 0),
    (24, 3, 7, [8, 4], [50, 50], 'tanh', 2),
    (40, 11, 9, [8, 4], [50, 50], 'htanh', 4),
    (550, 8, 12, [32, 4], [75, 62.5], 'tanh', 8),
    (550, 13, 6, [32, 4], [75, 62.5], 'relu', 0),
    (70, 5, 8, [16, 2], [50, 50], 'sigmoid', 1),
    (1024, 9, 5, [32, 4], [75, 62.5], 'tanh', 0),
])
def test_recurrent_kernels(H, B, T, blocks, drops, act, cl):
    from b200 import native
    torch.manual_seed(H + B + T)
    mask = random_mask(H, H, blocks, drops, 3)
    plan = native.Plan(mask, recurrent=True, cluster_size=cl)
    assert plan.cluster_size in (1, 2, 4, 8, 16) and plan.w_in_smem
    us = [(torch.randn(H, H) * (1.5 / np.sqrt(max(mask.sum(1).mean().item(), 1.0))) * mask).requires_grad_(True)
          for _ in range(4)]
    pre = torch.randn(T, B, 4 * H).requires_grad_(True)
    drop = torch.bernoulli(torch.full((B, H), 0.8))
    href = _lstm_layer_oracle(None, us, pre, drop, act)
    dh = torch.randn(T, B, H)
    href.backward(dh)
    d = dev()
    ud = [u.detach().to(d) for u in us]
    fp = torch.empty(plan.fwd_pack_floats, device=d)
    bp = torch.empty(max(plan.bwd_pack_floats, 1), device=d)
    native.lstm_pack(plan, ud, fp, bp)
    h = torch.empty(T, B, H, device=d)
    c = torch.empty(T, B, H, device=d)
    gates = torch.empty(T, B, 4 * H, device=d)
    code = native.ACT_CODES[act]
    native.lstm_fwd(plan, fp, pre.detach().to(d), drop.to(d), 1.0, code, h, c, gates)
    torch.cuda.synchronize()
    close(h, href, what='h')
    dpre = torch.empty(T, B, 4 * H, device=d)
    native.lstm_bwd(plan, bp, dh.to(d), gates, c, drop.to(d), 1.0, code, dpre)
    torch.cuda.synchronize()
    close(dpre, pre.grad, rtol=3e-5, what='dpre')
    # weight gradient through the SDDMM on the same plan: dU_g = sum_t dpre_g[t]^T h[t-1]
    dus = [torch.zeros(H, H, device=d) for _ in range(4)]
    M = T * B
    native.bs_sddmm(plan, dpre.view(M, 4 * H)[B:], h.view(M, H)[:M - B], dus, mask.to(d))
    for g in range(4):
        close(dus[g], us[g].grad * mask, rtol=5e-5, what='dU')
    # eval variant: scalar drop, no saved state
    h2 = torch.empty(T, B, H, device=d)
    native.lstm_fwd(plan, fp, pre.detach().to(d), None, 0.8, code, h2, None, None)
    href2 = _lstm_layer_oracle(None, [u.detach() for u in us], pre.detach(), torch.tensor([0.8]), act)
    close(h2, href2, what='h eval')
    plan.destroy()


@pytest.mark.parametrize('H,B,T,blocks,drops', [(24, 75, 9, [8, 4], [50, 50]), (550, 100, 6, [32, 4], [75, 62.5]),
                                                 (550, 8, 6, [32, 4], [75, 62.5])])
def test_recurrent_quantised_h_scale_spans_clusters(H, B, T, blocks, drops):
    """act_bits: Quantize_inp of h_t scales by max |h_t| over the WHOLE [B, H] tensor (quantized_modules.py:104-108).
    A cluster of the kernel holds 8 sequences, so for B > 8 the clusters exchange their maxima at every step.  With 4
    bits the stored h_t (t < T - 1: the reference's `.data` swap) must lie on the grid j * var_t / 8 of the GLOBAL
    var_t = max over all sequences, and that maximum is itself a grid point; the sequences are scaled so that the
    per-cluster maxima differ by up to 20x (a per-cluster scale would put the quiet clusters off the global grid)."""
    from b200 import native
    torch.manual_seed(B)
    mask = random_mask(H, H, blocks, drops, 3)
    plan = native.Plan(mask, recurrent=True)
    d = dev()
    us = [(torch.randn(H, H) * (1.5 / np.sqrt(max(mask.sum(1).mean().item(), 1.0))) * mask).to(d) for _ in range(4)]
    pre = torch.randn(T, B, 4 * H) * torch.linspace(0.05, 1.0, B).view(1, B, 1)
    fp = torch.empty(plan.fwd_pack_floats, device=d)
    native.lstm_pack(plan, us, fp, None)
    h = torch.empty(T, B, H, device=d)
    native.lstm_fwd(plan, fp, pre.to(d), None, 1.0, native.ACT_CODES['tanh'], h, None, None, act_bits=4)
    torch.cuda.synchronize()
    h = h.cpu()
    assert torch.isfinite(h).all()
    for t in range(T - 1):
        var = h[t].abs().max()
        assert var > 0
        r = h[t].abs().double() * 8 / var.double()
        off = (r - r.round()).abs().max().item()
        assert off <= 1e-4, 'step %d: stored h is not on the grid of the whole-batch scale (off by %.3e levels)' % (t, off)
        if B > 8:
            quiet = h[t, :8].abs().max()
            assert quiet < 0.7 * var                    # the first cluster alone would have chosen another scale
    # the last step is never re-quantised (nothing consumes it inside the loop)
    r = h[T - 1].abs().double() * 8 / h[T - 1].abs().max().double()
    assert (r - r.round()).abs().max().item() > 1e-2
    if H == 550:
        big = 8 * 19                                     # 19 clusters of 8 CTAs do not fit 148 SMs
        pre2 = torch.zeros(2, big, 4 * H, device=d)
        with pytest.raises(native.NativeError, match='co-resident'):
            native.lstm_fwd(plan, fp, pre2, None, 1.0, native.ACT_CODES['tanh'], torch.empty(2, big, H, device=d), None,
                            None, act_bits=4)
    plan.destroy()


def test_recurrent_dense_masks():
    """No HCGS (lstm_hcgs=False): the all-ones mask is just a very wide block list.  It runs while the
    compressed state fits the cluster's shared memory and is refused loudly (no silent fallback) beyond."""
    from b200 import native
    H, B, T = 256, 4, 4
    torch.manual_seed(0)
    plan = native.Plan(torch.ones(H, H), recurrent=True)
    us = [torch.randn(H, H) * 0.05 for _ in range(4)]
    pre = torch.randn(T, B, 4 * H)
    href = _lstm_layer_oracle(None, us, pre, torch.ones(B, H), 'tanh')
    d = dev()
    fp = torch.empty(plan.fwd_pack_floats, device=d)
    native.lstm_pack(plan, [u.to(d) for u in us], fp, None)
    h = torch.empty(T, B, H, device=d)
    native.lstm_fwd(plan, fp, pre.to(d), None, 1.0, native.ACT_CODES['tanh'], h, None, None)
    close(h, href, what='h dense')
    with pytest.raises(native.NativeError, match='shared memory'):
        native.Plan(torch.ones(1024, 1024), recurrent=True)


def test_softmax_nll_and_activations():
    from b200 import native
    torch.manual_seed(5)
    d = dev()
    M, P = 517, 1944
    z = torch.randn(M, P) * 3
    lp = torch.randn(P)
    out = torch.empty(M, P, device=d)
    native.logsoftmax_fwd(z.to(d), out, None)
    ref = F.log_softmax(z, 1)
    close(out, ref, what='logsoftmax')
    out2 = torch.empty(M, P, device=d)
    native.logsoftmax_fwd(z.to(d), out2, lp.to(d))
    assert (out2.cpu() - (ref - lp)).abs().max() < 1e-5
    dl = torch.randn(M, P)
    zz = z.clone().requires_grad_(True)
    F.log_softmax(zz, 1).backward(dl)
    dz = torch.empty(M, P, device=d)
    native.logsoftmax_bwd(dl.to(d), out, dz)
    close(dz, zz.grad, rtol=2e-5, what='logsoftmax bwd')
    labels = torch.randint(0, P, (M,)).float()
    inp = torch.zeros(M, 3)
    inp[:, 2] = labels
    stats = torch.zeros(2, device=d)
    dzz = torch.empty(M, P, device=d)
    native.nll_fwd_bwd(out, inp.to(d)[:, 2:], 3, 1.0 / M, dzz, stats)
    zz = z.clone().requires_grad_(True)
    lsm = F.log_softmax(zz, 1)
    loss = F.nll_loss(lsm, labels.long())
    loss.backward()
    err = (lsm.argmax(1) != labels.long()).float().sum()
    s = stats.cpu()
    assert abs(s[0].item() / M - loss.item()) < 1e-5 * abs(loss.item())
    assert s[1].item() == err.item()
    close(dzz, zz.grad, rtol=2e-5, what='nll dlogits')
    for name in ('relu', 'tanh', 'htanh', 'sigmoid', 'leaky_relu', 'elu', 'linear'):
        x = (torch.randn(1000) * 2).requires_grad_(True)
        m = torch.bernoulli(torch.full((1000,), 0.85)) / 0.85
        y = model_oracle.act(name, x) * m
        g = torch.randn(1000)
        y.backward(g)
        yd = torch.empty(1000, device=d)
        native.act_fwd(x.detach().to(d), yd, m.to(d), native.ACT_CODES[name])
        close(yd, y, what='act ' + name)
        dx = torch.empty(1000, device=d)
        native.act_bwd(g.to(d), yd, m.to(d), dx, native.ACT_CODES[name])
        close(dx, x.grad, what='act bwd ' + name)


def test_quantisation_kernels_bit_exact():
    from b200 import native
    from helpers import load_golden
    q = load_golden('quant')
    d = dev()
    master = torch.from_numpy(q['w'].copy()).to(d)
    out = torch.empty_like(master)
    native.quantize_w(master, out, 8)
    assert np.array_equal(master.cpu().numpy(), q['master']) and np.array_equal(out.cpu().numpy(), q['q'])
    x = torch.from_numpy(q['x'].copy()).to(d)
    y = torch.empty_like(x)
    sc = torch.zeros(2, device=d)
    native.quantize_act(x, y, 16, 1, sc)
    assert np.array_equal(y.cpu().numpy(), q['xq1'])
    native.quantize_act(x, y, 16, 4, sc)
    assert np.array_equal(y.cpu().numpy(), q['xq4'])
    z = torch.zeros(4, 5, device=d)
    native.quantize_act(z, z, 16, 1, sc)          # all-zero input: skipped (var == 0)
    assert z.abs().sum().item() == 0
    w = torch.from_numpy(q['w'].copy()).to(d)
    m = torch.from_numpy(q['prune70'].copy()).to(d)
    native.mask_apply([w], m)
    assert np.array_equal(w.cpu().numpy(), q['w'] * q['prune70'])


@pytest.mark.parametrize('kind', ['sgd', 'sgd_mom', 'rmsprop', 'rmsprop_mom', 'adam'])
def test_optimizer_steps_match_torch(kind):
    from b200 import native
    torch.manual_seed(3)
    d = dev()
    p0 = torch.randn(5000)
    p = torch.nn.Parameter(p0.clone())
    if kind == 'sgd':
        opt, k, hp = torch.optim.SGD([p], lr=0.08), 'sgd', [0.08, 0, 0, 0, 0, 0, 0, 0]
    elif kind == 'sgd_mom':
        opt = torch.optim.SGD([p], lr=0.05, momentum=0.9, weight_decay=1e-3, nesterov=True)
        k, hp = 'sgd', [0.05, 0.9, 0.0, 0, 1e-3, 1.0, 0, 0]
    elif kind == 'rmsprop':
        opt = torch.optim.RMSprop([p], lr=0.0016, alpha=0.95, eps=1e-8)
        k, hp = 'rmsprop', [0.0016, 0.95, 0.0, 1e-8, 0, 0, 0, 0]
    elif kind == 'rmsprop_mom':
        opt = torch.optim.RMSprop([p], lr=0.001, alpha=0.9, eps=1e-8, momentum=0.5, weight_decay=1e-4)
        k, hp = 'rmsprop', [0.001, 0.9, 0.5, 1e-8, 1e-4, 0, 0, 0]
    else:
        opt = torch.optim.Adam([p], lr=0.002, betas=(0.9, 0.999), eps=1e-8)
        k, hp = 'adam', [0.002, 0.9, 0.999, 1e-8, 0, 0, 0, 0]
    pd, s1, s2 = p0.clone().to(d), torch.zeros(5000, device=d), torch.zeros(5000, device=d)
    for step in range(1, 4):
        g = torch.randn(5000)
        p.grad = g.clone()
        opt.step()
        hp[7] = float(step)
        native.opt_step(k, pd, g.to(d), s1, s2, hp)
        close(pd, p.detach(), rtol=2e-6, what='%s step %d' % (kind, step))


@pytest.mark.parametrize('kind,kw', [
    ('sgd', dict(lr=0.08)),
    ('sgd', dict(lr=0.05, momentum=0.9, weight_decay=1e-3, nesterov=True)),
    ('rmsprop', dict(lr=0.0016, alpha=0.95, eps=1e-8)),
    ('rmsprop', dict(lr=0.001, alpha=0.9, eps=1e-8, momentum=0.5, weight_decay=1e-4)),
    ('adam', dict(lr=0.002, betas=(0.9, 0.999), eps=1e-8)),
])
def test_fused_optimizer_matches_torch_optim(kind, kw):
    """b200/optim.py (hcgs_opt_step_multi, one launch per 48 tensors) == torch.optim on the same parameters, incl. the
    per-parameter state names, over three steps and more tensors than one launch holds."""
    from b200.optim import FusedOptimizer
    torch.manual_seed(5)
    d = dev()
    shapes = [(7, 5), (1,), (33,), (4, 1944)] * 13 + [(550, 64)]     # 53 tensors: two launches
    ref = [torch.nn.Parameter(torch.randn(*s, device=d)) for s in shapes]
    ours = [torch.nn.Parameter(r.detach().clone()) for r in ref]
    cls = {'sgd': torch.optim.SGD, 'rmsprop': torch.optim.RMSprop, 'adam': torch.optim.Adam}[kind]
    o_ref, o_our = cls(ref, **kw), FusedOptimizer(ours, kind, **kw)
    for step in range(3):
        for r, o in zip(ref, ours):
            g = torch.randn_like(r)
            r.grad, o.grad = g.clone(), g.clone()
        if step == 1:
            ours[3].grad = None      # a parameter without a gradient is skipped, like torch.optim does
            ref[3].grad = None
        o_ref.step()
        o_our.step()
        for i, (r, o) in enumerate(zip(ref, ours)):
            close(o.detach(), r.detach(), rtol=3e-6, what='%s tensor %d step %d' % (kind, i, step))
    names = {'sgd': ['momentum_buffer'] if kw.get('momentum') else [], 'rmsprop': ['square_avg'],
             'adam': ['exp_avg', 'exp_avg_sq']}[kind]
    for n in names:
        close(o_our.state[ours[0]][n], o_ref.state[ref[0]][n], rtol=3e-6, what='state ' + n)


@pytest.mark.parametrize('M,P', [(517, 1944), (300, 48), (70, 3480), (9, 4)])
def test_logsoftmax_bwd_fused(M, P):
    """hcgs_logsoftmax_bwd_fused = logsoftmax_bwd + colsum + cast_bf16 of the unfused path (same arithmetic)."""
    from b200 import native
    d = torch.device('cuda', 0)
    torch.manual_seed(M + P)
    logp = torch.log_softmax(torch.randn(M, P) * 3, 1).to(d)
    dlogp = (torch.randn(M, P) * (torch.rand(M, P) < 0.1)).to(d)      # mostly zeros, like the NLL gradient
    ld = native.pad8(P)
    dzb = torch.full((M, ld), 7.0, device=d, dtype=torch.bfloat16)
    db = torch.zeros(P, device=d)
    native.logsoftmax_bwd_fused(dlogp, logp, dzb, db)
    dz = torch.empty(M, P, device=d)
    native.logsoftmax_bwd(dlogp, logp, dz)
    torch.cuda.synchronize()
    ref = dlogp.double() - logp.double().exp() * dlogp.double().sum(1, keepdim=True)
    assert (dz.double() - ref).abs().max().item() <= 1e-5 * max(ref.abs().max().item(), 1e-30)
    want = dz.to(torch.bfloat16).float()
    got = dzb[:, :P].float()
    # same fp32 value rounded to bf16; __expf vs expf may flip a rounding of the last bf16 bit
    assert (got - want).abs().max().item() <= 2 ** -7 * max(want.abs().max().item(), 1e-30)
    assert (got - want).abs().mean().item() <= 1e-4 * max(want.abs().max().item(), 1e-30)
    if ld > P:
        assert dzb[:, P:].float().abs().max().item() == 0
    assert (db.double() - ref.sum(0)).abs().max().item() <= 1e-4 * max(ref.sum(0).abs().max().item(), 1e-30)


@pytest.mark.parametrize('M,P', [(517, 1944), (300, 48), (70, 3480), (9, 4)])
def test_logsoftmax_bwd_fused_onehot_equals_dense_gradient(M, P):
    """The NLL gradient kept as one entry per row (hcgs_logsoftmax_bwd_fused_onehot) against the same gradient scattered
    into a dense matrix and pushed through hcgs_logsoftmax_bwd_fused: identical bf16 dz (same arithmetic per element), same
    bias gradient; a label outside [0, P) (ignore_index) gives a zero row."""
    from b200 import native
    d = torch.device('cuda', 0)
    torch.manual_seed(M * 3 + P)
    logp = torch.log_softmax(torch.randn(M, P) * 3, 1).to(d)
    cols = torch.randint(0, P, (M,))
    cols[M // 2] = -100
    vals = -torch.rand(M) / M
    dense = torch.zeros(M, P)
    ok = cols >= 0
    dense[torch.arange(M)[ok], cols[ok]] = vals[ok]
    ld = native.pad8(P)
    dz_d = torch.full((M, ld), 7.0, device=d, dtype=torch.bfloat16)
    dz_h = torch.full((M, ld), 7.0, device=d, dtype=torch.bfloat16)
    db_d = torch.zeros(P, device=d)
    db_h = torch.zeros(P, device=d)
    native.logsoftmax_bwd_fused(dense.to(d), logp, dz_d, db_d)
    native.logsoftmax_bwd_fused_onehot(torch.arange(M, device=d), cols.to(d), vals.to(d), logp, dz_h, db_h)
    torch.cuda.synchronize()
    assert torch.equal(dz_h, dz_d)
    assert (db_h - db_d).abs().max().item() <= 1e-6 * max(db_d.abs().max().item(), 1e-30)
    assert dz_h[M // 2].float().abs().max().item() == 0


@pytest.mark.parametrize('M,P,pad', [(517, 1944, 0), (300, 48, 0), (33, 10, 6)])
def test_nll_err_rows_matches_torch(M, P, pad):
    from b200 import native
    d = torch.device('cuda', 0)
    torch.manual_seed(M + P)
    full = torch.log_softmax(torch.randn(M, P + pad) * 3, 1).to(d)
    logp = full[:, :P]
    logp[5, 3] = logp[5, 7] = 1.0            # a tie: torch.max takes the first index
    lab = torch.randint(0, P, (M,), device=d)
    lab[5] = 3
    lr, er = native.nll_err_rows(logp, lab)
    want_l = -logp[torch.arange(M, device=d), lab]
    want_e = (torch.max(logp, dim=1)[1] != lab).float()
    assert torch.equal(lr, want_l) and torch.equal(er, want_e) and er[5].item() == 0.0


def test_nll_sparse_gradient_through_the_fused_output_layer():
    """graph._NLLSparseGrad + MLPLayerTCFunction: loss and every gradient equal to the dense nn.NLLLoss path."""
    import neural_networks as nb
    from helpers import MLP_OPTS
    from b200 import graph as bgraph
    d = torch.device('cuda', 0)
    torch.manual_seed(4)
    np.random.seed(4)
    opts = dict(MLP_OPTS, dnn_lay='1944', hcgs_block='64,4', hcgs_sparse='50,25', b200_precision='bf16')
    net = nb.MLP(opts, 552).to(d).train()
    x = torch.randn(300, 552, device=d)
    lab = torch.randint(0, 1944, (300,), device=d)
    res = []
    for sparse in (True, False):
        xi = x.clone().requires_grad_(True)
        for p in net.parameters():
            p.grad = None
        o = net(xi)
        assert type(o.grad_fn).__name__ == 'MLPLayerTCFunctionBackward'
        loss = bgraph._nll(torch.nn.NLLLoss(), o, lab) if sparse else torch.nn.NLLLoss()(o, lab)
        (loss * 1.5).backward()
        res.append((loss.item(), xi.grad.clone(), {k: p.grad.clone() for k, p in net.named_parameters() if p.grad is not None}))
    (l0, gx0, g0), (l1, gx1, g1) = res
    assert abs(l0 - l1) <= 1e-6 * abs(l1)
    close(gx0, gx1, rtol=1e-6, what='dx')
    assert set(g0) == set(g1)
    for k in g1:
        close(g0[k], g1[k], rtol=1e-6, what=k)


@pytest.mark.parametrize('shape', [(7, 3, 12), (300, 550), (65, 1)])
def test_reference_layernorm_fwd_bwd(shape):
    """hcgs_layernorm_fwd / _bwd = the reference's LayerNorm class (neural_networks.py:40-51: unbiased std, divide by
    std + eps) and its autograd, against torch on the CPU in float64."""
    from b200.functions import LayerNormFunction
    if shape[-1] == 1:
        pytest.skip('n = 1: the unbiased std is nan in the reference too')
    torch.manual_seed(sum(shape))
    d = torch.device('cuda', 0)
    F = shape[-1]
    x = torch.randn(*shape) * 2 + 0.3
    gamma, beta = torch.rand(F) + 0.5, torch.randn(F)
    w = torch.randn(*shape)
    xr, gr, br = [t.double().requires_grad_(True) for t in (x, gamma, beta)]
    yr = gr * (xr - xr.mean(-1, keepdim=True)) / (xr.std(-1, keepdim=True) + 1e-6) + br
    (yr * w.double()).sum().backward()
    xg, gg, bg = [t.to(d).requires_grad_(True) for t in (x, gamma, beta)]
    y = LayerNormFunction.apply(xg, gg, bg, 1e-6)
    (y * w.to(d)).sum().backward()
    torch.cuda.synchronize()

    def close_(a, b, what):
        err = (a.detach().cpu().double() - b.detach()).abs().max().item()
        assert err <= 1e-5 * max(b.detach().abs().max().item(), 1e-30), '%s: %.3e' % (what, err)
    close_(y, yr, 'y')
    close_(xg.grad, xr.grad, 'dx')
    close_(gg.grad, gr.grad, 'dgamma')
    close_(bg.grad, br.grad, 'dbeta')
