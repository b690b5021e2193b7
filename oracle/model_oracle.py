"""ORACLE (test infrastructure only) -- torch-CPU fp32 restatement of the reference's floating-point
hot path: LSTM.forward (neural_networks.py:823-1112), MLP.forward (:245-319), the cost lines of
utils.forward_model (utils.py:1935-2011) and the posterior normalisation (core.py:241-245).

The arithmetic of the reference lives in PyTorch itself, so the oracle is "the reference's operator
sequence on the installed torch CPU kernels": the same Python loop over T with four F.linear calls
per step on masked-dense weights.  Gradients come from torch autograd, exactly as in the
reference.  It is pinned against tensors dumped from the real reference (tests/golden/*.npz,
produced by tests/golden/make_golden.py) in tests/test_oracle_pinned.py.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) import this.

Parameters are passed as flat dicts keyed like the reference's state_dict
(``wfx.0.weight``, ``hcgsx.0.mask``, ``bn_wfx.0.running_mean`` ...), so a state_dict of either
implementation can be fed in directly.
"""
import torch
import torch.nn.functional as F

GATES = ('f', 'i', 'o', 'c')


def act(name, x):
    """act_fun (neural_networks.py:54-78)."""
    if name == 'relu':
        return torch.relu(x)
    if name == 'tanh':
        return torch.tanh(x)
    if name == 'htanh':
        return F.hardtanh(x)
    if name == 'sigmoid':
        return torch.sigmoid(x)
    if name == 'leaky_relu':
        return F.leaky_relu(x, 0.2)
    if name == 'elu':
        return F.elu(x)
    if name == 'linear':
        return x
    raise ValueError(name)


def _ste(x, xq):
    """value of xq, gradient of x (what `.data = quantised` does to the autograd graph)."""
    return x + (xq - x).detach()


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


class _RoundGrad(torch.autograd.Function):
    """identity whose incoming gradient is rounded to bf16"""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return _bf(g)


class _SplitLinear(torch.autograd.Function):
    """forward: x_hi w_hi^T + x_lo w_hi^T + x_hi w_lo^T (hi = bf16(v), lo = bf16(v - hi)), fp32 accumulate;
    backward: the single-term bf16 products of the tensor-core build (dx = bf16(g) w_hi, dw = bf16(g)^T x_hi)."""

    @staticmethod
    def forward(ctx, x, w):
        xh, wh = _bf(x), _bf(w)
        xl, wl = _bf(x - xh), _bf(w - wh)
        ctx.save_for_backward(xh, wh)
        return F.linear(xh, wh) + F.linear(xl, wh) + F.linear(xh, wl)

    @staticmethod
    def backward(ctx, g):
        xh, wh = ctx.saved_tensors
        gb = _bf(g)
        dx = gb @ wh
        dw = gb.reshape(-1, gb.shape[-1]).t() @ xh.reshape(-1, xh.shape[-1])
        return dx, dw


def linear(x, w, b=None, bf16=False):
    """F.linear; with bf16=True the three matrix products of the layer (forward, dX, dW) take bf16-rounded
    operands (x, w, and the incoming gradient) and accumulate in fp32 -- the arithmetic of the tensor-core
    build (spec key 'bf16_operands').  bf16='split': the forward product takes two-term bf16 operands (the
    projections and hidden MLP layers of the tensor-core build, hcgs_tc_gemm_nt_split), backward as bf16=True.
    Bias and its gradient stay fp32."""
    if not bf16:
        return F.linear(x, w, b)
    if bf16 == 'split':
        y = _SplitLinear.apply(x, w)
    else:
        y = _RoundGrad.apply(F.linear(_ste(x, _bf(x.detach())), _ste(w, _bf(w.detach()))))
    return y if b is None else y + b


def quantize_w_t(w, bits):
    """Quantize(balanced=False) (quantized_modules.py:77-97) on a tensor; returns STE tensor."""
    wc = w.detach().clamp(-1, 1)
    s = 2 ** (bits - 1)
    q = wc.abs().mul(s).ceil().div(s) * wc.sign()
    return _ste(w, q)


def prune_t(w, perc):
    """quantized_modules.prune (:15-28) on one weight matrix: threshold = np.percentile(|w|, perc) (float64, linear
    interpolation), mask = |w| > threshold; the reference multiplies the mask into weight.data (the gradient
    stays dense)."""
    import numpy as np
    thr = np.percentile(w.detach().abs().numpy().flatten(), perc)
    mask = (w.detach().abs() > thr).float()
    return _ste(w, w.detach() * mask) if w.requires_grad else w * mask, mask


def quantize_inp_t(x, bits, repeat=1):
    """Quantize_inp (quantized_modules.py:99-119) applied `repeat` times; STE."""
    v = x.detach()
    for _ in range(repeat):
        mx, mn = v.max().abs(), v.min().abs()
        var = mx if mx > mn else mn
        if var == 0.:
            break
        sgn = v.sign()
        t = v.div(var)
        t = t.abs().mul(2 ** (bits - 1)).ceil().div(2 ** (bits - 1))
        v = t.mul(var) * sgn
    return _ste(x, v)


def lstm_forward(p, spec, x, drop_masks, training):
    """LSTM.forward.  p: dict of tensors; spec keys: lay, act, use_bn, drop, hcgs, quant (bits or 0),
    quant_inp (bits or 0).  drop_masks: list of [B,H] tensors (train) -- the reference draws them
    with torch.bernoulli on the CPU generator (:843-845); in eval the scalar (1-p) is used (:847).
    BN running statistics in `p` are updated in place when training (momentum 0.05)."""
    bfo = bool(spec.get('bf16_operands', False))
    bfp = ('split' if spec.get('proj_terms', 2) == 2 else True) if bfo else False   # projections: two-term operands
    bidir = bool(spec.get('bidir', False))      # the reference's dormant branch (:836-838, 1104-1107), override :835 lifted
    for i, H in enumerate(spec['lay']):
        if bidir:
            x = torch.cat([x, torch.flip(x, [0])], 1)
        T, B, _ = x.shape
        drop = drop_masks[i] if training else torch.tensor([1 - spec['drop'][i]])
        w, u = {}, {}
        for g in GATES:
            w[g] = p['w%sx.%d.weight' % (g, i)]
            u[g] = p['u%sh.%d.weight' % (g, i)]
            if spec.get('hcgs', False):
                w[g] = w[g] * p['hcgsx.%d.mask' % i]                      # :858-861
                u[g] = u[g] * p['hcgsh.%d.mask' % i]                      # :980-983
            if spec.get('prune', 0):                                      # :887-896, 997-1005
                w[g] = w[g] * prune_t(w[g], spec['prune'])[1]
                u[g] = u[g] * prune_t(u[g], spec['prune'])[1]
            if spec.get('quant', 0):
                w[g] = quantize_w_t(w[g], spec['quant'])
                u[g] = quantize_w_t(u[g], spec['quant'])
        xin = x
        if spec.get('quant_inp', 0):
            xin = quantize_inp_t(x, spec['quant_inp'], 4)                 # four gate linears re-quantise x
        pre = {}
        for g in GATES:
            out = linear(xin, w[g], None, bfp)                            # :951-954
            if spec['use_bn'][i]:
                k = 'bn_w%sx.%d.' % (g, i)
                out = F.batch_norm(out.view(T * B, H), p[k + 'running_mean'], p[k + 'running_var'], p[k + 'weight'],
                                   p[k + 'bias'], training, 0.05, 1e-5).view(T, B, H)   # :957-976
            pre[g] = out
        ht = torch.zeros(B, H)
        ct = ht
        hs = []
        a = spec['act'][i]
        for k in range(T):                                                # :1077-1097
            hq = ht
            if spec.get('quant_inp', 0) and k > 0:
                hq = quantize_inp_t(ht, spec['quant_inp'], 4)
                hs[-1] = hq                                               # the stacked copy is the swapped tensor
            ft = torch.sigmoid(pre['f'][k] + linear(hq, u['f'], None, bfo))
            it = torch.sigmoid(pre['i'][k] + linear(hq, u['i'], None, bfo))
            ot = torch.sigmoid(pre['o'][k] + linear(hq, u['o'], None, bfo))
            ct = it * act(a, pre['c'][k] + linear(hq, u['c'], None, bfo)) * drop + ft * ct
            ht = ot * act(a, ct)
            if spec.get('use_ln', None) and spec['use_ln'][i]:            # :1092-1095 with the reference's class :40-51
                mean = ht.mean(-1, keepdim=True)
                std = ht.std(-1, keepdim=True)                            # unbiased
                ht = p['ln.%d.gamma' % i] * (ht - mean) / (std + 1e-6) + p['ln.%d.beta' % i]
            hs.append(ht)
        x = torch.stack(hs)                                               # :1100
        if bidir:
            x = torch.cat([x[:, :B // 2], torch.flip(x[:, B // 2:], [0])], 2)
    return x


CELL_GATES = {'gru': ('h', 'z', 'r'), 'ligru': ('h', 'z'), 'minimalgru': ('h', 'z'), 'rnn': ('h',)}


def gated_rnn_forward(p, spec, x, drop_masks, training):
    """GRU.forward (neural_networks.py:1343-1426), liGRU (:1523-1599), minimalGRU (:1701-1777), RNN (:1863-1931).
    spec keys: cell ('gru' | 'ligru' | 'minimalgru' | 'rnn'), lay, act, use_bn, drop, bidir, hcgs (masks hcgsx.N.mask /
    hcgsh.N.mask shared by the gates -- not in the reference, which has these cells dense)."""
    cell = spec['cell']
    gates = CELL_GATES[cell]
    bidir = bool(spec.get('bidir', False))
    for i, H in enumerate(spec['lay']):
        if bidir:
            x = torch.cat([x, torch.flip(x, [0])], 1)
        T, B, _ = x.shape
        drop = drop_masks[i] if training else torch.tensor([1 - spec['drop'][i]])
        w, u, b, pre = {}, {}, {}, {}
        for g in gates:
            w[g], u[g] = p['w%s.%d.weight' % (g, i)], p['u%s.%d.weight' % (g, i)]
            b[g] = p.get('w%s.%d.bias' % (g, i))
            if spec.get('hcgs', False):
                w[g] = w[g] * p['hcgsx.%d.mask' % i]
                u[g] = u[g] * p['hcgsh.%d.mask' % i]
            out = F.linear(x, w[g], b[g])
            if spec['use_bn'][i]:
                k = 'bn_w%s.%d.' % (g, i)
                out = F.batch_norm(out.view(T * B, H), p[k + 'running_mean'], p[k + 'running_var'], p[k + 'weight'],
                                   p[k + 'bias'], training, 0.05, 1e-5).view(T, B, H)
            pre[g] = out
        ht = torch.zeros(B, H)
        hs = []
        a = spec['act'][i]
        for k in range(T):
            if cell == 'rnn':
                ht = act(a, pre['h'][k] + F.linear(ht, u['h'])) * drop
            else:
                zt = torch.sigmoid(pre['z'][k] + F.linear(ht, u['z']))
                if cell == 'gru':
                    rt = torch.sigmoid(pre['r'][k] + F.linear(ht, u['r']))
                    at = pre['h'][k] + F.linear(rt * ht, u['h'])
                elif cell == 'minimalgru':
                    at = pre['h'][k] + F.linear(zt * ht, u['h'])
                else:
                    at = pre['h'][k] + F.linear(ht, u['h'])
                hcand = act(a, at) * drop
                ht = zt * ht + (1 - zt) * hcand
            hs.append(ht)
        x = torch.stack(hs)
        if bidir:
            x = torch.cat([x[:, :B // 2], torch.flip(x[:, B // 2:], [0])], 2)
    return x


def mlp_forward(p, spec, x, training, drop_masks=None):
    """MLP.forward (:245-319).  spec keys: lay, act, use_bn, drop, hcgs, quant, quant_inp."""
    for i, O in enumerate(spec['lay']):
        w = p['wx.%d.weight' % i]
        b = p['wx.%d.bias' % i]
        if spec.get('hcgs', False):
            w = w * p['hcgs.%d.mask' % i]                                 # :258
        if spec.get('prune', 0):                                          # :275-277
            w = w * prune_t(w, spec['prune'])[1]
        if spec.get('quant', 0):
            w = quantize_w_t(w, spec['quant'])
        if spec.get('quant_inp', 0):
            x = quantize_inp_t(x, spec['quant_inp'], 1)
        bfo = bool(spec.get('bf16_operands', False))
        if bfo and spec['act'][i] != 'softmax' and spec.get('proj_terms', 2) == 2:
            bfo = 'split'                                                 # hidden layers: two-term operands
        z = linear(x, w, b, bfo)
        if spec['use_bn'][i]:
            k = 'bn.%d.' % i
            z = F.batch_norm(z, p[k + 'running_mean'], p[k + 'running_var'], p[k + 'weight'], p[k + 'bias'],
                             training, 0.05, 1e-5)
        if spec['act'][i] == 'softmax':
            z = F.log_softmax(z, dim=1)                                   # :73-74
        else:
            z = act(spec['act'][i], z)
        if training and spec['drop'][i] > 0:
            z = z * drop_masks[i]                                         # nn.Dropout: mask holds 1/(1-p)
        x = z
    return x


def costs(logp, labels):
    """cost_nll + cost_err (utils.py:1935-1952, 1993-2011): mean over ALL rows incl. padding."""
    lab = labels.reshape(-1).long()
    loss = F.nll_loss(logp, lab)
    err = torch.mean((torch.max(logp, dim=1)[1] != lab).float())
    return loss, err


def model_forward(params, specs, inp, n_fea, training, drop_masks, mono_weight=1.0):
    """The shipped two-head graph (cfg/TIMIT_CGS/*.cfg [model]):
        out1 = LSTM(fea); out2 = MLP_cd(out1); out3 = MLP_mono(out1)
        loss_final = nll(out2, lab_cd) + w * nll(out3, lab_mono); err_final = err(out2, lab_cd)
    inp: [T, B, n_fea + n_labels] with the labels as trailing float columns (core.py:183-200).
    params/specs: dicts keyed 'lstm', 'cd', optionally 'mono'."""
    T, B, _ = inp.shape
    recurrent = gated_rnn_forward if specs['lstm'].get('cell') else lstm_forward
    h = recurrent(params['lstm'], specs['lstm'], inp[:, :, :n_fea], drop_masks, training)
    flat = h.view(T * B, -1)
    out = {'out_dnn1': h}
    out['out_dnn2'] = mlp_forward(params['cd'], specs['cd'], flat, training)
    if specs['cd'].get('quant_inp', 0):
        # the cd head swapped out_dnn1.data for its quantised copy; the mono head starts from that
        flat = quantize_inp_t(flat, specs['cd']['quant_inp'], 1)
    loss, err = costs(out['out_dnn2'], inp[:, :, n_fea])
    out['loss_cd'], out['err_final'] = loss, err
    out['loss_final'] = loss
    if 'mono' in params:
        out['out_dnn3'] = mlp_forward(params['mono'], specs['mono'], flat, training)
        lm, _ = costs(out['out_dnn3'], inp[:, :, n_fea + 1])
        out['loss_mono'] = lm
        out['loss_final'] = loss + lm * mono_weight
    return out


def posterior(logp, counts):
    """core.py:241-245 in float32 NumPy."""
    import numpy as np
    counts = np.asarray(counts, dtype=np.float32)
    return logp - np.log(counts / np.sum(counts))
