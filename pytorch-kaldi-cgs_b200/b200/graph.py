"""Host-side mirror of the reference's graph interpreter for the hot path
(utils.model_init :1749-1830, utils.forward_model :1884-2050, utils.optimizer_init :1833-1881).

In a pytorch-kaldi checkout the reference's own ``utils.py``/``core.py`` drive the plugin classes
unchanged (cfg ``arch_library = neural_networks``).  This module is the same interpreter for use
*outside* such a checkout (tests, bench.py, export tooling): it understands the ``[model]`` lines
the CGS cfgs use -- compute, cost_nll, cost_err, sum, mult_constant, sum_constant, mult, avg,
concatenate, mse -- with the same tensor reshaping rules between sequential and frame-level nets.
``cost_l1/l2/gl`` need the regularisation helpers of utils.py and are not on the hot path.
"""
import re
import warnings

import torch
import torch.nn as nn
import torch.optim as optim

_LINE = re.compile(r'(.*)=(.*)\((.*),(.*)\)')


def parse_model(model):
    lines = [l.strip() for l in (model.split('\n') if isinstance(model, str) else model) if l.strip()]
    return [tuple(s.strip() for s in _LINE.findall(l)[0]) for l in lines]


class _NLLSparseGrad(torch.autograd.Function):
    """nn.NLLLoss() (mean over the rows of -logp[i, label_i]; utils.py:1931-1940) for the fused output layer of the bf16
    build.  Forward: ONE pass over the log-posteriors gives the per-row NLL terms and the per-row error flags of cost_err
    (utils.py:1941-1952) -- `err_rows` is handed back through ``stash`` for the cost_err line of the same output.
    Backward: the gradient is handed on as what it is -- one entry per row -- instead of being scattered into a
    zero-filled dense [rows, classes] matrix; functions.MLPLayerTCFunction.backward consumes that form directly
    (hcgs_logsoftmax_bwd_fused_onehot), and autograd densifies it by itself if it has to be added to another gradient."""

    @staticmethod
    def forward(ctx, logp, labels, stash):
        from . import native
        ctx.save_for_backward(labels)
        ctx.shape = tuple(logp.shape)
        loss_rows, err_rows = native.nll_err_rows(logp, labels)
        stash['err_rows'] = err_rows
        return loss_rows.mean()

    @staticmethod
    def backward(ctx, g):
        labels, = ctx.saved_tensors
        m, p = ctx.shape
        idx = torch.stack([torch.arange(m, device=labels.device), labels])
        vals = (g.reshape(1) * (-1.0 / m)).expand(m).contiguous()
        with warnings.catch_warnings():
            warnings.simplefilter('ignore', UserWarning)   # "invariant checks are implicitly disabled": that is the intent
            return torch.sparse_coo_tensor(idx, vals, (m, p), check_invariants=False, is_coalesced=True), None, None


class _SharedInput(torch.autograd.Function):
    """One tensor read by several nets (the LSTM output feeding the pdf and the monophone head): hands out one alias per
    reader and adds their gradients in ONE pass straight into rows padded to 16 bytes -- the layout the recurrent layers
    read their incoming gradient in (functions._padded_rows) -- instead of autograd's out-of-place sum followed by a
    repacking copy."""

    @staticmethod
    def forward(ctx, x, n):
        return tuple(x.view_as(x) for _ in range(n))

    @staticmethod
    def backward(ctx, *gs):
        gs = [g for g in gs if g is not None]
        if not gs:
            return None, None
        g0 = gs[0]
        if len(gs) == 1 or g0.dim() != 2 or not g0.is_cuda or any(g.is_sparse or g.shape != g0.shape for g in gs):
            tot = gs[0]
            for g in gs[1:]:
                tot = tot + g
            return tot, None
        m, d = g0.shape
        out = torch.empty(m, (d + 7) // 8 * 8, device=g0.device, dtype=g0.dtype)[:, :d]
        torch.add(gs[0], gs[1], out=out)
        for g in gs[2:]:
            out.add_(g)
        return out, None


def _fused_head(o):
    fn = o.grad_fn
    return (fn is not None and type(fn).__name__ == 'MLPLayerTCFunctionBackward' and o.is_cuda and o.dim() == 2 and
            o.dtype == torch.float32 and o.stride(1) == 1 and torch.is_grad_enabled())


def _nll(nll_module, o, labels, stash=None):
    if _fused_head(o):
        return _NLLSparseGrad.apply(o, labels.contiguous(), stash if stash is not None else {})
    return nll_module(o, labels)


class Graph:
    """nets: name -> module; seq: name -> bool (arch_seq_model); fea: name -> (start, end) column
    range in the minibatch tensor; lab: name -> label column (core.py packs labels as float columns)."""

    def __init__(self, model, nets, seq, fea, lab):
        self.ops = parse_model(model)
        self.nets, self.seq, self.fea, self.lab = nets, seq, fea, lab
        self.nll = nn.NLLLoss()
        # how many nets read each tensor (a tensor with several readers is handed out through _SharedInput)
        self.readers = {}
        for _, op, _, b in self.ops:
            if op == 'compute':
                self.readers[b] = self.readers.get(b, 0) + 1

    def _labels(self, inp, name):
        col = self.lab[name]
        lab = inp[:, :, col] if inp.dim() == 3 else inp[:, col]
        return lab.reshape(-1).long()

    def forward(self, inp, max_len, batch_size, to_do='train', forward_outs=()):
        outs = {}
        for f, (a, b) in self.fea.items():
            outs[f] = inp[:, :, a:b] if inp.dim() == 3 else inp[:, a:b]
        last = forward_outs[-1] if len(forward_outs) else None
        fused_err = {}
        shared = {}
        for out_name, op, a, b in self.ops:
            stop = to_do == 'forward' and out_name == last
            if op == 'compute':
                x = outs[b]
                if b in self.fea:
                    if inp.dim() == 3 and not self.seq[a]:
                        x = x.reshape(max_len * batch_size, -1)
                    if inp.dim() == 2 and self.seq[a]:
                        x = x.view(max_len, batch_size, -1)
                else:
                    if not self.seq[a] and x.dim() == 3:
                        x = outs[b] = x.view(max_len * batch_size, -1)
                    if self.seq[a] and x.dim() == 2:
                        x = outs[b] = x.view(max_len, batch_size, -1)
                if self.readers.get(b, 0) > 1 and b not in self.fea and x.is_cuda and x.dim() == 2 and x.requires_grad:
                    if b not in shared:
                        shared[b] = list(_SharedInput.apply(x, self.readers[b]))
                    x = shared[b].pop()
                outs[out_name] = self.nets[a](x)
            elif op == 'cost_nll':
                if to_do != 'forward':
                    o = outs[a]
                    if o.dim() == 3:
                        o = o.view(max_len * batch_size, -1)
                    lab = self._labels(inp, b)
                    st = {}
                    outs[out_name] = _nll(self.nll, o, lab, st)
                    if 'err_rows' in st:       # cost_err of the same output / labels comes out of the same pass
                        fused_err[(a, b)] = st['err_rows']
            elif op == 'cost_err':
                if to_do != 'forward':
                    o = outs[a]
                    if o.dim() == 3:
                        o = o.view(max_len * batch_size, -1)
                    if (a, b) in fused_err:
                        outs[out_name] = fused_err[(a, b)].mean()
                    else:
                        pred = torch.max(o, dim=1)[1]
                        outs[out_name] = torch.mean((pred != self._labels(inp, b)).float())
            elif op == 'concatenate':
                outs[out_name] = torch.cat((outs[a], outs[b]), outs[a].dim() - 1)
            elif op == 'mult':
                outs[out_name] = outs[a] * outs[b]
            elif op == 'sum':
                outs[out_name] = outs[a] + outs[b]
            elif op == 'mult_constant':
                outs[out_name] = outs[a] * float(b)
            elif op == 'sum_constant':
                outs[out_name] = outs[a] + float(b)
            elif op == 'avg':
                outs[out_name] = (outs[a] + outs[b]) / 2
            elif op == 'mse':
                outs[out_name] = torch.mean((outs[a] - outs[b]) ** 2)
            else:
                raise NotImplementedError('model operation %r is not on the HCGS hot path' % op)
            if stop and op != 'cost_nll' and op != 'cost_err':
                break
        return outs


def make_optimizer(net, opts, fused=None):
    """utils.optimizer_init for one net; ``opts`` is the cfg section (string values).  For CUDA nets the update runs
    as one fused launch per net (b200/optim.py, same arithmetic and state names as torch.optim); ``fused=False`` or the
    cfg key ``b200_fused_opt = False`` builds the torch.optim instance the reference would build."""
    from neural_networks import strtobool
    lr = float(opts['arch_lr'])
    kind = opts['arch_opt']
    params = list(net.parameters())
    if fused is None:
        fused = bool(strtobool(opts.get('b200_fused_opt', 'True')))
    fused = fused and bool(params) and all(p.is_cuda and p.dtype == torch.float32 for p in params)
    if fused and kind == 'rmsprop' and bool(strtobool(opts['opt_centered'])):
        fused = False
    if fused and kind == 'adam' and bool(strtobool(opts['opt_amsgrad'])):
        fused = False
    if fused and kind in ('sgd', 'adam', 'rmsprop'):
        from .optim import FusedOptimizer
        if kind == 'sgd':
            return FusedOptimizer(params, 'sgd', lr, momentum=float(opts['opt_momentum']),
                                  weight_decay=float(opts['opt_weight_decay']), dampening=float(opts['opt_dampening']),
                                  nesterov=bool(strtobool(opts['opt_nesterov'])))
        if kind == 'adam':
            return FusedOptimizer(params, 'adam', lr, betas=list(map(float, opts['opt_betas'].split(','))),
                                  eps=float(opts['opt_eps']), weight_decay=float(opts['opt_weight_decay']))
        return FusedOptimizer(params, 'rmsprop', lr, momentum=float(opts['opt_momentum']), alpha=float(opts['opt_alpha']),
                              eps=float(opts['opt_eps']), weight_decay=float(opts['opt_weight_decay']))
    if kind == 'sgd':
        return optim.SGD(net.parameters(), lr=lr, momentum=float(opts['opt_momentum']),
                         weight_decay=float(opts['opt_weight_decay']), dampening=float(opts['opt_dampening']),
                         nesterov=bool(strtobool(opts['opt_nesterov'])))
    if kind == 'adam':
        betas = list(map(float, opts['opt_betas'].split(',')))
        return optim.Adam(net.parameters(), lr=lr, betas=betas, eps=float(opts['opt_eps']),
                          weight_decay=float(opts['opt_weight_decay']), amsgrad=bool(strtobool(opts['opt_amsgrad'])))
    if kind == 'rmsprop':
        return optim.RMSprop(net.parameters(), lr=lr, momentum=float(opts['opt_momentum']),
                             alpha=float(opts['opt_alpha']), eps=float(opts['opt_eps']),
                             centered=bool(strtobool(opts['opt_centered'])),
                             weight_decay=float(opts['opt_weight_decay']))
    raise ValueError('unknown arch_opt %r' % kind)
