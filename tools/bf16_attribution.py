"""Which bf16 operand rounding dominates the posterior error of the tensor-core build?  (CPU, oracle only.)

Runs oracle/model_oracle.py (the reference-pinned fp32 operator sequence) once in fp32 and then with bf16 rounding
switched on for ONE class of matrix product at a time:
    proj   W{f,i,o,c}x . x      (time-batched projections)
    rec    U{f,i,o,c}h . h_{t-1}   (recurrent product)            rec_h / rec_u: only that operand rounded
    out    output layers
    *_2t   the operand kept as a 2-term bf16 split (hi + lo), i.e. ~16 mantissa bits
and prints max-abs / mean-abs of the log-posterior difference against fp32.  Tool, not product.

    python tools/bf16_attribution.py --T 256 --B 8 --H 550 --layers 4
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, 'pytorch-kaldi-cgs_b200'), ROOT, os.path.join(ROOT, 'tests')):
    sys.path.insert(0, p)

from oracle import model_oracle  # noqa: E402


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


def split2(x):
    hi = bf(x)
    return hi, bf(x - hi)


MODE = {'proj': None, 'rec': None, 'out': None}   # None | 'both' | 'x' | 'w' | '2t' (both 2-term) | 'x2t' (x 2-term, w bf16)


def classify(x, w):
    if x.dim() == 3:
        return 'proj'
    if w.shape[0] == w.shape[1] and x.shape[1] == w.shape[1] and x.shape[0] < 4096 and CUR['in_rec']:
        return 'rec'
    return 'out'


CUR = {'in_rec': True}


def linear(x, w, b=None, bf16=False):
    kind = classify(x, w)
    m = MODE[kind]
    if m is None:
        y = F.linear(x, w)
    elif m == 'both':
        y = F.linear(bf(x), bf(w))
    elif m == 'x':
        y = F.linear(bf(x), w)
    elif m == 'w':
        y = F.linear(x, bf(w))
    elif m == '2t':
        xh, xl = split2(x)
        wh, wl = split2(w)
        y = F.linear(xh, wh) + F.linear(xl, wh) + F.linear(xh, wl)
    elif m == 'x2t':
        xh, xl = split2(x)
        wh = bf(w)
        y = F.linear(xh, wh) + F.linear(xl, wh)
    elif m == 'w2t':
        xh = bf(x)
        wh, wl = split2(w)
        y = F.linear(xh, wh) + F.linear(xh, wl)
    else:
        raise ValueError(m)
    return y if b is None else y + b


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--T', type=int, default=256)
    ap.add_argument('--B', type=int, default=8)
    ap.add_argument('--H', type=int, default=550)
    ap.add_argument('--layers', type=int, default=4)
    ap.add_argument('--P', type=int, default=1944)
    ap.add_argument('--eval', action='store_true')
    ap.add_argument('--out', default='')
    a = ap.parse_args()
    from helpers import build_product_nets
    from test_oracle_pinned import golden_specs
    D, H, L, T, B, P = 40, a.H, a.layers, a.T, a.B, a.P
    lay = ','.join([str(H)] * L)
    lstm_over = dict(lstm_lay=lay, lstm_drop=','.join(['0.2'] * L), lstm_use_laynorm=','.join(['False'] * L),
                     lstm_use_batchnorm=','.join(['True'] * L), lstm_act=','.join(['tanh'] * L),
                     hcgsx_block='32,4', hcgsx_sparse='75,62.5', hcgsh_block='32,4', hcgsh_sparse='75,62.5')
    mlp_over = dict(dnn_lay=str(P), hcgs_block='64,4', hcgs_sparse='50,25')
    nets, _ = build_product_nets('train', D=D, seed=2234, lstm_over=lstm_over, mlp_over=mlp_over, device='cpu')
    params = {k: {n: v.detach().clone() for n, v in nets[m].state_dict().items()}
              for k, m in (('lstm', 'LSTM_layers'), ('cd', 'MLP_layers'), ('mono', 'MLP_layers2'))}
    gen = torch.Generator().manual_seed(1)
    inp = torch.zeros(T, B, D + 2)
    inp[:, :, :D] = torch.randn(T, B, D, generator=gen)
    inp[:, :, D] = torch.randint(0, P, (T, B), generator=gen).float()
    inp[:, :, D + 1] = torch.randint(0, 5, (T, B), generator=gen).float()
    drops = [torch.bernoulli(torch.full((B, H), 0.8), generator=gen) for _ in range(L)]
    specs = golden_specs(dict(lay=[H] * L, act=['tanh'] * L, use_bn=[True] * L, drop=[0.2] * L))
    specs['cd']['lay'] = [P]
    model_oracle.linear = linear
    training = not a.eval

    def run(**modes):
        for k in MODE:
            MODE[k] = modes.get(k)
        p2 = {k: {n: v.clone() for n, v in d.items()} for k, d in params.items()}
        with torch.no_grad():
            return model_oracle.model_forward(p2, specs, inp, D, training, drops)['out_dnn2']

    ref = run()
    rows = {}
    cases = [('all bf16', dict(proj='both', rec='both', out='both')),
             ('proj only', dict(proj='both')), ('proj x only', dict(proj='x')), ('proj w only', dict(proj='w')),
             ('rec only', dict(rec='both')), ('rec h only', dict(rec='x')), ('rec U only', dict(rec='w')),
             ('out only', dict(out='both')), ('out x only', dict(out='x')), ('out w only', dict(out='w')),
             ('rec h 2-term, rest bf16', dict(proj='both', rec='x2t', out='both')),
             ('rec h,U 2-term, rest bf16', dict(proj='both', rec='2t', out='both')),
             ('all x 2-term (h, x), w bf16', dict(proj='x2t', rec='x2t', out='x2t')),
             ('all w 2-term, x bf16', dict(proj='w2t', rec='w2t', out='w2t')),
             ('all 2-term', dict(proj='2t', rec='2t', out='2t')),
             ('proj 2-term, rec + out bf16', dict(proj='2t', rec='both', out='both')),
             ('proj 2-term, out x 2-term, rec bf16', dict(proj='2t', rec='both', out='x2t')),
             ('proj + out 2-term, rec bf16', dict(proj='2t', rec='both', out='2t'))]
    for name, modes in cases:
        e = (run(**modes) - ref).abs()
        rows[name] = (e.max().item(), e.mean().item())
        print('%-34s max-abs %.3e  mean-abs %.3e' % (name, *rows[name]), flush=True)
    if a.out:
        with open(a.out, 'w') as f:
            json.dump({'T': T, 'B': B, 'H': H, 'layers': L, 'P': P, 'training': training, 'rows': rows}, f, indent=1)


if __name__ == '__main__':
    main()
