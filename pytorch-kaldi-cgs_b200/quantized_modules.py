"""Weight / activation quantisation and magnitude pruning (reference: quantized_modules.py).

Only the symbols the hot path uses are provided: ``Quantize``, ``Quantize_inp``,
``QuantizeLinear``, ``prune``.  ``BinarizeLinear``/``QuantizeVar``/``HingeLoss``
are unused by every shipped cfg and are out of scope.

Semantics that define parity (and look like bugs):
  * ``Quantize`` clamps the *master* tensor to [-1, 1] in place, then returns
    ``sign(w) * ceil(|w| * 2^(b-1)) / 2^(b-1)`` (magnitudes round up).
  * ``Quantize_inp`` scales by the max |x| of the *whole* tensor, is skipped when that is 0,
    and is applied by ``QuantizeLinear`` through ``input.data = ...`` (straight-through).
  * ``prune`` thresholds at ``np.percentile(|w|, perc)`` (float64 linear interpolation).
All arithmetic on CUDA tensors goes through the native library.
"""
import math

import numpy as np
import torch
import torch.nn as nn
from torch.nn.parameter import Parameter

from b200 import native


def _scratch(dev):
    return torch.zeros(2, device=dev, dtype=torch.float32)


def Quantize(tensor, numBits=3, if_forward=False, balanced=True):
    """quantized_modules.py:77-97.  ``balanced=True`` differs from ``False`` only by an extra clamp
    of the (already <= 1) magnitudes when ``if_forward`` is set, which is the only way the
    reference calls it (neural_networks.py:933-948); the ``balanced and not if_forward`` scaling
    branch is never reached by the hot path and is rejected."""
    if balanced and not if_forward:
        raise NotImplementedError('Quantize(balanced=True, if_forward=False) is unused by the reference hot path')
    native._require_cuda(tensor)
    if not tensor.is_contiguous():
        raise native.NativeError('Quantize needs a contiguous tensor')
    if if_forward:
        native.quantize_w(tensor, tensor, numBits)      # in place on the caller's tensor
        return tensor
    out = torch.empty_like(tensor)
    native.quantize_w(tensor, out, numBits)             # master clamped in place, q returned
    return out


def Quantize_inp(inp, Bits, if_forward=False, repeat=1):
    """quantized_modules.py:99-119.  ``repeat`` applies the map several times with the same scale
    (max|x| is a fixed point of the map), which is what four gate linears sharing one input do."""
    native._require_cuda(inp)
    x2d = inp.reshape(-1, inp.shape[-1]) if inp.dim() > 1 else inp.reshape(1, -1)
    if x2d.stride(-1) != 1:
        raise native.NativeError('Quantize_inp needs a view whose last dimension is contiguous')
    if if_forward:
        native.quantize_act(x2d, x2d, Bits, repeat, _scratch(inp.device))
        return inp
    out = torch.empty(x2d.shape, device=inp.device, dtype=torch.float32)
    native.quantize_act(x2d, out, Bits, repeat, _scratch(inp.device))
    return out.view(inp.shape)


def percentile_threshold(values, perc):
    """np.percentile(values, perc) of a 1-D float32 tensor WITHOUT leaving its device, as the installed numpy (2.x)
    evaluates it for a float32 array and a Python-number percentile (method 'linear'): everything in float32 --
    q = float32(perc) / float32(100), virtual index (n-1)*q, gamma = its fractional part, then
    a + (b-a)*gamma, or b - (b-a)*(1-gamma) when gamma >= 0.5, on the two neighbouring order statistics.
    Only the index arithmetic (scalars) runs on the host; returns a 0-dim float32 tensor."""
    n = values.numel()
    q = np.float32(perc) / np.float32(100)
    vidx = np.float32(n - 1) * q
    if vidx >= n - 1:
        lo = hi = n - 1
    elif vidx < 0:
        lo = hi = 0
    else:
        lo = int(np.floor(vidx))
        hi = lo + 1
    gamma = np.float32(vidx - np.floor(vidx))
    srt = torch.sort(values.reshape(-1))[0]
    a, b = srt[lo], srt[hi]
    diff = b - a
    if gamma >= np.float32(0.5):
        return b - diff * float(np.float32(1) - gamma)
    return a + diff * float(gamma)


def prune(model, pruning_perc):
    """quantized_modules.py:15-28: magnitude mask per >=2-D parameter, one global percentile.  The reference moves
    every matrix to the host for np.percentile on every forward (8 round trips per LSTM layer); here the order
    statistics, the interpolation and the comparison stay on the device (no synchronisation), bit-identical to the
    host result (tests/test_host_cpu.py::test_percentile_threshold_matches_numpy)."""
    mats = [p for p in model.parameters() if p.dim() != 1]
    flat = torch.cat([p.detach().abs().reshape(-1) for p in mats]) if len(mats) != 1 else mats[0].detach().abs().reshape(-1)
    threshold = percentile_threshold(flat, pruning_perc)
    return [(p.data.abs() > threshold).float() for p in mats]


class QuantizeLinear(nn.Module):
    """Parameter container + standalone forward (quantized_modules.py:182-222).

    Inside LSTM/MLP the owning layer drives the native kernels directly with the same
    semantics; this forward exists for scripts that call the module on its own."""

    def __init__(self, in_features, out_features, numBits=8, bias=True, if_forward=False, if_inp_quant=False,
                 inp_quant=16):
        super(QuantizeLinear, self).__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.weight = Parameter(torch.Tensor(out_features, in_features))
        self.weight_org = torch.Tensor(out_features, in_features)
        if bias:
            self.bias = Parameter(torch.Tensor(out_features))
        else:
            self.register_parameter('bias', None)
        self.numBits = numBits
        self.if_forward = if_forward
        self.if_inp_quant = if_inp_quant
        self.inp_quant = inp_quant
        self._plan = None
        self.reset_parameters()

    def reset_parameters(self):
        stdv = 1. / math.sqrt(self.weight.size(1))
        self.weight.data.uniform_(-stdv, stdv)
        if self.bias is not None:
            self.bias.data.uniform_(-stdv, stdv)

    def forward(self, input):
        from b200.functions import LayerRuntime, MLPLayerFunction
        if self._plan is None:
            self._plan = native.Plan(torch.ones(self.out_features, self.in_features), recurrent=False)
        rt = LayerRuntime()
        rt.plan_x = self._plan
        rt.act = native.ACT_CODES['linear']
        if self.if_inp_quant:
            input.data = Quantize_inp(input.data, self.inp_quant, self.if_forward)
        x2d = input.reshape(-1, self.in_features)
        if self.if_forward:
            self.weight.data = Quantize(self.weight.data, numBits=self.numBits, if_forward=True, balanced=False)
            out = MLPLayerFunction.apply(x2d, self.weight, self.bias, None, None, rt)
        else:
            self.weight_org = self.weight.data
            self.weight.data = Quantize(self.weight.data, numBits=self.numBits, balanced=False)
            out = MLPLayerFunction.apply(x2d, self.weight, self.bias, None, None, rt)
            self.weight.data = self.weight_org
        return out.view(*input.shape[:-1], self.out_features)

    def extra_repr(self):
        return 'in_features={}, out_features={}, bias={}'.format(
            self.in_features, self.out_features, self.bias is not None)
