"""``HCGS`` / ``guidedHCGS`` mask modules (reference: HCGS.py:12-28, 43-59).

The mask is an ``nn.Parameter`` so that it travels in ``state_dict()`` under the
reference's keys (``hcgsx.N.mask`` ...).  It never receives a gradient from the
hot path (only ``.data`` is used), matching the reference.
"""
from torch.nn.modules.module import Module
from torch.nn.parameter import Parameter

import guided_hcgs
import hcgs


class HCGS(Module):
    def __init__(self, in_features, out_features, block_sizes, drop_ratios, des='xyz'):
        super(HCGS, self).__init__()
        self.in_features = in_features
        self.out_features = out_features
        # argument order as in the reference: rows = out_features (HCGS.py:28)
        self.mask = Parameter(hcgs.conn_mat(out_features, in_features, block_sizes[:], drop_ratios[:], des))


class guidedHCGS(Module):
    def __init__(self, in_features, out_features, block_sizes, drop_ratios, w_mat, des='xyz'):
        super(guidedHCGS, self).__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.mask = Parameter(
            guided_hcgs.conn_mat(out_features, in_features, block_sizes[:], drop_ratios[:], w_mat, des))
