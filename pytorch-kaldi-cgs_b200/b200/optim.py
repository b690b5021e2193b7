"""torch.optim drop-ins on the fused update kernel (csrc/optimizer.cu, hcgs_opt_step_multi).

``utils.optimizer_init`` (utils.py:1833-1881) builds ``torch.optim.SGD / RMSprop / Adam`` over all parameters of a net;
torch then walks the group with a handful of multi-tensor launches per update (21 ATen launches and 0.8 ms per train
step for the three RMSprop instances of the TIMIT_CGS nets).  ``FusedOptimizer`` has the same constructor arguments,
per-parameter state names (``momentum_buffer`` / ``square_avg`` / ``exp_avg`` / ``exp_avg_sq`` / ``step``) and arithmetic
(tests/test_gpu_kernels.py compares every kind with torch.optim), and updates a whole group in one launch.
``b200.graph.make_optimizer`` uses it for CUDA nets; inside a pytorch-kaldi-CGS checkout the reference keeps building
torch.optim itself -- both update the same ``nn.Parameter`` objects, checkpoints stay interchangeable.
"""
import torch

from . import native


class FusedOptimizer(torch.optim.Optimizer):
    def __init__(self, params, kind, lr, momentum=0.0, dampening=0.0, nesterov=False, alpha=0.99, eps=1e-8,
                 betas=(0.9, 0.999), weight_decay=0.0, centered=False, amsgrad=False):
        if kind not in native.OPT_KINDS:
            raise ValueError('unknown optimizer kind %r' % kind)
        if centered or amsgrad:
            raise NotImplementedError('centered RMSprop / amsgrad Adam are not fused; use torch.optim')
        super().__init__(params, dict(kind=kind, lr=lr, momentum=momentum, dampening=dampening, nesterov=nesterov,
                                      alpha=alpha, eps=eps, betas=tuple(betas), weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            kind = group['kind']
            ps = [p for p in group['params'] if p.grad is not None]
            if not ps:
                continue
            # torch.optim counts the steps per parameter (a parameter without a gradient skips a step): parameters with
            # the same count share a launch
            by_step = {}
            for p in ps:
                if p.grad.is_sparse or not p.is_contiguous():
                    raise RuntimeError('FusedOptimizer needs dense contiguous parameters')
                st = self.state[p]
                if not st:
                    st['step'] = 0
                    if kind == 'sgd' and group['momentum'] != 0:
                        st['momentum_buffer'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    if kind == 'rmsprop':
                        st['square_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                        if group['momentum'] > 0:
                            st['momentum_buffer'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    if kind == 'adam':
                        st['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                        st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st['step'] += 1
                by_step.setdefault(st['step'], []).append(p)
            for step, plist in by_step.items():
                gs = [p.grad if p.grad.is_contiguous() else p.grad.contiguous() for p in plist]
                if kind == 'sgd':
                    s1, s2 = [self.state[p].get('momentum_buffer') for p in plist], None
                    hp = [group['lr'], group['momentum'], group['dampening'], 0.0, group['weight_decay'],
                          1.0 if group['nesterov'] else 0.0, 0.0, float(step)]
                elif kind == 'rmsprop':
                    s1 = [self.state[p]['square_avg'] for p in plist]
                    s2 = [self.state[p].get('momentum_buffer') for p in plist]
                    hp = [group['lr'], group['alpha'], group['momentum'], group['eps'], group['weight_decay'], 0.0, 0.0,
                          float(step)]
                else:
                    s1 = [self.state[p]['exp_avg'] for p in plist]
                    s2 = [self.state[p]['exp_avg_sq'] for p in plist]
                    hp = [group['lr'], group['betas'][0], group['betas'][1], group['eps'], group['weight_decay'], 0.0,
                          0.0, float(step)]
                s1 = s1 if s1 and s1[0] is not None else None
                s2 = s2 if s2 and s2[0] is not None else None
                native.opt_step_multi(kind, [p.data for p in plist], gs, s1, s2, hp)
        return loss
