"""autograd glue between the plugin classes (neural_networks.LSTM / MLP) and the C ABI.

Each Function is one *layer*: its forward/backward are straight sequences of native calls
(no torch arithmetic on the data path; torch only allocates buffers).  Gradients are produced
in the dense ``[out, in]`` fp32 layout the reference's optimizers expect (``.grad`` of the same
``nn.Parameter`` objects), but they are only *computed* inside the non-empty tiles of the mask
and re-masked on the way out (observationally identical for every kept weight).
"""
import torch

from . import native


class LayerRuntime:
    """Per-layer mutable state shared between forward and backward (plans, flags, BN buffers)."""

    def __init__(self):
        self.plan_x = None        # native.Plan of the input-projection mask
        self.plan_h = None        # native.Plan (recurrent) of the hidden-hidden mask
        self.mask_x = None        # device mask tensors (or None = dense)
        self.mask_h = None
        self.use_bn = False
        self.training = False
        self.bn_running = None    # list of (running_mean, running_var) per group
        self.bn_eps = 1e-5
        self.bn_momentum = 0.05
        self.act = 0
        self.drop = None          # [B,H] mask tensor (train) or None
        self.drop_scalar = 1.0    # eval: (1-p) ; train without tensor: 1
        self.mlp_drop_mask = None  # [N,out] scaled keep mask for MLP dropout (or None)
        self.softmax = False
        self.log_prior = None
        # quantised compute copies (QuantizeLinear semantics: forward AND backward run on the
        # quantised weights, the gradient lands on the fp32 master -- straight-through)
        self.q_ws = None
        self.q_us = None
        self.ln_eps = 1e-6        # eps of the in-loop LayerNorm (lstm_use_laynorm)
        self.xl_in = None         # bf16 low term of the input rows when the layer below produced it (two-term projections)
        self.want_hl = False      # emit the low term of this layer's output for the layer above
        self.hl_out = None
        self.act_bits = 0         # > 0: per-step Quantize_inp of h_t inside the recurrence (lstm_quant_inp)
        self.cell = None          # GRU family: 'gru' | 'minimalgru' | 'ligru' | 'rnn'
        # tensor-core build: bf16 copy of the layer input handed over by the previous layer, and of
        # this layer's output for the next one (saves a cast pass per layer)
        self.xb_in = None
        self.hb_out = None
        # forward products of the projections / hidden MLP layers with two-term bf16 operands (hi + lo, three
        # tcgen05 products per k tile): keeps the posteriors within 1e-4 of the fp32 reference
        self.split_fwd = True


def _ws(device, n_cols):
    """BN reduction scratch: enough for 64 row-splits."""
    return torch.empty(2 * n_cols * 64, device=device, dtype=torch.float32)


class LayerNormFunction(torch.autograd.Function):
    """The reference's LayerNorm (neural_networks.py:40-51) over the last dimension of any-rank input."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        F = x.shape[-1]
        x2d = x.reshape(-1, F)
        if x2d.stride(-1) != 1:
            x2d = x2d.contiguous()
        M = x2d.shape[0]
        y = torch.empty(M, F, device=x.device, dtype=torch.float32)
        need = any(ctx.needs_input_grad)
        mean = torch.empty(M, device=x.device, dtype=torch.float32) if need else None
        std = torch.empty(M, device=x.device, dtype=torch.float32) if need else None
        native.layernorm_fwd(x2d, y, gamma, beta, eps, mean, std)
        if need:
            ctx.eps = eps
            ctx.shape = x.shape
            ctx.save_for_backward(x2d, gamma, mean, std)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2d, gamma, mean, std = ctx.saved_tensors
        M, F = x2d.shape
        dy2d = dy.reshape(M, F)
        if dy2d.stride(-1) != 1:
            dy2d = dy2d.contiguous()
        dx = torch.empty(M, F, device=dy.device, dtype=torch.float32) if ctx.needs_input_grad[0] else None
        dg = torch.zeros(F, device=dy.device, dtype=torch.float32)
        db = torch.zeros(F, device=dy.device, dtype=torch.float32)
        native.layernorm_bwd(dy2d, x2d, gamma, ctx.eps, mean, std, dx, dg, db)
        return (dx.view(ctx.shape) if dx is not None else None), dg, db, None


class BatchNormInputFunction(torch.autograd.Function):
    """nn.BatchNorm1d(momentum=0.05) on the layer input, all rows at once (*_use_batchnorm_inp,
    neural_networks.py:828-830, 251-252)."""

    @staticmethod
    def forward(ctx, x2d, gamma, beta, running_mean, running_var, eps, momentum, training):
        x2d = x2d if x2d.stride(-1) == 1 else x2d.contiguous()
        M, F = x2d.shape
        y = torch.empty(M, F, device=x2d.device, dtype=torch.float32)
        sm = torch.empty(F, device=x2d.device, dtype=torch.float32)
        si = torch.empty(F, device=x2d.device, dtype=torch.float32)
        native.bn_fwd(x2d, y, 1, F, [gamma], [beta], [running_mean], [running_var], eps, momentum, training, sm, si,
                      _ws(x2d.device, F))
        if any(ctx.needs_input_grad):
            if not training:
                raise native.NativeError('backward through eval-mode batch norm is not built')
            ctx.save_for_backward(x2d, gamma, sm, si)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2d, gamma, sm, si = ctx.saved_tensors
        M, F = x2d.shape
        dy = dy.contiguous()
        dx = torch.empty(M, F, device=dy.device, dtype=torch.float32)
        dg = torch.empty(F, device=dy.device, dtype=torch.float32)
        db = torch.empty(F, device=dy.device, dtype=torch.float32)
        native.bn_bwd(dy, x2d, 1, F, [gamma], sm, si, dx, [dg], [db], _ws(dy.device, F))
        return dx, dg, db, None, None, None, None, None


class ActDropFunction(torch.autograd.Function):
    """y = drop_mask * act(x) (or LogSoftmax(dim=1) - log_prior) as a stand-alone step: used when a LayerNorm sits
    between the linear and the activation of an MLP layer (neural_networks.py:306-307)."""

    @staticmethod
    def forward(ctx, x, act, drop_mask, softmax, log_prior):
        x = x.contiguous()
        y = torch.empty_like(x)
        if softmax:
            native.logsoftmax_fwd(x, y, log_prior)
        else:
            native.act_fwd(x, y, drop_mask, act)
        ctx.act, ctx.softmax, ctx.had_prior, ctx.drop_mask = act, softmax, log_prior is not None, drop_mask
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        y, = ctx.saved_tensors
        dy = dy.contiguous()
        dx = torch.empty_like(y)
        if ctx.softmax:
            if ctx.had_prior:
                raise native.NativeError('backward through a prior-normalised posterior is not defined')
            native.logsoftmax_bwd(dy, y, dx)
        else:
            native.act_bwd(dy, y, ctx.drop_mask, dx, ctx.act)
        return dx, None, None, None, None


class LSTMLayerFunction(torch.autograd.Function):
    """One HCGS LSTM layer: projections -> BN -> persistent recurrence (neural_networks.py:950-1100)."""

    @staticmethod
    def forward(ctx, x, wf, wi, wo, wc, uf, ui, uo, uc, gf, bf, gi, bi, go, bo, gc, bc, rt, lng, lnb):
        T, B, D = x.shape
        H = wf.shape[0]
        M = T * B
        dev = x.device
        x2d = x.reshape(M, D)
        if x2d.stride(-1) != 1:
            x2d = x2d.contiguous()
        ws = rt.q_ws if rt.q_ws is not None else [wf, wi, wo, wc]
        us = rt.q_us if rt.q_us is not None else [uf, ui, uo, uc]
        need_grad = any(ctx.needs_input_grad)
        pre_raw = torch.empty(M, 4 * H, device=dev, dtype=torch.float32)
        native.bsgemm_nt(rt.plan_x, ws, x2d, pre_raw)
        save_mean = save_invstd = None
        if rt.use_bn:
            pre = torch.empty_like(pre_raw)
            save_mean = torch.empty(4 * H, device=dev, dtype=torch.float32)
            save_invstd = torch.empty(4 * H, device=dev, dtype=torch.float32)
            rms = [r[0] for r in rt.bn_running]
            rvs = [r[1] for r in rt.bn_running]
            native.bn_fwd(pre_raw, pre, 4, H, [gf, gi, go, gc], [bf, bi, bo, bc], rms, rvs, rt.bn_eps,
                          rt.bn_momentum, rt.training, save_mean, save_invstd, _ws(dev, 4 * H))
        else:
            pre = pre_raw
        plan = rt.plan_h
        fwd_pack = torch.empty(plan.fwd_pack_floats, device=dev, dtype=torch.float32)
        bwd_pack = torch.empty(max(plan.bwd_pack_floats, 1), device=dev, dtype=torch.float32) if need_grad else None
        native.lstm_pack(plan, us, fwd_pack, bwd_pack)
        h = torch.empty(T, B, H, device=dev, dtype=torch.float32)
        c = gates = None
        if need_grad:
            c = torch.empty(T, B, H, device=dev, dtype=torch.float32)
            gates = torch.empty(T, B, 4 * H, device=dev, dtype=torch.float32)
        ln_stats = None
        if lng is not None:
            # LayerNorm of h_t inside the time loop (lstm_use_laynorm, neural_networks.py:1092-1095)
            ln_stats = torch.empty(T, B, 2, device=dev, dtype=torch.float32) if need_grad else None
            native.lstm_fwd_ln(plan, fwd_pack, pre.view(T, B, 4 * H), rt.drop, rt.drop_scalar, rt.act, h, c, gates,
                               rt.act_bits, lng.detach().contiguous(), lnb.detach().contiguous(), rt.ln_eps, ln_stats)
        else:
            native.lstm_fwd(plan, fwd_pack, pre.view(T, B, 4 * H), rt.drop, rt.drop_scalar, rt.act, h, c, gates,
                            act_bits=rt.act_bits)
        if need_grad:
            ctx.rt = rt
            ctx.dims = (T, B, D, H)
            ctx.drop = rt.drop
            ctx.drop_scalar = rt.drop_scalar
            ctx.h_store = h.detach()   # private alias: survives a later `.data =` swap on the output
            ctx.ln = lng is not None
            ctx.save_for_backward(x2d, ws[0], ws[1], ws[2], ws[3], gf, gi, go, gc, pre_raw, save_mean,
                                  save_invstd, c, gates, bwd_pack, lng, ln_stats)
        return h

    @staticmethod
    def backward(ctx, dh):
        rt = ctx.rt
        T, B, D, H = ctx.dims
        M = T * B
        (x2d, wf, wi, wo, wc, gf, gi, go, gc, pre_raw, save_mean, save_invstd, c, gates, bwd_pack, lng,
         ln_stats) = ctx.saved_tensors
        dev = dh.device
        dh = dh.contiguous()
        dpre = torch.empty(T, B, 4 * H, device=dev, dtype=torch.float32)
        dlng = dlnb = None
        if ctx.ln:
            dlng, dlnb = native.lstm_bwd_ln(rt.plan_h, bwd_pack, dh, gates, c, ctx.drop, ctx.drop_scalar, rt.act, dpre,
                                            lng.detach().contiguous(), ln_stats, rt.ln_eps)
        else:
            native.lstm_bwd(rt.plan_h, bwd_pack, dh, gates, c, ctx.drop, ctx.drop_scalar, rt.act, dpre)
        dpre2d = dpre.view(M, 4 * H)
        dgam = [None] * 4
        dbet = [None] * 4
        if rt.use_bn:
            dproj = torch.empty_like(dpre2d)
            dgam = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(4)]
            dbet = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(4)]
            native.bn_bwd(dpre2d, pre_raw, 4, H, [gf, gi, go, gc], save_mean, save_invstd, dproj, dgam, dbet,
                          _ws(dev, 4 * H))
        else:
            dproj = dpre2d
        # recurrent weight gradients: dU_g = sum_{t>=1} dpre_g[t]^T h[t-1]
        dus = [torch.zeros(H, H, device=dev, dtype=torch.float32) for _ in range(4)]
        if T > 1:
            h2d = ctx.h_store.view(M, H)
            native.bs_sddmm(rt.plan_h, dpre2d[B:], h2d[:M - B], dus, rt.mask_h)
        dws = [torch.zeros(H, D, device=dev, dtype=torch.float32) for _ in range(4)]
        native.bs_sddmm(rt.plan_x, dproj, x2d, dws, rt.mask_x)
        dx = None
        if ctx.needs_input_grad[0]:
            dx2d = torch.empty(M, D, device=dev, dtype=torch.float32)
            native.bsgemm_nn(rt.plan_x, [wf, wi, wo, wc], dproj, dx2d)
            dx = dx2d.view(T, B, D)
        return (dx, dws[0], dws[1], dws[2], dws[3], dus[0], dus[1], dus[2], dus[3],
                dgam[0], dbet[0], dgam[1], dbet[1], dgam[2], dbet[2], dgam[3], dbet[3], None, dlng, dlnb)


class GatedRNNLayerFunction(torch.autograd.Function):
    """One layer of the GRU family on the fp32 build: projections -> BN -> persistent recurrence
    (neural_networks.py GRU :1376-1410, liGRU :1553-1587, minimalGRU :1731-1765, RNN :1890-1918).
    Gates in cell order (native.CELL_GATES): ws / us / gammas / betas / biases are lists of that length."""

    @staticmethod
    def forward(ctx, x, rt, n_g, *params):
        ws, us = list(params[:n_g]), list(params[n_g:2 * n_g])
        gammas, betas = list(params[2 * n_g:3 * n_g]), list(params[3 * n_g:4 * n_g])
        biases = list(params[4 * n_g:5 * n_g])
        cell = rt.cell
        gates = native.CELL_GATES[cell]
        T, B, D = x.shape
        H = ws[0].shape[0]
        M = T * B
        dev = x.device
        x2d = x.reshape(M, D)
        if x2d.stride(-1) != 1:
            x2d = x2d.contiguous()
        need_grad = any(ctx.needs_input_grad)
        pre_raw = torch.empty(M, n_g * H, device=dev, dtype=torch.float32)
        native.bsgemm_nt(rt.plan_x, ws, x2d, pre_raw, biases if biases[0] is not None else None)
        save_mean = save_invstd = None
        if rt.use_bn:
            pre = torch.empty_like(pre_raw)
            save_mean = torch.empty(n_g * H, device=dev, dtype=torch.float32)
            save_invstd = torch.empty(n_g * H, device=dev, dtype=torch.float32)
            native.bn_fwd(pre_raw, pre, n_g, H, gammas, betas, [r[0] for r in rt.bn_running],
                          [r[1] for r in rt.bn_running], rt.bn_eps, rt.bn_momentum, rt.training, save_mean, save_invstd,
                          _ws(dev, n_g * H))
        else:
            pre = pre_raw
        plan = rt.plan_h
        fwd_pack = torch.empty(plan.fwd_pack_floats, device=dev, dtype=torch.float32)
        bwd_pack = torch.empty(max(plan.bwd_pack_floats, 1), device=dev, dtype=torch.float32) if need_grad else None
        native.gru_pack(plan, cell, dict(zip(gates, us)), fwd_pack, bwd_pack)
        h = torch.empty(T, B, H, device=dev, dtype=torch.float32)
        two_phase = cell in ('gru', 'minimalgru')
        m = torch.empty(T, B, H, device=dev, dtype=torch.float32) if (need_grad and two_phase) else None
        g_out = torch.empty(T, B, n_g * H, device=dev, dtype=torch.float32) if need_grad else None
        native.gru_fwd(plan, cell, fwd_pack, pre.view(T, B, n_g * H), rt.drop, rt.drop_scalar, rt.act, h, m, g_out)
        if need_grad:
            ctx.rt, ctx.n_g, ctx.dims = rt, n_g, (T, B, D, H)
            ctx.drop, ctx.drop_scalar = rt.drop, rt.drop_scalar
            ctx.has_bias = biases[0] is not None
            ctx.h_store = h.detach()
            ctx.save_for_backward(x2d, pre_raw if rt.use_bn else None, save_mean, save_invstd, g_out, bwd_pack, m,
                                  *ws, *[g for g in gammas if g is not None])
        return h

    @staticmethod
    def backward(ctx, dh):
        rt, n_g = ctx.rt, ctx.n_g
        T, B, D, H = ctx.dims
        M = T * B
        cell = rt.cell
        gates = native.CELL_GATES[cell]
        saved = ctx.saved_tensors
        x2d, pre_raw, save_mean, save_invstd, g_out, bwd_pack, m = saved[:7]
        ws = list(saved[7:7 + n_g])
        gammas = list(saved[7 + n_g:])
        dev = dh.device
        dh = dh.contiguous()
        h = ctx.h_store
        dpre = torch.empty(T, B, n_g * H, device=dev, dtype=torch.float32)
        native.gru_bwd(rt.plan_h, cell, bwd_pack, dh, g_out, h, ctx.drop, ctx.drop_scalar, rt.act, dpre)
        dpre2d = dpre.view(M, n_g * H)
        dgam = [None] * n_g
        dbet = [None] * n_g
        if rt.use_bn:
            dproj = torch.empty_like(dpre2d)
            dgam = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(n_g)]
            dbet = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(n_g)]
            native.bn_bwd(dpre2d, pre_raw, n_g, H, gammas, save_mean, save_invstd, dproj, dgam, dbet, _ws(dev, n_g * H))
        else:
            dproj = dpre2d
        # recurrent weight gradients: dU_g = sum_{t>=1} dpre_g[t]^T (operand of U_g at step t); the candidate's operand
        # is the gated copy m[t] for the two-phase cells, h[t-1] otherwise (and for the update / reset gates)
        dus = [torch.zeros(H, H, device=dev, dtype=torch.float32) for _ in range(n_g)]
        if T > 1:
            h2d = h.view(M, H)
            hi = gates.index('h')
            if m is not None:
                if hi > 0:
                    native.bs_sddmm(rt.plan_h, dpre2d[B:, :hi * H], h2d[:M - B], dus[:hi], rt.mask_h)
                native.bs_sddmm(rt.plan_h, dpre2d[B:, hi * H:(hi + 1) * H], m.view(M, H)[B:], [dus[hi]], rt.mask_h)
            else:
                native.bs_sddmm(rt.plan_h, dpre2d[B:], h2d[:M - B], dus, rt.mask_h)
        dws = [torch.zeros(H, D, device=dev, dtype=torch.float32) for _ in range(n_g)]
        native.bs_sddmm(rt.plan_x, dproj, x2d, dws, rt.mask_x)
        dbias = [None] * n_g
        if ctx.has_bias:
            dbias = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(n_g)]
            native.colsum(dproj, n_g, H, dbias)
        dx = None
        if ctx.needs_input_grad[0]:
            dx2d = torch.empty(M, D, device=dev, dtype=torch.float32)
            native.bsgemm_nn(rt.plan_x, ws, dproj, dx2d)
            dx = dx2d.view(T, B, D)
        return (dx, None, None, *dws, *dus, *dgam, *dbet, *dbias)


class MLPLayerFunction(torch.autograd.Function):
    """One MLP layer: y = drop(act(bn(x W^T + b))) with act possibly LogSoftmax (neural_networks.py:306-317)."""

    @staticmethod
    def forward(ctx, x, w, b, gamma, beta, rt):
        N, D = x.shape
        O = w.shape[0]
        dev = x.device
        x2d = x if x.stride(-1) == 1 else x.contiguous()
        z = torch.empty(N, O, device=dev, dtype=torch.float32)
        cw = rt.q_ws[0] if rt.q_ws is not None else w
        native.bsgemm_nt(rt.plan_x, [cw], x2d, z, [b] if b is not None else None)
        save_mean = save_invstd = None
        if rt.use_bn:
            zb = torch.empty_like(z)
            save_mean = torch.empty(O, device=dev, dtype=torch.float32)
            save_invstd = torch.empty(O, device=dev, dtype=torch.float32)
            native.bn_fwd(z, zb, 1, O, [gamma], [beta], [rt.bn_running[0][0]], [rt.bn_running[0][1]], rt.bn_eps,
                          rt.bn_momentum, rt.training, save_mean, save_invstd, _ws(dev, O))
        else:
            zb = z
        out = torch.empty_like(z)
        if rt.softmax:
            native.logsoftmax_fwd(zb, out, rt.log_prior)
        else:
            native.act_fwd(zb, out, rt.mlp_drop_mask, rt.act)
        if any(ctx.needs_input_grad):
            ctx.rt = rt
            ctx.softmax = rt.softmax
            ctx.had_prior = rt.log_prior is not None
            ctx.drop_mask = rt.mlp_drop_mask
            ctx.out_store = out.detach()
            ctx.save_for_backward(x2d, cw, gamma, z, save_mean, save_invstd)
        return out

    @staticmethod
    def backward(ctx, dout):
        rt = ctx.rt
        x2d, w, gamma, z, save_mean, save_invstd = ctx.saved_tensors
        N, D = x2d.shape
        O = w.shape[0]
        dev = dout.device
        dout = dout.contiguous()
        dzb = torch.empty(N, O, device=dev, dtype=torch.float32)
        if ctx.softmax:
            if ctx.had_prior:
                raise native.NativeError('backward through a prior-normalised posterior is not defined')
            native.logsoftmax_bwd(dout, ctx.out_store, dzb)
        else:
            native.act_bwd(dout, ctx.out_store, ctx.drop_mask, dzb, rt.act)
        dgamma = dbeta = None
        if rt.use_bn:
            dz = torch.empty_like(dzb)
            dgamma = torch.empty(O, device=dev, dtype=torch.float32)
            dbeta = torch.empty(O, device=dev, dtype=torch.float32)
            native.bn_bwd(dzb, z, 1, O, [gamma], save_mean, save_invstd, dz, [dgamma], [dbeta], _ws(dev, O))
        else:
            dz = dzb
        db = None
        if ctx.needs_input_grad[2]:
            db = torch.empty(O, device=dev, dtype=torch.float32)
            native.colsum(dz, 1, O, [db])
        dw = torch.zeros(O, D, device=dev, dtype=torch.float32)
        native.bs_sddmm(rt.plan_x, dz, x2d, [dw], rt.mask_x)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(N, D, device=dev, dtype=torch.float32)
            native.bsgemm_nn(rt.plan_x, [w], dz, dx)
        return dx, dw, db, dgamma, dbeta, None


# ---------------------------------------------------------------------------------------------
# bf16 tensor-core build (b200_precision = bf16): tcgen05 block-sparse GEMMs + mma.sync recurrence.
# Operands (x, W, U, h_{t-1}, dpre) are rounded to bf16, every accumulation and all gate / cell /
# BN / softmax math stays fp32.  Buffers are gate-interleaved and padded (column u*4+g, ldH = pad8(H)).
# ---------------------------------------------------------------------------------------------

def tc_config_for(plan, batch):
    """Cluster shape of the tensor-core recurrence.  Cluster 8 has the lower step latency; cluster 4 keeps
    twice as many groups of 16 sequences co-resident, so it wins as soon as cluster 8 would run in waves."""
    cache = plan.__dict__.setdefault('_tc_cfgs', {})

    def get(cl):
        if cl not in cache:
            try:
                cfg = native.TcConfig(plan, cl, 16 if cl else 0)
                cache[cl] = cfg if (cl == 0 or cfg.cluster_size == cl) else None
            except native.NativeError:
                if cl == 0:
                    raise
                cache[cl] = None
        return cache[cl]

    groups = (batch + 15) // 16
    c8, c4 = get(8), get(4)
    if c8 is not None and (c4 is None or groups <= min(c8.max_clusters_fwd, c8.max_clusters_bwd)):
        return c8
    if c4 is not None:
        return c4
    return get(0)   # raises with the library's message if no cluster shape fits


def _padded_rows(g, T, B, H, ldH):
    """[T,B,H] gradient -> [T*B, ldH] fp32 with 16-byte aligned rows.  A gradient produced by the layer above is
    already a [:, :, :H] view of such a buffer and is used in place; anything else is copied once."""
    M = T * B
    if ldH == H and g.is_contiguous():
        return g.view(M, H)
    if (g.dim() == 3 and g.stride() == (B * ldH, ldH, 1) and (g.storage_offset() * 4) % 16 == 0 and
            g.untyped_storage().nbytes() >= (g.storage_offset() + M * ldH) * 4):
        return g.as_strided((M, ldH), (ldH, 1))
    out = torch.empty(M, ldH, device=g.device, dtype=torch.float32)   # pad columns are ignored by the consumer
    out[:, :H].copy_(g.reshape(M, H))
    return out


_out_info = native.out_info


# The bf16 copy of the most recent LSTM layer output (written by the recurrent kernel next to the fp32 h).  The MLP
# heads that consume that output are separate modules; they recognise their input as an unmodified view of the cached
# buffer (same storage, offset, strides and version counter) and reuse the copy instead of casting h again.
_LAST_H = [None]


def _cached_bf16_of(x, ld):
    c = _LAST_H[0]
    if c is None:
        return None
    h_pad, hb, version = c
    if (x.dim() == 2 and x.stride() == (ld, 1) and hb.shape == (x.shape[0], ld) and x._version == version and
            x.storage_offset() == h_pad.storage_offset() and
            x.untyped_storage().data_ptr() == h_pad.untyped_storage().data_ptr()):
        return hb
    return None


def _bf16_rows(x2d, ld):
    out = torch.empty(x2d.shape[0], ld, device=x2d.device, dtype=torch.bfloat16)
    native.cast_bf16(x2d, out)
    return out


def _rows_2d(x, M, D):
    x2d = x.reshape(M, D)     # a [:, :, :D] view of a padded buffer stays a view (row pitch = the padded width)
    return x2d if x2d.stride(-1) == 1 else x2d.contiguous()


def _split_rows(x2d, ld, hi=None):
    """(hi, lo) bf16 [rows, ld] with hi = bf16(x), lo = bf16(x - hi); an existing hi copy is reused."""
    lo = torch.empty(x2d.shape[0], ld, device=x2d.device, dtype=torch.bfloat16)
    if hi is None:
        hi = torch.empty_like(lo)
        native.split_bf16(x2d, hi, lo)
    else:
        native.split_bf16(x2d, None, lo)
    return hi, lo


def _gemm_nt_fwd(plan, ws, xb, xl, y, biases=None, stats=None):
    """forward product of one layer: two-term operands when xl is given, else the plain bf16 product"""
    n_w = len(ws)
    dev = y.device
    if xl is not None:
        wpack = torch.empty(2 * plan.nz_tiles * 32 * n_w * 32, device=dev, dtype=torch.bfloat16)
        native.tc_pack_nt_split(plan, ws, wpack)
        native.tc_gemm_nt_split(plan, n_w, wpack, xb, xl, y, biases, stats)
    else:
        wpack = torch.empty(plan.nz_tiles * 32 * n_w * 32, device=dev, dtype=torch.bfloat16)
        native.tc_pack_nt(plan, ws, wpack)
        native.tc_gemm_nt(plan, n_w, wpack, xb, y, biases, stats)


class LSTMLayerTCFunction(torch.autograd.Function):
    """One HCGS LSTM layer on the tensor-core build (same math as LSTMLayerFunction)."""

    @staticmethod
    def forward(ctx, x, wf, wi, wo, wc, uf, ui, uo, uc, gf, bf, gi, bi, go, bo, gc, bc, rt):
        T, B, D = x.shape
        H = wf.shape[0]
        M = T * B
        ldH = native.pad8(H)
        ldD = native.pad8(D)
        dev = x.device
        ws = rt.q_ws if rt.q_ws is not None else [wf, wi, wo, wc]
        us = rt.q_us if rt.q_us is not None else [uf, ui, uo, uc]
        need_grad = any(ctx.needs_input_grad)
        xb = rt.xb_in
        if xb is not None and xb.shape != (M, ldD):
            xb = None
        xl = None
        if rt.split_fwd:
            xl = rt.xl_in if (xb is not None and rt.xl_in is not None and rt.xl_in.shape == (M, ldD)) else None
            if xl is None:      # first layer (or a foreign input): one pass over the fp32 rows
                xb, xl = _split_rows(_rows_2d(x, M, D), ldD, xb)
        elif xb is None:
            xb = _bf16_rows(_rows_2d(x, M, D), ldD)
        px = rt.plan_x
        pre_raw = torch.empty(M, 4 * ldH, device=dev, dtype=torch.float32)
        save_mean = save_invstd = bn = None
        if rt.use_bn:
            # statistics come out of the GEMM epilogue (per 128-row tile), the normalisation itself is
            # applied inside the recurrent kernel: the projections are written once and read once
            part = torch.empty((M + 127) // 128, 4 * ldH, 2, device=dev, dtype=torch.float32) if rt.training else None
            _gemm_nt_fwd(px, ws, xb, xl, pre_raw, None, part)
            save_mean = torch.empty(4 * ldH, device=dev, dtype=torch.float32)
            save_invstd = torch.empty(4 * ldH, device=dev, dtype=torch.float32)
            rms = [r[0] for r in rt.bn_running]
            rvs = [r[1] for r in rt.bn_running]
            native.bn_stats_il(part, 128, M, 4, H, ldH, rms, rvs, rt.bn_eps, rt.bn_momentum, rt.training,
                               save_mean, save_invstd)
            bn = (save_mean, save_invstd, [gf, gi, go, gc], [bf, bi, bo, bc])
        else:
            _gemm_nt_fwd(px, ws, xb, xl, pre_raw)
        del xl
        ph = rt.plan_h
        cfg = tc_config_for(ph, B)
        fwd_pack = torch.empty(cfg.fwd_pack_u32, device=dev, dtype=torch.int32)
        bwd_pack = torch.empty(max(cfg.bwd_pack_u32, 1), device=dev, dtype=torch.int32) if need_grad else None
        native.lstm_tc_pack(ph, cfg, us, fwd_pack, bwd_pack)
        h_pad = torch.empty(T, B, ldH, device=dev, dtype=torch.float32)   # rows 16-byte aligned for the bulk stores
        h = h_pad[:, :, :H]
        hb = torch.empty(M, ldH, device=dev, dtype=torch.bfloat16)
        c = gates = None
        if need_grad:
            c = torch.empty(M, ldH, device=dev, dtype=torch.float32)
            gates = torch.empty(M, 4 * ldH, device=dev, dtype=torch.float32)
        # the layer above multiplies with two-term operands: its low term comes out of this recurrence as well
        hl = torch.empty(M, ldH, device=dev, dtype=torch.bfloat16) if rt.want_hl else None
        native.lstm_tc_fwd(ph, cfg, fwd_pack, pre_raw, rt.drop, rt.drop_scalar, rt.act, T, B, h_pad.view(M, ldH), hb, c, gates,
                           bn, hl)
        rt.hb_out = hb
        rt.hl_out = hl
        _LAST_H[0] = (h_pad, hb, h_pad._version)
        if need_grad:
            ctx.rt = rt
            ctx.cfg = cfg
            ctx.dims = (T, B, D, H)
            ctx.drop = rt.drop
            ctx.drop_scalar = rt.drop_scalar
            ctx.save_for_backward(xb, ws[0], ws[1], ws[2], ws[3], gf, gi, go, gc, pre_raw, save_mean, save_invstd,
                                  c, gates, bwd_pack, hb)
        return h

    @staticmethod
    def backward(ctx, dh):
        rt = ctx.rt
        T, B, D, H = ctx.dims
        M = T * B
        ldH = native.pad8(H)
        (xb, wf, wi, wo, wc, gf, gi, go, gc, pre_raw, save_mean, save_invstd, c, gates, bwd_pack,
         hb) = ctx.saved_tensors
        dev = dh.device
        dh_pad = _padded_rows(dh, T, B, H, ldH)
        # only the bf16 dpre exists: it is the operand of the dU product and (rounded once more after the batch-norm
        # backward) of the dW / dx products, so the fp32 copy would be written and read for nothing
        dpre_b = torch.empty(M, 4 * ldH, device=dev, dtype=torch.bfloat16)
        native.lstm_tc_bwd(rt.plan_h, ctx.cfg, bwd_pack, dh_pad, gates, c, ctx.drop, ctx.drop_scalar, rt.act,
                           T, B, None, dpre_b)
        dgam = [None] * 4
        dbet = [None] * 4
        if rt.use_bn:
            dproj_b = torch.empty(M, 4 * ldH, device=dev, dtype=torch.bfloat16)
            dgam = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(4)]
            dbet = [torch.empty(H, device=dev, dtype=torch.float32) for _ in range(4)]
            native.bn_bwd_il(dpre_b, pre_raw, 4, H, ldH, [gf, gi, go, gc], save_mean, save_invstd, None, dproj_b,
                             dgam, dbet, _ws(dev, 4 * ldH))
        else:
            dproj_b = dpre_b
        # the eight weight gradients are views of ONE zero-filled buffer (one fill instead of eight)
        wflat = torch.zeros(4 * H * H + 4 * H * D, device=dev, dtype=torch.float32)
        dus = [wflat[g * H * H:(g + 1) * H * H].view(H, H) for g in range(4)]
        dws = [wflat[4 * H * H + g * H * D:4 * H * H + (g + 1) * H * D].view(H, D) for g in range(4)]
        if T > 1:
            native.tc_sddmm(rt.plan_h, 4, dpre_b[B:], hb[:M - B], dus, rt.mask_h)
        native.tc_sddmm(rt.plan_x, 4, dproj_b, xb, dws, rt.mask_x)
        dx = None
        if ctx.needs_input_grad[0]:
            px = rt.plan_x
            wpack = torch.empty(px.nz_tiles * 32 * 4 * 32, device=dev, dtype=torch.bfloat16)
            native.tc_pack_nn(px, [wf, wi, wo, wc], wpack)
            # rows padded to 16 bytes: the layer below reads its dh through TMA boxes (no repacking pass)
            # (pad columns stay uninitialised: the recurrent backward of the layer below ignores them, include/hcgs_b200.h)
            dx_pad = torch.empty(M, native.pad8(D), device=dev, dtype=torch.float32)
            native.tc_gemm_nn(px, 4, wpack, dproj_b, dx_pad)
            dx = dx_pad[:, :D].view(T, B, D)
        return (dx, dws[0], dws[1], dws[2], dws[3], dus[0], dus[1], dus[2], dus[3],
                dgam[0], dbet[0], dgam[1], dbet[1], dgam[2], dbet[2], dgam[3], dbet[3], None)


class MLPLayerTCFunction(torch.autograd.Function):
    """One MLP layer on the tensor-core build: tcgen05 GEMM with bf16 operands, fp32 everything else."""

    @staticmethod
    def forward(ctx, x, w, b, gamma, beta, rt):
        N, D = x.shape
        O = w.shape[0]
        dev = x.device
        xb = rt.xb_in
        if xb is None or xb.shape != (N, native.pad8(D)):
            xb = _cached_bf16_of(x, native.pad8(D))
        if xb is None:
            xb = _bf16_rows(x if x.stride(-1) == 1 else x.contiguous(), native.pad8(D))
        cw = rt.q_ws[0] if rt.q_ws is not None else w
        px = rt.plan_x
        if rt.softmax and not rt.use_bn and _out_info(px)['cluster'] > 0:
            # output layer: GEMM + log-softmax (+ prior) in one kernel, the logits never reach HBM
            oi = _out_info(px)
            opack = torch.empty(max(oi['pack_bytes'] // 2, 1), device=dev, dtype=torch.bfloat16)
            native.out_pack(px, cw, opack)
            out = torch.empty(N, O, device=dev, dtype=torch.float32)
            native.out_logsoftmax_prior_fwd(px, opack, xb, b, rt.log_prior, out)
            if any(ctx.needs_input_grad):
                ctx.rt = rt
                ctx.softmax = True
                ctx.had_prior = rt.log_prior is not None
                ctx.drop_mask = None
                ctx.out_store = out.detach()
                ctx.save_for_backward(xb, cw, gamma, None, None, None)
                ctx.dims = (N, D, O)
            return out
        z = torch.empty(N, O, device=dev, dtype=torch.float32)
        xl = None
        if rt.split_fwd and not rt.softmax:
            # hidden layers: two-term operands; output (softmax) layers stay single-term like the fused kernel
            _, xl = _split_rows(x if x.stride(-1) == 1 else x.contiguous(), native.pad8(D), xb)
        _gemm_nt_fwd(px, [cw], xb, xl, z, [b] if b is not None else None)
        del xl
        save_mean = save_invstd = None
        if rt.use_bn:
            zb = torch.empty_like(z)
            save_mean = torch.empty(O, device=dev, dtype=torch.float32)
            save_invstd = torch.empty(O, device=dev, dtype=torch.float32)
            native.bn_fwd(z, zb, 1, O, [gamma], [beta], [rt.bn_running[0][0]], [rt.bn_running[0][1]], rt.bn_eps,
                          rt.bn_momentum, rt.training, save_mean, save_invstd, _ws(dev, O))
        else:
            zb = z
        out = torch.empty_like(z)
        if rt.softmax:
            native.logsoftmax_fwd(zb, out, rt.log_prior)
        else:
            native.act_fwd(zb, out, rt.mlp_drop_mask, rt.act)
        if any(ctx.needs_input_grad):
            ctx.rt = rt
            ctx.softmax = rt.softmax
            ctx.had_prior = rt.log_prior is not None
            ctx.drop_mask = rt.mlp_drop_mask
            ctx.out_store = out.detach()
            ctx.save_for_backward(xb, cw, gamma, z if rt.use_bn else None, save_mean, save_invstd)
            ctx.dims = (N, D, O)
        return out

    @staticmethod
    def backward(ctx, dout):
        rt = ctx.rt
        xb, w, gamma, z, save_mean, save_invstd = ctx.saved_tensors
        N, D, O = ctx.dims
        dev = dout.device
        dgamma = dbeta = db = None
        if ctx.softmax and ctx.had_prior:
            raise native.NativeError('backward through a prior-normalised posterior is not defined')
        fused_ok = (ctx.softmax and not rt.use_bn and O % 4 == 0 and O <= 4096 and ctx.out_store.data_ptr() % 16 == 0)
        if dout.is_sparse and fused_ok and dout._nnz() == N:
            # the gradient of an NLL loss in its one-entry-per-row form (graph._NLLSparseGrad): no dense [N, O] gradient
            # is written or read; the kernel traps if entry i is not in row i
            idx = dout._indices()
            dz_b = torch.empty(N, native.pad8(O), device=dev, dtype=torch.bfloat16)
            if ctx.needs_input_grad[2]:
                db = torch.zeros(O, device=dev, dtype=torch.float32)
            native.logsoftmax_bwd_fused_onehot(idx[0].contiguous(), idx[1].contiguous(), dout._values().contiguous(),
                                               ctx.out_store, dz_b, db)
            dout = None
        else:
            if dout.is_sparse:
                dout = dout.to_dense()
            dout = dout.contiguous()
        if dout is None:
            pass
        elif fused_ok and dout.data_ptr() % 16 == 0:
            # log-softmax backward, bias gradient and the bf16 operand of the dW / dX products in one pass
            dz_b = torch.empty(N, native.pad8(O), device=dev, dtype=torch.bfloat16)
            if ctx.needs_input_grad[2]:
                db = torch.zeros(O, device=dev, dtype=torch.float32)
            native.logsoftmax_bwd_fused(dout, ctx.out_store, dz_b, db)
        else:
            dzb = torch.empty(N, O, device=dev, dtype=torch.float32)
            if ctx.softmax:
                native.logsoftmax_bwd(dout, ctx.out_store, dzb)
            else:
                native.act_bwd(dout, ctx.out_store, ctx.drop_mask, dzb, rt.act)
            if rt.use_bn:
                dz = torch.empty_like(dzb)
                dgamma = torch.empty(O, device=dev, dtype=torch.float32)
                dbeta = torch.empty(O, device=dev, dtype=torch.float32)
                native.bn_bwd(dzb, z, 1, O, [gamma], save_mean, save_invstd, dz, [dgamma], [dbeta], _ws(dev, O))
            else:
                dz = dzb
            if ctx.needs_input_grad[2]:
                db = torch.empty(O, device=dev, dtype=torch.float32)
                native.colsum(dz, 1, O, [db])
            dz_b = _bf16_rows(dz, native.pad8(O))
            del dz, dzb
        dw = torch.zeros(O, D, device=dev, dtype=torch.float32)
        native.tc_sddmm(rt.plan_x, 1, dz_b, xb, [dw], rt.mask_x)
        dx = None
        if ctx.needs_input_grad[0]:
            px = rt.plan_x
            wpack = torch.empty(px.nz_tiles * 32 * 32, device=dev, dtype=torch.bfloat16)
            native.tc_pack_nn(px, [w], wpack)
            # rows padded to 16 bytes like the recurrent layers' own buffers: the gradients of several heads are added in
            # place by autograd and the sum is then read through TMA boxes by the layer below without a repacking copy
            dx_pad = torch.empty(N, native.pad8(D), device=dev, dtype=torch.float32)
            native.tc_gemm_nn(px, 1, wpack, dz_b, dx_pad)
            dx = dx_pad[:, :D]
        return dx, dw, db, dgamma, dbeta, None
