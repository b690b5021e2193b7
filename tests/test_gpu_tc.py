"""Tensor-core (bf16 operand) recurrent kernels against a torch reference that rounds the SAME
operands to bf16 (U, h_{t-1} in the forward product; dpre in the backward product) and accumulates
in fp32 -- the "bf16-accumulate mode" of the north star.  Differences are then summation order
plus the occasional bf16 rounding flip of an h element, hence a tight mean and a loose max bound."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import hcgs_oracle, model_oracle  # noqa: E402


def dev():
    return torch.device('cuda', 0)


def bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


class _RoundGrad(torch.autograd.Function):
    """identity whose incoming gradient is rounded to bf16 (the backward kernel's A operand)"""

    @staticmethod
    def forward(ctx, x):
        return x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return bf(g)


def lstm_layer_oracle_bf16(us_r, pre, drop, act):
    T, B, H4 = pre.shape
    H = H4 // 4
    h = torch.zeros(B, H)
    c = h
    hs = []
    for k in range(T):
        hr = h + (bf(h) - h).detach()
        z = [pre[k, :, g * H:(g + 1) * H] + _RoundGrad.apply(hr @ us_r[g].t()) for g in range(4)]
        c = torch.sigmoid(z[1]) * model_oracle.act(act, z[3]) * drop + torch.sigmoid(z[0]) * c
        h = torch.sigmoid(z[2]) * model_oracle.act(act, c)
        hs.append(h)
    return torch.stack(hs)


def check(a, b, mean_tol, max_tol, what):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    scale = max(b.abs().max().item(), 1e-30)
    err = (a - b).abs()
    assert err.mean().item() <= mean_tol * scale, '%s: mean err %.3e (scale %.3e)' % (what, err.mean().item(), scale)
    assert err.max().item() <= max_tol * scale, '%s: max err %.3e (scale %.3e)' % (what, err.max().item(), scale)


def random_mask(rows, cols, blocks, drops, seed):
    np.random.seed(seed)
    return torch.from_numpy(hcgs_oracle.conn_mat(rows, cols, list(blocks), list(drops)).astype(np.float32))


def to_inter(x, T, B, H, ldH):
    """gate-major [T,B,4H] -> gate-interleaved, padded [T*B, 4*ldH] (pad columns zero)"""
    out = torch.zeros(T * B, ldH, 4)
    out[:, :H, :] = x.reshape(T * B, 4, H).transpose(1, 2)
    return out.reshape(T * B, 4 * ldH).contiguous()


def from_inter(x, T, B, H, ldH):
    return x.reshape(T * B, ldH, 4)[:, :H, :].transpose(1, 2).reshape(T, B, 4 * H)


@pytest.mark.parametrize('H,B,T,blocks,drops,act,cl,nb', [
    (24, 3, 7, [8, 4], [50, 50], 'tanh', 4, 16),
    (24, 3, 1, [8, 4], [50, 50], 'tanh', 4, 16),     # T = 1: c_{-1} and the second c buffer are never loaded
    (550, 35, 3, [32, 4], [75, 62.5], 'tanh', 4, 16),  # three clusters, the last one with 3 of 16 sequences
    (40, 19, 9, [8, 4], [50, 50], 'htanh', 4, 16),
    (550, 8, 12, [32, 4], [75, 62.5], 'tanh', 8, 16),
    (550, 20, 6, [32, 4], [75, 62.5], 'tanh', 0, 0),
    (550, 37, 5, [32, 4], [75, 62.5], 'relu', 16, 16),
    (512, 16, 8, [32, 4], [75, 62.5], 'tanh', 0, 0),
    (1024, 20, 5, [32, 4], [75, 62.5], 'tanh', 0, 0),
    (70, 5, 8, [16, 2], [50, 50], 'sigmoid', 4, 16),
])
def test_tc_recurrent_kernels(H, B, T, blocks, drops, act, cl, nb):
    from b200 import native
    torch.manual_seed(H + B + T)
    mask = random_mask(H, H, blocks, drops, 3)
    plan = native.Plan(mask, recurrent=True)
    cfg = native.TcConfig(plan, cl, nb)
    assert cfg.cluster_size in (4, 8, 16) and cfg.nb == 16
    us = [bf(torch.randn(H, H) * (1.5 / np.sqrt(max(mask.sum(1).mean().item(), 1.0))) * mask).requires_grad_(True)
          for _ in range(4)]
    pre = torch.randn(T, B, 4 * H).requires_grad_(True)
    drop = torch.bernoulli(torch.full((B, H), 0.8))
    href = lstm_layer_oracle_bf16(us, pre, drop, act)
    dh = torch.randn(T, B, H)
    href.backward(dh)
    d = dev()
    ldH = native.pad8(H)
    ud = [u.detach().to(d) for u in us]
    fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32)
    bp = torch.empty(max(cfg.bwd_pack_u32, 1), device=d, dtype=torch.int32)
    native.lstm_tc_pack(plan, cfg, ud, fp, bp)
    pre_d = to_inter(pre.detach(), T, B, H, ldH).to(d)
    h_pad = torch.full((T * B, ldH), 5.0, device=d)
    h = h_pad[:, :H]
    c = torch.zeros(T * B, ldH, device=d)
    gates = torch.zeros(T * B, 4 * ldH, device=d)
    hb = torch.full((T * B, ldH), 7.0, device=d, dtype=torch.bfloat16)
    code = native.ACT_CODES[act]
    native.lstm_tc_fwd(plan, cfg, fp, pre_d, drop.to(d), 1.0, code, T, B, h_pad, hb, c, gates)
    torch.cuda.synchronize()
    # fast transcendental forms (abs err ~1e-6) + occasional bf16 rounding flips of h
    check(h.reshape(T, B, H), href, 3e-5, 1e-2, 'h')
    assert torch.equal(hb[:, :H].float().cpu(), bf(h.cpu())), 'bf16 copy of h'
    if ldH > H:
        assert h_pad[:, H:].abs().max().item() == 0, 'pad columns of h must be zero'
    if ldH > H:
        assert hb[:, H:].float().abs().max().item() == 0, 'pad columns of hb must be zero'
    dpre = torch.full((T * B, 4 * ldH), 3.0, device=d)
    dpb = torch.full((T * B, 4 * ldH), 3.0, device=d, dtype=torch.bfloat16)
    dh_pad = torch.full((T * B, ldH), float('nan'), device=d)   # pad columns must be ignored
    dh_pad[:, :H] = dh.to(d).view(T * B, H)
    native.lstm_tc_bwd(plan, cfg, bp, dh_pad, gates, c, drop.to(d), 1.0, code, T, B, dpre, dpb)
    torch.cuda.synchronize()
    check(from_inter(dpre.cpu(), T, B, H, ldH), pre.grad, 1e-4, 3e-2, 'dpre')
    assert torch.equal(dpb.float().cpu(), bf(dpre.cpu())), 'bf16 copy of dpre'
    dpb2 = torch.full((T * B, 4 * ldH), 3.0, device=d, dtype=torch.bfloat16)   # the product path: bf16 dpre only
    native.lstm_tc_bwd(plan, cfg, bp, dh_pad, gates, c, drop.to(d), 1.0, code, T, B, None, dpb2)
    torch.cuda.synchronize()
    assert torch.equal(dpb2, dpb), 'bf16-only dpre'
    if ldH > H:
        assert dpre.view(T * B, ldH, 4)[:, H:, :].abs().max().item() == 0, 'pad columns of dpre must be zero'
    # fused batch norm: (pre - mean) * invstd * gamma + beta inside the kernel == normalising pre beforehand
    mean = torch.randn(4 * ldH) * 0.3
    invstd = torch.rand(4 * ldH) + 0.5
    gam = [torch.rand(H) + 0.5 for _ in range(4)]
    bet = [torch.randn(H) * 0.2 for _ in range(4)]
    raw = torch.randn(T, B, 4 * H)
    raw_i = to_inter(raw, T, B, H, ldH)
    gi = to_inter(torch.cat(gam).view(1, 1, 4 * H), 1, 1, H, ldH).view(-1)
    bi = to_inter(torch.cat(bet).view(1, 1, 4 * H), 1, 1, H, ldH).view(-1)
    normed = ((raw_i - mean) * invstd * gi + bi) * (gi != 0)
    h_a = torch.empty(T * B, ldH, device=d)
    h_b = torch.empty(T * B, ldH, device=d)
    native.lstm_tc_fwd(plan, cfg, fp, normed.to(d), drop.to(d), 1.0, code, T, B, h_a)
    native.lstm_tc_fwd(plan, cfg, fp, raw_i.to(d), drop.to(d), 1.0, code, T, B, h_b,
                       bn=(mean.to(d), invstd.to(d), [g.to(d) for g in gam], [b.to(d) for b in bet]))
    torch.cuda.synchronize()
    check(h_b, h_a, 3e-5, 1e-2, 'fused batch norm')
    # two-term output: hl = bf16(h - float(bf16(h))) next to hb, identical to the separate split pass, same h / hb
    h_c = torch.empty(T * B, ldH, device=d)
    hb_c = torch.full((T * B, ldH), 7.0, device=d, dtype=torch.bfloat16)
    hl_c = torch.full((T * B, ldH), 7.0, device=d, dtype=torch.bfloat16)
    native.lstm_tc_fwd(plan, cfg, fp, pre_d, drop.to(d), 1.0, code, T, B, h_c, hb_c, hl_out=hl_c)
    hi_ref = torch.empty_like(hb_c)
    lo_ref = torch.empty_like(hl_c)
    native.split_bf16(h_c, hi_ref, lo_ref)
    torch.cuda.synchronize()
    assert torch.equal(h_c, h_pad) and torch.equal(hb_c, hb), 'h / hb with the low part requested'
    assert torch.equal(hb_c, hi_ref) and torch.equal(hl_c, lo_ref), 'low part of h == split_bf16(h)'
    # eval variant: scalar drop, only the bf16 output
    hb2 = torch.zeros(T * B, ldH, device=d, dtype=torch.bfloat16)
    native.lstm_tc_fwd(plan, cfg, fp, pre_d, None, 0.8, code, T, B, None, hb2)
    href2 = lstm_layer_oracle_bf16([u.detach() for u in us], pre.detach(), torch.tensor([0.8]), act)
    check(hb2[:, :H].float().view(T, B, H), href2, 2e-3, 2e-2, 'h eval (bf16 output)')
    plan.destroy()


def test_tc_matches_fp32_kernels_on_bf16_exact_inputs():
    """With weights on the 8-bit grid (bf16-exact, cfg 3) and T=1 (h_0 = 0) the two builds must agree to fp32."""
    from b200 import native
    H, B, T = 64, 16, 1
    mask = random_mask(H, H, [8, 4], [50, 50], 5)
    plan = native.Plan(mask, recurrent=True)
    cfg = native.TcConfig(plan)
    d = dev()
    torch.manual_seed(1)
    us = [(torch.randint(-128, 129, (H, H)).float() / 128 * mask).to(d) for _ in range(4)]
    pre = torch.randn(T, B, 4 * H, device=d)
    fp = torch.empty(cfg.fwd_pack_u32, device=d, dtype=torch.int32)
    native.lstm_tc_pack(plan, cfg, us, fp, None)
    h = torch.empty(T * B, H, device=d)   # H = 64: no padding
    native.lstm_tc_fwd(plan, cfg, fp, to_inter(pre.cpu(), T, B, H, H).to(d), None, 1.0, 2, T, B, h)
    fp32 = torch.empty(plan.fwd_pack_floats, device=d)
    native.lstm_pack(plan, us, fp32, None)
    h32 = torch.empty(T, B, H, device=d)
    native.lstm_fwd(plan, fp32, pre, None, 1.0, 2, h32)
    assert (h.view(T, B, H) - h32).abs().max().item() <= 5e-6   # fast exp forms vs libm
