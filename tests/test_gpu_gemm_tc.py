"""tcgen05 block-sparse GEMM family against torch on the SAME bf16-rounded operands (fp32 accumulate)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import hcgs_oracle  # noqa: E402
from test_gpu_tc import bf, dev, random_mask  # noqa: E402


def close(a, b, rtol, what):
    a = a.detach().cpu().double()
    b = b.detach().cpu().double()
    scale = max(b.abs().max().item(), 1e-30)
    err = (a - b).abs().max().item()
    assert err <= rtol * scale, '%s: max err %.3e (scale %.3e)' % (what, err, scale)


CASES = [
    # n_out, n_in, blocks, drops, n_w, M
    (64, 64, [32, 4], [50, 50], 4, 128),
    (64, 96, [32, 4], [50, 50], 1, 200),
    (550, 550, [32, 4], [75, 62.5], 4, 1000),
    (550, 40, [32, 4], [75, 62.5], 4, 333),
    (1944, 550, [64, 4], [50, 25], 1, 517),
    (48, 550, None, None, 1, 300),
]


@pytest.mark.parametrize('n_out,n_in,blocks,drops,n_w,M', CASES)
def test_tc_gemm_nt(n_out, n_in, blocks, drops, n_w, M):
    from b200 import native
    torch.manual_seed(n_out + n_in + M)
    mask = random_mask(n_out, n_in, blocks, drops, 7) if blocks else torch.ones(n_out, n_in)
    plan = native.Plan(mask)
    d = dev()
    ws = [torch.randn(n_out, n_in) * mask for _ in range(n_w)]
    x = torch.randn(M, n_in)
    bias = [torch.randn(n_out) for _ in range(n_w)] if n_w == 1 else None
    ldx = native.pad8(n_in)
    xb = torch.full((M, ldx), 5.0, device=d, dtype=torch.bfloat16)
    native.cast_bf16(x.to(d), xb)
    assert torch.equal(xb[:, :n_in].float().cpu(), bf(x)) and (ldx == n_in or xb[:, n_in:].float().abs().max() == 0)
    wpack = torch.empty(max(plan.nz_tiles, 1) * 32 * n_w * 32, device=d, dtype=torch.bfloat16)
    native.tc_pack_nt(plan, [w.to(d) for w in ws], wpack)
    ldo = native.pad8(n_out) if n_w == 4 else n_out
    y = torch.full((M, ldo * n_w), 9.0, device=d)
    native.tc_gemm_nt(plan, n_w, wpack, xb, y, [b.to(d) for b in bias] if bias else None)
    torch.cuda.synchronize()
    ref = torch.stack([bf(x) @ bf(w).t() + (bias[g] if bias else 0) for g, w in enumerate(ws)], 2)  # [M, n_out, n_w]
    got = y.cpu().view(M, ldo, n_w)[:, :n_out, :]
    close(got, ref, 2e-5, 'nt')
    if ldo > n_out and n_w == 4:
        assert y.view(M, ldo, n_w)[:, n_out:, :].abs().max().item() == 0, 'pad columns of a padded output must be zero'
    if n_w == 4:
        # fused batch-norm statistics: per-tile partials from the epilogue, merged by bn_stats_il
        part = torch.full(((M + 127) // 128, ldo * 4, 2), 7.0, device=d)
        y2 = torch.empty_like(y)
        native.tc_gemm_nt(plan, 4, wpack, xb, y2, None, part)
        assert torch.equal(y2, y)
        sm = torch.empty(ldo * 4, device=d)
        si = torch.empty(ldo * 4, device=d)
        rm = [torch.zeros(n_out, device=d) for _ in range(4)]
        rv = [torch.ones(n_out, device=d) for _ in range(4)]
        native.bn_stats_il(part, 128, M, 4, n_out, ldo, rm, rv, 1e-5, 0.05, True, sm, si)
        torch.cuda.synchronize()
        yy = y.cpu().view(M, ldo, 4)[:, :n_out, :].double()
        mean, var = yy.mean(0), yy.var(0, unbiased=False)
        close(sm.cpu().view(ldo, 4)[:n_out], mean, 1e-5, 'bn mean')
        close(si.cpu().view(ldo, 4)[:n_out], 1.0 / torch.sqrt(var + 1e-5), 1e-5, 'bn invstd')
        for g in range(4):
            close(rm[g].cpu(), 0.05 * mean[:, g], 1e-5, 'running mean')
            close(rv[g].cpu(), 0.95 + 0.05 * yy.var(0, unbiased=True)[:, g], 1e-5, 'running var')
        # eval: the running statistics come back as mean / invstd
        native.bn_stats_il(None, 128, M, 4, n_out, ldo, rm, rv, 1e-5, 0.05, False, sm, si)
        close(sm.cpu().view(ldo, 4)[:n_out, 2], rm[2].cpu(), 1e-7, 'eval mean')
        close(si.cpu().view(ldo, 4)[:n_out, 1], 1.0 / torch.sqrt(rv[1].cpu() + 1e-5), 1e-6, 'eval invstd')
    plan.destroy()


@pytest.mark.parametrize('n_out,n_in,blocks,drops,n_w,M', CASES)
def test_tc_gemm_nt_split(n_out, n_in, blocks, drops, n_w, M):
    """two-term operands: x_hi w_hi + x_lo w_hi + x_hi w_lo against float64 on the same split operands (2e-5) and
    against the un-rounded fp32 product (the point of the split: 2^-15 instead of 2^-8 per operand)."""
    from b200 import native
    torch.manual_seed(n_out + n_in + M + 2)
    mask = random_mask(n_out, n_in, blocks, drops, 7) if blocks else torch.ones(n_out, n_in)
    plan = native.Plan(mask)
    d = dev()
    ws = [torch.randn(n_out, n_in) * mask for _ in range(n_w)]
    x = torch.randn(M, n_in)
    bias = [torch.randn(n_out) for _ in range(n_w)] if n_w == 1 else None
    ldx = native.pad8(n_in)
    xh = torch.full((M, ldx), 5.0, device=d, dtype=torch.bfloat16)
    xl = torch.full((M, ldx), 5.0, device=d, dtype=torch.bfloat16)
    native.split_bf16(x.to(d), xh, xl)
    xh_ref = bf(x)
    xl_ref = bf(x - xh_ref)
    assert torch.equal(xh[:, :n_in].float().cpu(), xh_ref) and torch.equal(xl[:, :n_in].float().cpu(), xl_ref)
    assert ldx == n_in or (xh[:, n_in:].float().abs().max() == 0 and xl[:, n_in:].float().abs().max() == 0)
    # lo only (the hi copy already exists: the recurrent kernel's bf16 h)
    xl2 = torch.empty_like(xl)
    native.split_bf16(x.to(d), None, xl2)
    assert torch.equal(xl2, xl)
    wpack = torch.empty(2 * max(plan.nz_tiles, 1) * 32 * n_w * 32, device=d, dtype=torch.bfloat16)
    native.tc_pack_nt_split(plan, [w.to(d) for w in ws], wpack)
    ldo = native.pad8(n_out) if n_w == 4 else n_out
    y = torch.full((M, ldo * n_w), 9.0, device=d)
    part = torch.full(((M + 127) // 128, ldo * 4, 2), 7.0, device=d) if n_w == 4 else None
    native.tc_gemm_nt_split(plan, n_w, wpack, xh, xl, y, [b.to(d) for b in bias] if bias else None, part)
    torch.cuda.synchronize()
    refs, exact = [], []
    for g, w in enumerate(ws):
        wh = bf(w)
        wl = bf(w - wh)
        b = bias[g].double() if bias else 0
        refs.append(xh_ref.double() @ wh.double().t() + xl_ref.double() @ wh.double().t() +
                    xh_ref.double() @ wl.double().t() + b)
        exact.append(x.double() @ w.double().t() + b)
    got = y.cpu().view(M, ldo, n_w)[:, :n_out, :]
    close(got, torch.stack(refs, 2), 2e-5, 'nt split vs split operands')
    close(got, torch.stack(exact, 2), 6e-5, 'nt split vs exact product')
    if ldo > n_out and n_w == 4:
        assert y.view(M, ldo, n_w)[:, n_out:, :].abs().max().item() == 0
    if n_w == 4:
        sm = torch.empty(ldo * 4, device=d)
        si = torch.empty(ldo * 4, device=d)
        rm = [torch.zeros(n_out, device=d) for _ in range(4)]
        rv = [torch.ones(n_out, device=d) for _ in range(4)]
        native.bn_stats_il(part, 128, M, 4, n_out, ldo, rm, rv, 1e-5, 0.05, True, sm, si)
        yy = y.cpu().view(M, ldo, 4)[:, :n_out, :].double()
        close(sm.cpu().view(ldo, 4)[:n_out], yy.mean(0), 1e-5, 'bn mean (split)')
    plan.destroy()


def _dy_layout(dy_list, M, n_out, n_w, ld_units):
    """list of per-gate [M, n_out] -> bf16-able fp32 [M, ld_units*n_w], column u*n_w+g, pad columns zero"""
    out = torch.zeros(M, ld_units, n_w)
    for g, t in enumerate(dy_list):
        out[:, :n_out, g] = t
    return out.view(M, ld_units * n_w)


@pytest.mark.parametrize('n_out,n_in,blocks,drops,n_w,M', CASES)
def test_tc_gemm_nn_and_sddmm(n_out, n_in, blocks, drops, n_w, M):
    from b200 import native
    torch.manual_seed(n_out + n_in + M + 1)
    mask = random_mask(n_out, n_in, blocks, drops, 7) if blocks else torch.ones(n_out, n_in)
    plan = native.Plan(mask)
    d = dev()
    ws = [torch.randn(n_out, n_in) * mask for _ in range(n_w)]
    x = torch.randn(M, n_in)
    dys = [torch.randn(M, n_out) for _ in range(n_w)]
    ldu = native.pad8(n_out)
    dyb = torch.empty(M, ldu * n_w, device=d, dtype=torch.bfloat16)
    native.cast_bf16(_dy_layout(dys, M, n_out, n_w, ldu).to(d), dyb)
    xb = torch.empty(M, native.pad8(n_in), device=d, dtype=torch.bfloat16)
    native.cast_bf16(x.to(d), xb)
    # data gradient
    wpack = torch.empty(max(plan.nz_tiles, 1) * 32 * n_w * 32, device=d, dtype=torch.bfloat16)
    native.tc_pack_nn(plan, [w.to(d) for w in ws], wpack)
    dx = torch.full((M, n_in), 9.0, device=d)
    native.tc_gemm_nn(plan, n_w, wpack, dyb, dx)
    torch.cuda.synchronize()
    ref = sum(bf(dys[g]) @ bf(ws[g]) for g in range(n_w))
    close(dx, ref, 3e-5, 'nn')
    # weight gradient (two launches accumulate: atomics into a caller-zeroed buffer)
    dws = [torch.zeros(n_out, n_in, device=d) for _ in range(n_w)]
    native.tc_sddmm(plan, n_w, dyb, xb, dws, mask.to(d))
    torch.cuda.synchronize()
    for g in range(n_w):
        refw = (bf(dys[g]).t() @ bf(x)) * mask
        close(dws[g], refw, 3e-5, 'sddmm gate %d' % g)
    # shifted operands = the recurrent weight gradient dU = sum_t dpre[t]^T h[t-1]
    if M > 40:
        dws2 = [torch.zeros(n_out, n_in, device=d) for _ in range(n_w)]
        native.tc_sddmm(plan, n_w, dyb[8:], xb[:M - 8], dws2, None)
        torch.cuda.synchronize()
        keep = torch.zeros(n_out, n_in)
        tr, tc_ = (n_out + 31) // 32, (n_in + 31) // 32
        for i in range(tr):
            for j in range(tc_):
                if mask[i * 32:(i + 1) * 32, j * 32:(j + 1) * 32].any():
                    keep[i * 32:(i + 1) * 32, j * 32:(j + 1) * 32] = 1
        for g in range(n_w):
            refw = (bf(dys[g])[8:].t() @ bf(x)[:M - 8]) * keep
            close(dws2[g], refw, 3e-5, 'sddmm shifted gate %d' % g)
    plan.destroy()


OUT_CASES = [
    # n_out (pdfs), n_in, blocks, drops, M, prior, bias
    (1944, 550, [64, 4], [50, 25], 517, True, True),     # TIMIT cd head: cluster 8 x 4 strips
    (48, 550, None, None, 300, False, True),             # monophone head: one CTA, one strip
    (3480, 1024, [64, 4], [50, 25], 260, True, True),    # LibriSpeech head: cluster 8 x 7 strips (448 TMEM columns)
    (1000, 72, [32, 4], [75, 62.5], 129, True, False),   # ragged strips, rows of the mask with no kept tile
    (130, 40, [64, 4], [50, 25], 7, False, False),       # in=40 < block 64: the reference's all-zero mask
    (4096, 64, None, None, 128, True, True),             # widest supported: 8 x 8 strips, all 512 TMEM columns
]


@pytest.mark.parametrize('n_out,n_in,blocks,drops,M,prior,use_bias', OUT_CASES)
def test_out_logsoftmax_prior_fused(n_out, n_in, blocks, drops, M, prior, use_bias):
    """Fused output layer (hcgs_out_logsoftmax_prior_fwd) vs torch on the same bf16-rounded operands and vs the
    unfused pair tc_gemm_nt + logsoftmax_fwd.  Tolerance: max-abs 1e-4 on the log-posteriors (north star)."""
    from b200 import native
    torch.manual_seed(n_out + n_in + M + 2)
    mask = random_mask(n_out, n_in, blocks, drops, 11) if blocks else torch.ones(n_out, n_in)
    plan = native.Plan(mask)
    info = native.out_info(plan)
    assert info['cluster'] in (1, 2, 4, 8) and 1 <= info['strips'] <= 8
    assert info['cluster'] * info['strips'] * 64 >= n_out
    d = dev()
    w = torch.randn(n_out, n_in) * mask * (2.0 / n_in ** 0.5)
    x = torch.randn(M, n_in)
    bias = torch.randn(n_out) if use_bias else None
    counts = torch.rand(n_out) + 0.01
    log_prior = torch.log(counts / counts.sum()) if prior else None
    xb = torch.empty(M, native.pad8(n_in), device=d, dtype=torch.bfloat16)
    native.cast_bf16(x.to(d), xb)
    opack = torch.empty(max(info['pack_bytes'] // 2, 1), device=d, dtype=torch.bfloat16)
    native.out_pack(plan, w.to(d), opack)
    out = torch.full((M, n_out), 9.0, device=d)
    native.out_logsoftmax_prior_fwd(plan, opack, xb, bias.to(d) if use_bias else None,
                                    log_prior.to(d) if prior else None, out)
    torch.cuda.synchronize()
    z = (bf(x).double() @ bf(w).double().t()) + (bias.double() if use_bias else 0)
    ref = torch.log_softmax(z, 1) - (log_prior.double() if prior else 0)
    err = (out.cpu().double() - ref).abs().max().item()
    assert err <= 1e-4, 'fused output layer: max-abs %.3e' % err
    # the unfused pair computes the same thing through HBM
    wpack = torch.empty(max(plan.nz_tiles, 1) * 32 * 32, device=d, dtype=torch.bfloat16)
    native.tc_pack_nt(plan, [w.to(d)], wpack)
    z2 = torch.empty(M, n_out, device=d)
    native.tc_gemm_nt(plan, 1, wpack, xb, z2, [bias.to(d)] if use_bias else None)
    out2 = torch.empty_like(z2)
    native.logsoftmax_fwd(z2, out2, log_prior.to(d) if prior else None)
    torch.cuda.synchronize()
    assert (out - out2).abs().max().item() <= 2e-5
    # every row is a normalised distribution once the prior is added back
    p = out.cpu().double() + (log_prior.double() if prior else 0)
    assert (p.exp().sum(1) - 1).abs().max().item() <= 1e-5
    plan.destroy()


def test_out_fused_too_wide_reports():
    from b200 import native
    plan = native.Plan(torch.ones(4100, 32))
    assert native.out_info(plan)['cluster'] == 0
    with pytest.raises(native.NativeError):
        native.out_pack(plan, torch.zeros(4100, 32, device=dev()), torch.empty(8, device=dev(), dtype=torch.bfloat16))
    plan.destroy()
