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
    plan.destroy()
