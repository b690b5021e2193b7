"""hcgs_bn_fwd_static: the allocation-free variant of hcgs_bn_fwd (statistics scratch in a static device array of the
library; include/hcgs_b200.h).  It has no counterpart in the reference -- it exists for callers that cannot call
cudaMalloc -- so the check is against the existing entry point hcgs_bn_fwd on the same inputs, through the C ABI,
on two calls in a row (the second call finds the first call's contents in the shared scratch), within the tolerance
of the other batch-norm tests (2e-5 of the tensor's scale).

Collected last on purpose (file name)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _err(a, b):
    a, b = a.detach().cpu().double(), b.detach().cpu().double()
    return (a - b).abs().max().item(), max(b.abs().max().item(), 1e-30)


def _inputs(M, groups, cols, seed, d):
    g = torch.Generator().manual_seed(seed)
    C = groups * cols
    x = (torch.randn(M, C, generator=g) * 2 + 0.5).to(d)
    gam = [(torch.rand(cols, generator=g) + 0.5).to(d) for _ in range(groups)]
    bet = [torch.randn(cols, generator=g).to(d) for _ in range(groups)]
    return x, gam, bet


def _run(fn_static, native, x, gam, bet, groups, cols, training, rm0, rv0):
    """one call of either entry point into fresh output buffers; returns everything the call writes"""
    d = x.device
    M, C = x.shape
    y = torch.full((M, C), 9.0, device=d)
    sm = torch.full((C,), 9.0, device=d)
    si = torch.full((C,), 9.0, device=d)
    rm = [t.clone() for t in rm0]
    rv = [t.clone() for t in rv0]
    if fn_static:
        native.bn_fwd_static(x, y, groups, cols, gam, bet, rm, rv, 1e-5, 0.05, training, sm, si)
    else:
        ws = torch.empty(2 * C * 64, device=d)
        native.bn_fwd(x, y, groups, cols, gam, bet, rm, rv, 1e-5, 0.05, training, sm, si, ws)
    return y, sm, si, rm, rv


def test_bn_fwd_static_matches_bn_fwd_on_two_calls_in_a_row():
    from b200 import native
    d = torch.device('cuda', 0)
    lib = native.load()
    assert hasattr(lib, 'hcgs_bn_fwd_static')
    # call 1 fills most of the scratch (4 x 550 columns, 64 row splits); call 2 is smaller, with other row splits, and
    # finds call 1's partials in the array; call 3 is the eval form (running statistics, no scratch)
    cases = [(4097, 4, 550, True), (333, 1, 1944, True), (200, 4, 24, False)]
    prepared = []
    for k, (M, groups, cols, training) in enumerate(cases):
        x, gam, bet = _inputs(M, groups, cols, 10 + k, d)
        rm0 = [torch.randn(cols, device=d) * 0.1 for _ in range(groups)]
        rv0 = [torch.rand(cols, device=d) + 0.5 for _ in range(groups)]
        want = _run(False, native, x, gam, bet, groups, cols, training, rm0, rv0)
        prepared.append((x, gam, bet, groups, cols, training, rm0, rv0, want))
    torch.cuda.synchronize()
    got_all = []
    for (x, gam, bet, groups, cols, training, rm0, rv0, _) in prepared:
        M, C = x.shape
        # every buffer the call writes exists before the call: inside it nothing may be allocated
        y = torch.full((M, C), 9.0, device=d)
        sm = torch.full((C,), 9.0, device=d)
        si = torch.full((C,), 9.0, device=d)
        rm = [t.clone() for t in rm0]
        rv = [t.clone() for t in rv0]
        torch.cuda.synchronize()
        n_alloc = torch.cuda.memory_stats(d)['allocation.all.allocated']
        n_bytes = torch.cuda.memory_allocated(d)
        native.bn_fwd_static(x, y, groups, cols, gam, bet, rm, rv, 1e-5, 0.05, training, sm, si)
        assert torch.cuda.memory_stats(d)['allocation.all.allocated'] == n_alloc, 'the call allocated through torch'
        assert torch.cuda.memory_allocated(d) == n_bytes
        got_all.append((y, sm, si, rm, rv))
    torch.cuda.synchronize()
    for k, ((_, _, _, groups, cols, training, _, _, want), got) in enumerate(zip(prepared, got_all)):
        names = ['y'] + (['save_mean', 'save_invstd'] if training else [])
        for name, a, b in zip(names, got[:3], want[:3]):
            mx, sc = _err(a, b)
            assert mx <= 2e-5 * sc + 1e-6, 'call %d %s: %.3e (scale %.3e)' % (k, name, mx, sc)
        for g in range(groups):
            for name, a, b in (('running_mean', got[3][g], want[3][g]), ('running_var', got[4][g], want[4][g])):
                mx, sc = _err(a, b)
                assert mx <= 2e-5 * sc + 1e-6, 'call %d %s[%d]: %.3e (scale %.3e)' % (k, name, g, mx, sc)
        assert torch.isfinite(got[0]).all() and (got[0] != 9.0).any()


def test_bn_fwd_static_rejects_more_columns_than_the_static_scratch_holds():
    from b200 import native
    d = torch.device('cuda', 0)
    groups, cols = 4, native.BN_STATIC_MAX_COLS // 4 + 1          # 4100 columns > 4096
    x, gam, bet = _inputs(64, groups, cols, 3, d)
    rm = [torch.zeros(cols, device=d) for _ in range(groups)]
    rv = [torch.ones(cols, device=d) for _ in range(groups)]
    y = torch.empty_like(x)
    sm = torch.empty(groups * cols, device=d)
    si = torch.empty(groups * cols, device=d)
    with pytest.raises(native.NativeError, match='exceed the static scratch'):
        native.bn_fwd_static(x, y, groups, cols, gam, bet, rm, rv, 1e-5, 0.05, True, sm, si)
    # exactly the maximum is accepted
    cols = native.BN_STATIC_MAX_COLS // 4
    x, gam, bet = _inputs(130, groups, cols, 4, d)
    rm = [torch.zeros(cols, device=d) for _ in range(groups)]
    rv = [torch.ones(cols, device=d) for _ in range(groups)]
    want = _run(False, native, x, gam, bet, groups, cols, True, rm, rv)
    got = _run(True, native, x, gam, bet, groups, cols, True, rm, rv)
    torch.cuda.synchronize()
    mx, sc = _err(got[0], want[0])
    assert mx <= 2e-5 * sc + 1e-6
