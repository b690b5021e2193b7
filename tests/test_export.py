"""Posterior export (b200/export.py): batched, right-padded forward == the reference's batch_size=1 forward
(core.py:183-249), records identical to data_io.write_mat, utterance sharding without a collective."""
import numpy as np
import pytest
import torch

from helpers import build_product_nets, make_graph


def test_plan_groups_and_padding_cpu():
    from b200 import export
    lengths = [7, 3, 9, 3, 5, 1]
    groups = export.plan_groups(lengths, 4)
    assert sorted(i for g in groups for i in g) == list(range(6))
    assert all(len(g) <= 4 for g in groups)
    flat = [lengths[i] for g in groups for i in g]
    assert flat == sorted(lengths)                       # ascending length like the reference's chunk order
    capped = export.plan_groups(lengths, 4, max_frames=10)
    assert all(len(g) * max(lengths[i] for i in g) <= 10 or len(g) == 1 for g in capped)
    feats = [torch.full((n, 2), float(i + 1)) for i, n in enumerate(lengths)]
    x = export.pack_group(feats, groups[0])
    assert x.shape == (max(lengths[i] for i in groups[0]), len(groups[0]), 2)
    for b, i in enumerate(groups[0]):
        assert torch.equal(x[:lengths[i], b], feats[i]) and x[lengths[i]:, b].abs().sum() == 0   # right padding
    assert export.plan_groups([], 4) == []


@pytest.mark.gpu
@pytest.mark.parametrize('precision', ['fp32', 'bf16'])
def test_batched_export_equals_single_utterance_forward(precision, tmp_path):
    import posterior_io
    from b200 import export
    D = 12
    over = {'b200_precision': precision}
    nets, _ = build_product_nets('forward', D=D, seed=77, device='cuda', out_folder=str(tmp_path),
                                 lstm_over=over, mlp_over=over)
    torch.manual_seed(5)
    for net in nets.values():                 # non-trivial running statistics so that eval BN matters
        for m in net.modules():
            if isinstance(m, torch.nn.BatchNorm1d):
                m.running_mean.normal_(0, 0.3)
                m.running_var.uniform_(0.5, 1.5)
    counts = np.random.RandomState(3).rand(nets['MLP_layers'].out_dim).astype(np.float32) + 0.1
    nets['MLP_layers'].log_prior = posterior_io.log_prior_tensor(counts, 'cuda')
    graph = make_graph(nets, D)
    lengths = [17, 5, 33, 1, 12, 33, 8]
    feats = [torch.randn(n, D) for n in lengths]
    keys = ['utt_%04d' % i for i in range(len(feats))]

    def fwd(x):
        T, B = x.shape[0], x.shape[1]
        return graph.forward(x, T, B, 'forward', ['out_dnn2'])['out_dnn2']

    # the reference's way: one utterance per forward, batch_size = 1
    single = []
    with torch.no_grad():
        for f in feats:
            single.append(fwd(f.view(-1, 1, D).cuda()).cpu().numpy())
    ex = export.PosteriorExporter(fwd, torch.device('cuda', 0), batch=3)
    got = {i: m.copy() for i, m in ex.run(feats)}     # views of the pinned double buffer: copy to keep
    assert sorted(got) == list(range(len(feats))) and ex.frames == sum(lengths)
    tol = 1e-5 if precision == 'fp32' else 1e-4
    for i, ref in enumerate(single):
        assert got[i].shape == ref.shape and got[i].dtype == np.float32
        assert np.abs(got[i] - ref).max() <= tol, (i, np.abs(got[i] - ref).max())
    # ark records: bytes identical to write_mat of the same matrices, readable back
    ark = str(tmp_path / 'p_to_decode.ark')
    n = export.export_ark(fwd, feats, keys, ark, torch.device('cuda', 0), batch=4)
    assert n == sum(lengths)
    back = posterior_io.read_ark(ark)
    order = sorted(range(len(feats)), key=lambda i: (lengths[i], i))      # ascending length (data_io.py:78)
    assert [k for k, _ in back] == [keys[i] for i in order]
    for (k, m), i in zip(back, order):
        assert m.shape == single[i].shape and np.abs(m - single[i]).max() <= tol
    # two "ranks": disjoint shards that together cover every utterance once
    seen = []
    for r in range(2):
        p = str(tmp_path / ('r%d.ark' % r))
        export.export_ark(fwd, feats, keys, p, torch.device('cuda', 0), batch=4, rank=r, world=2)
        seen += [k for k, _ in posterior_io.read_ark(p)]
    assert sorted(seen) == keys
