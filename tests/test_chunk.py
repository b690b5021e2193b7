"""Chunk pipeline: minibatch assembly with the reference's random left padding (core.py:183-200).

tests/golden/chunk_batches.npz holds the `inp` tensors the reference's own core.run_nn passed to forward_model
(tests/golden/make_golden.py gen_chunk_batches).  CPU: the oracle restatement and the host tables against them, bit for
bit, including the position of Python's `random` stream afterwards.  GPU: hcgs_assemble_batch against the same."""
import os
import random
import sys

import numpy as np
import pytest
import torch

import ref_harness as rh
from helpers import build_product_nets, load_golden, make_graph
from oracle import chunk_oracle

from b200 import chunk as chunk_mod
from b200 import native

CASES = ('a', 'b', 'c')


def _case(g, tag):
    n_snt, bs, seed = (int(v) for v in g['%s/n_snt' % tag])
    ch = rh.synthetic_chunk(n_snt=n_snt, seed=seed)
    want = [g['%s/inp%d' % (tag, i)] for i in range(n_snt // bs)]
    return ch, bs, want, float(g['%s/next_random' % tag][0])


@pytest.mark.parametrize('tag', CASES)
def test_oracle_assembly_matches_reference_batches(tag):
    ch, bs, want, nxt = _case(load_golden('chunk_batches'), tag)
    random.seed(2234)                       # core.py:40
    got = chunk_oracle.assemble_minibatches(ch['data_set'], ch['data_end_index'], bs)
    assert len(got) == len(want)
    for (inp, max_len, _), w in zip(got, want):
        assert inp.shape == w.shape and max_len == w.shape[0]
        assert np.array_equal(inp, w)
    assert random.random() == nxt


@pytest.mark.parametrize('tag', CASES)
def test_chunk_plan_draws_the_reference_offsets(tag):
    ch, bs, want, nxt = _case(load_golden('chunk_batches'), tag)
    random.seed(2234)
    ora = chunk_oracle.assemble_minibatches(ch['data_set'], ch['data_end_index'], bs)
    random.seed(2234)
    plan = chunk_mod.ChunkPlan(ch['data_end_index'], bs)
    assert random.random() == nxt
    assert plan.n_batches == len(want) and plan.rows == ch['data_set'].shape[0]
    for i, (inp, max_len, lefts) in enumerate(ora):
        assert int(plan.max_len[i]) == max_len and list(plan.left[i]) == lefts
        for k in range(bs):                  # the tables describe exactly the reference's tensor
            b, n, l = int(plan.beg[i, k]), int(plan.length[i, k]), int(plan.left[i, k])
            assert np.array_equal(want[i][l:l + n, k], ch['data_set'][b:b + n])
            assert not want[i][:l, k].any() and not want[i][l + n:, k].any()


def test_chunk_plan_edge_cases():
    assert chunk_mod.ChunkPlan([], 4).n_batches == 0
    p = chunk_mod.ChunkPlan([5, 9], 4)          # fewer sentences than one minibatch: nothing to do (core.py:158)
    assert p.n_batches == 0 and p.frames() == 0
    p = chunk_mod.ChunkPlan([3, 3, 10], 3, random.Random(1))   # an empty sentence
    assert list(p.length[0]) == [3, 0, 7] and int(p.max_len[0]) == 7 and list(p.beg[0]) == [0, 3, 3]
    with pytest.raises(ValueError):
        chunk_mod.ChunkPlan([5, 4], 1)
    with pytest.raises(ValueError):
        chunk_mod.ChunkPlan([5, 9], 0)


def test_resident_chunk_has_no_cpu_path():
    ch = rh.synthetic_chunk()
    plan = chunk_mod.ChunkPlan(ch['data_end_index'], 2, random.Random(0))
    with pytest.raises(native.NativeError):
        chunk_mod.ResidentChunk(ch['data_set'], plan, 'cpu')


@pytest.mark.gpu
@pytest.mark.parametrize('tag', CASES)
def test_device_assembly_matches_reference_batches(tag):
    ch, bs, want, nxt = _case(load_golden('chunk_batches'), tag)
    random.seed(2234)
    plan = chunk_mod.ChunkPlan(ch['data_end_index'], bs)
    assert random.random() == nxt
    res = chunk_mod.ResidentChunk(ch['data_set'], plan, 'cuda:0')
    for i, w in enumerate(want):
        got = res.minibatch(i).cpu().numpy()
        assert got.shape == w.shape and np.array_equal(got, w)


@pytest.mark.gpu
@pytest.mark.parametrize('n_snt,bs,D,lo,hi', [(40, 8, 40, 1, 300), (64, 16, 41, 100, 1000), (9, 3, 7, 1, 1),
                                              (2112, 528, 40, 128, 256)])
def test_device_assembly_matches_oracle_ragged(n_snt, bs, D, lo, hi):
    """Vector (dim % 4 == 0) and scalar row widths, single-frame sentences, the bench's minibatch."""
    ch = rh.synthetic_chunk(n_snt=n_snt, D=D - 2, min_len=lo, max_len=hi, seed=n_snt)
    random.seed(7)
    ora = chunk_oracle.assemble_minibatches(ch['data_set'], ch['data_end_index'], bs)
    s = random.getstate()
    random.seed(7)
    plan = chunk_mod.ChunkPlan(ch['data_end_index'], bs)
    assert random.getstate() == s
    res = chunk_mod.ResidentChunk(ch['data_set'], plan, 'cuda:0')
    for i, (inp, max_len, _) in enumerate(ora):
        assert torch.equal(res.minibatch(i).cpu(), torch.from_numpy(inp))
    # rows that would leave the chunk are written as zeros, never read
    bad = torch.tensor([ch['data_set'].shape[0] - 1, -5], device='cuda:0')
    out = native.assemble_batch(res.data, bad, torch.tensor([4, 4], dtype=torch.int32, device='cuda:0'),
                                torch.zeros(2, dtype=torch.int32, device='cuda:0'), 4)
    assert torch.equal(out[0, 0].cpu(), torch.from_numpy(ch['data_set'][-1])) and not out[1:].any() and not out[:, 1].any()


@pytest.mark.gpu
def test_run_chunk_matches_the_loop_over_oracle_batches():
    """b200.chunk.run_chunk == the reference's loop body (core.py:203-232, 250-273) over minibatches assembled by the
    oracle on the host: same losses, same parameters after the chunk."""
    from b200.graph import make_optimizer
    D = 12
    ch = rh.synthetic_chunk(n_snt=10, D=D, seed=3)
    results = []
    for mode in ('device', 'host'):
        nets, secs = build_product_nets('train', D=D, device='cuda:0')
        graph = make_graph(nets, D)
        opts = {n: make_optimizer(nets[n], s) for n, s in zip(nets, secs)}
        random.seed(2234)
        torch.manual_seed(5)
        if mode == 'device':
            loss, err = chunk_mod.run_chunk(graph, opts, ch['data_set'], ch['data_end_index'], 3)
        else:
            ls, es = 0.0, 0.0
            batches = chunk_oracle.assemble_minibatches(ch['data_set'], ch['data_end_index'], 3)
            for inp, max_len, _ in batches:
                outs = graph.forward(torch.from_numpy(inp).cuda(), max_len, 3, 'train')
                for o in opts.values():
                    o.zero_grad()
                outs['loss_final'].backward()
                for o in opts.values():
                    o.step()
                ls += float(outs['loss_final']); es += float(outs['err_final'])
            loss, err = ls / len(batches), es / len(batches)
        results.append((loss, err, {a + '/' + k: v.detach().cpu().clone() for a, n in nets.items() for k, v in n.state_dict().items()},
                        random.random()))
    (l0, e0, p0, r0), (l1, e1, p1, r1) = results
    assert r0 == r1
    assert abs(l0 - l1) <= 1e-5 * abs(l1) and abs(e0 - e1) <= 1e-6
    for k in p1:
        assert torch.allclose(p0[k].float(), p1[k].float(), rtol=1e-4, atol=1e-6), k
    with torch.no_grad():
        lv, ev = chunk_mod.run_chunk(graph, None, ch['data_set'], ch['data_end_index'], 3, to_do='valid')
    assert np.isfinite(lv) and 0.0 <= ev <= 1.0


@pytest.mark.gpu
def test_plans_are_shared_across_model_rebuilds():
    """core.run_nn rebuilds the nets for every chunk (core.py:107-121); with the same masks the second build must find
    every plan in the process-wide store, and a different mask must not."""
    D = 12
    x = torch.randn(6, 3, D + 2, device='cuda:0')
    x[:, :, D:] = 0

    def build_and_run(seed):
        nets, _ = build_product_nets('train', D=D, device='cuda:0', seed=seed)
        torch.manual_seed(1)
        outs = make_graph(nets, D).forward(x, 6, 3, 'train')
        return nets, float(outs['loss_final'])
    nets0, l0 = build_and_run(2234)
    ids0 = {id(h[1]) for n in nets0.values() for h in n._plans._store.values()}
    before = dict(native.PLAN_STORE_STATS)
    nets1, l1 = build_and_run(2234)
    ids1 = {id(h[1]) for n in nets1.values() for h in n._plans._store.values()}
    assert ids0 and ids1 == ids0 and l1 == l0
    assert native.PLAN_STORE_STATS['misses'] == before['misses'] and native.PLAN_STORE_STATS['hits'] > before['hits']
    nets2, _ = build_and_run(7)                      # other masks: new plans
    assert native.PLAN_STORE_STATS['misses'] > before['misses']
    del nets0, nets1, nets2


@pytest.mark.gpu
@pytest.mark.parametrize('t_max,batch,cols,pad', [(9, 5, 1944, 0), (17, 33, 10, 3), (4, 1, 7, 0), (256, 64, 48, 16)])
def test_unpad_rows_packs_every_utterance(t_max, batch, cols, pad):
    """hcgs_unpad_rows: rows t * batch + k (t < len[k]) -> one contiguous [len[k], cols] block per utterance, padding rows
    dropped; bit-exact against indexing, for 16-byte and scalar row widths and a strided source."""
    g = torch.Generator().manual_seed(t_max)
    lens = torch.randint(0 if batch > 4 else 1, t_max + 1, (batch,), generator=g)
    lens[batch // 2] = t_max
    src_full = torch.randn(t_max * batch, cols + pad, generator=g).cuda()
    src = src_full[:, :cols]
    off = torch.cat([torch.zeros(1, dtype=torch.int64), torch.cumsum(lens, 0)[:-1]])
    out = torch.full((int(lens.sum()), cols), float('nan'), device='cuda:0')
    native.unpad_rows(src, off.cuda(), lens.to(torch.int32).cuda(), t_max, out)
    s3 = src.cpu().view(t_max, batch, cols)
    want = torch.cat([s3[:int(lens[k]), k] for k in range(batch)], 0)
    assert torch.equal(out.cpu(), want)
