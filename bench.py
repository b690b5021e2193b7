#!/usr/bin/env python
"""Headline benchmark: HCGS-LSTM train frames/s on B200 (BASELINE.json configs[1], TIMIT_CGS LSTM).

    python bench.py --gpus N --steps K --warmup W            # our arm (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...   # reference arm: CPU path on the host cores

One "step" = one minibatch through the whole hot path: forward (4x550 HCGS LSTM + 1944-way and
48-way MLP heads), loss, backward, RMSprop step -- the body of core.run_nn's training loop.
Prints ONE JSON line (rank 0).
"""
import argparse
import gc
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, 'pytorch-kaldi-cgs_b200'), ROOT, os.path.join(ROOT, 'tests')):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

D_FEA, N_PDF, N_MONO = 40, 1944, 48
LAYERS, HID = 4, 550
DROP, CONFIG = '0.2', 'timit'


def select_config(name):
    """timit = BASELINE configs[1] (the headline workload); libri = configs[3]: 5x1024 HCGS LSTM, 3480 pdfs, single
    cd head, no dropout (cfg/LibriSpeech_CGS/libri_LSTM_fmllr_460.cfg:93-94, 157-160; BN on)."""
    global N_PDF, N_MONO, LAYERS, HID, DROP, CONFIG, MODEL
    CONFIG = name
    if name == 'timit':
        N_PDF, N_MONO, LAYERS, HID, DROP = 1944, 48, 4, 550, '0.2'
        MODEL = list(MODEL_TIMIT)
    if name == 'libri':
        N_PDF, N_MONO, LAYERS, HID, DROP = 3480, 0, 5, 1024, '0.0'
        MODEL = ['out_dnn1=compute(LSTM_layers,fbank)', 'out_dnn2=compute(MLP_layers,out_dnn1)',
                 'loss_final=cost_nll(out_dnn2,lab_cd)', 'err_final=cost_err(out_dnn2,lab_cd)']


def cfg_sections(to_do='train', precision='bf16'):
    lay = ','.join([str(HID)] * LAYERS)
    rep = lambda v: ','.join([v] * LAYERS)  # noqa: E731
    lstm = dict(arch_name='LSTM_layers', lstm_lay=lay, lstm_drop=rep(DROP), lstm_use_laynorm_inp='False',
                lstm_use_batchnorm_inp='False', lstm_use_laynorm=rep('False'), lstm_use_batchnorm=rep('True'),
                lstm_bidir='False', lstm_act=rep('tanh'), if_hsigmoid='True', lstm_orthinit='True', arch_lr='0.0016',
                arch_opt='rmsprop', opt_momentum='0.0', opt_alpha='0.95', opt_eps='1e-8', opt_centered='False',
                opt_weight_decay='0.0', out_folder='', lstm_hcgs='True', guided_hcgs='False',
                apply_guided_hcgs='False', hcgsx_block='32,4', hcgsx_sparse='75,62.5', hcgsh_block='32,4',
                hcgsh_sparse='75,62.5', lstm_quant='False', param_quant=rep('8'), lstm_quant_inp='False',
                inp_quant='16', lstm_prune='False', lstm_prune_perc='70', skip_regularization='True', to_do=to_do,
                use_cuda='True', b200_precision=precision)
    mlp = dict(arch_name='MLP_layers', dnn_lay=str(N_PDF), dnn_drop='0.0', dnn_use_laynorm_inp='False',
               dnn_use_batchnorm_inp='False', dnn_use_batchnorm='False', dnn_use_laynorm='False', dnn_act='softmax',
               arch_lr='0.0016', arch_opt='rmsprop', opt_momentum='0.0', opt_alpha='0.95', opt_eps='1e-8',
               opt_centered='False', opt_weight_decay='0.0', out_folder='', mlp_hcgs='True', guided_hcgs='False',
               apply_guided_hcgs='False', hcgs_block='64,4', hcgs_sparse='50,25', mlp_quant='False', param_quant='8',
               mlp_quant_inp='False', inp_quant='16', mlp_prune='False', mlp_prune_perc='70',
               skip_regularization='True', to_do=to_do, use_cuda='True', b200_precision=precision)
    mono = dict(mlp, arch_name='MLP_layers2', dnn_lay=str(max(N_MONO, 1)), mlp_hcgs='False', arch_lr='0.0004')
    return lstm, mlp, mono


MODEL_TIMIT = ['out_dnn1=compute(LSTM_layers,fbank)', 'out_dnn2=compute(MLP_layers,out_dnn1)',
         'out_dnn3=compute(MLP_layers2,out_dnn1)', 'loss_mono=cost_nll(out_dnn3,lab_mono)',
         'loss_mono_w=mult_constant(loss_mono,1.0)', 'loss_cd=cost_nll(out_dnn2,lab_cd)',
         'loss_final=sum(loss_cd,loss_mono_w)', 'err_final=cost_err(out_dnn2,lab_cd)']
MODEL = list(MODEL_TIMIT)


def synthetic_minibatch(T, B, seed):
    """N(0,1) features, uniform pdf labels as trailing float columns, sequence lengths
    U[T/2, T] sorted ascending, zero padding (random left offset like core.py:193-197)."""
    g = torch.Generator().manual_seed(seed)
    inp = torch.zeros(T, B, D_FEA + 2)
    lens = torch.sort(torch.randint(max(1, T // 2), T + 1, (B,), generator=g))[0]
    frames = 0
    for b in range(B):
        n = int(lens[b])
        left = int(torch.randint(0, T - n + 1, (1,), generator=g))
        inp[left:left + n, b, :D_FEA] = torch.randn(n, D_FEA, generator=g)
        inp[left:left + n, b, D_FEA] = torch.randint(0, N_PDF, (n,), generator=g).float()
        inp[left:left + n, b, D_FEA + 1] = torch.randint(0, max(N_MONO, 1), (n,), generator=g).float()
        frames += n
    return inp, frames


def synthetic_chunk(n_snt, T, seed):
    """One chunk as data_io.read_lab_fea hands it to core.run_nn (data_io.py:276-282): sentences of U[T/2, T] frames
    sorted by length, [rows, D_FEA + 2] fp32 (features, then the pdf / monophone labels as float columns)."""
    g = np.random.RandomState(seed)
    lens = np.sort(g.randint(max(1, T // 2), T + 1, n_snt))
    rows = int(lens.sum())
    data = np.empty((rows, D_FEA + 2), dtype=np.float32)
    data[:, :D_FEA] = g.standard_normal((rows, D_FEA)).astype(np.float32)
    data[:, D_FEA] = g.randint(0, N_PDF, rows)
    data[:, D_FEA + 1] = g.randint(0, max(N_MONO, 1), rows)
    return data, np.cumsum(lens)


def run_chunk_leg(graph, opts, dev, steps, T, B):
    """one training chunk of `steps` minibatches through b200.chunk.run_chunk -- the chunk uploaded from
    pinned host memory once, every minibatch assembled on the device (hcgs_assemble_batch), loss / error read once."""
    import random
    from b200 import chunk as bchunk
    data, end = synthetic_chunk(steps * B, T, 4321)
    host = torch.from_numpy(data).pin_memory()
    random.seed(2234)
    # warm-up: the whole chunk once -- its minibatches all have different lengths, and the first time a length is seen the
    # caching allocator goes to cudaMalloc (16 calls, 9-140 ms per chunk depending on what it holds); epochs repeat lengths
    bchunk.run_chunk(graph, opts, host, end, B)
    torch.cuda.synchronize()
    random.seed(2234)
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    gc.collect()
    gc.disable()       # like the other timed regions: the host is within 15 % of the GPU time on the short minibatches
    t0 = time.perf_counter()
    c0.record()
    loss, err = bchunk.run_chunk(graph, opts, host, end, B)
    c1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3
    gc.enable()
    ms = c0.elapsed_time(c1)
    assert np.isfinite(loss) and 0.0 <= err <= 1.0, (loss, err)
    plan = bchunk.ChunkPlan(end, B, random.Random(0))
    return {'workload': 'one training chunk (core.py:156-273) of %d minibatches x %d sentences of U[%d, %d] frames sorted by '
                        'length: chunk uploaded once, minibatches assembled on the device with the random left padding of '
                        'core.py:183-200, loss / error read once per chunk; second pass over the chunk (allocator warm)' % (steps, B, max(1, T // 2), T),
            'value': int(end[-1]) / (ms * 1e-3), 'unit': 'frames/s', 'ms_per_minibatch': ms / steps, 'ms_wall': wall,
            'minibatches': steps, 'frames': int(end[-1]), 'padded_rows': int(plan.max_len.sum()) * B,
            'h2d_bytes_per_chunk': int(data.nbytes), 'd2h_bytes_per_chunk': 8, 'loss': loss}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.idx), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', '100'],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for l in self.lines:
            f = [x.strip() for x in l.split(',')]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx = float(f[2])
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith('active'):
                    reasons.add(n)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': mx, 'reasons': sorted(reasons),
                'samples': len(sm)}


def build_workload(device, seed=2234, precision='bf16'):
    import neural_networks as nb
    from b200.graph import Graph, make_optimizer
    torch.manual_seed(seed)
    np.random.seed(seed)
    s_lstm, s_cd, s_mono = cfg_sections('train', precision)
    lstm = nb.LSTM(s_lstm, D_FEA)
    cd = nb.MLP(s_cd, lstm.out_dim)
    nets = {'LSTM_layers': lstm, 'MLP_layers': cd}
    if N_MONO:
        nets['MLP_layers2'] = nb.MLP(s_mono, lstm.out_dim)
    for n in nets.values():
        n.to(device).train()
    graph = Graph(MODEL, nets, {'LSTM_layers': True, 'MLP_layers': False, 'MLP_layers2': False},
                  {'fbank': (0, D_FEA)}, {'lab_cd': D_FEA, 'lab_mono': D_FEA + 1})
    opts = {k: make_optimizer(nets[k], s) for k, s in zip(('LSTM_layers', 'MLP_layers', 'MLP_layers2'),
                                                          (s_lstm, s_cd, s_mono)) if k in nets}
    return nets, graph, opts


def train_step(graph, opts, inp, T, B, bucket=None):
    out = graph.forward(inp, T, B, 'train', ['out_dnn2'])
    for o in opts.values():
        o.zero_grad(set_to_none=True)
    out['loss_final'].backward()
    if bucket is not None:
        bucket.all_reduce_mean()
    for o in opts.values():
        o.step()
    return out['loss_final']


def algorithmic_work(name, a):
    """(bytes, flops) one call of a native entry point has to move / compute:
    every operand and result counted once at its stored width, products only inside kept 32x32 tiles."""
    try:
        if name == 'lstm_tc_fwd':        # (plan, cfg, pack, pre, drop, scalar, act, T, B, h, hb, c, gates)
            plan, pre, T, B = a[0], a[3], a[7], a[8]
            ld, H, rows = pre.shape[-1] // 4, plan.n_out, a[7] * a[8]
            byts = rows * (16 * ld + sum(n for t, n in zip(a[9:13], (4 * H, 2 * ld, 4 * ld, 16 * ld)) if t is not None))
            return byts, 2 * 4 * 16 * plan.n_blocks * rows
        if name == 'lstm_tc_bwd':        # (plan, cfg, pack, dh, gates, c, drop, scalar, act, T, B, dpre, dpre_b)
            plan, gates = a[0], a[4]
            ld, H, rows = gates.shape[-1] // 4, plan.n_out, a[9] * a[10]
            byts = rows * (4 * H + 16 * ld + 4 * ld + (16 * ld if a[11] is not None else 0) +
                           (8 * ld if len(a) > 12 and a[12] is not None else 0))
            return byts, 2 * 4 * 16 * plan.n_blocks * rows
        if name in ('tc_gemm_nt', 'tc_gemm_nn'):   # (plan, n_w, wpack, x_bf16 | dy_bf16, y | dx)
            plan, n_w, src, dst = a[0], a[1], a[3], a[4]
            m = src.shape[0]
            return m * (src.shape[1] * 2 + dst.shape[1] * 4), 2 * m * plan.nz_tiles * 1024 * n_w
        if name == 'tc_gemm_nt_split':   # (plan, n_w, wpack2, x_hi, x_lo, y): algorithmic FLOPs of ONE product (3 issued)
            plan, n_w, xh, dst = a[0], a[1], a[3], a[5]
            m = xh.shape[0]
            return m * (2 * xh.shape[1] * 2 + dst.shape[1] * 4), 2 * m * plan.nz_tiles * 1024 * n_w
        if name == 'split_bf16':         # (x2d, hi, lo)
            x, hi, lo = a[0], a[1], a[2]
            return x.shape[0] * (x.shape[1] * 4 + lo.shape[1] * 2 * (2 if hi is not None else 1)), 0
        if name == 'tc_sddmm':           # (plan, n_w, dy_bf16, x_bf16, dws, mask)
            plan, n_w, dy, x = a[0], a[1], a[2], a[3]
            m = dy.shape[0]
            return m * (dy.shape[1] + x.shape[1]) * 2, 2 * m * plan.nz_tiles * 1024 * n_w
        if name == 'out_logsoftmax_prior_fwd':   # (plan, wpack, x_bf16, bias, log_prior, out)
            plan, x, out = a[0], a[2], a[5]
            m = x.shape[0]
            return m * (x.shape[1] * 2 + out.shape[1] * 4), 2 * m * plan.nz_tiles * 1024
        if name in ('lstm_fwd', 'lstm_bwd'):
            plan = a[0]
            t = a[2] if name == 'lstm_fwd' else a[3]
            rows = t.shape[0] * t.shape[1]
            return rows * 10 * plan.n_out * 4, 2 * 4 * 16 * plan.n_blocks * rows
    except Exception:
        return None
    return None


class KernelTimers:
    """CUDA-event timing of every native entry point inside the timed region (same stream)."""

    NAMES = ['mask_apply', 'bsgemm_nt', 'bsgemm_nn', 'bs_sddmm', 'colsum', 'bn_fwd', 'bn_bwd', 'lstm_pack',
             'lstm_fwd', 'lstm_bwd', 'logsoftmax_fwd', 'logsoftmax_bwd', 'act_fwd', 'act_bwd', 'quantize_w',
             'quantize_act', 'nll_fwd_bwd', 'bn_fwd_il', 'bn_bwd_il', 'lstm_tc_pack', 'lstm_tc_fwd', 'lstm_tc_bwd',
             'tc_pack_nt', 'tc_gemm_nt', 'tc_pack_nn', 'tc_gemm_nn', 'tc_sddmm', 'cast_bf16', 'out_pack',
             'out_logsoftmax_prior_fwd', 'bn_stats_il', 'logsoftmax_bwd_fused', 'logsoftmax_bwd_fused_onehot', 'tc_pack_nt_split',
             'tc_gemm_nt_split', 'split_bf16']

    def __init__(self, names=None):
        from b200 import native
        self.native, self.orig, self.events, self.on = native, {}, [], False
        for n in (names or self.NAMES):
            self.orig[n] = getattr(native, n)
            setattr(native, n, self._wrap(n))

    def _wrap(self, name):
        fn = self.orig[name]

        def timed(*a, **k):
            if not self.on:
                return fn(*a, **k)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            self.events.append((name, e0, e1, algorithmic_work(name, a)))
            return r
        return timed

    def summary(self):
        tot, cnt, work = {}, {}, {}
        for n, e0, e1, w in self.events:
            tot[n] = tot.get(n, 0.0) + e0.elapsed_time(e1)
            cnt[n] = cnt.get(n, 0) + 1
            if w is not None:
                b, f = work.get(n, (0, 0))
                work[n] = (b + w[0], f + w[1])
        return tot, cnt, work

    def restore(self):
        for n, f in self.orig.items():
            setattr(self.native, n, f)


EXPORT_MODEL = ['out_dnn1=compute(LSTM_layers,fbank)', 'out_dnn2=compute(MLP_layers,out_dnn1)']


def synthetic_utterances(n, T, seed):
    """Evaluation set: n utterances of U[T/2, T] frames of N(0,1) features."""
    g = torch.Generator().manual_seed(seed)
    lens = torch.randint(max(1, T // 2), T + 1, (n,), generator=g).tolist()
    return [torch.randn(l, D_FEA, generator=g) for l in lens]


def run_export(dev, nets_train, args, T, B, rank, world, hbm_peak, peak_kind):
    """Forward-only posterior export (BASELINE configs[4]; reference: core.run_nn to_do='forward', core.py:183-249):
    the trained nets in eval mode, log prior fused into the output kernel, utterances batched right-padded."""
    import tempfile
    import neural_networks as nb
    import posterior_io
    from b200 import export as bexport
    from b200 import native
    from b200.graph import Graph
    s_lstm, s_cd, _ = cfg_sections('forward', args.precision)
    tmp = tempfile.mkdtemp(prefix='hcgs_export_')
    s_lstm['out_folder'] = s_cd['out_folder'] = tmp
    torch.manual_seed(2234)
    np.random.seed(2234)
    lstm = nb.LSTM(s_lstm, D_FEA)
    cd = nb.MLP(s_cd, lstm.out_dim)
    lstm.load_state_dict(nets_train['LSTM_layers'].state_dict())
    cd.load_state_dict(nets_train['MLP_layers'].state_dict())
    lstm.to(dev).eval()
    cd.to(dev).eval()
    counts = np.random.RandomState(7).rand(N_PDF).astype(np.float32) + 0.05
    cd.log_prior = posterior_io.log_prior_tensor(counts, dev)
    graph = Graph(EXPORT_MODEL, {'LSTM_layers': lstm, 'MLP_layers': cd}, {'LSTM_layers': True, 'MLP_layers': False},
                  {'fbank': (0, D_FEA)}, {})

    def fwd(x):
        return graph.forward(x, x.shape[0], x.shape[1], 'forward', ['out_dnn2'])['out_dnn2']

    n_groups = 8     # a "large synthetic eval set" (BASELINE configs[4]): 8 x B utterances per GPU, ~0.8 M frames at T = 256
    feats = synthetic_utterances(n_groups * B, T, 4242 + rank)
    lengths = [int(f.shape[0]) for f in feats]
    groups = bexport.plan_groups(lengths, B)
    resident = [bexport.pack_group(feats, g).to(dev) for g in groups]
    real = [sum(lengths[i] for i in g) for g in groups]

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
            torch.cuda.synchronize()

    steps = max(args.steps, 3)
    with torch.no_grad():
        for s in range(3):
            fwd(resident[s % len(resident)])
        names = (['out_logsoftmax_prior_fwd', 'lstm_tc_fwd', 'tc_gemm_nt', 'tc_gemm_nt_split', 'split_bf16']
                 if args.precision == 'bf16' else ['lstm_fwd'])
        timers = KernelTimers(names)
        sync_all()
        native.reset_launch_count()
        timers.on = True
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        frames = padded = 0
        for s in range(steps):
            post = fwd(resident[s % len(resident)])
            frames += real[s % len(resident)]
            padded += post.shape[0]
        e1.record()
        sync_all()
        timers.on = False
        launches = native.LAUNCHES['total']
        ms = e0.elapsed_time(e1)
        ktot, kcnt, kwork = timers.summary()
        timers.restore()
        # end to end through the exporter: host features -> pinned pack -> H2D -> forward -> D2H of the posteriors
        ex = bexport.PosteriorExporter(fwd, dev, batch=B)
        for _ in ex.run(feats):              # warm-up pass: the pinned staging buffers are allocated once
            pass
        ex.frames = ex.h2d_bytes = ex.d2h_bytes = 0
        sync_all()
        t0 = time.perf_counter()
        checksum = 0.0
        for _, m in ex.run(feats):
            checksum += float(m[0, 0])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    assert np.isfinite(checksum)
    t = torch.tensor([ms, dt * 1e3, float(frames), float(ex.frames), float(launches), float(ex.h2d_bytes),
                      float(ex.d2h_bytes)], device=dev, dtype=torch.float64)
    if world > 1:
        tmax = t.clone()
        torch.distributed.all_reduce(tmax, op=torch.distributed.ReduceOp.MAX)
        tsum = t.clone()
        torch.distributed.all_reduce(tsum, op=torch.distributed.ReduceOp.SUM)
        t = torch.stack([tmax[0], tmax[1], tsum[2], tsum[3], tsum[4], tsum[5], tsum[6]])
    ms, ms_e2e, frames, frames_e2e, launches, h2d, d2h = [v.item() for v in t]
    res = {'metric': 'hcgs_lstm_posterior_frames_per_sec', 'value': frames / (ms * 1e-3), 'unit': 'frames/s',
           'steps': steps, 'ms_per_step': ms / steps, 'gpu_launches': int(launches),
           'workload': 'forward-only export of the same 4x550 HCGS LSTM + 1944-pdf head, eval batch norm, fused '
                       'log-softmax - log-prior; %d utterances of U[%d,%d] frames per GPU, %d per right-padded group, '
                       'utterances sharded over ranks without a collective' % (len(feats), T // 2, T, B),
           'frames_counted': 'real (unpadded) frames; %.0f %% of the padded rows' % (100.0 * frames / max(padded * world, 1)),
           'e2e': {'value': frames_e2e / (ms_e2e * 1e-3), 'unit': 'frames/s', 'h2d_bytes_per_step': h2d / len(groups),
                   'd2h_bytes_per_step': d2h / len(groups), 'ms_per_step': ms_e2e / len(groups),
                   'timed': 'wall clock around PosteriorExporter.run (host packing, H2D, forward, padding rows dropped on the '
                            'device, D2H of every posterior matrix into pinned memory, double buffered)'}}
    k = 'out_logsoftmax_prior_fwd'
    if k in ktot and k in kwork:
        # the 1944-pdf head only (this export graph has no monophone head)
        avg_ms = ktot[k] / kcnt[k]
        byts = kwork[k][0] / kcnt[k]
        traffic = None
        tpath = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get('%s@T%d_B%d' % (k, T, B))
            except Exception:
                traffic = None
        res['roofline'] = {'kernel': k, 'bound': 'hbm', 'achieved': byts / (avg_ms * 1e-3) / 1e9, 'peak': hbm_peak,
                           'unit': 'GB/s', 'frac': byts / (avg_ms * 1e-3) / 1e9 / hbm_peak, 'traffic': traffic,
                           'peak_source': peak_kind, 'avg_launch_ms': avg_ms, 'share_of_step': ktot[k] / ms,
                           'algorithmic_bytes_per_launch': byts,
                           'traffic_note': 'ncu capture of the same kernel on a full 131072-row group (T*B rows); '
                                           'export groups are shorter on average',
                           'effective_tflops': kwork[k][1] / kcnt[k] / (avg_ms * 1e-3) / 1e12}
    res['kernels'] = {n: {'ms_per_step': round(ktot[n] / steps, 4), 'launches_per_step': kcnt[n] / steps}
                      for n in sorted(ktot, key=lambda x: -ktot[x])}
    return res


def _reference_modules():
    """Import the UNMODIFIED reference (baseline/_ref, else /root/reference) in THIS process: its directory goes first on
    sys.path and the plugin package is taken off, so `neural_networks`, `utils`, `hcgs` ... are the reference's own.
    Only the reference arm calls this (a process of its own: module names collide with the plugin's by design)."""
    import ref_harness as rh
    ref = rh.reference_dir()
    if ref is None:
        return None, None
    sys.path[:] = [p for p in sys.path if not p.rstrip('/').endswith('pytorch-kaldi-cgs_b200')]
    sys.path.insert(0, ref)
    import warnings
    warnings.filterwarnings('ignore')
    import neural_networks as ref_nn   # noqa: F401
    import utils as ref_utils
    assert os.path.dirname(os.path.abspath(ref_utils.__file__)) == ref and 'pytorch-kaldi-cgs_b200' not in ref_nn.__file__
    rh.cpu_shims(sys.modules)
    return ref_utils, ref


def _reference_nets(ref_utils, to_do, model):
    import configparser
    s_lstm, s_cd, s_mono = cfg_sections(to_do, 'fp32')
    cfg = configparser.ConfigParser()
    cfg['exp'] = dict(use_cuda='False', to_do=to_do, multi_gpu='False', seed='2234', out_folder='')
    extra = dict(arch_proto='', arch_library='neural_networks', arch_pretrain_file='none', arch_freeze='False',
                 arch_halving_factor='0.5', arch_improvement_threshold='0.001')
    import tempfile
    tmp = tempfile.mkdtemp(prefix='hcgs_ref_')
    for k, (sec, cls, seq) in enumerate(((s_lstm, 'LSTM', 'True'), (s_cd, 'MLP', 'False'), (s_mono, 'MLP', 'False'))):
        sec = {a: b for a, b in sec.items() if not a.startswith('b200_') and a not in ('to_do', 'use_cuda')}
        cfg['architecture%d' % (k + 1)] = dict(sec, arch_class=cls, arch_seq_model=seq, out_folder=tmp, **extra)
    fea_dict = {'fbank': ['scp', 'opts', 0, 0, 0, 0, D_FEA, D_FEA]}
    lab_dict = {'lab_cd': ['f', 'o', 'c', D_FEA], 'lab_mono': ['f', 'o', 'c', D_FEA + 1]}
    arch_dict = {'LSTM_layers': ['architecture1', 'LSTM_layers', True], 'MLP_layers': ['architecture2', 'MLP_layers', False]}
    if N_MONO and any('MLP_layers2' in line for line in model):
        arch_dict['MLP_layers2'] = ['architecture3', 'MLP_layers2', False]
    torch.manual_seed(2234)
    np.random.seed(2234)
    inp_out = dict(fea_dict)
    nns, costs = ref_utils.model_init(inp_out, model, cfg, arch_dict, False, False, to_do)
    opts = ref_utils.optimizer_init(nns, cfg, arch_dict) if to_do == 'train' else {}
    return nns, costs, opts, (fea_dict, lab_dict, arch_dict, inp_out)


def time_reference_train(ref_utils, T, B, steps, warmup, threads):
    """forward_model + backward + optimizer step of the reference's own classes (body of core.py:216-232)."""
    torch.set_num_threads(threads)
    nns, costs, opts, (fea, lab, arch, io) = _reference_nets(ref_utils, 'train', MODEL)
    times, frames = [], 0
    for s in range(warmup + steps):
        inp, fr = synthetic_minibatch(T, B, 100 + s)
        t0 = time.perf_counter()
        out = ref_utils.forward_model(fea, lab, arch, MODEL, nns, costs, inp, io, T, B, 'train', ['out_dnn2'])
        for o in opts.values():
            o.zero_grad()
        out['loss_final'].backward()
        for o in opts.values():
            o.step()
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
            frames += fr
    return frames / sum(times), sum(times) / len(times) * 1e3


def time_reference_export(ref_utils, T, n_utt, warmup, threads):
    """to_do = forward at batch size 1 (core.py:234-249): forward_model under no_grad, D2H (a no-op on the host),
    `out - log(counts / sum(counts))` in NumPy per utterance."""
    torch.set_num_threads(threads)
    nns, costs, _, (fea, lab, arch, io) = _reference_nets(ref_utils, 'forward', EXPORT_MODEL)
    counts = np.random.RandomState(7).rand(N_PDF).astype(np.float32) + 0.05
    feats = synthetic_utterances(warmup + n_utt, T, 4242)
    t_sum, frames = 0.0, 0
    for i, f in enumerate(feats):
        inp = torch.zeros(f.shape[0], 1, D_FEA + 2)
        inp[:, 0, :D_FEA] = f
        t0 = time.perf_counter()
        with torch.no_grad():
            out = ref_utils.forward_model(fea, lab, arch, EXPORT_MODEL, nns, costs, inp, io, f.shape[0], 1, 'forward',
                                          ['out_dnn2'])
        post = out['out_dnn2'].data.cpu().numpy()
        post = post - np.log(counts / np.sum(counts))
        dt = time.perf_counter() - t0
        if i >= warmup:
            t_sum += dt
            frames += f.shape[0]
    assert np.isfinite(post).all()
    return frames / t_sum, t_sum / n_utt * 1e3


REF_BATCH = 8    # the reference cfg's minibatch (cfg/TIMIT_CGS/TIMIT_LSTM_fmllr_hcgs.cfg:124 batch_size_train)


def run_reference_arm(args, rank):
    """Reference arm: the UNMODIFIED reference's own classes and forward_model on the box's host cores, all threads.
    One step = one optimizer step of the reference on a bounded sample of the workload: B = 8 sequences (the cfg's own
    minibatch size) of the same padded length T as the GPU arm (a full 528-sequence step takes about a minute of
    CPU time)."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    ref_utils, ref_dir = _reference_modules()
    base = {'impl': 'reference', 'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup}
    if ref_utils is None:
        print(json.dumps(dict(base, unavailable='reference sources not found (run baseline/install_reference.py)')))
        return
    T, B = args.seq_len, args.ref_batch
    if args.what == 'export':
        fps, ms = time_reference_export(ref_utils, T, max(args.steps, 1), args.warmup, threads)
        metric, sample = 'hcgs_lstm_posterior_frames_per_sec', (
            '%d utterances of U[%d,%d] frames, batch size 1 (core.py:183-249): forward_model under no_grad + '
            'prior normalisation in NumPy' % (max(args.steps, 1), T // 2, T))
    else:
        fps, ms = time_reference_train(ref_utils, T, B, max(args.steps, 1), args.warmup, threads)
        metric, sample = 'hcgs_lstm_train_frames_per_sec', (
            '%d train steps (forward_model, backward, RMSprop) of one [T=%d, B=%d] minibatch each: B = the reference '
            "cfg's batch_size_train; the GPU arm's step holds %s such sequences" % (max(args.steps, 1), T, B, args.batch))
    line = dict(base, metric=metric, value=fps, unit='frames/s', ms_per_step=ms, higher_is_better=True, scaling='weak',
                vs_baseline=None, dtype='f32', data='synthetic', config=workload_config(args, T, B),
                cpu_baseline={'value': fps, 'unit': 'frames/s', 'cores': threads, 'kind': 'reference',
                              'sample': sample, 'source': ref_dir},
                e2e={'value': fps, 'unit': 'frames/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0})
    print(json.dumps(line))


def reference_subprocess(what, T, B, steps, warmup, config):
    """cpu_baseline legs of the GPU arm: the reference arm in a process of its own (module names collide)."""
    cmd = [sys.executable, os.path.abspath(__file__), '--impl', 'reference', '--what', what, '--seq-len', str(T),
           '--ref-batch', str(B), '--steps', str(steps), '--warmup', str(warmup), '--config', config]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=900,
                             env=dict(os.environ, CUDA_VISIBLE_DEVICES='')).stdout
        line = json.loads([l for l in out.splitlines() if l.startswith('{')][-1])
    except Exception as e:   # noqa: BLE001
        return {'unavailable': 'reference arm failed: %r' % (e,)}
    if 'unavailable' in line:
        return {'unavailable': line['unavailable']}
    cb = line['cpu_baseline']
    cb['ms_per_step'] = line['ms_per_step']
    return cb


def workload_config(args, T, B):
    wl = ('TIMIT_CGS LSTM train step: 40-dim fbank -> 4x550 HCGS LSTM (32,4 / 75,62.5 masks, BN, tanh, '
          'drop 0.2) -> 1944-pdf (64,4 / 50,25) + 48-mono softmax heads, NLL, RMSprop')
    if CONFIG == 'libri':
        wl = ('LibriSpeech_CGS LSTM train step: 40-dim fbank -> 5x1024 HCGS LSTM (32,4 / 75,62.5 masks, BN, tanh, '
              'no dropout) -> 3480-pdf head (64,4 / 50,25), NLL, RMSprop')
    return {'workload': wl,
            'precision': getattr(args, 'precision', 'bf16') if getattr(args, 'impl', 'b200') == 'b200' else 'fp32',
            'seq_len': T, 'batch_per_gpu': B, 'frames_counted': 'real (unpadded) frames',
            'parallelism': 'dp%d' % args.gpus,
            'l2_policy': 'per-step working set (~%.1f GB of activations) exceeds the 126 MB L2; no explicit flush'
                         % (T * B * HID * 4 * 4 * 4 * 10 / 1e9)}


def measured_peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get('hbm_gbs', 6650.0), d.get('bf16_tflops_sustained', 1400.0), 'measured'
    return 6650.0, 1590.0, 'fallback'


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--seq-len', type=int, default=256, help='padded minibatch length T (cfg max_seq_length_train <= 1000)')
    ap.add_argument('--batch', type=int, default=None,
                    help='sequences per GPU per optimizer step (default 528 = 33 clusters of 16 sequences: every '
                         'co-resident cluster slot of the recurrent kernels is used; libri: 224)')
    ap.add_argument('--precision', default='bf16', choices=['bf16', 'fp32'],
                    help='bf16 = tensor-core build (BASELINE configs[1] "bf16 train"), fp32 = CUDA-core parity build')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-export', action='store_true', help='skip the posterior-export measurement')
    ap.add_argument('--no-libri', action='store_true', help='skip the LibriSpeech-shape (configs[3]) object of the line')
    ap.add_argument('--no-chunk', action='store_true', help='skip the chunk-pipeline (device minibatch assembly) object')
    ap.add_argument('--no-cfg-batch', action='store_true', help='skip the T = 1000, B = 8 (reference cfg minibatch) line')
    ap.add_argument('--what', default='train', choices=['train', 'export'], help='reference arm: which leg to time')
    ap.add_argument('--ref-batch', type=int, default=REF_BATCH, help='reference arm: sequences per step (cfg batch size)')
    ap.add_argument('--config', default='timit', choices=['timit', 'libri'],
                    help='timit = BASELINE configs[1] (headline); libri = configs[3], 5x1024 / 3480 pdfs (no CPU leg)')
    args = ap.parse_args()
    select_config(args.config)
    if args.config != 'timit':
        args.no_cpu_baseline = True
    if args.batch is None:
        # whole waves of co-resident clusters of the recurrent kernels: 33 clusters of 4 CTAs at H = 550 (16 sequences
        # each), 7 clusters of 16 CTAs at H = 1024 -> 2 waves = 224 sequences
        args.batch = 528 if args.config == 'timit' else 224

    from b200 import dist as bdist
    if args.impl == 'reference':
        rank = int(os.environ.get('RANK', '0'))
        run_reference_arm(args, rank)
        return
    rank, world, local = bdist.init_from_env('nccl')
    assert torch.cuda.is_available(), 'bench.py needs a CUDA device (no CPU fallback)'
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    from b200 import native
    T, B = args.seq_len, args.batch
    W = max(args.warmup, 3)
    nets, graph, opts = build_workload(dev, precision=args.precision)
    bucket = None
    if world > 1:
        # one compressed-gradient bucket per LSTM layer / head; each all-reduce starts from an autograd hook as soon as
        # its gradients exist and overlaps the backward of the layers below
        bucket = bdist.OverlappedGradBuckets(bdist.lstm_mlp_groups(list(nets.values())), dev)
    n_batches = 4
    host = []
    frames_per = []
    for i in range(n_batches):
        inp, fr = synthetic_minibatch(T, B, 1000 * (rank + 1) + i)
        host.append(inp.pin_memory())
        frames_per.append(fr)
    resident = [h.to(dev) for h in host]

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
            torch.cuda.synchronize()

    # ---- device-resident throughput ("value") ---------------------------------------------------
    # Inside the timed region only the recurrent kernels (the dominant kernel is one of them) are bracketed by
    # CUDA events -- 8 launches per step; the full per-kernel table comes from a second, untimed pass.
    for s in range(W):
        train_step(graph, opts, resident[s % n_batches], T, B, bucket)
    # untimed pass with every native entry point bracketed: the per-kernel table of the report.  It runs BEFORE the
    # timed region (it also lets the caching allocator reach its steady state: with 3 warm-up steps alone the first
    # timed steps still paid cudaMalloc calls, 32.6 vs 30.4 ms per step)
    full = KernelTimers()
    full.on = True
    full_steps = min(max(args.steps, 8), 10)   # at least 8: with fewer the first timed steps still see allocator growth
    for s in range(full_steps):
        train_step(graph, opts, resident[s % n_batches], T, B, bucket)
    sync_all()
    full.on = False
    ftot, fcnt, fwork = full.summary()
    full.restore()
    rec = ['lstm_tc_fwd', 'lstm_tc_bwd'] if args.precision == 'bf16' else ['lstm_fwd', 'lstm_bwd']
    timers = KernelTimers(rec)
    sampler = ClockSampler(local)
    # no cyclic-GC pass inside the timed regions: a generation-2 collection stalls the launching thread for tens of
    # milliseconds (seen once as a 33 ms step among 30 ms ones)
    import gc
    gc.collect()
    gc.disable()
    sync_all()
    sampler.start()
    native.reset_launch_count()
    timers.on = True
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    frames = 0
    for s in range(args.steps):
        train_step(graph, opts, resident[s % n_batches], T, B, bucket)
        frames += frames_per[s % n_batches]
    e1.record()
    sync_all()
    timers.on = False
    clocks = sampler.stop()
    launches = native.LAUNCHES['total']
    ms_total = e0.elapsed_time(e1)
    ktot, kcnt, kwork = timers.summary()
    timers.restore()

    # ---- end to end through the public API: pinned host minibatch -> GPU -> loss on the host ------
    # Every step copies its minibatch from pinned host memory and its loss back to the host inside the timed
    # region.  Like a production loop the copies are pipelined: the next minibatch is uploaded on a side stream
    # while the current step runs, and the loss of step i is read (pinned buffer + event) after step i+3 has been
    # queued, so the host never drains the GPU and a late host thread (8 ranks share the box's cores) does not stall it.
    copy_stream = torch.cuda.Stream(device=dev)
    main_stream = torch.cuda.current_stream()
    stage_in = [torch.empty_like(resident[0]) for _ in range(2)]
    LAG = 4   # the loss of step i is read once step i + LAG - 1 has been queued (a ring of LAG pinned scalars)
    loss_host = [torch.zeros(1, dtype=torch.float32).pin_memory() for _ in range(LAG)]
    loss_ready = [torch.cuda.Event() for _ in range(LAG)]
    in_ready = [torch.cuda.Event() for _ in range(2)]
    in_free = [torch.cuda.Event() for _ in range(2)]

    def upload(i):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(in_free[i % 2])
            stage_in[i % 2].copy_(host[i % n_batches], non_blocking=True)
            in_ready[i % 2].record(copy_stream)

    def e2e_pass(n_steps):
        """n_steps through the pipelined loop; returns (losses read on the host, frames)."""
        got, frames_n = [], 0
        upload(0)
        for s in range(n_steps):
            if s + 1 < n_steps:
                upload(s + 1)
            main_stream.wait_event(in_ready[s % 2])
            loss = train_step(graph, opts, stage_in[s % 2], T, B, bucket)
            in_free[s % 2].record(main_stream)
            loss_host[s % LAG].copy_(loss.detach().reshape(1), non_blocking=True)    # D2H read of the step's result
            loss_ready[s % LAG].record(main_stream)
            if s >= LAG - 1:
                r = s - (LAG - 1)
                loss_ready[r % LAG].synchronize()
                got.append(float(loss_host[r % LAG][0]))
            frames_n += frames_per[s % n_batches]
        for r in range(max(0, n_steps - (LAG - 1)), n_steps):
            loss_ready[r % LAG].synchronize()
            got.append(float(loss_host[r % LAG][0]))
        return got, frames_n

    for ev in in_free:
        ev.record(main_stream)
    sync_all()
    # two untimed steps through the same loop: the first use of the side stream, the staging buffers and the pinned loss
    # ring costs a one-off ~10 ms on some runs (seen as one 23.4 ms average among 22.3 ms ones)
    e2e_pass(2)
    for ev in in_free:
        ev.record(main_stream)
    sync_all()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    losses, frames_e2e = e2e_pass(args.steps)
    e3.record()
    sync_all()
    ms_e2e = e2.elapsed_time(e3)
    gc.enable()
    assert len(losses) == args.steps and all(np.isfinite(losses)), losses

    # ---- the cfg-faithful minibatch: T = 1000, B = 8 sequences per optimizer step, same nets -----------
    cfg_batch = None
    if world == 1 and not args.no_cfg_batch and args.config == 'timit':
        Tc, Bc = 1000, REF_BATCH
        mbs = [synthetic_minibatch(Tc, Bc, 777 + i) for i in range(2)]
        res_c = [m[0].to(dev) for m in mbs]
        for s in range(3):
            train_step(graph, opts, res_c[s % 2], Tc, Bc, None)
        torch.cuda.synchronize()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        fr_c = 0
        for s in range(args.steps):
            train_step(graph, opts, res_c[s % 2], Tc, Bc, None)
            fr_c += mbs[s % 2][1]
        c1.record()
        torch.cuda.synchronize()
        ms_c = c0.elapsed_time(c1)
        cfg_batch = {'seq_len': Tc, 'batch_per_gpu': Bc, 'value': fr_c / (ms_c * 1e-3), 'unit': 'frames/s',
                     'ms_per_step': ms_c / args.steps, 'steps': args.steps,
                     'note': "the reference cfg's own minibatch (batch_size_train = 8, max_seq_length_train = 1000): "
                             'one cluster group of sequences, the recurrence is latency bound'}

    chunk_leg = None
    if world == 1 and not args.no_chunk and args.config == 'timit':
        chunk_leg = run_chunk_leg(graph, opts, dev, args.steps, T, B)

    hbm_peak, tf_peak, peak_kind = measured_peaks()
    export = None
    if not args.no_export:
        export = run_export(dev, nets, args, T, B, rank, world, hbm_peak, peak_kind)

    libri = None
    if args.config == 'timit' and not args.no_libri and args.precision == 'bf16':
        del resident, stage_in
        torch.cuda.empty_cache()
        libri = run_libri_leg(dev, args, rank, world,
                              lambda n: bdist.OverlappedGradBuckets(bdist.lstm_mlp_groups(list(n.values())), dev))
    lib_ms, lib_frames = (libri[0], float(libri[1])) if libri else (0.0, 0.0)
    t = torch.tensor([ms_total, ms_e2e, float(frames), float(frames_e2e), float(launches), lib_ms, lib_frames],
                     device=dev, dtype=torch.float64)
    if world > 1:
        tmax = t.clone()
        torch.distributed.all_reduce(tmax, op=torch.distributed.ReduceOp.MAX)
        tsum = t.clone()
        torch.distributed.all_reduce(tsum, op=torch.distributed.ReduceOp.SUM)
        ms_total, ms_e2e = tmax[0].item(), tmax[1].item()
        frames, frames_e2e, launches = tsum[2].item(), tsum[3].item(), int(tsum[4].item())
        lib_ms, lib_frames = tmax[5].item(), tsum[6].item()
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
    if rank != 0:
        return

    value = frames / (ms_total * 1e-3)
    e2e = frames_e2e / (ms_e2e * 1e-3)
    # dominant kernel = largest share of the timed region among our kernels
    dom = max(ktot, key=ktot.get)
    avg_ms = ktot[dom] / kcnt[dom]
    roof = {'kernel': dom, 'bound': 'hbm', 'achieved': None, 'peak': hbm_peak, 'unit': 'GB/s', 'frac': None,
            'traffic': None, 'peak_source': peak_kind, 'avg_launch_ms': avg_ms, 'launches_per_step': kcnt[dom] / args.steps,
            'share_of_step': ktot[dom] / ms_total}
    if dom in kwork:
        byts, flops = kwork[dom][0] / kcnt[dom], kwork[dom][1] / kcnt[dom]
        roof['algorithmic_bytes_per_launch'] = byts
        roof['algorithmic_flops_per_launch'] = flops
        if dom in ('tc_gemm_nt', 'tc_gemm_nt_split', 'tc_gemm_nn', 'tc_sddmm'):
            ach = flops / (avg_ms * 1e-3) / 1e12
            roof.update(bound='tensor', achieved=ach, peak=tf_peak, unit='TFLOP/s', frac=ach / tf_peak)
        else:
            ach = byts / (avg_ms * 1e-3) / 1e9
            roof.update(achieved=ach, frac=ach / hbm_peak)
            if dom.startswith('lstm'):
                roof['us_per_time_step'] = avg_ms * 1e3 / T
                roof['effective_tflops'] = flops / (avg_ms * 1e-3) / 1e12
                roof['note'] = ('persistent recurrent kernel: streams its operands once; the floor is the serial '
                                'chain of T time steps (cluster exchange of h_t per step), not HBM -- DESIGN.md 3.3')
    traffic = None
    tpath = os.path.join(ROOT, 'profiles', 'ncu_traffic.json')   # dram bytes per launch from `ncu --set full`
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get('%s@T%d_B%d' % (dom, T, B))
        except Exception:
            traffic = None
    roof['traffic'] = float(traffic) if isinstance(traffic, (int, float)) else None
    per_kernel = {}
    for k in sorted(ftot, key=lambda x: -ftot[x]):
        e = {'ms_per_step': round(ftot[k] / full_steps, 4), 'launches_per_step': fcnt[k] / full_steps}
        if k in fwork:
            e['GB_per_s'] = round(fwork[k][0] / (ftot[k] * 1e-3) / 1e9, 1)
            e['TFLOP_per_s'] = round(fwork[k][1] / (ftot[k] * 1e-3) / 1e12, 2)
            # fractions of the measured peaks (MEASURED_PEAKS.json: HBM copy bandwidth, sustained dense bf16): algorithmic
            # bytes / FLOPs of the entry point over its event-timed duration inside a step
            e['hbm_frac'] = round(e['GB_per_s'] / hbm_peak, 3)
            if e['TFLOP_per_s'] > 0:
                e['tensor_frac'] = round(e['TFLOP_per_s'] / tf_peak, 3)
        per_kernel[k] = e
    line = {'metric': 'hcgs_lstm_train_frames_per_sec', 'value': value, 'unit': 'frames/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': W, 'ms_per_step': ms_total / args.steps, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16' if args.precision == 'bf16' else 'f32',
            'data': 'synthetic', 'config': workload_config(args, T, B), 'clocks': clocks, 'gpu_launches': launches,
            'e2e': {'value': e2e, 'unit': 'frames/s', 'h2d_bytes_per_step': host[0].numel() * 4 * world,
                    'd2h_bytes_per_step': 4 * world, 'ms_per_step': ms_e2e / args.steps,
                    'pipelining': 'next minibatch uploaded on a side stream; loss of step i read after step i+3 is queued (every loss is read inside the timed region)'},
            'roofline': roof,
            'kernels_untimed_pass': per_kernel}
    if export is not None:
        line['export'] = export
        # top-level copies so that per-N export numbers survive in the driver's SCALE records
        line['export_value'] = export['value']
        line['export_e2e_value'] = export['e2e']['value']
    if libri is not None:
        line['libri'] = {'workload': 'LibriSpeech_CGS LSTM train step (BASELINE configs[3]): 40-dim fbank -> 5x1024 HCGS LSTM '
                                     '(32,4 / 75,62.5 masks, BN, tanh, no dropout) -> 3480-pdf (64,4 / 50,25) softmax head, '
                                     'NLL, RMSprop', 'value': lib_frames / (lib_ms * 1e-3), 'unit': 'frames/s',
                         'ms_per_step': lib_ms / args.steps, 'steps': args.steps, 'seq_len': libri[2],
                         'batch_per_gpu': libri[3], 'n_gpus': world, 'data': 'synthetic',
                         'note': 'H = 1024 needs clusters of 16 CTAs for the recurrent kernels (7 co-resident = 112 '
                                 'sequences per wave, two waves per step)'}
        line['libri_value'] = line['libri']['value']   # top-level copy for the driver's SCALE records
    if chunk_leg is not None:
        line['chunk'] = chunk_leg
    if cfg_batch is not None:
        line['cfg_batch'] = cfg_batch
        if not args.no_cpu_baseline:
            cfg_batch['cpu_baseline'] = reference_subprocess('train', cfg_batch['seq_len'], REF_BATCH, 2, 1, args.config)
    if not args.no_cpu_baseline and world == 1:
        # the reference's own classes on the host cores (baseline/_ref), bounded samples of the same workloads
        line['cpu_baseline'] = reference_subprocess('train', T, REF_BATCH, 3, 1, args.config)
        if export is not None:
            export['cpu_baseline'] = reference_subprocess('export', T, 1, 12, 2, args.config)
    print(json.dumps(line))


def run_libri_leg(dev, args, rank, world, bucket_factory):
    """BASELINE configs[3] as a second object of the default line: 5x1024 HCGS LSTM, 3480 pdfs, B = 224 sequences per GPU
    (two waves of the 7 co-resident clusters of 16 CTAs), same T, device-resident minibatches, the same train step.
    Every rank runs it (data parallel like the headline); returns (ms_total, frames) of this rank."""
    select_config('libri')
    try:
        T, B = args.seq_len, 224
        nets, graph, opts = build_workload(dev, precision=args.precision)
        bucket = bucket_factory(nets) if world > 1 else None
        mbs = [synthetic_minibatch(T, B, 5000 * (rank + 1) + i) for i in range(2)]
        res = [m[0].to(dev) for m in mbs]
        for s in range(6):   # warm-up + allocator steady state
            train_step(graph, opts, res[s % 2], T, B, bucket)
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
            torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        frames = 0
        for s in range(args.steps):
            train_step(graph, opts, res[s % 2], T, B, bucket)
            frames += mbs[s % 2][1]
        e1.record()
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        del nets, graph, opts, res, bucket
        torch.cuda.empty_cache()
        return ms, frames, T, B
    finally:
        select_config('timit')


if __name__ == '__main__':
    main()
