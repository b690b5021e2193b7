"""Builder-run measurements of the two BASELINE.json configurations that are not bench.py lines:

  configs[0]  TIMIT_MLP_mfcc_basic: 39-dim MFCC, +-5 frames of context (429 in), 4x1024 ReLU MLP with batch norm and
              dropout 0.15, 1944-pdf softmax, SGD, minibatch 128 frames (cfg/TIMIT_baselines/TIMIT_MLP_mfcc_basic.cfg:78-104)
  configs[2]  cfg/20200910_QuantExp/TIMIT_LSTM_fmllr_L2_8bw_16ba_whcgs.cfg: 2x512 HCGS LSTM, 8-bit weights, 16-bit
              Quantize_inp of layer inputs and of h_t at every step, prune 70, 1944 + 48 quantised HCGS heads, RMSprop,
              minibatch = 8 sequences of <= 1000 frames (fp32 build: the per-step whole-tensor scale needs one cluster)

    python tools/bench_configs.py [mlp|quant|both] [steps]      -> one JSON line per configuration
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from b200.graph import Graph, make_optimizer

dev = torch.device('cuda', 0)
torch.cuda.set_device(0)
which = sys.argv[1] if len(sys.argv) > 1 else 'both'
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 20


def timed_steps(fn, n, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def run_mlp(precision):
    import neural_networks as nb
    torch.manual_seed(1234); np.random.seed(1234)
    D, P, N = 429, 1944, 128
    sec = dict(arch_name='MLP_layers1', dnn_lay='1024,1024,1024,1024,%d' % P, dnn_drop='0.15,0.15,0.15,0.15,0.0',
               dnn_use_laynorm_inp='False', dnn_use_batchnorm_inp='False', dnn_use_batchnorm='True,True,True,True,False',
               dnn_use_laynorm='False,False,False,False,False', dnn_act='relu,relu,relu,relu,softmax', arch_lr='0.08',
               arch_opt='sgd', opt_momentum='0.0', opt_weight_decay='0.0', opt_dampening='0.0', opt_nesterov='False',
               out_folder='', mlp_hcgs='False', guided_hcgs='False', apply_guided_hcgs='False', hcgs_block='64,4',
               hcgs_sparse='50,25', mlp_quant='False', param_quant='8,8,8,8,8', mlp_quant_inp='False', inp_quant='16',
               mlp_prune='False', mlp_prune_perc='70', skip_regularization='True', to_do='train', use_cuda='True',
               b200_precision=precision)
    net = nb.MLP(sec, D).to(dev).train()
    graph = Graph(['out_dnn1=compute(MLP_layers1,mfcc)', 'loss_final=cost_nll(out_dnn1,lab_cd)',
                   'err_final=cost_err(out_dnn1,lab_cd)'], {'MLP_layers1': net}, {'MLP_layers1': False},
                  {'mfcc': (0, D)}, {'lab_cd': D})
    opt = make_optimizer(net, sec)
    g = torch.Generator().manual_seed(7)
    out = {}
    for nb_ in (N, 128 * 256):   # the cfg minibatch, and a minibatch that fills the GPU
        inp = torch.cat([torch.randn(nb_, D, generator=g), torch.randint(0, P, (nb_, 1), generator=g).float()], 1).to(dev)

        def step():
            o = graph.forward(inp, nb_, 1, 'train', ['out_dnn1'])
            opt.zero_grad(set_to_none=True)
            o['loss_final'].backward()
            opt.step()
        ms = timed_steps(step, steps)
        out['minibatch_%d' % nb_] = {'frames_per_s': nb_ / (ms * 1e-3), 'ms_per_step': ms}
    return {'config': 'BASELINE configs[0]: TIMIT_MLP_mfcc_basic, 429 -> 4x1024 ReLU (BN, drop 0.15) -> 1944 softmax, SGD',
            'precision': precision, 'metric': 'train frames/s (synthetic, device-resident)', **out}


def run_quant():
    import neural_networks as nb
    torch.manual_seed(2234); np.random.seed(2234)
    bench.select_config('timit')
    bench.LAYERS, bench.HID, bench.DROP = 2, 512, '0.0'
    try:
        s_lstm, s_cd, s_mono = bench.cfg_sections('train', 'fp32')
    finally:
        bench.select_config('timit')
    s_lstm.update(lstm_quant='True', param_quant='8,8', lstm_quant_inp='True', inp_quant='16', lstm_prune='True',
                  lstm_prune_perc='70,70')  # the shipped cfg has a single 70 and the reference indexes it per layer (:888)
    for s in (s_cd, s_mono):
        s.update(mlp_quant='True', param_quant='8', mlp_quant_inp='True', inp_quant='16', mlp_prune='True',
                 mlp_prune_perc='70', mlp_hcgs='True')
    lstm = nb.LSTM(s_lstm, bench.D_FEA)
    nets = {'LSTM_layers': lstm, 'MLP_layers': nb.MLP(s_cd, lstm.out_dim), 'MLP_layers2': nb.MLP(s_mono, lstm.out_dim)}
    for n in nets.values():
        n.to(dev).train()
    graph = Graph(bench.MODEL_TIMIT, nets, {'LSTM_layers': True, 'MLP_layers': False, 'MLP_layers2': False},
                  {'fbank': (0, bench.D_FEA)}, {'lab_cd': bench.D_FEA, 'lab_mono': bench.D_FEA + 1})
    opts = {k: make_optimizer(nets[k], s) for k, s in zip(('LSTM_layers', 'MLP_layers', 'MLP_layers2'), (s_lstm, s_cd, s_mono))}
    out = {'config': 'BASELINE configs[2]: cfg/20200910_QuantExp (2x512 HCGS LSTM, 8-bit weights, 16-bit Quantize_inp '
                     'of layer inputs and of h_t per step, prune 70, quantised HCGS heads)',
           'precision': 'fp32 build (the per-step whole-batch activation scale is exchanged between the clusters of the '
                        'fp32 recurrent kernel; every cluster must be co-resident)',
           'metric': 'train frames/s (synthetic, device-resident, real frames)'}
    from b200 import native
    for T, B in ((1000, 8), (256, 64), (256, 144), (256, 288)):
        mbs = [bench.synthetic_minibatch(T, B, 40 + i) for i in range(2)]
        res = [m[0].to(dev) for m in mbs]
        i = [0]

        def step():
            bench.train_step(graph, opts, res[i[0] % 2], T, B, None)
            i[0] += 1
        try:
            ms = timed_steps(step, steps, warm=4)
        except native.NativeError as e:
            out['T%d_B%d' % (T, B)] = {'refused': str(e)[-160:]}
            continue
        fr = sum(mbs[k % 2][1] for k in range(steps)) / steps
        out['T%d_B%d' % (T, B)] = {'frames_per_s': fr / (ms * 1e-3), 'ms_per_step': ms}
    return out


if which in ('mlp', 'both'):
    for prec in ('bf16', 'fp32'):
        print(json.dumps(run_mlp(prec)), flush=True)
if which in ('quant', 'both'):
    print(json.dumps(run_quant()), flush=True)
