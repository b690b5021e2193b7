"""Generate the golden fixtures in this directory by RUNNING THE REFERENCE ITSELF.

Run once in the build container (the only place /root/reference exists):

    python tests/golden/make_golden.py

It imports the unmodified reference modules from /root/reference (read-only), applies the four
harness-side shims listed in SURVEY.md 8(c) (no-op the forced ``.to("cuda")`` of mask creation on a
GPU-less host; nothing else is touched), drives them through the reference's own
``utils.model_init`` / ``optimizer_init`` / ``forward_model`` and ``core``-style posterior
normalisation + ``data_io.write_mat``, and stores small ``.npz`` fixtures.  Nothing under tests/,
bench.py or __graft_entry__.py reads /root/reference at run time; they read these fixtures.
"""
import configparser
import io
import os
import sys
import tempfile
import types
import warnings

import numpy as np
import torch

warnings.filterwarnings('ignore')
REF = '/root/reference'
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REF)

import hcgs  # noqa: E402  (reference)


class _TorchProxy(types.ModuleType):
    """`torch` as seen by the reference's hcgs.py: device("cuda") resolves to the CPU here."""

    def __getattr__(self, k):
        return getattr(torch, k)

    def device(self, *_a, **_k):
        return torch.device('cpu')


hcgs.torch = _TorchProxy('torch')
import guided_hcgs  # noqa: E402
guided_hcgs.torch = _TorchProxy('torch')
import neural_networks as ref_nn  # noqa: E402
import utils as ref_utils  # noqa: E402
import quantized_modules as ref_q  # noqa: E402
import data_io as ref_io  # noqa: E402

MASK_CASES = [
    (550, 550, [32, 4], [75, 62.5], 2234), (550, 40, [32, 4], [75, 62.5], 2234),
    (1944, 550, [64, 4], [50, 25], 2234), (48, 550, [64, 4], [50, 25], 2234),
    (40, 64, [64, 4], [50, 25], 3), (1024, 1024, [128, 4], [25, 62.5], 7),
    (512, 440, [32, 2], [75, 75], 2234), (3480, 1024, [64, 4], [50, 25], 1234),
    (100, 70, [16], [50], 9), (33, 65, [8, 2], [50, 50], 11), (7, 5, [4, 2, 1], [50, 50, 0], 4),
    (512, 512, [32, 4], [75, 62.5], 2234), (1024, 1024, [32, 4], [75, 62.5], 1234),
]


def gen_masks():
    d = tempfile.mkdtemp()
    out = {'cases': np.array([repr(c) for c in MASK_CASES])}
    for k, (r, c, bs, dr, seed) in enumerate(MASK_CASES):
        np.random.seed(seed)
        m = hcgs.conn_mat(r, c, list(bs), list(dr), dir=d, for_test=True)
        out['bits_%d' % k] = np.packbits(m.astype(np.uint8))
        out['shape_%d' % k] = np.array(m.shape)
        out['next_rand_%d' % k] = np.array([np.random.randint(0, 2 ** 31 - 1)])  # RNG stream position
    # guided masks: deterministic given W
    torch.manual_seed(5)
    w = torch.randn(48, 40)
    g = guided_hcgs.conn_mat(48, 40, [8, 4], [50, 50], w, dir=d, for_test=True)
    out['guided_w'] = w.numpy()
    out['guided_bits'] = np.packbits(g.astype(np.uint8))
    np.savez_compressed(os.path.join(HERE, 'masks.npz'), **out)
    print('masks.npz', len(MASK_CASES), 'cases')


LSTM_OPTS = dict(
    arch_name='LSTM_layers', arch_proto='proto/LSTM.proto', arch_library='neural_networks', arch_class='LSTM',
    arch_pretrain_file='none', arch_freeze='False', arch_seq_model='True', lstm_lay='24,24', lstm_drop='0.2,0.2',
    lstm_use_laynorm_inp='False', lstm_use_batchnorm_inp='False', lstm_use_laynorm='False,False',
    lstm_use_batchnorm='True,True', lstm_bidir='False', lstm_act='tanh,tanh', if_hsigmoid='True',
    lstm_orthinit='True', arch_lr='0.0016', arch_halving_factor='0.5', arch_improvement_threshold='0.001',
    arch_opt='rmsprop', opt_momentum='0.0', opt_alpha='0.95', opt_eps='1e-8', opt_centered='False',
    opt_weight_decay='0.0', out_folder='', lstm_hcgs='True', guided_hcgs='False', apply_guided_hcgs='False',
    hcgsx_block='8,4', hcgsx_sparse='50,50', hcgsh_block='8,4', hcgsh_sparse='50,50', lstm_quant='False',
    param_quant='8,8', lstm_quant_inp='False', inp_quant='16', lstm_prune='False', lstm_prune_perc='70',
    skip_regularization='True')

MLP_OPTS = dict(
    arch_name='MLP_layers', arch_proto='proto/MLP.proto', arch_library='neural_networks', arch_class='MLP',
    arch_pretrain_file='none', arch_freeze='False', arch_seq_model='False', dnn_lay='10', dnn_drop='0.0',
    dnn_use_laynorm_inp='False', dnn_use_batchnorm_inp='False', dnn_use_batchnorm='False', dnn_use_laynorm='False',
    dnn_act='softmax', arch_lr='0.0016', arch_halving_factor='0.5', arch_improvement_threshold='0.001',
    arch_opt='rmsprop', opt_momentum='0.0', opt_alpha='0.95', opt_eps='1e-8', opt_centered='False',
    opt_weight_decay='0.0', out_folder='', mlp_hcgs='True', guided_hcgs='False', apply_guided_hcgs='False',
    hcgs_block='8,4', hcgs_sparse='50,25', mlp_quant='False', param_quant='8', mlp_quant_inp='False',
    inp_quant='16', mlp_prune='False', mlp_prune_perc='70', skip_regularization='True')

MODEL_2HEAD = ['out_dnn1=compute(LSTM_layers,fbank)', 'out_dnn2=compute(MLP_layers,out_dnn1)',
               'out_dnn3=compute(MLP_layers2,out_dnn1)', 'loss_mono=cost_nll(out_dnn3,lab_mono)',
               'loss_mono_w=mult_constant(loss_mono,1.0)', 'loss_cd=cost_nll(out_dnn2,lab_cd)',
               'loss_final=sum(loss_cd,loss_mono_w)', 'err_final=cost_err(out_dnn2,lab_cd)']


def build_config(to_do, lstm_over=None, mlp_over=None, out_folder='', mono_over=None, arch1=None):
    cfg = configparser.ConfigParser()
    cfg['exp'] = dict(use_cuda='False', to_do=to_do, multi_gpu='False', seed='2234', out_folder=out_folder)
    a1 = dict(arch1 if arch1 is not None else LSTM_OPTS, out_folder=out_folder)
    a1.update(lstm_over or {})
    a2 = dict(MLP_OPTS, out_folder=out_folder)
    a2.update(mlp_over or {})
    a3 = dict(a2, arch_name='MLP_layers2', dnn_lay='5', mlp_hcgs='False', arch_lr='0.0004')
    a3.update(mono_over or {})
    cfg['architecture1'], cfg['architecture2'], cfg['architecture3'] = a1, a2, a3
    return cfg


def run_model(tag, to_do, T=7, B=3, D=12, lstm_over=None, mlp_over=None, steps=1, seed=2234, mono_over=None,
              extra=None, arch1=None):
    """One (or a few) reference train steps / one eval forward; returns a dict of arrays."""
    tmp = tempfile.mkdtemp()
    cfg = build_config(to_do, lstm_over, mlp_over, out_folder=tmp, mono_over=mono_over, arch1=arch1)
    torch.manual_seed(seed)
    np.random.seed(seed)
    fea_dict = {'fbank': ['scp', 'opts', 0, 0, 0, 0, D, D]}
    lab_dict = {'lab_cd': ['f', 'o', 'c', D], 'lab_mono': ['f', 'o', 'c', D + 1]}
    arch_dict = {'LSTM_layers': ['architecture1', 'LSTM_layers', True], 'MLP_layers': ['architecture2', 'MLP_layers', False],
                 'MLP_layers2': ['architecture3', 'MLP_layers2', False]}
    inp_out_dict = dict(fea_dict)
    nns, costs = ref_utils.model_init(inp_out_dict, MODEL_2HEAD, cfg, arch_dict, False, False, to_do)
    opts = ref_utils.optimizer_init(nns, cfg, arch_dict)
    out = {}
    for net in nns:
        for k, v in nns[net].state_dict().items():
            out['init/%s/%s' % (net, k)] = v.detach().numpy().copy()
    gen = torch.Generator().manual_seed(seed + 1)
    drops = []
    real_bernoulli = torch.bernoulli

    def rec_bernoulli(p, *a, **k):
        r = real_bernoulli(p, *a, **k)
        drops.append(r.clone())
        return r
    torch.bernoulli = rec_bernoulli
    try:
        for s in range(steps):
            inp = torch.zeros(T, B, D + 2)
            inp[:, :, :D] = torch.randn(T, B, D, generator=gen)
            inp[:, :, D] = torch.randint(0, 10, (T, B), generator=gen).float()
            inp[:, :, D + 1] = torch.randint(0, 5, (T, B), generator=gen).float()
            inp[T - 2:, B - 1, :] = 0   # right padding of the shortest sequence, as core.py builds it
            out['inp/%d' % s] = inp.numpy().copy()
            if to_do == 'train':
                o = ref_utils.forward_model(fea_dict, lab_dict, arch_dict, MODEL_2HEAD, nns, costs, inp, inp_out_dict,
                                            T, B, to_do, ['out_dnn2'])
                for opt in opts.values():
                    opt.zero_grad()
                o['loss_final'].backward()
                if s == steps - 1:
                    for net in nns:
                        for k, p in nns[net].named_parameters():
                            if p.grad is not None:
                                out['grad/%s/%s' % (net, k)] = p.grad.numpy().copy()
                for opt in opts.values():
                    opt.step()
                out['loss/%d' % s] = np.array([o['loss_final'].item(), o['err_final'].item(), o['loss_cd'].item(),
                                               o['loss_mono'].item()], dtype=np.float64)
                out['out_dnn1/%d' % s] = o['out_dnn1'].detach().numpy().copy()
                out['out_dnn2/%d' % s] = o['out_dnn2'].detach().numpy().copy()
            else:
                with torch.no_grad():
                    o = ref_utils.forward_model(fea_dict, lab_dict, arch_dict, MODEL_2HEAD, nns, costs, inp,
                                                inp_out_dict, T, B, to_do, ['out_dnn2'])
                out['out_dnn1/%d' % s] = o['out_dnn1'].detach().numpy().copy()
                out['out_dnn2/%d' % s] = o['out_dnn2'].detach().numpy().copy()
                if to_do == 'valid':
                    out['loss/%d' % s] = np.array([o['loss_final'].item(), o['err_final'].item()], dtype=np.float64)
    finally:
        torch.bernoulli = real_bernoulli
    for i, d in enumerate(drops):
        out['drop/%d' % i] = d.numpy().copy()
    for net in nns:
        for k, v in nns[net].state_dict().items():
            out['final/%s/%s' % (net, k)] = v.detach().numpy().copy()
    out['dims'] = np.array([T, B, D, steps])
    if extra is not None:
        extra(out, nns)
    np.savez_compressed(os.path.join(HERE, tag + '.npz'), **out)
    print(tag + '.npz', 'loss', out.get('loss/0'))
    return out, nns


def gen_posterior():
    """Posterior export: eval forward of one utterance (batch 1), prior normalisation in NumPy
    (core.py:241-249) and the Kaldi ark record bytes (data_io.write_mat)."""
    res, nns = run_model('eval_forward', 'forward', T=9, B=1, D=12, seed=77)
    rng = np.random.RandomState(3)
    counts = (rng.rand(10) * 1000 + 1).astype(np.float32)
    with tempfile.NamedTemporaryFile('w', suffix='.count', delete=False) as f:
        f.write('[ ' + ' '.join('%g' % v for v in counts) + ' ]\n')
        cpath = f.name
    loaded = ref_io.load_counts(cpath)
    post = res['out_dnn2/0'] - np.log(loaded / np.sum(loaded))
    ark = tempfile.NamedTemporaryFile(suffix='.ark', delete=False).name
    with open(ark, 'wb') as fd:
        ref_io.write_mat('.', fd, post, 'utt_0001')
    blob = np.frombuffer(open(ark, 'rb').read(), dtype=np.uint8)
    np.savez_compressed(os.path.join(HERE, 'posterior.npz'), counts_text=np.array(open(cpath).read()),
                        counts=loaded, post=post, ark=blob, logp=res['out_dnn2/0'])
    print('posterior.npz', post.shape, post.dtype, len(blob), 'bytes')


def gen_quant():
    torch.manual_seed(9)
    w = torch.randn(37, 23) * 0.8
    x = torch.randn(11, 23) * 3
    master = w.clone()
    q = ref_q.Quantize(master, numBits=8, balanced=False)
    xq = x.clone()
    for _ in range(4):
        xq = ref_q.Quantize_inp(xq, 16)
    x1 = ref_q.Quantize_inp(x.clone(), 16)
    wf = w.clone()
    qf = ref_q.Quantize(wf, numBits=8, if_forward=True)          # balanced=True, in place
    lin = torch.nn.Linear(23, 37)
    lin.weight.data = w.clone()
    pm = ref_q.prune(lin, 70)[0]
    np.savez_compressed(os.path.join(HERE, 'quant.npz'), w=w.numpy(), x=x.numpy(), master=master.numpy(),
                        q=q.numpy(), xq4=xq.numpy(), xq1=x1.numpy(), qf=qf.numpy(), prune70=pm.numpy())
    print('quant.npz')


def gen_quant_inp():
    """The shipped QuantExp configuration (cfg/20200910_QuantExp/...whcgs.cfg:166-171, 202-207): 8-bit weights,
    16-bit activations (Quantize_inp of the layer input AND of h_t at every time step), magnitude pruning 70."""
    run_model('train_quant_inp', 'train', T=6, B=3, D=12,
              lstm_over=dict(lstm_quant='True', lstm_quant_inp='True', lstm_prune='True', lstm_prune_perc='70,70',
                             lstm_lay='24,24'),
              mlp_over=dict(mlp_quant='True', mlp_quant_inp='True', mlp_prune='True', mlp_prune_perc='70'),
              steps=1, seed=6)


def gen_norm_inp():
    """Input normalisations and the reference's own LayerNorm: LSTM with lstm_use_laynorm_inp + lstm_use_batchnorm_inp
    (neural_networks.py:825-830), MLP heads with dnn_use_laynorm_inp + dnn_use_batchnorm_inp (:248-252) and
    dnn_use_laynorm between the linear and the activation (:306-307)."""
    run_model('train_norm_inp', 'train', T=6, B=4, D=12,
              lstm_over=dict(lstm_use_laynorm_inp='True', lstm_use_batchnorm_inp='True', lstm_lay='24,24'),
              mlp_over=dict(dnn_use_laynorm_inp='True', dnn_use_batchnorm_inp='True', dnn_use_laynorm='True'),
              steps=2, seed=8)


# ---- pattern masks -----------------------------------------------------------------------
def gen_laynorm_loop():
    """lstm_use_laynorm: the reference's LayerNorm class (:40-51) applied to h_t INSIDE the time loop (:1092-1095);
    two train steps, so the second runs with gamma / beta that the optimizer has moved."""
    run_model('train_laynorm', 'train', T=6, B=3, D=12, steps=2, seed=17,
              lstm_over=dict(lstm_use_laynorm='True,True', lstm_lay='24,20'))
    run_model('train_laynorm_nobn', 'train', T=9, B=11, D=12, steps=2, seed=18,
              lstm_over=dict(lstm_use_laynorm='True,False', lstm_use_batchnorm='False,True', lstm_act='tanh,relu',
                             lstm_drop='0.0,0.2'))

def gen_mlp_ln_bn():
    """dnn_use_laynorm and dnn_use_batchnorm on the same MLP layer: drop(act(bn(ln(wx(x))))) (neural_networks.py:312-314),
    a hidden ReLU layer with both in front of the softmax layer of each head."""
    run_model('train_mlp_ln_bn', 'train', T=6, B=4, D=12, steps=2, seed=9,
              mlp_over=dict(dnn_lay='16,10', dnn_drop='0.0,0.0', dnn_use_batchnorm='True,False',
                            dnn_use_laynorm='True,False', dnn_act='relu,softmax', param_quant='8,8', mlp_prune_perc='70,70'),
              mono_over=dict(dnn_lay='16,5'))

def _cpu_cuda_shim():
    """SURVEY 8(c) shim 3: sparsity/sparsity.py:1045, 1125, 1131 call .cuda() unconditionally; on this GPU-less
    host Tensor.cuda() becomes the identity (nothing else of the reference is touched)."""
    torch.Tensor.cuda = lambda self, *a, **k: self


def gen_patterns():
    """Pattern.Pattern (pattern / coo / pattern_coo, drawn and file-backed pattern sets), sparsity.apply_patterns
    (incl. ties: overlapping patterns add up to mask values of 2) and sparsity.find_top_k_by_kmeans."""
    _cpu_cuda_shim()
    import Pattern as ref_pattern
    from sparsity import sparsity as ref_sp
    out = {}
    cases = []
    k = 0
    for (r, c, seed) in ((64, 40, 21), (32, 48, 22)):
        for mode in ('pattern', 'coo', 'pattern_coo'):
            for from_file in ('', 'b08b08_k04_n16_pattern.npy', 'b08b08_k04_n16_pattern_v1.npy',
                              'b08b08_k04_n16_pattern_v2.npy'):
                if from_file and mode == 'coo':
                    continue
                torch.manual_seed(seed)
                w = torch.randn(r, c)
                np.random.seed(seed)
                if from_file:
                    # Pattern.get_pattern tests and loads the module-level name `pattern_from_file`, which the
                    # reference never defines (Pattern.py:63-67, a NameError as shipped); the harness provides it
                    ref_pattern.pattern_from_file = os.path.join(REF, 'pattern_file', from_file)
                m = ref_pattern.Pattern(w.numpy(), mode, [8, 8], 4, 16, '', pattern_from_file=bool(from_file))
                mask = m.mask.detach().numpy()
                assert mask.dtype == np.float64 and set(np.unique(mask)) <= {0.0, 1.0}
                out['w_%d' % k] = w.numpy()
                out['bits_%d' % k] = np.packbits(mask.astype(np.uint8))
                out['next_rand_%d' % k] = np.array([np.random.randint(0, 2 ** 31 - 1)])
                cases.append(repr((r, c, mode, from_file, seed)))
                k += 1
    out['cases'] = np.array(cases)
    for f in ('b08b08_k04_n16_pattern.npy', 'b08b08_k04_n16_pattern_v1.npy', 'b08b08_k04_n16_pattern_v2.npy'):
        out['file/' + f] = np.packbits(np.load(os.path.join(REF, 'pattern_file', f)).astype(np.uint8))
    # apply_patterns on fixed kernels; kernel set B contains a duplicated and an overlapping pattern => ties => 2s
    pats = np.split(np.load(os.path.join(REF, 'pattern_file', 'b08b08_k04_n16_pattern.npy')), 16, 0)
    kern_a = torch.from_numpy(np.stack(pats)[:, None].astype(np.float32))
    kern_b = kern_a.clone()
    kern_b[1] = kern_b[0]
    torch.manual_seed(31)
    w = torch.randn(48, 64)
    w[8:16, 0:8] = 1.0        # a constant block: every pattern with 4 ones ties
    for name, kern in (('a', kern_a), ('b', kern_b)):
        mask = ref_sp.apply_patterns(w, kern)
        out['apply_%s_kernel' % name] = kern.numpy()
        out['apply_%s_mask' % name] = mask.numpy()
    assert out['apply_b_mask'].max() >= 2.0
    out['apply_w'] = w.numpy()
    # k-means pattern mining (sklearn draws its initialisation from the global numpy RNG)
    torch.manual_seed(32)
    w = torch.randn(64, 48)
    np.random.seed(32)
    kern = ref_sp.find_top_k_by_kmeans(w, 16, [8, 8], 4, [8, 8])
    out['kmeans_w'] = w.numpy()
    out['kmeans_kernel'] = kern.numpy()
    out['kmeans_next_rand'] = np.array([np.random.randint(0, 2 ** 31 - 1)])
    np.savez_compressed(os.path.join(HERE, 'patterns.npz'), **out)
    print('patterns.npz', k, 'Pattern cases; tie mask max', out['apply_b_mask'].max())


def gen_pattern_model():
    """LSTM + cd head with if_pattern=True (neural_networks.py:876-884, 265-272, 1162-1237): two train steps; the
    mined kernels and the pattern masks are dumped so that the product can be checked with k-means out of the loop."""
    _cpu_cuda_shim()

    def extra(out, nns):
        lstm, cd = nns['LSTM_layers'], nns['MLP_layers']
        for name, lst in lstm.pattern.items():
            for i, kern in enumerate(lst):
                out['pattern/LSTM_layers/%s/%d' % (name, i)] = kern.numpy()
        for name, lst in lstm.pattern_mask.items():
            for i, m in enumerate(lst):
                out['pattern_mask/LSTM_layers/%s/%d' % (name, i)] = m.numpy()
        for i, kern in enumerate(cd.pattern):
            out['pattern/MLP_layers/%d' % i] = kern.numpy()
        for i, m in enumerate(cd.pattern_mask):
            out['pattern_mask/MLP_layers/%d' % i] = m.numpy()

    pat = dict(if_pattern='True', pattern_mode='pattern', pattern_shape='8,8')
    run_model('train_pattern', 'train', T=6, B=3, D=16,
              lstm_over=dict(pat, pattern_nnz='4,4', pattern_num='4,6', lstm_lay='32,40', lstm_hcgs='False'),
              mlp_over=dict(pat, pattern_nnz='8', pattern_num='5', dnn_lay='16', mlp_hcgs='False'),
              mono_over=dict(if_pattern='False'), steps=2, seed=12, extra=extra)
    # HCGS and pattern masks together (the pattern search then sees HCGS-masked W but not yet masked U, :858-884)
    run_model('train_pattern_hcgs', 'train', T=5, B=2, D=16,
              lstm_over=dict(pat, pattern_nnz='4', pattern_num='4', lstm_lay='32', lstm_drop='0.2',
                             lstm_use_laynorm='False', lstm_use_batchnorm='True', lstm_act='tanh',
                             hcgsx_block='16,4', hcgsh_block='16,4'),
              mlp_over=dict(dnn_lay='10'), steps=1, seed=13, extra=lambda out, nns: [
                  out.__setitem__('pattern_mask/LSTM_layers/%s/%d' % (n, i), m.numpy())
                  for n, lst in nns['LSTM_layers'].pattern_mask.items() for i, m in enumerate(lst)] and None)


def gen_bidir():
    """bidirectional LSTM.  The reference's LSTM carries the whole bidirectional branch (:836-838, 1104-1107)
    but switches it off with `self.bidir = 0` at :835 -- after __init__ has sized the layers for 2H -- so every
    lstm_bidir = True cfg crashes as shipped.  The fixture comes from the reference source with exactly that ONE line
    removed (in memory; nothing is written), driven through the reference's own model_init / forward_model."""
    src = open(os.path.join(REF, 'neural_networks.py')).read()
    marker = ('            self.bidir = 0\n            if self.bidir:\n'
              '                h_init = torch.zeros(2 * x.shape[1], self.lstm_lay[i])')
    assert src.count(marker) == 1
    src = src.replace(marker, marker.replace('            self.bidir = 0\n', ''))
    mod = types.ModuleType('neural_networks_bidir')
    mod.__file__ = os.path.join(REF, 'neural_networks.py')
    exec(compile(src, mod.__file__ + ' [line 835 removed]', 'exec'), mod.__dict__)
    sys.modules['neural_networks_bidir'] = mod
    run_model('train_bidir', 'train', T=7, B=3, D=12,
              lstm_over=dict(lstm_bidir='True', arch_library='neural_networks_bidir', lstm_lay='24,16'),
              steps=2, seed=21)


def rnn_family_opts(prefix, cls, **over):
    """cfg section of one of the reference's dense recurrent classes (cfg/TIMIT_baselines/TIMIT_{GRU,liGRU,RNN}_*.cfg)."""
    p = prefix
    o = {'arch_name': 'LSTM_layers', 'arch_proto': 'proto/%s.proto' % cls, 'arch_library': 'neural_networks',
         'arch_class': cls, 'arch_pretrain_file': 'none', 'arch_freeze': 'False', 'arch_seq_model': 'True',
         p + '_lay': '24,16', p + '_drop': '0.2,0.2', p + '_use_laynorm_inp': 'False', p + '_use_batchnorm_inp': 'False',
         p + '_use_laynorm': 'False,False', p + '_use_batchnorm': 'True,True', p + '_bidir': 'False',
         p + '_act': 'tanh,tanh', p + '_orthinit': 'True', 'arch_lr': '0.0016', 'arch_halving_factor': '0.5',
         'arch_improvement_threshold': '0.001', 'arch_opt': 'rmsprop', 'opt_momentum': '0.0', 'opt_alpha': '0.95',
         'opt_eps': '1e-8', 'opt_centered': 'False', 'opt_weight_decay': '0.0', 'out_folder': ''}
    o.update(over)
    return o


RNN_FAMILY = [   # (fixture, option prefix, class, overrides)
    ('train_gru', 'gru', 'GRU', {}),
    ('train_gru_bidir_nobn', 'gru', 'GRU', {'gru_bidir': 'True', 'gru_use_batchnorm': 'False,False', 'gru_drop': '0.0,0.1'}),
    ('train_ligru', 'ligru', 'liGRU', {'ligru_act': 'relu,relu'}),
    ('train_minimalgru', 'minimalgru', 'minimalGRU', {}),
    ('train_rnn', 'rnn', 'RNN', {'rnn_act': 'relu,relu'}),
]


def gen_gru_family():
    """the reference's own GRU / liGRU / minimalGRU / RNN (dense; neural_networks.py:1240-1931) through
    model_init / forward_model: two train steps each.  The plugin's block-sparse versions must reproduce them with
    all-ones masks."""
    for k, (tag, prefix, cls, over) in enumerate(RNN_FAMILY):
        run_model(tag, 'train', T=6, B=3, D=12, arch1=rnn_family_opts(prefix, cls, **over), steps=2, seed=30 + k)


def gen_run_nn():
    """SURVEY a12 / 8(c) shim 4: the reference's OWN chunk loop core.run_nn (train chunk, then a forward chunk from the
    training checkpoints) over a synthetic chunk, with the reference's own nets on the CPU.  tests/test_reference_loop.py
    runs the same loop with the plugin classes on the GPU and compares with what is stored here."""
    sys.path.insert(0, os.path.join(os.path.dirname(HERE)))
    import ref_harness as rh
    import core as ref_core
    rh.cpu_shims(sys.modules)
    chunk = rh.synthetic_chunk()
    a1, a2, a3 = (dict(LSTM_OPTS), dict(MLP_OPTS), dict(MLP_OPTS, arch_name='MLP_layers2', dnn_lay='5',
                                                        mlp_hcgs='False', arch_lr='0.0004'))
    work = tempfile.mkdtemp()
    res = rh.run_chunk_pair(ref_core, chunk, [a1, a2, a3], work, use_cuda=False)
    out = {'loss': np.array([res['loss'], res['err']]), 'ark': np.frombuffer(res['ark'], dtype=np.uint8),
           'files': np.array([f for f in res['files'] if not f.endswith('.cfg')])}
    for arch, sd in res['model_par'].items():
        for k, v in sd.items():
            out['model_par/%s/%s' % (arch, k)] = v.numpy()
    np.savez_compressed(os.path.join(HERE, 'run_nn_chunk.npz'), **out)
    print('run_nn_chunk.npz loss', res['loss'], 'err', res['err'], len(res['ark']), 'ark bytes', res['files'])


def gen_chunk_batches():
    """the minibatch tensors the reference's OWN loop (core.py:183-200) hands to forward_model for a
    synthetic chunk -- random left padding drawn from Python's `random` after core.run_nn's random.seed(seed) -- and the
    next random.random() after the chunk (position of the stream).  oracle/chunk_oracle.py and the device assembly
    (b200/chunk.py, hcgs_assemble_batch) are compared with these."""
    import random
    sys.path.insert(0, os.path.join(os.path.dirname(HERE)))
    import ref_harness as rh
    import core as ref_core
    rh.cpu_shims(sys.modules)
    a = [dict(LSTM_OPTS), dict(MLP_OPTS), dict(MLP_OPTS, arch_name='MLP_layers2', dnn_lay='5', mlp_hcgs='False',
                                                arch_lr='0.0004')]
    out = {}
    for tag, n_snt, bs in (('a', 7, 2), ('b', 9, 4), ('c', 5, 1)):
        chunk = rh.synthetic_chunk(n_snt=n_snt, seed=20 + n_snt)
        seen = []
        orig = ref_core.forward_model

        def rec(fea_dict, lab_dict, arch_dict, model, nns, costs, inp, *rest):
            seen.append(inp.detach().clone().numpy())
            return orig(fea_dict, lab_dict, arch_dict, model, nns, costs, inp, *rest)
        ref_core.forward_model = rec
        try:
            work = tempfile.mkdtemp()
            rh.patch_read_lab_fea(ref_core, chunk)
            cfg = os.path.join(work, 'train_chunk.cfg')
            rh.write_chunk_cfg(cfg, 'train', work, a, False, seed=2234, batch_size=bs)
            ref_core.run_nn(None, None, None, None, None, None, cfg, True, cfg)
            nxt = random.random()
        finally:
            ref_core.forward_model = orig
        out['%s/n_snt' % tag] = np.array([n_snt, bs, 20 + n_snt])
        out['%s/next_random' % tag] = np.array([nxt])
        for i, x in enumerate(seen):
            out['%s/inp%d' % (tag, i)] = x
        print('chunk_batches', tag, [x.shape for x in seen], nxt)
    np.savez_compressed(os.path.join(HERE, 'chunk_batches.npz'), **out)


if __name__ == '__main__':
    if len(sys.argv) > 1:
        for name in sys.argv[1:]:
            globals()[name]()
        sys.exit(0)
    gen_patterns()
    gen_pattern_model()
    gen_masks()
    run_model('train_small', 'train', steps=2)
    run_model('train_htanh', 'train', T=5, B=9, D=12, lstm_over=dict(lstm_act='htanh,htanh', lstm_lay='20,28',
                                                                     lstm_drop='0.0,0.1'), steps=1, seed=11)
    run_model('valid_small', 'valid', steps=1)
    gen_posterior()
    gen_quant()
    run_model('train_quant', 'train', lstm_over=dict(lstm_quant='True', lstm_lay='24,24'),
              mlp_over=dict(mlp_quant='True'), steps=1, seed=5)
    gen_quant_inp()
    gen_norm_inp()
