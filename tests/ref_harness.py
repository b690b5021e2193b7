"""Harness for driving the REFERENCE's own chunk loop (core.run_nn, core.py:24-362) over synthetic data
(SURVEY 8(c) shim 4).  Test / bench infrastructure only.

The reference reads a chunk with data_io.read_lab_fea (Kaldi binaries); here `core.read_lab_fea` is replaced by a
function that appends a synthetic `shared_list` with the layout of data_io.py:276-282:
    [data_name, data_end_index, fea_dict, lab_dict, arch_dict, data_set]
Everything after that -- model_init, optimizer_init, the minibatch loop with its random left padding, forward_model,
backward, optimizer step, checkpoint pickles, posterior normalisation, data_io.write_mat, the .info file -- is the
reference's code, unchanged.  Which `neural_networks` module it instantiates is decided by sys.path alone
(`arch_library = neural_networks`, utils.py:1768-1770): the reference's own, or the B200 plugin.

Used by tests/golden/make_golden.py (reference nets on the CPU -> fixture), tests/test_reference_loop.py (plugin nets
on the GPU, same loop) and bench.py --impl reference.
"""
import configparser
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_INSTALLED = os.path.join(ROOT, 'baseline', '_ref')

MODEL_2HEAD = ['out_dnn1=compute(LSTM_layers,fbank)', 'out_dnn2=compute(MLP_layers,out_dnn1)',
               'out_dnn3=compute(MLP_layers2,out_dnn1)', 'loss_mono=cost_nll(out_dnn3,lab_mono)',
               'loss_mono_w=mult_constant(loss_mono,1.0)', 'loss_cd=cost_nll(out_dnn2,lab_cd)',
               'loss_final=sum(loss_cd,loss_mono_w)', 'err_final=cost_err(out_dnn2,lab_cd)']


def reference_dir():
    """/root/reference in the build container, else the copy made by baseline/install_reference.py, else None."""
    for d in ('/root/reference', REF_INSTALLED):
        if os.path.exists(os.path.join(d, 'core.py')):
            return d
    return None


class _TorchProxy(types.ModuleType):
    """`torch` as seen by the reference's hcgs.py / guided_hcgs.py: device("cuda") resolves to the CPU (their
    unconditional mask.to("cuda"), hcgs.py:71-72, would otherwise split a CPU model across devices)."""

    def __getattr__(self, k):
        return getattr(torch, k)

    def device(self, *_a, **_k):
        return torch.device('cpu')


def cpu_shims(modules):
    """SURVEY 8(c) shims 1 and 3 for running the reference's own classes on the host."""
    for name in ('hcgs', 'guided_hcgs'):
        if name in modules:
            modules[name].torch = _TorchProxy('torch')
    torch.Tensor.cuda = lambda self, *a, **k: self


def synthetic_chunk(n_snt=6, D=12, n_cd=10, n_mono=5, min_len=9, max_len=20, seed=5):
    """One chunk: sentences sorted by length (data_io.py:78), features N(0,1), labels as trailing float columns."""
    rng = np.random.RandomState(seed)
    lens = np.sort(rng.randint(min_len, max_len + 1, n_snt))
    rows = []
    for n in lens:
        x = rng.randn(n, D).astype(np.float32)
        lab = np.stack([rng.randint(0, n_cd, n), rng.randint(0, n_mono, n)], 1).astype(np.float32)
        rows.append(np.concatenate([x, lab], 1))
    data_set = np.concatenate(rows, 0)
    return {'data_name': ['utt_%04d' % i for i in range(n_snt)], 'data_end_index': np.cumsum(lens),
            'data_set': data_set, 'D': D}


def shared_list_for(chunk):
    D = chunk['D']
    fea_dict = {'fbank': ['scp', 'opts', 0, 0, 0, 0, D, D]}
    lab_dict = {'lab_cd': ['f', 'o', 'c', D], 'lab_mono': ['f', 'o', 'c', D + 1]}
    arch_dict = {'LSTM_layers': ['architecture1', 'LSTM_layers', True],
                 'MLP_layers': ['architecture2', 'MLP_layers', False],
                 'MLP_layers2': ['architecture3', 'MLP_layers2', False]}
    return [list(chunk['data_name']), chunk['data_end_index'].copy(), fea_dict, lab_dict, arch_dict,
            chunk['data_set'].copy()]


def patch_read_lab_fea(core_module, chunk):
    """shim 4: the chunk reader of core.run_nn (called on a thread for this and for the next chunk)."""
    def fake(cfg_file, is_production, shared_list, output_folder):
        shared_list.extend(shared_list_for(chunk))
    core_module.read_lab_fea = fake


def write_chunk_cfg(path, to_do, out_folder, sections, use_cuda, seed=2234, batch_size=2, model=MODEL_2HEAD,
                    count_file='', pretrain=None):
    """The chunk-specific cfg run_exp.py writes for core.run_nn (only the keys run_nn and model_init read)."""
    cfg = configparser.ConfigParser()
    info = os.path.join(out_folder, 'exp_files', '%s_chunk.info' % to_do)
    os.makedirs(os.path.dirname(info), exist_ok=True)
    cfg['exp'] = dict(seed=str(seed), out_folder=out_folder, use_cuda=str(use_cuda), multi_gpu='False', to_do=to_do,
                      out_info=info, save_gpumem='False', production='False')
    cfg['model'] = {'model': '\n'.join(model)}
    cfg['forward'] = dict(forward_out='out_dnn2', normalize_posteriors='True', normalize_with_counts_from=count_file,
                          require_decoding='True')
    cfg['batches'] = dict(batch_size_train=str(batch_size), batch_size_valid=str(batch_size))
    for k, sec in enumerate(sections):
        sec = dict(sec, out_folder=out_folder)
        sec.pop('to_do', None)
        sec.pop('use_cuda', None)
        if pretrain:
            sec['arch_pretrain_file'] = pretrain[k]
        cfg['architecture%d' % (k + 1)] = sec
    with open(path, 'w') as f:
        cfg.write(f)
    return info


def write_counts(path, n, seed=3):
    counts = (np.random.RandomState(seed).rand(n) * 1000 + 1).astype(np.float32)
    with open(path, 'w') as f:
        f.write('[ ' + ' '.join('%g' % v for v in counts) + ' ]\n')
    return path


def run_chunk_pair(core_module, chunk, sections, work_dir, use_cuda, seed=2234):
    """One training chunk followed by a forward chunk that starts from the training chunk's checkpoints, both through
    core.run_nn.  Returns {'loss', 'err', 'model_par': {arch: state_dict}, 'ark': bytes, 'info': {...}}."""
    patch_read_lab_fea(core_module, chunk)
    counts = write_counts(os.path.join(work_dir, 'cd.count'), int(sections[1]['dnn_lay'].split(',')[-1]))
    tr_cfg = os.path.join(work_dir, 'train_chunk.cfg')
    info = write_chunk_cfg(tr_cfg, 'train', work_dir, sections, use_cuda, seed=seed, count_file=counts)
    core_module.run_nn(None, None, None, None, None, None, tr_cfg, True, tr_cfg)
    res = {'model_par': {}}
    parser = configparser.ConfigParser()
    parser.read(info)
    res['loss'], res['err'] = float(parser['results']['loss']), float(parser['results']['err'])
    pkls = [info.replace('.info', '_architecture%d.pkl' % (k + 1)) for k in range(len(sections))]
    for k, p in enumerate(pkls):
        ck = torch.load(p, map_location='cpu', weights_only=False)
        res['model_par']['architecture%d' % (k + 1)] = {a: b.detach().cpu() for a, b in ck['model_par'].items()}
    fw_cfg = os.path.join(work_dir, 'forward_chunk.cfg')
    finfo = write_chunk_cfg(fw_cfg, 'forward', work_dir, sections, use_cuda, seed=seed, count_file=counts,
                            pretrain=pkls)
    core_module.run_nn(None, None, None, None, None, None, fw_cfg, True, fw_cfg)
    with open(finfo.replace('.info', '_out_dnn2_to_decode.ark'), 'rb') as f:
        res['ark'] = f.read()
    res['files'] = sorted(os.listdir(work_dir))
    return res
