"""the REFERENCE's own chunk loop (core.run_nn + utils.model_init / optimizer_init / forward_model +
data_io.write_mat, unchanged) driving the B200 plugin classes -- the drop-in claim shown, not asserted.

sys.path order is the whole integration: the plugin package precedes the reference directory, so
`arch_library = neural_networks` (utils.py:1768-1770) resolves to the plugin while core / utils / data_io stay the
reference's.  The chunk reader is replaced by a synthetic one (tests/ref_harness.py, SURVEY 8(c) shim 4).

Expected values: tests/golden/run_nn_chunk.npz = the same loop run with the reference's own nets on the CPU
(tests/golden/make_golden.py gen_run_nn).  The reference sources come from /root/reference in the build container
and from baseline/_ref (copied by baseline/install_reference.py, git-ignored) on the GPU box."""
import os
import sys

import numpy as np
import pytest
import torch

import ref_harness as rh
from helpers import LSTM_OPTS, MLP_OPTS, load_golden

REF = rh.reference_dir()
needs_ref = pytest.mark.skipif(REF is None, reason='reference sources not present (baseline/install_reference.py)')


def _import_ref_core():
    if REF not in sys.path:
        sys.path.append(REF)            # AFTER the plugin package: neural_networks, hcgs, ... stay the plugin's
    import core
    import neural_networks
    import utils
    assert os.path.dirname(os.path.abspath(core.__file__)) == REF and os.path.dirname(utils.__file__) == REF
    assert 'pytorch-kaldi-cgs_b200' in neural_networks.__file__
    return core, utils, neural_networks


def _sections():
    return [dict(LSTM_OPTS), dict(MLP_OPTS),
            dict(MLP_OPTS, arch_name='MLP_layers2', dnn_lay='5', mlp_hcgs='False', arch_lr='0.0004')]


@needs_ref
def test_reference_model_init_builds_plugin_classes(tmp_path):
    core, utils, plugin = _import_ref_core()
    import configparser
    chunk = rh.synthetic_chunk()
    cfg_path = str(tmp_path / 'c.cfg')
    rh.write_chunk_cfg(cfg_path, 'train', str(tmp_path), _sections(), use_cuda=False)
    config = configparser.ConfigParser()
    config.read(cfg_path)
    _, _, fea_dict, _, arch_dict, _ = rh.shared_list_for(chunk)
    torch.manual_seed(2234)
    np.random.seed(2234)
    nns, costs = utils.model_init(fea_dict, rh.MODEL_2HEAD, config, arch_dict, False, False, 'train')
    opts = utils.optimizer_init(nns, config, arch_dict)
    assert type(nns['LSTM_layers']) is plugin.LSTM and type(nns['MLP_layers']) is plugin.MLP
    assert set(opts) == set(nns) and isinstance(opts['LSTM_layers'], torch.optim.RMSprop)
    assert sorted(costs) == ['loss_cd', 'loss_mono']
    # the attributes core.run_nn reads (core.py:123-131, 292-306)
    for net in nns.values():
        assert net.prune in (0, 1, False, True) and hasattr(net, 'guided_hcgs') and hasattr(net, 'if_pattern')


@needs_ref
@pytest.mark.skipif(torch.cuda.is_available(), reason='CPU-only check')
def test_reference_loop_reaches_the_plugin_forward_and_has_no_cpu_fallback(tmp_path):
    """Without a GPU the reference loop runs up to the first forward of the plugin, which must raise (no CPU path)."""
    core, _, _ = _import_ref_core()
    from b200 import native
    rh.patch_read_lab_fea(core, rh.synthetic_chunk())
    cfg_path = str(tmp_path / 'train_chunk.cfg')
    rh.write_chunk_cfg(cfg_path, 'train', str(tmp_path), _sections(), use_cuda=False)
    with pytest.raises(native.NativeError):
        core.run_nn(None, None, None, None, None, None, cfg_path, True, cfg_path)


@needs_ref
@pytest.mark.gpu
def test_reference_run_nn_drives_plugin_train_and_forward_chunks(tmp_path):
    import posterior_io
    core, _, _ = _import_ref_core()
    g = load_golden('run_nn_chunk')
    from b200 import native
    native.reset_launch_count()
    res = rh.run_chunk_pair(core, rh.synthetic_chunk(), _sections(), str(tmp_path), use_cuda=True)
    assert native.LAUNCHES.get('hcgs_lstm_fwd', 0) >= 2 * (3 + 6) and native.LAUNCHES.get('hcgs_lstm_bwd', 0) == 2 * 3
    # running loss / error of the training chunk, as written to the .info file by the reference
    assert abs(res['loss'] - g['loss'][0]) <= 2e-5 * abs(g['loss'][0]), (res['loss'], g['loss'][0])
    assert abs(res['err'] - g['loss'][1]) <= 1e-6
    # checkpoints (three RMSprop steps): masks bit-equal, kept weights close, BN statistics close
    for key in g.files:
        if not key.startswith('model_par/'):
            continue
        _, arch, k = key.split('/')
        sd = res['model_par'][arch]
        ref, mine = torch.from_numpy(g[key]), sd[k]
        if k.endswith('mask'):
            assert torch.equal(mine, ref), key
            continue
        if k.endswith('num_batches_tracked'):
            assert int(mine) == int(ref), key
            continue
        mk = None
        if arch == 'architecture1' and k[0] in 'wu' and k.endswith('.weight'):
            mk = sd[('hcgsx.%s.mask' if k[0] == 'w' else 'hcgsh.%s.mask') % k.split('.')[1]]
        if arch == 'architecture2' and k == 'wx.0.weight':
            mk = sd['hcgs.0.mask']
        if mk is not None:
            mine, ref = mine * mk, ref * mk
        err = (mine.double() - ref.double()).abs().max().item()
        assert err <= 2e-4 * max(ref.abs().max().item(), 1e-30) + 1e-7, (key, err)
    # posteriors: same utterance keys, shapes and Kaldi framing; values follow three optimizer steps
    ark = str(tmp_path / 'ref.ark')
    with open(ark, 'wb') as f:
        f.write(bytes(g['ark']))
    mine_path = str(tmp_path / 'mine.ark')
    with open(mine_path, 'wb') as f:
        f.write(res['ark'])
    a, b = posterior_io.read_ark(mine_path), posterior_io.read_ark(ark)
    assert [k for k, _ in a] == [k for k, _ in b] and len(res['ark']) == len(bytes(g['ark']))
    for (_, ma), (_, mb) in zip(a, b):
        assert ma.shape == mb.shape and ma.dtype == np.float32
        assert np.abs(ma - mb).max() <= 1e-4 * np.abs(mb).max()
    # side effects of the first forward call (save_cgs_mat.py) -- the same files the reference wrote
    want = set(str(f) for f in g['files']) - {'exp_files'}
    assert want <= set(res['files']), sorted(want - set(res['files']))
