"""Recipe: put the UNMODIFIED reference (hellboywyh/pytorch-kaldi-CGS) under baseline/_ref so that it travels to the
GPU box (baseline/_ref is git-ignored -- reference sources never enter the history -- but not gpurun-ignored).

The reference has no setup.py / pyproject.toml (`pip install /root/reference` fails: "neither 'setup.py' nor
'pyproject.toml' found"), and it needs none: it is a flat directory of importable Python modules that run_exp.py uses
through sys.path.  "Installing" it is therefore copying its modules where they lie:

    python baseline/install_reference.py            # /root/reference -> baseline/_ref (no-op if the source is absent)

What uses baseline/_ref (never the product path):
  * bench.py --impl reference and bench.py's cpu_baseline legs: the reference's own neural_networks.LSTM / MLP,
    utils.model_init / optimizer_init / forward_model on the host cores (kind = "reference");
  * tests/test_reference_loop.py: the reference's own core.run_nn / utils.forward_model driving the plugin classes.
"""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(ROOT, 'baseline', '_ref')
MODULES = ['core.py', 'utils.py', 'data_io.py', 'neural_networks.py', 'hcgs.py', 'HCGS.py', 'cgs_base.py',
           'guided_hcgs.py', 'guided_cgs_base.py', 'guided_choices.py', 'quantized_modules.py', 'save_cgs_mat.py',
           'Pattern.py', 'config_train.py', 'run_exp.py']
TREES = ['sparsity', 'pattern_file', 'proto']


def install(src='/root/reference', dest=DEST):
    if not os.path.isdir(src):
        return os.path.isdir(dest)
    os.makedirs(dest, exist_ok=True)
    for m in MODULES:
        shutil.copy2(os.path.join(src, m), os.path.join(dest, m))
    for t in TREES:
        d = os.path.join(dest, t)
        if os.path.isdir(d):
            shutil.rmtree(d)
        shutil.copytree(os.path.join(src, t), d, ignore=shutil.ignore_patterns('*.ipynb', '*.mat', '*.sh'))
    with open(os.path.join(dest, 'INSTALLED_FROM'), 'w') as f:
        f.write('%s (unmodified copy made by baseline/install_reference.py)\n' % src)
    return True


if __name__ == '__main__':
    ok = install(*(sys.argv[1:2]))
    print('baseline/_ref %s' % ('installed' if ok else 'unavailable (no reference source here)'))
