"""ms per step of every native entry point of the default train step (the untimed pass of bench.py), one line."""
import json, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run([sys.executable, os.path.join(ROOT, 'bench.py'), '--no-cpu-baseline', '--no-export', '--no-libri',
                      '--no-cfg-batch', '--steps', '8'], capture_output=True, text=True)
d = json.loads(out.stdout.strip().splitlines()[-1])
print('%.2f ms/step | ' % d['ms_per_step'] + ' '.join('%s=%.3f' % (k, v['ms_per_step']) for k, v in d['kernels_untimed_pass'].items() if v['ms_per_step'] > 0.1))
