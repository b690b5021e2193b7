#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py into a per-step table.

    python tools/launch_summary.py gpurun_out/launches.csv [marker-substring] [launches-of-marker-per-step]

One step = the launches between the n-th and (n+per_step)-th occurrence of the marker kernel
(default: the forward recurrent kernel, 4 per step), taken from the LAST complete step in the list."""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    marker = sys.argv[2] if len(sys.argv) > 2 else 'lstm_tc_fwd'
    per_step = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    with open(path) as f:
        lines = [l for l in f if not l.startswith('==')]
    rows = list(csv.DictReader(lines))
    idx = [i for i, r in enumerate(rows) if marker in r['Kernel Name']]
    n_steps = len(idx) // per_step
    if n_steps < 2:
        raise SystemExit('need at least two steps of %s in the list' % marker)
    a, b = idx[(n_steps - 2) * per_step], idx[(n_steps - 1) * per_step]
    tot, cnt = collections.OrderedDict(), {}
    for r in rows[a:b]:
        n = re.sub(r'\(.*', '', re.sub(r'<.*', '', r['Kernel Name'])).replace('void ', '')
        v = float(r['Metric Value']) / 1e6
        tot[n] = tot.get(n, 0.0) + v
        cnt[n] = cnt.get(n, 0) + 1
    s = sum(tot.values())
    print('| kernel | launches/step | ms/step | share |')
    print('|---|---|---|---|')
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        print('| `%s` | %d | %.3f | %.1f %% |' % (k, cnt[k], v, 100 * v / s))
    print('| **total** | %d | %.3f | |' % (b - a, s))


if __name__ == '__main__':
    main()
