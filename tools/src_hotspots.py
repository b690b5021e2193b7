#!/usr/bin/env python
"""Per-source-line hot spots from `ncu -i X.ncu-rep --page source --csv --print-source sass,cuda`.

    python tools/src_hotspots.py combined.csv [top_n]
Prints, per CUDA source line: stall samples, executed warp instructions and the top stall reasons."""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    cur_file, hdr, lines = None, None, []
    for r in rows:
        if not r:
            continue
        if r[0] == 'File Path':
            cur_file = r[1].split('/')[-1]
            continue
        if r[0] == 'Line No':
            hdr = r
            continue
        if hdr is None or r[0] in ('', 'Function Name'):
            continue
        d = dict(zip(hdr[4:], r[4:]))
        def num(k):
            try:
                return float(d.get(k, 0))
            except ValueError:
                return 0.0
        stalls = {k: num(k) for k in d if k.startswith('stall_') and '(Not Issued)' not in k}
        lines.append((num('# Samples'), num('Instructions Executed'), cur_file, r[0], r[1].strip(), stalls))
    tot_s = sum(l[0] for l in lines) or 1
    tot_i = sum(l[1] for l in lines) or 1
    print('total samples %d, total warp instructions %d' % (tot_s, tot_i))
    for s, i, f, ln, src, st in sorted(lines, key=lambda x: -x[0])[:top]:
        why = ', '.join('%s %.0f%%' % (k[6:], 100 * v / max(s, 1)) for k, v in sorted(st.items(), key=lambda x: -x[1])[:3] if v)
        print('%5.1f%% smp %5.1f%% inst  %s:%s  %s\n        [%s]' % (100 * s / tot_s, 100 * i / tot_i, f, ln, src[:110], why))


if __name__ == '__main__':
    main()
