#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list:
per-kernel launch count, summed device time and share (cold-cache, serialised:
compare SHARES, not absolutes).  usage: summarize_launches.py launches.csv [out.md]"""
import collections
import csv
import re
import sys


def main():
    path = sys.argv[1]
    lines = [l for l in open(path) if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    agg = collections.defaultdict(lambda: [0, 0.0])
    for row in r:
        name = re.sub(r'^void\s+', '', row[ki]).replace('<unnamed>::', '')
        name = re.sub(r'\((?!anonymous).*$', '', name)
        v = float(row[vi].replace(',', ''))
        v *= {'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 's': 1e6}.get(row[ui], 1.0)
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    out = ['| kernel | launches | device time (ms) | share |', '|---|---:|---:|---:|']
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        out.append('| `%s` | %d | %.3f | %.1f%% |' % (k[:90], v[0], v[1] / 1e3, 100 * v[1] / tot))
    out.append('| **total** | %d | %.3f | 100%% |' % (sum(v[0] for v in agg.values()), tot / 1e3))
    text = '\n'.join(out)
    if len(sys.argv) > 2:
        open(sys.argv[2], 'a').write(text + '\n')
    print(text)


if __name__ == '__main__':
    main()
