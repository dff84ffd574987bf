"""Summarise an ncu launch list (gpu__time_duration.sum) per (stream, kernel) for the second
half of the launches of la:: kernels (the timed repetition of tools/factor_once.py)."""
import csv
import sys
from collections import defaultdict


def us(r):
    v = float(r['Metric Value'].replace(',', ''))
    u = r['Metric Unit']
    return v * 1e-3 if u in ('ns', 'nsecond') else v * 1e3 if u in ('ms', 'msecond') else v


def main(path):
    lines = [l for l in open(path) if not l.startswith('==')]
    rows = [r for r in csv.DictReader(lines) if r['Metric Name'] == 'gpu__time_duration.sum']
    la = [r for r in rows if 'la::' in r['Kernel Name']]
    half = la[len(la) // 2:]
    agg = defaultdict(lambda: [0, 0.0])
    per_stream = defaultdict(float)
    for r in half:
        k = (r['Stream'], r['Kernel Name'].split('(')[0].split('::')[-1][:34])
        agg[k][0] += 1
        agg[k][1] += us(r)
        per_stream[r['Stream']] += us(r)
    print(path, len(half), 'launches')
    for st, t in per_stream.items():
        print(f'  stream {st}: {t / 1e3:.2f} ms')
    for k, v in sorted(agg.items()):
        print(f'  {k[0]:>3s} {k[1]:36s} {v[0]:5d} x {v[1] / v[0]:8.1f} us = {v[1] / 1e3:8.2f} ms')


if __name__ == '__main__':
    for p in sys.argv[1:]:
        main(p)
