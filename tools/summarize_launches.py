"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel (+grid) -> markdown table."""
import csv
import sys
from collections import defaultdict


def main(path, by_grid=True):
  rows = []
  with open(path) as f:
    lines = [l for l in f if not l.startswith('==')]
  rd = csv.DictReader(lines)
  for r in rd:
    if r.get('Metric Name') != 'gpu__time_duration.sum':
      continue
    v = float(r['Metric Value'].replace(',', ''))
    unit = r['Metric Unit']
    us = v / 1e3 if unit in ('ns', 'nsecond') else v if unit in ('us', 'usecond') else v * 1e3 if unit.startswith('ms') else v
    name = r['Kernel Name']
    key = (name.split('(')[0][:70], r['Grid Size'] if by_grid else '')
    rows.append((key, us))
  agg = defaultdict(lambda: [0, 0.0])
  for k, us in rows:
    agg[k][0] += 1
    agg[k][1] += us
  tot = sum(v[1] for v in agg.values())
  print(f'total {tot/1e3:.2f} ms over {len(rows)} launches\n')
  print('| kernel | grid | launches | total ms | share | avg us |')
  print('|---|---|---|---|---|---|')
  for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f'| {k[0]} | {k[1]} | {n} | {us/1e3:.3f} | {100*us/tot:.1f}% | {us/n:.1f} |')


if __name__ == '__main__':
  main(sys.argv[1], by_grid=len(sys.argv) < 3)
