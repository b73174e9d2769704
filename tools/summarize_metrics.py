"""Aggregate an `ncu --metrics gpu__time_duration.sum,dram__bytes_*.sum,lts__t_bytes.sum,sm__pipe_tensor_cycles_active...,
gpu__dram_throughput...,sm__throughput... --csv` capture of one register() by kernel (+grid) into a markdown table:
time share, DRAM GB/s achieved (read+write bytes / duration), L2 request bytes, tensor-pipe / DRAM / SM utilisation.
Durations are ncu's (cold cache, serialised, replayed): compare shares and utilisation, not absolute times.

  python tools/summarize_metrics.py gpurun_out/kmetrics.csv [hbm_peak_GBs]"""
import csv
import json
import os
import sys
from collections import defaultdict

SHORT = {
    'gpu__time_duration.sum': 'ns',
    'dram__bytes_read.sum': 'rd',
    'dram__bytes_write.sum': 'wr',
    'lts__t_bytes.sum': 'l2',
    'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed': 'tensor',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed': 'dram',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed': 'sm',
}
SCALE = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'ns': 1.0, 'nsecond': 1.0, 'us': 1e3, 'usecond': 1e3,
         'ms': 1e6, 'msecond': 1e6, '%': 1.0}


def main(path, hbm_peak):
  with open(path) as f:
    lines = [l for l in f if not l.startswith('==')]
  per = defaultdict(dict)
  for r in csv.DictReader(lines):
    m = SHORT.get(r['Metric Name'])
    if m is None:
      continue
    v = float(r['Metric Value'].replace(',', '')) * SCALE.get(r['Metric Unit'], 1.0)
    d = per[int(r['ID'])]
    d[m] = v
    d['key'] = (r['Kernel Name'].split('(')[0].replace('void ', '').replace('<unnamed>::', '')[:48], r['Grid Size'])
  agg = defaultdict(lambda: defaultdict(float))
  for d in per.values():
    a = agg[d['key']]
    a['n'] += 1
    for k in ('ns', 'rd', 'wr', 'l2'):
      a[k] += d.get(k, 0.0)
    for k in ('tensor', 'dram', 'sm'):          # time-weighted utilisation
      a[k] += d.get(k, 0.0) * d.get('ns', 0.0)
  tot = sum(a['ns'] for a in agg.values())
  print(f'total {tot/1e6:.2f} ms of kernel time over {len(per)} launches; HBM peak used for the fraction: {hbm_peak:.0f} GB/s\n')
  print('| kernel | grid | launches | total ms | share | avg us | DRAM MB/launch (rd+wr) | DRAM GB/s | frac of HBM peak | '
        'L2 req MB/launch | tensor pipe % | dram % | sm % |')
  print('|---|---|---|---|---|---|---|---|---|---|---|---|---|')
  for k, a in sorted(agg.items(), key=lambda kv: -kv[1]['ns']):
    n, ns = a['n'], a['ns']
    gbs = (a['rd'] + a['wr']) / ns if ns else 0.0
    print(f"| {k[0]} | {k[1]} | {int(n)} | {ns/1e6:.3f} | {100*ns/tot:.1f}% | {ns/n/1e3:.1f} | "
          f"{a['rd']/n/1e6:.1f}+{a['wr']/n/1e6:.1f} | {gbs:.0f} | {gbs/hbm_peak:.2f} | {a['l2']/n/1e6:.0f} | "
          f"{a['tensor']/ns:.1f} | {a['dram']/ns:.1f} | {a['sm']/ns:.1f} |")


if __name__ == '__main__':
  peak = float(sys.argv[2]) if len(sys.argv) > 2 else None
  if peak is None:
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')
    peak = json.load(open(p))['hbm_gbs'] if os.path.exists(p) else 6650.0
  main(sys.argv[1], peak)
