"""Stem kernel timing experiments (dev tool): FPB_STEM_DBG bits switch off stores (1), MMAs (8), input loads (128).
  python tools/stem_bench.py [--nhyp 252] [--reps 8]      runs modes 4 (im2col kernel) and 5 (net_stem.cu)"""
import argparse
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def one(a):
  import torch
  from foundationpose_b200._lib import lib, check, ptr, stream
  n = 2 * a.nhyp
  flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
  x = torch.randn(n, 160, 168, 8, device='cuda').bfloat16()
  w = (torch.randn(64, 448, device='cuda') / 20).bfloat16()
  out = torch.empty(n * 6400, 64, device='cuda', dtype=torch.bfloat16)
  sc = torch.ones(64, device='cuda'); sh = torch.zeros(64, device='cuda')
  for mode in (4, 5):
    def launch():
      check(lib().fpb_debug_igemm_ex(mode, ptr(x), ptr(w), ptr(out), n * 6400, 64, 448, 160, 161, 64, 80, 80, 7, 1, 2, 3, -1, 168, 8,
                                     ptr(sc), ptr(sh), None, 1, 1, stream()))
    for _ in range(2):
      launch()
    ts = []
    for _ in range(a.reps):
      flush.zero_()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record(); launch(); e1.record()
      torch.cuda.synchronize()
      ts.append(e0.elapsed_time(e1))
    print(f'dbg={os.environ.get("FPB_STEM_DBG", "0"):>4s} mode {mode}: {sorted(ts)[len(ts) // 2] * 1e3:8.1f} us', flush=True)


if __name__ == '__main__':
  ap = argparse.ArgumentParser()
  ap.add_argument('--nhyp', type=int, default=252)
  ap.add_argument('--reps', type=int, default=8)
  ap.add_argument('--dbg', default='0,1,8,128,9,129,136')
  ap.add_argument('--child', action='store_true')
  a = ap.parse_args()
  if a.child:
    one(a)
  else:
    for d in a.dbg.split(','):
      env = dict(os.environ, FPB_STEM_DBG=d)
      subprocess.run([sys.executable, __file__, '--child', '--nhyp', str(a.nhyp), '--reps', str(a.reps)], env=env)
