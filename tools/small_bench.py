#!/usr/bin/env python
"""CUDA-event timing of the small rank-redundant kernels of register(): guess_translation statistics (cluster kernel vs
FPB_MASK_STATS_1CTA=1) and ScoreNet's cross-hypothesis attention at L = 252.   python tools/small_bench.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from foundationpose_b200 import nets, synthetic  # noqa: E402
from foundationpose_b200._lib import lib, check, ptr, stream  # noqa: E402


def timed(fn, reps=50):
  for _ in range(5):
    fn()
  ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
  for a, b in ev:
    a.record(); fn(); b.record()
  torch.cuda.synchronize()
  return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3


def main():
  sc = synthetic.make_scene(2, n_points=300000)
  depth = torch.as_tensor(sc['depth']).cuda()
  mask = torch.as_tensor(sc['mask']).cuda().view(torch.uint8)
  oi = torch.empty(6, dtype=torch.int32, device='cuda'); of = torch.empty(2, dtype=torch.float, device='cuda')
  out = {}
  out['mask_stats_us'] = timed(lambda: check(lib().fpb_mask_depth_stats(ptr(depth), ptr(mask), 480, 640, ptr(oi), ptr(of), stream())))
  out['mask_stats_variant'] = '1cta' if os.environ.get('FPB_MASK_STATS_1CTA') == '1' else 'cluster8'
  out['stats'] = oi.cpu().tolist() + of.cpu().tolist()
  net = nets.Net(nets.SCORE, nets.synthetic_state_dict(nets.SCORE), 'fp16')
  feats = torch.randn(252, 512, device='cuda')
  out['score_cross_252_us'] = timed(lambda: net.score_cross(feats, 252, 1))
  print(json.dumps(out))


if __name__ == '__main__':
  main()
