#!/usr/bin/env python
"""Per-frame latency of the video path (BASELINE config 3): register() on frame 0, then track_one(iteration=2) on a
stream of 640x480 RGB-D frames held in pinned host memory (H2D of the frame and D2H of the pose inside the timed call).

  python tools/track_bench.py [--frames 200] [--iters 2] [--graph 1]

--graph 1 additionally replays the frame body (stencils, xyz map, 2 x [crop, render, warp, RefineNet, pose update])
as one captured CUDA graph over static buffers (FoundationPose.track_one_graphed) and reports both latencies."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import make_estimator  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--frames', type=int, default=200)
  ap.add_argument('--iters', type=int, default=2)
  ap.add_argument('--graph', type=int, default=1)
  args = ap.parse_args()
  scene = synthetic.make_scene(2)
  est, _, _ = make_estimator(scene, 'bf16')
  K = scene['K']
  h_rgb = torch.from_numpy(scene['rgb']).pin_memory()
  h_depth = torch.from_numpy(scene['depth'].astype(np.float32)).pin_memory()
  est.register(K=K, rgb=scene['rgb'], depth=scene['depth'], ob_mask=scene['mask'], iteration=5)
  pose0 = est.pose_last.clone()
  out = {}
  modes = [('eager', est.track_one)]
  if args.graph and hasattr(est, 'track_one_graphed'):
    modes.append(('graph', est.track_one_graphed))
  for name, fn in modes:
    est.pose_last = pose0.clone()
    for _ in range(10):
      fn(rgb=h_rgb, depth=h_depth, K=K, iteration=args.iters)
    lat = []
    for _ in range(args.frames):
      torch.cuda.synchronize()
      t0 = time.perf_counter()
      pose = fn(rgb=h_rgb, depth=h_depth, K=K, iteration=args.iters)
      lat.append((time.perf_counter() - t0) * 1e3)
    lat = np.array(lat)
    out[name] = dict(p50_ms=float(np.percentile(lat, 50)), p95_ms=float(np.percentile(lat, 95)), min_ms=float(lat.min()),
                     pose_t=[float(x) for x in np.asarray(pose)[:3, 3]])
  print(json.dumps(dict(workload='cfg3: track_one, 1 hypothesis, %d refine iterations, 640x480 host frames' % args.iters,
                        frames=args.frames, **out)))


if __name__ == '__main__':
  main()
