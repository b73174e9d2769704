#!/usr/bin/env python
"""register() latency of BASELINE config 4: bent-sheet CAD part (251,001 vertices / 500,000 faces), 1920x1200 frame,
1008 hypotheses (42 views x 24 in-plane steps of 15 degrees, unclustered as SURVEY 8(d) prescribes), 5 refine iterations.

  python tools/cfg_bench.py [--steps 5]                                   one GPU
  torchrun --nproc-per-node N tools/cfg_bench.py                          hypotheses sharded over N GPUs (dist.register_sharded)

Prints one JSON line on rank 0: register p50 (host call, synchronous), hypotheses/s and the per-stage device times."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from bench import make_estimator  # noqa: E402
from foundationpose_b200 import synthetic, dist as fdist  # noqa: E402
from foundationpose_b200.estimater import make_rotation_grid_host  # noqa: E402


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--steps', type=int, default=5)
  ap.add_argument('--iters', type=int, default=5)
  ap.add_argument('--precision', default='fp16', choices=['fp16', 'bf16'])
  args = ap.parse_args()
  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  torch.cuda.set_device(local)
  if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  sc = synthetic.make_scene(4, n_points=400000)
  est, _, _ = make_estimator(sc, args.precision)
  grid = make_rotation_grid_host(min_n_views=40, inplane_step=15)
  est.rot_grid = torch.as_tensor(grid, device='cuda', dtype=torch.float)
  n = len(grid)
  rgb, depth, mask = (torch.as_tensor(sc[k]).cuda() for k in ('rgb', 'depth', 'mask'))

  def step():
    if world > 1:
      return fdist.register_sharded(est, sc['K'], rgb, depth, mask, iteration=args.iters)
    return est.register(K=sc['K'], rgb=rgb, depth=depth, ob_mask=mask, iteration=args.iters)

  for _ in range(2):
    step()
  torch.cuda.synchronize()
  if world > 1:
    dist.barrier()
  per = []
  for _ in range(args.steps):
    t0 = time.perf_counter()
    step()
    per.append((time.perf_counter() - t0) * 1e3)
  p50 = float(np.median(per))
  if world > 1:
    t = torch.tensor([p50], device='cuda', dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    p50 = float(t[0])
  match = None
  if world > 1:                       # the sharded result against the unsharded register() on every rank
    ps = step()
    bs = int(est.best_id)
    pu = est.register(K=sc['K'], rgb=rgb, depth=depth, ob_mask=mask, iteration=args.iters)
    ok = torch.tensor([1.0 if (np.abs(ps - pu).max() <= 1e-4 and bs == int(est.best_id)) else 0.0], device='cuda')
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    match = bool(ok[0] > 0.5)
  if rank == 0:
    V, F = sc['mesh'].vertices.shape[0], sc['mesh'].faces.shape[0]
    print(json.dumps(dict(workload=f'cfg4: {V} vertices / {F} faces, {sc["W"]}x{sc["H"]}, {n} hypotheses, {args.iters} refine iterations',
                          n_gpus=world, precision=args.precision, matches_unsharded=match, register_ms_p50=p50, hypotheses_per_s=n / (p50 / 1e3), steps=args.steps,
                          sharding='hypotheses (dist.register_sharded, p2p exchange)' if world > 1 else 'none')))
  if world > 1:
    dist.destroy_process_group()


if __name__ == '__main__':
  main()
