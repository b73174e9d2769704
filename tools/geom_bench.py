#!/usr/bin/env python
"""CUDA-event timing of the geometry stage at the bench shapes (cfg 2: 252 hypotheses, 20,480 faces, 160x160 crops):
crop windows, render (z-buffer clear + vertex + raster + shade -> network input A) and crop/warp (-> input B), each
call timed on its own with the previous call's outputs still in L2 (as inside register()), plus their roofline figures.

  python tools/geom_bench.py [--reps 50]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import peaks  # noqa: E402
from foundationpose_b200 import synthetic, utils as U  # noqa: E402
from foundationpose_b200.estimater import make_rotation_grid_host  # noqa: E402


def timed(fn, reps):
  for _ in range(5):
    fn()
  ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
  for a, b in ev:
    a.record(); fn(); b.record()
  torch.cuda.synchronize()
  return float(np.median([a.elapsed_time(b) for a, b in ev])) * 1e3    # us


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--reps', type=int, default=50)
  ap.add_argument('--cfg', type=int, default=2)
  ap.add_argument('--inplane', type=int, default=60, help='in-plane rotation step in degrees (15 -> 1008 hypotheses)')
  args = ap.parse_args()
  sc = synthetic.make_scene(args.cfg, n_points=300000)
  mesh = sc['mesh'].copy()
  center = (mesh.vertices.min(0) + mesh.vertices.max(0)) / 2
  mesh.vertices = mesh.vertices - center.reshape(1, 3)      # as FoundationPose.reset_object (src/estimater.py:44-51)
  mt = U.make_mesh_tensors(mesh)
  S, H, W = 160, sc['H'], sc['W']
  poses = make_rotation_grid_host(40, args.inplane).astype(np.float32)
  poses[:, :3, 3] = (sc['gt_pose'][:3, :3] @ center + sc['gt_pose'][:3, 3]).astype(np.float32)
  N = len(poses)
  poses_t = torch.as_tensor(poses, device='cuda')
  rgb = torch.as_tensor(sc['rgb'], device='cuda')
  depth = torch.as_tensor(sc['depth'].astype(np.float32), device='cuda')
  xyz = U.depth2xyzmap_cuda(depth, sc['K'], math_f64=False)
  AB = U.net_input_buffer(N, S, torch.bfloat16)
  params = U.crop_params(poses_t, sc['K'], sc['diameter'], 1.2, S, H, W)
  V, F = mt['pos'].shape[0], mt['faces'].shape[0]
  out = dict(N=N, V=V, F=F, S=S)
  out['crop_params_us'] = timed(lambda: U.crop_params(poses_t, sc['K'], sc['diameter'], 1.2, S, H, W), args.reps)
  out['render_us'] = timed(lambda: U.render_crops(mt, poses_t, params, S, sc['diameter'], 0.001, out_A=AB[:N]), args.reps)
  out['warp_refine_us'] = timed(lambda: U.warp_crops(rgb, xyz, depth, poses_t, params, sc['K'], S, sc['diameter'], 0,
                                                    out_B=AB[N:]), args.reps)
  out['warp_score_us'] = timed(lambda: U.warp_crops(rgb, xyz, depth, poses_t, params, sc['K'], S, sc['diameter'], 1,
                                                   out_B=AB[N:]), args.reps)
  pk = peaks()
  # algorithmic bytes: render V*36 + F*12 + S^2*6*e per hypothesis; warp S^2*6*e per hypothesis + H*W*24 per pass
  e = 2
  rb = N * (V * 36 + F * 12 + S * S * 6 * e)
  wb = N * S * S * 6 * e + H * W * 24
  out['render_alg_GBs'] = rb / out['render_us'] / 1e3
  out['warp_alg_GBs'] = wb / out['warp_refine_us'] / 1e3
  out['render_frac_hbm'] = out['render_alg_GBs'] / pk['hbm']
  out['warp_frac_hbm'] = out['warp_alg_GBs'] / pk['hbm']
  # bytes the kernels actually store (8 channels x bf16 per pixel incl. the two zero channels)
  out['warp_store_GBs'] = N * S * S * 16 / out['warp_refine_us'] / 1e3
  print(json.dumps(out))


if __name__ == '__main__':
  main()
