"""CPU ORACLE (test infrastructure, NOT product code) -- the whole register()/track_one() path on the host:
C restatement of the stencil / raster / warp arithmetic (oracle/fp_oracle.c) + fp32 torch-CPU networks
(oracle/nets.py).  reference: src/estimater.py:159-268, learning/training/predict_pose_refine.py:150-234,
learning/training/predict_score.py:161-212.  PARITY UNPINNED (see oracle/geometry.py header).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
import time

import numpy as np
import torch

from . import geometry as og
from . import nets as on
from . import ops as oo

S = 160


class OracleEstimator:
  def __init__(self, mesh_np, diameter, refine_sd, score_sd, symmetry_tfs=None, rot_normalizer=0.3490658503988659,
               refine_crop_ratio=1.2, score_crop_ratio=1.1, net_chunk=16):
    """mesh_np: dict of numpy arrays in make_mesh_tensors layout, already centred (src/estimater.py:44-51)."""
    self.mesh = mesh_np
    self.diameter = float(diameter)
    self.refine = on.build('refine', refine_sd)
    self.score = on.build('score', score_sd)
    self.rot_grid = og.make_rotation_grid(symmetry_tfs=symmetry_tfs)
    self.rot_normalizer = rot_normalizer
    self.rc, self.sc = refine_crop_ratio, score_crop_ratio
    self.net_chunk = net_chunk
    self.timings = {}
    self.trace = {}

  def preprocess(self, depth):
    return oo.bilateral_filter_depth(oo.erode_depth(depth, radius=2), radius=2)

  def refine_inputs(self, rgb, xyz_map, poses, K):
    H, W = rgb.shape[:2]
    cp = oo.crop_params(poses, K, self.diameter, self.rc, S, H, W)
    color, xyz, tri = oo.render(poses, cp, self.mesh, S)
    A = oo.make_A(color, xyz, poses, self.diameter, 0.001)
    B = oo.make_B(rgb.astype(np.float32), xyz_map, None, poses, cp, K, S, self.diameter, 0)
    return A, B, cp

  def score_inputs(self, rgb, depth, poses, K):
    H, W = rgb.shape[:2]
    cp = oo.crop_params(poses, K, self.diameter, self.sc, S, H, W)
    color, xyz, tri = oo.render(poses, cp, self.mesh, S)
    A = oo.make_A(color, xyz, poses, self.diameter, 0.1)
    B = oo.make_B(rgb.astype(np.float32), None, depth, poses, cp, K, S, self.diameter, 1)
    return A, B, cp

  def refine_predict(self, rgb, xyz_map, poses, K, iteration, keep_trace=False):
    poses = np.asarray(poses, np.float32).reshape(-1, 4, 4).copy()
    for it in range(iteration):
      t0 = time.perf_counter()
      A, B, _ = self.refine_inputs(rgb, xyz_map, poses, K)
      t1 = time.perf_counter()
      trans, rot = on.forward_chunked(self.refine, 'refine', A, B, self.net_chunk)
      t2 = time.perf_counter()
      new = oo.pose_update(poses, trans, rot, self.diameter, self.rot_normalizer)
      self.timings['refine_inputs'] = self.timings.get('refine_inputs', 0) + t1 - t0
      self.timings['refine_net'] = self.timings.get('refine_net', 0) + t2 - t1
      if keep_trace:
        self.trace.setdefault('refine', []).append(dict(poses_in=poses, trans=trans, rot=rot, poses_out=new))
      poses = new
    return poses

  def score_predict(self, rgb, depth, poses, K):
    t0 = time.perf_counter()
    A, B, _ = self.score_inputs(rgb, depth, poses, K)
    t1 = time.perf_counter()
    logits, feats = on.forward_chunked(self.score, 'score', A, B, self.net_chunk)
    self.timings['score_inputs'] = self.timings.get('score_inputs', 0) + t1 - t0
    self.timings['score_net'] = self.timings.get('score_net', 0) + time.perf_counter() - t1
    return logits + np.float32(100), feats

  def register(self, K, rgb, depth, ob_mask, iteration=5, n_hyp=None, keep_trace=False):
    """Returns (best pose [4,4] of the centred mesh, poses sorted, scores sorted).  n_hyp truncates the
    hypothesis set (bounded CPU samples)."""
    t0 = time.perf_counter()
    depth = self.preprocess(depth)
    valid = (depth >= 0.001) & (ob_mask > 0)
    center = og.guess_translation(depth, ob_mask, K)
    if valid.sum() < 4:
      pose = np.eye(4); pose[:3, 3] = center
      return pose, None, None
    poses = self.rot_grid.copy()
    if n_hyp is not None:
      poses = poses[:n_hyp]
    poses[:, :3, 3] = center.astype(np.float32).reshape(1, 3)
    xyz_map = oo.depth2xyzmap(depth, K)
    self.timings['preprocess'] = self.timings.get('preprocess', 0) + time.perf_counter() - t0
    poses = self.refine_predict(rgb, xyz_map, poses, K, iteration, keep_trace)
    scores, feats = self.score_predict(rgb, depth, poses, K)
    if keep_trace:
      self.trace['score'] = dict(poses=poses, scores=scores, feats=feats)
    ids = np.argsort(-scores, kind='stable')
    return poses[ids[0]], poses[ids], scores[ids]

  def track_one(self, rgb, depth, K, pose_last, iteration):
    depth = self.preprocess(depth)
    xyz_map = oo.depth2xyzmap_batch(depth[None], np.asarray(K, np.float32)[None], np.inf)[0]
    return self.refine_predict(rgb, xyz_map, np.asarray(pose_last).reshape(1, 4, 4), K, iteration)[0]
