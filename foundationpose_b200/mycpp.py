"""Drop-in for the reference's pybind module `mycpp` (mycpp/src/app/pybind_api.cpp:24-76):
`cluster_poses(angle_diff, dist_diff, poses_in, symmetry_tfs) -> list of float32 4x4`, evaluated by a CUDA
kernel that preserves the sequential greedy semantics (kept set identical to the C++ loop)."""
import numpy as np
import torch

from . import _lib
from ._lib import lib, check, ptr, stream

_last_keep = None


def cluster_poses(angle_diff, dist_diff, poses_in, symmetry_tfs):
  global _last_keep
  _lib.require_cuda()
  poses = np.ascontiguousarray(np.asarray(poses_in, dtype=np.float64).reshape(-1, 4, 4), dtype=np.float32)
  sym = np.ascontiguousarray(np.asarray(symmetry_tfs, dtype=np.float64).reshape(-1, 4, 4), dtype=np.float32)
  if len(poses) == 0:
    raise IndexError('cluster_poses: poses_in is empty')   # the reference dereferences poses_in[0] unguarded
  print(f'num original candidates = {len(poses)}')
  p = torch.from_numpy(poses).cuda()
  s = torch.from_numpy(sym).cuda()
  keep = torch.empty(len(poses), dtype=torch.int32, device='cuda')
  nk = torch.empty(1, dtype=torch.int32, device='cuda')
  check(lib().fpb_cluster_poses_f32(ptr(p), len(poses), ptr(s), len(sym), float(angle_diff), float(dist_diff),
                                    ptr(keep), ptr(nk), stream()))
  n = int(nk.item())
  idx = keep[:n].cpu().numpy()
  _last_keep = idx
  print(f'num of pose after clustering: {n}')
  return [poses[i].copy() for i in idx]


def last_keep_indices():
  """Indices (into poses_in) kept by the most recent cluster_poses call."""
  return _last_keep
