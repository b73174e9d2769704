"""GPU parity of the 8(f) rows: ADD / ADD-S kernels against the oracle restatement of src/Utils.py:232-253 (numpy +
cKDTree), symmetry-aware clustering with symmetry_tfs_from_info output, and the pinned frame streamer on the device."""
import numpy as np
import pytest
import torch

from oracle import geometry as og
from foundationpose_b200 import utils as U

pytestmark = pytest.mark.gpu


def _poses(n, seed):
  rng = np.random.default_rng(seed)
  out = []
  for _ in range(n):
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    w, x, y, z = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    T = np.eye(4); T[:3, :3] = R; T[:3, 3] = rng.normal(size=3) * 0.1 + [0, 0, 0.6]
    out.append(T)
  return np.stack(out)


@pytest.mark.parametrize('P', [1, 255, 1000, 2049])
def test_add_and_adds_match_oracle(P):
  rng = np.random.default_rng(P)
  pts = (rng.random((P, 3)) - 0.5) * [0.1, 0.2, 0.05]
  pred, gt = _poses(7, 1), _poses(7, 2)
  pred[0] = gt[0]                                              # identical pose: both errors are exactly 0
  a = U.add_err(pred, gt, pts).cpu().numpy()
  s = U.adds_err(pred, gt, pts).cpu().numpy()
  a_ref = np.array([og.add_err(p, g, pts) for p, g in zip(pred, gt)])
  s_ref = np.array([og.adds_err(p, g, pts) for p, g in zip(pred, gt)])
  assert a[0] == 0 and s[0] == 0
  assert np.allclose(a, a_ref, rtol=2e-5, atol=1e-7), np.abs(a - a_ref).max()
  assert np.allclose(s, s_ref, rtol=2e-5, atol=1e-7), np.abs(s - s_ref).max()
  assert (s <= a + 1e-7).all()                                 # ADD-S <= ADD by construction
  # single pose API returns a float; one gt for the whole batch broadcasts
  assert abs(U.add_err(pred[1], gt[1], pts) - a_ref[1]) < 1e-6
  b = U.adds_err(pred, gt[3], pts).cpu().numpy()
  assert abs(b[3] - og.adds_err(pred[3], gt[3], pts)) < 1e-6


def test_adds_is_symmetry_invariant_where_add_is_not():
  """A cylinder-like point set rotated about its symmetry axis: ADD-S ~ 0 (up to sampling), ADD large."""
  t = np.linspace(0, 2 * np.pi, 360, endpoint=False)
  pts = np.stack([0.05 * np.cos(t), 0.05 * np.sin(t), np.zeros_like(t)], 1)
  pts = np.concatenate([pts + [0, 0, z] for z in np.linspace(-0.1, 0.1, 9)])
  gt = np.eye(4); gt[2, 3] = 0.5
  sym = U.symmetry_tfs_from_info({'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]}, rot_angle_discrete=5)
  pred = gt @ sym[10]
  assert U.adds_err(pred, gt, pts) < 1e-6
  assert U.add_err(pred, gt, pts) > 0.02


def test_cluster_poses_with_symmetry_tfs_from_info():
  """cluster_poses(symmetry_tfs) merges hypotheses that differ by a model symmetry (src/estimater.py:73-76,128-136)."""
  from foundationpose_b200 import mycpp
  sym = U.symmetry_tfs_from_info({'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]}, rot_angle_discrete=30)
  base = _poses(3, 5)
  hyps = np.concatenate([[b @ sym[k] for k in (1, 4, 7)] for b in base])      # 3 symmetric copies of 3 poses
  kept_sym = mycpp.cluster_poses(10, 99999, hyps.astype(np.float32), sym)
  kept_plain = mycpp.cluster_poses(10, 99999, hyps.astype(np.float32), np.eye(4)[None])
  assert len(kept_sym) == 3 and len(kept_plain) == 9


def test_frame_streamer_device(tmp_path):
  from tests.test_io_cpu import _write_video
  from foundationpose_b200.datareader import FrameStreamer, YcbineoatReader
  d = str(tmp_path / 'v')
  frames = _write_video(d, n=6)
  r = YcbineoatReader(d)
  assert r.get_xyz_map(0).shape == (48, 64, 3)
  n = 0
  for i, rgb, dep in FrameStreamer(r):
    assert rgb.is_cuda and dep.is_cuda and rgb.dtype == torch.uint8 and dep.dtype == torch.float32
    assert np.array_equal(rgb.cpu().numpy(), frames[i][0])
    assert np.allclose(dep.cpu().numpy(), r.get_depth(i).astype(np.float32))
    n += 1
  assert n == 6
