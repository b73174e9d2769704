"""CPU tests of the 8(f) host pieces: PoseTransformer (src/transform.py), symmetry_tfs_from_info (src/Utils.py:806-834),
the YcbInEOAT directory reader + frame streamer (src/datareader.py:85-180), oriented bounds / headless main flow."""
import math
import os

import numpy as np
import pytest

from foundationpose_b200 import utils as U
from foundationpose_b200.datareader import FrameStreamer, YcbineoatReader, track_video
from foundationpose_b200.transform import METERS_TO_INCHES, PoseTransformer, oriented_bounds, project_box_edges


def _euler_zyx(roll, pitch, yaw):
  cr, sr, cp, sp, cy, sy = math.cos(roll), math.sin(roll), math.cos(pitch), math.sin(pitch), math.cos(yaw), math.sin(yaw)
  Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
  Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
  Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
  return Rz @ Ry @ Rx


def test_pose_transformer_units_and_round_trip():
  P = np.eye(4)
  P[:3, :3] = _euler_zyx(0.3, -0.4, 1.1)
  P[:3, 3] = [0.1, -0.2, 0.5]
  x, y, z, r, p, yw = PoseTransformer().transform_pose(P)                 # defaults: inches, degrees (src/transform.py:7)
  assert np.allclose([x, y, z], np.array([0.1, -0.2, 0.5]) * 39.3701)
  assert METERS_TO_INCHES == 39.3701
  assert np.allclose([r, p, yw], np.degrees([0.3, -0.4, 1.1]), atol=1e-9)
  x, y, z, r, p, yw = PoseTransformer(to_inches=False, to_degrees=False).transform_pose(P)
  assert np.allclose([x, y, z, r, p, yw], [0.1, -0.2, 0.5, 0.3, -0.4, 1.1], atol=1e-12)


@pytest.mark.parametrize('sign', [1, -1])
def test_pose_transformer_gimbal_lock_branch(sign):
  """|R[2,0]| > 0.9999: yaw = 0, pitch = -+pi/2, roll from R[0,1], R[1,1] (src/transform.py:52-60)."""
  R = _euler_zyx(0.7, sign * math.pi / 2, 0.0)
  roll, pitch, yaw = PoseTransformer(False, False)._rotation_matrix_to_euler_angles(R)
  assert yaw == 0
  assert math.isclose(pitch, sign * math.pi / 2)
  exp = math.atan2(R[0, 1], R[1, 1]) if R[2, 0] < 0 else -math.atan2(R[0, 1], R[1, 1])
  assert math.isclose(roll, exp)


def test_symmetry_tfs_from_info():
  assert U.symmetry_tfs_from_info({}).shape == (1, 4, 4)
  T = np.eye(4); T[:3, :3] = _euler_zyx(0, 0, math.pi); T[:3, 3] = [10, 20, 30]
  s = U.symmetry_tfs_from_info({'symmetries_discrete': [T.reshape(-1).tolist()]})
  assert s.shape == (2, 4, 4) and np.allclose(s[0], np.eye(4))
  assert np.allclose(s[1, :3, 3], [0.01, 0.02, 0.03]) and np.allclose(s[1, :3, :3], T[:3, :3])
  s = U.symmetry_tfs_from_info({'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]}, rot_angle_discrete=90)
  assert s.shape == (5, 4, 4)                                   # identity + 4 samples (0 deg appears twice, as in the reference)
  assert np.allclose(s[2, :3, :3], _euler_zyx(0, 0, math.pi / 2), atol=1e-12)
  for m in s:
    assert np.allclose(m[:3, :3] @ m[:3, :3].T, np.eye(3), atol=1e-12)


def _write_video(d, n=4, H=48, W=64):
  import cv2
  for sub in ('rgb', 'depth', 'masks', 'annotated_poses'):
    os.makedirs(os.path.join(d, sub), exist_ok=True)
  rng = np.random.default_rng(0)
  frames = []
  for i in range(n):
    rgb = rng.integers(0, 255, (H, W, 3), dtype=np.uint8)
    dep = rng.integers(0, 3000, (H, W)).astype(np.uint16)
    mask = np.zeros((H, W, 3), np.uint8); mask[10:30, 20:40, 1] = 255
    cv2.imwrite(f'{d}/rgb/{i:06d}.png', rgb[..., ::-1])
    cv2.imwrite(f'{d}/depth/{i:06d}.png', dep)
    cv2.imwrite(f'{d}/masks/{i:06d}.png', mask)
    np.savetxt(f'{d}/annotated_poses/{i:06d}.txt', np.eye(4) * (i + 1))
    frames.append((rgb, dep))
  np.savetxt(f'{d}/cam_K.txt', np.array([[500.0, 0, 32], [0, 500.0, 24], [0, 0, 1]]))
  return frames


def test_ycbineoat_reader(tmp_path):
  d = str(tmp_path / 'mustard0')
  frames = _write_video(d)
  r = YcbineoatReader(d, zfar=2.5)
  assert len(r) == 4 and r.id_strs[2] == '000002' and (r.H, r.W) == (48, 64) and r.get_video_name() == 'mustard0'
  assert np.array_equal(r.get_color(1), frames[1][0])
  dep = r.get_depth(1)
  exp = frames[1][1] / 1e3
  exp[(exp < 0.001) | (exp >= 2.5)] = 0
  assert np.array_equal(dep, exp)
  m = r.get_mask(0)
  assert m.dtype == np.uint8 and m.sum() == 20 * 20 and m[15, 25] == 1
  assert np.allclose(r.get_gt_pose(2), np.eye(4) * 3) and r.get_gt_pose(99) is None
  assert r.get_occ_mask(0).sum() == 0
  half = YcbineoatReader(d, shorter_side=24)
  assert (half.H, half.W) == (24, 32) and np.allclose(half.K[:2], r.K[:2] * 0.5)
  assert half.get_color(0).shape == (24, 32, 3) and half.get_depth(0).shape == (24, 32)


def test_frame_streamer_host_order_and_content(tmp_path):
  d = str(tmp_path / 'v')
  frames = _write_video(d, n=5)
  r = YcbineoatReader(d)
  seen = []
  for i, rgb, dep in FrameStreamer(r, device=None):
    seen.append(i)
    assert np.array_equal(rgb.numpy(), frames[i][0])
    assert np.allclose(dep.numpy(), r.get_depth(i).astype(np.float32))
  assert seen == [0, 1, 2, 3, 4]
  assert [i for i, _, _ in FrameStreamer(r, start=1, stop=3, device=None)] == [1, 2]


def test_track_video_calls_register_then_track(tmp_path):
  d = str(tmp_path / 'v')
  _write_video(d, n=3)

  class Est:
    calls = []

    def register(self, K, rgb, depth, ob_mask, iteration):
      self.calls.append(('register', iteration, int(ob_mask.sum())))
      return np.eye(4)

    def track_one(self, rgb, depth, K, iteration):
      self.calls.append(('track', iteration))
      return 2 * np.eye(4)

  est = Est()
  poses = track_video(est, YcbineoatReader(d), device=None)
  assert est.calls == [('register', 5, 400), ('track', 2), ('track', 2)]
  assert poses.shape == (3, 4, 4) and poses[1, 0, 0] == 2


def test_oriented_bounds_of_a_rotated_box():
  rng = np.random.default_rng(1)
  ext = np.array([0.05, 0.12, 0.30])
  pts = (rng.random((400, 3)) - 0.5) * ext
  corners = np.array([[sx, sy, sz] for sx in (-.5, .5) for sy in (-.5, .5) for sz in (-.5, .5)]) * ext
  pts = np.concatenate([pts, corners])
  R = _euler_zyx(0.4, -0.7, 1.3)
  V = pts @ R.T + np.array([0.3, -0.1, 0.9])
  T, e = oriented_bounds(V)
  assert np.allclose(e, ext, atol=1e-9)                          # ascending extents of the true box
  assert np.allclose(T[:3, :3] @ T[:3, :3].T, np.eye(3), atol=1e-12) and np.linalg.det(T[:3, :3]) > 0
  Vc = V @ T[:3, :3].T + T[:3, 3]
  assert np.allclose(Vc.max(0), e / 2, atol=1e-9) and np.allclose(Vc.min(0), -e / 2, atol=1e-9)
  edges = project_box_edges(np.array([[500.0, 0, 320], [0, 500.0, 240], [0, 0, 1]]), np.linalg.inv(T), np.stack([-e / 2, e / 2]))
  assert edges.shape == (12, 2, 2) and np.isfinite(edges).all()
