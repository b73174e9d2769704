"""CPU tests of the oracle itself (no GPU): internal consistency, known answers, edge cases."""
import numpy as np
import pytest

from oracle import geometry as og
from oracle import ops as oo


def test_icosphere_counts_and_unit_norm():
  for sub, (nv, nf) in {0: (12, 20), 1: (42, 80), 2: (162, 320)}.items():
    v, f = og.icosphere(sub)
    assert v.shape == (nv, 3) and f.shape == (nf, 3)
    assert np.allclose(np.linalg.norm(v, axis=1), 1.0, atol=1e-12)
    assert f.min() == 0 and f.max() == nv - 1


def test_rotation_grid_252_and_rigid():
  g = og.make_rotation_grid_f64()
  assert g.shape == (252, 4, 4)
  R = g[:, :3, :3]
  assert np.allclose(R @ R.transpose(0, 2, 1), np.eye(3)[None], atol=1e-12)
  assert np.allclose(np.linalg.det(R), 1.0, atol=1e-12)
  # camera sits at distance 1 looking at the origin: object origin is at (0,0,1) in camera frame
  assert np.allclose(g[:, :3, 3], np.array([0, 0, 1.0])[None], atol=1e-12)


def test_cluster_poses_reference_probe():
  """all 252 survive 30 deg / identity symmetry; the 42x24 grid is cut to 384."""
  g = og.make_rotation_grid_f64()
  kept, idx = og.cluster_poses(30, 99999, g, np.eye(4)[None])
  assert len(kept) == 252 and (idx == np.arange(252)).all()
  g24 = og.make_rotation_grid_f64(inplane_step=15)
  assert len(g24) == 1008
  kept, idx = og.cluster_poses(30, 99999, g24, np.eye(4)[None])
  assert len(kept) == 384
  kept10, _ = og.cluster_poses(10, 99999, g24, np.eye(4)[None])
  assert len(kept10) == 1008


def test_cluster_poses_symmetry_and_distance():
  g = og.make_rotation_grid_f64()
  rz180 = og.euler_matrix_z(np.pi)
  kept, _ = og.cluster_poses(30, 99999, g, np.stack([np.eye(4), rz180]))
  assert len(kept) < 252
  far = g.copy(); far[:, 0, 3] = np.arange(len(g)) * 10.0
  kept, _ = og.cluster_poses(180, 1.0, far, np.eye(4)[None])   # translations differ by >= dist: nothing merges
  assert len(kept) == 252
  kept, _ = og.cluster_poses(181, 99999, g, np.eye(4)[None])   # everything within 181 deg of pose 0
  assert len(kept) == 1


def test_erode_and_bilateral_edge_cases():
  z = np.zeros((7, 9), np.float32)
  assert (oo.erode_depth(z) == 0).all() and (oo.bilateral_filter_depth(z) == 0).all()
  c = np.full((12, 10), 0.5, np.float32)
  assert (oo.erode_depth(c) == c).all()
  assert np.allclose(oo.bilateral_filter_depth(c), 0.5, atol=1e-6)
  c2 = c.copy(); c2[5, 5] = 0.9     # isolated outlier: all 24 neighbours differ -> eroded
  assert oo.erode_depth(c2)[5, 5] == 0
  far = np.full((6, 6), 150.0, np.float32)   # >= zfar is invalid
  assert (oo.erode_depth(far) == 0).all() and (oo.bilateral_filter_depth(far) == 0).all()
  one = np.array([[0.7]], np.float32)        # 1x1 image: only the centre tap is in bounds
  assert oo.erode_depth(one)[0, 0] == np.float32(0.7)


def test_depth2xyz_matches_numpy_formula():
  rng = np.random.default_rng(0)
  d = rng.uniform(0, 2, (48, 64)).astype(np.float32); d[d < 0.2] = 0
  K = np.array([[600.0, 0, 31.5], [0, 590.0, 24.2], [0, 0, 1]])
  vs, us = np.meshgrid(np.arange(48), np.arange(64), indexing='ij')
  ref = np.stack([(us - K[0, 2]) * d / K[0, 0], (vs - K[1, 2]) * d / K[1, 1], d], -1).astype(np.float32)
  ref[d < 0.001] = 0
  assert np.array_equal(oo.depth2xyzmap(d, K), ref)
  K32 = K.astype(np.float32)
  ref32 = np.stack([(us.astype(np.float32) - K32[0, 2]) * d / K32[0, 0], (vs.astype(np.float32) - K32[1, 2]) * d / K32[1, 1], d], -1)
  ref32[d < 0.001] = 0
  assert np.array_equal(oo.depth2xyzmap_batch(d[None], K32[None], np.inf)[0], ref32)


def test_guess_translation_edge_cases():
  K = np.array([[600.0, 0, 320], [0, 600.0, 240], [0, 0, 1]])
  d = np.full((480, 640), 0.6, np.float32)
  m = np.zeros((480, 640), bool)
  assert (og.guess_translation(d, m, K) == 0).all()
  m[100:201, 300:341] = True
  t = og.guess_translation(d, m, K)
  assert np.allclose(t, [0, -0.09, 0.6], atol=1e-6)
  assert (og.guess_translation(np.zeros_like(d), m, K) == 0).all()


def test_crop_params_window_is_square_and_centered():
  poses = np.tile(np.eye(4, dtype=np.float32)[None], (3, 1, 1))
  poses[:, :3, 3] = [[0, 0, 0.6], [0.05, -0.03, 0.7], [-0.1, 0.08, 0.5]]
  K = np.array([[600.0, 0, 320], [0, 600.0, 240], [0, 0, 1]])
  cp = oo.crop_params(poses, K, 0.2, 1.2, 160, 480, 640)
  tf = cp['tf'].reshape(-1, 3, 3)
  for n in range(3):
    c = K @ poses[n, :3, 3]; c = c[:2] / c[2]
    crop_c = tf[n] @ np.array([c[0], c[1], 1.0])
    assert abs(crop_c[0] - 80) < 1.0 and abs(crop_c[1] - 80) < 1.0
    assert abs(tf[n, 0, 0] - tf[n, 1, 1]) < 1e-6
    inv = np.linalg.inv(tf[n].astype(np.float64))
    assert np.allclose([inv[0, 0], inv[0, 2], inv[1, 1], inv[1, 2]], cp['inv'][n], rtol=1e-5, atol=1e-3)


def test_render_cube_covers_and_is_consistent(scene1):
  from tests.helpers import mesh_np, centered, hypo_poses
  mesh, _ = centered(scene1)
  poses, _ = hypo_poses(scene1, n=4)
  cp = oo.crop_params(poses, scene1['K'], scene1['diameter'], 1.2, 160, 480, 640)
  color, xyz, tri = oo.render(poses, cp, mesh_np(mesh), 160)
  cov = (tri > 0).mean(axis=(1, 2))
  assert (cov > 0.15).all() and (cov < 0.8).all()
  assert tri.max() <= 12
  z = xyz[..., 2]
  assert (z[tri > 0] > 0.4).all() and (z[tri == 0] == 0).all()
  assert color[tri == 0].max() == 0 and color.max() <= 1.0
  # xyz of covered pixels re-projects to the pixel centre through the crop window
  n = 0
  tf = cp['tf'][n].reshape(3, 3).astype(np.float64)
  K = scene1['K']
  ii, jj = np.nonzero(tri[n] > 0)
  p = xyz[n, ii, jj].astype(np.float64)
  uv = (K @ p.T).T; uv = uv[:, :2] / uv[:, 2:3]
  # rendered window quirk: u = l + (r-l)(px+0.5)/S
  l, t, r, b = cp['bbox'][n]
  assert np.allclose(uv[:, 0], l + (r - l) * (jj + 0.5) / 160, atol=0.02)
  assert np.allclose(uv[:, 1], t + (b - t) * (ii + 0.5) / 160, atol=0.02)


def test_so3_exp_and_pose_update():
  r = np.array([[0, 0, np.pi / 2], [0.1, -0.2, 0.05], [0, 0, 0]], np.float32)
  R = oo.so3_exp_map(r)
  assert np.allclose(R @ R.transpose(0, 2, 1), np.eye(3)[None], atol=1e-5)
  assert np.allclose(R[0], [[0, -1, 0], [1, 0, 0], [0, 0, 1]], atol=1e-6)
  poses = np.tile(np.eye(4, dtype=np.float32)[None], (3, 1, 1))
  out = oo.pose_update(poses, np.ones((3, 3), np.float32) * 0.1, np.zeros((3, 3), np.float32), 0.2, 0.349)
  assert np.allclose(out[:, :3, 3], 0.01, atol=1e-7)
  assert np.allclose(out[:, :3, :3], np.eye(3)[None], atol=1e-6)


def test_pose_update_deepim_and_6d_known_answers():
  """Config-gated representations (predict_pose_refine.py:201-224): a zero image-plane shift with depth ratio 1 leaves
  the translation unchanged; the 6-D rotation of (e1, e2) is the identity; Gram-Schmidt output is a rotation."""
  g = og.make_rotation_grid()[:16].copy()
  g[:, :3, 3] = np.array([0.03, -0.02, 0.6], np.float32)
  K = np.array([[600, 0, 320], [0, 600, 240], [0, 0, 1.0]])
  cp = oo.crop_params(g, K, 0.2, 1.2, 160, 480, 640)
  tf = cp['tf'].reshape(-1, 3, 3)
  trans = np.zeros((16, 3), np.float32); trans[:, 2] = 1.0
  rot6 = np.tile(np.array([1, 0, 0, 0, 1, 0], np.float32), (16, 1))
  out = oo.pose_update(g, trans, rot6, 0.2, 0.349, [0.02, 0.02, 0.05], False, trans_rep='deepim', rot_rep='6d',
                       tf_to_crops=tf, K=K, S=160)
  assert np.abs(out - g).max() < 2e-6
  # a pure depth change moves the object along its viewing ray
  trans[:, 2] = 1.1
  out = oo.pose_update(g, trans, rot6, 0.2, 0.349, [0.02, 0.02, 0.05], False, trans_rep='deepim', rot_rep='6d',
                       tf_to_crops=tf, K=K, S=160)
  assert np.allclose(out[:, :3, 3], g[:, :3, 3] * 1.1, atol=1e-5)
  rng = np.random.default_rng(0)
  R = oo.rotation_6d_to_matrix(rng.normal(size=(64, 6)).astype(np.float32)).astype(np.float64)
  assert np.abs(R @ R.transpose(0, 2, 1) - np.eye(3)).max() < 1e-5 and np.allclose(np.linalg.det(R), 1, atol=1e-5)
  with pytest.raises(RuntimeError):
    oo.pose_update(g, trans, rot6, 0.2, 0.349, None, True, rot_rep='quat')
