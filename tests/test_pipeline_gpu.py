"""GPU: the README pipeline processor end to end, and the largest BASELINE configuration (cfg 4: 251 k-vertex mesh,
1920x1200 frame, 1008 hypotheses = 4 network chunks) as a maximum-size robustness case."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _estimator(scene, precision='fp16'):
  from foundationpose_b200 import nets
  from foundationpose_b200.estimater import FoundationPose
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from foundationpose_b200.predict_score import ScorePredictor
  sd_r, sd_s = nets.synthetic_state_dict(nets.REFINE), nets.synthetic_state_dict(nets.SCORE)
  return FoundationPose(model_pts=scene['mesh'].vertices, model_normals=scene['mesh'].vertex_normals, mesh=scene['mesh'],
                        refiner=PoseRefinePredictor(state_dict=sd_r, precision=precision),
                        scorer=ScorePredictor(state_dict=sd_s, precision=precision), mesh_diameter=scene['diameter'])


def test_pipeline_processor_equals_direct_register(scene2):
  from foundationpose_b200 import pipeline as P
  sc = scene2
  est = _estimator(sc)
  est.rot_grid = est.rot_grid[::21].contiguous()                 # 12 hypotheses
  direct = est.register(K=sc['K'], rgb=sc['rgb'], depth=sc['depth'], ob_mask=sc['mask'], iteration=2)
  proc = P.FoundationPoseEstimator(mesh_file=sc['mesh'], cam_intrinsics_file=sc['K'], est_refine_iter=2, estimator=est)
  out = P.Pipeline('pose', stop_on_error=True).add_processor(proc).run(
      P.PipelineData(rgb=sc['rgb'], depth=sc['depth'], mask=sc['mask'].astype(np.uint8) * 255))
  assert np.array_equal(out.pose, direct)                        # same inputs, deterministic kernels
  assert len(out.pose_6d) == 6 and np.allclose(out.pose_6d[:3], direct[:3, 3])


def test_register_cfg4_maximum_size():
  """1008 hypotheses (42 views x 24 in-plane steps of 15 degrees, unclustered as SURVEY 8(d) prescribes), 500 k faces,
  1920x1200: exercises the chunked network path (4 chunks of <= 256), the large render scratch and the wide frame."""
  from foundationpose_b200 import synthetic
  from foundationpose_b200.estimater import make_rotation_grid_host
  sc = synthetic.make_scene(4, n_points=400000)
  est = _estimator(sc)
  grid = make_rotation_grid_host(min_n_views=40, inplane_step=15)
  assert grid.shape == (1008, 4, 4)
  est.rot_grid = torch.as_tensor(grid, device='cuda', dtype=torch.float)
  pose = est.register(K=sc['K'], rgb=sc['rgb'], depth=sc['depth'], ob_mask=sc['mask'], iteration=2)
  assert pose.shape == (4, 4) and np.isfinite(pose).all()
  R = pose[:3, :3].astype(np.float64)
  assert np.abs(R @ R.T - np.eye(3)).max() < 1e-3 and abs(np.linalg.det(R) - 1) < 1e-3
  scores = est.scores.cpu().numpy()
  assert scores.shape == (1008,) and np.isfinite(scores).all() and (np.diff(scores) <= 0).all()
  assert est.poses.shape == (1008, 4, 4) and torch.isfinite(est.poses).all()
  # translation stays near the guessed object centre (updates are bounded by trans_normalizer * radius per iteration)
  assert np.linalg.norm(pose[:3, 3] - sc['gt_pose'][:3, 3]) < 0.5
  assert 0 <= int(est.best_id) < 1008
  # tracking from the registered pose on the same frame (one hypothesis, cfg-3 style call)
  pose2 = est.track_one(rgb=sc['rgb'], depth=sc['depth'], K=sc['K'], iteration=1)
  assert pose2.shape == (4, 4) and np.isfinite(pose2).all()


def test_cfg4_render_and_refine_iteration_match_oracle():
  """Parity AT cfg-4 size (500 k-face bent sheet, 1920x1200, the auto-selected lean + cull-then-rasterise kernels): for 4 of
  the 1008 hypotheses the render is bit-exact against the oracle (colour, xyz, triangle ids) and one fp16 refine iteration
  stays inside the north-star bounds (1e-3 rad, 0.1 mm) on the oracle's input poses."""
  from foundationpose_b200 import nets, synthetic, utils as U
  from foundationpose_b200.estimater import make_rotation_grid_host
  from oracle import ops as oo
  from oracle.pipeline import OracleEstimator
  from tests.helpers import centered, mesh_np
  sc = synthetic.make_scene(4, n_points=400000)
  mesh, c = centered(sc)
  grid = make_rotation_grid_host(min_n_views=40, inplane_step=15).astype(np.float32)
  poses = grid[[3, 250, 611, 1000]].copy()
  poses[:, :3, 3] = (sc['gt_pose'][:3, :3] @ c + sc['gt_pose'][:3, 3]).astype(np.float32)
  S, H, W = 160, sc['H'], sc['W']
  mnp = mesh_np(mesh)
  cp = oo.crop_params(poses, sc['K'], sc['diameter'], 1.2, S, H, W)
  color, xyz, tri = oo.render(poses, cp, mnp, S)
  mt = U.make_mesh_tensors(mesh)
  pt = torch.as_tensor(poses).cuda()
  params = U.crop_params(pt, sc['K'], sc['diameter'], 1.2, S, H, W)
  import os
  os.environ.setdefault('FPB_RENDER_LEAN', '1')          # what N = 1008 selects automatically (read once per process)
  _, ex = U.render_crops(mt, pt, params, S, sc['diameter'], 0.001, bf16=False, debug=True)
  assert np.array_equal(ex['tri'].cpu().numpy(), tri) and (tri > 0).mean() > 0.05
  assert np.array_equal(ex['color'].cpu().numpy(), color) and np.array_equal(ex['xyz'].cpu().numpy(), xyz)
  sd_r, sd_s = nets.synthetic_state_dict(nets.REFINE), nets.synthetic_state_dict(nets.SCORE)
  orc = OracleEstimator(mnp, sc['diameter'], sd_r, sd_s)
  d = orc.preprocess(sc['depth'])
  xyz_map = oo.depth2xyzmap(d, sc['K'])
  ref = orc.refine_predict(sc['rgb'], xyz_map, poses, sc['K'], 1)
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  got, _ = PoseRefinePredictor(state_dict=sd_r, precision='fp16').predict(
      rgb=sc['rgb'], depth=d, K=sc['K'], ob_in_cams=poses, xyz_map=xyz_map, mesh=mesh, mesh_tensors=mt, mesh_diameter=sc['diameter'],
      iteration=1)
  got = got.cpu().numpy()
  rot = 2 * np.arcsin(np.clip(np.linalg.norm((got[:, :3, :3].astype(np.float64) - ref[:, :3, :3]).reshape(4, -1), axis=-1) / (2 * np.sqrt(2.0)), 0, 1))
  trn = np.abs(got[:, :3, 3] - ref[:, :3, 3]).max()
  print(f'cfg-4 size, 4 hypotheses: rot err {rot.max():.2e} rad, trans err {trn * 1e3:.4f} mm')
  assert rot.max() < 1e-3 and trn < 1e-4
