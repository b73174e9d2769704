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
