"""CPU: pins of the restated pieces to the reference's own code / to real library samplers.

  * tests/golden/pins.npz (tests/golden/make_golden.py, generated where /root/reference exists): the imported
    src/transform.py PoseTransformer and AST-extracted add_err / adds_err / to_homo_torch / transform_pts /
    transform_dirs / symmetry_tfs_from_info  -> the package's host functions and the oracle's restatements.
  * oracle make_B (kornia.warp_perspective + transform_batch restated as closed-form index math) against
    tests/kornia_ref.py, which restates only kornia's coordinate normalisation and calls the REAL
    torch.nn.functional.grid_sample.  Bit-equality is not attainable here: kornia reaches the sampling position through
    a 3x3 LU inverse and normalised coordinates (whose rounding even differs between torch-CPU and torch-CUDA), the
    oracle / CUDA path through the algebraically identical p*W/(W-1)-0.5.  The bound asserted: bilinear rgb within 1e-4
    (of [0,1]) and nearest-neighbour picks identical except where the source coordinate is an exact .5 tie, which
    takes out at most one crop row + one crop column per hypothesis.
"""
import json
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


@pytest.fixture(scope='module')
def pins():
  return np.load(os.path.join(GOLD, 'pins.npz'))


def test_pose_transformer_matches_reference_module(pins):
  from foundationpose_b200.transform import PoseTransformer
  import contextlib, io
  with contextlib.redirect_stdout(io.StringIO()):
    a = np.array([PoseTransformer().transform_pose(P) for P in pins['pt_poses']], dtype=np.float64)
    b = np.array([PoseTransformer(False, False).transform_pose(P) for P in pins['pt_poses']], dtype=np.float64)
  assert np.array_equal(a, pins['pt_inch_deg']) and np.array_equal(b, pins['pt_m_rad'])
  assert abs(abs(pins['pt_m_rad'][8, 4]) - np.pi / 2) < 1e-12 and pins['pt_m_rad'][8, 5] == 0      # the gimbal rows are in the set
  assert abs(abs(pins['pt_m_rad'][9, 4]) - np.pi / 2) < 1e-12 and pins['pt_m_rad'][9, 5] == 0


def test_pose_to_6d_of_the_pipeline_matches_reference_module(pins):
  from foundationpose_b200.pipeline import pose_to_6d
  got = np.array([pose_to_6d(P) for P in pins['pt_poses']], dtype=np.float64)
  assert np.allclose(got, pins['pt_m_rad'], atol=1e-12, rtol=0)


def test_point_transform_helpers_match_reference(pins):
  from foundationpose_b200 import utils as U
  tp, tf1, tfb = (torch.from_numpy(pins[k]) for k in ('tp_pts', 'tp_tf1', 'tp_tfb'))
  assert np.array_equal(U.to_homo_torch(tp).numpy(), pins['tp_homo'])
  assert np.array_equal(U.transform_pts(tp, tf1).numpy(), pins['tp_single'])
  assert np.array_equal(U.transform_pts(tp, tfb).numpy(), pins['tp_batched'])
  assert np.array_equal(U.transform_dirs(tp, tf1).numpy(), pins['td_single'])
  assert np.array_equal(U.transform_dirs(tp, tfb).numpy(), pins['td_batched'])


def test_symmetry_tfs_from_info_matches_reference(pins):
  from foundationpose_b200 import utils as U
  infos = json.loads(str(pins['sym_infos']))
  for i, info in enumerate(infos):
    got = U.symmetry_tfs_from_info(info, rot_angle_discrete=5 if i != 4 else 30)
    assert got.shape == pins[f'sym_{i}'].shape and np.array_equal(got, pins[f'sym_{i}']), i


def test_oracle_add_adds_match_reference(pins):
  from oracle import geometry as og
  a = og.add_err(pins['err_pred'], pins['err_gt'], pins['err_pts'])
  s = og.adds_err(pins['err_pred'], pins['err_gt'], pins['err_pts'])
  assert abs(a - float(pins['add'])) < 1e-12 and abs(s - float(pins['adds'])) < 1e-12
  assert float(pins['adds']) < float(pins['add'])


@pytest.mark.parametrize('mode', [0, 1])
def test_oracle_make_B_against_real_grid_sample(scene1, mode):
  """predict_pose_refine.py:63,72 + h5_dataset.py:79-127 (mode 0) / predict_score.py:89-93 + h5_dataset.py:137-179
  (mode 1: depth crop -> full frame -> xyz -> crop) evaluated with the real ATen sampler."""
  from oracle import ops as oo
  from tests import kornia_ref as kr
  from tests.helpers import hypo_poses
  sc = scene1
  poses, d = hypo_poses(sc, n=8, jitter_seed=3)
  poses[1, :3, 3] += np.array([0.25, 0.0, 0.0], np.float32)          # crop partly outside the image (zero padding)
  poses[2, :3, 3] += np.array([0.0, -0.2, 0.05], np.float32)
  K, S, H, W = sc['K'], 160, 480, 640
  N = len(poses)
  xyz_map = oo.depth2xyzmap(d, K)
  cp = oo.crop_params(poses, K, sc['diameter'], 1.2 if mode == 0 else 1.1, S, H, W)
  B = oo.make_B(sc['rgb'].astype(np.float32), xyz_map if mode == 0 else None, d if mode == 1 else None, poses, cp, K, S,
                sc['diameter'], mode)
  tf = torch.from_numpy(cp['tf'].reshape(-1, 3, 3))
  rgb = torch.from_numpy(sc['rgb'].astype(np.float32)).permute(2, 0, 1)[None].expand(N, -1, -1, -1)
  rgbBs = kr.warp_perspective(rgb, tf, (S, S), 'bilinear') / 255.0
  poseA, rad = torch.from_numpy(poses), torch.full((N,), sc['diameter'] / 2)
  if mode == 0:
    xyzBs = kr.warp_perspective(torch.from_numpy(xyz_map).permute(2, 0, 1)[None].expand(N, -1, -1, -1), tf, (S, S), 'nearest')
    xyzBs = kr.transform_xyz(xyzBs, poseA, rad, 0.001)
  else:
    depthBs = kr.warp_perspective(torch.from_numpy(d)[None, None].expand(N, -1, -1, -1), tf, (S, S), 'nearest')
    ori = kr.warp_perspective(depthBs, tf.inverse(), (H, W), 'nearest')
    Ks = torch.from_numpy(K.astype(np.float32)).reshape(1, 3, 3).expand(N, 3, 3)
    xm = kr.depth2xyzmap_batch(ori[:, 0], Ks, np.inf).permute(0, 3, 1, 2)
    xyzBs = kr.transform_xyz(kr.warp_perspective(xm, tf, (S, S), 'nearest'), poseA, rad, 0.1)
  ref = torch.cat([rgbBs, xyzBs], 1).permute(0, 2, 3, 1).numpy()
  assert np.abs(ref[..., :3] - B[..., :3]).max() < 1e-4
  bad = np.abs(ref[..., 3:] - B[..., 3:]).max(-1) > 0                  # [N,S,S] pixels whose nearest pick differs
  assert (np.abs(B[..., 3:]).max(-1) > 0).mean() > 0.5                 # the crops are not empty
  assert bad.mean() < 5e-3
  for n in range(N):                                                   # mismatches sit on whole tie rows / columns
    rows, cols = np.unique(np.nonzero(bad[n])[0]), np.unique(np.nonzero(bad[n])[1])
    assert bad[n].sum() <= 2 * S and (len(rows) <= 1 or len(cols) <= 1), (n, int(bad[n].sum()))


def test_missing_weights_raise_instead_of_random_init():
  """The reference crashes on a missing checkpoint (predict_pose_refine.py:133-137); so must the drop-in, unless
  synthetic weights are requested explicitly."""
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from foundationpose_b200.predict_score import ScorePredictor
  with pytest.raises(FileNotFoundError):
    PoseRefinePredictor()
  with pytest.raises(FileNotFoundError):
    ScorePredictor()


def test_run_config_is_read_and_backward_compat_defaults_apply(tmp_path):
  from foundationpose_b200 import predict_pose_refine as pr
  (tmp_path / 'config.yml').write_text('c_in: 6\nuse_BN: true\nnormalize_xyz: true\ncrop_ratio: null\nzfar: .inf\n'
                                       'rot_normalizer: 0.25\ntrans_normalizer: [0.01, 0.01, 0.03]\n')
  cfg, sd = pr.resolve_weights('none', pr.default_refine_cfg(), pr.REFINE_COMPAT_DEFAULTS, None, {'x': 1},
                               str(tmp_path / 'model_best.pth'), False, 'test')
  assert sd == {'x': 1}
  assert cfg['rot_normalizer'] == 0.25 and cfg['trans_normalizer'] == [0.01, 0.01, 0.03]
  assert cfg['crop_ratio'] == 1.2 and cfg['trans_rep'] == 'tracknet' and cfg['zfar'] == np.inf     # :113-125
  (tmp_path / 'config.yml').write_text('rot_normalizer: 0.25\n')
  cfg, _ = pr.resolve_weights('none', pr.default_refine_cfg(), pr.REFINE_COMPAT_DEFAULTS, None, {'x': 1},
                              str(tmp_path / 'model_best.pth'), False, 'test')
  assert cfg['c_in'] == 4 and cfg['normalize_xyz'] is False and cfg['use_BN'] is False           # old-run defaults :107-128


def test_batch_pose_data_bag():
  from foundationpose_b200.datasets import BatchPoseData, PoseRefinePairH5Dataset
  b = BatchPoseData(rgbAs=torch.arange(24.).reshape(4, 3, 1, 2), poseA=torch.eye(4)[None].repeat(4, 1, 1))
  calls = []
  b.set_lazy('depthAs', lambda: (calls.append(1), torch.ones(4, 1, 1, 2))[1])
  sel = b.select_by_indices(torch.tensor([2, 0]))
  assert sel.rgbAs.shape == (2, 3, 1, 2) and torch.equal(sel.rgbAs[1], b.rgbAs[0]) and sel.depthAs.shape == (2, 1, 1, 2)
  assert sel.rgbBs is None and len(calls) == 1 and b.depthAs.shape == (4, 1, 1, 2) and len(calls) == 1
  with pytest.raises(NotImplementedError):
    PoseRefinePairH5Dataset(cfg={}).transform_batch(BatchPoseData(), 480, 640)
  b.transformed = True
  assert PoseRefinePairH5Dataset(cfg={}).transform_batch(b, 480, 640) is b
