"""GPU: the drop-in surface around the hot path -- checkpoint + config.yml loading, BatchPoseData, last_*_update
semantics, nvdiffrast_render's remaining arguments (normal_map, use_light, weights), the mesh_tensors-dict path and
batches above the 65535-hypothesis launch limit."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')

RELEASED_REFINE_YML = ('input_resize: [160, 160]\nc_in: 6\nuse_BN: true\nuse_normal: false\nuse_mask: false\nnormalize_xyz: true\n'
                       'rot_rep: axis_angle\ntrans_rep: tracknet\ntrans_normalizer: [0.019999999552965164, 0.019999999552965164, '
                       '0.05000000074505806]\nrot_normalizer: 0.3490658503988659\ncrop_ratio: 1.2\nn_view: 1\nzfar: .inf\n')
RELEASED_SCORE_YML = 'input_resize: [160, 160]\nc_in: 6\nuse_BN: true\nuse_normal: false\nnormalize_xyz: true\ncrop_ratio: 1.1\nzfar: .inf\n'


def _golden_AB(seed, n, dtype):
  from foundationpose_b200 import utils as U
  from foundationpose_b200.nets import _calibration_batch
  x = _calibration_batch(torch.Generator().manual_seed(seed), n)            # [2n,6,160,160]
  AB = U.net_input_buffer(n, 160, dtype)
  AB[:, :, U.IN_XOFF:U.IN_XOFF + 160, :6] = x.permute(0, 2, 3, 1).to(dtype).cuda()
  return AB


@pytest.mark.parametrize('wrap', [True, False])
def test_predictors_load_model_best_pth_and_config_yml(tmp_path, wrap):
  """predict_pose_refine.py:133-143 / predict_score.py:146-156: torch.load -> ckpt['model'] -> load_state_dict, run config
  beside it.  The checkpoint is the state_dict of a real torch module with the reference's key set (BatchNorm
  num_batches_tracked, the pos_embed.pe buffer); outputs must equal the golden vectors of the reference's own modules."""
  from foundationpose_b200 import nets
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from foundationpose_b200.predict_score import ScorePredictor
  from oracle import nets as on
  for kind, name, yml in (('refine', nets.REFINE, RELEASED_REFINE_YML), ('score', nets.SCORE, RELEASED_SCORE_YML)):
    d = tmp_path / kind
    d.mkdir()
    sd = on.build(kind, nets.synthetic_state_dict(name)).state_dict()
    assert any(k.endswith('num_batches_tracked') for k in sd) and 'pos_embed.pe' in sd
    torch.save({'model': sd, 'epoch': 3} if wrap else sd, str(d / 'model_best.pth'))
    (d / 'config.yml').write_text(yml)
  ref = PoseRefinePredictor(ckpt_path=str(tmp_path / 'refine' / 'model_best.pth'), precision='fp32')
  sco = ScorePredictor(ckpt_path=str(tmp_path / 'score' / 'model_best.pth'), precision='fp32')
  assert ref.cfg['crop_ratio'] == 1.2 and sco.cfg['crop_ratio'] == 1.1 and ref.cfg['zfar'] == np.inf
  g = np.load(os.path.join(GOLD, 'refine_net.npz'))
  t, r = ref.model.refine(_golden_AB(int(g['seed']), int(g['n']), torch.float), int(g['n']))
  assert np.abs(t.cpu().numpy() - g['trans']).max() < 2e-4 and np.abs(r.cpu().numpy() - g['rot']).max() < 2e-4
  g = np.load(os.path.join(GOLD, 'score_net.npz'))
  f = sco.model.score_features(_golden_AB(int(g['seed']), int(g['n']), torch.float), int(g['n']))
  assert np.abs(f.cpu().numpy() - g['feats']).max() < 2e-4 * max(1.0, float(np.abs(g['feats']).max()))
  # a run whose config.yml lacks the keys gets the reference's old-run defaults (c_in=4 ...), which this build rejects loudly
  (tmp_path / 'refine' / 'config.yml').write_text('rot_normalizer: 0.3\n')
  with pytest.raises(NotImplementedError):
    PoseRefinePredictor(ckpt_path=str(tmp_path / 'refine' / 'model_best.pth'))


def _predictors(precision='fp32'):
  from foundationpose_b200 import nets
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from foundationpose_b200.predict_score import ScorePredictor
  return (PoseRefinePredictor(state_dict=nets.synthetic_state_dict(nets.REFINE), precision=precision),
          ScorePredictor(state_dict=nets.synthetic_state_dict(nets.SCORE), precision=precision))


def test_make_crop_data_batch_returns_batch_pose_data(scene1):
  """learning/datasets/pose_dataset.py:67-134 fields as zero-copy views of the fused network input."""
  from foundationpose_b200 import predict_pose_refine as pr, predict_score as ps, utils as U
  from foundationpose_b200.datasets import BatchPoseData
  from oracle import ops as oo
  from tests.helpers import centered, hypo_poses, mesh_np
  from tests import kornia_ref as kr
  sc = scene1
  mesh, _ = centered(sc)
  poses, d = hypo_poses(sc, n=6, jitter_seed=5)
  xyz = oo.depth2xyzmap(d, sc['K'])
  rgb_t, xyz_t, d_t = torch.as_tensor(sc['rgb']).cuda(), torch.as_tensor(xyz).cuda(), torch.as_tensor(d).cuda()
  mt = U.make_mesh_tensors(mesh)
  N, S = len(poses), 160
  pd = pr.make_crop_data_batch([S, S], torch.as_tensor(poses).cuda(), mesh, rgb_t, None, sc['K'], 1.2, xyz_t,
                               mesh_diameter=sc['diameter'], mesh_tensors=mt, bf16=False)
  assert isinstance(pd, BatchPoseData) and pd.transformed
  for name in ('rgbAs', 'rgbBs', 'xyz_mapAs', 'xyz_mapBs'):
    v = getattr(pd, name)
    assert v.shape == (N, 3, S, S) and v.untyped_storage().data_ptr() == pd.AB.untyped_storage().data_ptr()   # a view
  assert pd.poseA.shape == (N, 4, 4) and pd.tf_to_crops.shape == (N, 3, 3) and pd.Ks.shape == (1, 3, 3)
  assert pd.mesh_diameters.shape == (N,) and pd.depthAs is None and pd.depthBs is None                        # :83 passes None
  cp = oo.crop_params(poses, sc['K'], sc['diameter'], 1.2, S, 480, 640)
  color, oxyz, _ = oo.render(poses, cp, mesh_np(mesh), S)
  A = oo.make_A(color, oxyz, poses, sc['diameter'], 0.001)
  B = oo.make_B(sc['rgb'].astype(np.float32), xyz, None, poses, cp, sc['K'], S, sc['diameter'], 0)
  assert np.array_equal(pd.rgbAs.permute(0, 2, 3, 1).cpu().numpy(), A[..., :3])
  assert np.array_equal(pd.xyz_mapAs.permute(0, 2, 3, 1).cpu().numpy(), A[..., 3:])
  assert np.array_equal(pd.rgbBs.permute(0, 2, 3, 1).cpu().numpy(), B[..., :3])
  assert np.array_equal(pd.xyz_mapBs.permute(0, 2, 3, 1).cpu().numpy(), B[..., 3:])
  assert np.array_equal(pd.tf_to_crops.cpu().numpy().reshape(N, 9), cp['tf'].reshape(N, 9))
  sel = pd.select_by_indices(torch.tensor([4, 1]))
  assert torch.equal(sel.rgbAs[0], pd.rgbAs[4]) and torch.equal(sel.poseA[1], pd.poseA[1])
  assert sel.rgbAs.untyped_storage().data_ptr() != pd.AB.untyped_storage().data_ptr() and sel.Ks.shape == (1, 3, 3)
  refiner, scorer = _predictors()
  assert refiner.dataset.transform_batch(pd, H_ori=480, W_ori=640, bound=1) is pd
  # scorer flavour: raw crop depths on demand (predict_score.py:84-93)
  pds = ps.make_crop_data_batch([S, S], torch.as_tensor(poses).cuda(), mesh, rgb_t, d_t, sc['K'], 1.1,
                                mesh_diameter=sc['diameter'], mesh_tensors=mt, bf16=False)
  assert pds.Ks.shape == (N, 3, 3) and pds.xyz_mapBs.shape == (N, 3, S, S)
  cps = oo.crop_params(poses, sc['K'], sc['diameter'], 1.1, S, 480, 640)
  _, sxyz, _ = oo.render(poses, cps, mesh_np(mesh), S)
  assert pds.depthAs.shape == (N, 1, S, S) and np.array_equal(pds.depthAs[:, 0].cpu().numpy(), sxyz[..., 2])
  want = kr.warp_perspective(torch.from_numpy(d)[None, None].expand(N, -1, -1, -1), torch.from_numpy(cps['tf'].reshape(-1, 3, 3)),
                             (S, S), 'nearest').numpy()
  got = pds.depthBs.cpu().numpy()
  assert got.shape == (N, 1, S, S) and (got != want).mean() < 5e-3 and (got > 0).mean() > 0.5          # .5-tie rows only
  assert scorer.dataset.transform_batch(pds, 480, 640) is pds


def test_last_updates_hold_trans_delta_and_rot_mat_delta(scene1):
  """predict_pose_refine.py:238-239: last_trans_update = trans_delta [N,3], last_rot_update = rot_mat_delta [N,3,3]."""
  from foundationpose_b200 import utils as U
  from oracle import ops as oo
  from tests.helpers import centered, hypo_poses
  sc = scene1
  mesh, _ = centered(sc)
  poses, d = hypo_poses(sc, n=5, jitter_seed=2)
  refiner, _ = _predictors()
  out, _ = refiner.predict(rgb=sc['rgb'], depth=d, K=sc['K'], ob_in_cams=poses, xyz_map=oo.depth2xyzmap(d, sc['K']), mesh=mesh,
                           mesh_tensors=U.make_mesh_tensors(mesh), mesh_diameter=sc['diameter'], iteration=1)
  td, rd = refiner.last_trans_update.cpu().numpy(), refiner.last_rot_update.cpu().numpy()
  raw_t, raw_r = refiner.last_trans_raw.cpu().numpy(), refiner.last_rot_raw.cpu().numpy()
  assert td.shape == (5, 3) and rd.shape == (5, 3, 3) and raw_t.shape == (5, 3) and raw_r.shape == (5, 3)
  assert np.array_equal(td, raw_t * np.float32(sc['diameter'] / 2))                                    # :195-199, :226-227
  want = oo.so3_exp_map(np.tanh(raw_r) * np.float32(refiner.cfg['rot_normalizer'])).transpose(0, 2, 1)
  assert np.abs(rd - want).max() < 1e-6
  out = out.cpu().numpy()
  assert np.array_equal(out[:, :3, 3], poses[:, :3, 3] + td)
  assert np.abs(out[:, :3, :3] - rd @ poses[:, :3, :3]).max() < 1e-6


def test_nvdiffrast_render_normal_map_and_light_arguments(scene1):
  """src/Utils.py:133-219: normal_map (:192-197), use_light on / off, w_ambient / w_diffuse (:211); unsupported light
  arguments raise instead of being ignored."""
  from foundationpose_b200 import utils as U
  from oracle import ops as oo
  from tests.helpers import centered, hypo_poses, mesh_np
  sc = scene1
  mesh, _ = centered(sc)
  poses, _ = hypo_poses(sc, n=4, jitter_seed=7)
  S = 160
  cp = oo.crop_params(poses, sc['K'], sc['diameter'], 1.2, S, 480, 640)
  mt = U.make_mesh_tensors(mesh)
  bbox = torch.as_tensor(cp['bbox']).cuda()
  for kw in (dict(use_light=True), dict(use_light=False, get_normal=True), dict(use_light=True, w_ambient=0.6, w_diffuse=0.9)):
    extra = {}
    color, depth, normal = U.nvdiffrast_render(K=sc['K'], H=480, W=640, ob_in_cams=torch.as_tensor(poses).cuda(), mesh_tensors=mt,
                                               bbox2d=bbox, output_size=[S, S], extra=extra, **kw)
    oc, ox, _, on_ = oo.render(poses, cp, mesh_np(mesh), S, use_light=kw['use_light'], w_ambient=kw.get('w_ambient', 0.8),
                               w_diffuse=kw.get('w_diffuse', 0.5), get_normal=True)
    # this API path builds its matrices with torch matmul (not the fused crop-parameter kernel): tolerance, as in
    # test_geometry_gpu.test_nvdiffrast_render_api; the bit-exact check of the normal map follows below
    assert (np.abs(color.cpu().numpy() - oc) > 1e-3).mean() < 2e-3 and (np.abs(extra['xyz_map'].cpu().numpy() - ox) > 1e-4).mean() < 2e-3
    assert np.array_equal(depth.cpu().numpy(), extra['xyz_map'][..., 2].cpu().numpy())
    assert normal is not None and (np.abs(normal.cpu().numpy() - on_) > 1e-3).mean() < 2e-3   # use_light implies get_normal (:190-191)
    nn = np.linalg.norm(on_, axis=-1)
    assert np.all((np.abs(nn - 1) < 1e-5) | (nn == 0)) and (nn > 0).mean() > 0.2
  # bit-exact through the kernel's own crop parameters: colour (both light settings) and normal map
  pr = U.crop_params(torch.as_tensor(poses).cuda(), sc['K'], sc['diameter'], 1.2, S, 480, 640)
  for ul, wa, wd in ((True, 0.8, 0.5), (False, 0.8, 0.5), (True, 0.6, 0.9)):
    _, ex = U.render_crops(mt, torch.as_tensor(poses).cuda(), pr, S, sc['diameter'], 0.001, bf16=False, debug=True, use_light=ul,
                           w_ambient=wa, w_diffuse=wd, get_normal=True)
    oc, ox, _, on_ = oo.render(poses, cp, mesh_np(mesh), S, use_light=ul, w_ambient=wa, w_diffuse=wd, get_normal=True)
    assert np.array_equal(ex['color'].cpu().numpy(), oc) and np.array_equal(ex['xyz'].cpu().numpy(), ox)
    assert np.array_equal(ex['normal'].cpu().numpy(), on_)
  color, depth, normal = U.nvdiffrast_render(K=sc['K'], H=480, W=640, ob_in_cams=torch.as_tensor(poses).cuda(), mesh_tensors=mt,
                                             bbox2d=bbox, output_size=[S, S], use_light=False)
  assert normal is None
  with pytest.raises(NotImplementedError):
    U.nvdiffrast_render(K=sc['K'], H=480, W=640, ob_in_cams=torch.as_tensor(poses).cuda(), mesh_tensors=mt, bbox2d=bbox,
                        output_size=[S, S], use_light=True, light_dir=np.array([0, 1, 0]))
  with pytest.raises(NotImplementedError):
    U.nvdiffrast_render(K=sc['K'], H=480, W=640, ob_in_cams=torch.as_tensor(poses).cuda(), mesh_tensors=mt, bbox2d=bbox,
                        output_size=[S, S], use_light=True, light_dir=None)
  with pytest.raises(NotImplementedError):
    U.nvdiffrast_render(K=sc['K'], H=480, W=640, ob_in_cams=torch.as_tensor(poses).cuda(), mesh_tensors=mt, bbox2d=bbox,
                        output_size=[S, S], use_light=True, light_color=[1, 1, 1])
  # full-frame render (bbox2d=None): the object lands where the synthetic frame has it
  gt = (sc['gt_pose'] @ np.linalg.inv(np.block([[np.eye(3), -centered(sc)[1].reshape(3, 1)], [np.zeros((1, 3)), 1]]))).astype(np.float32)
  color, depth, _ = U.nvdiffrast_render(K=sc['K'], H=480, W=640, ob_in_cams=torch.as_tensor(gt[None]).cuda(), mesh_tensors=mt,
                                        use_light=True)
  assert color.shape == (1, 480, 640, 3) and depth.shape == (1, 480, 640)
  m = depth[0].cpu().numpy() > 0
  inter = (m & sc['mask']).sum() / max(1, (m | sc['mask']).sum())
  assert inter > 0.9, inter


def test_pipeline_accepts_a_mesh_tensors_dict(scene1):
  from foundationpose_b200 import utils as U
  from foundationpose_b200.estimater import FoundationPose
  from foundationpose_b200.pipeline import FoundationPoseEstimator, PipelineData
  sc = scene1
  refiner, scorer = _predictors('bf16')
  mt = {k: v for k, v in U.make_mesh_tensors(sc['mesh']).items()}
  proc = FoundationPoseEstimator(mesh_file=mt, cam_intrinsics_file=sc['K'], est_refine_iter=1, scorer=scorer, refiner=refiner)
  out = proc.process(PipelineData(rgb=sc['rgb'], depth=sc['depth'], mask=sc['mask']))
  est = FoundationPose(model_pts=sc['mesh'].vertices, model_normals=sc['mesh'].vertex_normals, mesh=sc['mesh'], scorer=scorer,
                       refiner=refiner)
  want = est.register(K=sc['K'], rgb=sc['rgb'], depth=sc['depth'], ob_mask=sc['mask'], iteration=1)
  assert out.pose.shape == (4, 4) and np.abs(out.pose - want).max() < 1e-5
  assert proc.est.diameter == pytest.approx(est.diameter, rel=1e-6)


def test_render_and_warp_above_the_launch_limit(scene1):
  """N > 65535 hypotheses run in chunks (the reference batches renders in chunks of 512 for any N)."""
  from foundationpose_b200 import utils as U
  from tests.helpers import centered, hypo_poses
  sc = scene1
  mesh, _ = centered(sc)
  base, d = hypo_poses(sc, n=7, jitter_seed=9)
  N, S = 65535 + 9, 8
  idx = np.arange(N) % len(base)
  poses = torch.as_tensor(base[idx]).cuda()
  mt = U.make_mesh_tensors(mesh)
  params = U.crop_params(poses, sc['K'], sc['diameter'], 1.2, S, 480, 640)
  A, _ = U.render_crops(mt, poses, params, S, sc['diameter'], 0.001, bf16=False)
  rgb_t = torch.as_tensor(sc['rgb']).cuda()
  from oracle import ops as oo
  xyz_t = torch.as_tensor(oo.depth2xyzmap(d, sc['K'])).cuda()
  B = U.warp_crops(rgb_t, xyz_t, None, poses, params, sc['K'], S, sc['diameter'], 0, bf16=False)
  A7, _ = U.render_crops(mt, poses[:7], params[:7].contiguous(), S, sc['diameter'], 0.001, bf16=False)
  B7 = U.warp_crops(rgb_t, xyz_t, None, poses[:7], params[:7].contiguous(), sc['K'], S, sc['diameter'], 0, bf16=False)
  sel = torch.as_tensor(idx).cuda()
  assert torch.equal(A, A7[sel]) and torch.equal(B, B7[sel]) and float(A[-1].abs().sum()) > 0


@pytest.mark.parametrize('N', [1, 7, 252, 1008, 4096])
def test_rank_poses_matches_stable_descending_sort(N):
  """src/estimater.py:221-236: argsort(descending) + gathers + best pose @ tf_to_center in one kernel; ties keep input order."""
  from foundationpose_b200 import utils as U
  g = torch.Generator(device='cuda').manual_seed(N)
  scores = torch.randn(N, device='cuda', generator=g)
  if N > 4:
    scores[N // 2] = scores[1]; scores[N - 1] = scores[1]           # ties
  poses = torch.randn(N, 4, 4, device='cuda', generator=g)
  tf = torch.eye(4, device='cuda'); tf[:3, 3] = torch.tensor([0.1, -0.2, 0.3])
  ids, s, p, best = U.rank_poses(scores, poses, tf)
  want = torch.sort(scores, descending=True, stable=True).indices
  assert torch.equal(ids, want) and torch.equal(s, scores[want]) and torch.equal(p, poses[want])
  assert torch.allclose(best, poses[want[0]] @ tf, atol=1e-6)


@pytest.mark.parametrize('precision,N', [('fp32', 12), ('fp16', 252), ('bf16', 40), ('fp16', 300)])
def test_first_iteration_shared_observed_crop_is_bit_identical(scene1, precision, N):
  """register()'s hypotheses share one translation => one crop window => one observed crop B in the first refine iteration
  (src/estimater.py:127-134, src/Utils.py:577-621).  fpb_refine_forward_shared_b warps / encodes B once and broadcasts its
  features: poses and raw head outputs must equal the replicated path bit for bit (N = 300 crosses the 256-hypothesis chunk)."""
  from foundationpose_b200 import nets, utils as U
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from oracle import ops as oo
  from tests.helpers import centered, hypo_poses
  sc = scene1
  mesh, _ = centered(sc)
  base, d = hypo_poses(sc, n=None)
  poses = base[np.arange(N) % len(base)].copy()
  assert (poses[:, :3, 3] == poses[0, :3, 3]).all()
  refiner = PoseRefinePredictor(state_dict=nets.synthetic_state_dict(nets.REFINE), precision=precision)
  kw = dict(rgb=sc['rgb'], depth=d, K=sc['K'], xyz_map=oo.depth2xyzmap(d, sc['K']), mesh=mesh, mesh_tensors=U.make_mesh_tensors(mesh),
            mesh_diameter=sc['diameter'], iteration=2)
  a, _ = refiner.predict(ob_in_cams=poses, same_translation=False, **kw)
  a, ta, ra = a.clone(), refiner.last_trans_raw.clone(), refiner.last_rot_raw.clone()
  from foundationpose_b200._lib import lib
  lib().fpb_launch_count(1)
  b, _ = refiner.predict(ob_in_cams=poses, **kw)                      # host poses: detected automatically
  assert torch.equal(a, b) and torch.equal(ta, refiner.last_trans_raw) and torch.equal(ra, refiner.last_rot_raw)
  c, _ = refiner.predict(ob_in_cams=torch.as_tensor(poses).cuda(), same_translation=True, **kw)
  assert torch.equal(a, c)
  jit = poses.copy(); jit[1, 0, 3] += 1e-4                           # not shared: the detection must say no
  e, _ = refiner.predict(ob_in_cams=jit, **kw)
  f, _ = refiner.predict(ob_in_cams=jit, same_translation=False, **kw)
  assert torch.equal(e, f) and not torch.equal(e[1], a[1])


def test_reset_object_voxel_downsample_follows_open3d_rule(scene2):
  """src/estimater.py:59-64: pcd.voxel_down_sample(vox_size) -- grid anchored at min_bound - vox/2, per-voxel averages of
  points and normals; est.pts / est.normals / min_xyz / max_xyz derive from it (brute-force restatement of the rule)."""
  from foundationpose_b200 import nets
  from foundationpose_b200.estimater import FoundationPose
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from foundationpose_b200.predict_score import ScorePredictor
  sc = scene2
  est = FoundationPose(model_pts=sc['mesh'].vertices, model_normals=sc['mesh'].vertex_normals, mesh=sc['mesh'],
                       refiner=PoseRefinePredictor(state_dict=nets.synthetic_state_dict(nets.REFINE)),
                       scorer=ScorePredictor(state_dict=nets.synthetic_state_dict(nets.SCORE)), mesh_diameter=sc['diameter'])
  assert est.vox_size == max(sc['diameter'] / 20.0, 0.003) and est.dist_bin == est.vox_size / 2 and est.angle_bin == 20
  V = np.asarray(sc['mesh'].vertices) - est.model_center.reshape(1, 3)
  Nn = np.asarray(sc['mesh'].vertex_normals)
  origin = V.min(0) - est.vox_size / 2
  cells = {}
  for p, n in zip(V, Nn):
    cells.setdefault(tuple(np.floor((p - origin) / est.vox_size).astype(int)), []).append((p, n))
  want_p = np.array([np.mean([q[0] for q in cells[k]], axis=0) for k in sorted(cells)])
  want_n = np.array([np.mean([q[1] for q in cells[k]], axis=0) for k in sorted(cells)])
  want_n /= np.linalg.norm(want_n, axis=1, keepdims=True)
  assert est.pts.shape == (len(cells), 3) and 100 < len(cells) < len(V)
  assert np.abs(est.pts.cpu().numpy() - want_p).max() < 1e-6 and np.abs(est.normals.cpu().numpy() - want_n).max() < 1e-5
  assert np.allclose(est.max_xyz, want_p.max(0)) and np.allclose(est.min_xyz, want_p.min(0))
