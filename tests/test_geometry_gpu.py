"""GPU parity tests (through the C ABI) for depth pre-processing, hypothesis generation, crop windows,
the rasteriser and the crop/warp kernels.  Oracle = checker only.  Bars: bit-exact for everything without a
transcendental; <= 4 ulp where expf / sinf / tanhf are involved (stated per test)."""
import numpy as np
import pytest
import torch

from oracle import geometry as og
from oracle import ops as oo
from tests.helpers import mesh_np, centered, hypo_poses

pytestmark = pytest.mark.gpu


def _depth(H, W, seed):
  rng = np.random.default_rng(seed)
  yy, xx = np.mgrid[0:H, 0:W]
  d = 0.6 + 0.1 * np.sin(xx / 40.0) * np.cos(yy / 30.0) + rng.normal(0, 0.0007, (H, W))
  d[rng.random((H, W)) < 0.03] = 0
  d[(xx - W // 2) ** 2 + (yy - H // 2) ** 2 < (H // 6) ** 2] += 0.25    # depth discontinuity
  d[: H // 10] = 120.0                                                  # beyond zfar
  return d.astype(np.float32)


@pytest.mark.parametrize('shape', [(480, 640), (1200, 1920), (37, 53), (1, 1), (5, 3)])
def test_erode_bit_exact(shape):
  from foundationpose_b200 import utils as U
  d = _depth(*shape, seed=1)
  got = U.erode_depth(d)
  assert isinstance(got, np.ndarray)
  assert np.array_equal(got, oo.erode_depth(d))
  t = torch.as_tensor(d, device='cuda')
  got_t = U.erode_depth(t, radius=1, depth_diff_thres=0.002, ratio_thres=0.5, zfar=3)
  assert torch.is_tensor(got_t) and got_t.is_cuda
  assert np.array_equal(got_t.cpu().numpy(), oo.erode_depth(d, 1, 0.002, 0.5, 3))


@pytest.mark.parametrize('shape', [(480, 640), (1200, 1920), (37, 53), (1, 1)])
def test_bilateral_within_4ulp(shape):
  from foundationpose_b200 import utils as U
  d = _depth(*shape, seed=2)
  got = U.bilateral_filter_depth(d)
  ref = oo.bilateral_filter_depth(d)
  assert np.array_equal(got == 0, ref == 0)
  ulp = np.abs(got.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64))
  assert ulp.max() <= 4, ulp.max()      # expf differs by <= 2 ulp between glibc and CUDA


def test_depth2xyz_bit_exact():
  from foundationpose_b200 import utils as U
  d = _depth(480, 640, seed=3)
  K = np.array([[600.3, 0, 320.7], [0, 599.1, 241.2], [0, 0, 1.0]])
  assert np.array_equal(U.depth2xyzmap(d, K), oo.depth2xyzmap(d, K))
  t = torch.as_tensor(d, device='cuda')
  Kt = torch.as_tensor(K, dtype=torch.float, device='cuda')
  got = U.depth2xyzmap_batch(t[None], Kt[None], zfar=np.inf)[0].cpu().numpy()
  assert np.array_equal(got, oo.depth2xyzmap_batch(d[None], K.astype(np.float32)[None], np.inf)[0])
  got = U.depth2xyzmap_batch(t[None], Kt[None], zfar=0.7)[0].cpu().numpy()
  assert np.array_equal(got, oo.depth2xyzmap_batch(d[None], K.astype(np.float32)[None], 0.7)[0])


@pytest.mark.parametrize('case', ['scene', 'empty_mask', 'no_valid', 'odd', 'even', 'full1200'])
def test_guess_translation_bit_exact(case, scene2):
  from foundationpose_b200.estimater import guess_translation_cuda
  K = scene2['K']
  d, m = scene2['depth'].copy(), scene2['mask'].copy()
  if case == 'empty_mask':
    m[:] = False
  elif case == 'no_valid':
    d[m] = 0
  elif case == 'odd':
    m[:] = False; m[10, 10:15] = True; d[10, 10:15] = [0.5, 0.7, 0.6, 0.9, 0.8]
  elif case == 'even':
    m[:] = False; m[10, 10:14] = True; d[10, 10:14] = [0.5, 0.7, 0.6, 0.9]
  elif case == 'full1200':
    d = _depth(1200, 1920, 5); m = d > 0.55
  got = guess_translation_cuda(torch.as_tensor(d, device='cuda'), torch.as_tensor(m, device='cuda'), K)
  ref = og.guess_translation(d, m, K)
  assert got.dtype == np.float64 and np.array_equal(got, ref)


def test_rot_grid_and_cluster_bit_exact():
  from foundationpose_b200 import mycpp
  from foundationpose_b200.estimater import make_rotation_grid_host
  g = make_rotation_grid_host(40, 60)
  assert np.array_equal(g, og.make_rotation_grid_f64())
  for step, ang, sym in [(60, 30, None), (15, 30, None), (15, 10, None), (60, 30, 'rz180'), (30, 45, 'rz90')]:
    grid = make_rotation_grid_host(40, step)
    syms = np.eye(4)[None]
    if sym == 'rz180':
      syms = np.stack([np.eye(4), og.euler_matrix_z(np.pi)])
    if sym == 'rz90':
      syms = np.stack([og.euler_matrix_z(a) for a in (0, np.pi / 2, np.pi, 3 * np.pi / 2)])
    ref, ref_idx = og.cluster_poses(ang, 99999, grid, syms)
    out = mycpp.cluster_poses(ang, 99999, grid, syms)
    assert len(out) == len(ref)
    assert np.array_equal(np.asarray(out), ref)
    assert np.array_equal(mycpp.last_keep_indices(), ref_idx)
  # distance gate
  far = make_rotation_grid_host(40, 60).copy(); far[:, 0, 3] = np.arange(len(far)) % 7 * 0.5
  ref, _ = og.cluster_poses(60, 0.4, far, np.eye(4)[None])
  assert np.array_equal(np.asarray(mycpp.cluster_poses(60, 0.4, far, np.eye(4)[None])), ref)


@pytest.mark.parametrize('cfg', [1, 2])
def test_crop_params_bit_exact(cfg, scene1, scene2):
  from foundationpose_b200 import utils as U
  scene = scene1 if cfg == 1 else scene2
  poses, _ = hypo_poses(scene, jitter_seed=7)
  for ratio in (1.2, 1.1):
    ref = oo.crop_params(poses, scene['K'], scene['diameter'], ratio, 160, 480, 640)
    got = U.crop_params(poses, scene['K'], scene['diameter'], ratio, 160, 480, 640).cpu().numpy()
    assert np.array_equal(got[:, 0:9], ref['tf'])
    assert np.array_equal(got[:, 9:13], ref['bbox'])
    assert np.array_equal(got[:, 13:29], ref['mtx'])
    assert np.array_equal(got[:, 29:33], ref['rtf'])
    assert np.array_equal(got[:, 33:37], ref['inv'])
  tf = U.compute_crop_window_tf_batch(H=480, W=640, poses=torch.as_tensor(poses, device='cuda'), K=scene['K'],
                                      crop_ratio=1.2, out_size=(160, 160), method='box_3d',
                                      mesh_diameter=scene['diameter'])
  assert tf.shape == (len(poses), 3, 3)


@pytest.mark.parametrize('cfg,n', [(1, 252), (2, 24)])
def test_render_bit_exact(cfg, n, scene1, scene2):
  from foundationpose_b200 import utils as U
  scene = scene1 if cfg == 1 else scene2
  mesh, _ = centered(scene)
  mnp = mesh_np(mesh)
  poses, _ = hypo_poses(scene, n=n, jitter_seed=3)
  S = 160
  cp = oo.crop_params(poses, scene['K'], scene['diameter'], 1.2, S, 480, 640)
  color, xyz, tri = oo.render(poses, cp, mnp, S)
  mt = {k: torch.as_tensor(v, device='cuda') for k, v in mnp.items()}
  params = U.crop_params(poses, scene['K'], scene['diameter'], 1.2, S, 480, 640)
  for zthr in (0.001, 0.1):
    A, ex = U.render_crops(mt, poses, params, S, scene['diameter'], zthr, bf16=False, debug=True)
    assert np.array_equal(ex['tri'].cpu().numpy(), tri)          # triangle ids: bit-exact everywhere
    assert np.array_equal(ex['xyz'].cpu().numpy(), xyz)
    assert np.array_equal(ex['color'].cpu().numpy(), color)
    refA = oo.make_A(color, xyz, poses, scene['diameter'], zthr)
    A = A.cpu().numpy()
    assert np.array_equal(A[..., :6], refA) and (A[..., 6:] == 0).all()
  Ab, _ = U.render_crops(mt, poses, params, S, scene['diameter'], 0.001, bf16=True)
  ref16 = torch.from_numpy(oo.make_A(color, xyz, poses, scene['diameter'], 0.001)).bfloat16().float().numpy()
  assert np.array_equal(Ab[..., :6].float().cpu().numpy(), ref16)


def test_nvdiffrast_render_api(scene1):
  from foundationpose_b200 import utils as U
  mesh, _ = centered(scene1)
  mnp = mesh_np(mesh)
  poses, _ = hypo_poses(scene1, n=6)
  S = 160
  cp = oo.crop_params(poses, scene1['K'], scene1['diameter'], 1.2, S, 480, 640)
  color, xyz, tri = oo.render(poses, cp, mnp, S)
  extra = {}
  mt = {k: torch.as_tensor(v, device='cuda') for k, v in mnp.items()}
  c, d, nrm = U.nvdiffrast_render(K=scene1['K'], H=480, W=640, ob_in_cams=torch.as_tensor(poses, device='cuda'),
                                  mesh_tensors=mt, bbox2d=torch.as_tensor(cp['bbox'], device='cuda'), output_size=(S, S),
                                  use_light=True, extra=extra)
  # matrices come from torch matmul here (not the fused crop kernel): compare with a tolerance
  assert (np.abs(c.cpu().numpy() - color) > 1e-3).mean() < 2e-3
  assert (np.abs(extra['xyz_map'].cpu().numpy() - xyz) > 1e-4).mean() < 2e-3
  assert nrm.shape == (6, S, S, 3) and d.shape == (6, S, S)      # use_light implies get_normal (src/Utils.py:190-191)


@pytest.mark.parametrize('cfg', [1, 2])
@pytest.mark.parametrize('mode', [0, 1])
def test_warp_bit_exact(cfg, mode, scene1, scene2):
  from foundationpose_b200 import utils as U
  scene = scene1 if cfg == 1 else scene2
  poses, d = hypo_poses(scene, n=40, jitter_seed=9)
  poses[3, :3, 3] += np.float32(0.25)      # crop partly outside the image -> zero padding path
  S = 160
  K = scene['K']
  xyz_map = oo.depth2xyzmap(d, K)
  cp = oo.crop_params(poses, K, scene['diameter'], 1.2 if mode == 0 else 1.1, S, 480, 640)
  ref = oo.make_B(scene['rgb'].astype(np.float32), xyz_map, d, poses, cp, K, S, scene['diameter'], mode)
  params = U.crop_params(poses, K, scene['diameter'], 1.2 if mode == 0 else 1.1, S, 480, 640)
  rgb_u8 = torch.as_tensor(scene['rgb'], device='cuda')
  for rgb_t in (rgb_u8, rgb_u8.float()):
    B = U.warp_crops(rgb_t, torch.as_tensor(xyz_map, device='cuda'), torch.as_tensor(d, device='cuda'), poses, params,
                     K, S, scene['diameter'], mode, bf16=False).cpu().numpy()
    assert np.array_equal(B[..., :6], ref) and (B[..., 6:] == 0).all()
  assert (ref[3] == 0).mean() > 0.2 and np.abs(ref[..., 3:]).max() < 2


def test_pose_update_close():
  from foundationpose_b200 import utils as U
  rng = np.random.default_rng(4)
  poses, _ = None, None
  g = og.make_rotation_grid()
  g[:, :3, 3] = rng.normal(0, 0.3, (len(g), 3)).astype(np.float32)
  trans = rng.normal(0, 0.5, (len(g), 3)).astype(np.float32)
  rot = rng.normal(0, 1.5, (len(g), 3)).astype(np.float32)
  rot[0] = 0
  for nx in (True, False):
    ref = oo.pose_update(g, trans, rot, 0.21, 0.3490658504, [0.02, 0.02, 0.05], nx)
    got = U.pose_update(g, torch.as_tensor(trans, device='cuda'), torch.as_tensor(rot, device='cuda'), 0.21,
                        0.3490658504, [0.02, 0.02, 0.05], nx).cpu().numpy()
    assert np.abs(got - ref).max() < 2e-6     # sinf/cosf/tanhf: device vs numpy


@pytest.mark.parametrize('trans_rep,rot_rep', [('deepim', 'axis_angle'), ('tracknet', '6d'), ('deepim', '6d')])
def test_pose_update_config_gated_representations(trans_rep, rot_rep, scene2):
  """trans_rep='deepim' (predict_pose_refine.py:201-215) and rot_rep='6d' (:223-224) against the numpy restatement."""
  from foundationpose_b200 import utils as U
  rng = np.random.default_rng(5)
  poses, _ = hypo_poses(scene2, n=40, jitter_seed=2)
  K = scene2['K']
  S = 160
  cp = oo.crop_params(poses, K, scene2['diameter'], 1.2, S, 480, 640)
  params = U.crop_params(poses, K, scene2['diameter'], 1.2, S, 480, 640)
  trans = rng.normal(0, 0.05, (len(poses), 3)).astype(np.float32)
  trans[:, 2] = 1 + rng.normal(0, 0.02, len(poses)).astype(np.float32)          # deepim: depth ratio
  rot = rng.normal(0, 1.0, (len(poses), 6 if rot_rep == '6d' else 3)).astype(np.float32)
  for nx in (True, False):
    ref = oo.pose_update(poses, trans, rot, scene2['diameter'], 0.3490658504, [0.02, 0.02, 0.05], nx, trans_rep=trans_rep,
                         rot_rep=rot_rep, tf_to_crops=cp['tf'].reshape(-1, 3, 3), K=K, S=S)
    got = U.pose_update(poses, torch.as_tensor(trans, device='cuda'), torch.as_tensor(rot, device='cuda'),
                        scene2['diameter'], 0.3490658504, [0.02, 0.02, 0.05], nx, trans_rep=trans_rep, rot_rep=rot_rep,
                        crop_params=params, K=K, S=S).cpu().numpy()
    assert np.abs(got[:, :3, :3] - ref[:, :3, :3]).max() < 3e-6
    assert np.abs(got[:, :3, 3] - ref[:, :3, 3]).max() < 2e-5      # fp32 back-projection of ~300 px coordinates
    R = got[:, :3, :3].astype(np.float64)
    assert np.abs(R @ R.transpose(0, 2, 1) - np.eye(3)).max() < 1e-5


@pytest.mark.parametrize('env', [dict(FPB_RASTER_CULL_MIN='0'), dict(FPB_RENDER_LEAN='1'),
                                 dict(FPB_RASTER_CULL_MIN='0', FPB_RENDER_LEAN='1')])
def test_render_bit_exact_dense_mesh_paths(env):
  """Large / dense meshes take other kernels: cull-then-rasterise (raster_cull_kernel) and the lean vertex pass (only
  snapped positions are stored, rasteriser and shader recompute the corners they need).  Force them for every mesh (the
  switches are read once per process) and repeat the bit-exact render comparison against the oracle."""
  import os, subprocess, sys
  root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
  r = subprocess.run([sys.executable, '-m', 'pytest', os.path.join(root, 'tests', 'test_geometry_gpu.py'), '-m', 'gpu', '-q', '-x',
                      '-k', 'test_render_bit_exact and not dense_mesh_paths'], capture_output=True, text=True,
                     env=dict(os.environ, **env), cwd=root, timeout=900)
  assert r.returncode == 0 and '2 passed' in r.stdout, r.stdout[-1500:] + r.stderr[-500:]


@pytest.mark.parametrize('cfg', [1, 2])
def test_render_triangles_crossing_the_camera_plane(cfg, scene1, scene2):
  """Hypotheses so close that part of the mesh is behind the camera plane (w <= 0): nvdiffrast clips those triangles
  against the frustum; here they take the clip-free homogeneous coverage test (render.cu homog_cover) instead of being
  dropped.  Bit-exact against the oracle's restatement; the crossing triangles do contribute pixels."""
  from foundationpose_b200 import utils as U
  scene = scene1 if cfg == 1 else scene2
  mesh, _ = centered(scene)
  mnp = mesh_np(mesh)
  poses, _ = hypo_poses(scene, n=6)
  poses = poses.copy()
  poses[:, :3, 3] = np.array([0.01, -0.005, 0.02], np.float32)           # camera inside the object's bounding sphere
  poses[3, :3, 3] = np.array([0.0, 0.0, -0.2], np.float32)               # entirely behind the camera: nothing to draw
  S = 160
  cp = oo.crop_params(poses, scene['K'], scene['diameter'], 1.2, S, 480, 640)
  color, xyz, tri = oo.render(poses, cp, mnp, S)
  zc = (np.asarray(mnp['pos']) @ poses[0, :3, :3].T + poses[0, :3, 3])[:, 2]
  assert (zc <= 0).any() and (zc > 0).any()                               # the mesh does straddle the camera plane
  mt = {k: torch.as_tensor(v, device='cuda') for k, v in mnp.items()}
  params = U.crop_params(torch.as_tensor(poses).cuda(), scene['K'], scene['diameter'], 1.2, S, 480, 640)
  _, ex = U.render_crops(mt, torch.as_tensor(poses).cuda(), params, S, scene['diameter'], 0.001, bf16=False, debug=True)
  assert np.array_equal(ex['tri'].cpu().numpy(), tri)
  assert np.array_equal(ex['color'].cpu().numpy(), color) and np.array_equal(ex['xyz'].cpu().numpy(), xyz)
  assert (tri[3] == 0).all() and (tri[0] > 0).mean() > 0.2
  # faces that cross the plane are among the visible ones
  faces = np.asarray(mnp['faces'])
  crossing = ((zc[faces] <= 0).any(1)) & ((zc[faces] > 0).any(1))
  vis = np.unique(tri[0][tri[0] > 0]) - 1
  assert crossing.any()
  if cfg == 1:                       # the cube's 5 cm faces reach far beyond the near plane; the blob's millimetre triangles that
    assert crossing[vis].any()       # cross w = 0 lie entirely in front of the near plane (z < 1 mm) and are clipped away
