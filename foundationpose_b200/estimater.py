"""FoundationPose -- drop-in for src/estimater.py:18-268 on top of libfpb200.so.

register()/track_one() keep every intermediate on the device: depth stencils, masked-depth statistics, xyz map,
render/warp/refine loop and scoring; the only device->host reads are the 8 statistics of guess_translation and the
final 4x4 pose (the reference crosses the PCIe bus ~10 times per call).
"""
import logging
import os

import numpy as np
import torch

from . import mycpp
from . import utils as U
from .predict_pose_refine import PoseRefinePredictor
from .predict_score import ScorePredictor


def make_rotation_grid_host(min_n_views=40, inplane_step=60):
  """src/estimater.py:106-118 before clustering: float64 [n_views*n_inplane, 4, 4] (host numpy, init-time)."""
  cam_in_obs = U.sample_views_icosphere(n_views=min_n_views)
  rot_grid = []
  for i in range(len(cam_in_obs)):
    for inplane_rot in np.deg2rad(np.arange(0, 360, inplane_step)):
      cam_in_ob = cam_in_obs[i] @ U.euler_matrix(0, 0, inplane_rot)
      rot_grid.append(np.linalg.inv(cam_in_ob))
  return np.asarray(rot_grid)


def guess_translation_cuda(depth_t, mask_t, K):
  """src/estimater.py:137-156 with the reductions (bbox of the mask, exact median of the valid depths) on the
  device; the 3x3 float64 tail stays on the host.  Returns np.float64 [3]."""
  if mask_t.dtype != torch.uint8:
    mask_t = (mask_t > 0).to(torch.uint8)
  return _center_from_stats(U.mask_depth_stats(depth_t.contiguous(), mask_t.contiguous()), K)


def _center_from_stats(stats, K):
  umin, umax, vmin, vmax, n_mask, n_valid, lo, hi = stats
  if n_mask == 0:
    logging.info('mask is all zero')
    return np.zeros((3))
  uc = (umin + umax) / 2.0
  vc = (vmin + vmax) / 2.0
  if n_valid == 0:
    logging.info('valid is empty')
    return np.zeros((3))
  zc = np.float32(np.float32(lo + hi) / np.float32(2.0))      # np.median of float32 data (mean of the two middles)
  center = (np.linalg.inv(K) @ np.asarray([uc, vc, 1]).reshape(3, 1)) * zc
  return center.reshape(3)


def _h2d(x, dtype=None):
  """Frame buffer -> cuda tensor.  Page-locked host memory is copied asynchronously on the current stream (the kernels
  that read it are enqueued behind the copy and register() / track_one() end with a synchronisation, so the caller's
  buffer is free again when they return); pageable memory takes torch's synchronous path."""
  t = torch.as_tensor(x)
  if not t.is_cuda:
    t = t.to('cuda', non_blocking=t.is_pinned())
  if dtype is not None and t.dtype != dtype:
    t = t.to(dtype)
  return t.contiguous()


class FoundationPose:
  def __init__(self, model_pts, model_normals, symmetry_tfs=None, mesh=None, scorer: ScorePredictor = None,
               refiner: PoseRefinePredictor = None, glctx=None, debug=0, debug_dir='/tmp/fpb200_debug/',
               mesh_diameter=None):
    self.gt_pose = None
    self.ignore_normal_flip = True
    self.debug = debug
    self.debug_dir = debug_dir
    if debug >= 1:
      os.makedirs(debug_dir, exist_ok=True)
    self._diameter_override = mesh_diameter
    self.reset_object(model_pts, model_normals, symmetry_tfs=symmetry_tfs, mesh=mesh)
    self.make_rotation_grid(min_n_views=40, inplane_step=60)
    self.glctx = glctx      # accepted and ignored: there is no nvdiffrast context
    self.scorer = scorer if scorer is not None else ScorePredictor()
    self.refiner = refiner if refiner is not None else PoseRefinePredictor()
    self.pose_last = None   # used for tracking; w.r.t. the centred mesh

  def reset_object(self, model_pts, model_normals, symmetry_tfs=None, mesh=None):
    """src/estimater.py:44-78 (open3d's voxel down-sampling, which only feeds utilities off the hot path, restated in
    numpy)."""
    if isinstance(mesh, dict):                 # a ready mesh_tensors dict
      mesh = U.mesh_from_tensors(mesh)
    elif not hasattr(mesh, 'copy'):            # duck-typed mesh without .copy()
      mesh = U.MeshLike(mesh.vertices, mesh.faces, mesh.vertex_normals, getattr(mesh, 'visual', None))
    verts = np.asarray(mesh.vertices)
    max_xyz = verts.max(axis=0)
    min_xyz = verts.min(axis=0)
    self.model_center = (min_xyz + max_xyz) / 2
    self.mesh_ori = mesh.copy()
    mesh = mesh.copy()
    mesh.vertices = np.asarray(mesh.vertices) - self.model_center.reshape(1, 3)
    model_pts = np.asarray(mesh.vertices)
    if self._diameter_override is not None:
      self.diameter = float(self._diameter_override)
    else:
      self.diameter = U.compute_mesh_diameter(model_pts=model_pts, n_sample=10000)
    self.vox_size = max(self.diameter / 20.0, 0.003)
    logging.info(f'self.diameter:{self.diameter}, vox_size:{self.vox_size}')
    self.dist_bin = self.vox_size / 2
    self.angle_bin = 20
    # open3d 0.18 PointCloud::VoxelDownSample (src/estimater.py:59-60): voxel grid anchored half a voxel below the minimum
    # bound, index = floor((p - (min_bound - voxel/2)) / voxel), points and normals averaged per occupied voxel.  open3d emits
    # the voxels in hash-map order (unspecified); here they come in lexicographic voxel order.
    vmin = model_pts.min(axis=0) - 0.5 * self.vox_size
    keys = np.floor((model_pts - vmin) / self.vox_size).astype(np.int64)
    _, inv = np.unique(keys, axis=0, return_inverse=True)
    inv = inv.reshape(-1)
    cnt = np.bincount(inv).astype(np.float64)
    pts = np.stack([np.bincount(inv, weights=model_pts[:, k]) / cnt for k in range(3)], 1)
    nrm_src = np.asarray(model_normals) if model_normals is not None else np.asarray(mesh.vertex_normals)
    nrm = np.stack([np.bincount(inv, weights=nrm_src[:, k]) / cnt for k in range(3)], 1)
    self.max_xyz = pts.max(axis=0)
    self.min_xyz = pts.min(axis=0)
    dev = 'cuda' if torch.cuda.is_available() else 'cpu'
    self.pts = torch.tensor(pts, dtype=torch.float32, device=dev)
    self.normals = torch.nn.functional.normalize(torch.tensor(nrm, dtype=torch.float32, device=dev), dim=-1)
    self.mesh_path = None
    self.mesh = mesh
    self.mesh_tensors = U.make_mesh_tensors(self.mesh, device=dev)
    if symmetry_tfs is None:
      self.symmetry_tfs = torch.eye(4).float().to(dev)[None]
    else:
      self.symmetry_tfs = torch.as_tensor(symmetry_tfs, device=dev, dtype=torch.float)
    logging.info('reset done')

  def get_tf_to_centered_mesh(self, clone=True):
    # cached per model_center: building it costs a pageable host-to-device copy, which would synchronise register()
    key = tuple(np.asarray(self.model_center, dtype=np.float64).reshape(-1).tolist())
    if getattr(self, '_tf_to_center', None) is None or self._tf_to_center[0] != key:
      tf_to_center = torch.eye(4, dtype=torch.float, device='cuda')
      tf_to_center[:3, 3] = -torch.as_tensor(self.model_center, device='cuda', dtype=torch.float)
      self._tf_to_center = (key, tf_to_center)
    return self._tf_to_center[1].clone() if clone else self._tf_to_center[1]

  def to_device(self, s='cuda:0'):
    for k in self.__dict__:
      if torch.is_tensor(self.__dict__[k]):
        self.__dict__[k] = self.__dict__[k].to(s)
    for k in self.mesh_tensors:
      if torch.is_tensor(self.mesh_tensors[k]):
        self.mesh_tensors[k] = self.mesh_tensors[k].to(s)

  def make_rotation_grid(self, min_n_views=40, inplane_step=60):
    """src/estimater.py:106-124"""
    rot_grid = make_rotation_grid_host(min_n_views, inplane_step)
    logging.info(f'rot_grid:{rot_grid.shape}')
    rot_grid = mycpp.cluster_poses(30, 99999, rot_grid, self.symmetry_tfs.data.cpu().numpy())
    rot_grid = np.asarray(rot_grid)
    logging.info(f'after cluster, rot_grid:{rot_grid.shape}')
    self.rot_grid = torch.as_tensor(rot_grid, device='cuda', dtype=torch.float)

  def generate_random_pose_hypo(self, K, rgb, depth, mask, scene_pts=None):
    """src/estimater.py:127-134"""
    ob_in_cams = self.rot_grid.clone()
    center = self.guess_translation(depth=depth, mask=mask, K=K)
    ob_in_cams[:, :3, 3] = torch.tensor(center, device='cuda', dtype=torch.float).reshape(1, 3)
    return ob_in_cams

  def guess_translation(self, depth, mask, K):
    """src/estimater.py:137-156 (numpy or cuda inputs)."""
    depth_t = torch.as_tensor(depth, device='cuda', dtype=torch.float)
    mask_t = torch.as_tensor(mask, device='cuda')
    return guess_translation_cuda(depth_t, mask_t, np.asarray(K, dtype=np.float64))

  @torch.inference_mode()
  def register(self, K, rgb, depth, ob_mask, ob_id=None, glctx=None, iteration=5):
    """src/estimater.py:159-240 -> np.ndarray [4,4] float32 (object pose in camera, un-centred mesh frame)."""
    U.set_seed(0)
    K = np.asarray(K, dtype=np.float64)
    depth_t = _h2d(depth, torch.float)
    rgb_t = _h2d(rgb)
    mask_t = _h2d(ob_mask)
    if mask_t.dtype == torch.bool:
      mask_t = mask_t.view(torch.uint8)          # bool is one byte holding 0 / 1: reinterpret, no kernel
    elif mask_t.dtype != torch.uint8:
      mask_t = (mask_t > 0).to(torch.uint8)
    mask_t = mask_t.contiguous()
    # The whole call is enqueued without a host round trip (the reference crosses the bus ~10 times): the translation
    # guess and the hypotheses are produced on the device, the mask statistics travel to pinned host memory
    # asynchronously and the "too few valid pixels" early exit (src/estimater.py:176-180) is evaluated after the one
    # synchronisation at the end (in that case the enqueued work ran on a dummy translation and is discarded).
    depth_t, xyz_map, poses, si, sf, _ = U.register_prologue(depth_t, mask_t, K, self.rot_grid, min_valid=4)
    if getattr(self, '_stats_host', None) is None:
      self._stats_host = (torch.empty(6, dtype=torch.int32, pin_memory=True), torch.empty(2, dtype=torch.float, pin_memory=True))
    self._stats_host[0].copy_(si, non_blocking=True)
    self._stats_host[1].copy_(sf, non_blocking=True)
    poses, _ = self.refiner.predict(mesh=self.mesh, mesh_tensors=self.mesh_tensors, rgb=rgb_t, depth=depth_t, K=K,
                                    ob_in_cams=poses, normal_map=None, xyz_map=xyz_map, glctx=self.glctx,
                                    mesh_diameter=self.diameter, iteration=iteration, get_vis=False,
                                    same_translation=True)      # hypotheses = rot_grid x ONE guessed translation
    scores, _ = self.scorer.predict(mesh=self.mesh, rgb=rgb_t, depth=depth_t, K=K, ob_in_cams=poses, normal_map=None,
                                    mesh_tensors=self.mesh_tensors, glctx=self.glctx, mesh_diameter=self.diameter,
                                    get_vis=False)
    ids, scores, poses, best = U.rank_poses(scores, poses, self.get_tf_to_centered_mesh(clone=False))
    best_pose = best.data.cpu().numpy()                                                # the only synchronisation
    oi, of = self._stats_host[0].numpy(), self._stats_host[1].numpy()
    stats = (int(oi[0]), int(oi[1]), int(oi[2]), int(oi[3]), int(oi[4]), int(oi[5]), of[0], of[1])
    if stats[5] < 4:
      logging.info('valid too small, return')
      pose = np.eye(4)
      pose[:3, 3] = _center_from_stats(stats, K)
      return pose
    self.H, self.W = depth_t.shape[:2]
    self.K = K
    self.ob_id = ob_id
    self.ob_mask = ob_mask
    self.pose_last = poses[0]
    self.best_id = ids[0]
    self.poses = poses
    self.scores = scores
    return best_pose

  def compute_add_err_to_gt_pose(self, poses):
    return -torch.ones(len(poses), device='cuda', dtype=torch.float)

  @torch.inference_mode()
  def track_one(self, rgb, depth, K, iteration, extra={}):
    """src/estimater.py:250-268"""
    if self.pose_last is None:
      logging.info('Please init pose by register first')
      raise RuntimeError
    K = np.asarray(K, dtype=np.float64)
    depth_t = _h2d(depth, torch.float)
    depth_t = U.erode_depth(depth_t, radius=2, device='cuda')
    depth_t = U.bilateral_filter_depth(depth_t, radius=2, device='cuda')
    xyz_map = U.depth2xyzmap_cuda(depth_t, K.astype(np.float32).astype(np.float64), math_f64=False, zfar=np.inf)
    pose, _ = self.refiner.predict(mesh=self.mesh, mesh_tensors=self.mesh_tensors, rgb=_h2d(rgb),
                                   depth=depth_t, K=K, ob_in_cams=self.pose_last.reshape(1, 4, 4), normal_map=None,
                                   xyz_map=xyz_map, mesh_diameter=self.diameter, glctx=self.glctx, iteration=iteration,
                                   get_vis=False)
    self.pose_last = pose
    return (pose @ self.get_tf_to_centered_mesh()).data.cpu().numpy().reshape(4, 4)

  # ------------------------------------------------------------------------------------------------------------
  # track_one as one CUDA graph: the video path (1 hypothesis, 2 iterations) is ~90 small launches whose device time
  # is far below the host time needed to enqueue them, so the frame body is captured once over static buffers and
  # replayed per frame.  Same arithmetic and same C-ABI calls as track_one (captured, not re-implemented).
  # ------------------------------------------------------------------------------------------------------------
  def _capture_track_graph(self, key, H, W, K, iteration, rgb_dtype):
    dev = 'cuda'
    g = dict(key=key)
    g['rgb'] = torch.zeros((H, W, 3), dtype=rgb_dtype, device=dev)
    g['depth'] = torch.zeros((H, W), dtype=torch.float, device=dev)
    g['pose_in'] = torch.eye(4, dtype=torch.float, device=dev)
    g['pose_out'] = torch.eye(4, dtype=torch.float, device=dev)
    g['result'] = torch.eye(4, dtype=torch.float, device=dev)
    g['pose_in'].copy_(self.pose_last.reshape(4, 4))
    tf = g['tf'] = self.get_tf_to_centered_mesh()     # read by the graph at every replay: must stay alive with it
    K32 = K.astype(np.float32).astype(np.float64)

    def body():
      d = U.erode_depth(g['depth'], radius=2, device='cuda')
      d = U.bilateral_filter_depth(d, radius=2, device='cuda')
      xyz_map = U.depth2xyzmap_cuda(d, K32, math_f64=False, zfar=np.inf)
      pose, _ = self.refiner.predict(mesh=self.mesh, mesh_tensors=self.mesh_tensors, rgb=g['rgb'], depth=d, K=K,
                                     ob_in_cams=g['pose_in'].reshape(1, 4, 4), normal_map=None, xyz_map=xyz_map,
                                     mesh_diameter=self.diameter, glctx=self.glctx, iteration=iteration, get_vis=False)
      g['pose_out'].copy_(pose.reshape(4, 4))
      g['result'].copy_(pose.reshape(4, 4) @ tf)

    # buffers the graph reads must belong to it: the predictor's cached input buffer / workspace may be replaced by a
    # later register() with another batch size, so the capture allocates its own from the graph's memory pool
    saved = (self.refiner._ab, self.refiner.model._ws)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    try:
      with torch.cuda.stream(side):
        for _ in range(2):          # eager warm-up on the capture stream: lazy one-time initialisation happens here
          body()
      torch.cuda.current_stream().wait_stream(side)
      torch.cuda.synchronize()
      self.refiner._ab, self.refiner.model._ws = None, None
      graph = torch.cuda.CUDAGraph()
      with torch.cuda.graph(graph, stream=side):
        body()
    finally:
      self.refiner._ab, self.refiner.model._ws = saved
    g['graph'] = graph
    self._track_graph = g
    return g

  @torch.inference_mode()
  def track_one_graphed(self, rgb, depth, K, iteration, extra={}):
    """track_one (src/estimater.py:250-268) with the frame body replayed from a captured CUDA graph.  rgb / depth:
    numpy, pinned or device tensors of a fixed [H,W] and dtype; the graph is re-captured when the shape, the
    intrinsics, the iteration count or the object changes."""
    if self.pose_last is None:
      logging.info('Please init pose by register first')
      raise RuntimeError
    K = np.asarray(K, dtype=np.float64)
    rgb_t = torch.as_tensor(rgb)
    depth_t = torch.as_tensor(depth)
    H, W = depth_t.shape[:2]
    rgb_dtype = torch.uint8 if rgb_t.dtype == torch.uint8 else torch.float
    key = (H, W, int(iteration), K.tobytes(), id(self.mesh_tensors), rgb_dtype, float(self.diameter))
    g = getattr(self, '_track_graph', None)
    if g is None or g['key'] != key:
      g = self._capture_track_graph(key, H, W, K, int(iteration), rgb_dtype)
    g['rgb'].copy_(rgb_t, non_blocking=True)
    g['depth'].copy_(depth_t, non_blocking=True)
    g['pose_in'].copy_(self.pose_last.reshape(4, 4), non_blocking=True)
    g['graph'].replay()
    out = g['result'].cpu().numpy().reshape(4, 4)          # the only synchronisation
    self.pose_last = g['pose_out'].clone().reshape(1, 4, 4)
    return out

