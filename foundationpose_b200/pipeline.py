"""Pipeline surface of the reference's README (readme.md:126-185): `Pipeline`, `Processor`, `PipelineData` and the
`FoundationPoseEstimator` processor.  The sources of `src/pipeline/*` and `src/modules/*` are not part of the
reference snapshot, so this is the minimal contract the README shows -- a thin, headless adapter over
`FoundationPose.register`; detection / segmentation / NetworkTables processors are out of scope.

  pipeline = Pipeline("pose")
  pipeline.add_processor(FoundationPoseEstimator(mesh_or_obj_path, K_or_intrinsics_file))
  result = pipeline.run(PipelineData(rgb=rgb, depth=depth, mask=mask))
  result.pose      # np.ndarray [4,4]  object pose in the camera frame
  result.pose_6d   # (x, y, z, roll, pitch, yaw) metres / radians, Euler convention of src/transform.py:50-68
"""
import logging
import math
import re
from types import SimpleNamespace

import numpy as np


class PipelineData:
  """Mutable attribute bag passed from processor to processor."""

  def __init__(self, rgb=None, depth=None, mask=None, K=None, mesh=None, pose=None, pose_6d=None, **extra):
    self.rgb, self.depth, self.mask, self.K, self.mesh = rgb, depth, mask, K, mesh
    self.pose, self.pose_6d = pose, pose_6d
    self.errors = []
    for k, v in extra.items():
      setattr(self, k, v)


class Processor:
  """Base class: `process(data) -> data`, optional `visualize(data)` (readme.md:161-178)."""

  def process(self, data):
    raise NotImplementedError

  def visualize(self, data):
    return None


class Pipeline:
  def __init__(self, name='pipeline', stop_on_error=False, visualize=False):
    self.name, self.stop_on_error, self.show = name, stop_on_error, visualize
    self.processors = []

  def add_processor(self, p):
    if not callable(getattr(p, 'process', None)):
      raise TypeError('a processor needs a process(data) method')
    self.processors.append(p)
    return self

  def run(self, data):
    for p in self.processors:
      try:
        out = p.process(data)
        data = data if out is None else out
        if self.show:
          p.visualize(data)
      except Exception as e:  # --stop-on-error (readme.md:120): otherwise record and continue
        if self.stop_on_error:
          raise
        logging.exception('%s: processor %s failed', self.name, type(p).__name__)
        data.errors.append((type(p).__name__, e))
    return data


def load_camera_intrinsics(path, camera_section='LEFT_CAM_FHD1200'):
  """Both formats of src/file_processing.py:11-80: a 3x3 matrix (three whitespace-separated rows) or the ZED style
  `[SECTION] fx= fy= cx= cy=` file.  Raises ValueError instead of returning None."""
  with open(path, 'r') as f:
    text = f.read()
  if text.lstrip().startswith('['):
    m = re.search(r'\[' + re.escape(camera_section) + r'\](.*?)(?=\[|$)', text, re.DOTALL)
    if not m:
      raise ValueError(f'camera section {camera_section} not found in {path}')
    v = {}
    for key in ('fx', 'fy', 'cx', 'cy'):
      mm = re.search(key + r'\s*=\s*([-+\d.eE]+)', m.group(1))
      if not mm:
        raise ValueError(f'{key} missing in section {camera_section} of {path}')
      v[key] = float(mm.group(1))
    return np.array([[v['fx'], 0, v['cx']], [0, v['fy'], v['cy']], [0, 0, 1]], dtype=np.float64)
  rows = [ln.split() for ln in text.splitlines() if ln.strip()]
  if len(rows) < 3 or any(len(r) < 3 for r in rows[:3]):
    raise ValueError(f'{path}: expected a 3x3 matrix')
  return np.array([[float(x) for x in r[:3]] for r in rows[:3]], dtype=np.float64)


def load_obj(path):
  """Minimal Wavefront reader (v / vn / f, polygons fan-triangulated) -> object with the attributes the drop-in
  reads from a trimesh.Trimesh: vertices, faces, vertex_normals, visual (trimesh itself is absent offline).
  Vertex normals are area-weighted face-normal averages when the file has none per vertex."""
  vs, fs = [], []
  with open(path, 'r') as f:
    for ln in f:
      t = ln.split()
      if not t:
        continue
      if t[0] == 'v':
        vs.append([float(x) for x in t[1:4]])
      elif t[0] == 'f':
        idx = [int(w.split('/')[0]) for w in t[1:]]
        idx = [i - 1 if i > 0 else len(vs) + i for i in idx]
        for k in range(1, len(idx) - 1):
          fs.append([idx[0], idx[k], idx[k + 1]])
  V = np.asarray(vs, dtype=np.float64).reshape(-1, 3)
  F = np.asarray(fs, dtype=np.int64).reshape(-1, 3)
  if len(V) == 0 or len(F) == 0:
    raise ValueError(f'{path}: no geometry')
  fn = np.cross(V[F[:, 1]] - V[F[:, 0]], V[F[:, 2]] - V[F[:, 0]])
  vn = np.zeros_like(V)
  for k in range(3):
    np.add.at(vn, F[:, k], fn)
  vn /= np.maximum(np.linalg.norm(vn, axis=1, keepdims=True), 1e-12)
  grey = np.tile(np.array([[128, 128, 128, 255]], dtype=np.uint8), (len(V), 1))   # untextured CAD part
  from .utils import MeshLike
  return MeshLike(V, F, vn, SimpleNamespace(vertex_colors=grey, uv=None, material=None))


def pose_to_6d(pose):
  """4x4 pose -> (x, y, z, roll, pitch, yaw), metres / radians; Euler angles as in src/transform.py:50-68."""
  R = np.asarray(pose, dtype=np.float64)[:3, :3]
  x, y, z = (float(v) for v in np.asarray(pose)[:3, 3])
  if abs(R[2, 0]) > 0.9999:
    yaw = 0.0
    if R[2, 0] < 0:
      pitch, roll = math.pi / 2, math.atan2(R[0, 1], R[1, 1])
    else:
      pitch, roll = -math.pi / 2, -math.atan2(R[0, 1], R[1, 1])
  else:
    pitch = -math.asin(R[2, 0])
    c = math.cos(pitch)
    roll = math.atan2(R[2, 1] / c, R[2, 2] / c)
    yaw = math.atan2(R[1, 0] / c, R[0, 0] / c)
  return (x, y, z, roll, pitch, yaw)


class FoundationPoseEstimator(Processor):
  """`FoundationPoseEstimator(mesh_file, cam_intrinsics_file)` (readme.md:143): reads data.rgb / data.depth /
  data.mask (and data.K / data.mesh when the constructor got none), writes data.pose and data.pose_6d.
  `mesh_file` may be an .obj path, a trimesh-like object or a ready mesh_tensors dict; `cam_intrinsics_file` a path
  or a 3x3 array.  `estimator` injects a ready FoundationPose (tests use a stub); otherwise one is built lazily on
  the first frame with the default ScorePredictor / PoseRefinePredictor."""

  def __init__(self, mesh_file=None, cam_intrinsics_file=None, camera_section='LEFT_CAM_FHD1200', est_refine_iter=5,
               debug=0, estimator=None, scorer=None, refiner=None):
    self.mesh = load_obj(mesh_file) if isinstance(mesh_file, str) else mesh_file
    if isinstance(cam_intrinsics_file, str):
      self.K = load_camera_intrinsics(cam_intrinsics_file, camera_section)
    else:
      self.K = None if cam_intrinsics_file is None else np.asarray(cam_intrinsics_file, dtype=np.float64).reshape(3, 3)
    self.iteration, self.debug = int(est_refine_iter), debug
    self.est, self.scorer, self.refiner = estimator, scorer, refiner

  def _estimator(self, mesh):
    if self.est is None:
      from .estimater import FoundationPose
      from .predict_pose_refine import PoseRefinePredictor
      from .predict_score import ScorePredictor
      if isinstance(mesh, dict):              # a ready mesh_tensors dict: rebuild the trimesh-like view reset_object needs
        from .utils import mesh_from_tensors
        mesh = mesh_from_tensors(mesh)
      pts, nrm = np.asarray(mesh.vertices), np.asarray(mesh.vertex_normals)
      self.est = FoundationPose(model_pts=pts, model_normals=nrm, mesh=mesh, scorer=self.scorer or ScorePredictor(),
                                refiner=self.refiner or PoseRefinePredictor(), glctx=None, debug=self.debug)
    return self.est

  def process(self, data):
    mesh = self.mesh if self.mesh is not None else data.mesh
    K = self.K if self.K is not None else data.K
    if mesh is None or K is None:
      raise ValueError('FoundationPoseEstimator needs a mesh and camera intrinsics')
    if data.rgb is None or data.depth is None or data.mask is None:
      raise ValueError('FoundationPoseEstimator needs data.rgb, data.depth and data.mask')
    est = self._estimator(mesh)
    mask = np.asarray(data.mask).astype(bool) if isinstance(data.mask, np.ndarray) else data.mask
    data.pose = np.asarray(est.register(K=np.asarray(K, dtype=np.float64), rgb=data.rgb, depth=data.depth, ob_mask=mask,
                                        iteration=self.iteration)).reshape(4, 4)
    data.pose_6d = pose_to_6d(data.pose)
    data.K, data.mesh = K, mesh
    return data
