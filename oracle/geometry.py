"""CPU ORACLE (test infrastructure, NOT product code) -- numpy restatement of the
host-side geometry on FoundationPose's hypothesis path.

PARITY UNPINNED: the reference has no tests / golden vectors, and trimesh,
transformations, pytorch3d are absent offline; each function restates the
published algorithm of the pinned version and cites the reference call site.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this package.
"""
import ctypes
import math
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build_lib(force=False):
  """Compile oracle/fp_oracle.c -> oracle/_build/libfporacle.so (gcc, OpenMP if available)."""
  import subprocess
  out = os.path.join(_HERE, '_build', 'libfporacle.so')
  src = os.path.join(_HERE, 'fp_oracle.c')
  if not force and os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(src):
    return out
  os.makedirs(os.path.dirname(out), exist_ok=True)
  base = ['-O2', '-fPIC', '-shared', '-ffp-contract=off', '-fno-fast-math', '-o', out, src, '-lm']
  for cc in ('/usr/bin/gcc', 'gcc', 'cc'):
    for omp in (['-fopenmp'], []):
      try:
        subprocess.run([cc] + omp + base, check=True, capture_output=True)
        return out
      except Exception:
        continue
  raise RuntimeError('could not build the oracle library')


def lib():
  global _LIB
  if _LIB is None:
    _LIB = ctypes.CDLL(build_lib())
    _LIB.orc_cluster_poses.restype = ctypes.c_int
  return _LIB


def _p(a):
  return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _f32(a):
  return np.ascontiguousarray(a, dtype=np.float32)


# ----------------------------------------------------------------------------
# trimesh 4.2.2 creation.icosphere (requirements.txt:43), SURVEY appendix A
# ----------------------------------------------------------------------------
def icosahedron():
  t = (1.0 + 5.0 ** 0.5) / 2.0
  v = np.array([-1, t, 0, 1, t, 0, -1, -t, 0, 1, -t, 0, 0, -1, t, 0, 1, t, 0, -1, -t, 0, 1, -t,
                t, 0, -1, t, 0, 1, -t, 0, -1, -t, 0, 1], dtype=np.float64).reshape(-1, 3) / np.sqrt(2.0 + t)
  f = np.array([0, 11, 5, 0, 5, 1, 0, 1, 7, 0, 7, 10, 0, 10, 11, 1, 5, 9, 5, 11, 4, 11, 10, 2, 10, 7, 6,
                7, 1, 8, 3, 9, 4, 3, 4, 2, 3, 2, 6, 3, 6, 8, 3, 8, 9, 4, 9, 5, 2, 4, 11, 6, 2, 10,
                8, 6, 7, 9, 8, 1], dtype=np.int64).reshape(-1, 3)
  return v, f


def _subdivide(vertices, faces):
  edges = np.sort(faces[:, [0, 1, 1, 2, 2, 0]].reshape(-1, 2), axis=1)
  key = edges[:, 0] ^ (edges[:, 1] << 32)
  _, unique, inverse = np.unique(key, return_index=True, return_inverse=True)
  mid = vertices[edges[unique]].mean(axis=1)
  mid_idx = inverse.reshape(-1, 3) + len(vertices)
  f = np.column_stack([faces[:, 0], mid_idx[:, 0], mid_idx[:, 2],
                       mid_idx[:, 0], faces[:, 1], mid_idx[:, 1],
                       mid_idx[:, 2], mid_idx[:, 1], faces[:, 2],
                       mid_idx[:, 0], mid_idx[:, 1], mid_idx[:, 2]]).reshape(-1, 3)
  return np.vstack((vertices, mid)), f


def icosphere(subdivisions=1, radius=1.0):
  v, f = icosahedron()
  for _ in range(int(subdivisions)):
    v, f = _subdivide(v, f)
    scalar = np.sqrt(np.dot(v ** 2, [1, 1, 1]))
    unit = v / scalar.reshape(-1, 1)
    v = v + unit * (radius - scalar).reshape(-1, 1)
  return v, f


def sample_views_icosphere(n_views, subdivisions=None, radius=1):
  """reference: src/Utils.py:483-507"""
  if subdivisions is not None:
    verts, _ = icosphere(subdivisions, radius)
  else:
    sub = 1
    while True:
      verts, _ = icosphere(sub, radius)
      if verts.shape[0] >= n_views:
        break
      sub += 1
  cam_in_obs = np.tile(np.eye(4)[None], (len(verts), 1, 1))
  cam_in_obs[:, :3, 3] = verts
  up = np.array([0, 0, 1])
  z_axis = -cam_in_obs[:, :3, 3]
  z_axis /= np.linalg.norm(z_axis, axis=-1).reshape(-1, 1)
  x_axis = np.cross(up.reshape(1, 3), z_axis)
  invalid = (x_axis == 0).all(axis=-1)
  x_axis[invalid] = [1, 0, 0]
  x_axis /= np.linalg.norm(x_axis, axis=-1).reshape(-1, 1)
  y_axis = np.cross(z_axis, x_axis)
  y_axis /= np.linalg.norm(y_axis, axis=-1).reshape(-1, 1)
  cam_in_obs[:, :3, 0] = x_axis
  cam_in_obs[:, :3, 1] = y_axis
  cam_in_obs[:, :3, 2] = z_axis
  return cam_in_obs


def euler_matrix_z(a):
  """transformations 2024.6.1 euler_matrix(0,0,a,'sxyz') (src/estimater.py:113)."""
  si, sj, sk = math.sin(0.0), math.sin(0.0), math.sin(a)
  ci, cj, ck = math.cos(0.0), math.cos(0.0), math.cos(a)
  cc, cs = ci * ck, ci * sk
  sc, ss = si * ck, si * sk
  M = np.identity(4)
  M[0, 0] = cj * ck
  M[0, 1] = sj * sc - cs
  M[0, 2] = sj * cc + ss
  M[1, 0] = cj * sk
  M[1, 1] = sj * ss + cc
  M[1, 2] = sj * cs - sc
  M[2, 0] = -sj
  M[2, 1] = cj * si
  M[2, 2] = cj * ci
  return M


def make_rotation_grid_f64(min_n_views=40, inplane_step=60):
  """reference: src/estimater.py:106-118 (before clustering); float64 [n,4,4]."""
  cam_in_obs = sample_views_icosphere(n_views=min_n_views)
  rot_grid = []
  for i in range(len(cam_in_obs)):
    for inplane_rot in np.deg2rad(np.arange(0, 360, inplane_step)):
      cam_in_ob = cam_in_obs[i] @ euler_matrix_z(inplane_rot)
      rot_grid.append(np.linalg.inv(cam_in_ob))
  return np.asarray(rot_grid)


def cluster_poses(angle_diff, dist_diff, poses_in, symmetry_tfs):
  """reference: mycpp/src/app/pybind_api.cpp:24-68; returns (kept poses f32, kept indices)."""
  poses = _f32(np.asarray(poses_in).reshape(-1, 4, 4))   # pybind narrows f64 -> Matrix4f
  sym = _f32(np.asarray(symmetry_tfs).reshape(-1, 4, 4))
  keep = np.zeros(len(poses), dtype=np.int32)
  n = lib().orc_cluster_poses(ctypes.c_float(angle_diff), ctypes.c_float(dist_diff), _p(poses),
                              len(poses), _p(sym), len(sym), _p(keep))
  keep = keep[:n].copy()
  return poses[keep], keep


def make_rotation_grid(min_n_views=40, inplane_step=60, symmetry_tfs=None, angle_diff=30.0):
  """reference: src/estimater.py:106-124; float32 [n,4,4] after cluster_poses(30, 99999)."""
  g = make_rotation_grid_f64(min_n_views, inplane_step)
  if symmetry_tfs is None:
    symmetry_tfs = np.eye(4)[None]
  kept, _ = cluster_poses(angle_diff, 99999, g, symmetry_tfs)
  return kept.astype(np.float32)


def guess_translation(depth, mask, K):
  """reference: src/estimater.py:137-156"""
  vs, us = np.where(mask > 0)
  if len(us) == 0:
    return np.zeros((3))
  uc = (us.min() + us.max()) / 2.0
  vc = (vs.min() + vs.max()) / 2.0
  valid = mask.astype(bool) & (depth >= 0.001)
  if not valid.any():
    return np.zeros((3))
  zc = np.median(depth[valid])
  center = (np.linalg.inv(K) @ np.asarray([uc, vc, 1]).reshape(3, 1)) * zc
  return center.reshape(3)


def projection_matrix_from_intrinsics(K, height, width, znear, zfar):
  """reference: src/Utils.py:752-802, window_coords='y_down' branch (float64)."""
  depth = float(zfar - znear)
  q = -(zfar + znear) / depth
  qn = -2 * (zfar * znear) / depth
  w, h = width, height
  return np.array([[2 * K[0, 0] / w, -2 * K[0, 1] / w, (-2 * K[0, 2] + w) / w, 0],
                   [0, 2 * K[1, 1] / h, (2 * K[1, 2] - h) / h, 0],
                   [0, 0, q, qn],
                   [0, 0, -1, 0]])


def add_err(pred, gt, model_pts):
  """src/Utils.py:232-240."""
  pred, gt, pts = (np.asarray(a, dtype=np.float64) for a in (pred, gt, model_pts))
  a = pts @ pred[:3, :3].T + pred[:3, 3]
  b = pts @ gt[:3, :3].T + gt[:3, 3]
  return float(np.linalg.norm(a - b, axis=-1).mean())


def adds_err(pred, gt, model_pts):
  """src/Utils.py:242-253 (cKDTree nearest neighbour, k=1)."""
  from scipy.spatial import cKDTree
  pred, gt, pts = (np.asarray(a, dtype=np.float64) for a in (pred, gt, model_pts))
  a = pts @ pred[:3, :3].T + pred[:3, 3]
  b = pts @ gt[:3, :3].T + gt[:3, 3]
  d, _ = cKDTree(a).query(b, k=1)
  return float(d.mean())
