"""CPU ORACLE (test infrastructure, NOT product code) -- ctypes front-end of
oracle/fp_oracle.c plus the small torch/numpy pieces of the refine loop.
PARITY UNPINNED (see oracle/geometry.py header).  May be imported only by
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference.
"""
import ctypes

import numpy as np

from .geometry import lib, _p, _f32, projection_matrix_from_intrinsics

c_f = ctypes.c_float
c_i = ctypes.c_int


def erode_depth(depth, radius=2, depth_diff_thres=0.001, ratio_thres=0.8, zfar=100):
  """reference: src/Utils.py:359-395"""
  d = _f32(depth)
  out = np.empty_like(d)
  lib().orc_erode_depth(_p(d), _p(out), d.shape[0], d.shape[1], c_i(radius), c_f(depth_diff_thres),
                        c_f(ratio_thres), c_f(zfar))
  return out


def bilateral_filter_depth(depth, radius=2, zfar=100, sigmaD=2, sigmaR=100000):
  """reference: src/Utils.py:304-356"""
  d = _f32(depth)
  out = np.empty_like(d)
  lib().orc_bilateral_depth(_p(d), _p(out), d.shape[0], d.shape[1], c_i(radius), c_f(zfar), c_f(sigmaD),
                            c_f(sigmaR))
  return out


def depth2xyzmap(depth, K):
  """reference: src/Utils.py:399-417 (float64 math, float32 result)"""
  d = _f32(depth)
  K9 = np.ascontiguousarray(K, dtype=np.float64).reshape(9)
  out = np.empty(d.shape + (3,), dtype=np.float32)
  lib().orc_depth2xyz_f64(_p(d), _p(K9), _p(out), d.shape[0], d.shape[1])
  return out


def depth2xyzmap_batch(depths, Ks, zfar):
  """reference: src/Utils.py:420-438 (fp32 math)"""
  depths = _f32(depths)
  Ks = _f32(Ks).reshape(-1, 9)
  out = np.empty(depths.shape + (3,), dtype=np.float32)
  zf = np.float32(min(zfar, 3.0e38))
  for b in range(depths.shape[0]):
    lib().orc_depth2xyz_f32(_p(depths[b]), _p(Ks[b if len(Ks) > 1 else 0]), c_f(zf), _p(out[b]),
                            depths.shape[1], depths.shape[2])
  return out


def crop_params(poses, K, mesh_diameter, crop_ratio, S, H, W):
  """compute_crop_window_tf_batch (src/Utils.py:577-621) + bbox2d_ori + render matrices."""
  poses = _f32(poses).reshape(-1, 16)
  N = len(poses)
  K9 = _f32(np.asarray(K).reshape(9))
  radius = np.float32(mesh_diameter * crop_ratio / 2)
  P = _f32(projection_matrix_from_intrinsics(np.asarray(K, dtype=np.float64), H, W, 0.001, 100).reshape(16))
  out = dict(tf=np.empty((N, 9), np.float32), bbox=np.empty((N, 4), np.float32),
             mtx=np.empty((N, 16), np.float32), rtf=np.empty((N, 4), np.float32),
             inv=np.empty((N, 4), np.float32))
  lib().orc_crop_params(_p(poses), N, _p(K9), c_f(radius), S, H, W, _p(P), _p(out['tf']), _p(out['bbox']),
                        _p(out['mtx']), _p(out['rtf']), _p(out['inv']))
  return out


def render(poses, cp, mesh, S, use_light=True, w_ambient=0.8, w_diffuse=0.5, get_normal=False):
  """nvdiffrast_render restatement (src/Utils.py:133-219); mesh = dict of numpy arrays
  (pos, faces, vnormals, and vertex_color | tex+uv+uv_idx, like make_mesh_tensors).
  Returns (color, xyz, tri) and, with get_normal, a 4th array: the normal_map (:192-197)."""
  poses = _f32(poses).reshape(-1, 16)
  N = len(poses)
  pos = _f32(mesh['pos']); vn = _f32(mesh['vnormals'])
  faces = np.ascontiguousarray(mesh['faces'], dtype=np.int32)
  vc = uv = uvi = tex = None
  Ht = Wt = 0
  if 'tex' in mesh:
    tex = _f32(mesh['tex']).reshape(mesh['tex'].shape[-3], mesh['tex'].shape[-2], 3)
    Ht, Wt = tex.shape[:2]
    uv = _f32(mesh['uv']); uvi = np.ascontiguousarray(mesh['uv_idx'], dtype=np.int32)
  else:
    vc = _f32(mesh['vertex_color'])
  color = np.empty((N, S, S, 3), np.float32)
  xyz = np.empty((N, S, S, 3), np.float32)
  tri = np.empty((N, S, S), np.int32)
  normal = np.empty((N, S, S, 3), np.float32) if get_normal else None
  lib().orc_render(_p(poses), _p(cp['mtx']), _p(cp['rtf']), N, S, _p(pos), _p(vn), _p(vc), len(pos),
                   _p(faces), len(faces), _p(uv), _p(uvi), _p(tex), Ht, Wt, _p(color), _p(xyz), _p(tri),
                   1 if use_light else 0, c_f(w_ambient), c_f(w_diffuse), _p(normal))
  if get_normal:
    return color, xyz, tri, normal
  return color, xyz, tri


def make_A(color, xyz, poses, mesh_diameter, zthr):
  poses = _f32(poses).reshape(-1, 16)
  N, S = color.shape[0], color.shape[1]
  A = np.empty((N, S, S, 6), np.float32)
  lib().orc_make_A(_p(_f32(color)), _p(_f32(xyz)), _p(poses), N, S, c_f(mesh_diameter), c_f(zthr), _p(A))
  return A


def make_B(rgb, xyz_map, depth, poses, cp, K, S, mesh_diameter, mode):
  """mode 0 = refiner (nearest xyz_map), mode 1 = scorer (depth round trip)."""
  poses = _f32(poses).reshape(-1, 16)
  N = len(poses)
  rgb = _f32(rgb); H, W = rgb.shape[:2]
  xyz_map = _f32(xyz_map) if xyz_map is not None else None
  depth = _f32(depth) if depth is not None else None
  K9 = _f32(np.asarray(K).reshape(9))
  B = np.empty((N, S, S, 6), np.float32)
  lib().orc_make_B(_p(rgb), _p(xyz_map), _p(depth), H, W, _p(poses), _p(cp['tf']), _p(cp['inv']), _p(K9),
                   N, S, c_f(mesh_diameter), mode, _p(B))
  return B


def so3_exp_map(log_rot, eps=1e-4):
  """pytorch3d so3_exp_map (predict_pose_refine.py:222); fp32 numpy, [n,3] -> [n,3,3]."""
  r = _f32(log_rot)
  nrms = (r * r).sum(1)
  theta = np.sqrt(np.clip(nrms, eps, None)).astype(np.float32)
  fac1 = (np.sin(theta) / theta).astype(np.float32)
  fac2 = ((np.float32(1.0) - np.cos(theta)) / (theta * theta)).astype(np.float32)
  Km = np.zeros((len(r), 3, 3), np.float32)
  Km[:, 0, 1] = -r[:, 2]; Km[:, 0, 2] = r[:, 1]
  Km[:, 1, 0] = r[:, 2]; Km[:, 1, 2] = -r[:, 0]
  Km[:, 2, 0] = -r[:, 1]; Km[:, 2, 1] = r[:, 0]
  K2 = Km @ Km
  return (fac1[:, None, None] * Km + fac2[:, None, None] * K2 + np.eye(3, dtype=np.float32)[None]).astype(np.float32)


def rotation_6d_to_matrix(d6):
  """pytorch3d.transforms.rotation_6d_to_matrix (Zhou et al. 2019; third-party, source absent: restated)."""
  d6 = _f32(d6)
  a1, a2 = d6[:, :3], d6[:, 3:]
  b1 = a1 / np.maximum(np.linalg.norm(a1, axis=-1, keepdims=True), np.float32(1e-12))
  b2 = a2 - (b1 * a2).sum(-1, keepdims=True) * b1
  b2 = b2 / np.maximum(np.linalg.norm(b2, axis=-1, keepdims=True), np.float32(1e-12))
  b3 = np.cross(b1, b2)
  return np.stack([b1, b2, b3], axis=-2).astype(np.float32)


def pose_update(poses, trans, rot, mesh_diameter, rot_normalizer, trans_normalizer=None, normalize_xyz=True,
                trans_rep='tracknet', rot_rep='axis_angle', tf_to_crops=None, K=None, S=160):
  """predict_pose_refine.py:195-231 + egocentric_delta_pose_to_pose (src/Utils.py:848-855)."""
  poses = _f32(poses).reshape(-1, 4, 4)
  trans = _f32(trans); rot = _f32(rot)
  if trans_rep == 'tracknet':
    if not normalize_xyz:
      trans_delta = np.tanh(trans) * _f32(trans_normalizer).reshape(1, 3)
    else:
      trans_delta = trans.copy()
  elif trans_rep == 'deepim':                                   # :201-215
    K32 = _f32(K).reshape(3, 3)
    tf = _f32(tf_to_crops).reshape(-1, 3, 3)
    centers = poses[:, :3, 3]
    uvs = (K32[None] @ centers.reshape(-1, 3, 1)).reshape(-1, 3)
    uvs = uvs / uvs[:, 2:3]
    uvA_crop = (tf @ uvs.reshape(-1, 3, 1)).reshape(-1, 3)[:, :2]
    z_pred = trans[:, 2] * poses[:, 2, 3]
    uv_pred_crop = uvA_crop + trans[:, :2] * np.float32(S)
    tf_inv = np.linalg.inv(tf.astype(np.float64)).astype(np.float32)
    uv_pred = (tf_inv[:, :2, :2] @ uv_pred_crop[..., None] + tf_inv[:, :2, 2:])[..., 0]      # transform_pts
    center_pred = np.concatenate([uv_pred, np.ones((len(uv_pred), 1), np.float32)], -1)
    K_inv = np.linalg.inv(K32.astype(np.float64)).astype(np.float32)
    center_pred = (K_inv[None] @ center_pred.reshape(-1, 3, 1)).reshape(-1, 3) * z_pred.reshape(-1, 1)
    trans_delta = center_pred - centers
  else:
    trans_delta = trans.copy()
  if rot_rep == 'axis_angle':
    rot_delta = so3_exp_map(np.tanh(rot) * np.float32(rot_normalizer)).transpose(0, 2, 1)
  elif rot_rep == '6d':
    rot_delta = rotation_6d_to_matrix(rot).transpose(0, 2, 1)
  else:
    raise RuntimeError
  if normalize_xyz:
    trans_delta = trans_delta * np.float32(mesh_diameter / 2)
  out = np.tile(np.eye(4, dtype=np.float32)[None], (len(poses), 1, 1))
  out[:, :3, 3] = poses[:, :3, 3] + trans_delta
  out[:, :3, :3] = rot_delta @ poses[:, :3, :3]
  return out
