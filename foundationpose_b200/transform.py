"""`PoseTransformer` of the reference (src/transform.py:4-67): 4x4 pose -> (x, y, z, roll, pitch, yaw), by default in
inches and degrees, with the gimbal-lock branch at |R[2,0]| > 0.9999.  Plus the headless part of main.py:35-79
(`run_foundation_pose`): register, write the pose file, return the pose of the oriented-bounding-box centre frame.
No window / cv2.imshow (graphics are unavailable on the box); the 3-D box overlay is a pure numpy projection."""
import datetime
import logging
import math
import os

import numpy as np

from .pipeline import pose_to_6d

METERS_TO_INCHES = 39.3701    # src/transform.py:43


class PoseTransformer:
  """Same constructor, methods and return order as the reference class."""

  def __init__(self, to_inches=True, to_degrees=True):
    self.to_inches = to_inches
    self.to_degrees = to_degrees

  def transform_pose(self, center_pose):
    pose_6d = self._convert_pose_matrix_to_6d(center_pose)
    x, y, z, roll, pitch, yaw = pose_6d
    logging.info('object position (%s): x=%.4f, y=%.4f, z=%.4f', 'inches' if self.to_inches else 'meters', x, y, z)
    logging.info('object rotation (%s): roll=%.4f, pitch=%.4f, yaw=%.4f', 'degrees' if self.to_degrees else 'radians',
                 roll, pitch, yaw)
    return pose_6d

  def _convert_pose_matrix_to_6d(self, pose_matrix):
    x, y, z, roll, pitch, yaw = pose_to_6d(pose_matrix)
    if self.to_degrees:
      roll, pitch, yaw = math.degrees(roll), math.degrees(pitch), math.degrees(yaw)
    if self.to_inches:
      x, y, z = x * METERS_TO_INCHES, y * METERS_TO_INCHES, z * METERS_TO_INCHES
    return x, y, z, roll, pitch, yaw

  def _rotation_matrix_to_euler_angles(self, R):
    P = np.eye(4)
    P[:3, :3] = np.asarray(R, dtype=np.float64)
    return pose_to_6d(P)[3:]


def oriented_bounds(vertices, n_dirs=None):
  """Stand-in for trimesh.bounds.oriented_bounds (trimesh 4.2.2, absent offline; main.py:39): minimum-volume box over
  the candidate frames given by the convex-hull face normals (each with the minimum-area rectangle of the projected
  hull, rotating calipers over the 2-D hull edges).  Returns (to_origin [4,4], extents [3]) with the same meaning:
  `to_origin` moves the mesh so that the box is centred at the origin and axis aligned.  The search space is the one
  trimesh uses; its tie-breaking and axis ordering are not reproduced (parity unpinned for this helper only -- it
  affects the frame the 6-D pose is reported in, never the registration itself)."""
  from scipy.spatial import ConvexHull
  V = np.asarray(vertices, dtype=np.float64).reshape(-1, 3)
  hull = ConvexHull(V)
  H = V[hull.vertices]
  normals = hull.equations[:, :3]
  # one normal per direction (hemisphere + rounding)
  sign = np.sign(normals[np.arange(len(normals)), np.argmax(np.abs(normals), axis=1)])
  normals = np.unique(np.round(normals * sign[:, None], 6), axis=0)
  best = (np.inf, None, None)
  for n in normals:
    n = n / np.linalg.norm(n)
    a = np.cross(n, [1.0, 0, 0] if abs(n[0]) < 0.9 else [0, 1.0, 0])
    a /= np.linalg.norm(a)
    b = np.cross(n, a)
    P2 = H @ np.stack([a, b], 1)                      # projected hull vertices
    h2 = P2[ConvexHull(P2).vertices]
    e = np.roll(h2, -1, 0) - h2
    ang = np.unique(np.round(np.mod(np.arctan2(e[:, 1], e[:, 0]), np.pi / 2), 9))
    height = np.ptp(H @ n)
    for t in ang:
      c, s = math.cos(t), math.sin(t)
      u = h2 @ np.array([c, s]); v = h2 @ np.array([-s, c])
      vol = np.ptp(u) * np.ptp(v) * height
      if vol < best[0] - 1e-15:
        best = (vol, (n, a, b), t)
  _, (n, a, b), t = best
  c, s = math.cos(t), math.sin(t)
  ax = np.stack([c * a + s * b, -s * a + c * b, n], 0)            # rows = box axes in the mesh frame
  ext = np.ptp(V @ ax.T, axis=0)
  order = np.argsort(ext)                                         # extents ascending, as trimesh reports them
  ax, ext = ax[order], ext[order]
  if np.linalg.det(ax) < 0:
    ax[0] = -ax[0]
  lo = (V @ ax.T).min(0)
  T = np.eye(4)
  T[:3, :3] = ax
  T[:3, 3] = -(lo + ext / 2)
  return T, ext


def project_box_edges(K, ob_in_cam, bbox):
  """Image-plane end points [12,2,2] of the edges of the posed 3-D box (what draw_posed_3d_box draws, main.py:65)."""
  lo, hi = np.asarray(bbox, dtype=np.float64).reshape(2, 3)
  c = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])])
  pc = c @ np.asarray(ob_in_cam)[:3, :3].T + np.asarray(ob_in_cam)[:3, 3]
  uv = pc @ np.asarray(K, dtype=np.float64).T
  uv = uv[:, :2] / uv[:, 2:3]
  edges = [(i, j) for i in range(8) for j in range(i + 1, 8) if bin(i ^ j).count('1') == 1]
  return np.stack([np.stack([uv[i], uv[j]]) for i, j in edges])


def run_foundation_pose(rgb_img, depth_img, mask_img, mesh, K, output_dir='output_dir', est=None, iteration=5, scorer=None,
                        refiner=None):
  """Headless main.py:35-79: returns `center_pose` (pose of the oriented-box centre frame) and writes
  `<output_dir>/<timestamp>.txt` with the raw 4x4 pose."""
  from .estimater import FoundationPose
  from .predict_pose_refine import PoseRefinePredictor
  from .predict_score import ScorePredictor
  to_origin, extents = oriented_bounds(mesh.vertices)
  if est is None:
    est = FoundationPose(model_pts=np.asarray(mesh.vertices), model_normals=np.asarray(mesh.vertex_normals), mesh=mesh,
                         scorer=scorer or ScorePredictor(), refiner=refiner or PoseRefinePredictor(), glctx=None, debug=0)
  pose = np.asarray(est.register(K=np.asarray(K, dtype=np.float64), rgb=rgb_img, depth=depth_img,
                                 ob_mask=np.asarray(mask_img).astype(bool), iteration=iteration)).reshape(4, 4)
  os.makedirs(output_dir, exist_ok=True)
  stamp = datetime.datetime.now().strftime('%Y%m%d_%H%M%S')
  np.savetxt(os.path.join(output_dir, f'{stamp}.txt'), pose)
  return pose @ np.linalg.inv(to_origin)
