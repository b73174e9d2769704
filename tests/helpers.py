"""Shared helpers for the parity tests (oracle = checker only)."""
import numpy as np

# make_mesh_tensors divides by 255.0 on a CUDA tensor (src/Utils.py:112,121): ATen multiplies by the fp32 reciprocal
INV255 = np.float32(1.0) / np.float32(255.0)


def mesh_np(mesh):
  """SimpleMesh -> dict of numpy arrays in make_mesh_tensors layout (src/Utils.py:104-130)."""
  d = dict(pos=np.asarray(mesh.vertices, np.float32), faces=np.asarray(mesh.faces, np.int32),
           vnormals=np.asarray(mesh.vertex_normals, np.float32))
  vis = mesh.visual
  if vis.uv is not None and vis.material is not None:
    img = np.asarray(vis.material.image)[..., :3]
    d['tex'] = (img.astype(np.float32) * INV255)[None]
    uv = np.asarray(vis.uv, np.float32).copy()
    uv[:, 1] = 1 - uv[:, 1]
    d['uv'] = uv
    d['uv_idx'] = d['faces']
  else:
    d['vertex_color'] = np.asarray(vis.vertex_colors)[..., :3].astype(np.float32) * INV255
  return d


def centered(scene):
  """Centre the mesh like FoundationPose.reset_object (src/estimater.py:44-51)."""
  m = scene['mesh'].copy()
  c = (m.vertices.min(0) + m.vertices.max(0)) / 2
  m.vertices = m.vertices - c.reshape(1, 3)
  return m, c


def hypo_poses(scene, n=None, jitter_seed=None):
  """rot grid + guessed translation, as generate_random_pose_hypo (src/estimater.py:127-134)."""
  from oracle import geometry as og
  from oracle import ops as oo
  d = oo.bilateral_filter_depth(oo.erode_depth(scene['depth']))
  grid = og.make_rotation_grid()
  t = og.guess_translation(d, scene['mask'], scene['K'])
  poses = grid.copy()
  poses[:, :3, 3] = t.astype(np.float32)
  if n is not None:
    poses = poses[:: max(1, len(poses) // n)][:n].copy()
  if jitter_seed is not None:
    rng = np.random.default_rng(jitter_seed)
    poses[:, :3, 3] += rng.normal(0, 0.004, (len(poses), 3)).astype(np.float32)
  return poses, d
