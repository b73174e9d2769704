"""Deterministic synthetic meshes, RGB-D frames and network weights for the BASELINE.json configs
(SURVEY.md 8(d)).  No dataset or checkpoint is reachable offline, so tests and bench.py use these.

Everything here is seeded numpy / torch-CPU and produces identical arrays in the build container and on the
GPU box.  The frame generator is a dense point-splat z-buffer (input data only -- it is not the renderer
under test).
"""
import math
from types import SimpleNamespace

import numpy as np
import torch

from .utils import icosphere


class SimpleMesh:
  """Minimal trimesh-like container (trimesh is absent offline)."""

  def __init__(self, vertices, faces, vertex_normals=None, vertex_colors=None, uv=None, image=None):
    self.vertices = np.asarray(vertices, dtype=np.float64)
    self.faces = np.asarray(faces, dtype=np.int64)
    self.vertex_normals = (np.asarray(vertex_normals, dtype=np.float64) if vertex_normals is not None
                           else compute_vertex_normals(self.vertices, self.faces))
    material = SimpleNamespace(image=image) if image is not None else None
    self.visual = SimpleNamespace(vertex_colors=vertex_colors, uv=uv, material=material)

  def copy(self):
    v = self.visual
    return SimpleMesh(self.vertices.copy(), self.faces.copy(), self.vertex_normals.copy(),
                      None if v.vertex_colors is None else np.array(v.vertex_colors),
                      None if v.uv is None else np.array(v.uv),
                      None if v.material is None else np.array(v.material.image))


def compute_vertex_normals(vertices, faces):
  fn = np.cross(vertices[faces[:, 1]] - vertices[faces[:, 0]], vertices[faces[:, 2]] - vertices[faces[:, 0]])
  vn = np.zeros_like(vertices)
  for k in range(3):
    np.add.at(vn, faces[:, k], fn)
  n = np.linalg.norm(vn, axis=1, keepdims=True)
  return vn / np.maximum(n, 1e-12)


def cube_mesh(side=0.10, seed=10, tex_size=256):
  """cfg 1: textured cube, 24 vertices (4 per face), 12 faces, per-face normals, low-passed noise texture."""
  h = side / 2
  normals = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
  V, Fc, Nn, UV = [], [], [], []
  for fi, n in enumerate(normals):
    n = np.array(n, dtype=np.float64)
    a = np.array([0, 0, 1.0]) if abs(n[2]) < 0.5 else np.array([1.0, 0, 0])
    u = np.cross(a, n); u /= np.linalg.norm(u)
    v = np.cross(n, u)
    base = len(V)
    for (su, sv) in [(-1, -1), (1, -1), (1, 1), (-1, 1)]:
      V.append(h * (n + su * u + sv * v))
      Nn.append(n)
      cu, cv = fi % 3, fi // 3   # 3x2 atlas
      UV.append(((cu + 0.5 + 0.45 * su) / 3.0, (cv + 0.5 + 0.45 * sv) / 2.0))
    Fc += [(base, base + 1, base + 2), (base, base + 2, base + 3)]
  rng = np.random.default_rng(seed)
  tex = rng.integers(0, 256, size=(tex_size // 4, tex_size // 4, 3)).astype(np.float64)
  tex = np.kron(tex, np.ones((4, 4, 1)))
  k = np.ones(5) / 5.0
  for ax in (0, 1):
    tex = np.apply_along_axis(lambda m: np.convolve(np.pad(m, 2, mode='wrap'), k, mode='valid'), ax, tex)
  tex = np.clip(tex, 0, 255).astype(np.uint8)
  return SimpleMesh(np.array(V), np.array(Fc), np.array(Nn), uv=np.array(UV), image=tex)


def blob_mesh(subdivisions=5, radius=0.1, seed=20, coef=(0.15, 3.0, 2.0, 0.1)):
  """cfg 2: displaced icosphere (subdiv 5 -> 10,242 v / 20,480 f), smooth vertex colours + noise."""
  v, f = icosphere(subdivisions, 1.0)
  a, fx, fy, b = coef
  disp = 1 + a * np.sin(fx * v[:, 0]) * np.cos(fy * v[:, 1]) + b * v[:, 2]
  verts = v * disp[:, None] * radius
  rng = np.random.default_rng(seed)
  col = np.stack([0.5 + 0.5 * np.sin(4 * v[:, 0] + 1), 0.5 + 0.5 * np.cos(3 * v[:, 1]), 0.5 + 0.5 * np.sin(5 * v[:, 2] + 2)], 1)
  col = np.clip(col * 255 + rng.normal(0, 12, col.shape), 0, 255).astype(np.uint8)
  return SimpleMesh(verts, f, vertex_colors=col)


def sheet_mesh(n=501, size=(0.30, 0.20), seed=40):
  """cfg 4: bent sheet, n x n grid (501 -> 251,001 v / 500,000 f), folded 90 deg along a rounded line, grey."""
  xs = np.linspace(-size[0] / 2, size[0] / 2, n)
  ys = np.linspace(-size[1] / 2, size[1] / 2, n)
  X, Y = np.meshgrid(xs, ys, indexing='xy')
  r = 0.02
  s = X.copy()
  # arc-length parameterisation: flat for s<0, quarter circle of radius r, then vertical
  arc = np.clip(s, 0, r * math.pi / 2)
  ang = arc / r
  Xn = np.where(s < 0, s, r * np.sin(ang) + np.where(s > r * math.pi / 2, 0.0, 0.0))
  Zn = np.where(s < 0, 0.0, r * (1 - np.cos(ang)) + np.clip(s - r * math.pi / 2, 0, None))
  verts = np.stack([Xn, Y, Zn], -1).reshape(-1, 3)
  idx = np.arange(n * n).reshape(n, n)
  a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, 1:].ravel(), idx[1:, :-1].ravel()
  faces = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], 0)
  col = np.full((len(verts), 3), 128, dtype=np.uint8)
  return SimpleMesh(verts, faces, vertex_colors=col)


def random_rotation(seed):
  rng = np.random.default_rng(seed)
  q = rng.normal(size=4); q /= np.linalg.norm(q)
  w, x, y, z = q
  return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                   [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                   [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def _mesh_colors_at(mesh, fidx, bary):
  vis = mesh.visual
  if vis.uv is not None and vis.material is not None:
    uv = (np.asarray(vis.uv)[mesh.faces[fidx]] * bary[..., None]).sum(1)
    img = np.asarray(vis.material.image)
    Ht, Wt = img.shape[:2]
    px = np.clip((uv[:, 0] * Wt).astype(int), 0, Wt - 1)
    py = np.clip(((1 - uv[:, 1]) * Ht).astype(int), 0, Ht - 1)
    return img[py, px, :3].astype(np.float64)
  col = np.asarray(vis.vertex_colors)[..., :3].astype(np.float64)
  return (col[mesh.faces[fidx]] * bary[..., None]).sum(1)


def render_frame(mesh, pose, K, H, W, n_points=1500000, seed=0, bg_depth=1.2, noise_sigma=0.001, dropout=0.02):
  """Synthetic RGB-D observation of `mesh` at `pose` (object-in-camera 4x4): dense surface samples are
  projected and z-buffered.  Returns rgb uint8 [H,W,3], depth float32 [H,W] (metres), mask bool [H,W]."""
  rng = np.random.default_rng(seed)
  v, f = mesh.vertices, mesh.faces
  tri = v[f]
  area = 0.5 * np.linalg.norm(np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0]), axis=1)
  fidx = rng.choice(len(f), size=n_points, p=area / area.sum())
  r1, r2 = np.sqrt(rng.random(n_points)), rng.random(n_points)
  bary = np.stack([1 - r1, r1 * (1 - r2), r1 * r2], 1)
  pts = (tri[fidx] * bary[..., None]).sum(1)
  col = _mesh_colors_at(mesh, fidx, bary)
  nrm = (mesh.vertex_normals[f[fidx]] * bary[..., None]).sum(1)
  pc = pts @ pose[:3, :3].T + pose[:3, 3]
  nc = nrm @ pose[:3, :3].T
  nc /= np.maximum(np.linalg.norm(nc, axis=1, keepdims=True), 1e-12)
  shade = 0.8 + 0.5 * np.clip(-nc[:, 2], 0, 1)
  col = np.clip(col * shade[:, None], 0, 255)
  u = np.round(K[0, 0] * pc[:, 0] / pc[:, 2] + K[0, 2]).astype(int)
  w_ = np.round(K[1, 1] * pc[:, 1] / pc[:, 2] + K[1, 2]).astype(int)
  ok = (u >= 0) & (u < W) & (w_ >= 0) & (w_ < H) & (pc[:, 2] > 0.01)
  u, w_, z, col = u[ok], w_[ok], pc[ok, 2], col[ok]
  lin = w_ * W + u
  order = np.lexsort((z, lin))
  lin_s = lin[order]
  first = np.ones(len(lin_s), dtype=bool); first[1:] = lin_s[1:] != lin_s[:-1]
  sel = order[first]
  depth = np.full(H * W, bg_depth, dtype=np.float64)
  rgb = np.zeros((H * W, 3), dtype=np.float64)
  yy, xx = np.mgrid[0:H, 0:W]
  bg = np.stack([90 + 40 * np.sin(xx / 37.0), 100 + 30 * np.cos(yy / 29.0), 110 + 35 * np.sin((xx + yy) / 53.0)], -1)
  rgb[:] = bg.reshape(-1, 3)
  mask = np.zeros(H * W, dtype=bool)
  depth[lin[sel]] = z[sel]; rgb[lin[sel]] = col[sel]; mask[lin[sel]] = True
  depth = depth + rng.normal(0, noise_sigma, depth.shape)
  depth[rng.random(depth.shape) < dropout] = 0.0
  return (np.clip(rgb, 0, 255).astype(np.uint8).reshape(H, W, 3), depth.astype(np.float32).reshape(H, W),
          mask.reshape(H, W))


def default_K(W=640, H=480):
  if (W, H) == (640, 480):
    return np.array([[600.0, 0, 320.0], [0, 600.0, 240.0], [0, 0, 1.0]])
  return np.array([[1400.0, 0, W / 2.0], [0, 1400.0, H / 2.0], [0, 0, 1.0]])


def make_scene(cfg=2, n_points=1500000):
  """Returns dict(mesh, K, rgb, depth, mask, gt_pose, H, W, diameter) for BASELINE config 1, 2 or 4."""
  if cfg == 1:
    mesh, seeds, (W, H) = cube_mesh(), (11, 12), (640, 480)
  elif cfg == 2:
    mesh, seeds, (W, H) = blob_mesh(), (21, 22), (640, 480)
  elif cfg == 4:
    mesh, seeds, (W, H) = sheet_mesh(), (41, 42), (1920, 1200)
  else:
    raise ValueError(cfg)
  K = default_K(W, H)
  pose = np.eye(4)
  pose[:3, :3] = random_rotation(seeds[0])
  center = (mesh.vertices.min(0) + mesh.vertices.max(0)) / 2
  pose[:3, 3] = np.array([0.03, -0.02, 0.60]) - pose[:3, :3] @ center
  rgb, depth, mask = render_frame(mesh, pose, K, H, W, n_points=n_points, seed=seeds[1])
  ext = mesh.vertices - center
  diameter = float(2 * np.linalg.norm(ext, axis=1).max())
  return dict(mesh=mesh, K=K, rgb=rgb, depth=depth, mask=mask, gt_pose=pose, H=H, W=W, diameter=diameter)
