"""Host-side mirror of the hot-path functions of the reference's `src/Utils.py`, backed by libfpb200.so.

Same names, argument meaning and numpy-in => numpy-out / torch-in => torch-out behaviour as the reference
(file:line cited per function).  There is no CPU fallback: every compute function needs a CUDA device.
"""
import logging
import math
import random

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, ptr, hptr, stream, f32_host

glcam_in_cvcam = np.array([[1, 0, 0, 0], [0, -1, 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]]).astype(float)
CROP_STRIDE = 40


def set_seed(random_seed):
  """reference: src/Utils.py:222-229"""
  np.random.seed(random_seed)
  random.seed(random_seed)
  torch.manual_seed(random_seed)
  if torch.cuda.is_available():
    torch.cuda.manual_seed_all(random_seed)
  torch.backends.cudnn.deterministic = True
  torch.backends.cudnn.benchmark = False


def _as_cuda_f32(x, device='cuda'):
  return torch.as_tensor(x, dtype=torch.float, device=device).contiguous()


# ------------------------------------------------------------------------------------------------
# depth pre-processing
# ------------------------------------------------------------------------------------------------
def erode_depth(depth, radius=2, depth_diff_thres=0.001, ratio_thres=0.8, zfar=100, device='cuda'):
  """reference: src/Utils.py:387-395 (kernel :359-384)"""
  _lib.require_cuda()
  d = _as_cuda_f32(depth, device)
  out = torch.empty_like(d)
  check(lib().fpb_erode_depth_f32(ptr(d), ptr(out), d.shape[0], d.shape[1], radius, depth_diff_thres, ratio_thres,
                                  zfar, stream()))
  if isinstance(depth, np.ndarray):
    return out.data.cpu().numpy()
  return out


def bilateral_filter_depth(depth, radius=2, zfar=100, sigmaD=2, sigmaR=100000, device='cuda'):
  """reference: src/Utils.py:345-356 (kernel :304-343)"""
  _lib.require_cuda()
  d = _as_cuda_f32(depth, device)
  out = torch.empty_like(d)
  check(lib().fpb_bilateral_depth_f32(ptr(d), ptr(out), d.shape[0], d.shape[1], radius, zfar, sigmaD, sigmaR,
                                      stream()))
  if isinstance(depth, np.ndarray):
    return out.data.cpu().numpy()
  return out


def depth2xyzmap_cuda(depth_t, K, math_f64=True, zfar=np.inf):
  """Device-resident depth -> xyz map ([H,W] fp32 -> [H,W,3] fp32)."""
  K9 = np.ascontiguousarray(np.asarray(K, dtype=np.float64).reshape(9))
  out = torch.empty(depth_t.shape + (3,), dtype=torch.float, device=depth_t.device)
  zf = float(min(zfar, 3.0e38))
  check(lib().fpb_depth_to_xyz(ptr(depth_t), hptr(K9), zf, ptr(out), depth_t.shape[0], depth_t.shape[1],
                               1 if math_f64 else 0, stream()))
  return out


def depth2xyzmap(depth, K, uvs=None):
  """reference: src/Utils.py:399-417 -- float64 math evaluated on the GPU, float32 numpy result."""
  _lib.require_cuda()
  d = _as_cuda_f32(depth)
  xyz = depth2xyzmap_cuda(d, K, math_f64=True).data.cpu().numpy()
  if uvs is not None:
    uvs = np.asarray(uvs).round().astype(int)
    keep = np.zeros(xyz.shape[:2], dtype=bool)
    keep[uvs[:, 1], uvs[:, 0]] = True
    xyz[~keep] = 0
  return xyz


def depth2xyzmap_batch(depths, Ks, zfar):
  """reference: src/Utils.py:420-438 -- fp32 math; depths (B,H,W), Ks (B,3,3) torch tensors."""
  _lib.require_cuda()
  depths = _as_cuda_f32(depths)
  Ks_h = torch.as_tensor(Ks).detach().cpu().double().numpy().reshape(-1, 3, 3)
  outs = []
  for b in range(depths.shape[0]):
    Kb = Ks_h[b if len(Ks_h) > 1 else 0].astype(np.float32).astype(np.float64)
    outs.append(depth2xyzmap_cuda(depths[b].contiguous(), Kb, math_f64=False, zfar=zfar))
  return torch.stack(outs, dim=0)


def mask_depth_stats(depth_t, mask_t):
  """Device statistics behind guess_translation (src/estimater.py:137-156).
  Returns (umin, umax, vmin, vmax, n_mask, n_valid, mid_lo, mid_hi)."""
  oi = torch.empty(6, dtype=torch.int32, device=depth_t.device)
  of = torch.empty(2, dtype=torch.float, device=depth_t.device)
  check(lib().fpb_mask_depth_stats(ptr(depth_t), ptr(mask_t), depth_t.shape[0], depth_t.shape[1], ptr(oi), ptr(of),
                                   stream()))
  oi = oi.cpu().numpy(); of = of.cpu().numpy()
  return int(oi[0]), int(oi[1]), int(oi[2]), int(oi[3]), int(oi[4]), int(oi[5]), of[0], of[1]


def register_prologue(depth_t, mask_t, K, rot_grid, min_valid=4):
  """FoundationPose.register() up to the first refiner call (src/estimater.py:159-205) in one C-ABI call and without a
  host synchronisation: erode_depth -> bilateral_filter_depth -> guess_translation -> hypotheses -> depth2xyzmap.
  Returns (depth [H,W], xyz_map [H,W,3], poses [N,4,4], stats_i32 [6], stats_f32 [2], center [3] float64) -- all
  device tensors; the statistics are those of mask_depth_stats."""
  H, W = depth_t.shape
  K = np.asarray(K, dtype=np.float64)
  K9 = np.ascontiguousarray(K.reshape(9))
  Ki9 = np.ascontiguousarray(np.linalg.inv(K).reshape(9))
  dev = depth_t.device
  tmp = torch.empty_like(depth_t)
  out = torch.empty_like(depth_t)
  xyz = torch.empty((H, W, 3), dtype=torch.float, device=dev)
  N = rot_grid.shape[0]
  poses = torch.empty((N, 4, 4), dtype=torch.float, device=dev)
  si = torch.empty(6, dtype=torch.int32, device=dev)
  sf = torch.empty(2, dtype=torch.float, device=dev)
  center = torch.empty(3, dtype=torch.double, device=dev)
  check(lib().fpb_register_prologue(ptr(depth_t), ptr(mask_t), H, W, hptr(K9), hptr(Ki9), ptr(tmp), ptr(out), ptr(xyz),
                                    ptr(rot_grid), N, int(min_valid), ptr(poses), ptr(si), ptr(sf), ptr(center), stream()))
  return out, xyz, poses, si, sf, center


# ------------------------------------------------------------------------------------------------
# hypothesis generation helpers (host, float64 -- init-time only in the reference too)
# ------------------------------------------------------------------------------------------------
_ICO_T = (1.0 + math.sqrt(5.0)) / 2.0
_ICO_V = [(-1, _ICO_T, 0), (1, _ICO_T, 0), (-1, -_ICO_T, 0), (1, -_ICO_T, 0), (0, -1, _ICO_T), (0, 1, _ICO_T),
          (0, -1, -_ICO_T), (0, 1, -_ICO_T), (_ICO_T, 0, -1), (_ICO_T, 0, 1), (-_ICO_T, 0, -1), (-_ICO_T, 0, 1)]
_ICO_F = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
          (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
          (8, 6, 7), (9, 8, 1)]


def icosphere(subdivisions=1, radius=1.0):
  """trimesh 4.2.2 `creation.icosphere` vertex/face order (requirements.txt:43; used at src/Utils.py:485-489):
  every level appends the unique edge mid-points, ordered by the (lo, hi) edge key with hi most significant,
  then re-projects all vertices onto the sphere."""
  verts = np.asarray(_ICO_V, dtype=np.float64) / np.sqrt(2.0 + _ICO_T)
  faces = np.asarray(_ICO_F, dtype=np.int64)
  for _ in range(int(subdivisions)):
    nv = len(verts)
    e = np.stack([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]], axis=1).reshape(-1, 2)
    e.sort(axis=1)
    code = e[:, 0] ^ (e[:, 1] << 32)
    _, first, inv = np.unique(code, return_index=True, return_inverse=True)
    mids = verts[e[first]].mean(axis=1)
    m = inv.reshape(-1, 3) + nv
    a, b, c = faces[:, 0], faces[:, 1], faces[:, 2]
    faces = np.stack([a, m[:, 0], m[:, 2], m[:, 0], b, m[:, 1], m[:, 2], m[:, 1], c, m[:, 0], m[:, 1], m[:, 2]],
                     axis=1).reshape(-1, 3)
    verts = np.vstack((verts, mids))
    norm = np.sqrt(np.dot(verts ** 2, [1, 1, 1]))
    verts = verts + (verts / norm.reshape(-1, 1)) * (radius - norm).reshape(-1, 1)
  return verts, faces


def sample_views_icosphere(n_views, subdivisions=None, radius=1):
  """reference: src/Utils.py:483-507 -> cam_in_obs float64 [n,4,4]"""
  if subdivisions is not None:
    verts, _ = icosphere(subdivisions, radius)
  else:
    level = 1
    while True:
      verts, _ = icosphere(level, radius)
      if verts.shape[0] >= n_views:
        break
      level += 1
  n = len(verts)
  cam_in_obs = np.tile(np.eye(4)[None], (n, 1, 1))
  cam_in_obs[:, :3, 3] = verts
  up = np.array([0, 0, 1])
  z_axis = -cam_in_obs[:, :3, 3]
  z_axis /= np.linalg.norm(z_axis, axis=-1).reshape(-1, 1)
  x_axis = np.cross(up.reshape(1, 3), z_axis)
  degenerate = (x_axis == 0).all(axis=-1)
  x_axis[degenerate] = [1, 0, 0]
  x_axis /= np.linalg.norm(x_axis, axis=-1).reshape(-1, 1)
  y_axis = np.cross(z_axis, x_axis)
  y_axis /= np.linalg.norm(y_axis, axis=-1).reshape(-1, 1)
  cam_in_obs[:, :3, 0] = x_axis
  cam_in_obs[:, :3, 1] = y_axis
  cam_in_obs[:, :3, 2] = z_axis
  return cam_in_obs


def euler_matrix(ai, aj, ak):
  """`transformations.euler_matrix` (2024.6.1), static 'sxyz' axes only (src/estimater.py:113)."""
  si, sj, sk = math.sin(ai), math.sin(aj), math.sin(ak)
  ci, cj, ck = math.cos(ai), math.cos(aj), math.cos(ak)
  cc, cs = ci * ck, ci * sk
  sc, ss = si * ck, si * sk
  M = np.identity(4)
  M[0, 0] = cj * ck; M[0, 1] = sj * sc - cs; M[0, 2] = sj * cc + ss
  M[1, 0] = cj * sk; M[1, 1] = sj * ss + cc; M[1, 2] = sj * cs - sc
  M[2, 0] = -sj; M[2, 1] = cj * si; M[2, 2] = cj * ci
  return M


def projection_matrix_from_intrinsics(K, height, width, znear, zfar, window_coords='y_down'):
  """reference: src/Utils.py:752-802 (float64)"""
  w, h = width, height
  depth = float(zfar - znear)
  q = -(zfar + znear) / depth
  qn = -2 * (zfar * znear) / depth
  if window_coords == 'y_up':
    row1 = [0, -2 * K[1, 1] / h, (-2 * K[1, 2] + h) / h, 0]
  elif window_coords == 'y_down':
    row1 = [0, 2 * K[1, 1] / h, (2 * K[1, 2] - h) / h, 0]
  else:
    raise NotImplementedError
  return np.array([[2 * K[0, 0] / w, -2 * K[0, 1] / w, (-2 * K[0, 2] + w) / w, 0], row1, [0, 0, q, qn], [0, 0, -1, 0]])


def compute_mesh_diameter(model_pts=None, mesh=None, n_sample=1000):
  """reference: src/Utils.py:559-574 (host numpy; the mesh= branch of the reference is broken and unused)."""
  if mesh is not None:
    model_pts = np.asarray(mesh.vertices)
  model_pts = np.asarray(model_pts)
  if n_sample is None:
    pts = model_pts
  else:
    ids = np.random.choice(len(model_pts), size=min(n_sample, len(model_pts)), replace=False)
    pts = model_pts[ids]
  best = 0.0
  for b in range(0, len(pts), 512):   # blocked pairwise distances (bounded memory)
    d = np.linalg.norm(pts[None, b:b + 512] - pts[:, None], axis=-1)
    best = max(best, float(d.max()))
  return best


# ------------------------------------------------------------------------------------------------
# crop windows
# ------------------------------------------------------------------------------------------------
def crop_params(poses, K, mesh_diameter, crop_ratio, S, H, W):
  """All per-hypothesis window quantities in one launch -> cuda float tensor [N, 40]
  (layout: include/fpb200.h FPB_CROP_STRIDE)."""
  _lib.require_cuda()
  poses = _as_cuda_f32(poses).reshape(-1, 4, 4)
  N = len(poses)
  K64 = np.asarray(torch.as_tensor(K).detach().cpu().numpy() if torch.is_tensor(K) else K, dtype=np.float64)
  K9 = f32_host(K64, 9)
  P16 = f32_host(projection_matrix_from_intrinsics(K64, height=H, width=W, znear=0.001, zfar=100), 16)
  radius = float(np.float32(mesh_diameter * crop_ratio / 2))
  params = torch.empty((N, CROP_STRIDE), dtype=torch.float, device=poses.device)
  check(lib().fpb_crop_params(ptr(poses), N, hptr(K9), radius, S, H, W, hptr(P16), ptr(params), stream()))
  return params


def compute_crop_window_tf_batch(pts=None, H=None, W=None, poses=None, K=None, crop_ratio=1.2, out_size=None, rgb=None,
                                 uvs=None, method='min_box', mesh_diameter=None):
  """reference: src/Utils.py:577-626 -> tf_to_crops (B,3,3) cuda float.  Only method='box_3d' exists upstream."""
  if method != 'box_3d':
    raise RuntimeError
  assert out_size[0] == out_size[1], 'square crops only'
  params = crop_params(poses, K, mesh_diameter, crop_ratio, int(out_size[0]), H, W)
  return params[:, :9].reshape(-1, 3, 3).contiguous()


# ------------------------------------------------------------------------------------------------
# mesh
# ------------------------------------------------------------------------------------------------
class MeshHandle:
  """Owns an fpb_mesh (device copies of the mesh arrays)."""

  def __init__(self, mesh_tensors):
    import ctypes
    _lib.require_cuda()
    mt = mesh_tensors
    pos = _as_cuda_f32(mt['pos']); vn = _as_cuda_f32(mt['vnormals'])
    faces = torch.as_tensor(mt['faces'], dtype=torch.int32, device='cuda').contiguous()
    vc = uv = uvi = tex = None
    Vt = Ht = Wt = 0
    if 'tex' in mt:
      tex = _as_cuda_f32(mt['tex']).reshape(mt['tex'].shape[-3], mt['tex'].shape[-2], 3).contiguous()
      Ht, Wt = tex.shape[:2]
      uv = _as_cuda_f32(mt['uv']); Vt = len(uv)
      uvi = torch.as_tensor(mt['uv_idx'], dtype=torch.int32, device='cuda').contiguous()
    else:
      vc = _as_cuda_f32(mt['vertex_color'])
    self.V, self.F = len(pos), len(faces)
    h = ctypes.c_void_p()
    check(lib().fpb_mesh_create(ctypes.byref(h), ptr(pos), ptr(vn), self.V, ptr(faces), self.F, ptr(vc), ptr(uv), Vt,
                                ptr(uvi), ptr(tex), Ht, Wt, stream()))
    self.handle = h

  def __del__(self):
    try:
      if getattr(self, 'handle', None):
        lib().fpb_mesh_destroy(self.handle)
        self.handle = None
    except Exception:
      pass


def make_mesh_tensors(mesh, device='cuda', max_tex_size=None):
  """reference: src/Utils.py:104-130.  `mesh` is trimesh-like (duck-typed: .vertices, .faces, .vertex_normals and
  either .visual.uv + .visual.material.image, or .visual.vertex_colors)."""
  mesh_tensors = {}
  visual = getattr(mesh, 'visual', None)
  uv = getattr(visual, 'uv', None) if visual is not None else None
  if uv is not None and getattr(visual, 'material', None) is not None:
    img = visual.material.image
    img = np.array(img.convert('RGB')) if hasattr(img, 'convert') else np.asarray(img)
    img = img[..., :3]
    if max_tex_size is not None and max(img.shape[0], img.shape[1]) > max_tex_size:
      import cv2
      scale = 1 / max(img.shape[0], img.shape[1]) * max_tex_size
      img = cv2.resize(img, fx=scale, fy=scale, dsize=None)
    mesh_tensors['tex'] = torch.as_tensor(np.ascontiguousarray(img), device=device, dtype=torch.float)[None] / 255.0
    mesh_tensors['uv_idx'] = torch.as_tensor(np.asarray(mesh.faces), device=device, dtype=torch.int)
    uv_t = torch.as_tensor(np.asarray(uv), device=device, dtype=torch.float).clone()
    uv_t[:, 1] = 1 - uv_t[:, 1]
    mesh_tensors['uv'] = uv_t
  else:
    vcol = getattr(visual, 'vertex_colors', None) if visual is not None else None
    if vcol is None:
      logging.info("WARN: mesh doesn't have vertex_colors, assigning a pure color")
      vcol = np.tile(np.array([128, 128, 128]).reshape(1, 3), (len(mesh.vertices), 1))
    mesh_tensors['vertex_color'] = torch.as_tensor(np.asarray(vcol)[..., :3], device=device, dtype=torch.float) / 255.0
  mesh_tensors.update({
    'pos': torch.tensor(np.asarray(mesh.vertices), device=device, dtype=torch.float),
    'faces': torch.tensor(np.asarray(mesh.faces), device=device, dtype=torch.int),
    'vnormals': torch.tensor(np.asarray(mesh.vertex_normals), device=device, dtype=torch.float),
  })
  return mesh_tensors


class MeshLike:
  """Minimal duck-typed stand-in for trimesh.Trimesh (trimesh is absent offline): what reset_object /
  make_mesh_tensors touch -- .vertices, .faces, .vertex_normals, .visual and .copy()."""

  def __init__(self, vertices, faces, vertex_normals, visual):
    self.vertices, self.faces, self.vertex_normals, self.visual = vertices, faces, vertex_normals, visual

  def copy(self):
    import copy
    return MeshLike(np.array(self.vertices), np.array(self.faces), np.array(self.vertex_normals), copy.deepcopy(self.visual))


def mesh_from_tensors(mesh_tensors):
  """Inverse of make_mesh_tensors: a ready mesh_tensors dict (pos, faces, vnormals, vertex_color | tex + uv [+ uv_idx],
  torch or numpy) -> MeshLike, so that callers holding only the dict can build a FoundationPose."""
  from types import SimpleNamespace

  def host(x):
    return x.detach().cpu().numpy() if torch.is_tensor(x) else np.asarray(x)
  mt = mesh_tensors
  if 'tex' in mt:
    tex = host(mt['tex']).reshape(mt['tex'].shape[-3], mt['tex'].shape[-2], 3)
    uv = host(mt['uv']).astype(np.float64).copy()
    uv[:, 1] = 1 - uv[:, 1]                                   # make_mesh_tensors flips v
    img = np.clip(np.rint(tex * 255.0), 0, 255).astype(np.uint8)
    visual = SimpleNamespace(uv=uv, material=SimpleNamespace(image=img), vertex_colors=None)
  else:
    col = np.clip(np.rint(host(mt['vertex_color']) * 255.0), 0, 255).astype(np.uint8)
    visual = SimpleNamespace(uv=None, material=None, vertex_colors=col)
  return MeshLike(host(mt['pos']).astype(np.float64), host(mt['faces']).astype(np.int64),
                  host(mt['vnormals']).astype(np.float64), visual)


def mesh_handle_of(mesh_tensors):
  """Cache the device mesh handle inside the mesh_tensors dict (the dict is the drop-in's mesh interface)."""
  h = mesh_tensors.get('_fpb_handle')
  if h is None:
    h = MeshHandle(mesh_tensors)
    mesh_tensors['_fpb_handle'] = h
  return h


# ------------------------------------------------------------------------------------------------
# render / warp / pose update
# ------------------------------------------------------------------------------------------------
IN_PITCH, IN_XOFF = 168, 4      # FPB_IN_PITCH / FPB_IN_XOFF of include/fpb200.h (S = 160)
_DT_CODE = {torch.float: 0, torch.bfloat16: 1, torch.float16: 2}     # a_is_bf16 / b_is_bf16 of fpb_render_crops / fpb_warp_crops


def net_input_buffer(N, S, dtype):
  """Zeroed [2N,S,S+8,8] network input: rendered crops first, observed crops second; rows padded so the 7x7 stem can
  read 8 taps x 8 channels as one 128-byte vector (csrc/net.cu).  Pad pixels stay zero (kernels never write them)."""
  assert S + 2 * IN_XOFF == IN_PITCH
  return torch.zeros((2 * N, S, IN_PITCH, 8), dtype=dtype, device='cuda')


def render_crops(mesh_tensors, poses, params, S, mesh_diameter, zthr, out_A=None, bf16=True, debug=False, use_light=True,
                 w_ambient=0.8, w_diffuse=0.5, get_normal=False):
  """Render all N hypotheses at crop resolution and write network input A ([N,S,P,8]; P = S for a dense buffer,
  P = IN_PITCH for the row-padded layout the networks read, see net_input_buffer).
  Returns (A, extras) where extras holds fp32 color/xyz/tri when debug=True and `normal` when get_normal=True."""
  mh = mesh_handle_of(mesh_tensors)
  poses = _as_cuda_f32(poses).reshape(-1, 4, 4)
  N = len(poses)
  if out_A is None:
    out_A = torch.empty((N, S, S, 8), dtype=torch.bfloat16 if bf16 else torch.float, device='cuda')
  ws = torch.empty(lib().fpb_render_ws_bytes(N, mh.V, S), dtype=torch.uint8, device='cuda')
  extras = {}
  c = x = t = None
  if debug:
    c = torch.empty((N, S, S, 3), dtype=torch.float, device='cuda')
    x = torch.empty((N, S, S, 3), dtype=torch.float, device='cuda')
    t = torch.empty((N, S, S), dtype=torch.int32, device='cuda')
    extras = dict(color=c, xyz=x, tri=t)
  nm = None
  if get_normal:
    nm = extras['normal'] = torch.empty((N, S, S, 3), dtype=torch.float, device='cuda')
  check(lib().fpb_render_crops_ex(mh.handle, ptr(poses), ptr(params), N, S, float(mesh_diameter), float(zthr), ptr(ws),
                                  ptr(out_A), _DT_CODE[out_A.dtype], out_A.shape[2],
                                  (out_A.shape[2] - S) // 2, ptr(c), ptr(x), ptr(t), 1 if use_light else 0,
                                  float(w_ambient), float(w_diffuse), ptr(nm), stream()))
  return out_A, extras


def warp_crops(rgb_t, xyz_map_t, depth_t, poses, params, K, S, mesh_diameter, mode, out_B=None, bf16=True):
  """Observed-side crops -> network input B ([N,S,S,8]).  rgb_t: cuda uint8 or float [H,W,3] (0..255)."""
  poses = _as_cuda_f32(poses).reshape(-1, 4, 4)
  N = len(poses)
  H, W = rgb_t.shape[:2]
  if out_B is None:
    out_B = torch.empty((N, S, S, 8), dtype=torch.bfloat16 if bf16 else torch.float, device='cuda')
  K9 = f32_host(torch.as_tensor(K).detach().cpu().numpy() if torch.is_tensor(K) else K, 9)
  is_u8 = rgb_t.dtype == torch.uint8
  if not is_u8:
    rgb_t = rgb_t.float().contiguous()
  check(lib().fpb_warp_crops(ptr(rgb_t), 1 if is_u8 else 0, ptr(xyz_map_t), ptr(depth_t), H, W, ptr(poses), ptr(params),
                             hptr(K9), N, S, float(mesh_diameter), mode, ptr(out_B),
                             _DT_CODE[out_B.dtype], out_B.shape[2], (out_B.shape[2] - S) // 2,
                             stream()))
  return out_B


def warp_depth_nearest(depth, params, S):
  """kornia warp_perspective(depth, tf_to_crops, dsize=(S,S), mode='nearest', align_corners=False) of
  predict_score.py:93 -> [N,1,S,S] fp32.  Runs through the refiner-mode warp kernel on the map (0, 0, depth) with a
  zero translation and diameter 2^11: the kernel's (v - 0) * 2/d and |.| >= 2 test are then an exact power-of-two
  scaling (undone here) and a no-op for depths below 2048 m; depths below 0.001 come out as 0."""
  d = _as_cuda_f32(depth)
  H, W = d.shape
  N = len(params)
  xyz = torch.zeros((H, W, 3), dtype=torch.float, device=d.device)
  xyz[..., 2] = d
  rgb = torch.zeros((H, W, 3), dtype=torch.uint8, device=d.device)
  poses = torch.eye(4, dtype=torch.float, device=d.device).expand(N, 4, 4).contiguous()
  out = torch.empty((N, S, S, 8), dtype=torch.float, device=d.device)
  K9 = f32_host(np.eye(3), 9)
  check(lib().fpb_warp_crops(ptr(rgb), 1, ptr(xyz), None, H, W, ptr(poses), ptr(params.contiguous()), hptr(K9), N, S, 2048.0,
                             0, ptr(out), 0, S, 0, stream()))
  return (out[..., 5] * 1024.0).reshape(N, 1, S, S)


TRANS_REPS = {'tracknet': 0, 'deepim': 1}
ROT_REPS = {'axis_angle': 0, '6d': 1}


def pose_update(poses, trans, rot, mesh_diameter, rot_normalizer, trans_normalizer=None, normalize_xyz=True,
                trans_rep='tracknet', rot_rep='axis_angle', crop_params=None, K=None, S=160, return_deltas=False):
  """predict_pose_refine.py:195-231 + egocentric_delta_pose_to_pose (src/Utils.py:848-855) in one kernel.
  trans_rep='deepim' (:201-215) needs the iteration's crop parameters, K and the crop size; rot_rep='6d' (:223-224)
  takes rot [N,6].  return_deltas: also (trans_delta [N,3], rot_mat_delta [N,3,3]) as the reference stores them in
  last_trans_update / last_rot_update (:238-239)."""
  poses = _as_cuda_f32(poses).reshape(-1, 4, 4)
  out = torch.empty_like(poses)
  if return_deltas:
    if trans_rep not in TRANS_REPS or rot_rep not in ROT_REPS:
      raise RuntimeError(f'unknown trans_rep / rot_rep: {trans_rep} / {rot_rep}')
    rot = rot.contiguous()
    assert rot.shape == (len(poses), 6 if rot_rep == '6d' else 3)
    tn_ = f32_host(trans_normalizer, 3) if trans_normalizer is not None else None
    K9_ = f32_host(torch.as_tensor(K).detach().cpu().numpy() if torch.is_tensor(K) else K, 9) if K is not None else None
    td = torch.empty((len(poses), 3), dtype=torch.float, device=poses.device)
    rd = torch.empty((len(poses), 3, 3), dtype=torch.float, device=poses.device)
    check(lib().fpb_pose_update_deltas(ptr(poses), ptr(trans.contiguous()), ptr(rot), len(poses), float(mesh_diameter),
                                       float(rot_normalizer), hptr(tn_), 1 if normalize_xyz else 0, TRANS_REPS[trans_rep],
                                       ROT_REPS[rot_rep], ptr(crop_params), hptr(K9_), int(S), ptr(out), ptr(td), ptr(rd),
                                       stream()))
    return out, td, rd
  tn = f32_host(trans_normalizer, 3) if trans_normalizer is not None else None
  if trans_rep not in TRANS_REPS or rot_rep not in ROT_REPS:
    raise RuntimeError(f'unknown trans_rep / rot_rep: {trans_rep} / {rot_rep}')
  rot = rot.contiguous()
  assert rot.shape == (len(poses), 6 if rot_rep == '6d' else 3)
  if trans_rep == 'tracknet' and rot_rep == 'axis_angle':
    check(lib().fpb_pose_update(ptr(poses), ptr(trans.contiguous()), ptr(rot), len(poses),
                                float(mesh_diameter), float(rot_normalizer), hptr(tn), 1 if normalize_xyz else 0,
                                ptr(out), stream()))
    return out
  K9 = f32_host(torch.as_tensor(K).detach().cpu().numpy() if torch.is_tensor(K) else K, 9) if K is not None else None
  check(lib().fpb_pose_update_ex(ptr(poses), ptr(trans.contiguous()), ptr(rot), len(poses), float(mesh_diameter),
                                 float(rot_normalizer), hptr(tn), 1 if normalize_xyz else 0, TRANS_REPS[trans_rep],
                                 ROT_REPS[rot_rep], ptr(crop_params), hptr(K9), int(S), ptr(out), stream()))
  return out


def rank_poses(scores, poses, tf_to_center):
  """argsort(descending) + gathers + best pose of FoundationPose.register (src/estimater.py:221-236) in one launch.
  Returns (ids int64 [N], scores sorted [N], poses sorted [N,4,4], best [4,4] = poses[ids[0]] @ tf_to_center)."""
  scores = scores.contiguous()
  poses = poses.reshape(-1, 4, 4).contiguous()
  N = len(scores)
  if N > 4096:                    # beyond the one-block kernel: library sort
    ids = scores.argsort(descending=True, stable=True)
    return ids, scores[ids], poses[ids], poses[ids[0]] @ tf_to_center
  ids = torch.empty(N, dtype=torch.long, device=scores.device)
  s_out = torch.empty_like(scores)
  p_out = torch.empty_like(poses)
  best = torch.empty((4, 4), dtype=torch.float, device=scores.device)
  check(lib().fpb_rank_poses(ptr(scores), ptr(poses), ptr(tf_to_center.contiguous()), N, ptr(ids), ptr(s_out), ptr(p_out), ptr(best),
                             stream()))
  return ids, s_out, p_out, best


def nvdiffrast_render(K=None, H=None, W=None, ob_in_cams=None, glctx=None, context='cuda', get_normal=False,
                      mesh_tensors=None, mesh=None, projection_mat=None, bbox2d=None, output_size=None, use_light=False,
                      light_color=None, light_dir=np.array([0, 0, 1]), light_pos=np.array([0, 0, 0]), w_ambient=0.8,
                      w_diffuse=0.5, extra={}):
  """reference: src/Utils.py:133-219.  glctx / context are accepted and ignored (there is no nvdiffrast context).
  Returns (color [n,h,w,3], depth [n,h,w], normal_map [n,h,w,3] | None); extra['xyz_map'] = camera-frame xyz.
  Supported: bbox2d + square output_size (the hot path) or the full frame (bbox2d=None, rendered into a square of
  max(H, W) and cut back to [H,W]), use_light on/off with any w_ambient / w_diffuse, get_normal.  The default head
  light (light_dir=(0,0,1), light_color=None) is the only light implemented: other values raise instead of being
  silently ignored, as does a caller-supplied projection_mat."""
  if projection_mat is not None or light_color is not None:
    raise NotImplementedError('projection_mat / light_color are not supported')
  if use_light and (light_dir is None or not np.array_equal(np.asarray(light_dir, dtype=np.float64).reshape(-1), [0, 0, 1])):
    raise NotImplementedError('only the default head light light_dir=(0,0,1) is implemented (light_pos is unused with it)')
  if mesh_tensors is None:
    mesh_tensors = make_mesh_tensors(mesh)
  poses = _as_cuda_f32(ob_in_cams).reshape(-1, 4, 4)
  N = len(poses)
  full = bbox2d is None
  if full:
    # the full frame is the crop (0, 0, L, L), L = max(H, W): the same projection evaluated on a square raster whose
    # first H rows / W columns are the frame's pixels
    L = int(max(H, W))
    S = L
    bbox2d = torch.tensor([0.0, 0.0, float(L), float(L)], device='cuda').reshape(1, 4).expand(N, 4)
  else:
    S = int(output_size[0])
    assert int(output_size[1]) == S, 'square crops only'
  K64 = np.asarray(K, dtype=np.float64)
  # rebuild the parameter block from the caller's bbox2d (umin,vmin,umax,vmax)
  params = torch.zeros((N, CROP_STRIDE), dtype=torch.float, device='cuda')
  P = torch.as_tensor(projection_matrix_from_intrinsics(K64, height=H, width=W, znear=0.001, zfar=100),
                      dtype=torch.float, device='cuda')
  G = torch.as_tensor(glcam_in_cvcam, dtype=torch.float, device='cuda')
  params[:, 13:29] = (P[None] @ (G[None] @ poses)).reshape(N, 16)
  bb = _as_cuda_f32(bbox2d).reshape(N, 4)
  l, t, r, b = bb[:, 0], H - bb[:, 1], bb[:, 2], H - bb[:, 3]
  params[:, 29] = W / (r - l); params[:, 30] = H / (t - b)
  params[:, 31] = (W - r - l) / (r - l); params[:, 32] = (H - t - b) / (t - b)
  _, ex = render_crops(mesh_tensors, poses, params, S, 1.0, 0.0, bf16=False, debug=True, use_light=bool(use_light),
                       w_ambient=w_ambient, w_diffuse=w_diffuse, get_normal=bool(get_normal or use_light))
  color, xyz, normal = ex['color'], ex['xyz'], ex.get('normal')
  if full:
    color, xyz = color[:, :H, :W].contiguous(), xyz[:, :H, :W].contiguous()
    normal = normal[:, :H, :W].contiguous() if normal is not None else None
  extra['xyz_map'] = xyz
  return color, xyz[..., 2].contiguous(), normal


# ------------------------------------------------------------------------------------------------
# small torch helpers kept for API compatibility (src/Utils.py:520-546, 836-855)
# ------------------------------------------------------------------------------------------------
def to_homo_torch(pts):
  ones = torch.ones((*pts.shape[:-1], 1), dtype=torch.float, device=pts.device)
  return torch.cat((pts, ones), dim=-1)


def transform_pts(pts, tf):
  if len(tf.shape) >= 3 and tf.shape[-3] != pts.shape[-2]:
    tf = tf[..., None, :, :]
  return (tf[..., :-1, :-1] @ pts[..., None] + tf[..., :-1, -1:])[..., 0]


def transform_dirs(dirs, tf):
  if len(tf.shape) >= 3 and tf.shape[-3] != dirs.shape[-2]:
    tf = tf[..., None, :, :]
  return (tf[..., :3, :3] @ dirs[..., None])[..., 0]


def pose_to_egocentric_delta_pose(A_in_cam, B_in_cam):
  trans_delta = B_in_cam[:, :3, 3] - A_in_cam[:, :3, 3]
  rot_mat_delta = B_in_cam[:, :3, :3] @ A_in_cam[:, :3, :3].permute(0, 2, 1)
  return trans_delta, rot_mat_delta


def egocentric_delta_pose_to_pose(A_in_cam, trans_delta, rot_mat_delta):
  B_in_cam = torch.eye(4, dtype=torch.float, device=A_in_cam.device)[None].expand(len(A_in_cam), -1, -1).contiguous()
  B_in_cam[:, :3, 3] = A_in_cam[:, :3, 3] + trans_delta
  B_in_cam[:, :3, :3] = rot_mat_delta @ A_in_cam[:, :3, :3]
  return B_in_cam


def symmetry_tfs_from_info(info, rot_angle_discrete=5):
  """src/Utils.py:806-834: BOP models_info entry -> [S,4,4] symmetry transforms (identity first; discrete ones with
  millimetre translations converted to metres; one continuous axis sampled every `rot_angle_discrete` degrees)."""
  symmetry_tfs = [np.eye(4)]
  if 'symmetries_discrete' in info:
    tfs = np.array(info['symmetries_discrete'], dtype=np.float64).reshape(-1, 4, 4)
    tfs[..., :3, 3] *= 0.001
    symmetry_tfs = [np.eye(4)] + list(tfs)
  if 'symmetries_continuous' in info:
    axis = np.array(info['symmetries_continuous'][0]['axis']).reshape(3)
    offset = info['symmetries_continuous'][0]['offset']
    rxs, rys, rzs = [0], [0], [0]
    steps = np.arange(0, 360, rot_angle_discrete) / 180.0 * np.pi
    if axis[0] > 0:
      rxs = steps
    elif axis[1] > 0:
      rys = steps
    elif axis[2] > 0:
      rzs = steps
    for rx in rxs:
      for ry in rys:
        for rz in rzs:
          tf = euler_matrix(rx, ry, rz)
          tf[:3, 3] = offset
          symmetry_tfs.append(tf)
  return np.array(symmetry_tfs)


def _pose_errors(fn, pred, gt, model_pts):
  single = np.ndim(pred) == 2
  p = _as_cuda_f32(pred).reshape(-1, 4, 4).contiguous()
  g = _as_cuda_f32(gt).reshape(-1, 4, 4).contiguous()
  if g.shape[0] not in (1, p.shape[0]):
    raise ValueError('gt must be one pose or one per prediction')
  pts = _as_cuda_f32(model_pts).reshape(-1, 3).contiguous()
  out = torch.empty(p.shape[0], dtype=torch.float, device=p.device)
  check(getattr(lib(), fn)(ptr(p), ptr(g), int(g.shape[0] > 1), ptr(pts), pts.shape[0], p.shape[0], ptr(out), stream()))
  return float(out.item()) if single else out


def add_err(pred, gt, model_pts, symetry_tfs=None):
  """src/Utils.py:232-240 (the reference ignores `symetry_tfs` too).  A single 4x4 returns a float, [B,4,4] a CUDA
  tensor [B]."""
  return _pose_errors('fpb_add_err', pred, gt, model_pts)


def adds_err(pred, gt, model_pts):
  """src/Utils.py:242-253."""
  return _pose_errors('fpb_adds_err', pred, gt, model_pts)
