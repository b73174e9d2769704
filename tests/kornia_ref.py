"""TEST INFRASTRUCTURE: kornia 0.7.2 `geometry.transform.warp_perspective` restated around the REAL
torch.nn.functional.grid_sample (requirements.txt:81; call sites predict_pose_refine.py:63-76, predict_score.py:89-99,
h5_dataset.py:89-104,147-161).  kornia itself is absent offline; its sampler (ATen grid_sample) is not, so only the
coordinate normalisation is restated here (kornia/geometry/transform/imgwarp.py `warp_perspective`,
kornia/geometry/conversions.py `normalize_homography` / `normal_transform_pixel`,
kornia/geometry/linalg.py `transform_points`, kornia/utils/grid.py `create_meshgrid`)."""
import torch
import torch.nn.functional as F


def normal_transform_pixel(height, width, eps=1e-14, dtype=torch.float32):
  tr = torch.tensor([[1.0, 0.0, -1.0], [0.0, 1.0, -1.0], [0.0, 0.0, 1.0]], dtype=dtype)
  wd = eps if width == 1 else width - 1.0
  hd = eps if height == 1 else height - 1.0
  tr[0, 0] = tr[0, 0] * 2.0 / wd
  tr[1, 1] = tr[1, 1] * 2.0 / hd
  return tr[None]


def normalize_homography(M, dsize_src, dsize_dst):
  src_h, src_w = dsize_src
  dst_h, dst_w = dsize_dst
  src_norm_trans_src_pix = normal_transform_pixel(src_h, src_w, dtype=M.dtype)
  src_pix_trans_src_norm = torch.inverse(src_norm_trans_src_pix)
  dst_norm_trans_dst_pix = normal_transform_pixel(dst_h, dst_w, dtype=M.dtype)
  return dst_norm_trans_dst_pix @ (M @ src_pix_trans_src_norm)


def create_meshgrid(height, width, dtype=torch.float32):
  xs = torch.linspace(-1, 1, width, dtype=dtype)
  ys = torch.linspace(-1, 1, height, dtype=dtype)
  base = torch.stack(torch.meshgrid([xs, ys], indexing='ij'), dim=-1)     # W x H x 2
  return base.permute(1, 0, 2).unsqueeze(0)                                # 1 x H x W x 2


def transform_points(trans, points, eps=1e-8):
  ph = F.pad(points, [0, 1], 'constant', 1.0)
  out = (trans @ ph[..., None])[..., 0]
  z = out[..., -1:]
  scale = torch.where(z.abs() > eps, 1.0 / (z + eps), torch.ones_like(z))
  return scale * out[..., :-1]


def warp_perspective(src, M, dsize, mode='bilinear', padding_mode='zeros', align_corners=False):
  B, _, H, W = src.shape
  h_out, w_out = dsize
  dst_norm_trans_src_norm = normalize_homography(M, (H, W), (h_out, w_out))
  src_norm_trans_dst_norm = torch.inverse(dst_norm_trans_src_norm)
  grid = create_meshgrid(h_out, w_out, dtype=src.dtype).repeat(B, 1, 1, 1)
  grid = transform_points(src_norm_trans_dst_norm[:, None, None], grid)
  return F.grid_sample(src, grid, align_corners=align_corners, mode=mode, padding_mode=padding_mode)


def depth2xyzmap_batch(depths, Ks, zfar):
  """src/Utils.py:420-438 with the hard-coded .cuda() removed (device-agnostic restatement)."""
  bs = depths.shape[0]
  invalid = (depths < 0.001) | (depths > zfar)
  H, W = depths.shape[-2:]
  vs, us = torch.meshgrid(torch.arange(0, H), torch.arange(0, W), indexing='ij')
  vs = vs.reshape(-1).float()[None].expand(bs, -1)
  us = us.reshape(-1).float()[None].expand(bs, -1)
  zs = depths.reshape(bs, -1)
  Ks = Ks[:, None].expand(bs, zs.shape[-1], 3, 3)
  xs = (us - Ks[..., 0, 2]) * zs / Ks[..., 0, 0]
  ys = (vs - Ks[..., 1, 2]) * zs / Ks[..., 1, 1]
  pts = torch.stack([xs, ys, zs], dim=-1)
  xyz = pts.reshape(bs, H, W, 3).clone()
  xyz[invalid] = 0
  return xyz


def transform_xyz(xyz_map, poseA, mesh_radius, zthr):
  """h5_dataset.py:94-100 / :150-156 with normalize_xyz=True (zthr 0.001 for the pair, 0.1 for the triplet dataset)."""
  bs = len(xyz_map)
  invalid = xyz_map[:, 2:3] < zthr
  x = xyz_map - poseA[:, :3, 3].reshape(bs, 3, 1, 1)
  x = x * (1 / mesh_radius.reshape(bs, 1, 1, 1))
  invalid = invalid.expand(bs, 3, -1, -1) | (x.abs() >= 2)
  x[invalid] = 0
  return x
