"""PoseRefinePredictor -- drop-in for learning/training/predict_pose_refine.py:94-295.

predict() runs the whole render -> crop/warp -> RefineNet -> SE(3) update loop on the device; each
iteration is: 1 crop-parameter launch, 3 render launches (+1 memset), 1 warp launch, the network, 1 pose
update.  No nvdiffrast / kornia / pytorch3d; no CPU fallback.
"""
import logging
import os
from types import SimpleNamespace

import numpy as np
import torch

from . import nets
from . import utils as U


class Config(dict):
  """dict with attribute access (stands in for the OmegaConf object of the reference)."""
  __getattr__ = dict.__getitem__
  __setattr__ = dict.__setitem__


def default_refine_cfg():
  """Values of weights/2023-10-28-18-33-37/config.yml (absent offline) + reference defaults :102-131."""
  return Config(input_resize=[160, 160], c_in=6, use_BN=True, use_normal=False, use_mask=False, normalize_xyz=True,
                rot_rep='axis_angle', trans_rep='tracknet', trans_normalizer=[0.019999999552965164, 0.019999999552965164, 0.05000000074505806],
                rot_normalizer=0.3490658503988659, crop_ratio=1.2, n_view=1, zfar=np.inf, normal_uint8=False,
                enable_amp=True)


def load_checkpoint_state_dict(path):
  """reference :138-141 -- torch.load + unwrap ['model']."""
  ckpt = torch.load(path, map_location='cpu')
  if 'model' in ckpt:
    ckpt = ckpt['model']
  return ckpt


def make_crop_data_batch(render_size, ob_in_cams, mesh, rgb, depth, K, crop_ratio, xyz_map, normal_map=None,
                         mesh_diameter=None, cfg=None, glctx=None, mesh_tensors=None, dataset=None, out_AB=None,
                         bf16=True):
  """reference :26-89 fused: returns (AB [2N,S,S+8,8] with the rendered crops first and the observed crops second,
  params [N,40]).  rgb uint8/float cuda [H,W,3], xyz_map cuda [H,W,3]."""
  S = int(render_size[0])
  H, W = rgb.shape[:2]
  N = len(ob_in_cams)
  if mesh_tensors is None:
    mesh_tensors = U.make_mesh_tensors(mesh)
  params = U.crop_params(ob_in_cams, K, mesh_diameter, crop_ratio, S, H, W)
  if out_AB is None:
    out_AB = U.net_input_buffer(N, S, torch.bfloat16 if bf16 else torch.float)
  U.render_crops(mesh_tensors, ob_in_cams, params, S, mesh_diameter, 0.001, out_A=out_AB[:N])
  U.warp_crops(rgb, xyz_map, None, ob_in_cams, params, K, S, mesh_diameter, 0, out_B=out_AB[N:])
  return out_AB, params


class PoseRefinePredictor:
  def __init__(self, cfg=None, state_dict=None, ckpt_path=None, seed=1234, precision='bf16'):
    """Reference ctor takes no argument and reads weights/<run>/; offline there are no weights, so:
    state_dict / ckpt_path load real weights, otherwise seeded synthetic ones are generated."""
    self.amp = True
    self.run_name = '2023-10-28-18-33-37'
    self.cfg = default_refine_cfg()
    if cfg is not None:
      self.cfg.update(cfg)
    if ckpt_path is None:
      cand = os.path.join(os.path.dirname(os.path.realpath(__file__)), '..', 'weights', self.run_name, 'model_best.pth')
      if os.path.exists(cand):
        ckpt_path = cand
    if state_dict is None and ckpt_path is not None:
      state_dict = load_checkpoint_state_dict(ckpt_path)
    if self.cfg['trans_rep'] not in U.TRANS_REPS or self.cfg['c_in'] != 6:
      raise NotImplementedError("trans_rep must be 'tracknet' or 'deepim' and c_in must be 6")
    if self.cfg['rot_rep'] not in U.ROT_REPS:
      raise RuntimeError(f"unknown rot_rep {self.cfg['rot_rep']}")              # predict_pose_refine.py:225-226
    rot_dim = 6 if self.cfg['rot_rep'] == '6d' else 3
    if state_dict is None:
      state_dict = nets.synthetic_state_dict(nets.REFINE, seed, rot_out_dim=rot_dim)
    if int(state_dict['rot_head.1.weight'].shape[0]) != rot_dim:
      raise ValueError(f"rot_rep={self.cfg['rot_rep']} needs a {rot_dim}-output rot head, the weights have "
                       f"{int(state_dict['rot_head.1.weight'].shape[0])}")
    self.state_dict = state_dict
    self.precision = precision
    self.model = nets.Net(nets.REFINE, state_dict, precision)
    self.dataset = None
    self.last_trans_update = None
    self.last_rot_update = None
    self._ab = None

  def _ab_buffer(self, N, S):
    if self._ab is None or self._ab.shape[0] != 2 * N or self._ab.shape[1] != S:
      self._ab = U.net_input_buffer(N, S, self.model.act_dtype)
    return self._ab

  @torch.inference_mode()
  def predict(self, rgb, depth, K, ob_in_cams, xyz_map, normal_map=None, get_vis=False, mesh=None, mesh_tensors=None,
              glctx=None, mesh_diameter=None, iteration=5):
    """reference :150-295.  Returns (B_in_cams [N,4,4] float32 cuda, None)."""
    if get_vis:
      logging.info('get_vis is not supported headless; returning vis=None')
    if mesh_tensors is None:
      mesh_tensors = U.make_mesh_tensors(mesh)
    B_in_cams = torch.as_tensor(ob_in_cams, device='cuda', dtype=torch.float).reshape(-1, 4, 4).contiguous()
    N = len(B_in_cams)
    rgb_t = torch.as_tensor(rgb, device='cuda')
    if rgb_t.dtype != torch.uint8:
      rgb_t = rgb_t.float()
    rgb_t = rgb_t.contiguous()
    xyz_t = torch.as_tensor(xyz_map, device='cuda', dtype=torch.float).contiguous()
    S = int(self.cfg['input_resize'][0])
    AB = self._ab_buffer(N, S)
    tn = list(self.cfg['trans_normalizer']) if not isinstance(self.cfg['trans_normalizer'], float) else [self.cfg['trans_normalizer']] * 3
    trans = rot = None
    for _ in range(iteration):
      _, params = make_crop_data_batch(self.cfg['input_resize'], B_in_cams, mesh, rgb_t, None, K, self.cfg['crop_ratio'],
                                       xyz_t, mesh_diameter=mesh_diameter, cfg=self.cfg, mesh_tensors=mesh_tensors,
                                       out_AB=AB)
      trans, rot = self.model.refine(AB, N)
      B_in_cams = U.pose_update(B_in_cams, trans, rot, mesh_diameter, self.cfg['rot_normalizer'], tn,
                                self.cfg['normalize_xyz'], trans_rep=self.cfg['trans_rep'], rot_rep=self.cfg['rot_rep'],
                                crop_params=params, K=K, S=S)
    self.last_trans_update = trans
    self.last_rot_update = rot
    return B_in_cams, None
