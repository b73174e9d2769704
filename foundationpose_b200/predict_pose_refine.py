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
from .datasets import BatchPoseData, PoseRefinePairH5Dataset, views_of_AB


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


def load_run_config(ckpt_path):
  """weights/<run>/config.yml beside the checkpoint (the reference reads it with OmegaConf, :102); returns a dict or
  None when the file does not exist.  'inf' strings become np.inf like :124-125."""
  path = os.path.join(os.path.dirname(os.path.abspath(ckpt_path)), 'config.yml')
  if not os.path.exists(path):
    return None
  import yaml
  with open(path) as f:
    cfg = yaml.safe_load(f) or {}
  for k, v in list(cfg.items()):
    if isinstance(v, str) and 'inf' in v.lower() and k == 'zfar':
      cfg[k] = np.inf
  return cfg


# the reference's "defaults, to be backward compatible" for keys a run's config.yml may lack (:107-128); they differ
# from the values of the released run (e.g. c_in 4, normalize_xyz False), which is why a found config.yml must win
REFINE_COMPAT_DEFAULTS = dict(use_normal=False, use_mask=False, use_BN=False, c_in=4, crop_ratio=1.2, n_view=1,
                              trans_rep='tracknet', rot_rep='axis_angle', zfar=3, normalize_xyz=False, normal_uint8=False)


def resolve_weights(run_name, cfg_default, compat_defaults, cfg, state_dict, ckpt_path, allow_synthetic, kind):
  """Shared ctor logic of both predictors: locate weights/<run>/model_best.pth like the reference (:97-100), read
  the run's config.yml beside a checkpoint, and REFUSE to fall back to random weights unless the caller asked for
  them (the reference crashes on a missing checkpoint; silently returning plausible-looking poses from random
  weights would be worse).  Returns (cfg, state_dict | None)."""
  out = Config(cfg_default)
  if ckpt_path is None and state_dict is None:
    cand = os.path.join(os.path.dirname(os.path.realpath(__file__)), '..', 'weights', run_name, 'model_best.pth')
    if os.path.exists(cand):
      ckpt_path = cand
  if ckpt_path is not None:
    run_cfg = load_run_config(ckpt_path)
    if run_cfg is not None:
      merged = dict(compat_defaults)
      merged.update(run_cfg)
      if merged.get('crop_ratio') is None:
        merged['crop_ratio'] = compat_defaults.get('crop_ratio', 1.2)
      out.update(merged)
    if state_dict is None:
      state_dict = load_checkpoint_state_dict(ckpt_path)
  if cfg is not None:
    out.update(cfg)
  out['enable_amp'] = True
  if state_dict is None and not allow_synthetic:
    raise FileNotFoundError(
      f'{kind}: no checkpoint at weights/{run_name}/model_best.pth and neither state_dict nor ckpt_path was given. '
      f'Pass allow_synthetic=True (or seed=...) to run on seeded random weights (benchmarks / tests only).')
  if state_dict is None:
    logging.warning(f'{kind}: running on SYNTHETIC seeded random weights -- poses / scores are meaningless')
  return out, state_dict


def make_crop_data_batch(render_size, ob_in_cams, mesh, rgb, depth, K, crop_ratio, xyz_map, normal_map=None,
                         mesh_diameter=None, cfg=None, glctx=None, mesh_tensors=None, dataset=None, out_AB=None,
                         bf16=True, shared_b=False):
  """reference :26-89 with the dataset transform fused in.  Returns a BatchPoseData (datasets.py) whose rgbAs /
  xyz_mapAs / rgbBs / xyz_mapBs are zero-copy [N,3,S,S] views of the network input `pose_data.AB`
  ([2N,S,S+8,8], rendered crops first), already normalised like PairH5Dataset.transform_batch would (rgb / 255,
  xyz centred on the pose translation and divided by the mesh radius, invalid -> 0); `pose_data.params` is the
  [N,40] crop-parameter block.  rgb: uint8 / float cuda [H,W,3]; xyz_map: cuda [H,W,3]."""
  S = int(render_size[0])
  H, W = rgb.shape[:2]
  N = len(ob_in_cams)
  if mesh_tensors is None:
    mesh_tensors = U.make_mesh_tensors(mesh)
  params = U.crop_params(ob_in_cams, K, mesh_diameter, crop_ratio, S, H, W)
  if out_AB is None:
    out_AB = U.net_input_buffer(N, S, torch.bfloat16 if bf16 else torch.float)
  U.render_crops(mesh_tensors, ob_in_cams, params, S, mesh_diameter, 0.001, out_A=out_AB[:N])
  if shared_b:      # every hypothesis has the same translation => the same crop window => ONE observed crop (see predict)
    U.warp_crops(rgb, xyz_map, None, ob_in_cams[:1], params[:1], K, S, mesh_diameter, 0, out_B=out_AB[N:N + 1])
  else:
    U.warp_crops(rgb, xyz_map, None, ob_in_cams, params, K, S, mesh_diameter, 0, out_B=out_AB[N:])
  return batch_from_AB(out_AB, params, ob_in_cams, K, mesh_diameter, S, shared_b)


def batch_from_AB(AB, params, ob_in_cams, K, mesh_diameter, S, shared_b=False):
  N = len(ob_in_cams)
  rgbAs, xyzAs, rgbBs, xyzBs = views_of_AB(AB, N, S, U.IN_XOFF)
  if shared_b:
    rgbBs, xyzBs = rgbBs[:1].expand(N, -1, -1, -1), xyzBs[:1].expand(N, -1, -1, -1)
  poseA = torch.as_tensor(ob_in_cams, dtype=torch.float, device='cuda').reshape(-1, 4, 4)
  pd = BatchPoseData(rgbAs=rgbAs, rgbBs=rgbBs, xyz_mapAs=xyzAs, xyz_mapBs=xyzBs, poseA=poseA,
                     tf_to_crops=params[:, :9].reshape(N, 3, 3))
  # Ks / mesh_diameters cost a host-to-device copy / a fill launch: built on first access, not in the refine loop
  pd.set_lazy('Ks', lambda: torch.as_tensor(np.asarray(K, dtype=np.float32), device='cuda').reshape(1, 3, 3))
  pd.set_lazy('mesh_diameters', lambda: torch.full((N,), float(mesh_diameter), dtype=torch.float, device='cuda'))
  pd.AB, pd.params, pd.transformed = AB, params, True
  return pd


class PoseRefinePredictor:
  def __init__(self, cfg=None, state_dict=None, ckpt_path=None, seed=None, precision='fp16', allow_synthetic=False):
    """The reference ctor takes no argument and reads weights/2023-10-28-18-33-37/{model_best.pth,config.yml}
    (:94-143).  Same here when that directory exists next to the package; `ckpt_path` / `state_dict` name other
    weights, `cfg` overrides single keys.  Without any weights the ctor raises, unless allow_synthetic=True (or an
    explicit seed) asks for seeded random weights -- benchmarks and tests only."""
    self.amp = True
    self.run_name = '2023-10-28-18-33-37'
    synthetic_ok = allow_synthetic or seed is not None
    self.cfg, state_dict = resolve_weights(self.run_name, default_refine_cfg(), REFINE_COMPAT_DEFAULTS, cfg, state_dict,
                                           ckpt_path, synthetic_ok, 'PoseRefinePredictor')
    if self.cfg['trans_rep'] not in U.TRANS_REPS or self.cfg['c_in'] != 6:
      raise NotImplementedError("trans_rep must be 'tracknet' or 'deepim' and c_in must be 6")
    if self.cfg['rot_rep'] not in U.ROT_REPS:
      raise RuntimeError(f"unknown rot_rep {self.cfg['rot_rep']}")              # predict_pose_refine.py:225-226
    if not self.cfg['use_BN'] or self.cfg['use_normal'] or self.cfg.get('use_mask', False):
      raise NotImplementedError('only use_BN=True, use_normal=False, use_mask=False (the released configuration)')
    rot_dim = 6 if self.cfg['rot_rep'] == '6d' else 3
    if state_dict is None:
      state_dict = nets.synthetic_state_dict(nets.REFINE, 1234 if seed is None else seed, rot_out_dim=rot_dim)
    if int(state_dict['rot_head.1.weight'].shape[0]) != rot_dim:
      raise ValueError(f"rot_rep={self.cfg['rot_rep']} needs a {rot_dim}-output rot head, the weights have "
                       f"{int(state_dict['rot_head.1.weight'].shape[0])}")
    self.state_dict = state_dict
    self.precision = precision
    self.model = nets.Net(nets.REFINE, state_dict, precision)
    self.dataset = PoseRefinePairH5Dataset(cfg=self.cfg, h5_file='', mode='test')
    self.last_trans_update = None
    self.last_rot_update = None
    self.last_trans_raw = None        # raw network outputs of the last iteration (output['trans'] / output['rot'])
    self.last_rot_raw = None
    self._ab = None

  def _ab_buffer(self, N, S):
    if self._ab is None or self._ab.shape[0] != 2 * N or self._ab.shape[1] != S:
      self._ab = U.net_input_buffer(N, S, self.model.act_dtype)
    return self._ab

  @torch.inference_mode()
  def predict(self, rgb, depth, K, ob_in_cams, xyz_map, normal_map=None, get_vis=False, mesh=None, mesh_tensors=None,
              glctx=None, mesh_diameter=None, iteration=5, same_translation=None):
    """reference :150-295.  Returns (B_in_cams [N,4,4] float32 cuda, None).  last_trans_update / last_rot_update hold
    trans_delta [N,3] / rot_mat_delta [N,3,3] of the last iteration like the reference (:238-239).
    same_translation (not in the reference): all input poses share one translation, as the hypotheses of register() do
    (src/estimater.py:127-134).  Then the first iteration's box_3d crop window and with it the observed crop B are the same
    for every hypothesis: B is warped and encoded once instead of N times (bit-identical results).  None = detect it when
    the poses arrive as host data (the reference passes numpy); device tensors are not inspected (no synchronisation)."""
    if same_translation is None:
      same_translation = False
      if not (torch.is_tensor(ob_in_cams) and ob_in_cams.is_cuda):
        t = np.asarray(ob_in_cams, dtype=np.float32).reshape(-1, 4, 4)[:, :3, 3]
        same_translation = len(t) > 1 and bool((t == t[:1]).all())
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
    trans = rot = td = rd = None
    for it in range(iteration):
      shared_b = bool(same_translation) and it == 0 and N > 1
      pose_data = make_crop_data_batch(self.cfg['input_resize'], B_in_cams, mesh, rgb_t, None, K, self.cfg['crop_ratio'],
                                       xyz_t, mesh_diameter=mesh_diameter, cfg=self.cfg, mesh_tensors=mesh_tensors,
                                       dataset=self.dataset, out_AB=AB, shared_b=shared_b)
      trans, rot = self.model.refine(pose_data.AB, N, shared_b=shared_b)
      last = it == iteration - 1
      res = U.pose_update(B_in_cams, trans, rot, mesh_diameter, self.cfg['rot_normalizer'], tn,
                          self.cfg['normalize_xyz'], trans_rep=self.cfg['trans_rep'], rot_rep=self.cfg['rot_rep'],
                          crop_params=pose_data.params, K=K, S=S, return_deltas=last)
      if last:
        B_in_cams, td, rd = res
      else:
        B_in_cams = res
    self.last_trans_update, self.last_rot_update = td, rd
    self.last_trans_raw, self.last_rot_raw = trans, rot
    return B_in_cams, None
