"""ScorePredictor -- drop-in for learning/training/predict_score.py:118-226.

One render + one warp launch build the ScoreNet inputs (the reference's crop->full->crop depth round trip,
h5_dataset.py:137-170, is evaluated as index math inside the warp kernel), the per-hypothesis features come from
the tcgen05 encoder path and the cross-hypothesis attention + Linear run once over all N.
"""
import logging
import os

import numpy as np
import torch

from . import nets
from . import utils as U
from .datasets import ScoreMultiPairH5Dataset
from .predict_pose_refine import Config, batch_from_AB, resolve_weights


def default_score_cfg():
  """weights/2024-01-11-20-02-45/config.yml (absent offline) + reference defaults :126-143."""
  return Config(input_resize=[160, 160], c_in=6, use_BN=True, use_normal=False, normalize_xyz=True, crop_ratio=1.1,
                zfar=np.inf, enable_amp=True)


# predict_score.py:126-141: defaults for keys a run's config.yml may lack
SCORE_COMPAT_DEFAULTS = dict(use_normal=False, use_BN=False, zfar=np.inf, c_in=4, normalize_xyz=False, crop_ratio=1.2)


def make_crop_data_batch(render_size, ob_in_cams, mesh, rgb, depth, K, crop_ratio, normal_map=None, mesh_diameter=None,
                         glctx=None, mesh_tensors=None, dataset=None, cfg=None, out_AB=None, bf16=True):
  """reference :57-114 + TripletH5Dataset.transform_batch (h5_dataset.py:137-179) fused.  Returns a BatchPoseData
  of zero-copy views over the network input `pose_data.AB` ([2N,S,S+8,8]) as the refiner's function does;
  xyz_mapBs is the rebuilt (depth -> full frame -> xyz -> crop) map of h5_dataset.py:158-161.  depthAs / depthBs
  ([N,1,S,S], the raw crop depths of :84-99) are not needed by the fused path and are produced on first access."""
  S = int(render_size[0])
  H, W = rgb.shape[:2]
  N = len(ob_in_cams)
  if mesh_tensors is None:
    mesh_tensors = U.make_mesh_tensors(mesh)
  params = U.crop_params(ob_in_cams, K, mesh_diameter, crop_ratio, S, H, W)
  if out_AB is None:
    out_AB = U.net_input_buffer(N, S, torch.bfloat16 if bf16 else torch.float)
  U.render_crops(mesh_tensors, ob_in_cams, params, S, mesh_diameter, 0.1, out_A=out_AB[:N])
  U.warp_crops(rgb, None, depth, ob_in_cams, params, K, S, mesh_diameter, 1, out_B=out_AB[N:])
  pd = batch_from_AB(out_AB, params, ob_in_cams, K, mesh_diameter, S)
  pd.set_lazy('Ks', lambda: torch.as_tensor(np.asarray(K, dtype=np.float32), device='cuda').reshape(1, 3, 3).expand(N, 3, 3))   # :102
  poses = pd.poseA

  def depth_As():                                                   # :84,92: z of the rendered xyz map
    _, ex = U.render_crops(mesh_tensors, poses, params, S, 1.0, 0.0, bf16=False, debug=True)
    return ex['xyz'][..., 2].reshape(N, 1, S, S).contiguous()

  def depth_Bs():                                                   # :93: nearest warp of the observed depth
    return U.warp_depth_nearest(depth, params, S)

  pd.set_lazy('depthAs', depth_As)
  pd.set_lazy('depthBs', depth_Bs)
  return pd


class ScorePredictor:
  def __init__(self, amp=True, cfg=None, state_dict=None, ckpt_path=None, seed=None, precision='fp16',
               allow_synthetic=False):
    """reference :118-156 reads weights/2024-01-11-20-02-45/{model_best.pth,config.yml}; see PoseRefinePredictor for
    ckpt_path / state_dict / cfg / allow_synthetic."""
    self.amp = amp
    self.run_name = '2024-01-11-20-02-45'
    synthetic_ok = allow_synthetic or seed is not None
    self.cfg, state_dict = resolve_weights(self.run_name, default_score_cfg(), SCORE_COMPAT_DEFAULTS, cfg, state_dict,
                                           ckpt_path, synthetic_ok, 'ScorePredictor')
    if state_dict is None:
      state_dict = nets.synthetic_state_dict(nets.SCORE, 1234 if seed is None else seed)
    if not self.cfg['normalize_xyz'] or self.cfg['c_in'] != 6 or not self.cfg['use_BN'] or self.cfg['use_normal']:
      raise NotImplementedError('only normalize_xyz=True, c_in=6, use_BN=True, use_normal=False (the released configuration)')
    self.state_dict = state_dict
    self.precision = precision
    self.model = nets.Net(nets.SCORE, state_dict, precision)
    self.dataset = ScoreMultiPairH5Dataset(cfg=self.cfg, mode='test', h5_file=None, max_num_key=1)
    self._ab = None
    self.last_features = None

  def _ab_buffer(self, N, S):
    if self._ab is None or self._ab.shape[0] != 2 * N or self._ab.shape[1] != S:
      self._ab = U.net_input_buffer(N, S, self.model.act_dtype)
    return self._ab

  @torch.inference_mode()
  def extract_features(self, rgb, depth, K, ob_in_cams, mesh=None, mesh_tensors=None, mesh_diameter=None, dst_ptrs=None):
    """Per-hypothesis 512-d features (score_network.py:60-77); shardable across ranks.  With dst_ptrs (device
    addresses, e.g. every rank's feature table at this rank's row offset) the rows are stored there by the pooling
    kernel itself and nothing is returned."""
    poses = torch.as_tensor(ob_in_cams, dtype=torch.float, device='cuda').reshape(-1, 4, 4).contiguous()
    N = len(poses)
    rgb_t = torch.as_tensor(rgb, device='cuda')
    if rgb_t.dtype != torch.uint8:
      rgb_t = rgb_t.float()
    rgb_t = rgb_t.contiguous()
    depth_t = torch.as_tensor(depth, device='cuda', dtype=torch.float).contiguous()
    S = int(self.cfg['input_resize'][0])
    AB = self._ab_buffer(N, S)
    make_crop_data_batch(self.cfg['input_resize'], poses, mesh, rgb_t, depth_t, K, self.cfg['crop_ratio'],
                         mesh_diameter=mesh_diameter, mesh_tensors=mesh_tensors, cfg=self.cfg, out_AB=AB)
    if dst_ptrs is not None:
      return self.model.score_features_multi(AB, N, dst_ptrs)
    return self.model.score_features(AB, N)

  @torch.inference_mode()
  def score_from_features(self, feats):
    """Cross-hypothesis attention + Linear over all L = len(feats) hypotheses; scores = logits + 100 (:209)."""
    L = len(feats)
    return self.model.score_cross(feats, L, 1).reshape(-1) + 100

  @torch.inference_mode()
  def predict(self, rgb, depth, K, ob_in_cams, normal_map=None, get_vis=False, mesh=None, mesh_tensors=None, glctx=None,
              mesh_diameter=None):
    """reference :161-226.  Returns (scores [N] float32 cuda, None)."""
    if get_vis:
      logging.info('get_vis is not supported headless; returning vis=None')
    if mesh_tensors is None:
      mesh_tensors = U.make_mesh_tensors(mesh)
    feats = self.extract_features(rgb, depth, K, ob_in_cams, mesh=mesh, mesh_tensors=mesh_tensors,
                                  mesh_diameter=mesh_diameter)
    self.last_features = feats
    return self.score_from_features(feats), None
