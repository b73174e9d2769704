"""BatchPoseData and the inference-time dataset objects of the predictors.

  BatchPoseData                          learning/datasets/pose_dataset.py:67-134
  PairH5Dataset.transform_batch          learning/datasets/h5_dataset.py:118-127   (PoseRefinePairH5Dataset :210-219)
  TripletH5Dataset.transform_batch       learning/datasets/h5_dataset.py:173-179   (ScoreMultiPairH5Dataset :183-187)

In the reference `make_crop_data_batch` builds a BatchPoseData of NCHW fp32 tensors and hands it to
`dataset.transform_batch`, which normalises it in 8-10 eager passes.  Here the render / warp kernels write the
TRANSFORMED crops straight into the network input buffer AB ([2N,S,FPB_IN_PITCH,8] NHWC, rendered crops first), so
the BatchPoseData returned by this package's `make_crop_data_batch` holds zero-copy NCHW *views* of that buffer and
`transform_batch` is the identity on it.  Fields the fused path never materialises (the raw crop depths of the
scorer) are produced on first access.
"""
import torch


class BatchPoseData:
  """Attribute bag with the reference's field names (pose_dataset.py:67-107)."""
  _FIELDS = ('rgbAs', 'rgbBs', 'depthAs', 'depthBs', 'normalAs', 'normalBs', 'poseA', 'poseB', 'maskAs', 'maskBs',
             'xyz_mapAs', 'xyz_mapBs', 'tf_to_crops', 'crop_masks', 'Ks', 'model_pts', 'mesh_diameters', 'labels')

  def __init__(self, rgbAs=None, rgbBs=None, depthAs=None, depthBs=None, normalAs=None, normalBs=None, maskAs=None,
               maskBs=None, poseA=None, poseB=None, xyz_mapAs=None, xyz_mapBs=None, tf_to_crops=None, Ks=None,
               crop_masks=None, model_pts=None, mesh_diameters=None, labels=None):
    self.rgbAs = rgbAs
    self.rgbBs = rgbBs
    self.depthAs = depthAs
    self.depthBs = depthBs
    self.normalAs = normalAs
    self.normalBs = normalBs
    self.poseA = poseA
    self.poseB = poseB
    self.maskAs = maskAs
    self.maskBs = maskBs
    self.xyz_mapAs = xyz_mapAs
    self.xyz_mapBs = xyz_mapBs
    self.tf_to_crops = tf_to_crops
    self.crop_masks = crop_masks
    self.Ks = Ks
    self.model_pts = model_pts
    self.mesh_diameters = mesh_diameters
    self.labels = labels

  # fields computed on first access: name -> zero-argument callable (set by make_crop_data_batch)
  def set_lazy(self, name, thunk):
    self.__dict__.pop(name, None)
    self.__dict__.setdefault('_lazy', {})[name] = thunk

  def __getattr__(self, name):        # only reached when normal lookup fails
    lazy = self.__dict__.get('_lazy')
    if lazy and name in lazy:
      value = lazy.pop(name)()
      self.__dict__[name] = value
      return value
    raise AttributeError(name)

  def _materialise(self):
    for name in list(self.__dict__.get('_lazy', {})):
      getattr(self, name)

  def _tensor_fields(self):
    self._materialise()
    return [k for k in self._FIELDS if torch.is_tensor(self.__dict__.get(k))]

  def pin_memory(self):
    for k in self._tensor_fields():
      try:
        self.__dict__[k] = self.__dict__[k].pin_memory()
      except Exception:
        pass
    return self

  def cuda(self):
    for k in self._tensor_fields():
      self.__dict__[k] = self.__dict__[k].cuda()
    return self

  def select_by_indices(self, ids):
    """pose_dataset.py:129-134: gathers every non-None field (the result owns copies, not views of AB)."""
    out = BatchPoseData()
    ids = torch.as_tensor(ids)
    n = len(self.poseA) if torch.is_tensor(self.poseA) else None
    for k in self._tensor_fields():
      v = self.__dict__[k]
      # broadcast fields (the refiner's Ks is [1,3,3], predict_pose_refine.py:54) are shared, not gathered
      out.__dict__[k] = v if (n is not None and v.shape[0] == 1 and n != 1) else v[ids.to(v.device)]
    return out


class _FusedDataset:
  """Stands in for the H5 dataset objects the reference predictors construct only to call `transform_batch`
  (predict_pose_refine.py:131, predict_score.py:143)."""

  def __init__(self, cfg=None, h5_file=None, mode='test', max_num_key=None, cache_data=None):
    self.cfg = cfg
    self.mode = mode
    self.h5_file = h5_file

  def transform_batch(self, batch, H_ori=None, W_ori=None, bound=1):
    """Identity on a batch produced by this package's make_crop_data_batch (the normalisation of
    h5_dataset.py:79-127 / :137-179 already happened inside the render / warp kernels)."""
    if not getattr(batch, 'transformed', False):
      raise NotImplementedError('transform_batch expects the BatchPoseData of foundationpose_b200.make_crop_data_batch '
                                '(already normalised by the fused kernels)')
    return batch


class PoseRefinePairH5Dataset(_FusedDataset):
  pass


class ScoreMultiPairH5Dataset(_FusedDataset):
  pass


def views_of_AB(AB, N, S, xoff):
  """Zero-copy NCHW views (rgbAs, xyz_mapAs, rgbBs, xyz_mapBs) of the fused network input [2N,S,pitch,8]."""
  img = AB[:, :, xoff:xoff + S, :]
  rgb = img[..., 0:3].permute(0, 3, 1, 2)
  xyz = img[..., 3:6].permute(0, 3, 1, 2)
  return rgb[:N], xyz[:N], rgb[N:], xyz[N:]
