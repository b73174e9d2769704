"""ctypes binding of libfpb200.so (the C ABI in include/fpb200.h).

There is NO CPU fallback: if the shared library is missing, or a compute call is made without a CUDA device,
an exception is raised.
"""
import ctypes
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_PATH = os.environ.get('FPB_LIB_PATH') or os.path.join(_HERE, 'libfpb200.so')   # FPB_LIB_PATH: dev A/B of two builds
_lib = None

vp = ctypes.c_void_p
ci = ctypes.c_int
cf = ctypes.c_float
sz = ctypes.c_size_t

_SIGS = {
  'fpb_version': (ci, []),
  'fpb_launch_count': (ctypes.c_longlong, [ci]),
  'fpb_erode_depth_f32': (ci, [vp, vp, ci, ci, ci, cf, cf, cf, vp]),
  'fpb_bilateral_depth_f32': (ci, [vp, vp, ci, ci, ci, cf, cf, cf, vp]),
  'fpb_depth_to_xyz': (ci, [vp, vp, cf, vp, ci, ci, ci, vp]),
  'fpb_mask_depth_stats': (ci, [vp, vp, ci, ci, vp, vp, vp]),
  'fpb_hypotheses_from_stats': (ci, [vp, vp, vp, vp, ci, ci, vp, vp, vp]),
  'fpb_register_prologue': (ci, [vp, vp, ci, ci, vp, vp, vp, vp, vp, vp, ci, ci, vp, vp, vp, vp, vp]),
  'fpb_cluster_poses_f32': (ci, [vp, ci, vp, ci, cf, cf, vp, vp, vp]),
  'fpb_add_err': (ci, [vp, vp, ci, vp, ci, ci, vp, vp]),
  'fpb_adds_err': (ci, [vp, vp, ci, vp, ci, ci, vp, vp]),
  'fpb_crop_params': (ci, [vp, ci, vp, cf, ci, ci, ci, vp, vp, vp]),
  'fpb_mesh_create': (ci, [vp, vp, vp, ci, vp, ci, vp, vp, ci, vp, vp, ci, ci, vp]),
  'fpb_mesh_destroy': (ci, [vp]),
  'fpb_render_ws_bytes': (sz, [ci, ci, ci]),
  'fpb_render_crops': (ci, [vp, vp, vp, ci, ci, cf, cf, vp, vp, ci, ci, ci, vp, vp, vp, vp]),
  'fpb_render_crops_ex': (ci, [vp, vp, vp, ci, ci, cf, cf, vp, vp, ci, ci, ci, vp, vp, vp, ci, cf, cf, vp, vp]),
  'fpb_warp_crops': (ci, [vp, ci, vp, vp, ci, ci, vp, vp, vp, ci, ci, cf, ci, vp, ci, ci, ci, vp]),
  'fpb_pose_update': (ci, [vp, vp, vp, ci, cf, cf, vp, ci, vp, vp]),
  'fpb_pose_update_ex': (ci, [vp, vp, vp, ci, cf, cf, vp, ci, ci, ci, vp, vp, ci, vp, vp]),
  'fpb_pose_update_deltas': (ci, [vp, vp, vp, ci, cf, cf, vp, ci, ci, ci, vp, vp, ci, vp, vp, vp, vp]),
  'fpb_rank_poses': (ci, [vp, vp, vp, ci, vp, vp, vp, vp, vp]),
  'fpb_net_create': (ci, [vp, ci, ci, vp, sz, vp]),
  'fpb_net_destroy': (ci, [vp]),
  'fpb_net_ws_bytes': (sz, [ci, ci, ci]),
  'fpb_net_blob_floats': (sz, [ci]),
  'fpb_score_cross_ws_bytes': (sz, [ci, ci]),
  'fpb_debug_igemm': (ci, [ci, vp, vp, vp] + [ci] * 12 + [vp, vp, vp, ci, ci, vp]),
  'fpb_debug_igemm_ex': (ci, [ci, vp, vp, vp] + [ci] * 15 + [vp, vp, vp, ci, ci, vp]),
  'fpb_debug_conv3x3_halo': (ci, [ci, vp, vp, vp] + [ci] * 7 + [vp, vp, vp, ci, vp]),
  'fpb_debug_set_precision': (ci, [ci]),
  'fpb_debug_attention': (ci, [vp, vp, ctypes.c_longlong, vp, vp, ci, vp]),
  'fpb_debug_ffn': (ci, [ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ctypes.c_longlong, ci, vp]),
  'fpb_debug_ffn_stamps': (ci, [vp, ci]),
  'fpb_refine_forward': (ci, [vp, vp, ci, vp, vp, vp, vp]),
  'fpb_refine_forward_shared_b': (ci, [vp, vp, ci, vp, vp, vp, vp]),
  'fpb_score_features': (ci, [vp, vp, ci, vp, vp, vp]),
  'fpb_score_features_multi': (ci, [vp, vp, ci, vp, vp, ci, vp]),
  'fpb_store_multi_f32': (ci, [vp, ctypes.c_longlong, vp, ci, vp]),
  'fpb_score_cross': (ci, [vp, vp, ci, ci, vp, vp, vp]),
  'fpb_score_cross_static': (ci, [vp, vp, ci, vp, vp]),
}


def exported_symbols():
  """Names include/fpb200.h declares (used by the CPU-side symbol test)."""
  return ['fpb_last_error'] + list(_SIGS)


def lib():
  global _lib
  if _lib is None:
    if not os.path.exists(_PATH):
      raise RuntimeError(f'{_PATH} is missing: run `python -m foundationpose_b200.build` (no CPU fallback exists)')
    L = ctypes.CDLL(_PATH)
    L.fpb_last_error.restype = ctypes.c_char_p
    for name, (res, args) in _SIGS.items():
      fn = getattr(L, name, None)
      if fn is None:   # reported by tests/test_abi.py; calling it raises AttributeError
        continue
      fn.restype = res
      fn.argtypes = args
    _lib = L
  return _lib


def check(rc):
  if rc != 0:
    raise RuntimeError('libfpb200: ' + lib().fpb_last_error().decode())


def require_cuda():
  if not torch.cuda.is_available():
    raise RuntimeError('foundationpose_b200 needs a CUDA device (sm_100a); there is no CPU fallback')


def stream():
  return torch.cuda.current_stream().cuda_stream


def ptr(t):
  """Device pointer of a contiguous torch tensor (None -> NULL)."""
  if t is None:
    return None
  assert t.is_cuda and t.is_contiguous(), 'expected a contiguous CUDA tensor'
  return t.data_ptr()


def hptr(a):
  """Host pointer of a contiguous numpy array (kept alive by the caller)."""
  return a.ctypes.data_as(vp) if a is not None else None


def f32_host(x, n):
  a = np.ascontiguousarray(np.asarray(x, dtype=np.float64).reshape(-1), dtype=np.float32)
  assert a.size == n
  return a
