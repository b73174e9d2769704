"""Generates tests/golden/*.npz by running the REFERENCE's own code in the authoring container.

  python tests/golden/make_golden.py            (needs /root/reference; the GPU box never runs this)

What is pinned:
  * RefineNet / ScoreNetMultiPair  (learning/models/refine_network.py:26-93, score_network.py:27-90) imported from
    /root/reference with a stub `Utils` module, loaded (strict) with foundationpose_b200.nets.synthetic_state_dict
    and run on seeded pseudo-crops -> outputs frozen.  Inputs and weights are regenerated from their seeds by the
    tests; sha256 digests of both are stored so drift is detected.
  * pure functions AST-extracted from src/Utils.py: depth2xyzmap (:399-417), projection_matrix_from_intrinsics
    (:752-802), egocentric_delta_pose_to_pose (:848-855).
  * pins.npz: src/transform.py PoseTransformer imported as-is (both unit settings, both gimbal-lock branches), and
    AST-extracted add_err / adds_err (:232-253, scipy cKDTree), to_homo_torch / transform_pts / transform_dirs
    (:520-546), symmetry_tfs_from_info (:806-834; its `euler_matrix` is the absent `transformations` package, so the
    restatement foundationpose_b200.utils.euler_matrix is injected -- that one function stays unpinned).
"""
import ast
import hashlib
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference'
sys.path.insert(0, ROOT)


def digest(arrs):
  h = hashlib.sha256()
  for a in arrs:
    h.update(np.ascontiguousarray(a).tobytes())
  return h.hexdigest()


def net_inputs(seed, n):
  """Seeded pseudo-crops [n,6,160,160] x2 (A rendered-like, B observed-like), fp32 NCHW."""
  from foundationpose_b200.nets import _calibration_batch
  g = torch.Generator().manual_seed(seed)
  x = _calibration_batch(g, n)
  return x[:n].contiguous(), x[n:].contiguous()


def state_digest(sd):
  return digest([sd[k].detach().cpu().numpy() for k in sorted(sd)])


def reference_modules():
  sys.modules['Utils'] = types.ModuleType('Utils')          # the model files only do `from Utils import *`
  sys.path.insert(0, os.path.join(REF, 'learning', 'models'))
  import refine_network, score_network                      # noqa: E401
  return refine_network.RefineNet, score_network.ScoreNetMultiPair


class _Cfg(dict):
  use_BN = True


def extract_functions(names, extra_ns=None):
  src = open(os.path.join(REF, 'src', 'Utils.py')).read()
  tree = ast.parse(src)
  ns = {'np': np, 'torch': torch}
  ns.update(extra_ns or {})
  for node in tree.body:
    if isinstance(node, ast.FunctionDef) and node.name in names:
      exec(compile(ast.Module(body=[node], type_ignores=[]), 'Utils.py', 'exec'), ns)
  return [ns[n] for n in names]


def main():
  from foundationpose_b200 import nets
  torch.set_num_threads(8)
  RefineNet, ScoreNetMultiPair = reference_modules()
  out = {}
  with torch.no_grad():
    sd = nets.synthetic_state_dict(nets.REFINE)
    net = RefineNet(cfg=_Cfg(rot_rep='axis_angle'), c_in=6).eval()
    net.load_state_dict(sd, strict=True)
    A, B = net_inputs(101, 3)
    o = net(A, B)
    out['refine'] = dict(seed=101, n=3, trans=o['trans'].numpy(), rot=o['rot'].numpy(),
                         in_sha=digest([A.numpy(), B.numpy()]), sd_sha=state_digest(sd))
    sd = nets.synthetic_state_dict(nets.SCORE)
    net = ScoreNetMultiPair(cfg=_Cfg(), c_in=6).eval()
    net.load_state_dict(sd, strict=True)
    A, B = net_inputs(202, 5)
    o = net(A, B, L=5)
    feats = net.extract_feat(A, B)
    out['score'] = dict(seed=202, n=5, logits=o['score_logit'].numpy(), feats=feats.numpy(),
                        in_sha=digest([A.numpy(), B.numpy()]), sd_sha=state_digest(sd))
  for k, v in out.items():
    np.savez(os.path.join(HERE, f'{k}_net.npz'), **v)
    print(k, {kk: (vv.shape if hasattr(vv, 'shape') else vv) for kk, vv in v.items()})

  d2x, proj, ego = extract_functions(['depth2xyzmap', 'projection_matrix_from_intrinsics', 'egocentric_delta_pose_to_pose'])
  rng = np.random.default_rng(7)
  depth = rng.uniform(0.3, 1.5, (48, 64)).astype(np.float32)
  depth[rng.random((48, 64)) < 0.1] = 0
  K = np.array([[600.0, 0, 31.5], [0, 590.0, 23.25], [0, 0, 1]])
  xyz = d2x(depth, K)
  P = proj(K, 480, 640, 0.001, 100.0)
  Apose = np.tile(np.eye(4, dtype=np.float32), (4, 1, 1))
  q = rng.normal(size=(4, 3, 3)).astype(np.float32)
  Apose[:, :3, :3] = np.linalg.qr(q)[0]
  Apose[:, :3, 3] = rng.normal(size=(4, 3)).astype(np.float32)
  td = rng.normal(size=(4, 3)).astype(np.float32) * 0.01
  rd = np.linalg.qr(rng.normal(size=(4, 3, 3)).astype(np.float32))[0]
  Bp = ego(torch.from_numpy(Apose), torch.from_numpy(td), torch.from_numpy(rd)).numpy()
  np.savez(os.path.join(HERE, 'utils_fns.npz'), depth=depth, K=K, xyz=xyz, P=P, A=Apose, td=td, rd=rd, B=Bp)
  print('utils_fns ok')
  make_pins()


def rot_zyx(roll, pitch, yaw):
  import math
  cr, sr, cp, sp, cy, sy = math.cos(roll), math.sin(roll), math.cos(pitch), math.sin(pitch), math.cos(yaw), math.sin(yaw)
  Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
  Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
  Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
  return Rz @ Ry @ Rx


def make_pins():
  import contextlib
  import importlib.util
  import io
  import math
  from scipy.spatial import cKDTree
  from foundationpose_b200.utils import euler_matrix
  out = {}
  spec = importlib.util.spec_from_file_location('ref_transform', os.path.join(REF, 'src', 'transform.py'))
  mod = importlib.util.module_from_spec(spec)
  spec.loader.exec_module(mod)
  rng = np.random.default_rng(77)
  poses = []
  for k in range(10):
    P = np.eye(4)
    if k == 8:
      P[:3, :3] = rot_zyx(0.7, math.pi / 2, 0.0)          # gimbal lock, R[2,0] < 0
    elif k == 9:
      P[:3, :3] = rot_zyx(-0.4, -math.pi / 2, 0.0)        # gimbal lock, R[2,0] > 0
    else:
      P[:3, :3] = rot_zyx(*rng.uniform(-1.4, 1.4, 3))
    P[:3, 3] = rng.normal(size=3)
    poses.append(P)
  poses = np.array(poses)
  with contextlib.redirect_stdout(io.StringIO()):
    out['pt_poses'] = poses
    out['pt_inch_deg'] = np.array([mod.PoseTransformer().transform_pose(P) for P in poses], dtype=np.float64)
    out['pt_m_rad'] = np.array([mod.PoseTransformer(False, False).transform_pose(P) for P in poses], dtype=np.float64)
  names = ['to_homo_torch', 'transform_pts', 'transform_dirs', 'add_err', 'adds_err', 'symmetry_tfs_from_info']
  homo, tpts, tdirs, add_err, adds_err, sym = extract_functions(names, dict(cKDTree=cKDTree, euler_matrix=euler_matrix))
  pts = rng.normal(size=(200, 3)) * 0.05
  pred, gt = poses[0].copy(), poses[0].copy()
  gt[:3, :3] = gt[:3, :3] @ rot_zyx(0.05, -0.03, 0.08)
  gt[:3, 3] += [0.004, -0.002, 0.007]
  out['err_pts'], out['err_pred'], out['err_gt'] = pts, pred, gt
  out['add'] = np.float64(add_err(pred, gt, pts))
  out['adds'] = np.float64(adds_err(pred, gt, pts))
  tp = torch.from_numpy(rng.normal(size=(7, 3)).astype(np.float32))
  tf1 = torch.from_numpy(poses[1].astype(np.float32))
  tfb = torch.from_numpy(poses[:4].astype(np.float32))
  out['tp_pts'], out['tp_tf1'], out['tp_tfb'] = tp.numpy(), tf1.numpy(), tfb.numpy()
  out['tp_homo'] = homo(tp).numpy()
  out['tp_single'] = tpts(tp, tf1).numpy()
  out['tp_batched'] = tpts(tp, tfb).numpy()                 # [4,4,4] vs [7,3]: tf.shape[-3] != pts.shape[-2] -> broadcast
  out['td_single'] = tdirs(tp, tf1).numpy()
  out['td_batched'] = tdirs(tp, tfb).numpy()
  T = np.eye(4); T[:3, :3] = rot_zyx(0, 0, math.pi); T[:3, 3] = [10, 20, 30]
  infos = [{}, {'symmetries_discrete': [T.reshape(-1).tolist()]},
           {'symmetries_continuous': [{'axis': [0, 0, 1], 'offset': [0, 0, 0]}]},
           {'symmetries_continuous': [{'axis': [1, 0, 0], 'offset': [0.0, 0.01, 0.0]}],
            'symmetries_discrete': [T.reshape(-1).tolist()]},
           {'symmetries_continuous': [{'axis': [0, 1, 0], 'offset': [0, 0, 0]}]}]
  import json
  out['sym_infos'] = np.array(json.dumps(infos))
  for i, info in enumerate(infos):
    out[f'sym_{i}'] = sym(json.loads(json.dumps(info)), rot_angle_discrete=5 if i != 4 else 30)
  np.savez_compressed(os.path.join(HERE, 'pins.npz'), **out)
  print('pins ok', {k: getattr(v, 'shape', v) for k, v in out.items()})


if __name__ == '__main__':
  main()
