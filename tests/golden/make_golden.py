"""Generates tests/golden/*.npz by running the REFERENCE's own code in the authoring container.

  python tests/golden/make_golden.py            (needs /root/reference; the GPU box never runs this)

What is pinned:
  * RefineNet / ScoreNetMultiPair  (learning/models/refine_network.py:26-93, score_network.py:27-90) imported from
    /root/reference with a stub `Utils` module, loaded (strict) with foundationpose_b200.nets.synthetic_state_dict
    and run on seeded pseudo-crops -> outputs frozen.  Inputs and weights are regenerated from their seeds by the
    tests; sha256 digests of both are stored so drift is detected.
  * pure functions AST-extracted from src/Utils.py: depth2xyzmap (:399-417), projection_matrix_from_intrinsics
    (:752-802), egocentric_delta_pose_to_pose (:848-855).
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


def extract_functions(names):
  src = open(os.path.join(REF, 'src', 'Utils.py')).read()
  tree = ast.parse(src)
  ns = {'np': np, 'torch': torch}
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


if __name__ == '__main__':
  main()
