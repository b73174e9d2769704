"""Pins the oracle against the golden vectors produced by the REFERENCE's own modules
(tests/golden/make_golden.py, run once in the authoring container where /root/reference exists).  CPU only."""
import hashlib
import os

import numpy as np
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
TOL = 2e-5      # fp32 torch-CPU vs fp32 torch-CPU; only thread-count dependent summation order differs


def _inputs(seed, n):
  from foundationpose_b200.nets import _calibration_batch
  x = _calibration_batch(torch.Generator().manual_seed(seed), n)
  return x[:n].contiguous(), x[n:].contiguous()


def test_oracle_refinenet_matches_reference_golden():
  from foundationpose_b200 import nets
  from oracle import nets as on
  g = np.load(os.path.join(GOLD, 'refine_net.npz'))
  A, B = _inputs(int(g['seed']), int(g['n']))
  net = on.build('refine', nets.synthetic_state_dict(nets.REFINE))     # strict load: state_dict key parity
  with torch.no_grad():
    o = net(A, B)
  assert np.abs(o['trans'].numpy() - g['trans']).max() < TOL
  assert np.abs(o['rot'].numpy() - g['rot']).max() < TOL
  assert np.abs(g['trans']).max() > 1e-3 and np.abs(g['rot']).max() > 1e-3      # not a degenerate check


def test_oracle_scorenet_matches_reference_golden():
  from foundationpose_b200 import nets
  from oracle import nets as on
  g = np.load(os.path.join(GOLD, 'score_net.npz'))
  A, B = _inputs(int(g['seed']), int(g['n']))
  net = on.build('score', nets.synthetic_state_dict(nets.SCORE))
  with torch.no_grad():
    o = net(A, B, L=int(g['n']))
    f = net.extract_feat(A, B)
  assert np.abs(f.numpy() - g['feats']).max() < TOL
  assert np.abs(o['score_logit'].numpy() - g['logits']).max() < 10 * TOL
  assert np.ptp(g['logits']) > 1e-3


def test_oracle_pure_functions_match_reference_golden():
  from oracle import geometry as og
  from oracle import ops as oo
  g = np.load(os.path.join(GOLD, 'utils_fns.npz'))
  assert np.array_equal(oo.depth2xyzmap(g['depth'], g['K']), g['xyz'])            # bit-exact (float64 math -> f32)
  assert np.array_equal(og.projection_matrix_from_intrinsics(g['K'], 480, 640, 0.001, 100.0), g['P'])
  # egocentric_delta_pose_to_pose: B = [rd @ R_A | t_A + td]
  Bp = np.tile(np.eye(4, dtype=np.float32), (len(g['A']), 1, 1))
  Bp[:, :3, 3] = g['A'][:, :3, 3] + g['td']
  Bp[:, :3, :3] = (torch.from_numpy(g['rd']) @ torch.from_numpy(g['A'][:, :3, :3])).numpy()
  assert np.abs(Bp - g['B']).max() < 1e-6


def test_golden_files_have_generator():
  assert os.path.exists(os.path.join(GOLD, 'make_golden.py'))
  for f in ('refine_net.npz', 'score_net.npz', 'utils_fns.npz'):
    assert os.path.getsize(os.path.join(GOLD, f)) < 200_000
