"""Diagnostic for tests/test_register_cfg2_gpu.py::test_free_running_...: where the oracle's ranking and the GPU's
free-running ranking agree (rank of the oracle's winner, Spearman rho over the matched hypotheses, oracle rank of the
GPU's winner).  FPB_FFN=0/1 etc. are read by the library."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from foundationpose_b200 import nets, synthetic  # noqa: E402
from foundationpose_b200.estimater import FoundationPose  # noqa: E402
from foundationpose_b200.predict_pose_refine import PoseRefinePredictor  # noqa: E402
from foundationpose_b200.predict_score import ScorePredictor  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else 'fp16'
trace = np.load(os.path.join(ROOT, 'tests', 'golden', 'trace_cfg2.npz'))
sc = synthetic.make_scene(2, n_points=600000)
sd_r, sd_s = nets.synthetic_state_dict(nets.REFINE), nets.synthetic_state_dict(nets.SCORE)
est = FoundationPose(model_pts=sc['mesh'].vertices, model_normals=sc['mesh'].vertex_normals, mesh=sc['mesh'],
                     refiner=PoseRefinePredictor(state_dict=sd_r, precision=prec),
                     scorer=ScorePredictor(state_dict=sd_s, precision=prec), mesh_diameter=sc['diameter'])
est.register(K=sc['K'], rgb=sc['rgb'], depth=sc['depth'], ob_mask=sc['mask'], iteration=5)
got, ref = est.poses.cpu().numpy(), trace['it4_poses_out']
gs = est.scores.cpu().numpy().reshape(-1)
os_ = trace['score_scores'].reshape(-1)
dt = np.abs(got[:, None, :3, 3] - ref[None, :, :3, 3]).max(-1)
dr = 2 * np.arcsin(np.clip(np.linalg.norm((got[:, None, :3, :3].astype(np.float64) - ref[None, :, :3, :3]).reshape(len(got), len(ref), -1),
                                           axis=-1) / (2 * np.sqrt(2.0)), 0, 1))
cost = dr + dt * 10
j = cost.argmin(1)
er, et = dr[np.arange(len(got)), j], dt[np.arange(len(got)), j]
ok = (er < 5e-3) & (et < 5e-4)
tight = (er < 1e-3) & (et < 1e-4)
o_best = int(np.argmax(os_))
r = int(np.argmin(cost[:, o_best]))
rank = lambda a: np.argsort(np.argsort(-a))
sp = lambda a, b: float(np.corrcoef(np.argsort(np.argsort(a)), np.argsort(np.argsort(b)))[0, 1])
print(f'[{prec}] FPB_FFN={os.environ.get("FPB_FFN", "1")}: ok {int(ok.sum())}/252, tight {int(tight.sum())}/252')
print(f'  oracle winner {o_best} -> GPU rank {r}, dev {dr[r, o_best]:.2e} rad {dt[r, o_best]*1e3:.4f} mm')
print(f'  GPU winner matches oracle hyp {j[0]} (dev {er[0]:.2e} rad), oracle rank of it {int(rank(os_)[j[0]])}')
print(f'  spearman(GPU score, oracle score of matched hyp): all {sp(gs, os_[j]):.3f}, ok {sp(gs[ok], os_[j][ok]):.3f}, tight {sp(gs[tight], os_[j][tight]):.3f}')
top = np.argsort(-os_)[:10]
print('  oracle top-10 hyps -> GPU ranks:', [int(np.argmin(cost[:, h])) for h in top], 'devs', [f'{dr[int(np.argmin(cost[:, h])), h]:.1e}' for h in top])
print('  oracle top scores', np.round(np.sort(os_)[::-1][:6], 4), ' GPU top scores', np.round(gs[:6], 4))
