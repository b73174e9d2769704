"""Generates tests/golden/trace_cfg2.npz: the CPU oracle's full register() on BASELINE config 2 at the benchmarked
shape (252 hypotheses, 5 refine iterations, scoring), frozen per iteration so that the GPU tests can compare the CUDA
path stage by stage on the ORACLE's input poses (tests/test_register_cfg2_gpu.py).

  python tests/golden/make_trace_cfg2.py        (CPU only; ~5 min on 8 cores; does not need /root/reference)

The oracle's networks are oracle/nets.py, which tests/test_golden_cpu.py pins to the reference's own RefineNet /
ScoreNetMultiPair (tests/golden/make_golden.py).  Contents:
  it{k}_poses_in / _trans / _rot / _poses_out   k = 0..4, [252,...] (learning/training/predict_pose_refine.py:182-234)
  score_poses / score_feats / score_scores      scorer inputs / per-hypothesis features / logits + 100
  pool_poses / pool_feats                       a second set of 252 scorer evaluations (the poses after iteration 1):
                                                together with score_* a pool of 504 (pose, feature) pairs for the
                                                top-1-agreement trials
  it{k}_trans_ac16 / _rot_ac16, *_feats_ac16    the same modules under torch.autocast('cpu', fp16) on the same inputs: the
                                                reference's own fp16 deviation from fp32, for context
  sha_scene / sha_refine / sha_score            digests of the scene arrays and both state dicts (drift detection)
"""
import hashlib
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def digest(arrs):
  h = hashlib.sha256()
  for a in arrs:
    h.update(np.ascontiguousarray(a).tobytes())
  return h.hexdigest()


def scene_digest(sc):
  return digest([sc['rgb'], sc['depth'], sc['mask'].astype(np.uint8), sc['K'], np.float64(sc['diameter'])])


def state_digest(sd):
  return digest([sd[k].detach().cpu().numpy() for k in sorted(sd)])


def main():
  from foundationpose_b200 import nets, synthetic
  from oracle.pipeline import OracleEstimator
  from tests.helpers import centered, mesh_np
  torch.set_num_threads(len(os.sched_getaffinity(0)))
  sc = synthetic.make_scene(2, n_points=600000)
  sd_r, sd_s = nets.synthetic_state_dict(nets.REFINE), nets.synthetic_state_dict(nets.SCORE)
  mesh, _ = centered(sc)
  orc = OracleEstimator(mesh_np(mesh), sc['diameter'], sd_r, sd_s)
  t0 = time.time()
  orc.register(sc['K'], sc['rgb'], sc['depth'], sc['mask'], iteration=5, keep_trace=True)
  print('register', time.time() - t0, 's', orc.timings, flush=True)
  out = {}
  for k, st in enumerate(orc.trace['refine']):
    for name in ('poses_in', 'trans', 'rot', 'poses_out'):
      out[f'it{k}_{name}'] = np.asarray(st[name], np.float32)
  tr = orc.trace['score']
  out['score_poses'], out['score_feats'], out['score_scores'] = tr['poses'], tr['feats'], tr['scores']
  d = orc.preprocess(sc['depth'])
  pool = out['it0_poses_out']
  _, feats = orc.score_predict(sc['rgb'], d, pool, sc['K'])
  out['pool_poses'], out['pool_feats'] = pool, feats
  # The reference's OWN reduced-precision behaviour, for context: the same modules under torch.autocast fp16 (what
  # predict_pose_refine.py:190 / predict_score.py:193 do on CUDA), on the fp32 trajectory's inputs.
  from oracle import nets as on
  from oracle import ops as oo
  xyz_map = oo.depth2xyzmap(d, sc['K'])
  for k in range(5):
    A, B, _ = orc.refine_inputs(sc['rgb'], xyz_map, out[f'it{k}_poses_in'], sc['K'])
    t16, r16 = on.forward_chunked(orc.refine, 'refine', A, B, orc.net_chunk, autocast=torch.float16)
    out[f'it{k}_trans_ac16'], out[f'it{k}_rot_ac16'] = t16, r16
    print('autocast fp16 refine iter', k, np.abs(t16 - out[f'it{k}_trans']).max(), np.abs(r16 - out[f'it{k}_rot']).max(), flush=True)
  for name, poses in (('score', tr['poses']), ('pool', pool)):
    A, B, _ = orc.score_inputs(sc['rgb'], d, poses, sc['K'])
    _, f16 = on.forward_chunked(orc.score, 'score', A, B, orc.net_chunk, autocast=torch.float16)
    out[f'{name}_feats_ac16'] = f16
    print('autocast fp16 feats', name, np.abs(f16 - out[f'{name}_feats']).max(), np.abs(f16 - out[f'{name}_feats']).mean(), flush=True)
  out['sha_scene'], out['sha_refine'], out['sha_score'] = scene_digest(sc), state_digest(sd_r), state_digest(sd_s)
  np.savez_compressed(os.path.join(HERE, 'trace_cfg2.npz'), **out)
  print('done', time.time() - t0, 's', {k: getattr(v, 'shape', v) for k, v in out.items()})


if __name__ == '__main__':
  main()
