"""Host-side phase times of register() (perf_counter at the boundaries of the big calls; no profiler) -- dev tool."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import make_estimator, ITERS  # noqa: E402
from foundationpose_b200 import synthetic, utils as U  # noqa: E402

scene = synthetic.make_scene(2)
est, _, _ = make_estimator(scene, 'bf16')
d_rgb, d_depth, d_mask = (torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth', 'mask'))
marks = []


def wrap(obj, name, tag):
  f = getattr(obj, name)

  def g(*a, **k):
    marks.append((tag + ':in', time.perf_counter()))
    r = f(*a, **k)
    marks.append((tag + ':out', time.perf_counter()))
    return r
  setattr(obj, name, g)


wrap(U, 'set_seed', 'seed')
wrap(U, 'register_prologue', 'prologue')
wrap(est.refiner, 'predict', 'refine')
wrap(est.scorer, 'predict', 'score')
for _ in range(3):
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
torch.cuda.synchronize()
for rep in range(3):
  marks.clear()
  t0 = time.perf_counter()
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
  t1 = time.perf_counter()
  print(f'register {1e3 * (t1 - t0):.2f} ms: ' + ' '.join(f'{k}@{1e3 * (t - t0):.2f}' for k, t in marks))
