"""cProfile of the host side of one warmed-up register() (dev tool)."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import make_estimator, ITERS  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402

scene = synthetic.make_scene(2)
est, _, _ = make_estimator(scene, os.environ.get('FPB_PRECISION', 'bf16'))
d_rgb, d_depth, d_mask = (torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth', 'mask'))
for _ in range(3):
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
torch.cuda.synchronize()
ts = []
for _ in range(5):
  t0 = time.perf_counter()
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
  t1 = time.perf_counter()
  ts.append((t1 - t0) * 1e3)
print('register wall ms', ts)
pr = cProfile.Profile()
pr.enable()
est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(22)
import time as _t
from foundationpose_b200 import utils as U
t0 = _t.perf_counter(); U.set_seed(0); t1 = _t.perf_counter(); print('set_seed ms', (t1 - t0) * 1e3)
