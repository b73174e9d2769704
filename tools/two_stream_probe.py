"""Probe (dev tool): does software-pipelining two half batches on two streams beat one whole batch?
Two estimators with 126 hypotheses each, one thread per estimator (register() synchronises at its end), each on its own
CUDA stream, against one 252-hypothesis register().  Prints the wall time per round of both variants."""
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402
from bench import make_estimator, ITERS  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402

scene = synthetic.make_scene(2)
prec = os.environ.get('FPB_PRECISION', 'fp16')
whole, _, _ = make_estimator(scene, prec)
halves = [make_estimator(scene, prec)[0] for _ in range(2)]
n = len(whole.rot_grid)
halves[0].rot_grid = whole.rot_grid[: n // 2].contiguous()
halves[1].rot_grid = whole.rot_grid[n // 2:].contiguous()
d = {k: torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth', 'mask')}


def reg(est):
  est.register(K=scene['K'], rgb=d['rgb'], depth=d['depth'], ob_mask=d['mask'], iteration=ITERS)


for _ in range(3):
  reg(whole)
  for h in halves:
    reg(h)
torch.cuda.synchronize()
ts = []
for _ in range(10):
  t0 = time.perf_counter()
  reg(whole)
  ts.append(time.perf_counter() - t0)
print('one 252-hypothesis register(): median %.2f ms' % (np.median(ts) * 1e3))
ts = []
for _ in range(10):
  t0 = time.perf_counter()
  for h in halves:
    reg(h)
  ts.append(time.perf_counter() - t0)
print('two 126-hypothesis register() back to back: median %.2f ms' % (np.median(ts) * 1e3))
streams = [torch.cuda.Stream() for _ in range(2)]


def worker(i, barrier, out):
  with torch.cuda.stream(streams[i]):
    for r in range(13):
      barrier.wait()
      t0 = time.perf_counter()
      reg(halves[i])
      out[i].append(time.perf_counter() - t0)


out = [[], []]
bar = threading.Barrier(2)
th = [threading.Thread(target=worker, args=(i, bar, out)) for i in range(2)]
t_all = time.perf_counter()
for t in th:
  t.start()
for t in th:
  t.join()
t_all = time.perf_counter() - t_all
per_round = [max(a, b) for a, b in zip(out[0][3:], out[1][3:])]
print('two 126-hypothesis register() on two streams / threads: median %.2f ms per round (max of the two), total %.1f ms for 13 rounds'
      % (np.median(per_round) * 1e3, t_all * 1e3))
