"""Kernel timeline of one warmed-up register() via torch.profiler (CUPTI): GPU busy time vs wall span, largest gaps (dev tool)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import profile, ProfilerActivity  # noqa: E402
from bench import make_estimator, ITERS  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402

scene = synthetic.make_scene(2)
est, _, _ = make_estimator(scene, os.environ.get('FPB_PRECISION', 'bf16'))
d_rgb, d_depth, d_mask = (torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth', 'mask'))
for _ in range(3):
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
  with torch.profiler.record_function('fpb_register'):
    est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
  torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0, t1 = evs[0].time_range.start, max(e.time_range.end for e in evs)
busy = sum(e.time_range.end - e.time_range.start for e in evs)
reg = [e for e in prof.events() if e.name == 'fpb_register' and e.device_type == torch.autograd.DeviceType.CPU]
if reg:
  r = reg[0].time_range
  print(f'host call {(r.end - r.start) / 1e3:.2f} ms; first device activity {(t0 - r.start) / 1e3:.2f} ms after entry; return {(r.end - t1) / 1e3:.2f} ms after the last device activity')
  first = evs[:12]
  for e in first:
    print(f'   +{(e.time_range.start - r.start) / 1e3:7.3f} ms  {(e.time_range.end - e.time_range.start):7.1f} us  {e.name[:70]}')
print(f'{len(evs)} device activities, span {(t1 - t0) / 1e3:.2f} ms, sum of durations {busy / 1e3:.2f} ms')
gaps = []
end = evs[0].time_range.end
for a, b in zip(evs[:-1], evs[1:]):
  end = max(end, a.time_range.end)
  g = b.time_range.start - end
  if g > 0:
    gaps.append((g, a.name[:60], b.name[:60]))
print(f'total gap {sum(g[0] for g in gaps) / 1e3:.2f} ms over {len(gaps)} gaps')
for g in sorted(gaps, reverse=True)[:25]:
  print(f'  {g[0]:8.1f} us  after {g[1]}  before {g[2]}')
import collections
agg = collections.defaultdict(lambda: [0, 0.0])
for e in evs:
  agg[e.name[:70]][0] += 1; agg[e.name[:70]][1] += e.time_range.end - e.time_range.start
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:25]:
  print(f'{v[1] / 1e3:8.3f} ms  x{v[0]:4d}  {k}')
