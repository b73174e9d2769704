"""One warmed-up register() of a 32-hypothesis shard (what each rank of an 8-GPU sharded register() executes, cfg 2)
between cudaProfilerStart/Stop -- launch list of the strong-scaling per-rank step."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import make_estimator, ITERS  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402

scene = synthetic.make_scene(2)
est, _, _ = make_estimator(scene, os.environ.get('FPB_PRECISION', 'fp16'))
n = int(os.environ.get('FPB_SHARD', '32'))
est.rot_grid = est.rot_grid[:n].contiguous()
d_rgb, d_depth, d_mask = (torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth', 'mask'))
for _ in range(3):
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.cudart().cudaProfilerStart()
e0.record()
est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
e1.record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print('done: %d hypotheses, %.3f ms' % (n, e0.elapsed_time(e1)))
