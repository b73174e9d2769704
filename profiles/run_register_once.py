"""One warmed-up FoundationPose.register() of the bench workload (cfg2) between cudaProfilerStart/Stop.
Used under `ncu --profile-from-start off` to list every launch of one step (see profiles/README.md)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import make_estimator, ITERS  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402

scene = synthetic.make_scene(2)
est, _, _ = make_estimator(scene, os.environ.get('FPB_PRECISION', 'fp16'))
d_rgb, d_depth, d_mask = (torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth', 'mask'))
for _ in range(2):
  est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
est.register(K=scene['K'], rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print('done')
