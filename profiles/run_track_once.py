"""One warmed-up FoundationPose.track_one(iteration=2) frame (BASELINE config 3) between cudaProfilerStart/Stop.
Used under `ncu --profile-from-start off` to list every launch of one frame (see profiles/README.md)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from bench import make_estimator  # noqa: E402
from foundationpose_b200 import synthetic  # noqa: E402

scene = synthetic.make_scene(2)
est, _, _ = make_estimator(scene, os.environ.get('FPB_PRECISION', 'fp16'))
d_rgb, d_depth = (torch.as_tensor(scene[k]).cuda() for k in ('rgb', 'depth'))
est.register(K=scene['K'], rgb=scene['rgb'], depth=scene['depth'], ob_mask=scene['mask'], iteration=5)
for _ in range(3):
  est.track_one(rgb=d_rgb, depth=d_depth, K=scene['K'], iteration=2)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()
est.track_one(rgb=d_rgb, depth=d_depth, K=scene['K'], iteration=2)
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print('done')
