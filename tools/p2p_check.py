"""2-GPU check: register_sharded with the fused peer exchange equals the NCCL all_gather variant bit for bit."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
rank = int(os.environ['RANK']); local = int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
from foundationpose_b200 import synthetic, dist as fdist
from bench import make_estimator
sc = synthetic.make_scene(2, n_points=300000)
est, _, _ = make_estimator(sc, 'bf16')
est.rot_grid = est.rot_grid[::4][:63].contiguous()       # 63 hypotheses: ragged shards (32 + 31)
outs = {}
for mode in ('nccl', 'p2p', 'p2p', 'nccl', 'p2p'):
  pose = fdist.register_sharded(est, sc['K'], sc['rgb'], sc['depth'], sc['mask'], iteration=2, exchange=mode)
  outs.setdefault(mode, []).append((pose.copy(), est.scores.cpu().numpy().copy(), est.poses.cpu().numpy().copy()))
ref = outs['nccl'][0]
ok = True
# the sharded result against the unsharded register() of the same estimator (per-hypothesis arithmetic does not depend
# on the batch, up to the fp32 summation order of the pooling kernels)
full = est.register(K=sc['K'], rgb=sc['rgb'], depth=sc['depth'], ob_mask=sc['mask'], iteration=2)
ok &= bool(np.abs(full - ref[0]).max() < 1e-4) and bool(np.abs(est.scores.cpu().numpy() - ref[1]).max() < 1e-3)
for mode, lst in outs.items():
  for p, s, P in lst:
    ok &= np.array_equal(p, ref[0]) and np.array_equal(s, ref[1]) and np.array_equal(P, ref[2])
sys.stdout.write(f'rank{rank}:p2p_equals_nccl={bool(ok)}:best_score={float(ref[1][0]):.4f}\n')
sys.stdout.flush()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
