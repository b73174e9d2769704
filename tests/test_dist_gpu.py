"""Multi-GPU exchange of the sharded register() on real devices (needs >= 2 GPUs; skipped on a single-GPU box):
the fused peer-memory exchange (pooling kernel stores into every rank's table + one device barrier) gives bit-identical
poses and scores to the NCCL all_gather variant, including ragged shards and alternating phases."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs two GPUs')
def test_p2p_exchange_equals_nccl_on_two_gpus():
  cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
         '--master-port', '29541', os.path.join(ROOT, 'tools', 'p2p_check.py')]
  r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
  assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
  assert 'p2p_equals_nccl=False' not in r.stdout and r.stdout.count('p2p_equals_nccl=True') == 2, r.stdout[-2000:]
