"""Host-side logic of the multi-GPU path (foundationpose_b200/dist.py) on CPU: world_size-2 gloo."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def test_shard_bounds_cover_and_balance():
  from foundationpose_b200.dist import shard_bounds
  for n in (0, 1, 7, 252, 1008, 2016):
    for world in (1, 2, 3, 4, 8):
      spans = [shard_bounds(n, r, world) for r in range(world)]
      assert spans[0][0] == 0 and spans[-1][1] == n
      assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
      sizes = [hi - lo for lo, hi in spans]
      assert max(sizes) - min(sizes) <= 1


def _free_port():
  s = socket.socket()
  s.bind(('127.0.0.1', 0))
  p = s.getsockname()[1]
  s.close()
  return p


def _worker(rank, world, port, n_total, out):
  os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
  dist.init_process_group('gloo', rank=rank, world_size=world)
  try:
    from foundationpose_b200.dist import shard_bounds, all_gather_ragged
    lo, hi = shard_bounds(n_total, rank, world)
    full = torch.arange(n_total * 3, dtype=torch.float).reshape(n_total, 3)
    got = all_gather_ragged(full[lo:hi].clone(), n_total, world)
    ok = torch.equal(got, full)
    # global top-1 from gathered per-hypothesis scores is identical on every rank
    scores = torch.sin(torch.arange(n_total, dtype=torch.float) * 0.37)
    g = all_gather_ragged(scores[lo:hi].reshape(-1, 1).clone(), n_total, world).reshape(-1)
    best = int(g.argmax())
    out[rank] = (bool(ok), best, int(scores.argmax()))
  finally:
    dist.destroy_process_group()


@pytest.mark.parametrize('n_total', [252, 7, 1])
def test_all_gather_ragged_gloo_world2(n_total):
  world, port = 2, _free_port()
  mgr = mp.Manager()
  out = mgr.dict()
  mp.spawn(_worker, args=(world, port, n_total, out), nprocs=world, join=True)
  assert len(out) == world
  for r in range(world):
    ok, best, ref = out[r]
    assert ok and best == ref
