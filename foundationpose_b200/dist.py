"""Multi-GPU execution of register(): one process per GPU, torch.distributed (NCCL over NVLink / NVSwitch).

Two ways to shard:
  * by hypothesis (`register_sharded`): every rank renders / refines / encodes its contiguous slice of the N
    hypotheses; ScoreNet's cross-hypothesis attention needs all N feature vectors, so the [N/R,512] features and
    the [N/R,4,4] refined poses are all-gathered (two collectives, < 1 MB), then every rank ranks redundantly.
    On NVLink-connected GPUs the exchange is FUSED into the producing kernels (`PeerExchange`): ScoreNet's pooling
    kernel stores each feature row straight into every rank's feature table (peer-mapped symmetric memory), a small
    kernel does the same for the poses, and one device-side barrier orders the step -- no NCCL call on the data path.
  * by object (`gather_best`): each rank runs a full register() of its own object; only the best pose + score per
    object are all-gathered (BASELINE config 5).
The reference has no distributed code; this is new.
"""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import utils as U
from ._lib import lib, check, ptr, stream
from .estimater import _center_from_stats, _h2d

MAX_PEERS = 8      # FPB_MAX_PEERS (include/fpb200.h)


class PeerExchange:
  """Symmetric-memory tables for the sharded register(): per rank [2 phases] x ([N,512] features + [N,16] poses) fp32,
  mapped into every peer (torch.distributed._symmetric_memory: CUDA VMM handles exchanged once at rendezvous).
  Producers write rows of phase p into ALL ranks' tables; `barrier()` (device-side signal-pad barrier, stream
  ordered) makes them visible; consumers read their LOCAL table.  Two phases alternate per call, so a fast rank
  can start writing the next frame's rows while a slow one still reads this frame's: one barrier per frame."""

  def __init__(self, n_total, group=None):
    import torch.distributed._symmetric_memory as symm_mem
    group = group if group is not None else dist.group.WORLD
    self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
    if self.world > MAX_PEERS:
      raise RuntimeError(f'PeerExchange supports up to {MAX_PEERS} ranks (one NVSwitch box)')
    self.n = int(n_total)
    self.row = 512 + 16
    self.buf = symm_mem.empty((2, self.n * self.row), dtype=torch.float, device=torch.device('cuda', torch.cuda.current_device()))
    self.buf.zero_()
    self.hdl = symm_mem.rendezvous(self.buf, group.group_name)
    self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
    self.phase = 0
    torch.cuda.synchronize()
    self.hdl.barrier()

  def _base(self, r, phase):
    return self.ptrs[r] + phase * self.n * self.row * 4

  def feat_ptrs(self, lo):
    """Addresses of row `lo` of every rank's feature table (local first) for the current phase."""
    order = [self.rank] + [r for r in range(self.world) if r != self.rank]
    return [self._base(r, self.phase) + lo * 512 * 4 for r in order]

  def pose_ptrs(self, lo):
    order = [self.rank] + [r for r in range(self.world) if r != self.rank]
    return [self._base(r, self.phase) + (self.n * 512 + lo * 16) * 4 for r in order]

  def publish_poses(self, poses, lo):
    src = poses.reshape(-1).contiguous()
    arr = (ctypes.c_void_p * self.world)(*self.pose_ptrs(lo))
    check(lib().fpb_store_multi_f32(ptr(src), src.numel(), arr, self.world, stream()))

  def finish(self):
    """Barrier, then the local tables of this phase: (features [N,512], poses [N,4,4]); flips the phase."""
    self.hdl.barrier()
    tab = self.buf[self.phase]
    feats = tab[:self.n * 512].view(self.n, 512)
    poses = tab[self.n * 512:].view(self.n, 4, 4)
    self.phase ^= 1
    return feats, poses


_exchanges = {}


def peer_exchange(n_total, group=None):
  """Cached PeerExchange for (group, n_total); raises if symmetric memory cannot be set up."""
  key = (id(group), int(n_total))
  if key not in _exchanges:
    _exchanges[key] = PeerExchange(n_total, group)
  return _exchanges[key]


def shard_bounds(n, rank, world):
  """Contiguous, balanced slices: the first n % world ranks get one extra hypothesis."""
  base, extra = divmod(n, world)
  lo = rank * base + min(rank, extra)
  return lo, lo + base + (1 if rank < extra else 0)


def all_gather_ragged(x, n_total, world, group=None):
  """all_gather of per-rank slices produced by shard_bounds (pads to the largest slice)."""
  if world == 1:
    return x
  cap = -(-n_total // world)
  pad = torch.zeros((cap,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
  pad[:len(x)] = x
  out = torch.empty((world * cap,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
  dist.all_gather_into_tensor(out, pad, group=group)
  parts = []
  for r in range(world):
    lo, hi = shard_bounds(n_total, r, world)
    parts.append(out[r * cap:r * cap + (hi - lo)])
  return torch.cat(parts, dim=0)


@torch.inference_mode()
def register_sharded(est, K, rgb, depth, ob_mask, iteration=5, group=None, exchange='auto', ob_id=None):
  """FoundationPose.register (src/estimater.py:159-240) with the hypotheses sharded over the ranks of `group`.
  Every rank returns the same best pose (np.ndarray [4,4]).
  exchange: 'p2p' = stores into peer-mapped tables fused into the producing kernels + one device barrier
  (PeerExchange); 'nccl' = two all_gather collectives; 'auto' = 'p2p' on CUDA with an NCCL group of <= 8 ranks,
  else 'nccl' (the gloo CPU tests)."""
  world = dist.get_world_size(group) if dist.is_initialized() else 1
  rank = dist.get_rank(group) if dist.is_initialized() else 0
  U.set_seed(0)
  K = np.asarray(K, dtype=np.float64)
  depth_t = _h2d(depth, torch.float)
  rgb_t = _h2d(rgb)
  mask_t = _h2d(ob_mask)
  mask_t = (mask_t.view(torch.uint8) if mask_t.dtype == torch.bool else
            mask_t if mask_t.dtype == torch.uint8 else (mask_t > 0).to(torch.uint8)).contiguous()
  # one-call prologue without a host round trip (as FoundationPose.register): stencils, translation guess and all N
  # hypotheses on the device; the "too few valid pixels" early exit (src/estimater.py:176-180) is evaluated after the
  # single synchronisation at the end -- every rank takes the same branch, so the exchange below stays collective
  depth_t, xyz_map, poses_all, si, sf, _ = U.register_prologue(depth_t, mask_t, K, est.rot_grid, min_valid=4)
  if getattr(est, '_stats_host', None) is None:
    est._stats_host = (torch.empty(6, dtype=torch.int32, pin_memory=True), torch.empty(2, dtype=torch.float, pin_memory=True))
  est._stats_host[0].copy_(si, non_blocking=True)
  est._stats_host[1].copy_(sf, non_blocking=True)
  n_total = len(est.rot_grid)
  if exchange == 'auto':
    exchange = 'p2p' if (world > 1 and world <= MAX_PEERS and dist.get_backend(group) == 'nccl') else 'nccl'
  px = peer_exchange(n_total, group) if (exchange == 'p2p' and world > 1) else None
  lo, hi = shard_bounds(n_total, rank, world)
  poses = poses_all[lo:hi].contiguous()
  if hi > lo:
    poses, _ = est.refiner.predict(mesh=est.mesh, mesh_tensors=est.mesh_tensors, rgb=rgb_t, depth=depth_t, K=K,
                                   ob_in_cams=poses, xyz_map=xyz_map, mesh_diameter=est.diameter, iteration=iteration,
                                   same_translation=True)
    if px is not None:
      est.scorer.extract_features(rgb_t, depth_t, K, poses, mesh=est.mesh, mesh_tensors=est.mesh_tensors,
                                  mesh_diameter=est.diameter, dst_ptrs=px.feat_ptrs(lo))
      px.publish_poses(poses, lo)
    else:
      feats = est.scorer.extract_features(rgb_t, depth_t, K, poses, mesh=est.mesh, mesh_tensors=est.mesh_tensors,
                                          mesh_diameter=est.diameter)
  else:
    feats = torch.zeros((0, 512), dtype=torch.float, device='cuda')
  if px is not None:
    feats, poses = px.finish()
    feats, poses = feats.clone(), poses.clone()       # the tables are overwritten two frames later
  else:
    feats = all_gather_ragged(feats, n_total, world, group)
    poses = all_gather_ragged(poses.reshape(-1, 16), n_total, world, group).reshape(-1, 4, 4)
  scores = est.scorer.score_from_features(feats)
  ids, scores_sorted, poses_sorted, best_t = U.rank_poses(scores, poses, est.get_tf_to_centered_mesh(clone=False))
  best = best_t.data.cpu().numpy()                                                    # the only synchronisation
  oi, of = est._stats_host[0].numpy(), est._stats_host[1].numpy()
  stats = (int(oi[0]), int(oi[1]), int(oi[2]), int(oi[3]), int(oi[4]), int(oi[5]), of[0], of[1])
  if stats[5] < 4:
    pose = np.eye(4)
    pose[:3, 3] = _center_from_stats(stats, K)
    return pose
  est.H, est.W = depth_t.shape[:2]                   # the attributes register() leaves behind (src/estimater.py:192-240)
  est.K, est.ob_id, est.ob_mask = K, ob_id, ob_mask
  est.poses, est.scores = poses_sorted, scores_sorted
  est.pose_last, est.best_id = est.poses[0], ids[0]
  return best


def gather_best(pose, score, group=None):
  """Object-sharded mode: all_gather of each rank's best pose [4,4] and score -> ([R,4,4], [R])."""
  world = dist.get_world_size(group) if dist.is_initialized() else 1
  p = torch.as_tensor(pose, dtype=torch.float, device='cuda').reshape(1, 16)
  s = torch.as_tensor(score, dtype=torch.float, device='cuda').reshape(1)
  if world == 1:
    return p.reshape(1, 4, 4), s
  P = torch.empty((world, 16), dtype=torch.float, device='cuda')
  Sc = torch.empty((world,), dtype=torch.float, device='cuda')
  dist.all_gather_into_tensor(P, p, group=group)
  dist.all_gather_into_tensor(Sc, s, group=group)
  return P.reshape(world, 4, 4), Sc
