#!/usr/bin/env python
"""bench.py -- hypotheses/sec of the render-and-compare path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision fp16|bf16]
  torchrun --nproc-per-node N bench.py --gpus N ...        (one rank per GPU)

A step = one full FoundationPose.register(): depth stencils, hypothesis generation, `iteration` x
[crop windows, batched render, crop/warp, RefineNet, pose update], scorer render/warp, ScoreNet, ranking.
Workload = BASELINE config 2: synthetic 20,480-face mesh, 640x480 RGB-D, 252 hypotheses, 5 refine iterations,
160x160 crops, seeded random-init networks, tcgen05 tensor-core path (fp16 operands -- the reference's autocast dtype
-- by default, fp32 accumulation; --precision bf16 selects bf16 operands, same kernels and speed).
  N = 1 : value = hypotheses/s with the frame already resident in HBM (device tensors passed to register());
          e2e = the same call with HOST numpy buffers (pinned), H2D copies + the D2H read of the pose inside the timed region.
  N > 1 : the north-star partition -- ONE object, its 252 hypotheses sharded over the N ranks (dist.register_sharded;
          "scaling": "strong", 252 hypotheses per step whatever N); value / e2e as above.  The result is checked
          against the unsharded register() on every rank (`matches_unsharded`).  Extras: `replicas` (one full
          register() per rank = BASELINE config 5, 8 distinct meshes sharded by object, only best pose + score gathered).
--impl reference times the CPU restatement of the reference path (oracle/) on the host cores (full 252-hypothesis steps).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_HYP, ITERS, S = 252, 5, 160
FLOP_REFINE, FLOP_SCORE = 23.946e9, 21.938e9       # per hypothesis(-iteration), SURVEY 8(d)
METRIC = 'hypotheses/sec (render+crop+refine+score)'
WORKLOAD = ('cfg2: synthetic 20480-face mesh, 640x480 RGB-D, 252 hyps (42 views x 6 in-plane), est_refine_iter=5, '
            '160x160 crops, full register()')


_STDOUT = sys.stdout


def claim_stdout():
  """The ONE JSON line of the contract goes to the real stdout; everything else that libraries print while the estimator
  is built goes to stderr -- including what C code writes to file descriptor 1 directly (cluster_poses' two printf lines,
  mycpp/src/app/pybind_api.cpp:27,66; NCCL's "NCCL version ..." banner), hence the redirection at descriptor level."""
  global _STDOUT
  sys.stdout.flush()
  real = os.dup(1)
  os.dup2(2, 1)
  _STDOUT = os.fdopen(real, 'w')
  sys.stdout = sys.stderr


def emit(line):
  _STDOUT.write(json.dumps(line) + '\n')
  _STDOUT.flush()


def peaks():
  p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
  if os.path.exists(p):
    d = json.load(open(p))
    return dict(hbm=d['hbm_gbs'], tf_burst=d['bf16_tflops'], tf_sust=d.get('bf16_tflops_sustained', d['bf16_tflops']),
                src='measured')
  return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src='fallback')


class ClockSampler(threading.Thread):
  def __init__(self, index=0):
    super().__init__(daemon=True)
    self.index, self.samples, self.stop_flag = index, [], False

  def run(self):
    q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'
    while not self.stop_flag:
      try:
        out = subprocess.run(['nvidia-smi', '-i', str(self.index), f'--query-gpu={q}', '--format=csv,noheader,nounits'],
                             capture_output=True, text=True, timeout=5).stdout.strip().split(',')
        self.samples.append([x.strip() for x in out])
      except Exception:
        pass
      time.sleep(0.2)

  def summary(self):
    if not self.samples:
      return dict(sm_mhz=None, sm_max_mhz=None, reasons=[])
    sm = [float(s[0]) for s in self.samples if s[0].replace('.', '').isdigit()]
    mx = [float(s[1]) for s in self.samples if s[1].replace('.', '').isdigit()]
    names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
    reasons = sorted({names[i] for s in self.samples for i in range(4) if len(s) > 2 + i and s[2 + i].lower().startswith('active')})
    return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None, reasons=reasons)


def make_estimator(scene, precision='bf16'):
  from foundationpose_b200 import nets
  from foundationpose_b200.estimater import FoundationPose
  from foundationpose_b200.predict_pose_refine import PoseRefinePredictor
  from foundationpose_b200.predict_score import ScorePredictor
  sd_r, sd_s = nets.synthetic_state_dict(nets.REFINE), nets.synthetic_state_dict(nets.SCORE)
  est = FoundationPose(model_pts=scene['mesh'].vertices, model_normals=scene['mesh'].vertex_normals, mesh=scene['mesh'],
                       refiner=PoseRefinePredictor(state_dict=sd_r, precision=precision),
                       scorer=ScorePredictor(state_dict=sd_s, precision=precision), mesh_diameter=scene['diameter'])
  return est, sd_r, sd_s


def cpu_sample(scene, sd_r, sd_s, n_hyp, iters):
  """Oracle (CPU restatement of the reference path) on a bounded sample: n_hyp hypotheses, full register()."""
  import torch
  from oracle.pipeline import OracleEstimator
  from tests.helpers import mesh_np, centered
  cores = len(os.sched_getaffinity(0))
  torch.set_num_threads(cores)
  mesh, _ = centered(scene)
  orc = OracleEstimator(mesh_np(mesh), scene['diameter'], sd_r, sd_s)
  t0 = time.perf_counter()
  orc.register(scene['K'], scene['rgb'], scene['depth'], scene['mask'], iteration=iters, n_hyp=n_hyp)
  dt = time.perf_counter() - t0
  return dict(value=n_hyp / dt, unit='hypotheses/s', cores=cores, kind='port', seconds=dt,
              sample=f'{n_hyp} of {N_HYP} hypotheses through the full register() ({iters} refine iterations + scoring), '
                     f'torch-CPU fp32 networks + C/OpenMP raster/warp/stencil restatement')


def run_reference(args):
  """--impl reference: the CPU restatement of the reference path on all host cores (rank 0 only)."""
  rank = int(os.environ.get('RANK', '0'))
  if rank != 0:
    return
  from foundationpose_b200 import synthetic, nets
  scene = synthetic.make_scene(2)
  sd_r, sd_s = nets.synthetic_state_dict(nets.REFINE), nets.synthetic_state_dict(nets.SCORE)
  n_hyp = args.ref_hyps                 # default: all 252 hypotheses = the same workload as the GPU arm (~1 min per step)
  for _ in range(min(args.warmup, 1)):
    cpu_sample(scene, sd_r, sd_s, 2, 1)
  vals = []
  for _ in range(max(1, min(args.steps, 2))):
    vals.append(cpu_sample(scene, sd_r, sd_s, n_hyp, ITERS))
  best = sorted(vals, key=lambda d: d['seconds'])[len(vals) // 2]
  line = dict(metric=METRIC, value=best['value'], unit='hypotheses/s', n_gpus=args.gpus, steps=len(vals),
              warmup=min(args.warmup, 1), ms_per_step=best['seconds'] * 1e3, higher_is_better=True, scaling='weak',
              vs_baseline=None, dtype='f32', data='synthetic', impl='reference',
              config=dict(workload=WORKLOAD if n_hyp == N_HYP else f'cfg2 sample: {n_hyp} of {N_HYP} hypotheses, 20480-face mesh, '
                                   f'640x480, {ITERS} iters', hypotheses=n_hyp, iterations=ITERS, precision='fp32'),
              cpu_baseline=dict(value=best['value'], unit='hypotheses/s', cores=best['cores'], kind='port', sample=best['sample']),
              e2e=dict(value=best['value'], unit='hypotheses/s', h2d_bytes_per_step=0, d2h_bytes_per_step=0))
  emit(line)


def time_dominant_kernel(torch, n_rep=20):
  """CUDA-event timing of the dominant kernel in isolation: tc2_kernel<256,2> (persistent tcgen05 + TMA im2col,
  cta_group::2) on the 3x3 256->256 conv at 40x40 over all 252 hypotheses (M=403200, N=256, K=2304), the shape that
  runs 4x per network pass; L2 flushed between launches."""
  from foundationpose_b200._lib import lib, check, ptr, stream
  nimg, H, C = N_HYP, 40, 256
  x = torch.randn(nimg, H, H, C, device='cuda').bfloat16()
  w = (torch.randn(C, 9 * C, device='cuda') / 48).bfloat16()
  out = torch.empty(nimg * H * H, C, device='cuda', dtype=torch.bfloat16)
  sc = torch.ones(C, device='cuda'); sh = torch.zeros(C, device='cuda')
  flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
  M, K = nimg * H * H, 9 * C

  def launch():
    check(lib().fpb_debug_igemm(3, ptr(x), ptr(w), ptr(out), M, C, K, H, H, C, H, H, 3, 3, 1, 1, ptr(sc), ptr(sh), None, 1, 1,
                                stream()))
  for _ in range(3):
    launch()
  ts = []
  for _ in range(n_rep):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); launch(); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
  ms = float(np.mean(ts))
  flops = 2.0 * M * C * K
  return dict(kernel='tc2_kernel<256,2> conv3x3 256->256 @40x40 x252 hyps (M=403200,N=256,K=2304)', ms=ms, flops=flops, tflops=flops / ms / 1e9)


def main():
  claim_stdout()
  ap = argparse.ArgumentParser()
  ap.add_argument('--gpus', type=int, default=1)
  ap.add_argument('--exchange', default='auto', choices=['auto', 'p2p', 'nccl'], help='sharded register(): how features/poses are exchanged')
  ap.add_argument('--steps', type=int, default=10)
  ap.add_argument('--warmup', type=int, default=3)
  ap.add_argument('--impl', default='ours')
  ap.add_argument('--ref-hyps', type=int, default=N_HYP)
  ap.add_argument('--cpu-hyps', type=int, default=24)
  ap.add_argument('--no-cpu-baseline', action='store_true')
  ap.add_argument('--precision', default='fp16', choices=['fp16', 'bf16'])
  args = ap.parse_args()
  if args.impl == 'reference':
    return run_reference(args)

  import torch
  import torch.distributed as dist
  world = int(os.environ.get('WORLD_SIZE', '1'))
  rank = int(os.environ.get('RANK', '0'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  torch.cuda.set_device(local)
  if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
  from foundationpose_b200 import synthetic
  from foundationpose_b200 import dist as fdist
  from foundationpose_b200._lib import lib
  scene = synthetic.make_scene(2)
  est, sd_r, sd_s = make_estimator(scene, args.precision)
  K, rgb, depth, mask = scene['K'], scene['rgb'], scene['depth'], scene['mask']
  n_hyp = len(est.rot_grid)
  assert n_hyp == N_HYP
  # device-resident and pinned-host copies of the frame
  d_rgb, d_depth, d_mask = torch.as_tensor(rgb).cuda(), torch.as_tensor(depth).cuda(), torch.as_tensor(mask).cuda()
  p_rgb, p_depth, p_mask = torch.as_tensor(rgb).pin_memory(), torch.as_tensor(depth).pin_memory(), torch.as_tensor(mask).pin_memory()
  h_rgb, h_depth, h_mask = p_rgb.numpy(), p_depth.numpy(), p_mask.numpy()

  def barrier():
    if world > 1:
      dist.barrier()
    torch.cuda.synchronize()

  def timed(fn, steps, warmup, count_launches=False):
    for _ in range(warmup):
      fn()
    barrier()
    if count_launches:
      lib().fpb_launch_count(1)                  # reset AFTER the warm-up: gpu_launches counts the timed region only
    per = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
      s0 = time.perf_counter()
      fn()
      per.append(time.perf_counter() - s0)       # register() is synchronous (returns the host pose)
    e1.record()
    barrier()
    wall = time.perf_counter() - t0
    dev_ms = e0.elapsed_time(e1)
    tot = torch.tensor([max(dev_ms, 0.0), wall * 1e3], device='cuda', dtype=torch.float64)
    if world > 1:
      dist.all_reduce(tot, op=dist.ReduceOp.MAX)
    return float(tot[0]), float(tot[1]), per

  # N = 1: the full register(); N > 1: the same object, hypotheses sharded over the ranks (north-star partition)
  if world == 1:
    def step_device():
      return est.register(K=K, rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)

    def step_host():
      return est.register(K=K, rgb=h_rgb, depth=h_depth, ob_mask=h_mask, iteration=ITERS)
  else:
    def step_device():
      return fdist.register_sharded(est, K, d_rgb, d_depth, d_mask, iteration=ITERS, exchange=args.exchange)

    def step_host():
      return fdist.register_sharded(est, K, h_rgb, h_depth, h_mask, iteration=ITERS, exchange=args.exchange)

  sampler = ClockSampler(local)
  if rank == 0:
    sampler.start()
  dev_ms, wall_ms, per = timed(step_device, args.steps, max(args.warmup, 3), count_launches=True)
  launches = lib().fpb_launch_count(0)
  e_dev_ms, e_wall_ms, e_per = timed(step_host, args.steps, max(args.warmup, 3))
  sampler.stop_flag = True
  total_hyp = n_hyp * args.steps                 # strong scaling: 252 hypotheses per step whatever the number of ranks
  value = total_hyp / (dev_ms / 1e3)
  e2e_value = total_hyp / (e_wall_ms / 1e3)

  # ---- N > 1: correctness of the sharded result on every rank + the replica (object-sharded, cfg 5) number ----
  multi = None
  if world > 1:
    pose_s = step_host()
    best_s, top_s = int(est.best_id), est.poses[0].clone()
    pose_u = est.register(K=K, rgb=h_rgb, depth=h_depth, ob_mask=h_mask, iteration=ITERS)
    ok = float(np.abs(pose_s - pose_u).max()) <= 1e-4 and best_s == int(est.best_id)
    flag = torch.tensor([1.0 if ok else 0.0, float(np.abs(pose_s - pose_u).max())], device='cuda', dtype=torch.float64)
    mn = flag.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    mx = flag.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    # cfg 5: 8 distinct meshes (seeds 50..57), one object per rank, only best pose + score all-gathered
    obj_mesh = synthetic.blob_mesh(seed=50 + rank, coef=(0.15 + 0.01 * rank, 3.0 + 0.25 * rank, 2.0 - 0.1 * rank, 0.1))
    oc = (obj_mesh.vertices.min(0) + obj_mesh.vertices.max(0)) / 2
    obj_est, _, _ = make_estimator(dict(scene, mesh=obj_mesh, diameter=float(2 * np.linalg.norm(obj_mesh.vertices - oc, axis=1).max())),
                                   args.precision)

    def step_replica():
      pose = obj_est.register(K=K, rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
      return fdist.gather_best(pose, obj_est.scores[0])
    r_dev, r_wall, r_per = timed(step_replica, max(3, args.steps // 2), 2)
    r_steps = max(3, args.steps // 2)
    exch = args.exchange
    if exch == 'auto':
      exch = 'p2p' if world <= fdist.MAX_PEERS else 'nccl'      # what register_sharded resolves 'auto' to on NCCL
    how = ('NCCL all_gather of features + poses' if exch == 'nccl' else
           'feature rows / poses stored into every rank\'s table by the producing kernels (peer-mapped symmetric memory over '
           'NVLink) + one device barrier')
    multi = dict(matches_unsharded=bool(mn[0] > 0.5), max_pose_diff_vs_unsharded=float(mx[1]), exchange=exch, exchange_how=how,
                 replicas=dict(value=n_hyp * world * r_steps / (r_dev / 1e3), unit='hypotheses/s', scaling='weak',
                               register_ms_p50=float(np.median(r_per) * 1e3),
                               note='BASELINE cfg 5: one distinct mesh (blob seeds 50..57) per rank, full 252-hypothesis register() '
                                    'each, all_gather of the best pose + score only'))

  # per-stage device time: CUDA events at the boundaries of the three stages of one register(), median of 5
  stages = None
  extras = {}
  if rank == 0 and world == 1:
    from foundationpose_b200 import utils as U_
    marks = []

    def wrap(obj, name, tag):
      f = getattr(obj, name)

      def g(*a, **k):
        e_in = torch.cuda.Event(enable_timing=True); e_in.record()
        r = f(*a, **k)
        e_out = torch.cuda.Event(enable_timing=True); e_out.record()
        marks.append((tag, e_in, e_out))
        return r
      setattr(obj, name, g)
      return f
    saved = [(U_, 'register_prologue', wrap(U_, 'register_prologue', 'prologue')),
             (est.refiner, 'predict', wrap(est.refiner, 'predict', 'refine')),
             (est.scorer, 'predict', wrap(est.scorer, 'predict', 'score'))]
    acc = {'prologue': [], 'refine': [], 'score': []}
    for _ in range(5):
      marks.clear()
      est.register(K=K, rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS)
      torch.cuda.synchronize()
      for tag, a, b in marks:
        acc[tag].append(a.elapsed_time(b))
    for obj, name, f in saved:
      setattr(obj, name, f)
    stages = {k + '_ms': float(np.median(v)) for k, v in acc.items()}
    stages['note'] = f'device time between CUDA events; refine = {ITERS} x (crop windows, render, warp, RefineNet, pose update); score = render, warp, ScoreNet, ranking'
    # cfg 3 (track_one video path: 1 hypothesis, 2 refine iterations, host frames) as a driver-visible number
    lat = []
    for _ in range(60):
      s0 = time.perf_counter()
      est.track_one_graphed(rgb=h_rgb, depth=h_depth, K=K, iteration=2)
      lat.append(time.perf_counter() - s0)
    extras['track_ms_p50'] = float(np.median(lat[10:]) * 1e3)
    extras['track_note'] = 'cfg 3: track_one_graphed (CUDA-graph replay), 1 hypothesis x 2 iterations, pinned host frame in, pose out'
    # the other operand type on the same kernels
    other = 'bf16' if args.precision == 'fp16' else 'fp16'
    est_o, _, _ = make_estimator(scene, other)
    o_dev, _, o_per = timed(lambda: est_o.register(K=K, rgb=d_rgb, depth=d_depth, ob_mask=d_mask, iteration=ITERS), 5, 2)
    extras['other_precision'] = dict(dtype=other, ms_per_step=o_dev / 5, register_ms_p50=float(np.median(o_per) * 1e3))
    del est_o
    # cfg 1 (12-face textured cube, 252 hypotheses, 1 iteration): the large-triangle regime of the rasteriser
    sc1 = synthetic.make_scene(1)
    est1, _, _ = make_estimator(sc1, args.precision)
    d1 = [torch.as_tensor(sc1[k]).cuda() for k in ('rgb', 'depth', 'mask')]
    c_dev, _, c_per = timed(lambda: est1.register(K=sc1['K'], rgb=d1[0], depth=d1[1], ob_mask=d1[2], iteration=1), 5, 2)
    extras['cfg1'] = dict(ms_per_step=c_dev / 5, register_ms_p50=float(np.median(c_per) * 1e3),
                          note='BASELINE cfg 1: textured cube (12 faces), 252 hypotheses, 1 refine iteration + scoring')
    del est1
  if rank != 0:
    if world > 1:
      dist.destroy_process_group()
    return
  pk = peaks()
  dom = time_dominant_kernel(torch)
  step_flops = n_hyp * (ITERS * FLOP_REFINE + FLOP_SCORE)
  traffic = traffic_from = None
  for name in ('ncu_conv6_r02g.json', 'ncu_conv6_r02.json', 'ncu_conv6_r01.json'):      # dram__bytes_read+write of this kernel, one ncu --set full capture
    prof = os.path.join(ROOT, 'profiles', name)
    if os.path.exists(prof):
      pj = json.load(open(prof))
      traffic, traffic_from = pj.get('traffic_bytes_per_launch'), f"{name} (ncu --set full capture of build {pj.get('build', 'r01j')}, not of this run)"
      break
  roof = dict(bound='tensor', achieved=dom['tflops'], peak=pk['tf_burst'], unit='TFLOP/s', frac=dom['tflops'] / pk['tf_burst'],
              traffic=traffic, traffic_source=traffic_from, kernel=dom['kernel'], ms_per_launch=dom['ms'], flops_per_launch=dom['flops'], peak_source=pk['src'],
              step_tflops=step_flops * args.steps / (dev_ms / 1e3) / 1e12 / world,
              step_frac_of_sustained=step_flops * args.steps / (dev_ms / 1e3) / 1e12 / world / pk['tf_sust'])
  cpu = None
  if not args.no_cpu_baseline and world == 1:
    cpu = cpu_sample(scene, sd_r, sd_s, args.cpu_hyps, ITERS)
    cpu.pop('seconds')
  h2d = int(rgb.nbytes + depth.nbytes + mask.nbytes)
  line = dict(metric=METRIC, value=value, unit='hypotheses/s', n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
              ms_per_step=dev_ms / args.steps, higher_is_better=True, scaling='weak' if world == 1 else 'strong', vs_baseline=None,
              dtype=args.precision, data='synthetic',
              config=dict(workload=WORKLOAD + ('' if world == 1 else f'; ONE object, its 252 hypotheses sharded over {world} GPUs'),
                          hypotheses=n_hyp, iterations=ITERS, precision=args.precision,
                          l2='no flush needed: ~4 GB of activations stream through per step (>> 126 MB L2)'),
              register_ms_p50=float(np.median(per) * 1e3), register_ms_p95=float(np.percentile(per, 95) * 1e3),
              e2e=dict(value=e2e_value, unit='hypotheses/s', h2d_bytes_per_step=h2d, d2h_bytes_per_step=64 + 32,
                       register_ms_p50=float(np.median(e_per) * 1e3)),
              gpu_launches=int(launches), roofline=roof, cpu_baseline=cpu, clocks=sampler.summary())
  if stages:
    line['stages'] = stages
  line.update(extras)
  if multi:
    line.update(multi)
  emit(line)
  if world > 1:
    dist.destroy_process_group()


if __name__ == '__main__':
  main()
