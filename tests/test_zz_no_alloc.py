"""GPU: fpb_score_cross_static -- the allocation-free variant of fpb_score_cross (static __device__ scratch, no counterpart
in the reference; added for callers that cannot call cudaMalloc).  Collected last
(file name) so that a failure here cannot hide other tests under `pytest -x`.  Passes alone in a fresh process and after
the other tests: it creates its own network and touches no shared state except the library's static scratch array."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_score_cross_static_matches_score_cross_on_consecutive_calls():
  from foundationpose_b200 import nets
  from foundationpose_b200._lib import lib, check, ptr, stream
  net = nets.Net(nets.SCORE, nets.synthetic_state_dict(nets.SCORE), 'fp32')
  g = torch.Generator(device='cuda').manual_seed(77)
  # consecutive calls share the static array (it keeps the previous call's contents): different L back to back, then
  # the first problem again
  feats = {L: torch.randn(L, 512, device='cuda', generator=g) * 1.5 for L in (252, 7, 256, 1)}
  want = {L: net.score_cross(f, L, 1).reshape(-1).clone() for L, f in feats.items()}      # existing entry point, caller workspace
  torch.cuda.synchronize()
  for L in (252, 7, 252, 256, 1, 252):
    out = torch.full((L,), float('nan'), device='cuda')
    check(lib().fpb_score_cross_static(net.handle, ptr(feats[L]), L, ptr(out), stream()))
    torch.cuda.synchronize()
    diff = float((out - want[L]).abs().max())
    assert torch.isfinite(out).all() and diff <= 1e-6, (L, diff)          # same kernels on the same inputs (fp32 tests: 5e-4)
  assert float(np.ptp(want[252].cpu().numpy())) > 1e-3                   # not a degenerate comparison
  # beyond the static scratch: rejected with a clear error, nothing launched
  big = torch.randn(257, 512, device='cuda', generator=g)
  out = torch.zeros(257, device='cuda')
  rc = lib().fpb_score_cross_static(net.handle, ptr(big), 257, ptr(out), stream())
  assert rc != 0 and b'at most 256' in lib().fpb_last_error()
  with pytest.raises(RuntimeError):
    check(rc)
  torch.cuda.synchronize()
  assert float(out.abs().max()) == 0.0


def test_score_cross_static_allocates_nothing_and_captures_into_a_graph():
  """The entry point makes no allocation: torch's allocator statistics and the driver's free-memory figure do not move
  across calls, and the call can be captured into a CUDA graph (cudaMalloc during capture would fail) and replayed."""
  from foundationpose_b200 import nets
  from foundationpose_b200._lib import lib, check, ptr, stream
  net = nets.Net(nets.SCORE, nets.synthetic_state_dict(nets.SCORE), 'fp32')
  L = 64
  feat = torch.randn(L, 512, device='cuda', generator=torch.Generator(device='cuda').manual_seed(5))
  want = net.score_cross(feat, L, 1).reshape(-1).clone()
  out = torch.zeros(L, device='cuda')
  check(lib().fpb_score_cross_static(net.handle, ptr(feat), L, ptr(out), stream()))      # first call on this device: address set-up
  torch.cuda.synchronize()
  free0, _ = torch.cuda.mem_get_info()
  stats0 = torch.cuda.memory_stats()['allocation.all.allocated']
  for _ in range(3):
    check(lib().fpb_score_cross_static(net.handle, ptr(feat), L, ptr(out), stream()))
  torch.cuda.synchronize()
  assert torch.cuda.memory_stats()['allocation.all.allocated'] == stats0
  assert torch.cuda.mem_get_info()[0] == free0
  assert float((out - want).abs().max()) <= 1e-6
  side = torch.cuda.Stream()
  side.wait_stream(torch.cuda.current_stream())
  graph = torch.cuda.CUDAGraph()
  out.zero_()
  with torch.cuda.stream(side):
    torch.cuda.synchronize()
    with torch.cuda.graph(graph, stream=side):
      check(lib().fpb_score_cross_static(net.handle, ptr(feat), L, ptr(out), stream()))
  torch.cuda.current_stream().wait_stream(side)
  for _ in range(2):
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert float((out - want).abs().max()) <= 1e-6
