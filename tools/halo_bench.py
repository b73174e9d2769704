"""Halo-layout 3x3 conv vs im2col conv timing at the bench shapes (dev tool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from foundationpose_b200._lib import lib, check, ptr, stream
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def timeit(fn, reps=6):
  for _ in range(2): fn()
  ts = []
  for _ in range(reps):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); fn(); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
  return sorted(ts)[len(ts) // 2]
for (n, H, Cin, Cout) in [(504, 40, 128, 128), (252, 40, 256, 256), (252, 20, 512, 512)]:
  npos = n * (H + 1) * (H + 1)
  xp = torch.randn(npos, Cin, device='cuda').bfloat16()
  xd = torch.randn(n * H * H, Cin, device='cuda').bfloat16()
  w = (torch.randn(Cout, 9 * Cin, device='cuda') / (9 * Cin) ** 0.5).bfloat16()
  out = torch.empty(npos, Cout, device='cuda', dtype=torch.bfloat16)
  sc = torch.ones(Cout, device='cuda'); sh = torch.zeros(Cout, device='cuda')
  fl = 2.0 * n * H * H * Cout * 9 * Cin
  for cg in (1, 2, 3):
    if cg == 3 and Cout > 128:
      continue
    ms = timeit(lambda: check(lib().fpb_debug_conv3x3_halo(cg, ptr(xp), ptr(w), ptr(out), n, H, H, Cin, Cout, 1, 0, ptr(sc), ptr(sh), None, 1, stream())))
    print(f'halo   n={n} H={H} C={Cin}->{Cout} cg={cg}: {ms*1e3:7.1f} us  {fl/ms/1e9:7.1f} TFLOP/s (useful)')
    ms = timeit(lambda: check(lib().fpb_debug_igemm(1 + cg if cg < 3 else 4, ptr(xd), ptr(w), ptr(out), n * H * H, Cout, 9 * Cin, H, H, Cin, H, H, 3, 3, 1, 1, ptr(sc), ptr(sh), None, 1, 1, stream())))
    print(f'im2col n={n} H={H} C={Cin}->{Cout} cg={cg}: {ms*1e3:7.1f} us  {fl/ms/1e9:7.1f} TFLOP/s')
