"""Timing of the fused attention kernel at the bench shape (252 hypotheses x 4 heads x 400 tokens) -- dev tool."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from foundationpose_b200._lib import lib, check, ptr, stream  # noqa: E402
G = 252
R = G * 400
qk = torch.randn(R, 1024, device='cuda').bfloat16()
vt = torch.randn(512, R, device='cuda').bfloat16()
bv = torch.zeros(512, device='cuda')
ctx = torch.empty(R, 512, device='cuda', dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
ts = []
for i in range(8):
  flush.zero_()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record(); check(lib().fpb_debug_attention(ptr(qk), ptr(vt), R, ptr(bv), ptr(ctx), G, stream())); e1.record()
  torch.cuda.synchronize()
  ts.append(e0.elapsed_time(e1))
ms = sorted(ts[2:])[len(ts[2:]) // 2]
fl = 2.0 * G * 4 * 400 * 400 * 128 * 2
print(f'attention G={G}: {ms*1e3:.1f} us  {fl/ms/1e9:.1f} TFLOP/s')
if int(os.environ.get('FPB_ATTN_DBG', '0')) & 32:
  torch.cuda.synchronize()
  st = ctx.view(torch.int64).flatten()[:192].cpu().numpy()
  t0 = min(int(v) for v in st if v > 0)
  for name, b in (('mma', 0), ('softmax set0 (warp 2)', 64), ('softmax set1 (warp 6)', 128)):
    print(name, [int(v - t0) for v in st[b:b + 64] if v > 0])
