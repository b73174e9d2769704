"""Timing of the row-tile encoder-layer kernels (csrc/net_ffn.cu) at the bench shape (dev tool).
  python tools/ffn_bench.py [--nhyp 252] [--reps 10] [--dtype fp16]
Also times the unfused chain it replaces (operand-swapped K<=512 linears: fpb_debug_igemm mode 4)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from foundationpose_b200._lib import lib, check, ptr, stream  # noqa: E402


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--nhyp', type=int, default=252)
  ap.add_argument('--reps', type=int, default=10)
  ap.add_argument('--dtype', default='fp16', choices=['fp16', 'bf16'])
  ap.add_argument('--no-flush', action='store_true')
  ap.add_argument('--only', default='', help='proj | ffn | lin')
  ap.add_argument('--stamps', action='store_true', help='run with FPB_FFN_DBG=2048 in the environment: print the phase stamps of CTA 0')
  a = ap.parse_args()
  dt = torch.float16 if a.dtype == 'fp16' else torch.bfloat16
  check(lib().fpb_debug_set_precision(2 if a.dtype == 'fp16' else 1))
  M = a.nhyp * 400
  rn = lambda *s: torch.randn(*s, device='cuda')
  A, res = rn(M, 512).to(dt), rn(M, 512).to(dt)
  W1, W2 = (rn(512, 512) / 512 ** 0.5).to(dt), (rn(512, 512) / 512 ** 0.5).to(dt)
  b1, b2, gamma, beta = rn(512), rn(512), rn(512), rn(512)
  out = torch.empty(M, 512, device='cuda', dtype=dt)
  pooled = torch.empty(a.nhyp, 512, device='cuda')
  scratch = torch.empty(M // 16, 512, device='cuda')
  flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
  ones = torch.ones(512, device='cuda')

  def proj_ln():
    check(lib().fpb_debug_ffn(0, ptr(A), None, None, ptr(W2), ptr(b2), ptr(res), ptr(gamma), ptr(beta), ptr(out), None, M, 400, stream()))

  def ffn():
    check(lib().fpb_debug_ffn(1, ptr(A), ptr(W1), ptr(b1), ptr(W2), ptr(b2), None, ptr(gamma), ptr(beta), ptr(pooled), ptr(scratch), M, 400, stream()))

  def lin_old():
    check(lib().fpb_debug_igemm(4, ptr(A), ptr(W2), ptr(out), M, 512, 512, 1, 1, 512, 1, 1, 1, 1, 1, 0, ptr(ones), ptr(b2), ptr(res), 0, 1, stream()))

  vt = torch.empty(512, M, device='cuda', dtype=dt)

  def vt_proj():     # V^T = Wv . X^T + bv, stored channel-major (the attention kernel's K-major B operand)
    check(lib().fpb_debug_igemm_ex(4, ptr(A), ptr(W2), ptr(vt), M, 512, 512, 1, 1, 512, 1, 1, 1, 1, 1, 0, 0, 0, 0, None, ptr(b2), None, 0, 2,
                                   stream()))

  for name, fn in (('proj+LN (ffn_kernel<0>)', proj_ln), ('FFN+LN+sum (ffn_kernel<1> + pool)', ffn), ('lin: one K=512 linear, old kernel', lin_old),
                   ('vt: V^T projection (transposed output)', vt_proj)):
    if a.only and not name.lower().startswith(a.only):
      continue
    for _ in range(2):
      fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
      if not a.no_flush:
        flush.zero_()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record(); fn(); e1.record()
      torch.cuda.synchronize()
      ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    if a.stamps and 'lin' not in name:
      import ctypes
      import numpy as np
      buf = np.zeros(512, dtype=np.int64)
      check(lib().fpb_debug_ffn_stamps(buf.ctypes.data_as(ctypes.c_void_p), 512))
      per = 14 if 'FFN' in name else 10
      st = buf[:per * 6].reshape(-1, per)
      print('stamps (cycles, deltas within a tile; last column = gap to the next tile):')
      for i in range(len(st)):
        d = np.diff(st[i]).tolist()
        nxt = int(st[i + 1][0] - st[i][-1]) if i + 1 < len(st) and st[i + 1][0] > 0 else -1
        print('  tile', i, d, nxt)
    gf = 2.0 * M * 512 * 512 * (2 if 'FFN' in name else 1)
    print(f'{name}: median {ts[len(ts)//2]:.1f} us, min {ts[0]:.1f} us, {gf / ts[len(ts)//2] / 1e6:.0f} TFLOP/s (M={M})')


if __name__ == '__main__':
  main()
