"""Per-layer timing of the implicit-GEMM kernels at the bench workload's real shapes (dev tool).
  python tools/gemm_bench.py [--modes 2,3] [--layers conv6,conv11] [--reps 10] [--nhyp 252]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from foundationpose_b200._lib import lib, check, ptr, stream  # noqa: E402

# name: (imgs_per_hyp, Hin, Cin, Cout, k, stride, residual, relu)
LAYERS = {
  'conv1': (2, 80, 64, 128, 3, 2, 0, 1), 'conv2': (2, 40, 128, 128, 3, 1, 0, 1), 'conv3': (2, 40, 128, 128, 3, 1, 1, 1),
  'conv6': (1, 40, 256, 256, 3, 1, 0, 1), 'conv7': (1, 40, 256, 256, 3, 1, 1, 1), 'conv10': (1, 40, 256, 512, 3, 2, 0, 1),
  'conv11': (1, 20, 512, 512, 3, 1, 0, 1), 'conv12': (1, 20, 512, 512, 3, 1, 1, 1),
  'inproj': (400, 1, 512, 1536, 1, 1, 0, 0), 'lin': (400, 1, 512, 512, 1, 1, 1, 0),
}


DT = torch.float16


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument('--modes', default='2,3')
  ap.add_argument('--layers', default=','.join(LAYERS))
  ap.add_argument('--reps', type=int, default=10)
  ap.add_argument('--nhyp', type=int, default=252)
  ap.add_argument('--no-flush', action='store_true')
  ap.add_argument('--dtype', default='fp16', choices=['fp16', 'bf16'], help='operand type of the tcgen05 kernels')
  ap.add_argument('--custom', default='', help='extra layer: name:ipn,H,Cin,Cout,k,stride,res,relu (repeat with ;)')
  a = ap.parse_args()
  global DT
  DT = torch.float16 if a.dtype == 'fp16' else torch.bfloat16
  check(lib().fpb_debug_set_precision(2 if a.dtype == 'fp16' else 1))
  for spec in filter(None, a.custom.split(';')):
    nm, v = spec.split(':')
    LAYERS[nm] = tuple(int(t) for t in v.split(','))
  flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
  for name in a.layers.split(','):
    if name == 'stem':
      stem(a)
      continue
    ipn, H, Cin, Cout, k, stride, res, relu = LAYERS[name]
    nimg = ipn * a.nhyp
    pad = (k - 1) // 2
    Ho = (H + 2 * pad - k) // stride + 1
    M, K = nimg * Ho * Ho, k * k * Cin
    x = torch.randn(nimg, H, H, Cin, device='cuda').to(DT)
    w = (torch.randn(Cout, K, device='cuda') / K ** 0.5).to(DT)
    out = torch.empty(M, Cout, device='cuda', dtype=DT)
    r = torch.randn(M, Cout, device='cuda').to(DT) if res else None
    sc = torch.ones(Cout, device='cuda'); sh = torch.zeros(Cout, device='cuda')
    for mode in [int(m) for m in a.modes.split(',')]:
      if mode == 4 and Cout > 128 and not (k == 1 and K <= 512 and Cout % 128 == 0):
        continue
      def launch():
        check(lib().fpb_debug_igemm(mode, ptr(x), ptr(w), ptr(out), M, Cout, K, H, H, Cin, Ho, Ho, k, k, stride, pad, ptr(sc),
                                    ptr(sh), ptr(r), relu, 1, stream()))
      for _ in range(2):
        launch()
      ts = []
      for _ in range(a.reps):
        if not a.no_flush:
          flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); launch(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
      ms = sorted(ts)[len(ts) // 2]
      fl = 2.0 * M * Cout * K
      print(f'{name:7s} mode {mode} M={M} N={Cout} K={K}: {ms*1e3:8.1f} us  {fl/ms/1e9:7.1f} TFLOP/s', flush=True)
      if int(os.environ.get('FPB_TC2_DBG', '0')) & 2048:
        out.zero_(); launch(); torch.cuda.synchronize()
        st = out.view(torch.int64).flatten()[:192].cpu().numpy()
        t0 = min(int(v) for v in st if v > 0)
        for nm, b in (('mma', 64), ('epi warp 2', 128)):
          print('  ', nm, [int(v - t0) for v in st[b:b + 64] if v > 0][:40])


def stem(a):
  """7x7 stem as a 7x1 conv over wide pixels of the row-padded input (tests/test_nets_gpu.py::test_first_conv_7x7_padded_layout)"""
  flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
  n = 2 * a.nhyp
  x = torch.randn(n, 160, 168, 8, device='cuda').to(DT)
  w = (torch.randn(64, 448, device='cuda') / 20).to(DT)
  out = torch.empty(n * 6400, 64, device='cuda', dtype=DT)
  sc = torch.ones(64, device='cuda'); sh = torch.zeros(64, device='cuda')
  for mode in [int(m) for m in a.modes.split(',')]:
    def launch():   # mode 5 (dedicated stem kernel) includes the debug entry's weight repack + cudaMalloc: use tools/stem_bench.py
      check(lib().fpb_debug_igemm_ex(mode, ptr(x), ptr(w), ptr(out), n * 6400, 64, 448, 160, 161, 64, 80, 80, 7, 1, 2, 3, -1, 168, 8,
                                     ptr(sc), ptr(sh), None, 1, 1, stream()))
    for _ in range(2):
      launch()
    ts = []
    for _ in range(a.reps):
      flush.zero_()
      e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
      e0.record(); launch(); e1.record()
      torch.cuda.synchronize()
      ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(f'stem    mode {mode} M={n*6400} N=64 K=448: {ms*1e3:8.1f} us  {2.0*n*6400*64*448/ms/1e9:7.1f} TFLOP/s', flush=True)


if __name__ == '__main__':
  main()
