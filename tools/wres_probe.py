import os, sys, subprocess
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
CASES = [(700, 512, 512, 0), (700, 512, 512, 1), (2, 512, 512, 0), (300, 512, 1536, 0), (1000, 256, 256, 0)]
if len(sys.argv) > 1:
  import torch
  from foundationpose_b200._lib import lib, check, ptr, stream
  M, K, N, res = CASES[int(sys.argv[1])]
  x = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') / K ** 0.5).bfloat16()
  r = torch.randn(M, N, device='cuda').bfloat16() if res else None
  out = torch.zeros(M, N, device='cuda', dtype=torch.bfloat16)
  check(lib().fpb_debug_igemm(4, ptr(x), ptr(w), ptr(out), M, N, K, 1, 1, K, 1, 1, 1, 1, 1, 0, None, None, ptr(r), 0, 1, stream()))
  torch.cuda.synchronize()
  ref = x.float() @ w.float().t() + (r.float() if res else 0)
  print('case', sys.argv[1], CASES[int(sys.argv[1])], 'max err', (out.float() - ref).abs().max().item())
else:
  for i in range(len(CASES)):
    p = subprocess.run([sys.executable, __file__, str(i)], capture_output=True, text=True, timeout=120)
    print((p.stdout.strip().splitlines() or ['-'])[-1], '| rc', p.returncode, '|', (p.stderr.strip().splitlines() or [''])[-1][:160])
