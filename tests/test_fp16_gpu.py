"""GPU: the fp16-operand build of the tcgen05 kernels (csrc/act_type.cuh, namespace f16; FPB_PREC_FP16) -- the dtype
the reference's autocast path computes in (predict_pose_refine.py:190, predict_score.py:193).  Same kernels as the
bf16 build compiled with another element type, so each family is checked once on its representative shapes against
torch on the SAME fp16 operands (fp32 accumulation: tolerance = accumulation order + output rounding), then the whole
networks against the golden vectors of the reference's own modules."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')


@pytest.fixture()
def fp16_debug_entries():
  from foundationpose_b200._lib import lib, check
  check(lib().fpb_debug_set_precision(2))
  yield
  check(lib().fpb_debug_set_precision(1))


def _ref_conv(x, w, stride, pad, scale, shift, res, relu):
  y = F.conv2d(x.half().float().permute(0, 3, 1, 2).double(), w.half().float().double(), None, stride, pad).permute(0, 2, 3, 1)
  y = y * scale.double() + shift.double() + res.half().float().reshape(y.shape).double()
  return torch.relu(y).float() if relu else y.float()


# mode, (nimg, H, Cin, Cout, k, stride, pad)
@pytest.mark.parametrize('mode,case', [
  (1, (3, 20, 512, 512, 3, 1, 1)), (2, (2, 40, 256, 256, 3, 1, 1)), (3, (7, 40, 256, 256, 3, 1, 1)), (3, (2, 40, 256, 512, 3, 2, 1)),
  (3, (700, 1, 512, 512, 1, 1, 0)), (4, (5, 40, 128, 128, 3, 1, 1)), (4, (3, 80, 64, 128, 3, 2, 1)), (4, (333, 1, 256, 128, 1, 1, 0))])
def test_igemm_fp16_matches_torch(fp16_debug_entries, mode, case):
  from foundationpose_b200._lib import lib, check, ptr, stream
  nimg, H, Cin, Cout, k, stride, pad = case
  g = torch.Generator(device='cuda').manual_seed(sum(case))
  x = torch.randn(nimg, H, H, Cin, device='cuda', generator=g)
  w = torch.randn(Cout, Cin, k, k, device='cuda', generator=g) / (Cin * k * k) ** 0.5
  scale = torch.rand(Cout, device='cuda', generator=g) + 0.5
  shift = torch.randn(Cout, device='cuda', generator=g) * 0.1
  Ho = (H + 2 * pad - k) // stride + 1
  M, K = nimg * Ho * Ho, k * k * Cin
  res = torch.randn(M, Cout, device='cuda', generator=g)
  A = x.half().contiguous()
  W = w.permute(0, 2, 3, 1).reshape(Cout, -1).half().contiguous()
  C = torch.empty((M, Cout), dtype=torch.float16, device='cuda')
  check(lib().fpb_debug_igemm(mode, ptr(A), ptr(W), ptr(C), M, Cout, K, H, H, Cin, Ho, Ho, k, k, stride, pad, ptr(scale), ptr(shift),
                              ptr(res.half().contiguous()), 1, 1, stream()))
  torch.cuda.synchronize()
  ref = _ref_conv(x, w, stride, pad, scale, shift, res, 1).reshape(M, Cout)
  err = (C.float() - ref).abs().max().item()
  assert err < 6e-3, err                    # fp16 output rounding of O(4) values: 2^-10 relative (bf16 build: 3e-2)


def test_stem_and_attention_fp16(fp16_debug_entries):
  from foundationpose_b200._lib import lib, check, ptr, stream
  g = torch.Generator(device='cuda').manual_seed(5)
  x6 = torch.randn(3, 160, 160, 6, device='cuda', generator=g)
  w6 = torch.randn(64, 6, 7, 7, device='cuda', generator=g) / (6 * 49) ** 0.5
  x8 = torch.zeros(3, 160, 168, 8, device='cuda'); x8[:, :, 4:164, :6] = x6
  w8 = torch.zeros(64, 8, 7, 8, device='cuda'); w8[:, :6, :, :7] = w6
  A = x8.half().contiguous()
  W = w8.permute(0, 2, 3, 1).reshape(64, -1).half().contiguous()
  C = torch.empty((3 * 6400, 64), dtype=torch.float16, device='cuda')
  check(lib().fpb_debug_igemm_ex(5, ptr(A), ptr(W), ptr(C), 3 * 6400, 64, 448, 160, 161, 64, 80, 80, 7, 1, 2, 3, -1, 168, 8,
                                 None, None, None, 0, 1, stream()))
  torch.cuda.synchronize()
  ref = F.conv2d(x6.half().float().permute(0, 3, 1, 2), w6.half().float(), None, 2, 3).permute(0, 2, 3, 1)
  assert (C.float().reshape(3, 80, 80, 64) - ref).abs().max().item() < 4e-3
  groups, R = 2, 800
  qk = (torch.randn(R, 1024, device='cuda', generator=g) * 1.5).half()
  v = torch.randn(R, 512, device='cuda', generator=g).half()
  bv = torch.randn(512, device='cuda', generator=g) * 0.1
  vt = v.t().contiguous()
  ctx = torch.empty(R, 512, device='cuda', dtype=torch.float16)
  check(lib().fpb_debug_attention(ptr(qk), ptr(vt), R, ptr(bv), ptr(ctx), groups, stream()))
  torch.cuda.synchronize()
  q = qk[:, :512].float().reshape(groups, 400, 4, 128).permute(0, 2, 1, 3)
  k = qk[:, 512:].float().reshape(groups, 400, 4, 128).permute(0, 2, 1, 3)
  vv = v.float().reshape(groups, 400, 4, 128).permute(0, 2, 1, 3)
  p = torch.softmax(q @ k.transpose(-1, -2) / 128 ** 0.5, dim=-1)
  ref = (p @ vv).permute(0, 2, 1, 3).reshape(R, 512) + bv
  assert (ctx.float() - ref).abs().max().item() < 4e-3          # bf16 build: 2e-2


def _golden_AB(seed, n, dtype):
  from foundationpose_b200 import utils as U
  from foundationpose_b200.nets import _calibration_batch
  x = _calibration_batch(torch.Generator().manual_seed(seed), n)
  AB = U.net_input_buffer(n, 160, dtype)
  AB[:, :, U.IN_XOFF:U.IN_XOFF + 160, :6] = x.permute(0, 2, 3, 1).to(dtype).cuda()
  return AB


def test_precision_table_against_reference_golden():
  """bf16 vs fp16 vs fp32 deviation from the reference modules' golden outputs (the f2 precision table): fp16 must be
  well inside bf16's error and inside the bound the bf16 test uses."""
  from foundationpose_b200 import nets
  gr, gs = np.load(os.path.join(GOLD, 'refine_net.npz')), np.load(os.path.join(GOLD, 'score_net.npz'))
  rows = {}
  for prec in ('fp32', 'fp16', 'bf16'):
    dt = nets.ACT_DTYPE[prec]
    net = nets.Net(nets.REFINE, nets.synthetic_state_dict(nets.REFINE), prec)
    t, r = net.refine(_golden_AB(int(gr['seed']), int(gr['n']), dt), int(gr['n']))
    net = nets.Net(nets.SCORE, nets.synthetic_state_dict(nets.SCORE), prec)
    f = net.score_features(_golden_AB(int(gs['seed']), int(gs['n']), dt), int(gs['n']))
    rows[prec] = (np.abs(t.cpu().numpy() - gr['trans']).max(), np.abs(r.cpu().numpy() - gr['rot']).max(),
                  np.abs(f.cpu().numpy() - gs['feats']).max())
  print('\nprecision | max |trans - golden| | max |rot - golden| | max |feat - golden|  (golden scale: trans %.3f rot %.3f feat %.2f)'
        % (np.abs(gr['trans']).max(), np.abs(gr['rot']).max(), np.abs(gs['feats']).max()))
  for k, v in rows.items():
    print('  %s    |  %.2e  |  %.2e  |  %.2e' % ((k,) + v))
  assert rows['fp32'][0] < 2e-4 and rows['fp32'][1] < 2e-4
  assert rows['fp16'][0] < 1e-3 and rows['fp16'][1] < 3e-3 and rows['fp16'][2] < 2e-2
  assert rows['fp16'][1] < 0.5 * rows['bf16'][1] and rows['fp16'][2] < 0.5 * rows['bf16'][2]
