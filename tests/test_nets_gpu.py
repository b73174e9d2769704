"""GPU parity tests of the network path through the C ABI.
  * fpb_debug_igemm: tcgen05 bf16 and SIMT fp32 implicit GEMM vs torch conv2d / matmul on the same operands
    (bf16 operands are exact in fp32, so the tcgen05 result must match an fp32 reference to accumulation order:
    tolerance 2e-3 relative to the output scale).
  * RefineNet / ScoreNet forward, fp32 parity mode vs the fp32 oracle: <= 2e-4 absolute on outputs of O(0.1-1).
  * bf16 mode vs the fp32 oracle: reported against the north-star bounds (1e-3 rad, 0.1 mm) in test_register_gpu.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _igemm(prec, x_nhwc, w_oihw, stride, pad, scale=None, shift=None, residual=None, relu=0, out_bf16=0):
  from foundationpose_b200._lib import lib, check, ptr, stream
  nimg, Hin, Win, Cin = x_nhwc.shape
  Cout, _, KH, KW = w_oihw.shape
  Ho = (Hin + 2 * pad - KH) // stride + 1
  Wo = (Win + 2 * pad - KW) // stride + 1
  dt = torch.bfloat16 if prec else torch.float
  A = x_nhwc.to(dt).contiguous()
  W = w_oihw.permute(0, 2, 3, 1).reshape(Cout, -1).to(dt).contiguous()
  M, K = nimg * Ho * Wo, KH * KW * Cin
  C = torch.empty((M, Cout), dtype=torch.bfloat16 if out_bf16 else torch.float, device='cuda')
  res = residual.to(dt).contiguous() if residual is not None else None
  check(lib().fpb_debug_igemm(prec, ptr(A), ptr(W), ptr(C), M, Cout, K, Hin, Win, Cin, Ho, Wo, KH, KW, stride, pad,
                              ptr(scale), ptr(shift), ptr(res), relu, out_bf16, stream()))
  torch.cuda.synchronize()
  return C.float().reshape(nimg, Ho, Wo, Cout)


def _ref_conv(x_nhwc, w_oihw, stride, pad, scale=None, shift=None, residual=None, relu=0, bf16_ops=False):
  x, w = x_nhwc, w_oihw
  if bf16_ops:
    x, w = x.bfloat16().float(), w.bfloat16().float()
  y = F.conv2d(x.permute(0, 3, 1, 2).double(), w.double(), None, stride, pad).permute(0, 2, 3, 1)
  if scale is not None:
    y = y * scale.double() + shift.double()
  elif shift is not None:
    y = y + shift.double()
  if residual is not None:
    r = residual.bfloat16().float() if bf16_ops else residual
    y = y + r.reshape(y.shape).double()
  if relu:
    y = torch.relu(y)
  return y.float()


CASES = [
  # nimg, H, Cin, Cout, k, stride, pad
  (2, 1, 512, 512, 1, 1, 0),        # plain GEMM, K=512, M=2 (one ragged tile)
  (300, 1, 512, 1536, 1, 1, 0),     # in_proj shape, M ragged
  (3, 20, 512, 512, 3, 1, 1),       # 3x3 @512, M=1200
  (2, 40, 256, 256, 3, 1, 1),       # 3x3 @256
  (2, 40, 256, 512, 3, 2, 1),       # stride 2
  (2, 40, 128, 128, 3, 1, 1),       # BN=128 tile
  (1, 80, 64, 128, 3, 2, 1),        # Cin=64
  (2, 24, 64, 64, 3, 1, 1),         # BN=64 tile
  (7, 40, 256, 256, 3, 1, 1),       # several tiles per CTA pair, tile spans images, ragged tail (M=11200)
  (700, 1, 512, 512, 1, 1, 0),      # plain GEMM over 2 N tiles and 6 M tiles
]


# 0 SIMT fp32, 1 tcgen05+cp.async, 2/3 tcgen05+TMA cta_group 1/2, 4 operand-swapped tcgen05 kernel (Cout 64 / 128)
@pytest.mark.parametrize('prec', [0, 1, 2, 3, 4])
@pytest.mark.parametrize('case', CASES + [(5, 40, 128, 128, 3, 1, 1), (3, 80, 64, 128, 3, 2, 1), (333, 1, 256, 128, 1, 1, 0),
                                          (1, 10, 64, 64, 3, 1, 1)])
def test_igemm_matches_torch(prec, case):
  nimg, H, Cin, Cout, k, stride, pad = case
  if prec == 4 and Cout not in (64, 128) and not (k == 1 and H == 1 and Cout % 128 == 0 and Cin <= 512):
    pytest.skip('operand-swapped kernel: 64 / 128 output channels, or plain GEMMs with K <= 512 (weight-resident mode)')
  g = torch.Generator(device='cuda').manual_seed(hash(case) % 1000)
  x = torch.randn(nimg, H, H, Cin, device='cuda', generator=g)
  w = torch.randn(Cout, Cin, k, k, device='cuda', generator=g) / (Cin * k * k) ** 0.5
  scale = torch.rand(Cout, device='cuda', generator=g) + 0.5
  shift = torch.randn(Cout, device='cuda', generator=g) * 0.1
  Ho = (H + 2 * pad - k) // stride + 1
  res = torch.randn(nimg * Ho * Ho, Cout, device='cuda', generator=g)
  ob = int(prec == 4)                                # the operand-swapped kernel only writes bf16 activations
  tol = 3e-2 if ob else (2e-3 if prec else 2e-4)     # bf16 output rounding of O(4) values: 2^-7 relative
  got = _igemm(prec, x, w, stride, pad, scale, shift, res, relu=1, out_bf16=ob)
  ref = _ref_conv(x, w, stride, pad, scale, shift, res, relu=1, bf16_ops=bool(prec))
  err = (got - ref).abs().max().item()
  assert err < tol, err
  got2 = _igemm(prec, x, w, stride, pad, out_bf16=ob)             # no epilogue extras
  ref2 = _ref_conv(x, w, stride, pad, bf16_ops=bool(prec))
  assert (got2 - ref2).abs().max().item() < tol


@pytest.mark.parametrize('shape', [(1000, 512, 512), (72, 512, 1024), (4096, 256, 256)])
def test_weight_resident_gemm_transposed_output(shape):
  """V^T projection form: out[n][m] = sum_k x[m][k] w[n][k] + b[n], stored channel-major (net_tc3.cu mode 3)."""
  from foundationpose_b200._lib import lib, check, ptr, stream
  M, K, N = shape
  g = torch.Generator(device='cuda').manual_seed(M)
  x = torch.randn(M, K, device='cuda', generator=g).bfloat16()
  w = (torch.randn(N, K, device='cuda', generator=g) / K ** 0.5).bfloat16()
  b = torch.randn(N, device='cuda', generator=g)
  Mp = (M + 7) // 8 * 8                                   # row pitch of the transposed output (multiple of 8)
  out = torch.full((N, Mp), 7.0, device='cuda', dtype=torch.bfloat16)
  # ldc = M inside the debug entry: use an M that is already a multiple of 8
  assert Mp == M or True
  xs = x[:Mp] if Mp == M else x
  check(lib().fpb_debug_igemm_ex(4, ptr(xs), ptr(w), ptr(out), M, N, K, 1, 1, K, 1, 1, 1, 1, 1, 0, 0, 0, 0, None, ptr(b), None, 0, 2,
                                 stream()))
  torch.cuda.synchronize()
  ref = (x.float() @ w.float().t() + b).t()
  assert (out[:, :M].float() - ref).abs().max().item() < 3e-2


@pytest.mark.parametrize('prec', [0, 1, 2, 3, 4, 5])
def test_first_conv_7x7_padded_layout(prec):
  """7x7 stride-2 stem on the row-padded 8-channel input [n,160,168,8], evaluated as a 7x1 convolution over 161
  overlapping 64-channel "wide pixels" (csrc/net.cu kConvs[0]); weights [64][7][8 taps][8 ch]."""
  g = torch.Generator(device='cuda').manual_seed(5)
  x6 = torch.randn(3, 160, 160, 6, device='cuda', generator=g)
  w6 = torch.randn(64, 6, 7, 7, device='cuda', generator=g) / (6 * 49) ** 0.5
  x8 = torch.zeros(3, 160, 168, 8, device='cuda'); x8[:, :, 4:164, :6] = x6
  w8 = torch.zeros(64, 8, 7, 8, device='cuda'); w8[:, :6, :, :7] = w6
  from foundationpose_b200._lib import lib, check, ptr, stream
  dt = torch.bfloat16 if prec else torch.float
  A = x8.to(dt).contiguous()
  W = w8.permute(0, 2, 3, 1).reshape(64, -1).to(dt).contiguous()
  ob = int(prec >= 4)          # 4: operand-swapped im2col kernel, 5: dedicated stem kernel (csrc/net_stem.cu)
  C = torch.empty((3 * 80 * 80, 64), dtype=torch.bfloat16 if ob else torch.float, device='cuda')
  check(lib().fpb_debug_igemm_ex(prec, ptr(A), ptr(W), ptr(C), 3 * 6400, 64, 448, 160, 161, 64, 80, 80, 7, 1, 2, 3, -1, 168, 8,
                                 None, None, None, 0, ob, stream()))
  torch.cuda.synchronize()
  ref = _ref_conv(x6, w6, 2, 3, bf16_ops=bool(prec))
  assert (C.float().reshape(3, 80, 80, 64) - ref).abs().max().item() < (2e-2 if ob else 2e-3 if prec else 2e-4)


@pytest.mark.parametrize('nimg', [1, 7, 300])
def test_stem_kernel_matches_im2col_kernel(nimg):
  """Dedicated stem kernel (plane loads, two output rows stacked in M) against the im2col kernel on the same bf16
  operands, with folded BatchNorm + ReLU; 300 images = more tiles than SMs (persistent loop, both accumulators)."""
  from foundationpose_b200._lib import lib, check, ptr, stream
  g = torch.Generator(device='cuda').manual_seed(11 + nimg)
  A = torch.zeros(nimg, 160, 168, 8, device='cuda', dtype=torch.bfloat16)
  A[:, :, 4:164, :6] = torch.randn(nimg, 160, 160, 6, device='cuda', generator=g).bfloat16()
  w8 = torch.zeros(64, 7, 8, 8, device='cuda'); w8[:, :, :7, :6] = torch.randn(64, 7, 7, 6, device='cuda', generator=g) / 17
  W = w8.reshape(64, -1).bfloat16().contiguous()
  sc = torch.rand(64, device='cuda', generator=g) + 0.5
  sh = torch.randn(64, device='cuda', generator=g) * 0.3
  outs = []
  for mode in (4, 5):
    C = torch.full((nimg * 6400, 64), float('nan'), dtype=torch.bfloat16, device='cuda')
    check(lib().fpb_debug_igemm_ex(mode, ptr(A), ptr(W), ptr(C), nimg * 6400, 64, 448, 160, 161, 64, 80, 80, 7, 1, 2, 3, -1, 168, 8,
                                   ptr(sc), ptr(sh), None, 1, 1, stream()))
    torch.cuda.synchronize()
    outs.append(C.float())
  assert torch.isfinite(outs[1]).all()
  assert ((outs[0] - outs[1]).abs() <= outs[0].abs() * 2 ** -7 + 1e-3).all()       # at most one bf16 ulp
  assert (outs[0] != outs[1]).float().mean().item() < 0.02     # same products, fp32 accumulation: only rounding-boundary flips


def _inputs(n, seed=0):
  g = torch.Generator().manual_seed(seed)
  AB = torch.zeros(2 * n, 160, 160, 8)
  AB[..., :3] = torch.rand(2 * n, 160, 160, 3, generator=g)
  AB[..., 3:6] = torch.randn(2 * n, 160, 160, 3, generator=g).clamp(-1.9, 1.9) * 0.5
  # smooth it a little so it resembles image content
  AB[..., :6] = F.avg_pool2d(AB[..., :6].permute(0, 3, 1, 2), 5, 1, 2).permute(0, 2, 3, 1)
  return AB


@pytest.mark.parametrize('n', [3, 5])
def test_refine_fp32_matches_oracle(n, monkeypatch):
  from foundationpose_b200 import nets
  from oracle import nets as on
  sd = nets.synthetic_state_dict(nets.REFINE)
  AB = _inputs(n, 1)
  ref_t, ref_r = on.forward_chunked(on.build('refine', sd), 'refine', AB[:n, ..., :6].numpy(), AB[n:, ..., :6].numpy())
  net = nets.Net(nets.REFINE, sd, 'fp32')
  t, r = net.refine(AB.cuda(), n)
  assert np.abs(t.cpu().numpy() - ref_t).max() < 2e-4, np.abs(t.cpu().numpy() - ref_t).max()
  assert np.abs(r.cpu().numpy() - ref_r).max() < 2e-4, np.abs(r.cpu().numpy() - ref_r).max()


def test_refine_6d_rot_head_matches_oracle():
  """rot_rep='6d': Linear(512, 6) rot head (refine_network.py:61-69) through FPB_NET_REFINE_6D."""
  from foundationpose_b200 import nets
  from oracle import nets as on
  sd = nets.synthetic_state_dict(nets.REFINE, rot_out_dim=6)
  n = 3
  AB = _inputs(n, 6)
  ref_t, ref_r = on.forward_chunked(on.build('refine', sd), 'refine', AB[:n, ..., :6].numpy(), AB[n:, ..., :6].numpy())
  assert ref_r.shape == (n, 6)
  net = nets.Net(nets.REFINE, sd, 'fp32')
  t, r = net.refine(AB.cuda(), n)
  assert r.shape == (n, 6)
  assert np.abs(t.cpu().numpy() - ref_t).max() < 2e-4 and np.abs(r.cpu().numpy() - ref_r).max() < 2e-4
  net16 = nets.Net(nets.REFINE, sd, 'bf16')
  t16, r16 = net16.refine(AB.cuda().bfloat16(), n)
  assert np.abs(r16.cpu().numpy() - ref_r).max() < 0.05 * np.abs(ref_r).max() + 5e-3


def test_refine_bf16_close_to_oracle():
  from foundationpose_b200 import nets
  from oracle import nets as on
  sd = nets.synthetic_state_dict(nets.REFINE)
  n = 6
  AB = _inputs(n, 2)
  ref_t, ref_r = on.forward_chunked(on.build('refine', sd), 'refine', AB[:n, ..., :6].numpy(), AB[n:, ..., :6].numpy())
  net = nets.Net(nets.REFINE, sd, 'bf16')
  t, r = net.refine(AB.cuda().bfloat16(), n)
  et = np.abs(t.cpu().numpy() - ref_t).max(); er = np.abs(r.cpu().numpy() - ref_r).max()
  print('bf16 refine err trans/rot', et, er, 'scale', np.abs(ref_t).max(), np.abs(ref_r).max())
  assert et < 0.05 * max(1e-3, np.abs(ref_t).max()) + 5e-3 and er < 0.05 * max(1e-3, np.abs(ref_r).max()) + 5e-3


def test_score_fp32_matches_oracle():
  from foundationpose_b200 import nets
  from oracle import nets as on
  sd = nets.synthetic_state_dict(nets.SCORE)
  n = 5
  AB = _inputs(n, 3)
  ref_logit, ref_feat = on.forward_chunked(on.build('score', sd), 'score', AB[:n, ..., :6].numpy(), AB[n:, ..., :6].numpy())
  net = nets.Net(nets.SCORE, sd, 'fp32')
  feat = net.score_features(AB.cuda(), n)
  assert np.abs(feat.cpu().numpy() - ref_feat).max() < 2e-4
  logit = net.score_cross(feat, n, 1).reshape(-1)
  assert np.abs(logit.cpu().numpy() - ref_logit).max() < 5e-4
  # cross attention with two groups (bs=2) equals two independent calls
  f2 = torch.cat([feat, feat.flip(0)])
  l2 = net.score_cross(f2, n, 2)
  assert torch.allclose(l2[0], logit, atol=1e-5) and torch.allclose(l2[1], logit.flip(0), atol=1e-5)


def test_score_bf16_close_to_oracle():
  from foundationpose_b200 import nets
  from oracle import nets as on
  sd = nets.synthetic_state_dict(nets.SCORE)
  n = 6
  AB = _inputs(n, 4)
  ref_logit, ref_feat = on.forward_chunked(on.build('score', sd), 'score', AB[:n, ..., :6].numpy(), AB[n:, ..., :6].numpy())
  net = nets.Net(nets.SCORE, sd, 'bf16')
  feat = net.score_features(AB.cuda().bfloat16(), n)
  ef = np.abs(feat.cpu().numpy() - ref_feat).max()
  print('bf16 score feat err', ef, 'scale', np.abs(ref_feat).max())
  assert ef < 0.05 * np.abs(ref_feat).max() + 5e-3


def test_small_and_large_batch_pooling_paths_agree():
  """LayerNorm + token mean runs as 8-CTA clusters (DSMEM reduction) for <= 74 hypotheses and one block per hypothesis
  above, with the same summation order: a hypothesis gets BIT-IDENTICAL outputs in a small shard and in the full batch
  (what makes the hypothesis-sharded register() equal the unsharded one)."""
  from foundationpose_b200 import nets
  sd = nets.synthetic_state_dict(nets.REFINE)
  n, big = 6, 78
  AB = _inputs(n, 2).cuda().bfloat16()
  net = nets.Net(nets.REFINE, sd, 'bf16')
  t, r = net.refine(AB, n)
  rep = big // n
  ABb = torch.cat([AB[:n].repeat(rep, 1, 1, 1), AB[n:].repeat(rep, 1, 1, 1)]).contiguous()
  tb, rb = net.refine(ABb, big)
  for k in range(rep):
    assert torch.equal(tb[k * n:(k + 1) * n], t) and torch.equal(rb[k * n:(k + 1) * n], r)


def test_chunking_is_transparent(monkeypatch):
  """N larger than the hypothesis chunk gives the same result as chunk-sized calls."""
  import os, subprocess, sys, json
  code = ("import torch,sys;sys.path.insert(0,'.');from foundationpose_b200 import nets;from tests.test_nets_gpu import _inputs;"
          "sd=nets.synthetic_state_dict(0);n=5;AB=_inputs(n,1).cuda().bfloat16();net=nets.Net(0,sd,'bf16');"
          "t,r=net.refine(AB,n);print(torch.cat([t,r]).cpu().numpy().tolist())")
  outs = []
  for chunk in ('2', '42'):
    env = dict(os.environ, FPB_NET_CHUNK=chunk)
    out = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, env=env, check=True).stdout
    outs.append(np.array(json.loads(out.strip().splitlines()[-1])))
  assert np.abs(outs[0] - outs[1]).max() < 1e-6


def test_split_k_is_transparent():
  """Single-hypothesis batches (track_one) run the >= 256-channel convolutions split-K (K slices on separate CTAs, fp32 partial
  tiles summed by a second kernel): same outputs as the unsplit kernel up to fp32 summation order."""
  import os, subprocess, sys, json
  code = ("import torch,sys;sys.path.insert(0,'.');from foundationpose_b200 import nets;from tests.test_nets_gpu import _inputs;"
          "from foundationpose_b200._lib import lib;import json;"
          "sd=nets.synthetic_state_dict(0);out=[]\n"
          "for n in (1,):\n"
          "  AB=_inputs(n,1).cuda().bfloat16();net=nets.Net(0,sd,'bf16');lib().fpb_launch_count(1);t,r=net.refine(AB,n);"
          "out.append(torch.cat([t,r]).cpu().numpy().tolist());out.append(int(lib().fpb_launch_count(0)))\n"
          "print(json.dumps(out))")
  res = []
  for flag in ('0', '1'):
    env = dict(os.environ, FPB_SPLITK=flag)
    o = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, env=env, check=True).stdout
    res.append(json.loads(o.strip().splitlines()[-1]))
  a, b = np.array(res[0][0]), np.array(res[1][0])
  assert np.abs(a - b).max() < 2e-3 * max(1.0, np.abs(a).max()), (a, b)
  assert res[1][1] > res[0][1]          # the split path really ran (extra reduce launches)


def _golden_inputs(seed, n):
  from foundationpose_b200.nets import _calibration_batch
  x = _calibration_batch(torch.Generator().manual_seed(seed), n)            # [2n,6,160,160]
  AB = torch.zeros(2 * n, 160, 160, 8)
  AB[..., :6] = x.permute(0, 2, 3, 1)
  return AB


@pytest.mark.parametrize('prec,tol', [('fp32', 2e-4), ('bf16', 2e-2)])
def test_cuda_nets_match_reference_golden(prec, tol):
  """CUDA path vs outputs of the reference's own RefineNet / ScoreNetMultiPair (tests/golden/make_golden.py)."""
  import os
  from foundationpose_b200 import nets
  gold = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
  dt = torch.float if prec == 'fp32' else torch.bfloat16
  g = np.load(os.path.join(gold, 'refine_net.npz'))
  n = int(g['n'])
  net = nets.Net(nets.REFINE, nets.synthetic_state_dict(nets.REFINE), prec)
  t, r = net.refine(_golden_inputs(int(g['seed']), n).cuda().to(dt), n)
  assert np.abs(t.cpu().numpy() - g['trans']).max() < tol and np.abs(r.cpu().numpy() - g['rot']).max() < tol
  g = np.load(os.path.join(gold, 'score_net.npz'))
  n = int(g['n'])
  net = nets.Net(nets.SCORE, nets.synthetic_state_dict(nets.SCORE), prec)
  f = net.score_features(_golden_inputs(int(g['seed']), n).cuda().to(dt), n)
  assert np.abs(f.cpu().numpy() - g['feats']).max() < tol * max(1.0, float(np.abs(g['feats']).max()))
  if prec == 'fp32':
    logit = net.score_cross(f, n, 1).reshape(-1).cpu().numpy()
    assert np.abs(logit - g['logits'].reshape(-1)).max() < 5e-4


@pytest.mark.parametrize('groups', [1, 3])
def test_fused_attention_matches_torch(groups):
  """net_attn.cu vs torch softmax attention on the same bf16 operands (fp32 math): 400 tokens, 4 heads x 128."""
  from foundationpose_b200._lib import lib, check, ptr, stream
  g = torch.Generator(device='cuda').manual_seed(11 + groups)
  R = groups * 400
  qk = (torch.randn(R, 1024, device='cuda', generator=g) * 1.5).bfloat16()
  v = torch.randn(R, 512, device='cuda', generator=g).bfloat16()
  bv = torch.randn(512, device='cuda', generator=g) * 0.1
  vt = v.t().contiguous()                                   # [512, R]
  ctx = torch.empty(R, 512, device='cuda', dtype=torch.bfloat16)
  check(lib().fpb_debug_attention(ptr(qk), ptr(vt), R, ptr(bv), ptr(ctx), groups, stream()))
  torch.cuda.synchronize()
  q = qk[:, :512].float().reshape(groups, 400, 4, 128).permute(0, 2, 1, 3)
  k = qk[:, 512:].float().reshape(groups, 400, 4, 128).permute(0, 2, 1, 3)
  vv = v.float().reshape(groups, 400, 4, 128).permute(0, 2, 1, 3)
  p = torch.softmax(q @ k.transpose(-1, -2) / 128 ** 0.5, dim=-1)
  ref = (p @ vv).permute(0, 2, 1, 3).reshape(R, 512) + bv
  err = (ctx.float() - ref).abs().max().item()
  assert err < 2e-2, err                                    # bf16 probabilities + bf16 output rounding


def _to_pos(x):
  """dense [n,H,W,C] -> zero-bordered positions [n*(H+1)*(W+1), C] (csrc/net_kernels.cuh halo layout)."""
  n, H, W, C = x.shape
  P = torch.zeros(n, H + 1, W + 1, C, dtype=x.dtype, device=x.device)
  P[:, 1:, :W] = x
  return P.reshape(-1, C)


def _from_pos(P, n, H, W):
  return P.reshape(n, H + 1, W + 1, -1)[:, 1:, :W]


@pytest.mark.parametrize('cg', [1, 2, 3])          # 3: operand-swapped kernel (net_tc3.cu)
@pytest.mark.parametrize('shape', [(5, 40, 128, 128), (3, 20, 256, 256), (7, 40, 64, 128), (4, 40, 64, 64), (1, 7, 128, 128)])
def test_conv3x3_halo_layout(cg, shape):
  from foundationpose_b200._lib import lib, check, ptr, stream
  n, H, Cin, Cout = shape
  if cg == 3 and Cout > 128:
    pytest.skip('operand-swapped kernel covers 64 / 128 output channels')
  g = torch.Generator(device='cuda').manual_seed(n * 100 + H)
  x = torch.randn(n, H, H, Cin, device='cuda', generator=g).bfloat16()
  w = (torch.randn(Cout, Cin, 3, 3, device='cuda', generator=g) / (9 * Cin) ** 0.5).bfloat16()
  scale = torch.rand(Cout, device='cuda', generator=g) + 0.5
  shift = torch.randn(Cout, device='cuda', generator=g) * 0.1
  res = torch.randn(n, H, H, Cout, device='cuda', generator=g).bfloat16()
  Wk = w.permute(0, 2, 3, 1).reshape(Cout, -1).contiguous()
  ref = _ref_conv(x.float(), w.float(), 1, 1, scale, shift, res.float().reshape(-1, Cout), relu=1)
  npos = n * (H + 1) * (H + 1)
  xp, rp = _to_pos(x), _to_pos(res)          # keep alive: ptr() of a temporary would dangle
  # (a) positions -> positions, residual in position layout, borders must come out as zeros
  out = torch.full((npos, Cout), 7.0, device='cuda', dtype=torch.bfloat16)
  check(lib().fpb_debug_conv3x3_halo(cg, ptr(xp), ptr(Wk), ptr(out), n, H, H, Cin, Cout, 1, 0, ptr(scale), ptr(shift),
                                     ptr(rp), 1, stream()))
  torch.cuda.synchronize()
  got = _from_pos(out, n, H, H).float()
  assert (got - ref).abs().max().item() < 3e-2, (got - ref).abs().max().item()
  full = out.reshape(n, H + 1, H + 1, Cout).float()
  assert full[:, 0].abs().max().item() == 0 and full[:, :, H].abs().max().item() == 0
  # (b) positions -> dense rows
  out2 = torch.zeros(n * H * H, Cout, device='cuda', dtype=torch.bfloat16)
  check(lib().fpb_debug_conv3x3_halo(cg, ptr(xp), ptr(Wk), ptr(out2), n, H, H, Cin, Cout, 1, 2, ptr(scale), ptr(shift),
                                     ptr(rp), 1, stream()))
  torch.cuda.synchronize()
  assert (out2.reshape(n, H, H, Cout).float() - ref).abs().max().item() < 3e-2
  # (c) dense input through the im2col path, output rows remapped to positions (borders untouched)
  out3 = torch.zeros(npos, Cout, device='cuda', dtype=torch.bfloat16)
  rd = res.reshape(-1, Cout).contiguous()
  check(lib().fpb_debug_conv3x3_halo(cg, ptr(x), ptr(Wk), ptr(out3), n, H, H, Cin, Cout, 0, 1, ptr(scale), ptr(shift),
                                     ptr(rd), 1, stream()))
  torch.cuda.synchronize()
  assert (_from_pos(out3, n, H, H).float() - ref).abs().max().item() < 3e-2
  assert out3.reshape(n, H + 1, H + 1, Cout)[:, 0].abs().max().item() == 0


def _ffn_inputs(M, seed, dt):
  g = torch.Generator(device='cuda').manual_seed(seed)
  rn = lambda *s: torch.randn(*s, device='cuda', generator=g)
  A = (rn(M, 512) * 1.2 + 0.3).to(dt)                       # rows with a non-zero mean (shifted-sum statistics)
  res = rn(M, 512).to(dt)
  W1, W2 = (rn(512, 512) / 512 ** 0.5).to(dt), (rn(512, 512) / 512 ** 0.5).to(dt)
  b1, b2 = rn(512) * 0.2, rn(512) * 0.2
  gamma, beta = 1.0 + 0.3 * rn(512), 0.2 * rn(512)
  return A, res, W1, W2, b1, b2, gamma, beta


def _ln(v, gamma, beta):
  mu = v.mean(-1, keepdim=True)
  var = ((v - mu) ** 2).mean(-1, keepdim=True)
  return (v - mu) / torch.sqrt(var + 1e-5) * gamma.double() + beta.double()


@pytest.mark.parametrize('prec', ['bf16', 'fp16'])
@pytest.mark.parametrize('M', [400, 2000, 64 * 400, 252 * 400])
def test_row_tile_proj_layernorm(prec, M):
  """net_ffn.cu, proj+LN kernel: LayerNorm(A . W^T + b + residual) vs fp64 torch on the same 16-bit operands
  (one ragged tile, several tiles, more tiles than CTA pairs so every pipeline phase wraps)."""
  from foundationpose_b200._lib import lib, check, ptr, stream
  dt = torch.bfloat16 if prec == 'bf16' else torch.float16
  check(lib().fpb_debug_set_precision(1 if prec == 'bf16' else 2))
  try:
    A, res, _, W2, _, b2, gamma, beta = _ffn_inputs(M, 5 + M % 97, dt)
    out = torch.full((M + 8, 512), 7.0, device='cuda', dtype=dt)       # guard rows behind the tensor must stay untouched
    check(lib().fpb_debug_ffn(0, ptr(A), None, None, ptr(W2), ptr(b2), ptr(res), ptr(gamma), ptr(beta), ptr(out), None, M, 400,
                              stream()))
    torch.cuda.synchronize()
  finally:
    check(lib().fpb_debug_set_precision(1))
  ref = _ln(A.double() @ W2.double().t() + b2.double() + res.double(), gamma, beta)
  err = (out[:M].double() - ref).abs().max().item()
  assert (out[M:] == 7.0).all()
  assert err < (4e-2 if prec == 'bf16' else 5e-3), err      # one 16-bit rounding of outputs of magnitude <= ~6


@pytest.mark.parametrize('prec', ['bf16', 'fp16'])
@pytest.mark.parametrize('M', [400, 2000, 64 * 400, 252 * 400])
def test_row_tile_ffn_layernorm_token_mean(prec, M):
  """net_ffn.cu, fused linear1 + ReLU + linear2 + residual + LayerNorm2 + token mean vs fp64 torch (h rounded to the
  operand type like the kernel's shared-memory tile)."""
  from foundationpose_b200._lib import lib, check, ptr, stream
  dt = torch.bfloat16 if prec == 'bf16' else torch.float16
  check(lib().fpb_debug_set_precision(1 if prec == 'bf16' else 2))
  try:
    A, _, W1, W2, b1, b2, gamma, beta = _ffn_inputs(M, 17 + M % 89, dt)
    G = M // 400
    out = torch.full((G + 1, 512), 7.0, device='cuda')     # one guard row behind each output: must stay untouched
    scratch = torch.full((M // 16 + 16, 512), 7.0, device='cuda')
    for _ in range(2):                                       # twice: scratch / pipeline state of a first call must not matter
      check(lib().fpb_debug_ffn(1, ptr(A), ptr(W1), ptr(b1), ptr(W2), ptr(b2), None, ptr(gamma), ptr(beta), ptr(out),
                                ptr(scratch), M, 400, stream()))
    torch.cuda.synchronize()
  finally:
    check(lib().fpb_debug_set_precision(1))
  assert (out[G:] == 7.0).all() and (scratch[M // 16:] == 7.0).all()
  out = out[:G]
  h = torch.relu(A.double() @ W1.double().t() + b1.double()).to(dt).double()
  ref = _ln(h @ W2.double().t() + b2.double() + A.double(), gamma, beta).reshape(G, 400, 512).mean(1)
  err = (out.double() - ref).abs().max().item()
  assert err < (2e-3 if prec == 'bf16' else 3e-4), err


_FALLBACK_SCRIPT = r'''
import sys
import numpy as np
import torch
sys.path.insert(0, sys.argv[1])
from foundationpose_b200 import nets
from tests.test_nets_gpu import _inputs
sd = nets.synthetic_state_dict(nets.REFINE)
n = 5
AB = _inputs(n, 3)
net = nets.Net(nets.REFINE, sd, 'fp16')
t, r = net.refine(AB.cuda().half(), n)
np.save(sys.argv[2], np.concatenate([t.cpu().numpy().ravel(), r.cpu().numpy().ravel()]))
'''


def test_row_tile_kernels_agree_with_the_unfused_head_path(tmp_path):
  """FPB_FFN is read once per process, so the unfused path (one GEMM per linear + LayerNorm kernels, FPB_FFN=0) runs in a
  subprocess: RefineNet outputs of the two paths agree to the fp16 rounding of the intermediate rows."""
  import os
  import subprocess
  import sys
  root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
  outs = {}
  for flag in ('1', '0'):
    f = str(tmp_path / f'out{flag}.npy')
    env = dict(os.environ, FPB_FFN=flag)
    subprocess.run([sys.executable, '-c', _FALLBACK_SCRIPT, root, f], check=True, env=env, timeout=300)
    outs[flag] = np.load(f)
  d = np.abs(outs['1'] - outs['0']).max()
  scale = np.abs(outs['0']).max()
  print('row-tile vs unfused head path: max diff', d, 'output scale', scale)
  assert d > 0.0                       # they really are two different code paths
  assert d < 2e-4                      # measured 8.6e-6 on outputs of scale 0.19: only the rounding point of the LayerNorm inputs differs
