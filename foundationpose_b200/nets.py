"""Weight packing and device handles for RefineNet / ScoreNetMultiPair (reference:
learning/models/refine_network.py:26-93, score_network.py:27-90, network_modules.py:37-137).

A reference `state_dict` (134 / 116 entries, same keys as the released checkpoints) is folded (eval-mode
BatchNorm -> per-channel scale/shift), re-laid out K-major ([Cout][KH][KW][Cin], Cin 6->8 and the 7-wide kernel
row padded to 8 taps) and shipped to libfpb200.so as one flat fp32 blob; the layout is mirrored by
csrc/net.cu:layout().  Seeded synthetic weights stand in for the checkpoints, which are unavailable offline.
"""
import ctypes
import math

import numpy as np
import torch

from . import _lib
from ._lib import lib, check, ptr, stream

REFINE, SCORE = 0, 1
REFINE_6D = 2      # fpb_net_create kind for a RefineNet whose rot head has 6 outputs (include/fpb200.h)
PREC = {'fp32': 0, 'bf16': 1, 'fp16': 2}     # FPB_PREC_* (include/fpb200.h)
ACT_DTYPE = {'fp32': torch.float, 'bf16': torch.bfloat16, 'fp16': torch.float16}
D, TOK = 512, 400
# (Cin, Cout, k, stride) of the 15 convolutions, in execution order
_CONVS = [(6, 64, 7, 2), (64, 128, 3, 2)] + [(128, 128, 3, 1)] * 4 + [(256, 256, 3, 1)] * 4 + [(256, 512, 3, 2)] + \
         [(512, 512, 3, 1)] * 4


def _conv_keys(kind):
  a, ab = ('encodeA', 'encodeAB') if kind == REFINE else ('encoderA', 'encoderAB')
  keys = [(f'{a}.0.net.0', f'{a}.0.net.1'), (f'{a}.1.net.0', f'{a}.1.net.1')]
  for blk in (f'{a}.2', f'{a}.3', f'{ab}.0', f'{ab}.1'):
    keys += [(f'{blk}.conv1', f'{blk}.bn1'), (f'{blk}.conv2', f'{blk}.bn2')]
  keys.append((f'{ab}.2.net.0', f'{ab}.2.net.1'))
  for blk in (f'{ab}.3', f'{ab}.4'):
    keys += [(f'{blk}.conv1', f'{blk}.bn1'), (f'{blk}.conv2', f'{blk}.bn2')]
  return keys


def positional_embedding(d_model=D, max_len=TOK):
  """reference: learning/models/network_modules.py:115-137"""
  pe = torch.zeros(max_len, d_model).float()
  position = torch.arange(0, max_len).float().unsqueeze(1)
  div_term = (torch.arange(0, d_model, 2).float() * -(math.log(10000.0) / d_model)).exp()[None]
  pe[:, 0::2] = torch.sin(position * div_term)
  pe[:, 1::2] = torch.cos(position * div_term)
  return pe.unsqueeze(0)


class _Blob:
  def __init__(self):
    self.parts = []

  def add(self, t):
    a = np.ascontiguousarray(t.detach().cpu().float().numpy().reshape(-1), dtype=np.float32)
    pad = (-len(a)) % 8
    if pad:
      a = np.concatenate([a, np.zeros(pad, np.float32)])
    self.parts.append(a)

  def lin(self, sd, name, wkey='weight', bkey='bias'):
    self.add(sd[f'{name}.{wkey}'])
    self.add(sd[f'{name}.{bkey}'])

  def mha(self, sd, name):
    self.add(sd[f'{name}.in_proj_weight'])
    self.add(sd[f'{name}.in_proj_bias'])
    self.lin(sd, f'{name}.out_proj')

  def finish(self):
    return np.ascontiguousarray(np.concatenate(self.parts))


def pack_state_dict(kind, sd, eps=1e-5):
  """state_dict -> flat fp32 blob (BatchNorm folded when the bn keys exist, i.e. cfg.use_BN)."""
  b = _Blob()
  for (ck, bk), (cin, cout, k, _) in zip(_conv_keys(kind), _CONVS):
    w = sd[f'{ck}.weight'].detach().cpu().float()                      # [Cout,Cin,k,k]
    bias = sd[f'{ck}.bias'].detach().cpu().float()
    cinp, kwp = (8, 8) if k == 7 else (cin, k)
    wp = torch.zeros(cout, k, kwp, cinp)
    wp[:, :, :k, :cin] = w.permute(0, 2, 3, 1)
    if f'{bk}.weight' in sd:
      g, beta = sd[f'{bk}.weight'].cpu().float(), sd[f'{bk}.bias'].cpu().float()
      mean, var = sd[f'{bk}.running_mean'].cpu().float(), sd[f'{bk}.running_var'].cpu().float()
      scale = g / torch.sqrt(var + eps)
      shift = (bias - mean) * scale + beta
    else:
      scale, shift = torch.ones(cout), bias
    b.add(wp); b.add(scale); b.add(shift)
  pe = sd['pos_embed.pe'] if 'pos_embed.pe' in sd else positional_embedding()
  b.add(pe.reshape(-1)[:TOK * D])
  if kind == REFINE:
    for head in ('trans_head', 'rot_head'):
      b.mha(sd, f'{head}.0.self_attn')
      b.lin(sd, f'{head}.0.linear1')
      b.lin(sd, f'{head}.0.linear2')
      for nk in ('norm1', 'norm2'):
        b.add(sd[f'{head}.0.{nk}.weight']); b.add(sd[f'{head}.0.{nk}.bias'])
      fw, fb = sd[f'{head}.1.weight'], sd[f'{head}.1.bias']
      want = (3, 6) if head == 'rot_head' else (3,)       # rot_rep='6d' -> Linear(512, 6) (refine_network.py:61-69)
      if fw.shape[0] not in want:
        raise ValueError(f'{head}.1.weight has {fw.shape[0]} rows, expected one of {want}')
      b.add(fw); b.add(fb)
  else:
    b.mha(sd, 'att')
    b.mha(sd, 'att_cross')
    b.lin(sd, 'linear')
  return b.finish()


def synthetic_state_dict(kind, seed=1234, head_gain=None, calibrate=True, rot_out_dim=3):
  """Seeded random weights with the reference's keys and shapes: He-initialised convolutions,
  randomised eval-mode BatchNorm statistics, xavier attention projections; the last Linear of each head is
  scaled so pose updates are a few mm / degrees and score logits are separated."""
  g = torch.Generator().manual_seed(seed + 17 * kind)
  sd = {}

  def randn(*s):
    return torch.randn(*s, generator=g)

  def rand(*s):
    return torch.rand(*s, generator=g)

  for i, ((ck, bk), (cin, cout, k, _)) in enumerate(zip(_conv_keys(kind), _CONVS)):
    gain = 0.5 if ck.endswith('conv2') else 1.0
    sd[f'{ck}.weight'] = randn(cout, cin, k, k) * (gain * math.sqrt(2.0 / (cin * k * k)))
    sd[f'{ck}.bias'] = randn(cout) * 0.05
    sd[f'{bk}.weight'] = 0.5 + rand(cout)
    sd[f'{bk}.bias'] = randn(cout) * 0.1
    sd[f'{bk}.running_mean'] = randn(cout) * 0.1
    sd[f'{bk}.running_var'] = 0.5 + rand(cout)
    sd[f'{bk}.num_batches_tracked'] = torch.tensor(0, dtype=torch.long)
  if calibrate:
    _calibrate_batchnorm(kind, sd, g)
  sd['pos_embed.pe'] = positional_embedding()

  def lin(name, n, k, gain=1.0, bias_std=0.02):
    sd[f'{name}.weight'] = (rand(n, k) * 2 - 1) * (gain / math.sqrt(k))
    sd[f'{name}.bias'] = randn(n) * bias_std

  def mha(name, qk_gain=1.0):
    w = (rand(3 * D, D) * 2 - 1) * math.sqrt(6.0 / (4 * D))
    w[:2 * D] *= qk_gain      # sharper attention: random-init q.k logits are otherwise ~uniform
    sd[f'{name}.in_proj_weight'] = w
    sd[f'{name}.in_proj_bias'] = randn(3 * D) * 0.02
    lin(f'{name}.out_proj', D, D)

  if kind == REFINE:
    hg = head_gain if head_gain is not None else {'trans_head': 0.05, 'rot_head': 0.3}
    for head in ('trans_head', 'rot_head'):
      mha(f'{head}.0.self_attn')
      lin(f'{head}.0.linear1', D, D)
      lin(f'{head}.0.linear2', D, D)
      for nk in ('norm1', 'norm2'):
        sd[f'{head}.0.{nk}.weight'] = 0.75 + 0.5 * rand(D)
        sd[f'{head}.0.{nk}.bias'] = randn(D) * 0.05
      lin(f'{head}.1', rot_out_dim if head == 'rot_head' else 3, D, gain=hg[head], bias_std=0.0)
  else:
    mha('att')
    mha('att_cross', qk_gain=10.0)
    lin('linear', 1, D, gain=head_gain if head_gain is not None else 8.0, bias_std=0.0)
  return sd


def _calibration_batch(g, n=3):
  """A few smooth pseudo-crops (rgb in [0,1], xyz in [-1,1] with empty regions) as A-like and B-like inputs."""
  import torch.nn.functional as F
  x = torch.rand(2 * n, 6, 20, 20, generator=g)
  x = F.interpolate(x, size=(160, 160), mode='bilinear', align_corners=False)
  x = x + 0.05 * torch.randn(2 * n, 6, 160, 160, generator=g)
  x[:, 3:] = (x[:, 3:] - 0.5) * 2
  yy, xx = torch.meshgrid(torch.arange(160), torch.arange(160), indexing='ij')
  for i in range(2 * n):
    hole = ((xx - 80 - 10 * i) ** 2 + (yy - 80 + 7 * i) ** 2) > (55 + 3 * i) ** 2
    x[i, :, hole] = 0
  return x.clamp(-2, 2)


@torch.no_grad()
def _calibrate_batchnorm(kind, sd, g):
  """Data-dependent init: set every BatchNorm's running statistics to the statistics its input actually has on a
  small seeded batch, as a trained network's would be.  Without it the random-init features are dominated by a
  common-mode component and ScoreNet's ranking is degenerate."""
  import torch.nn.functional as F
  keys = _conv_keys(kind)
  x = _calibration_batch(g)
  n = x.shape[0] // 2

  def conv_bn(i, x, relu=True, residual=None):
    ck, bk = keys[i]
    cin, cout, k, stride = _CONVS[i]
    y = F.conv2d(x, sd[f'{ck}.weight'], sd[f'{ck}.bias'], stride, (k - 1) // 2)
    sd[f'{bk}.running_mean'] = y.mean(dim=(0, 2, 3)).clone()
    sd[f'{bk}.running_var'] = y.var(dim=(0, 2, 3), unbiased=False).clone() + 1e-3
    y = F.batch_norm(y, sd[f'{bk}.running_mean'], sd[f'{bk}.running_var'], sd[f'{bk}.weight'], sd[f'{bk}.bias'], False, 0.0, 1e-5)
    if residual is not None:
      y = y + residual
    return torch.relu(y) if relu else y

  x = conv_bn(0, x); x = conv_bn(1, x)
  for i in (2, 4):
    x = conv_bn(i + 1, conv_bn(i, x), residual=x)
  x = torch.cat((x[:n], x[n:]), 1)
  for i in (6, 8):
    x = conv_bn(i + 1, conv_bn(i, x), residual=x)
  x = conv_bn(10, x)
  for i in (11, 13):
    x = conv_bn(i + 1, conv_bn(i, x), residual=x)


def pad_input(AB):
  """Dense [n,160,160,8] -> row-padded [n,160,168,8] (no-op if already padded); test convenience."""
  if AB.shape[2] == 168:
    return AB
  out = torch.zeros((AB.shape[0], 160, 168, 8), dtype=AB.dtype, device=AB.device)
  out[:, :, 4:164] = AB
  return out


class Net:
  """Device-side network handle (fpb_net) plus its workspace."""

  def __init__(self, kind, state_dict, precision='fp16'):
    _lib.require_cuda()
    self.kind, self.precision = kind, PREC[precision]
    self.bf16 = precision == 'bf16'
    self.precision_name = precision
    blob = pack_state_dict(kind, state_dict)
    self.rot_dim = int(state_dict['rot_head.1.weight'].shape[0]) if kind == REFINE else 0
    h = ctypes.c_void_p()
    check(lib().fpb_net_create(ctypes.byref(h), REFINE_6D if self.rot_dim == 6 else kind, self.precision,
                               blob.ctypes.data_as(ctypes.c_void_p), blob.nbytes, stream()))
    self.handle = h
    self._ws = None

  @property
  def act_dtype(self):
    return ACT_DTYPE[self.precision_name]

  def workspace(self, N):
    need = lib().fpb_net_ws_bytes(self.kind, self.precision, N)
    if self._ws is None or self._ws.numel() < need:
      self._ws = None
      self._ws = torch.empty(need, dtype=torch.uint8, device='cuda')
    return self._ws

  def refine(self, AB, N, shared_b=False):
    """AB [2N,160,168,8] (utils.net_input_buffer layout) -> (trans [N,3], rot [N,3] or [N,6] for rot_rep='6d') fp32.
    shared_b: only AB[:N+1] is read -- N rendered crops and ONE observed crop shared by all hypotheses
    (fpb_refine_forward_shared_b; bit-identical to the replicated input)."""
    AB = pad_input(AB)
    assert AB.dtype == self.act_dtype and AB.shape[1:] == (160, 168, 8) and AB.shape[0] >= (N + 1 if shared_b else 2 * N)
    assert shared_b or AB.shape[0] == 2 * N
    trans = torch.empty((N, 3), dtype=torch.float, device='cuda')
    rot = torch.empty((N, self.rot_dim), dtype=torch.float, device='cuda')
    fn = lib().fpb_refine_forward_shared_b if shared_b else lib().fpb_refine_forward
    check(fn(self.handle, ptr(AB), N, ptr(self.workspace(N)), ptr(trans), ptr(rot), stream()))
    return trans, rot

  def score_features(self, AB, N):
    AB = pad_input(AB)
    assert AB.dtype == self.act_dtype and AB.shape == (2 * N, 160, 168, 8)
    feat = torch.empty((N, D), dtype=torch.float, device='cuda')
    check(lib().fpb_score_features(self.handle, ptr(AB), N, ptr(self.workspace(N)), ptr(feat), stream()))
    return feat

  def score_features_multi(self, AB, N, dst_ptrs):
    """score_features with the pooled [N,512] rows stored to every address of dst_ptrs (ints: local / peer-mapped
    device addresses already offset to this rank's first row) -- the fused all-gather of dist.register_sharded."""
    AB = pad_input(AB)
    assert AB.dtype == self.act_dtype and AB.shape == (2 * N, 160, 168, 8)
    arr = (ctypes.c_void_p * len(dst_ptrs))(*[int(p) for p in dst_ptrs])
    check(lib().fpb_score_features_multi(self.handle, ptr(AB), N, ptr(self.workspace(N)), arr, len(dst_ptrs), stream()))

  def score_cross(self, feat, L, bs=1):
    """feat [bs*L,512] fp32 -> logits [bs,L]"""
    need = lib().fpb_score_cross_ws_bytes(L, bs)
    ws = torch.empty(need, dtype=torch.uint8, device='cuda')
    out = torch.empty((bs, L), dtype=torch.float, device='cuda')
    check(lib().fpb_score_cross(self.handle, ptr(feat.contiguous()), L, bs, ptr(ws), ptr(out), stream()))
    return out

  def __del__(self):
    try:
      if getattr(self, 'handle', None):
        lib().fpb_net_destroy(self.handle)
        self.handle = None
    except Exception:
      pass
