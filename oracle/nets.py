"""CPU ORACLE (test infrastructure, NOT product code) -- fp32 torch restatement of
RefineNet / ScoreNetMultiPair with the reference's state_dict keys.

  reference: learning/models/network_modules.py:37-50,73-111,115-137
             learning/models/refine_network.py:26-93
             learning/models/score_network.py:27-90
PINNED in this container: tests/test_oracle_nets_vs_reference.py loads the same
state_dict into the reference's own modules (imported from /root/reference with a
stub `Utils`) and requires identical outputs; golden vectors produced that way are
committed under tests/golden/ with their generator (tests/golden/make_golden.py).
Only tests/, __graft_entry__.smoke(), bench.py cpu_baseline / --impl reference
may import this module.
"""
import math

import torch
import torch.nn as nn


class _ConvBNReLU(nn.Module):
  def __init__(self, cin, cout, k, stride):
    super().__init__()
    self.net = nn.Sequential(nn.Conv2d(cin, cout, k, stride, (k - 1) // 2, bias=True),
                             nn.BatchNorm2d(cout), nn.ReLU(inplace=True))

  def forward(self, x):
    return self.net(x)


class _Block(nn.Module):
  def __init__(self, c):
    super().__init__()
    self.conv1 = nn.Conv2d(c, c, 3, 1, 1, bias=True)
    self.bn1 = nn.BatchNorm2d(c)
    self.conv2 = nn.Conv2d(c, c, 3, 1, 1, bias=True)
    self.bn2 = nn.BatchNorm2d(c)

  def forward(self, x):
    out = torch.relu(self.bn1(self.conv1(x)))
    out = self.bn2(self.conv2(out))
    return torch.relu(out + x)


class _PosEmbed(nn.Module):
  def __init__(self, d_model=512, max_len=400):
    super().__init__()
    pe = torch.zeros(max_len, d_model).float()
    position = torch.arange(0, max_len).float().unsqueeze(1)
    div_term = (torch.arange(0, d_model, 2).float() * -(math.log(10000.0) / d_model)).exp()[None]
    pe[:, 0::2] = torch.sin(position * div_term)
    pe[:, 1::2] = torch.cos(position * div_term)
    self.register_buffer('pe', pe.unsqueeze(0))

  def forward(self, x):
    return x + self.pe[:, :x.size(1)]


def _encoder_a(c_in):
  return nn.Sequential(_ConvBNReLU(c_in, 64, 7, 2), _ConvBNReLU(64, 128, 3, 2), _Block(128), _Block(128))


def _encoder_ab():
  return nn.Sequential(_Block(256), _Block(256), _ConvBNReLU(256, 512, 3, 2), _Block(512), _Block(512))


class RefineNet(nn.Module):
  def __init__(self, c_in=6, rot_out_dim=3):
    super().__init__()
    self.encodeA = _encoder_a(c_in)
    self.encodeAB = _encoder_ab()
    self.pos_embed = _PosEmbed(512, 400)
    self.trans_head = nn.Sequential(
      nn.TransformerEncoderLayer(d_model=512, nhead=4, dim_feedforward=512, batch_first=True), nn.Linear(512, 3))
    self.rot_head = nn.Sequential(
      nn.TransformerEncoderLayer(d_model=512, nhead=4, dim_feedforward=512, batch_first=True),
      nn.Linear(512, rot_out_dim))

  def forward(self, A, B):
    bs = len(A)
    x = self.encodeA(torch.cat([A, B], dim=0))
    ab = torch.cat((x[:bs], x[bs:]), 1).contiguous()
    ab = self.encodeAB(ab)
    ab = self.pos_embed(ab.reshape(bs, ab.shape[1], -1).permute(0, 2, 1))
    return {'trans': self.trans_head(ab).mean(dim=1), 'rot': self.rot_head(ab).mean(dim=1)}


class ScoreNetMultiPair(nn.Module):
  def __init__(self, c_in=6):
    super().__init__()
    self.encoderA = _encoder_a(c_in)
    self.encoderAB = _encoder_ab()
    self.att = nn.MultiheadAttention(embed_dim=512, num_heads=4, bias=True, batch_first=True)
    self.att_cross = nn.MultiheadAttention(embed_dim=512, num_heads=4, bias=True, batch_first=True)
    self.pos_embed = _PosEmbed(512, 400)
    self.linear = nn.Linear(512, 1)

  def extract_feat(self, A, B):
    bs = A.shape[0]
    x = self.encoderA(torch.cat([A, B], dim=0))
    ab = self.encoderAB(torch.cat((x[:bs], x[bs:]), dim=1))
    ab = self.pos_embed(ab.reshape(bs, ab.shape[1], -1).permute(0, 2, 1))
    ab, _ = self.att(ab, ab, ab)
    return ab.mean(dim=1).reshape(bs, -1)

  def forward(self, A, B, L):
    bs = A.shape[0] // L
    x = self.extract_feat(A, B).reshape(bs, L, -1)
    x, _ = self.att_cross(x, x, x)
    return {'score_logit': self.linear(x).reshape(bs, L)}


def build(kind, state_dict):
  """kind in {'refine','score'}; strict load proves state_dict key compatibility."""
  net = RefineNet(rot_out_dim=int(state_dict['rot_head.1.weight'].shape[0])) if kind == 'refine' else ScoreNetMultiPair()
  net.load_state_dict(state_dict, strict=True)
  return net.eval()


@torch.no_grad()
def score_from_features(net, feats):
  """score_network.py:80-90 after extract_feat: feats [L,512] -> logits [L] (att_cross over the L hypotheses + linear)."""
  x = feats.reshape(1, feats.shape[0], -1)
  x, _ = net.att_cross(x, x, x)
  return net.linear(x).reshape(-1)


@torch.no_grad()
def forward_chunked(net, kind, A_nhwc, B_nhwc, chunk=16, autocast=None):
  """A/B numpy [N,S,S,6] -> refine: (trans [N,3], rot [N,3]); score: feats handled whole (L=N).
  autocast: None (fp32) or a torch dtype -- the model call runs under torch.autocast('cpu', dtype) and the outputs are
  cast back to fp32, as the reference does on CUDA (predict_pose_refine.py:190-193, predict_score.py:193-195)."""
  import contextlib
  import numpy as np
  if autocast is not None:
    with torch.autocast('cpu', dtype=autocast):
      out = forward_chunked(net, kind, A_nhwc, B_nhwc, chunk, None)
    return tuple(np.asarray(o, dtype=np.float32) for o in out)
  A = torch.from_numpy(np.ascontiguousarray(A_nhwc)).permute(0, 3, 1, 2).contiguous()
  B = torch.from_numpy(np.ascontiguousarray(B_nhwc)).permute(0, 3, 1, 2).contiguous()
  N = len(A)
  if kind == 'refine':
    tr, ro = [], []
    for b in range(0, N, chunk):
      o = net(A[b:b + chunk], B[b:b + chunk])
      tr.append(o['trans']); ro.append(o['rot'])
    return torch.cat(tr).float().numpy(), torch.cat(ro).float().numpy()
  feats = torch.cat([net.extract_feat(A[b:b + chunk], B[b:b + chunk]) for b in range(0, N, chunk)])
  x = feats.reshape(1, N, -1)
  x, _ = net.att_cross(x, x, x)
  return net.linear(x).reshape(N).float().numpy(), feats.float().numpy()
