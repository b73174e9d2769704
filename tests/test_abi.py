"""CPU: the C-ABI library loads and exports every symbol include/fpb200.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
  src = open(os.path.join(ROOT, 'include', 'fpb200.h')).read()
  src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
  return sorted(set(re.findall(r'\b(fpb_[a-z0-9_]+)\s*\(', src)))


def test_header_symbols_are_exported():
  import __graft_entry__ as ge
  so = os.path.join(ROOT, 'foundationpose_b200', 'libfpb200.so')
  if not os.path.exists(so):
    ge.build()
  L = ctypes.CDLL(so)
  names = _declared()
  assert len(names) >= 24
  missing = [n for n in names if not hasattr(L, n)]
  assert not missing, missing


def test_python_binding_covers_header():
  from foundationpose_b200 import _lib
  assert sorted(set(_lib.exported_symbols())) == _declared()


def test_host_only_helpers():
  from foundationpose_b200._lib import lib
  assert lib().fpb_version() == 100
  assert lib().fpb_net_blob_floats(0) == 17040784 and lib().fpb_net_blob_floats(1) == 15983496
  assert lib().fpb_render_ws_bytes(252, 10242, 160) > 252 * 160 * 160 * 8
  assert lib().fpb_net_ws_bytes(0, 1, 1000) == lib().fpb_net_ws_bytes(0, 1, 256)   # chunked (FPB_NET_CHUNK = 256)


def test_no_cpu_fallback():
  import numpy as np
  import pytest
  import torch
  if torch.cuda.is_available():
    pytest.skip('GPU present')
  from foundationpose_b200 import utils as U
  with pytest.raises(RuntimeError):
    U.erode_depth(np.zeros((4, 4), np.float32))
