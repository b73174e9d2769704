"""Pipeline surface of the reference README (readme.md:126-185): host logic only (stub estimator, file readers)."""
import math

import numpy as np
import pytest

from foundationpose_b200 import pipeline as P


class _StubEst:
  def __init__(self):
    self.calls = []

  def register(self, K, rgb, depth, ob_mask, iteration=5):
    self.calls.append((K.copy(), rgb.shape, depth.shape, ob_mask.dtype, iteration))
    pose = np.eye(4)
    pose[:3, 3] = [0.1, -0.2, 0.6]
    return pose


def test_pipeline_runs_processors_in_order_and_collects_errors():
  class Add(P.Processor):
    def __init__(self, k):
      self.k = k

    def process(self, data):
      data.trace = getattr(data, 'trace', []) + [self.k]
      return data

  class Boom(P.Processor):
    def process(self, data):
      raise RuntimeError('boom')

  pipe = P.Pipeline('t').add_processor(Add(1)).add_processor(Boom()).add_processor(Add(2))
  out = pipe.run(P.PipelineData())
  assert out.trace == [1, 2] and out.errors[0][0] == 'Boom'
  with pytest.raises(RuntimeError):
    P.Pipeline('t', stop_on_error=True).add_processor(Boom()).run(P.PipelineData())
  with pytest.raises(TypeError):
    P.Pipeline('t').add_processor(object())


def test_estimator_processor_with_stub(tmp_path):
  kf = tmp_path / 'cam_K.txt'
  kf.write_text('600 0 320\n0 600 240\n0 0 1\n')
  est = _StubEst()
  proc = P.FoundationPoseEstimator(mesh_file=dict(pos=np.zeros((3, 3)), vnormals=np.zeros((3, 3))), cam_intrinsics_file=str(kf),
                                   est_refine_iter=3, estimator=est)
  data = P.PipelineData(rgb=np.zeros((480, 640, 3), np.uint8), depth=np.zeros((480, 640), np.float32),
                        mask=np.ones((480, 640), np.uint8))
  out = P.Pipeline('pose').add_processor(proc).run(data)
  assert not out.errors
  K, rgb_shape, d_shape, mdt, it = est.calls[0]
  assert K[0, 0] == 600 and K[1, 2] == 240 and mdt == np.bool_ and it == 3
  assert out.pose.shape == (4, 4) and out.pose_6d[:3] == (0.1, -0.2, 0.6) and out.pose_6d[3:] == (0.0, 0.0, 0.0)
  # missing inputs are reported, not swallowed silently
  bad = P.Pipeline('pose').add_processor(proc).run(P.PipelineData())
  assert bad.errors and bad.pose is None


def test_intrinsics_section_format(tmp_path):
  f = tmp_path / 'zed.conf'
  f.write_text('[LEFT_CAM_2K]\nfx=1\nfy=2\ncx=3\ncy=4\n[LEFT_CAM_FHD1200]\nfx=1400.5\nfy=1401.25\ncx=960\ncy=6.0e2\nk1=0\n')
  K = P.load_camera_intrinsics(str(f))
  assert np.array_equal(K, np.array([[1400.5, 0, 960], [0, 1401.25, 600], [0, 0, 1]]))
  assert P.load_camera_intrinsics(str(f), 'LEFT_CAM_2K')[0, 0] == 1
  with pytest.raises(ValueError):
    P.load_camera_intrinsics(str(f), 'NOPE')


def test_obj_reader_and_euler(tmp_path):
  f = tmp_path / 'quad.obj'
  f.write_text('v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\n')
  m = P.load_obj(str(f))
  assert m.vertices.shape == (4, 3) and m.faces.tolist() == [[0, 1, 2], [0, 2, 3]]
  assert np.allclose(m.vertex_normals, [[0, 0, 1]] * 4) and m.visual.vertex_colors.shape == (4, 4)
  a = 0.3
  Rz = np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]])
  T = np.eye(4); T[:3, :3] = Rz; T[:3, 3] = [1, 2, 3]
  x, y, z, roll, pitch, yaw = P.pose_to_6d(T)
  assert (x, y, z) == (1, 2, 3) and abs(yaw - a) < 1e-12 and abs(roll) < 1e-12 and abs(pitch) < 1e-12


def test_bench_reference_arm_prints_one_json_line():
  """bench.py contract: exactly ONE JSON line on stdout (library chatter goes to stderr); the reference arm times the CPU
  port on a bounded sample and carries impl / cpu_baseline / e2e."""
  import json, os, subprocess, sys
  root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
  r = subprocess.run([sys.executable, os.path.join(root, 'bench.py'), '--impl', 'reference', '--steps', '1', '--warmup', '0',
                      '--ref-hyps', '2'], capture_output=True, text=True, cwd=root, timeout=600)
  assert r.returncode == 0, r.stderr[-1500:]
  lines = [l for l in r.stdout.splitlines() if l.strip()]
  assert len(lines) == 1
  d = json.loads(lines[0])
  assert d['impl'] == 'reference' and d['unit'] == 'hypotheses/s' and d['higher_is_better'] is True and d['value'] > 0
  assert d['cpu_baseline']['kind'] == 'port' and d['cpu_baseline']['cores'] >= 1 and d['cpu_baseline']['value'] == d['value']
  assert d['e2e']['value'] == d['value'] and d['e2e']['h2d_bytes_per_step'] == 0 and d['e2e']['d2h_bytes_per_step'] == 0
