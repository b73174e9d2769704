"""In-tree nvcc build of libfpb200.so (sm_100a only).

`python -m foundationpose_b200.build` or `build()`; the .so is git-ignored but travels to the GPU box.
Geometry translation units are compiled with -fmad=false so their fp32 arithmetic is bit-identical to the
CPU oracle; the network kernels use the default (fused) arithmetic.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
OUT = os.path.join(HERE, 'libfpb200.so')
OBJ = os.path.join(HERE, 'build')

ARCH = ['-gencode', 'arch=compute_100a,code=sm_100a']
COMMON = ['-O3', '-lineinfo', '-std=c++17', '-Xcompiler', '-fPIC,-fvisibility=hidden', '--expt-relaxed-constexpr']
# dev only (same-box A/B of compile-time variants): FPB_EXTRA_NVCC_FLAGS='-DFOO' FPB_LIB_OUT=/path/lib.so python -m foundationpose_b200.build
COMMON += os.environ.get('FPB_EXTRA_NVCC_FLAGS', '').split()
if os.environ.get('FPB_LIB_OUT'):
  OUT = os.environ['FPB_LIB_OUT']
  OBJ = OUT + '.build'
# (source, extra flags, object suffix).  The network kernel units are compiled twice: bf16 operands (namespace bf16) and,
# with -DFPB_ACT_FP16, IEEE fp16 operands (namespace f16) -- csrc/act_type.cuh.
NET_UNITS = ['net_ref.cu', 'net_tc.cu', 'net_tc2.cu', 'net_tc3.cu', 'net_stem.cu', 'net_attn.cu', 'net_ffn.cu']
UNITS = [
  ('api.cu', [], ''),
  ('depth.cu', ['-fmad=false'], ''),
  ('hypo.cu', ['-fmad=false'], ''),
  ('render.cu', ['-fmad=false'], ''),
  ('metrics.cu', ['-fmad=false'], ''),
] + [(u, [], '') for u in NET_UNITS] + [(u, ['-DFPB_ACT_FP16'], '_f16') for u in NET_UNITS] + [
  ('net.cu', [], ''),
]


def _nvcc():
  for p in (os.environ.get('NVCC'), '/usr/local/cuda/bin/nvcc', 'nvcc'):
    if p and (os.path.isabs(p) and os.path.exists(p) or not os.path.isabs(p)):
      return p
  raise RuntimeError('nvcc not found')


def _digest(paths, extra):
  h = hashlib.sha256(repr(extra).encode())
  for p in sorted(paths):
    with open(p, 'rb') as f:
      h.update(f.read())
  return h.hexdigest()


def build(force=False, verbose=False):
  os.makedirs(OBJ, exist_ok=True)
  nvcc = _nvcc()
  headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.h'))]
  headers.append(os.path.join(HERE, '..', 'include', 'fpb200.h'))
  objs = []
  rebuilt = False
  env = dict(os.environ)
  env.pop('CC', None); env.pop('CXX', None)
  jobs = []
  for name, flags, suffix in UNITS:
    src = os.path.join(CSRC, name)
    if not os.path.exists(src):
      continue
    obj = os.path.join(OBJ, name.replace('.cu', suffix + '.o'))
    stamp = obj + '.sha'
    dig = _digest([src] + headers, flags + COMMON + ARCH)
    if force or not os.path.exists(obj) or not os.path.exists(stamp) or open(stamp).read() != dig:
      cmd = [nvcc] + ARCH + COMMON + flags + ['-ccbin', '/usr/bin/g++', '-c', src, '-o', obj]
      jobs.append((name + suffix, cmd, stamp, dig))
    objs.append(obj)
  if jobs:                                    # independent translation units: compile them in parallel
    from concurrent.futures import ThreadPoolExecutor

    def run(job):
      name, cmd, stamp, dig = job
      if verbose:
        print(' '.join(cmd), flush=True)
      r = subprocess.run(cmd, capture_output=True, text=True, env=env)
      if r.returncode != 0:
        raise RuntimeError(f'nvcc failed for {name}:\n{r.stdout}\n{r.stderr}')
      with open(stamp, 'w') as f:
        f.write(dig)
    with ThreadPoolExecutor(max_workers=min(len(jobs), max(1, (os.cpu_count() or 2) - 1))) as ex:
      list(ex.map(run, jobs))
    rebuilt = True
  if rebuilt or force or not os.path.exists(OUT):
    cmd = [nvcc] + ARCH + ['-shared', '-ccbin', '/usr/bin/g++', '-o', OUT] + objs + ['-cudart', 'static']
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if r.returncode != 0:
      raise RuntimeError(f'link failed:\n{r.stdout}\n{r.stderr}')
  return OUT


if __name__ == '__main__':
  print(build(force='--force' in sys.argv, verbose=True))
