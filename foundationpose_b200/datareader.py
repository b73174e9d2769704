"""Input side of `track_one`: the YcbInEOAT-style directory reader of the reference
(src/datareader.py:85-180: rgb/*.png, depth/*.png in millimetres, masks/*.png, cam_K.txt, annotated_poses/*) and a
pinned, double-buffered host->device frame streamer so that decoding + the copy of frame i+1 overlap the tracking of
frame i on the GPU.  Decoding uses cv2 (present in the image); nothing here touches the CUDA kernels."""
import glob
import logging
import os
import threading

import numpy as np

from .utils import depth2xyzmap


def _cv2():
  import cv2
  return cv2


class YcbineoatReader:
  """Same constructor and accessors as the reference class."""

  def __init__(self, video_dir, downscale=1, shorter_side=None, zfar=np.inf):
    self.video_dir = video_dir
    self.downscale = downscale
    self.zfar = zfar
    self.color_files = sorted(glob.glob(f'{self.video_dir}/rgb/*.png'))
    if not self.color_files:
      raise FileNotFoundError(f'no rgb/*.png under {video_dir}')
    self.K = np.loadtxt(f'{video_dir}/cam_K.txt').reshape(3, 3)
    self.id_strs = [os.path.basename(f).replace('.png', '') for f in self.color_files]
    self.H, self.W = _cv2().imread(self.color_files[0]).shape[:2]
    if shorter_side is not None:
      self.downscale = shorter_side / min(self.H, self.W)
    self.H = int(self.H * self.downscale)
    self.W = int(self.W * self.downscale)
    self.K[:2] *= self.downscale
    self.gt_pose_files = sorted(glob.glob(f'{self.video_dir}/annotated_poses/*'))

  def get_video_name(self):
    return self.video_dir.rstrip('/').split('/')[-1]

  def __len__(self):
    return len(self.color_files)

  def get_gt_pose(self, i):
    try:
      return np.loadtxt(self.gt_pose_files[i]).reshape(4, 4)
    except Exception:
      logging.info('GT pose not found, return None')
      return None

  def _resize(self, img):
    if img.shape[0] == self.H and img.shape[1] == self.W:
      return img
    cv2 = _cv2()
    return cv2.resize(img, (self.W, self.H), interpolation=cv2.INTER_NEAREST)

  def get_color(self, i):
    cv2 = _cv2()
    bgr = cv2.imread(self.color_files[i], cv2.IMREAD_COLOR)       # imageio.imread(...)[..., :3] in the reference: RGB order
    return self._resize(np.ascontiguousarray(bgr[..., ::-1]))

  def get_mask(self, i):
    mask = _cv2().imread(self.color_files[i].replace('rgb', 'masks'), -1)
    if mask is None:
      raise FileNotFoundError(self.color_files[i].replace('rgb', 'masks'))
    if mask.ndim == 3:
      for c in range(3):
        if mask[..., c].sum() > 0:
          mask = mask[..., c]
          break
    return self._resize(mask).astype(bool).astype(np.uint8)

  def get_depth(self, i):
    d = _cv2().imread(self.color_files[i].replace('rgb', 'depth'), -1)
    if d is None:
      raise FileNotFoundError(self.color_files[i].replace('rgb', 'depth'))
    depth = self._resize(d / 1e3)
    depth[(depth < 0.001) | (depth >= self.zfar)] = 0
    return depth

  def get_xyz_map(self, i):
    return depth2xyzmap(self.get_depth(i), self.K)

  def get_occ_mask(self, i):
    cv2 = _cv2()
    occ = np.zeros((self.H, self.W), dtype=bool)
    for sub in ('masks_hand', 'masks_hand_right'):
      f = self.color_files[i].replace('rgb', sub)
      if os.path.exists(f):
        occ = occ | (self._resize(cv2.imread(f, -1)) > 0)
    return occ.astype(np.uint8)


class FrameStreamer:
  """Iterates `(i, rgb_cuda uint8 [H,W,3], depth_cuda float32 [H,W])` over a reader.  A worker thread decodes frame i+1
  into one of two pinned staging buffers and enqueues its host->device copy on a side stream while the caller tracks
  frame i; the consumer stream waits on the copy's event, never on the host.  `device=None` keeps everything on the
  host (CPU tests)."""

  def __init__(self, reader, start=0, stop=None, device='cuda', depth=2):
    import torch
    self.reader, self.start, self.stop = reader, start, len(reader) if stop is None else min(stop, len(reader))
    self.device = device
    self.nbuf = max(2, int(depth))
    pin = device is not None and torch.cuda.is_available()
    self._rgb_h = [torch.empty((reader.H, reader.W, 3), dtype=torch.uint8, pin_memory=pin) for _ in range(self.nbuf)]
    self._dep_h = [torch.empty((reader.H, reader.W), dtype=torch.float32, pin_memory=pin) for _ in range(self.nbuf)]
    self._rgb_d = self._dep_d = None
    self._copy_stream = None
    if device is not None:
      self._rgb_d = [torch.empty_like(t, device=device) for t in self._rgb_h]
      self._dep_d = [torch.empty_like(t, device=device) for t in self._dep_h]
      self._copy_stream = torch.cuda.Stream(device=device)
    self._free = threading.Semaphore(self.nbuf)
    self._ready = []
    self._cv = threading.Condition()
    self._err = None
    self.h2d_bytes_per_frame = reader.H * reader.W * (3 + 4)

  def _produce(self):
    import torch
    try:
      for n, i in enumerate(range(self.start, self.stop)):
        self._free.acquire()
        b = n % self.nbuf
        self._rgb_h[b].copy_(torch.from_numpy(np.ascontiguousarray(self.reader.get_color(i))))
        self._dep_h[b].copy_(torch.from_numpy(np.ascontiguousarray(self.reader.get_depth(i), dtype=np.float32)))
        ev = None
        if self.device is not None:
          with torch.cuda.stream(self._copy_stream):
            self._rgb_d[b].copy_(self._rgb_h[b], non_blocking=True)
            self._dep_d[b].copy_(self._dep_h[b], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        with self._cv:
          self._ready.append((i, b, ev))
          self._cv.notify()
    except Exception as e:   # surfaced to the consumer
      with self._cv:
        self._err = e
        self._cv.notify()
    with self._cv:
      self._ready.append(None)
      self._cv.notify()

  def __iter__(self):
    import torch
    t = threading.Thread(target=self._produce, daemon=True)
    t.start()
    held = None
    while True:
      with self._cv:
        while not self._ready and self._err is None:
          self._cv.wait()
        if self._err is not None:
          raise self._err
        item = self._ready.pop(0)
      if held is not None:            # the previous frame's buffers may be overwritten once the caller asks for the next one
        if self.device is not None:
          torch.cuda.current_stream().synchronize()
        self._free.release()
      if item is None:
        break
      i, b, ev = item
      held = b
      if self.device is not None:
        torch.cuda.current_stream().wait_event(ev)
        yield i, self._rgb_d[b], self._dep_d[b]
      else:
        yield i, self._rgb_h[b], self._dep_h[b]
    t.join()


def track_video(est, reader, mask0=None, iteration_register=5, iteration_track=2, device='cuda', start=0, stop=None):
  """run_demo-style loop of the reference family (register on the first frame, track_one afterwards) over a
  FrameStreamer; returns the [n,4,4] poses."""
  poses = []
  for i, rgb, depth in FrameStreamer(reader, start=start, stop=stop, device=device):
    if not poses:
      mask = reader.get_mask(i) if mask0 is None else mask0
      pose = est.register(K=reader.K, rgb=rgb, depth=depth, ob_mask=np.asarray(mask).astype(bool), iteration=iteration_register)
    else:
      pose = est.track_one(rgb=rgb, depth=depth, K=reader.K, iteration=iteration_track)
    poses.append(np.asarray(pose).reshape(4, 4))
  return np.stack(poses) if poses else np.zeros((0, 4, 4))
