"""Data pipeline: ``Generator`` (Keras ``Sequence`` equivalent), the train/val split, and a
synthetic crack dataset for boxes without the (unshipped) images.

Reference behaviour reproduced (``client_fit_model.py:19-43,54-90``):

* batch x = BGR->RGB, ``cv2.resize`` to (W,H) bilinear, ``/255``; batch y = mask read
  unchanged, resized, ``>0`` -> {0,1}, trailing channel axis; ``len = len(y)//batch``
  (remainder dropped); no augmentation;
* file lists: sorted ``*.jpg`` of the image and mask directories, both shuffled with
  ``random.Random(1337)``; client split = first 6213 train / rest val, the centralised
  script (``test/Segmentation.py:84-90``) uses last 6213 as val / rest train (quirk Q10).

Differences: batches are returned as uint8 images + uint8 masks (NHWC) so the ``/255``
and the float cast are fused into the first conv's load on the device;
``as_float=True`` gives the reference's float arrays.  Decoding can run in worker
threads with prefetch (the reference decodes 32 JPEGs per step synchronously inside
``fit``: SURVEY §3.3).
"""
from __future__ import annotations

import os
import random
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional, Sequence, Tuple

import numpy as np


class Generator:
    """``Generator(keras.utils.Sequence)`` of ``client_fit_model.py:19-43``."""

    _cv_threads_set = False

    def __init__(self, input_img_paths: Sequence[str], target_img_paths: Sequence[str], batch_size: int,
                 img_size: Tuple[int, int], as_float: bool = False, workers: int = 0):
        self.x, self.y = list(input_img_paths), list(target_img_paths)
        self.batch_size, self.img_size, self.as_float = batch_size, tuple(img_size), as_float
        self.pool = ThreadPoolExecutor(workers) if workers > 0 else None

    def __len__(self) -> int:
        return len(self.y) // self.batch_size

    @property
    def num_samples(self) -> int:
        return len(self) * self.batch_size

    def _load_x(self, fn: str) -> np.ndarray:
        import cv2
        if not Generator._cv_threads_set:  # batches are decoded in parallel by the caller's threads: no nested cv2 pool
            cv2.setNumThreads(1)
            Generator._cv_threads_set = True
        img = cv2.imread(fn, -1)
        if img is None:
            raise FileNotFoundError(fn)
        if img.ndim == 2:
            img = cv2.cvtColor(img, cv2.COLOR_GRAY2BGR)
        return cv2.resize(cv2.cvtColor(img, cv2.COLOR_BGR2RGB), (self.img_size[1], self.img_size[0]))

    def _load_y(self, fn: str) -> np.ndarray:
        import cv2
        m = cv2.imread(fn, -1)
        if m is None:
            raise FileNotFoundError(fn)
        if m.ndim == 3:
            m = m[..., 0]
        return (cv2.resize(m, (self.img_size[1], self.img_size[0])) > 0).astype(np.uint8)

    def __getitem__(self, idx: int):
        bx = self.x[idx * self.batch_size:(idx + 1) * self.batch_size]
        by = self.y[idx * self.batch_size:(idx + 1) * self.batch_size]
        mp = self.pool.map if self.pool else map
        x = np.stack(list(mp(self._load_x, bx)))
        y = np.stack(list(mp(self._load_y, by)))[..., None]
        if self.as_float:
            return x / 255, y / 1
        return x, y

    def images_only(self, idx: int) -> np.ndarray:
        """Predict path: the reference passes dummy mask paths it never reads the result of
        (``client_fit_model.py:201-207``)."""
        bx = self.x[idx * self.batch_size:(idx + 1) * self.batch_size]
        return np.stack([self._load_x(f) for f in bx])


def list_jpgs(d: str) -> List[str]:
    return sorted(os.path.join(d, f) for f in os.listdir(d) if f.endswith(".jpg") and not f.startswith("."))


def split_paths(images: List[str], masks: List[str], count: int = 6213, seed: int = 1337, mode: str = "client"):
    """Returns (train_x, train_y, val_x, val_y).  See module docstring for the two modes."""
    images, masks = list(images), list(masks)
    if len(images) != len(masks):
        # the two same-seed shuffles only keep pairs aligned for lists of equal length (the reference would train on
        # mismatched pairs without a word)
        raise ValueError(f"{len(images)} images but {len(masks)} masks")
    random.Random(seed).shuffle(images)
    random.Random(seed).shuffle(masks)
    if mode == "client":  # client_fit_model.py:78-81
        return images[:count], masks[:count], images[count:], masks[count:]
    if mode == "script":  # test/Segmentation.py:87-90
        return images[:-count], masks[:-count], images[-count:], masks[-count:]
    raise ValueError(mode)


def gen_train_val_data(image_dir: str, mask_dir: str, batch_size: int = 16, img_size=(128, 128), count: int = 6213,
                       seed: int = 1337, mode: str = "client", workers: int = 0):
    """``learning_fit.gen_train_val_data`` (client_fit_model.py:54-90) with the paths as arguments."""
    tx, ty, vx, vy = split_paths(list_jpgs(image_dir), list_jpgs(mask_dir), count, seed, mode)
    return (Generator(tx, ty, batch_size, img_size, workers=workers),
            Generator(vx, vy, batch_size, img_size, workers=workers))


# ---------------------------------------------------------------------------------------
def synth_crack_batch(n: int, h: int, w: int, seed: int) -> Tuple[np.ndarray, np.ndarray]:
    """Synthetic concrete-like RGB images with polyline "cracks" and their masks.

    Pure arithmetic (random-walk polylines rasterised by distance to segment); the device
    version of the same recipe is ``csrc/synth.cu``.  Returns uint8 (n,h,w,3), (n,h,w,1).
    """
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    imgs = np.empty((n, h, w, 3), np.uint8)
    masks = np.zeros((n, h, w, 1), np.uint8)
    for i in range(n):
        base = rng.uniform(110, 190)
        tex = rng.normal(0, 12, (h, w, 1)).astype(np.float32) + rng.normal(0, 4, (h, w, 3)).astype(np.float32)
        img = base + tex
        m = np.zeros((h, w), bool)
        if rng.random() > 0.15:  # ~15 % negatives
            for _ in range(rng.integers(1, 3)):
                px, py = rng.uniform(0, w), rng.uniform(0, h)
                ang = rng.uniform(0, 2 * np.pi)
                width = rng.uniform(0.8, 2.5) * max(h, w) / 128.0
                for _seg in range(rng.integers(4, 10)):
                    ang += rng.normal(0, 0.5)
                    ln = rng.uniform(0.08, 0.22) * max(h, w)
                    qx, qy = px + ln * np.cos(ang), py + ln * np.sin(ang)
                    dx, dy = qx - px, qy - py
                    t = np.clip(((xx - px) * dx + (yy - py) * dy) / (dx * dx + dy * dy + 1e-6), 0, 1)
                    d2 = (xx - (px + t * dx)) ** 2 + (yy - (py + t * dy)) ** 2
                    m |= d2 <= width * width
                    px, py = qx, qy
        img[m] = img[m] * 0.35
        imgs[i] = np.clip(img, 0, 255).astype(np.uint8)
        masks[i, ..., 0] = m
    return imgs, masks


class SyntheticCrackDataset:
    """Sequence-like dataset of synthetic batches, deterministic per (seed, idx); cached."""

    def __init__(self, num_samples: int, batch_size: int, img_size: Tuple[int, int], seed: int = 0,
                 cache: bool = True):
        self.n, self.batch_size, self.img_size, self.seed = num_samples, batch_size, tuple(img_size), seed
        self._cache = {} if cache else None
        self._lock = threading.Lock()

    def __len__(self) -> int:
        return self.n // self.batch_size

    @property
    def num_samples(self) -> int:
        return len(self) * self.batch_size

    def __getitem__(self, idx: int):
        if idx < 0 or idx >= len(self):
            raise IndexError(idx)
        if self._cache is not None and idx in self._cache:
            return self._cache[idx]
        b = synth_crack_batch(self.batch_size, self.img_size[0], self.img_size[1], self.seed * 1_000_003 + idx)
        if self._cache is not None:
            with self._lock:
                self._cache[idx] = b
        return b

    def images_only(self, idx: int) -> np.ndarray:
        return self[idx][0]


def make_datasets(cfg, client_index: int = 0):
    """Train/val datasets from a ``Config``: real directories when given, synthetic otherwise."""
    d, t = cfg.data, cfg.train
    if d.train_image_dir and d.train_mask_dir and not d.synthetic:
        return gen_train_val_data(d.train_image_dir, d.train_mask_dir, t.batch_size, t.image_size, d.split_count,
                                  d.split_seed, d.split_mode, d.workers)
    backend = t.backend
    if backend == "auto":
        try:
            import torch
            backend = "b200" if torch.cuda.is_available() else "torch_ref"
        except Exception:
            backend = "torch_ref"
    if backend == "b200":  # generate on the device (csrc/synth.cu): no host rasterisation, no H2D copies
        from .synthetic_device import DeviceSyntheticDataset
        return (DeviceSyntheticDataset(d.synthetic_train, t.batch_size, t.image_size, seed=1000 + client_index),
                DeviceSyntheticDataset(d.synthetic_val, t.batch_size, t.image_size, seed=5000 + client_index))
    return (SyntheticCrackDataset(d.synthetic_train, t.batch_size, t.image_size, seed=1000 + client_index),
            SyntheticCrackDataset(d.synthetic_val, t.batch_size, t.image_size, seed=5000 + client_index))
