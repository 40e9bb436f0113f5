"""Synthetic crack batches generated directly on the GPU (kernel K15, ``csrc/synth.cu``)."""
from __future__ import annotations

import torch

from ..ops import lib as L


def synth_crack_batch_device(batch: int, h: int, w: int, seed: int, device=None):
    """Returns (images uint8 (B,H,W,3), masks uint8 (B,H,W)) on ``device``; no host round trip."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    img = torch.empty((batch, h, w, 3), dtype=torch.uint8, device=dev)
    mask = torch.empty((batch, h, w), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        L.check(L.get_lib().crackfl_synth_cracks(img.data_ptr(), mask.data_ptr(), batch, h, w, seed & 0xFFFFFFFF,
                                                 torch.cuda.current_stream().cuda_stream), "synth_cracks")
    return img, mask


class DeviceSyntheticDataset:
    """Sequence-like dataset whose batches live on the device (for benchmarks and the on-box runtime)."""

    device_resident = True  # nothing to decode or copy: the trainer indexes it inline, no prefetch thread

    def __init__(self, num_samples: int, batch_size: int, img_size, seed: int = 0, device=None):
        self.n, self.batch_size, self.img_size, self.seed, self.device = num_samples, batch_size, tuple(img_size), seed, device
        self._cache = {}

    def __len__(self):
        return self.n // self.batch_size

    @property
    def num_samples(self):
        return len(self) * self.batch_size

    def __getitem__(self, idx: int):
        if idx not in self._cache:
            self._cache[idx] = synth_crack_batch_device(self.batch_size, self.img_size[0], self.img_size[1],
                                                        self.seed * 1_000_003 + idx, self.device)
        return self._cache[idx]
