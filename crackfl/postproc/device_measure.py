"""Device half of the crack measurement (kernel K17): per-image foreground area, boundary-pixel count and component
seeds straight from the probability map, without copying masks to the host.  The exact contour/arc-length numbers of
``test/Segmentation2.py:118-124`` come from ``postproc/measure.py`` (OpenCV); these are the triage statistics used to
decide which images are worth the CPU pass (no crack -> area 0)."""
from __future__ import annotations

import numpy as np
import torch

from ..ops import lib as L


def measure_masks_device(prob: torch.Tensor, threshold: int = 127) -> torch.Tensor:
    """prob: (B,H,W) fp32 on the GPU -> (B,4) [area_px, boundary_px, seeds, prob_sum]."""
    assert prob.is_cuda and prob.dtype == torch.float32 and prob.dim() == 3
    prob = prob.contiguous()
    out = torch.empty((prob.shape[0], 4), dtype=torch.float32, device=prob.device)
    with torch.cuda.device(prob.device):
        L.check(L.get_lib().crackfl_measure_masks(prob.data_ptr(), out.data_ptr(), prob.shape[0], prob.shape[1],
                                                  prob.shape[2], threshold, torch.cuda.current_stream().cuda_stream))
    return out


def measure_masks_reference(prob: np.ndarray, threshold: int = 127) -> np.ndarray:
    """NumPy restatement used as the test oracle."""
    out = np.zeros((prob.shape[0], 4), np.float64)
    for b, p in enumerate(prob):
        fg = (p * 255.0).astype(np.int32) > threshold
        pad = np.pad(fg, 1)
        n, s, w, e = pad[:-2, 1:-1], pad[2:, 1:-1], pad[1:-1, :-2], pad[1:-1, 2:]
        nw, ne = pad[:-2, :-2], pad[:-2, 2:]
        out[b, 0] = fg.sum()
        out[b, 1] = (fg & ~(n & s & w & e)).sum()
        out[b, 2] = (fg & ~n & ~w & ~nw & ~ne).sum()
        out[b, 3] = p.sum()
    return out
