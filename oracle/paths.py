"""Whole-path CPU restatements used as the checker in tests and as the timed CPU baseline in bench.py.
TEST INFRASTRUCTURE ONLY (see oracle.py header).  Each function strings together the per-stage restatements in the
order the reference runs them (ref: apps/model-runner/runtime_deployment.py:270-292 for the bioimageio order,
apps/fibsem-mito-analysis/analysis_deployment.py:241-252 for the fibsem order)."""
from __future__ import annotations

import time
from typing import Callable, Dict, Optional, Sequence, Tuple

import numpy as np

from . import oracle as O


def c2_pre(img: np.ndarray, tile: int, overlap: int, p_lo: float = 1.0, p_hi: float = 99.8) -> np.ndarray:
    """(1,1,*spatial) image → scale_range over the spatial axes → fibsem-convention tiles (T,1,*tile)."""
    norm = O.scale_range(img, p_lo, p_hi, axes=None)
    return O.fibsem_tiles(norm[0, 0], tile, overlap)[:, None]


def c2_post(logits: np.ndarray, spatial: Sequence[int], tile: int, overlap: int, connectivity: Optional[int] = None):
    """tile logits (T,1,*tile) → Gaussian blend → sigmoid → binarize 0.5 → connected components."""
    blended = O.fibsem_blend(logits[:, 0], spatial, tile, overlap)
    prob = O.sigmoid(blended)
    mask = O.binarize(prob, 0.5)
    return prob, mask, O.label(mask, connectivity)


def crop_pre(img: np.ndarray, block: Sequence[int], halo: Sequence[int], p_lo: float = 1.0, p_hi: float = 99.8):
    """bioimageio order: whole-sample scale_range, then halo blocks (Appendix A.1/A.4)."""
    return O.blocks_tensor(O.scale_range(img, p_lo, p_hi, axes=tuple(range(1, img.ndim))), block, halo)


def crop_post(logits: np.ndarray, shape: Sequence[int], block: Sequence[int], halo: Sequence[int],
              connectivity: Optional[int] = None):
    out = O.from_blocks(logits, shape, block, halo)
    prob = O.sigmoid(out)
    mask = O.binarize(prob, 0.5)
    labels = np.stack([O.label(mask[n, 0], connectivity) for n in range(shape[0])])
    return prob, mask, labels


def timed_c2(img: np.ndarray, logits: np.ndarray, tile: int, overlap: int,
             keep: Optional[dict] = None) -> Tuple[float, Dict[str, float]]:
    """Seconds for pre + post on one image (forward excluded: the logits are given).  `keep`, if given, receives the
    tiles and labels the pass produced (so a caller can compare them with another implementation's)."""
    t0 = time.perf_counter()
    tiles = c2_pre(img, tile, overlap)
    t1 = time.perf_counter()
    _, mask, labels = c2_post(logits, img.shape[2:], tile, overlap)
    t2 = time.perf_counter()
    if keep is not None:
        keep.update(tiles=tiles, mask=mask, labels=labels)
    return t2 - t0, {"pre_s": t1 - t0, "post_s": t2 - t1}
