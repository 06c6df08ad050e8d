"""Host-side tile geometry: the small per-axis index tables the gather / stitch kernels consume.

Two block conventions exist on the reference path (SURVEY.md §8 "tile convention"):
  * bioimageio.core blocking — block length includes the halo, inner = block - 2*halo, the last inner slice
    is shifted back inside the sample, missing halo is reflect-padded, reassembly copies inner regions and the
    last writer wins (Appendix A.4; `predict_sample_with_blocking`, ref: apps/model-runner/runtime_deployment.py:278);
  * fibsem overlap tiling — tiles at range(0, S, tile - overlap), bottom/right reflect padding relative to the
    valid part of the tile, Gaussian-weighted blending (ref: apps/fibsem-mito-analysis/analysis_deployment.py:120-153).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import numpy as np


def reflect_index(c: np.ndarray, v0: int, v1: int) -> np.ndarray:
    """Map coordinates c into [v0, v1) the way `np.pad(mode="reflect")` extends that range (edge not repeated,
    any number of bounces): with n = v1 - v0, r = (c - v0) mod 2(n-1), r >= n → 2(n-1) - r."""
    c = np.asarray(c, dtype=np.int64)
    n = v1 - v0
    if n <= 0:
        raise ValueError("empty valid range")
    if n == 1:
        return np.full_like(c, v0)
    period = 2 * (n - 1)
    r = np.mod(c - v0, period)
    r = np.where(r >= n, period - r, r)
    return r + v0


@dataclass(frozen=True)
class AxisBlocks:
    """Blocks of one axis: `src[k*B + i]` = sample coordinate read by offset i of block k;
    `sel[p]` = k*B + i of the block element that ends up at output coordinate p."""

    S: int
    B: int
    T: int
    src: np.ndarray  # int32 [T*B]
    sel: np.ndarray  # int32 [S] (bioimageio convention only; empty for fibsem)
    stride: int      # fibsem convention only (0 for bioimageio)


def bioimageio_axis(S: int, B: int, halo: int) -> AxisBlocks:
    """Appendix A.4: inner slices [(min(p, S-inner), min(p+inner, S)) for p in range(0, S, inner)]."""
    inner = B - 2 * halo
    if inner <= 0:
        raise ValueError(f"halo {halo} leaves no inner region in a block of {B}")
    if S < inner:
        raise ValueError(f"sample axis of {S} is shorter than the inner block {inner}")
    starts = [min(p, S - inner) for p in range(0, S, inner)]
    T = len(starts)
    i = np.arange(B, dtype=np.int64)
    src = np.empty(T * B, dtype=np.int32)
    sel = np.empty(S, dtype=np.int32)
    for k, p in enumerate(starts):
        src[k * B:(k + 1) * B] = reflect_index(p - halo + i, 0, S)
        # ascending k, so a later (shifted) block overwrites its predecessor: last writer wins
        sel[p:p + inner] = k * B + halo + np.arange(inner)
    return AxisBlocks(S, B, T, src, sel, 0)


def fibsem_axis(S: int, B: int, overlap: int) -> AxisBlocks:
    """ref: analysis_deployment.py:121,132-138 — tile k spans [k*stride, min(k*stride + B, S)); the part beyond the
    sample is `np.pad(tile, (0, B - th), mode="reflect")` of that valid span."""
    stride = B - overlap
    if stride <= 0:
        raise ValueError(f"overlap {overlap} must be smaller than the tile size {B}")
    starts = list(range(0, S, stride))
    T = len(starts)
    i = np.arange(B, dtype=np.int64)
    src = np.empty(T * B, dtype=np.int32)
    for k, o in enumerate(starts):
        src[k * B:(k + 1) * B] = reflect_index(o + i, o, min(o + B, S))
    return AxisBlocks(S, B, T, src, np.empty(0, np.int32), stride)


def identity_axis(S: int) -> AxisBlocks:
    """An axis that is not tiled (e.g. Z of a 2-D image): one block covering it."""
    a = np.arange(S, dtype=np.int32)
    return AxisBlocks(S, S, 1, a.copy(), a.copy(), S)


def gauss_window_1d(B: int) -> np.ndarray:
    """ref: analysis_deployment.py:125-127 — exp(-2 * linspace(-1, 1, B)**2), float64 (the fp32 cast happens after
    the outer product, inside the blend kernel)."""
    t = np.linspace(-1, 1, B)
    return np.exp(-2 * t**2)
