"""Instance-segmentation metric on the device: `average_precision(masks_true, masks_pred, threshold)`.

Drop-in for the `cellpose.metrics.average_precision` call at ref: apps/cellpose-finetuning/main.py:2002-2006 — same
arguments, same `(ap, tp, fp, fn)` float32 arrays of shape (n_images, n_thresholds).  The label-pair overlap histogram,
the areas, the IoU >= threshold test and the matching count all run in libbe_b200.so; the host only turns the integer
counts into the float32 ratios exactly as cellpose does.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple, Union

import numpy as np
import torch

from . import kernels as K

ArrayLike = Union[np.ndarray, torch.Tensor]


def _next_pow2(v: int) -> int:
    return 1 << max(int(v) - 1, 1).bit_length()


def _to_device_i32(a: ArrayLike, device) -> torch.Tensor:
    t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
    return t.to(device=device, dtype=torch.int32, non_blocking=True).contiguous()


def match_counts_device(lab_true: torch.Tensor, lab_pred: torch.Tensor, thr: torch.Tensor, cap_true: int,
                        cap_pred: int, capacity: int = 0) -> Tuple[torch.Tensor, torch.Tensor, dict]:
    """(N, ...) int32 CUDA label stacks → (counts int64 [N, n_thr, 2] {pairs, extra}, max_lbl int32 [N, 2], table).
    Grows the hash table and retries if any image overflowed it; raises on out-of-range labels."""
    capacity = capacity or _next_pow2(4 * (cap_true + cap_pred) + 64)
    while True:
        table = K.label_overlap(lab_true, lab_pred, capacity, cap_true, cap_pred)
        status = table["status"].cpu().numpy()
        if (status == 2).any():
            raise ValueError("label maps contain negative labels or labels beyond the declared capacity")
        if not (status == 1).any():
            break
        capacity *= 4
        if capacity > _next_pow2(2 * lab_true[0].numel()) * 2:
            raise RuntimeError("label-pair table overflow")  # cannot happen: pairs <= pixels
    return K.match_counts(table, thr), table["max_lbl"], table


def _assignment_tp(table: dict, i: int, th: float) -> int:
    """Exact cellpose `_true_positive` for one image from the device-built sparse table — only needed for thresholds
    below 0.5 when some label has several partners."""
    from scipy.optimize import linear_sum_assignment

    keys = table["keys"][i].cpu().numpy().astype(np.uint64)
    cnt = table["counts"][i].cpu().numpy()
    used = keys != np.uint64(0xFFFFFFFFFFFFFFFF)
    t = (keys[used] >> np.uint64(32)).astype(np.int64)
    p = (keys[used] & np.uint64(0xFFFFFFFF)).astype(np.int64)
    ml = table["max_lbl"][i].cpu().numpy()
    nt, npred = int(ml[0]) + 1, int(ml[1]) + 1
    overlap = np.zeros((nt, npred), np.float64)   # the table holds the pairs with both labels > 0 ...
    overlap[t, p] = cnt[used]
    # ... and the areas are the row / column sums of the full matrix (background overlaps included)
    n_true = table["area_true"][i].cpu().numpy()[:nt].astype(np.float64)[:, None]
    n_pred = table["area_pred"][i].cpu().numpy()[:npred].astype(np.float64)[None, :]
    with np.errstate(invalid="ignore", divide="ignore"):
        iou = overlap / (n_pred + n_true - overlap)
    iou[np.isnan(iou)] = 0.0
    iou = iou[1:, 1:]
    if iou.size == 0:
        return 0
    n_min = min(iou.shape)
    costs = -(iou >= th).astype(float) - iou / (2 * n_min)
    ti, pi = linear_sum_assignment(costs)
    return int((iou[ti, pi] >= th).sum())


def counts_to_tp(counts: np.ndarray, thresholds: Sequence[float], table: dict = None) -> np.ndarray:
    """{pairs, extra} → tp per (image, threshold).  Exact for thr >= 0.5; below 0.5 images with extra partners are solved
    by assignment on the host."""
    tp = counts[..., 0] - counts[..., 1]
    for k, th in enumerate(thresholds):
        if th < 0.5:
            for i in np.nonzero(counts[:, k, 1] > 0)[0]:
                if table is None:
                    raise ValueError("thresholds below 0.5 need the overlap table for the assignment step")
                tp[i, k] = _assignment_tp(table, int(i), float(th))
    return tp


def ap_from_counts(tp: np.ndarray, n_true: np.ndarray, n_pred: np.ndarray):
    """cellpose's last lines: float32 tp/fp/fn and ap = tp / (tp + fp + fn) (0/0 → NaN)."""
    tp = tp.astype(np.float32)
    tp = np.where(n_pred[:, None] > 0, tp, np.float32(0))
    fp = (n_pred[:, None] - tp).astype(np.float32)
    fn = (n_true[:, None] - tp).astype(np.float32)
    with np.errstate(invalid="ignore", divide="ignore"):
        ap = (tp / (tp + fp + fn)).astype(np.float32)
    return ap, tp.astype(np.float32), fp, fn


def average_precision(masks_true, masks_pred, threshold=(0.5, 0.75, 0.9), device="cuda:0"):
    """Same contract as cellpose.metrics.average_precision (lists of integer label images, or single arrays)."""
    not_list = False
    if not isinstance(masks_true, (list, tuple)):
        masks_true, masks_pred, not_list = [masks_true], [masks_pred], True
    if not isinstance(threshold, (list, tuple, np.ndarray)):
        threshold = [threshold]
    if len(masks_true) != len(masks_pred):
        raise ValueError("metrics.average_precision requires len(masks_true)==len(masks_pred)")
    dev = torch.device(device)
    n, m = len(masks_true), len(threshold)
    thr = torch.tensor([float(t) for t in threshold], dtype=torch.float64, device=dev)
    tp = np.zeros((n, m), np.int64)
    n_true = np.zeros(n, np.int64)
    n_pred = np.zeros(n, np.int64)
    # batch images of equal shape so one launch covers many of them
    by_shape = {}
    for i, (a, b) in enumerate(zip(masks_true, masks_pred)):
        if tuple(a.shape) != tuple(b.shape):
            raise ValueError(f"image {i}: true/pred shapes differ")
        by_shape.setdefault(tuple(a.shape), []).append(i)
    for shape, idx in by_shape.items():
        for s in range(0, len(idx), 64):
            chunk = idx[s:s + 64]
            t = torch.stack([_to_device_i32(masks_true[i], dev) for i in chunk])
            p = torch.stack([_to_device_i32(masks_pred[i], dev) for i in chunk])
            mt, mp = int(t.max().item()), int(p.max().item())
            counts, max_lbl, table = match_counts_device(t, p, thr, mt + 1, mp + 1)
            tp[chunk] = counts_to_tp(counts.cpu().numpy(), threshold, table)
            ml = max_lbl.cpu().numpy()
            n_true[chunk], n_pred[chunk] = ml[:, 0], ml[:, 1]
    ap, tpf, fp, fn = ap_from_counts(tp, n_true, n_pred)
    if not_list:
        return ap[0], tpf[0], fp[0], fn[0]
    return ap, tpf, fp, fn


def _safe_div(n: float, d: float) -> float:
    """Return n / d, or 0.0 when d is zero (ref: apps/cellpose-finetuning/main.py:1218-1220)."""
    return float(n / d) if d != 0 else 0.0


def binary_metrics(pred: ArrayLike, target: ArrayLike, device="cuda:0") -> dict:
    """Pixel-level foreground metrics of ref: apps/cellpose-finetuning/main.py:1225-1270 (`_compute_binary_metrics`): adaptive
    thresholds (0.5 if the values lie in [0, 1], else 0.0), TP / FP / FN / TN counted on the device, then the same ratios."""
    dev = torch.device(device)

    def dev_f32(a):
        t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a))
        return t.to(device=dev, dtype=torch.float32).contiguous()

    counts, _ = K.binary_confusion(dev_f32(pred), dev_f32(target))
    tp, fp, fn, tn = (int(v) for v in counts.cpu().tolist())
    precision = _safe_div(tp, tp + fp)
    recall = _safe_div(tp, tp + fn)
    return {
        "tp": tp, "fp": fp, "fn": fn, "tn": tn,
        "pixel_accuracy": _safe_div(tp + tn, tp + tn + fp + fn),
        "precision": precision,
        "recall": recall,
        "f1": _safe_div(2 * precision * recall, precision + recall),
        "iou": _safe_div(tp, tp + fp + fn),
    }
