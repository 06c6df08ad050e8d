"""Model screening: rank candidate models by mAP over a labelled image set, images sharded over the GPUs of one box.

The reference has no screening code of its own (SURVEY.md headline finding 4): in-tree it is repeated `infer()` calls
followed by `average_precision` and `np.nanmean` (ref: apps/cellpose-finetuning/main.py:2002-2008).  Here every rank
evaluates all models on its image shard (i % world == rank) entirely on its GPU and contributes integer TP / FP / FN
rows to a zero-initialised table; one all-reduce (sum ≡ gather, integer-exact) merges the shards and every rank computes
the same float32 AP, nanmean and ranking the single-process reference would.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, kernels as K
from .metrics import ap_from_counts, counts_to_tp, match_counts_device


def shard_indices(n_images: int, rank: int, world: int) -> List[int]:
    return list(range(rank, n_images, world))


def merge_and_rank(local_table: torch.Tensor, thresholds: Sequence[float]) -> Dict[str, np.ndarray]:
    """local_table: int64 [n_models, n_images, n_thr + 2] = per threshold tp, then n_true, n_pred; rows of images this
    rank did not process are zero.  All-reduces over the default process group (if initialised) and returns
    ap [n_models, n_images, n_thr] float32, mean_ap [n_models, n_thr], ranking (best first by mean AP at thresholds[0])."""
    table = local_table.clone()
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(table, op=dist.ReduceOp.SUM)
    t = table.cpu().numpy()
    n_thr = len(thresholds)
    n_models = t.shape[0]
    ap = np.zeros(t.shape[:2] + (n_thr,), np.float32)
    for m in range(n_models):
        ap[m], _, _, _ = ap_from_counts(t[m, :, :n_thr], t[m, :, n_thr], t[m, :, n_thr + 1])
    with np.errstate(invalid="ignore"):
        mean_ap = np.nanmean(ap, axis=1)  # ref: main.py:2008
    order = np.argsort(-np.nan_to_num(mean_ap[:, 0], nan=-1.0), kind="stable")
    return {"ap": ap, "mean_ap": mean_ap, "ranking": order, "table": t}


@torch.no_grad()
def screen_models(models: Sequence[Callable[[torch.Tensor], torch.Tensor]], images: Sequence[np.ndarray],
                  gt_labels: Sequence[np.ndarray], thresholds=(0.5, 0.75, 0.9), device="cuda:0", batch: int = 32,
                  rank: Optional[int] = None, world: Optional[int] = None, eps: float = 1e-6,
                  connectivity: Optional[int] = None) -> Dict[str, np.ndarray]:
    """images: list of (H, W) float arrays of one shape; gt_labels: matching int label images.
    Per (model, image): zero_mean_unit_variance → forward (untiled) → sigmoid → binarize 0.5 → CC labels → AP vs GT."""
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
        world = dist.get_world_size() if dist.is_initialized() else 1
    dev = torch.device(device)
    n_images, n_thr = len(images), len(thresholds)
    thr = torch.tensor([float(t) for t in thresholds], dtype=torch.float64, device=dev)
    table = torch.zeros((len(models), n_images, n_thr + 2), dtype=torch.int64, device=dev)
    mine = shard_indices(n_images, rank, world)
    ws = K.Workspace()
    for s in range(0, len(mine), batch):
        idx = mine[s:s + batch]
        x = torch.from_numpy(np.stack([np.asarray(images[i], np.float32) for i in idx])[:, None]).to(dev)
        gt = torch.from_numpy(np.stack([np.asarray(gt_labels[i], np.int32) for i in idx])).to(dev)
        cap_true = int(gt.max().item()) + 1
        _, norm = K.meanvar(x, len(idx), eps, ws=ws)
        xn = K.normalize(x, norm)
        for m, model in enumerate(models):
            logits = model(xn).contiguous()
            mask = K.binarize(K.sigmoid(logits), 0.5)
            labels, n_lab = K.ccl(mask[:, 0], connectivity, ws=ws)
            cap_pred = int(n_lab.max().item()) + 1
            counts, max_lbl, tab = match_counts_device(gt, labels, thr, cap_true, cap_pred)
            tp = counts_to_tp(counts.cpu().numpy(), thresholds, tab)
            rows = torch.tensor(idx, device=dev)
            table[m, rows, :n_thr] = torch.from_numpy(tp).to(dev)
            table[m, rows, n_thr:] = max_lbl.to(torch.int64)
    return merge_and_rank(table, thresholds)
