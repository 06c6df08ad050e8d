"""Model screening as a BioEngine app: an entry deployment that fans a labelled image set out over GPU replicas and merges
their integer tables, and the per-GPU worker deployment ("an 8-GPU screening app manifest").

The reference has no screening service; in-tree the same result takes `len(models) * len(images)` calls of
`infer()` followed by `cellpose.metrics.average_precision` and `np.nanmean` on one CPU core (ref:
apps/cellpose-finetuning/main.py:2002-2008).  The composition follows the reference's own two-deployment pattern
(ref: apps/model-runner/manifest.yaml:18-20, bioengine/apps/builder.py:1475-1508): the first class carries the schema
method and receives the second as a `DeploymentHandle`; work reaches a replica as `handle.score_shard.remote(...)`.

Every replica scores the images `i % n_shards == shard` for all candidate models on its GPU (`screening.screen_models`:
zero_mean_unit_variance → forward → sigmoid → binarize 0.5 → connected components → label overlap → IoU matching, a
device loop without host round trips) and returns an int64 table [models, images, thresholds + 2] = TP per threshold,
n_true, n_pred, zero for images of other shards.  Sums of integer tables are exact, so the merged AP / mean AP / ranking
are those of a single process (and of `torchrun` ranks merging by NCCL all-reduce, which is how `bench.py` runs it).
"""
import asyncio
from typing import Dict, List, Optional, Sequence, Union

import numpy as np
import torch

from ._serve_compat import DeploymentHandle, call_handle, deployment, schema_method
from .metrics import ap_from_counts

ArrayOrSource = Union[np.ndarray, str]


def rank_from_table(table: np.ndarray, thresholds: Sequence[float]) -> Dict[str, np.ndarray]:
    """Merged int64 table → float32 AP per (model, image, threshold), nanmean over images (ref: main.py:2008), ranking by
    the mean AP at thresholds[0], best first."""
    n_thr = len(thresholds)
    ap = np.zeros(table.shape[:2] + (n_thr,), np.float32)
    for m in range(table.shape[0]):
        ap[m], _, _, _ = ap_from_counts(table[m, :, :n_thr], table[m, :, n_thr], table[m, :, n_thr + 1])
    with np.errstate(invalid="ignore"):
        mean_ap = np.nanmean(ap, axis=1)
    order = np.argsort(-np.nan_to_num(mean_ap[:, 0], nan=-1.0), kind="stable")
    return {"ap": ap, "mean_ap": mean_ap, "ranking": order, "table": table}


def _resolve(source: ArrayOrSource) -> np.ndarray:
    if isinstance(source, str):
        from .entry_deployment import EntryDeployment

        return EntryDeployment._load_image_from_source(source)
    return np.asarray(source)


@deployment(
    ray_actor_options={"num_cpus": 1, "num_gpus": 1, "memory": 12 * 1024 * 1024 * 1024},
    max_ongoing_requests=1,
    autoscaling_config={"min_replicas": 8, "initial_replicas": 8, "max_replicas": 8,
                        "target_num_ongoing_requests_per_replica": 1},
    health_check_period_s=30.0, health_check_timeout_s=30.0, graceful_shutdown_timeout_s=120.0,
    graceful_shutdown_wait_loop_s=2.0,
)
class ScreeningWorker:
    """One GPU: all candidate models on one shard of the images."""

    def __init__(self, batch: int = 32):
        from . import models

        models.strict_fp32()
        self.batch = int(batch)
        self._nets: Dict[str, torch.nn.Module] = {}

    async def check_health(self) -> None:
        if not torch.cuda.is_available():
            raise RuntimeError("GPU runtime deployment is not available. No GPU could be reached by the runtime replica.")

    def _net(self, rdf_path: str, device: torch.device) -> torch.nn.Module:
        if rdf_path not in self._nets:
            from .rdf import load_model_spec
            from .runtime_deployment import _LoadedModel

            self._nets[rdf_path] = _LoadedModel(load_model_spec(rdf_path), device, None).net
        return self._nets[rdf_path]

    async def score_shard(self, rdf_paths: List[str], images: List[ArrayOrSource], gt_labels: List[ArrayOrSource],
                          thresholds: Sequence[float] = (0.5, 0.75, 0.9), shard: int = 0, n_shards: int = 1) -> np.ndarray:
        """int64 [models, images, thresholds + 2]: rows of the images `i % n_shards == shard`, zero elsewhere."""
        from .screening import screen_models, shard_indices

        if len(images) != len(gt_labels):
            raise ValueError(f"{len(images)} images but {len(gt_labels)} ground-truth label images")
        device = torch.device("cuda", torch.cuda.current_device())
        nets = [self._net(p, device) for p in rdf_paths]
        mine = set(shard_indices(len(images), shard, n_shards))
        # only this shard's sources are fetched / decoded; the others keep a placeholder that is never touched
        imgs = [_resolve(s) if i in mine else None for i, s in enumerate(images)]
        gts = [_resolve(s) if i in mine else None for i, s in enumerate(gt_labels)]
        res = screen_models(nets, imgs, gts, thresholds=tuple(thresholds), device=device, batch=self.batch, rank=shard,
                            world=n_shards)
        return res["table"]


@deployment(
    ray_actor_options={"num_cpus": 1, "num_gpus": 0, "memory": 4 * 1024 * 1024 * 1024},
    max_ongoing_requests=4,
    autoscaling_config={"min_replicas": 1, "initial_replicas": 1, "max_replicas": 1,
                        "target_num_ongoing_requests_per_replica": 1},
    health_check_period_s=30.0, health_check_timeout_s=30.0, graceful_shutdown_timeout_s=120.0,
    graceful_shutdown_wait_loop_s=2.0,
)
class ScreeningEntry:
    def __init__(self, screening_deployment: DeploymentHandle, n_shards: int = 8) -> None:
        self.screening_deployment = screening_deployment
        self.n_shards = int(n_shards)

    async def async_init(self) -> None:
        pass

    async def test_deployment(self) -> None:
        if self.n_shards < 1:
            raise RuntimeError("n_shards must be at least 1")

    async def check_health(self) -> None:
        if self.screening_deployment is None:
            raise RuntimeError("no screening worker deployment")

    @schema_method(arbitrary_types_allowed=True)
    async def screen(self, rdf_paths: List[str], images: List[ArrayOrSource], gt_labels: List[ArrayOrSource],
                     thresholds: Optional[List[float]] = None) -> dict:
        """Rank candidate models by mean average precision over a labelled image set (AP at IoU thresholds as
        `cellpose.metrics.average_precision` computes it).  Images / labels: arrays, or `.npy` / image sources
        (paths, URLs) that the GPU replicas resolve themselves."""
        thresholds = [0.5, 0.75, 0.9] if thresholds is None else [float(t) for t in thresholds]
        if not rdf_paths or not images:
            raise ValueError("nothing to screen: give at least one model and one image")
        if len(images) != len(gt_labels):
            raise ValueError(f"{len(images)} images but {len(gt_labels)} ground-truth label images")
        n = max(1, min(self.n_shards, len(images)))
        try:
            tables = await asyncio.gather(*[
                call_handle(self.screening_deployment, "score_shard", rdf_paths=list(rdf_paths), images=list(images),
                            gt_labels=list(gt_labels), thresholds=thresholds, shard=s, n_shards=n) for s in range(n)])
        except (FileNotFoundError, ValueError):
            raise
        except Exception as e:   # remote task errors surface as RuntimeError, as in the model-runner entry class
            raise RuntimeError(str(e)) from e
        merged = np.sum(np.stack([np.asarray(t, dtype=np.int64) for t in tables]), axis=0)
        res = rank_from_table(merged, thresholds)
        return {"thresholds": thresholds, "ap": res["ap"], "mean_ap": res["mean_ap"],
                "ranking": [rdf_paths[i] for i in res["ranking"]], "ranking_index": res["ranking"].tolist(),
                "table": merged, "n_shards": n}
