"""`MitoAnalysisDeployment` — B200 replacement for ref: apps/fibsem-mito-analysis/analysis_deployment.py.

Same `analyze(image, pixel_size_nm, tile_size, overlap)` arguments and the same result keys.  The reference pushes every
512-px tile through three network round trips to the model-runner service and blends / thresholds / labels with numpy,
scipy and skimage on one CPU core; here the image goes to the GPU once and p1/p99 normalisation, reflect-padded tiling,
the forward, the Gaussian blend, the 0.5 threshold, connected components and the per-label moments all run there.

`_prob_to_instances` (ref: analysis_deployment.py:160-176) runs on the device as well (`kernels.instances_from_mask`):
threshold, `remove_small_objects(min_size=300)`, `binary_closing(disk(4))`, the exact Euclidean distance transform,
`peak_local_max(min_distance=8, labels=closed)`, `measure.label` of the peaks and the marker `watershed` on the negated
distance.  Everything up to the flood matches scipy / skimage bit for bit; the flood matches skimage's order except between
seeds of exactly equal height inside one connected region, where skimage's order is an accident of its heap layout and the
kernel uses raster order (be_b200.h) — a few boundary pixels between twin objects.
`cleanup=False` labels the thresholded map directly (`measure.label(prob > 0.5)`) instead.
"""
from __future__ import annotations

import time
from datetime import datetime
from typing import Callable, Dict, Optional

import numpy as np
import torch

from . import kernels as K, models
from ._serve_compat import deployment, schema_method
from .pipeline import PipelineSpec, TiledPipeline

_MODEL_ID = "b200/random-init-unet2d"


def region_properties(sums: np.ndarray, pixel_size_nm: float) -> Dict[str, list]:
    """skimage regionprops quantities of ref: analysis_deployment.py:258-267 from exact integer raw moments.

    sums[l] = {area, Σz, Σy, Σx, Σzz, Σyy, Σxx, Σzy, Σzx, Σyx} (int64).  Central second moments / area give the inertia
    tensor; eigenvalues l1 >= l2; axis lengths 4·sqrt(l); eccentricity sqrt(1 - l2/l1).
    The scaled central moments A·Σyy − (Σy)² etc. are formed in integers (exact) wherever they fit in int64, and the
    small eigenvalue is taken as det / l1, so thin objects get l2 = 0 instead of cancellation noise."""
    pixel_um = pixel_size_nm / 1000.0
    s = np.asarray(sums)[1:]
    lab = np.nonzero(s[:, 0] > 0)[0]
    s = s[lab].astype(np.int64)
    A, Sy, Sx, Syy, Sxx, Syx = (s[:, j] for j in (0, 2, 3, 5, 6, 9))
    Af = A.astype(np.float64)
    fits = (Af * np.maximum(Syy, Sxx).astype(np.float64)) < 2.0**62
    with np.errstate(over="ignore"):
        m20_i, m02_i, m11_i = A * Syy - Sy * Sy, A * Sxx - Sx * Sx, A * Syx - Sy * Sx
    m20 = np.where(fits, m20_i.astype(np.float64), Af * Syy - Sy.astype(np.float64) ** 2) / (Af * Af)
    m02 = np.where(fits, m02_i.astype(np.float64), Af * Sxx - Sx.astype(np.float64) ** 2) / (Af * Af)
    m11 = np.where(fits, m11_i.astype(np.float64), Af * Syx - Sy.astype(np.float64) * Sx) / (Af * Af)
    tr, det = m20 + m02, np.maximum(m20 * m02 - m11 * m11, 0.0)
    l1 = tr / 2.0 + np.sqrt(np.maximum(tr * tr / 4.0 - det, 0.0))
    with np.errstate(invalid="ignore", divide="ignore"):
        l2 = np.where(l1 > 0, det / l1, 0.0)
        ecc = np.where(l1 > 0, np.sqrt(np.maximum(1.0 - l2 / l1, 0.0)), 0.0)
    major, minor = 4.0 * np.sqrt(np.maximum(l1, 0.0)), 4.0 * np.sqrt(l2)
    return {
        "label": (lab + 1).tolist(),
        "area_um2": (Af * pixel_um**2).tolist(),
        "aspect_ratio": (major / (minor + 1e-6)).tolist(),
        "eccentricity": ecc.tolist(),
        "centroid_y": (Sy / Af).tolist(),
        "centroid_x": (Sx / Af).tolist(),
    }


@deployment(ray_actor_options={"num_cpus": 2, "num_gpus": 1, "memory": 4 * 1024**3})
class MitoAnalysisDeployment:
    """Mitochondria instance segmentation from a 2-D EM image, tiled inference on one B200."""

    def __init__(self, model: Optional[Callable[[torch.Tensor], torch.Tensor]] = None, device="cuda:0"):
        self.device = torch.device(device)
        self.start_time = time.time()
        models.strict_fp32()
        self._model = model if model is not None else models.build_model("UNet2d", seed=0, final_sigmoid=True).to(self.device)
        self._pipes: Dict[tuple, TiledPipeline] = {}

    # ---- lifecycle hooks of the app builder (ref: analysis_deployment.py:53-80, bioengine/apps/builder.py:669-890) ----
    async def async_init(self) -> None:
        pass   # the reference connects to the model-runner service here; the model lives in this process

    async def test_deployment(self) -> None:
        """ref: analysis_deployment.py:72-77 — a 64x64 synthetic tile through the whole path."""
        res = await self.analyze(np.random.rand(64, 64).astype(np.float32), as_numpy=True)
        assert res["image_shape"] == [64, 64]

    async def check_health(self) -> None:
        if self._model is None:
            raise RuntimeError("model not loaded.")

    @schema_method
    async def ping(self) -> dict:
        """Return service status."""
        return {"status": "ok", "model": _MODEL_ID, "model_runner": "in-process (B200)",
                "uptime_s": round(time.time() - self.start_time, 1), "timestamp": datetime.now().isoformat()}

    def _pipeline(self, H: int, W: int, tile_size: int, overlap: int) -> TiledPipeline:
        if H <= tile_size and W <= tile_size:  # ref: analysis_deployment.py:247-248 — single call, no tiling
            key, block, ov = (H, W, 0, 0), (H, W), (0, 0)
        else:
            key, block, ov = (H, W, tile_size, overlap), (tile_size, tile_size), (overlap, overlap)
        if key not in self._pipes:
            spec = PipelineSpec(block=block, halo=(0, 0), convention="fibsem", overlap=ov, preprocess="fibsem",
                                sigmoid=False, threshold=0.5, connectivity=2, label=True, return_prob=False)
            self._pipes[key] = TiledPipeline((1, 1, H, W), spec, device=self.device)
        return self._pipes[key]

    @schema_method
    async def analyze(self, image, pixel_size_nm: float = 5.0, tile_size: int = 512, overlap: int = 64,
                      as_numpy: bool = False, cleanup: bool = True) -> dict:
        """Same contract as ref: analysis_deployment.py:192-286 (`as_numpy=True` skips the nested-list conversion;
        `cleanup=True`, the default, is the reference's `_prob_to_instances`: size filter, closing, distance peaks,
        watershed; `cleanup=False` returns the connected components of the thresholded map)."""
        t0 = time.time()
        if isinstance(image, np.ndarray) and image.dtype.name in K.NUMPY_DEVICE_DTYPES:
            image_np = image   # converted to float32 on the device (same values as the reference's host cast)
        else:
            image_np = np.array(image, dtype=np.float32)
        if image_np.ndim != 2:
            raise ValueError(f"Expected 2-D image, got shape {image_np.shape}.")
        H, W = image_np.shape
        pipe = self._pipeline(H, W, tile_size, overlap)
        res = pipe.run(image_np[None, None], self._model)
        if cleanup:
            labels_d, n_seeds = K.instances_from_mask(res["mask"][0, 0])
            n_mito = int(n_seeds.item())   # every seed keeps its own pixel, so labels.max() is the number of seeds
        else:
            labels_d = res["labels"][0, 0]
            n_mito = int(res["n_labels"].item())
        if n_mito > 0:
            sums, _, status = K.label_stats(labels_d[None], n_mito + 1)
            if int(status.item()) != 0:
                raise RuntimeError("label statistics overflow")
            properties = region_properties(sums.cpu().numpy()[0], pixel_size_nm)
        else:
            properties = {k: [] for k in ("label", "area_um2", "aspect_ratio", "eccentricity", "centroid_y", "centroid_x")}
        labels = labels_d.cpu().numpy()
        return {
            "labels": labels if as_numpy else labels.tolist(),
            "properties": properties,
            "n_mitochondria": n_mito,
            "image_shape": [H, W],
            "pixel_size_nm": pixel_size_nm,
            "model": _MODEL_ID,
            "processing_time_s": round(time.time() - t0, 2),
        }
