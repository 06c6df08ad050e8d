"""`EntryDeployment` — the model-runner's entry class, kept compatible with ref: apps/model-runner/entry_deployment.py
where the data path runs: the constructor the app builder calls (`runtime_deployment: DeploymentHandle,
cache_size_in_gb=50.0`, :1056-1060), `infer(...)` with the reference's signature (:1869-1907) as a schema method, the
call into the runtime replica as `runtime_deployment.predict.remote(...)` (:1969-1977), and the lifecycle hooks the builder
looks for (`async_init`, `test_deployment`, `check_health`; ref: bioengine/apps/builder.py:669-890).

Only the data path is mirrored: resolve the inputs (ndarray, dict of ndarrays, a direct URL or a local `.npy` / image path), find the
model package on disk, forward to `RuntimeDeployment.predict`, optionally hand the result back as a file instead of an
array.  Model download / LRU disk cache / test-report publishing / S3 presigned URLs (entry_deployment.py:72-1009,
1265-1300, 1821-1867) are control plane and storage — out of scope; a deployment that needs them keeps the
reference's EntryDeployment and swaps only the runtime deployment.  Model packages are looked up
under `models_dir` (keyword, or $BIOENGINE_B200_MODELS_DIR, default ~/.bioengine/models/<model_id>/rdf.yaml).
"""
import os
import tempfile
import uuid
from pathlib import Path
from typing import Dict, Literal, Optional, Union

import numpy as np
import torch

from ._serve_compat import DeploymentHandle, call_handle, deployment, schema_method


@deployment(
    ray_actor_options={"num_cpus": 1, "num_gpus": 0, "memory": 16 * 1024 * 1024 * 1024},
    max_ongoing_requests=10,
    max_queued_requests=30,
    autoscaling_config={"min_replicas": 1, "initial_replicas": 1, "max_replicas": 1,
                        "target_num_ongoing_requests_per_replica": 1},
    health_check_period_s=30.0, health_check_timeout_s=30.0, graceful_shutdown_timeout_s=300.0,
    graceful_shutdown_wait_loop_s=2.0,
)
class EntryDeployment:
    def __init__(self, runtime_deployment: DeploymentHandle, cache_size_in_gb: float = 50.0, *,
                 models_dir: Optional[str] = None) -> None:
        self.runtime_deployment = runtime_deployment
        self.cache_size_in_gb = cache_size_in_gb
        self.models_dir = Path(models_dir or os.environ.get("BIOENGINE_B200_MODELS_DIR")
                               or Path.home() / ".bioengine" / "models")
        self._tmp = Path(tempfile.mkdtemp(prefix="b200-model-runner-"))

    # ---- lifecycle hooks the builder calls when present (ref: bioengine/apps/builder.py:669-890) ----
    async def async_init(self) -> None:
        self.models_dir.mkdir(parents=True, exist_ok=True)

    async def test_deployment(self) -> None:
        """The runtime replica must answer (the reference runs a model test here; no zoo model is available offline)."""
        self._check_runtime_available()

    async def check_health(self) -> None:
        if not self._tmp.exists():
            raise RuntimeError("temporary result directory is gone")

    def _check_runtime_available(self) -> None:
        """ref: entry_deployment.py:1103-1118 — same message the reference's tests look for."""
        if not torch.cuda.is_available():
            raise RuntimeError("GPU runtime deployment is not available. No GPU could be reached by the runtime replica.")

    def _rdf_path(self, model_id: str) -> Path:
        p = self.models_dir / model_id / "rdf.yaml"
        if not p.exists():
            raise FileNotFoundError(f"model {model_id!r} is not available locally (expected {p})")
        return p

    @staticmethod
    def _load_image_from_source(source: str) -> np.ndarray:
        """ref: entry_deployment.py:1196-1263 for direct `http(s)://` URLs and local files (`file://` prefix accepted):
        `.npy`, `.png`, `.tif(f)`, `.jp(e)g`, decoded as `imageio.v3.imread` would; same exceptions (ValueError for an
        unsupported extension or unreadable content, FileNotFoundError for a missing / expired source).  Temporary paths of
        the reference's `get_upload_url` (presigned S3) are the storage service's business, not resolved here."""
        from .io import fetch_image, load_image

        if source.startswith(("http://", "https://")):
            return fetch_image(source)
        return load_image(source[7:] if source.startswith("file://") else source)

    @schema_method
    async def get_model_rdf(self, model_id: str) -> dict:
        """Return the resource description (rdf.yaml) of a locally available model."""
        import yaml

        return yaml.safe_load(self._rdf_path(model_id).read_text())

    @schema_method(arbitrary_types_allowed=True)
    async def infer(
        self,
        model_id: str,
        inputs: Union[np.ndarray, Dict[str, np.ndarray], str],
        weights_format: Optional[str] = None,
        device: Optional[Literal["cuda", "cpu"]] = None,
        default_blocksize_parameter: Optional[int] = None,
        sample_id: str = "sample",
        skip_cache: bool = False,
        return_download_url: bool = False,
    ) -> Dict[str, Union[np.ndarray, str]]:
        """Execute inference on a locally available bioimage.io model (same arguments and result as the reference's
        `infer`, ref: entry_deployment.py:1869-1990)."""
        self._check_runtime_available()
        import asyncio

        if isinstance(inputs, str):   # downloads and decoding off the event loop, as the reference does
            inputs = await asyncio.to_thread(self._load_image_from_source, inputs)
        elif isinstance(inputs, dict):
            inputs = {k: (await asyncio.to_thread(self._load_image_from_source, v) if isinstance(v, str) else v)
                      for k, v in inputs.items()}
        rdf_path = self._rdf_path(model_id)
        try:
            outputs = await call_handle(
                self.runtime_deployment, "predict",
                rdf_path=str(rdf_path), inputs=inputs, weights_format=weights_format, device=device,
                default_blocksize_parameter=default_blocksize_parameter, sample_id=sample_id,
                latest_remote_modified=rdf_path.stat().st_mtime if not skip_cache else None)
        except (FileNotFoundError, ValueError):
            raise
        except Exception as e:  # the reference re-wraps remote task errors as RuntimeError (entry_deployment.py:1978-1981)
            raise RuntimeError(str(e)) from e
        if return_download_url:
            out = {}
            for k, v in outputs.items():
                f = self._tmp / f"{sample_id}-{k}-{uuid.uuid4().hex[:8]}.npy"
                np.save(f, v)
                out[k] = f"file://{f}"
            return out
        return outputs
