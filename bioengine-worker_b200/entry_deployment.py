"""`EntryDeployment.infer` — the model-runner's public method, kept signature-compatible with
ref: apps/model-runner/entry_deployment.py:1869-1990 so callers (CLI, agents, the fibsem app) do not change.

Only the data path is mirrored: resolve the inputs (ndarray, dict of ndarrays, or a local `.npy` path), find the model
package on disk, forward to `RuntimeDeployment.predict`, optionally hand the result back as a file instead of an array.
Model download / LRU disk cache / test-report publishing / S3 presigned URLs (entry_deployment.py:72-1009, 1265-1300,
1821-1867) are control plane and storage — out of scope; a deployment that needs them keeps the reference's
EntryDeployment and swaps only the runtime deployment.
"""
from __future__ import annotations

import os
import tempfile
import uuid
from pathlib import Path
from typing import Dict, Literal, Optional, Union

import numpy as np
import torch


class EntryDeployment:
    def __init__(self, runtime_deployment, models_dir: str, cache_size_in_gb: float = 50.0):
        self.runtime_deployment = runtime_deployment
        self.models_dir = Path(models_dir)
        self.cache_size_in_gb = cache_size_in_gb
        self._tmp = Path(tempfile.mkdtemp(prefix="b200-model-runner-"))

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
        """ref: entry_deployment.py:1196-1263 restricted to local files (`file://` prefix accepted): `.npy`, `.png`,
        `.tif(f)`, `.jp(e)g`, decoded as `imageio.v3.imread` would; same exceptions (ValueError for an unsupported
        extension or unreadable content, FileNotFoundError for a missing file)."""
        from .io import load_image

        if source.startswith(("http://", "https://")):
            raise ValueError("remote inputs are resolved by the reference EntryDeployment; pass an array or a local file path")
        return load_image(source[7:] if source.startswith("file://") else source)

    async def get_model_rdf(self, model_id: str) -> dict:
        import yaml

        return yaml.safe_load(self._rdf_path(model_id).read_text())

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
        self._check_runtime_available()
        if isinstance(inputs, str):
            inputs = self._load_image_from_source(inputs)
        elif isinstance(inputs, dict):
            inputs = {k: (self._load_image_from_source(v) if isinstance(v, str) else v) for k, v in inputs.items()}
        rdf_path = self._rdf_path(model_id)
        try:
            outputs = await self.runtime_deployment.predict(
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
