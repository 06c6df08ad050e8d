"""Wire formats of the model-runner data path: `.npy` in, `.npy` out (minimal).

The reference moves arrays as `.npy` files over presigned-S3 HTTP (`np.load(BytesIO(...))` / `np.save`, ref:
apps/model-runner/entry_deployment.py:1196-1300, bioengine/cli/utils.py:93-212) and `analyze` even takes nested Python
lists.  Here a `.npy` file is memory-mapped, copied once into a reusable pinned staging buffer and uploaded
asynchronously, and results are written from a pinned download buffer — no intermediate pageable copies.
PNG / TIFF / JPEG buffers are decoded with OpenCV (`decode_image`), arranged the way `imageio.v3.imread` — what the
reference calls (entry_deployment.py:1250-1253) — would return them: (H, W) for grey, (H, W, 3|4) in RGB(A) order for
colour, (pages, H, W[, C]) for a multi-page TIFF, sample dtype preserved (uint8 / uint16 / float32).
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch


class PinnedStaging:
    """Grow-only pinned host buffer reused across requests (one per replica)."""

    def __init__(self):
        self._buf: Optional[torch.Tensor] = None

    def view(self, shape, dtype: torch.dtype) -> torch.Tensor:
        n = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
        if self._buf is None or self._buf.numel() < n:
            self._buf = torch.empty(max(n, 1 << 20), dtype=torch.uint8).pin_memory()
        return self._buf[:n].view(dtype).view(*shape)


_NP2TORCH = {np.dtype(k): v for k, v in {
    "float32": torch.float32, "float64": torch.float64, "uint8": torch.uint8, "int16": torch.int16,
    "int32": torch.int32, "int64": torch.int64, "float16": torch.float16, "bool": torch.bool}.items()}


def load_npy_to_device(path: str, device="cuda:0", staging: Optional[PinnedStaging] = None,
                       dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """`.npy` file → device tensor of `dtype` (uint16 microscope frames are widened on the way; torch has no uint16
    arithmetic, so they are staged as int32)."""
    arr = np.load(path, mmap_mode="r", allow_pickle=False)
    if arr.dtype == np.uint16:
        arr = arr.astype(np.int32)
    if arr.dtype not in _NP2TORCH:
        raise ValueError(f"unsupported .npy dtype {arr.dtype}")
    staging = staging or PinnedStaging()
    host = staging.view(arr.shape, _NP2TORCH[arr.dtype])
    np.copyto(host.numpy(), arr)  # the only host-side copy: file pages → pinned memory
    return host.to(device, non_blocking=True).to(dtype)


def save_npy_from_device(path: str, t: torch.Tensor, staging: Optional[PinnedStaging] = None) -> None:
    """device tensor → `.npy` file through a pinned download buffer."""
    staging = staging or PinnedStaging()
    host = staging.view(tuple(t.shape), t.dtype)
    host.copy_(t, non_blocking=True)
    torch.cuda.current_stream(t.device).synchronize()
    np.save(path, host.numpy())


SUPPORTED_FILE_TYPES = (".npy", ".png", ".tiff", ".tif", ".jpeg", ".jpg")  # ref: entry_deployment.py:29


def decode_image(data: bytes, ext: str) -> np.ndarray:
    """Image file content → ndarray (ref: entry_deployment.py:1246-1253).  Raises ValueError for anything unreadable."""
    import io as _io

    ext = ext.lower()
    if ext not in SUPPORTED_FILE_TYPES:
        raise ValueError(f"Unsupported file extension '{ext}'. Supported extensions: {SUPPORTED_FILE_TYPES}")
    if ext == ".npy":
        return np.load(_io.BytesIO(data), allow_pickle=False)
    import cv2

    buf = np.frombuffer(data, dtype=np.uint8)

    def to_rgb(a: np.ndarray) -> np.ndarray:   # OpenCV hands colour back as BGR(A)
        if a.ndim == 3 and a.shape[2] == 3:
            return np.ascontiguousarray(a[:, :, ::-1])
        if a.ndim == 3 and a.shape[2] == 4:
            return np.ascontiguousarray(a[:, :, [2, 1, 0, 3]])
        return a

    if ext in (".tif", ".tiff"):
        ok, pages = cv2.imdecodemulti(buf, cv2.IMREAD_UNCHANGED)
        if not ok or not pages:
            raise ValueError("Failed to parse image: not a readable TIFF")
        pages = [to_rgb(p) for p in pages]
        return pages[0] if len(pages) == 1 else np.stack(pages)
    img = cv2.imdecode(buf, cv2.IMREAD_UNCHANGED)
    if img is None:
        raise ValueError(f"Failed to parse image: not a readable {ext} file")
    return to_rgb(img)


def load_image(path: str) -> np.ndarray:
    """A local image file of a supported type → ndarray."""
    import os

    ext = os.path.splitext(path.split("?")[0])[1]
    if ext.lower() not in SUPPORTED_FILE_TYPES:
        raise ValueError(f"Unsupported file extension '{ext}' in source '{path}'. Supported extensions: {SUPPORTED_FILE_TYPES}")
    if not os.path.exists(path):
        raise FileNotFoundError(f"Source '{path}' does not exist or has expired.")
    with open(path, "rb") as f:
        return decode_image(f.read(), ext)


def fetch_image(url: str, timeout: float = 30.0, retries: int = 3) -> np.ndarray:
    """A direct `http(s)://` source → ndarray, as ref: entry_deployment.py:1217-1258 does it: the extension (query string
    stripped) must be a supported one (ValueError before anything is fetched); 404 → FileNotFoundError("... does not exist
    or has expired."); any other failure to download → FileNotFoundError("Failed to download source ..."); content that
    does not parse → ValueError.  Transient failures (connection errors, 5xx) are retried like the reference's
    `_get_url_with_retry`."""
    import os
    import time
    import urllib.error
    import urllib.request

    ext = os.path.splitext(url.split("?")[0])[1].lower()
    if ext not in SUPPORTED_FILE_TYPES:
        raise ValueError(f"Unsupported file extension '{ext}' in source '{url}'. Supported extensions: {SUPPORTED_FILE_TYPES}")
    if not url.startswith(("http://", "https://")):
        raise ValueError(f"not an http(s) source: '{url}'")
    data, last = None, None
    for attempt in range(max(1, retries)):
        try:
            with urllib.request.urlopen(url, timeout=timeout) as resp:
                data = resp.read()
            break
        except urllib.error.HTTPError as e:
            if e.code == 404:
                raise FileNotFoundError(f"Source '{url}' does not exist or has expired.") from e
            last = e
            if e.code < 500:   # a client error will not go away
                break
        except (urllib.error.URLError, OSError) as e:
            last = e
        time.sleep(0.2 * (attempt + 1))
    if data is None:
        raise FileNotFoundError(f"Failed to download source '{url}': {last}") from last
    try:
        return decode_image(data, ext)
    except ValueError as e:
        raise ValueError(f"Failed to parse image from source '{url}': {e}") from e
