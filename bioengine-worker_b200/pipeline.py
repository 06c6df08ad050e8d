"""Device-resident tiled inference pipeline: pre-process → halo tiles → forward → stitch → post → labels.

This is the B200 replacement for what `predict_sample_with_blocking` / `_infer_tiled` + `_prob_to_instances`'s
threshold-and-label step do on the CPU (ref: apps/model-runner/runtime_deployment.py:277-292,
apps/fibsem-mito-analysis/analysis_deployment.py:108-176,241-242).  The image is uploaded once, every stage is a
kernel from libbe_b200.so on torch's current stream, and only the requested results go back to the host.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib, geometry as G, kernels as K


@dataclass
class PipelineSpec:
    """What to run.  `preprocess`: "scale_range" (min/max percentile), "zero_mean_unit_variance", "fibsem" (p1/p99 +
    clip 0..1) or "none"; statistics are per (n, c) over the spatial axes.  `convention`: "bioimageio" (block includes
    halo, crop stitch) or "fibsem" (tile/overlap, Gaussian blend).  `sigmoid`/`threshold`: post-processing; `connectivity`
    None → ndim (skimage default)."""

    block: Sequence[int]
    halo: Sequence[int]                      # bioimageio: halo per side; fibsem: overlap = 2*halo... see `overlap`
    convention: str = "bioimageio"
    overlap: Optional[Sequence[int]] = None  # fibsem only
    preprocess: str = "scale_range"
    min_percentile: float = 1.0
    max_percentile: float = 99.8
    eps: float = 1e-6
    sigmoid: bool = True
    threshold: float = 0.5
    connectivity: Optional[int] = None
    label: bool = True
    return_prob: bool = False
    return_mask: bool = True                 # False: only the bit-packed mask moves between the stitcher and the labelling
    forward_batch: int = 64


class TiledPipeline:
    """Geometry tables, windows and buffers for one (sample shape, spec); reusable across calls."""

    def __init__(self, sample_shape: Sequence[int], spec: PipelineSpec, device="cuda:0"):
        self.device = torch.device(device)
        self.spec = spec
        self.shape = tuple(int(s) for s in sample_shape)
        if len(self.shape) not in (4, 5):
            raise ValueError(f"sample shape must be (N,C,Y,X) or (N,C,Z,Y,X), got {self.shape}")
        self.nd = len(self.shape) - 2
        spatial = self.shape[2:]
        if len(spec.block) != self.nd:
            raise ValueError("block rank does not match the sample")
        if spec.convention == "bioimageio":
            axes = [G.bioimageio_axis(s, b, h) for s, b, h in zip(spatial, spec.block, spec.halo)]
        elif spec.convention == "fibsem":
            ov = spec.overlap if spec.overlap is not None else [2 * h for h in spec.halo]
            axes = [G.fibsem_axis(s, b, o) for s, b, o in zip(spatial, spec.block, ov)]
        else:
            raise ValueError(f"unknown convention {spec.convention!r}")
        if self.nd == 2:
            axes = [G.identity_axis(1)] + axes
        self.axes = [K.DeviceAxis(a, self.device) for a in axes]
        self.n_tiles = self.shape[0] * axes[0].T * axes[1].T * axes[2].T
        self.tile_shape = (self.n_tiles, self.shape[1]) + tuple(spec.block)
        if spec.convention == "fibsem":
            wins = [torch.from_numpy(G.gauss_window_1d(b)).to(self.device) for b in spec.block]
            if self.nd == 2:
                wins = [torch.ones(1, dtype=torch.float64, device=self.device)] + wins
            self.windows = wins
            self.window_table = K.blend_window(wins) if self.nd == 2 else None
        if spec.preprocess == "scale_range":
            self.q = torch.tensor([spec.min_percentile / 100.0, spec.max_percentile / 100.0], dtype=torch.float64,
                                  device=self.device)
        elif spec.preprocess == "fibsem":
            self.q = torch.tensor(np.true_divide([1, 99], 100.0), dtype=torch.float64, device=self.device)
        self.ws = K.Workspace()
        self.tiles = torch.empty(self.tile_shape, dtype=torch.float32, device=self.device)

    # ---- stage 1: statistics + fused normalise/gather ----
    def preprocess_and_tile(self, x: torch.Tensor) -> torch.Tensor:
        sp = self.spec
        n_seg = self.shape[0] * self.shape[1]
        if sp.preprocess in ("scale_range", "fibsem"):
            _, norm = K.quantile(x, n_seg, self.q, sp.eps, ws=self.ws)
            clip = (0.0, 1.0) if sp.preprocess == "fibsem" else None
            return K.tile_gather(x, self.axes, norm=norm, norm_mode=1, clip=clip, out=self.tiles)
        if sp.preprocess == "zero_mean_unit_variance":
            _, norm = K.meanvar(x, n_seg, sp.eps, ws=self.ws)
            return K.tile_gather(x, self.axes, norm=norm, norm_mode=1, out=self.tiles)
        if sp.preprocess == "none":
            return K.tile_gather(x, self.axes, out=self.tiles)
        raise ValueError(f"unknown preprocess {sp.preprocess!r}")

    # ---- the unchanged torch forward, in chunks ----
    @torch.no_grad()
    def forward(self, model: Callable[[torch.Tensor], torch.Tensor], tiles: torch.Tensor,
                out: Optional[torch.Tensor] = None) -> torch.Tensor:
        fb = max(1, int(self.spec.forward_batch))
        first = model(tiles[:fb])
        if first.shape[0] == tiles.shape[0] and out is None:
            return first  # one chunk covered everything
        if out is None:
            out = torch.empty((tiles.shape[0],) + tuple(first.shape[1:]), dtype=torch.float32, device=tiles.device)
        out[:fb] = first
        for i in range(fb, tiles.shape[0], fb):
            out[i:i + fb] = model(tiles[i:i + fb])
        return out

    # ---- stage 2: stitch (+ sigmoid + binarize) + labels ----
    def stitch_and_label(self, logits: torch.Tensor, labels_out: Optional[torch.Tensor] = None,
                         n_out: Optional[torch.Tensor] = None) -> Dict[str, torch.Tensor]:
        sp = self.spec
        out_shape = (self.shape[0], logits.shape[1]) + self.shape[2:]
        epi = _lib.BE_EPI_SIGMOID if sp.sigmoid else _lib.BE_EPI_NONE
        want_mask = sp.return_mask or not sp.label
        # the labelling reads the mask bit-packed (one bit per pixel leaves the stitcher instead of a byte)
        if sp.convention == "bioimageio":
            prob, mask, bits = K.stitch_crop(logits, out_shape, self.axes, epi, sp.threshold, want_out=sp.return_prob,
                                             want_mask=want_mask, want_bits=sp.label)
        else:
            prob, mask, bits = K.stitch_blend(logits, out_shape, self.axes, self.windows, epi, sp.threshold,
                                              want_out=sp.return_prob, want_mask=want_mask, want_bits=sp.label,
                                              window=self.window_table)
        res = {}
        if mask is not None:
            res["mask"] = mask
        if prob is not None:
            res["prob"] = prob
        if sp.label:
            lshape = (out_shape[0] * out_shape[1],) + out_shape[2:]
            labels, n = K.ccl_bits(bits, lshape, sp.connectivity, ws=self.ws, out=labels_out, n_out=n_out)
            res["labels"] = labels.reshape(out_shape)
            res["n_labels"] = n
        return res

    def upload(self, image) -> torch.Tensor:
        """Host ndarray / CPU tensor / CUDA tensor of the sample shape → float32 CUDA tensor.  Integer, half and double
        inputs cross PCIe in their own element type and are converted on the device (`ensure_dtype(float32)`, the op
        bioimageio.core puts in front of every pre-processing chain)."""
        if isinstance(image, torch.Tensor):
            x = image
        else:
            a = np.asarray(image)
            if a.dtype.name not in K.NUMPY_DEVICE_DTYPES:
                a = a.astype(np.float32)
            x = torch.from_numpy(np.ascontiguousarray(a))
        if tuple(x.shape) != self.shape:
            raise ValueError(f"expected sample shape {self.shape}, got {tuple(x.shape)}")
        x = x.to(self.device, non_blocking=True).contiguous()
        return K.to_float32(x)

    def run(self, image, model: Callable[[torch.Tensor], torch.Tensor]) -> Dict[str, torch.Tensor]:
        """image: host ndarray / CPU tensor (uploaded here) or CUDA tensor of the sample shape, any of the element types
        `upload` takes."""
        x = self.upload(image)
        tiles = self.preprocess_and_tile(x)
        logits = self.forward(model, tiles)
        return self.stitch_and_label(logits)

    def _rawbuf(self, b: int, dtype: torch.dtype) -> torch.Tensor:
        if not hasattr(self, "_raw"):
            self._raw = {}
        if (b, dtype) not in self._raw:
            self._raw[(b, dtype)] = torch.empty(self.shape, dtype=dtype, device=self.device)
        return self._raw[(b, dtype)]

    def run_stream(self, host_images: Sequence[torch.Tensor], model: Callable[[torch.Tensor], torch.Tensor],
                   host_labels: Sequence[torch.Tensor], host_counts: Sequence[torch.Tensor]):
        """Throughput mode for a stream of same-shaped images (microscope frames, a screening shard).

        host_images[i] (pinned CPU tensor of the sample shape; fp32, or uint8 / uint16 / ... frames that are converted on
        the device after the upload) → host_labels[i] (pinned int32, sample shape), host_counts[i]
        (pinned int32 [N]).  Three streams: the upload of image i+1 and the download of labels i-1 overlap the kernels
        of image i; device buffers are double-buffered, so the only requirement on the caller is that host_labels[i]
        and host_labels[i+2] may alias but host_labels[i] and host_labels[i+1] may not.  Returns after everything has
        been enqueued; synchronise (or wait on the returned event) before reading the host buffers.
        """
        dev = self.device
        if not hasattr(self, "_streams"):
            self._streams = (torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev))
            self._xbuf = [torch.empty(self.shape, dtype=torch.float32, device=dev) for _ in range(2)]
            lshape = (self.shape[0], 1) + self.shape[2:]
            self._lbuf = [torch.empty(lshape, dtype=torch.int32, device=dev) for _ in range(2)]
            self._nbuf = [torch.empty((self.shape[0],), dtype=torch.int32, device=dev) for _ in range(2)]
        s_in, s_run, s_out = self._streams
        start = torch.cuda.current_stream(dev)
        for s in (s_in, s_run, s_out):
            s.wait_stream(start)
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_run = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]
        for i, img in enumerate(host_images):
            b = i & 1
            with torch.cuda.stream(s_in):
                if i >= 2:
                    s_in.wait_event(ev_run[b])      # the kernels that read this input buffer two images ago
                if img.dtype == torch.float32:
                    self._xbuf[b].copy_(img, non_blocking=True)
                else:   # raw frame up, converted on the device (same stream, so the event below covers both)
                    raw = self._rawbuf(b, img.dtype)
                    raw.copy_(img, non_blocking=True)
                    K.to_float32(raw, out=self._xbuf[b])
                ev_in[b].record(s_in)
            with torch.cuda.stream(s_run):
                s_run.wait_event(ev_in[b])
                if i >= 2:
                    s_run.wait_event(ev_out[b])     # the download that read this label buffer two images ago
                tiles = self.preprocess_and_tile(self._xbuf[b])
                logits = self.forward(model, tiles)
                self.stitch_and_label(logits, labels_out=self._lbuf[b], n_out=self._nbuf[b])
                ev_run[b].record(s_run)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_run[b])
                host_labels[i].copy_(self._lbuf[b], non_blocking=True)
                host_counts[i].copy_(self._nbuf[b], non_blocking=True)
                ev_out[b].record(s_out)
        done = torch.cuda.Event(enable_timing=True)
        start.wait_stream(s_out)
        start.wait_stream(s_run)
        start.wait_stream(s_in)
        done.record(start)
        return done
