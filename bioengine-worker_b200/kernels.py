"""Thin torch-tensor front end of the C-ABI: shape checks, workspace allocation, stream plumbing.

Every function takes CUDA tensors, launches on torch's current stream and returns CUDA tensors.  PyTorch is
used for device memory and streams only; all arithmetic happens in libbe_b200.so.
"""
from __future__ import annotations

import functools
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .geometry import AxisBlocks


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _tensor_devices(obj, found):
    if isinstance(obj, torch.Tensor):
        found.append(obj.device)
    elif isinstance(obj, DeviceAxis):
        found.append(obj.src.device)
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            _tensor_devices(o, found)
    elif isinstance(obj, dict):
        for o in obj.values():
            _tensor_devices(o, found)


def _on_tensor_device(fn):
    """The library launches on the CUDA device that is current and on the stream handed to it.  Every front-end function
    therefore makes the device of its tensors current for the duration of the call (so `_stream()` is that device's
    current stream) and refuses tensors that sit on different devices."""

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        found = []
        _tensor_devices(args, found)
        _tensor_devices(list(kwargs.values()), found)
        cuda = [d for d in found if d.type == "cuda"]
        if not cuda:
            return fn(*args, **kwargs)   # the function's own checks report the non-CUDA tensor
        if any(d != cuda[0] for d in cuda):
            raise ValueError(f"{fn.__name__}: tensors on different devices ({sorted({str(d) for d in cuda})})")
        if cuda[0].index is None or cuda[0].index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(cuda[0]):
            return fn(*args, **kwargs)

    return wrapper


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _check(t: torch.Tensor, dtype: torch.dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")
    if t.dtype != dtype:
        raise ValueError(f"{name} must be {dtype}, got {t.dtype}")
    if not t.is_contiguous():
        raise ValueError(f"{name} must be contiguous")
    return t


class Workspace:
    """Grow-only byte arena reused across calls on one stream (the replica is single-stream: max_ongoing_requests=1,
    ref: apps/model-runner/runtime_deployment.py:40)."""

    def __init__(self):
        self._buf: Optional[torch.Tensor] = None

    def get(self, nbytes: int, device) -> torch.Tensor:
        if self._buf is None or self._buf.numel() < nbytes or self._buf.device != torch.device(device):
            self._buf = torch.empty(max(nbytes, 256), dtype=torch.uint8, device=device)
        return self._buf


_default_ws = Workspace()


def _segments(x: torch.Tensor, n_seg: int) -> Tuple[int, int]:
    n = x.numel()
    if n_seg <= 0 or n % n_seg:
        raise ValueError(f"{n} elements do not split into {n_seg} segments")
    return n_seg, n // n_seg


@_on_tensor_device
def meanvar(x: torch.Tensor, n_seg: int, eps: float = 1e-6, ws: Workspace = _default_ws):
    """→ (stats [n_seg,2] fp64 {mean,var}, norm [n_seg,2] fp64 {mean, std+eps})."""
    _check(x, torch.float32, "x")
    n_seg, seg_len = _segments(x, n_seg)
    stats = torch.empty((n_seg, 2), dtype=torch.float64, device=x.device)
    norm = torch.empty((n_seg, 2), dtype=torch.float64, device=x.device)
    w = ws.get(_lib.query("be_meanvar_workspace_bytes", n_seg, seg_len), x.device)
    _lib.call("be_meanvar_f32", x.data_ptr(), n_seg, seg_len, float(eps), stats.data_ptr(), norm.data_ptr(),
              w.data_ptr(), _stream())
    return stats, norm


@_on_tensor_device
def quantile(x: torch.Tensor, n_seg: int, q: torch.Tensor, eps: float = 1e-6, want_norm: bool = True,
             ws: Workspace = _default_ws):
    """Exact nanquantile(method="linear") per segment.  q: fp64 CUDA tensor of 1 or 2 fractions.
    → (quant [n_seg, n_q] fp64, norm [n_seg,2] fp64 {q0, q1-q0+eps} or None)."""
    _check(x, torch.float32, "x")
    _check(q, torch.float64, "q")
    n_seg, seg_len = _segments(x, n_seg)
    n_q = q.numel()
    quant = torch.empty((n_seg, n_q), dtype=torch.float64, device=x.device)
    norm = torch.empty((n_seg, 2), dtype=torch.float64, device=x.device) if (want_norm and n_q == 2) else None
    w = ws.get(_lib.query("be_quantile_workspace_bytes", n_seg, seg_len), x.device)
    _lib.call("be_quantile_f32", x.data_ptr(), n_seg, seg_len, q.data_ptr(), n_q, float(eps), quant.data_ptr(),
              _p(norm), w.data_ptr(), _stream())
    return quant, norm


@_on_tensor_device
def normalize(x: torch.Tensor, norm: torch.Tensor, clip: Optional[Tuple[float, float]] = None,
              out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _check(x, torch.float32, "x")
    _check(norm, torch.float64, "norm")
    n_seg, seg_len = _segments(x, norm.shape[0])
    out = torch.empty_like(x) if out is None else _check(out, torch.float32, "out")
    lo, hi = clip if clip is not None else (0.0, 0.0)
    _lib.call("be_normalize_f32", x.data_ptr(), out.data_ptr(), n_seg, seg_len, norm.data_ptr(),
              int(clip is not None), float(lo), float(hi), _stream())
    return out


@_on_tensor_device
def elementwise(x: torch.Tensor, op: int, a: float = 0.0, b: float = 0.0, out: Optional[torch.Tensor] = None):
    _check(x, torch.float32, "x")
    out = torch.empty_like(x) if out is None else _check(out, torch.float32, "out")
    _lib.call("be_elementwise_f32", x.data_ptr(), out.data_ptr(), x.numel(), int(op), float(a), float(b), _stream())
    return out


def clip(x, lo, hi, out=None):
    return elementwise(x, _lib.BE_OP_CLIP, lo, hi, out)


def scale_linear(x, gain, offset, out=None):
    return elementwise(x, _lib.BE_OP_SCALE_LINEAR, gain, offset, out)


@_on_tensor_device
def scale_linear_per_channel(x: torch.Tensor, gain, offset, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bioimageio `scale_linear` with `gain` / `offset` given per channel (lists along the channel axis; a scalar is
    repeated): y[n, c] = x[n, c] * gain[c] + offset[c] in float32, for an (N, C, ...) tensor."""
    _check(x, torch.float32, "x")
    N, C = x.shape[:2]
    g = np.broadcast_to(np.asarray(gain, dtype=np.float64).reshape(-1), (C,)) if np.ndim(gain) else np.full(C, float(gain))
    o = np.broadcast_to(np.asarray(offset, dtype=np.float64).reshape(-1), (C,)) if np.ndim(offset) else np.full(C, float(offset))
    params = torch.from_numpy(np.tile(np.stack([o, g], axis=1), (N, 1))).to(x.device)   # [N*C][offset, gain]
    y = out if out is not None else torch.empty_like(x)
    n_seg, seg_len = _segments(x, N * C)
    _lib.call("be_scale_linear_seg_f32", x.data_ptr(), y.data_ptr(), n_seg, seg_len, params.data_ptr(), _stream())
    return y


@_on_tensor_device
def scale_mean_variance(x: torch.Tensor, ref_norm: torch.Tensor, n_seg: int, eps: float = 1e-6,
                        out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bioimageio `scale_mean_variance`: (x - mean) / (std + eps) * (ref_std + eps) + ref_mean per segment.
    ref_norm = `meanvar(reference, n_seg, eps)[1]` ({ref_mean, ref_std + eps}, fp64 on the device).  Three launches
    (statistics, normalise, per-segment scale_linear with the constants read from device memory), nothing on the host."""
    _check(ref_norm, torch.float64, "ref_norm")
    own = meanvar(x, n_seg, eps)[1]
    y = normalize(x, own, out=out)
    n_seg, seg_len = _segments(y, n_seg)
    _lib.call("be_scale_linear_seg_f32", y.data_ptr(), y.data_ptr(), n_seg, seg_len, ref_norm.data_ptr(), _stream())
    return y


_CAST_CODES = {torch.uint8: 0, torch.int8: 1, torch.uint16: 2, torch.int16: 3, torch.int32: 4, torch.uint32: 5,
               torch.float16: 6, torch.float64: 7}
NUMPY_DEVICE_DTYPES = ("uint8", "int8", "uint16", "int16", "int32", "uint32", "float16", "float32", "float64")


@_on_tensor_device
def to_float32(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """`ensure_dtype(float32)` on the device: an integer / half / double CUDA tensor → float32 (numpy astype rounding)."""
    if not x.is_cuda or not x.is_contiguous():
        raise ValueError("x must be a contiguous CUDA tensor")
    if x.dtype == torch.float32:
        return x
    if x.dtype not in _CAST_CODES:
        raise ValueError(f"unsupported input dtype {x.dtype}")
    out = torch.empty(x.shape, dtype=torch.float32, device=x.device) if out is None else _check(out, torch.float32, "out")
    _lib.call("be_cast_f32", x.data_ptr(), _CAST_CODES[x.dtype], out.data_ptr(), x.numel(), _stream())
    return out


def sigmoid(x, out=None):
    return elementwise(x, _lib.BE_OP_SIGMOID, 0.0, 0.0, out)


@_on_tensor_device
def binarize(x: torch.Tensor, thr: float) -> torch.Tensor:
    _check(x, torch.float32, "x")
    out = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    _lib.call("be_binarize_f32_u8", x.data_ptr(), out.data_ptr(), x.numel(), float(thr), _stream())
    return out


class DeviceAxis:
    """An AxisBlocks with its tables resident on the device."""

    def __init__(self, ax: AxisBlocks, device):
        self.ax = ax
        self.src = torch.from_numpy(ax.src).to(device)
        self.sel = torch.from_numpy(ax.sel).to(device) if ax.sel.size else None


def _as5d(img: torch.Tensor) -> Tuple[int, int, int, int, int]:
    if img.dim() == 4:
        N, C, Y, X = img.shape
        return N, C, 1, Y, X
    if img.dim() == 5:
        return tuple(img.shape)
    raise ValueError(f"expected (N,C,Y,X) or (N,C,Z,Y,X), got {tuple(img.shape)}")


@_on_tensor_device
def tile_gather(img: torch.Tensor, axes: Sequence[DeviceAxis], norm: Optional[torch.Tensor] = None, norm_mode: int = 0,
                clip: Optional[Tuple[float, float]] = None, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """img (N,C,[Z,]Y,X) fp32 → tiles (N*T, C, [Bz,] By, Bx) fp32.  axes = (z, y, x) DeviceAxis (z = identity for 2-D)."""
    _check(img, torch.float32, "img")
    N, C, Z, Y, X = _as5d(img)
    az, ay, ax = axes
    if (az.ax.S, ay.ax.S, ax.ax.S) != (Z, Y, X):
        raise ValueError("axis tables do not match the image shape")
    T = az.ax.T * ay.ax.T * ax.ax.T
    shape = (N * T, C, az.ax.B, ay.ax.B, ax.ax.B) if img.dim() == 5 else (N * T, C, ay.ax.B, ax.ax.B)
    out = torch.empty(shape, dtype=torch.float32, device=img.device) if out is None else _check(out, torch.float32, "out")
    lo, hi = clip if clip is not None else (0.0, 0.0)
    _lib.call("be_tile_gather_f32", img.data_ptr(), out.data_ptr(), N, C, Z, Y, X, az.ax.T, ay.ax.T, ax.ax.T,
              az.ax.B, ay.ax.B, ax.ax.B, az.src.data_ptr(), ay.src.data_ptr(), ax.src.data_ptr(), _p(norm),
              int(norm_mode if norm is not None else 0), int(clip is not None), float(lo), float(hi), _stream())
    return out


def packed_words(X: int) -> int:
    """64-bit words per row of a bit-packed mask."""
    return (int(X) + 63) // 64


def _stitch_outputs(shape, device, want_out, want_mask, want_bits):
    out = torch.empty(shape, dtype=torch.float32, device=device) if want_out else None
    mask = torch.empty(shape, dtype=torch.uint8, device=device) if want_mask else None
    bits = torch.empty(tuple(shape[:-1]) + (packed_words(shape[-1]),), dtype=torch.int64, device=device) if want_bits else None
    return out, mask, bits


@_on_tensor_device
def stitch_crop(tiles: torch.Tensor, sample_shape: Sequence[int], axes: Sequence[DeviceAxis], epilogue: int = 0,
                thr: float = 0.5, want_out: bool = True, want_mask: bool = False, want_bits: bool = False):
    """bioimageio inner-region reassembly.  → (out fp32 or None, mask uint8 or None), each of sample_shape; with
    want_bits a third element: the mask bit-packed, int64 (*sample_shape[:-1], packed_words(X))."""
    _check(tiles, torch.float32, "tiles")
    N, C, Z, Y, X = (sample_shape[0], sample_shape[1], 1, *sample_shape[2:]) if len(sample_shape) == 4 else sample_shape
    az, ay, ax = axes
    out, mask, bits = _stitch_outputs(tuple(sample_shape), tiles.device, want_out, want_mask, want_bits)
    _lib.call("be_stitch_crop_f32", tiles.data_ptr(), _p(out), _p(mask), _p(bits), N, C, Z, Y, X, az.ax.T, ay.ax.T, ax.ax.T,
              az.ax.B, ay.ax.B, ax.ax.B, az.sel.data_ptr(), ay.sel.data_ptr(), ax.sel.data_ptr(), int(epilogue),
              float(thr), _stream())
    return (out, mask, bits) if want_bits else (out, mask)


@_on_tensor_device
def blend_window(windows: Sequence[torch.Tensor]) -> torch.Tensor:
    """The reference's fp32 window table for 2-D tiles, (By, Bx) float32 = fp32(wz[0] * wy[iy] * wx[ix]); computed once per
    geometry and passed to `stitch_blend(..., window=...)`."""
    wz, wy, wx = windows
    table = torch.empty((wy.numel(), wx.numel()), dtype=torch.float32, device=wy.device)
    _lib.call("be_blend_window_f32", wz.data_ptr(), wy.data_ptr(), wx.data_ptr(), wy.numel(), wx.numel(), table.data_ptr(),
              _stream())
    return table


@_on_tensor_device
def stitch_blend(tiles: torch.Tensor, sample_shape: Sequence[int], axes: Sequence[DeviceAxis],
                 windows: Sequence[torch.Tensor], epilogue: int = 0, thr: float = 0.5, want_out: bool = True,
                 want_mask: bool = False, want_bits: bool = False, window: Optional[torch.Tensor] = None):
    """fibsem Gaussian blend.  windows = (wz, wy, wx) fp64 CUDA tensors (wz = [1.0] for 2-D); window = `blend_window(windows)`
    (optional: lets a 2-D mask-only call take the fp32 filter kernel).  Returns as stitch_crop."""
    _check(tiles, torch.float32, "tiles")
    N, C, Z, Y, X = (sample_shape[0], sample_shape[1], 1, *sample_shape[2:]) if len(sample_shape) == 4 else sample_shape
    az, ay, ax = axes
    out, mask, bits = _stitch_outputs(tuple(sample_shape), tiles.device, want_out, want_mask, want_bits)
    _lib.call("be_stitch_blend_f32", tiles.data_ptr(), _p(out), _p(mask), _p(bits), N, C, Z, Y, X, az.ax.T, ay.ax.T, ax.ax.T,
              az.ax.B, ay.ax.B, ax.ax.B, az.ax.stride, ay.ax.stride, ax.ax.stride, windows[0].data_ptr(),
              windows[1].data_ptr(), windows[2].data_ptr(), _p(window), int(epilogue), float(thr), _stream())
    return (out, mask, bits) if want_bits else (out, mask)


def _ccl_dims(shape):
    if len(shape) == 3:
        N, Y, X = shape
        return N, 1, Y, X
    if len(shape) == 4:
        return tuple(shape)
    raise ValueError(f"expected (N,Y,X) or (N,Z,Y,X), got {tuple(shape)}")


def _ccl_call(fn: str, lead_args, shape, device, connectivity, ws, out, n_out):
    N, Z, Y, X = _ccl_dims(shape)
    ndim = 3 if Z > 1 else 2
    conn = ndim if connectivity is None else int(connectivity)
    labels = torch.empty(tuple(shape), dtype=torch.int32, device=device) if out is None else _check(out, torch.int32, "out")
    n_labels = torch.empty((N,), dtype=torch.int32, device=device) if n_out is None else _check(n_out, torch.int32, "n_out")
    if labels.numel() != N * Z * Y * X or n_labels.numel() != N:
        raise ValueError("output buffers do not match the mask")
    w = ws.get(_lib.query("be_ccl_workspace_bytes", N, Z, Y, X), device)
    _lib.call(fn, *lead_args, labels.data_ptr(), N, Z, Y, X, conn, n_labels.data_ptr(), w.data_ptr(), _stream())
    return labels, n_labels


@_on_tensor_device
def ccl(mask: torch.Tensor, connectivity: Optional[int] = None, ws: Workspace = _default_ws,
        out: Optional[torch.Tensor] = None, n_out: Optional[torch.Tensor] = None):
    """mask (N,[Z,]Y,X) uint8 → (labels int32 same shape, n_labels int32 [N] on device)."""
    _check(mask, torch.uint8, "mask")
    return _ccl_call("be_ccl_u8_i32", [mask.data_ptr()], mask.shape, mask.device, connectivity, ws, out, n_out)


@_on_tensor_device
def ccl_bits(bits: torch.Tensor, shape: Sequence[int], connectivity: Optional[int] = None, ws: Workspace = _default_ws,
             out: Optional[torch.Tensor] = None, n_out: Optional[torch.Tensor] = None):
    """Bit-packed mask (int64 words, rows of `packed_words(X)`, lowest bit = first pixel) of an (N,[Z,]Y,X) image →
    (labels int32 of `shape`, n_labels int32 [N])."""
    _check(bits, torch.int64, "bits")
    N, Z, Y, X = _ccl_dims(tuple(shape))
    if bits.numel() != N * Z * Y * packed_words(X):
        raise ValueError("bits do not match the shape")
    return _ccl_call("be_ccl_bits_i32", [bits.data_ptr()], tuple(shape), bits.device, connectivity, ws, out, n_out)


@_on_tensor_device
def ccl_logits(x: torch.Tensor, epilogue: int = 0, thr: float = 0.5, connectivity: Optional[int] = None,
               ws: Workspace = _default_ws, out: Optional[torch.Tensor] = None, n_out: Optional[torch.Tensor] = None):
    """x (N,[Z,]Y,X) fp32 → labels of (sigmoid(x) if epilogue == BE_EPI_SIGMOID else x) > thr, no mask in between."""
    _check(x, torch.float32, "x")
    return _ccl_call("be_ccl_f32_i32", [x.data_ptr(), int(epilogue), float(thr)], x.shape, x.device, connectivity, ws,
                     out, n_out)


@_on_tensor_device
def label_overlap(lab_true: torch.Tensor, lab_pred: torch.Tensor, capacity: int, cap_true: int, cap_pred: int,
                  bounds: Optional[torch.Tensor] = None):
    """(N, ...) int32 label maps → dict of device tensors: keys, counts (pairs with both labels > 0), area_true,
    area_pred (background overlaps included), max_lbl, status.  `bounds`: optional int32 [N, 2] device tensor of per-image
    upper bounds of the true / pred labels — only the part of the tables they need is cleared and used."""
    _check(lab_true, torch.int32, "lab_true")
    _check(lab_pred, torch.int32, "lab_pred")
    if lab_true.shape != lab_pred.shape:
        raise ValueError("label maps differ in shape")
    N = lab_true.shape[0]
    n_px = lab_true.numel() // N
    row_len = lab_true.shape[-1] if lab_true.dim() >= 3 else 0
    dev = lab_true.device
    if bounds is not None:
        _check(bounds, torch.int32, "bounds")
        if bounds.numel() != 2 * N:
            raise ValueError("bounds must be [N, 2]")
    r = dict(
        keys=torch.empty((N, capacity), dtype=torch.int64, device=dev),
        counts=torch.empty((N, capacity), dtype=torch.int32, device=dev),
        area_true=torch.empty((N, cap_true), dtype=torch.int32, device=dev),
        area_pred=torch.empty((N, cap_pred), dtype=torch.int32, device=dev),
        max_lbl=torch.empty((N, 2), dtype=torch.int32, device=dev),
        status=torch.empty((N,), dtype=torch.int32, device=dev),
        bounds=bounds,
    )
    _lib.call("be_label_overlap_i32", lab_true.data_ptr(), lab_pred.data_ptr(), N, n_px, row_len, r["keys"].data_ptr(),
              r["counts"].data_ptr(), capacity, r["area_true"].data_ptr(), cap_true, r["area_pred"].data_ptr(),
              cap_pred, r["max_lbl"].data_ptr(), r["status"].data_ptr(), _p(bounds), _stream())
    return r


@_on_tensor_device
def match_counts(table: dict, thr: torch.Tensor, ws: Workspace = _default_ws) -> torch.Tensor:
    """→ int64 [N, n_thr, 2] {pairs with IoU >= thr, extra partners}."""
    _check(thr, torch.float64, "thr")
    N, capacity = table["keys"].shape
    cap_true, cap_pred = table["area_true"].shape[1], table["area_pred"].shape[1]
    n_thr = thr.numel()
    out = torch.empty((N, n_thr, 2), dtype=torch.int64, device=thr.device)
    w = ws.get(_lib.query("be_match_workspace_bytes", N, cap_true, cap_pred, n_thr), thr.device)
    _lib.call("be_match_counts", table["keys"].data_ptr(), table["counts"].data_ptr(), N, capacity,
              table["area_true"].data_ptr(), cap_true, table["area_pred"].data_ptr(), cap_pred, thr.data_ptr(), n_thr,
              out.data_ptr(), w.data_ptr(), _p(table.get("bounds")), _stream())
    return out


@_on_tensor_device
def screen_rows(counts: torch.Tensor, table_: dict, thr: torch.Tensor, rows: torch.Tensor, table_model: torch.Tensor,
                flags: torch.Tensor):
    """Write the batch's rows {tp per threshold, n_true, n_pred} into one model's [n_images, n_thr + 2] int64 block of the
    screening table at image indices `rows` (int64, device); sticky status flags go to `flags` (int32 [2])."""
    _check(counts, torch.int64, "counts")
    _check(rows, torch.int64, "rows")
    _check(table_model, torch.int64, "table_model")
    _check(flags, torch.int32, "flags")
    N, n_thr = counts.shape[0], counts.shape[1]
    if table_model.shape[-1] != n_thr + 2 or rows.numel() != N:
        raise ValueError("table block / rows do not match the counts")
    _lib.call("be_screen_rows_i64", counts.data_ptr(), table_["max_lbl"].data_ptr(), table_["status"].data_ptr(), N, n_thr,
              thr.data_ptr(), rows.data_ptr(), table_model.data_ptr(), flags.data_ptr(), _stream())


@_on_tensor_device
def label_stats(labels: torch.Tensor, cap: int):
    """labels (N,[Z,]Y,X) int32 → (sums int64 [N,cap,10], bbox int32 [N,cap,6], status int32 [N])."""
    _check(labels, torch.int32, "labels")
    if labels.dim() == 3:
        N, Y, X = labels.shape
        Z = 1
    else:
        N, Z, Y, X = labels.shape
    dev = labels.device
    sums = torch.empty((N, cap, 10), dtype=torch.int64, device=dev)
    bbox = torch.empty((N, cap, 6), dtype=torch.int32, device=dev)
    status = torch.empty((N,), dtype=torch.int32, device=dev)
    _lib.call("be_label_stats_i32", labels.data_ptr(), N, Z, Y, X, cap, sums.data_ptr(), bbox.data_ptr(),
              status.data_ptr(), _stream())
    return sums, bbox, status


@_on_tensor_device
def binary_confusion(pred: torch.Tensor, target: torch.Tensor, thresholds: Optional[Tuple[float, float]] = None):
    """→ (counts int64 [4] {tp, fp, fn, tn}, minmax fp32 [4]) on the device.  thresholds=None: the reference's adaptive rule."""
    _check(pred, torch.float32, "pred")
    _check(target, torch.float32, "target")
    if pred.numel() != target.numel():
        raise ValueError("pred and target differ in size")
    counts = torch.empty((4,), dtype=torch.int64, device=pred.device)
    minmax = torch.empty((4,), dtype=torch.float32, device=pred.device)
    tp, tt = thresholds if thresholds is not None else (0.0, 0.0)
    _lib.call("be_binary_confusion_f32", pred.data_ptr(), target.data_ptr(), pred.numel(), int(thresholds is None),
              float(tp), float(tt), minmax.data_ptr(), counts.data_ptr(), _stream())
    return counts, minmax


@_on_tensor_device
def remove_small_objects(mask: torch.Tensor, min_size: int, ws: Workspace = _default_ws) -> torch.Tensor:
    """skimage `remove_small_objects(mask, min_size)` (connectivity 1) for (N, Y, X) uint8 masks → uint8 mask."""
    _check(mask, torch.uint8, "mask")
    labels, n = ccl(mask, 1, ws=ws)
    N = mask.shape[0]
    n_px = mask.numel() // N
    cap = (n_px + 1) // 2 + 1   # most 4-connected components an image can hold (a checkerboard): no look at `n` on the host
    area = torch.empty((N, cap), dtype=torch.int32, device=mask.device)
    status = torch.empty((N,), dtype=torch.int32, device=mask.device)
    out = torch.empty_like(mask)
    _lib.call("be_label_area_i32", labels.data_ptr(), N, n_px, cap, area.data_ptr(), status.data_ptr(), _stream())
    _lib.call("be_keep_large_u8", labels.data_ptr(), area.data_ptr(), N, n_px, cap, int(min_size), out.data_ptr(), _stream())
    return out


def disk_half_widths(radius: int) -> np.ndarray:
    """Rows of skimage `morphology.disk(radius)`: half width of row dy = largest dx with dx^2 + dy^2 <= radius^2."""
    r = int(radius)
    return np.array([int(np.floor(np.sqrt(r * r - dy * dy))) for dy in range(-r, r + 1)], dtype=np.int32)


@_on_tensor_device
def binary_closing(mask: torch.Tensor, half_widths: np.ndarray, ws: Workspace = _default_ws) -> torch.Tensor:
    """scipy `binary_closing(mask, structure)` for (N, Y, X) uint8 masks; structure given by its row half widths."""
    import ctypes

    _check(mask, torch.uint8, "mask")
    N, Y, X = mask.shape
    hw = np.ascontiguousarray(half_widths, dtype=np.int32)
    radius = (hw.size - 1) // 2
    out = torch.empty_like(mask)
    w = ws.get(_lib.query("be_binary_closing_workspace_bytes", N, Y, X), mask.device)
    _lib.call("be_binary_closing_u8", mask.data_ptr(), out.data_ptr(), N, Y, X, hw.ctypes.data_as(ctypes.c_void_p), radius,
              w.data_ptr(), _stream())
    return out


@_on_tensor_device
def distance_transform_edt(mask: torch.Tensor, ws: Workspace = _default_ws) -> torch.Tensor:
    """scipy `distance_transform_edt(mask)` for (N, Y, X) uint8 masks → fp64 distances."""
    _check(mask, torch.uint8, "mask")
    N, Y, X = mask.shape
    dist = torch.empty(mask.shape, dtype=torch.float64, device=mask.device)
    w = ws.get(_lib.query("be_edt_workspace_bytes", N, Y, X), mask.device)
    _lib.call("be_edt_u8_f64", mask.data_ptr(), dist.data_ptr(), N, Y, X, w.data_ptr(), _stream())
    return dist


@_on_tensor_device
def distance_transform_edt_sq(mask: torch.Tensor, ws: Workspace = _default_ws) -> torch.Tensor:
    """Square of scipy `distance_transform_edt(mask)` as exact int32, for (N, Y, X) uint8 masks."""
    _check(mask, torch.uint8, "mask")
    N, Y, X = mask.shape
    d2 = torch.empty(mask.shape, dtype=torch.int32, device=mask.device)
    w = ws.get(_lib.query("be_edt_workspace_bytes", N, Y, X), mask.device)
    _lib.call("be_edt_sq_u8_i32", mask.data_ptr(), d2.data_ptr(), N, Y, X, w.data_ptr(), _stream())
    return d2


@_on_tensor_device
def peak_local_max(dist2: torch.Tensor, mask: torch.Tensor, min_distance: int, ws: Workspace = _default_ws) -> torch.Tensor:
    """skimage `peak_local_max(dist, min_distance=d, labels=mask)` on one (Y, X) int32 image (the squared distance) →
    uint8 image with 1 at the selected coordinates."""
    _check(dist2, torch.int32, "dist2")
    _check(mask, torch.uint8, "mask")
    if dist2.dim() != 2 or mask.shape != dist2.shape:
        raise ValueError("peak_local_max: dist2 and mask must be (Y, X) tensors of one shape")
    Y, X = dist2.shape
    peaks = torch.empty_like(mask)
    w = ws.get(_lib.query("be_peak_local_max_workspace_bytes", Y, X), mask.device)
    _lib.call("be_peak_local_max_i32", dist2.data_ptr(), mask.data_ptr(), peaks.data_ptr(), Y, X, int(min_distance),
              w.data_ptr(), _stream())
    return peaks


@_on_tensor_device
def watershed(height: torch.Tensor, markers: torch.Tensor, mask: torch.Tensor, negate: bool = False,
              ws: Workspace = _default_ws) -> torch.Tensor:
    """skimage `watershed(image, markers, mask=mask)` (connectivity 1) on one (Y, X) image: int32 flood order (`negate`
    floods by -height), int32 markers, uint8 mask → int32 labels."""
    _check(height, torch.int32, "height")
    _check(markers, torch.int32, "markers")
    _check(mask, torch.uint8, "mask")
    if height.dim() != 2 or markers.shape != height.shape or mask.shape != height.shape:
        raise ValueError("watershed: height, markers and mask must be (Y, X) tensors of one shape")
    Y, X = height.shape
    regions, n_regions = ccl(mask[None], 1, ws=ws)
    out = torch.empty_like(markers)
    w = ws.get(_lib.query("be_watershed_workspace_bytes", Y, X), mask.device)
    _lib.call("be_watershed_i32", height.data_ptr(), int(bool(negate)), markers.data_ptr(), regions.data_ptr(),
              n_regions.data_ptr(), out.data_ptr(), Y, X, w.data_ptr(), _stream())
    return out


def instances_from_mask(mask: torch.Tensor, min_size: int = 300, radius: int = 4, min_distance: int = 8,
                        ws: Workspace = _default_ws):
    """ref: analysis_deployment.py:160-176 `_prob_to_instances` after the threshold, on one (Y, X) uint8 mask:
    remove_small_objects → binary_closing(disk) → distance transform → peak_local_max → label(peaks) → watershed.
    Returns (labels int32 (Y, X), number of seeds as a device tensor).  An empty mask comes back as zeros without a look
    at it on the host: every stage maps nothing to nothing."""
    _check(mask, torch.uint8, "mask")
    if mask.dim() != 2:
        raise ValueError("instances_from_mask: mask must be a (Y, X) tensor")
    kept = remove_small_objects(mask[None], min_size, ws=ws)
    closed = binary_closing(kept, disk_half_widths(radius), ws=ws)
    d2 = distance_transform_edt_sq(closed, ws=ws)
    peaks = peak_local_max(d2[0], closed[0], min_distance, ws=ws)
    markers, n_seeds = ccl(peaks[None], 2, ws=ws)
    return watershed(d2[0], markers[0], closed[0], negate=True, ws=ws), n_seeds
