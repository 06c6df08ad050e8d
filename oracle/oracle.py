"""CPU oracle for the tiled bioimage inference path — TEST INFRASTRUCTURE ONLY.

This file restates, in numpy/scipy, what the reference (aicell-lab/bioengine-worker) computes on the hot
path.  Only `tests/`, `__graft_entry__.smoke()` and the CPU-baseline / `--impl reference` legs of `bench.py`
may import it; nothing under `bioengine-worker_b200/` does.

PARITY, row by row:
  * PINNED to the reference's own code: fibsem normalise (a10), overlap tiling + Gaussian blend (a11), pixel-level
    binary metrics (a16).  apps/fibsem-mito-analysis/analysis_deployment.py and apps/cellpose-finetuning/main.py import
    unmodified once `ray.serve` and `hypha_rpc.utils.schema` are stubbed; tests/golden/make_reference_golden.py runs
    `_infer_tiled`, `analyze` (up to `_prob_to_instances`) and `_binary_threshold`/`_compute_binary_metrics` themselves
    and commits their outputs as tests/golden/ref_*.npz; tests/test_reference_golden.py holds this oracle to them bit for
    bit, tests/test_gpu_reference_golden.py holds the CUDA path to the same files.
  * PARITY UNPINNED for the rows whose arithmetic lives in third-party packages that are not vendored and not
    installable here (bioimageio.core==0.10.0, xarray==2025.1.2, scikit-image, cellpose==4.0.7; pins at
    apps/model-runner/runtime_deployment.py:16-28 and apps/cellpose-finetuning/main.py:3608): scale_range /
    zero_mean_unit_variance / pointwise ops (a6-a9), halo blocking and reassembly (a5), labelling numbering (a13),
    regionprops (a14), average_precision (a15), and the skimage half of the fibsem post-processing (f3: peak_local_max,
    watershed — restated here and in oracle/watershed.c from skimage's published algorithm, including both orders in
    which seeds of equal height may leave its heap).  No reference test holds a numeric known answer for them (§8c).  They
    call the same numpy/scipy routines those packages call, follow their published behaviour, and
    are held to hand-derived known answers in tests/test_oracle_golden.py (6x6 masks, quantile lerp, shifted last block,
    reflect padding, IoU = 0.5 ties, rectangle moments) plus seeded vectors in tests/golden/ (make_golden.py).

Paths cited as `ref:` are relative to the reference tree root.
"""
from __future__ import annotations

import itertools
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import scipy.ndimage as ndi
from scipy.optimize import linear_sum_assignment

# --------------------------------------------------------------------------------------------------
# bioimageio.core pre-/post-processing ops (third-party)
# called through ref: apps/model-runner/runtime_deployment.py:278 / :280
# --------------------------------------------------------------------------------------------------


def _reduce_axes(ndim: int, axes: Optional[Sequence[int]]) -> Tuple[int, ...]:
    return tuple(range(ndim)) if axes is None else tuple(int(a) % ndim for a in axes)


def sample_percentiles(x: np.ndarray, percentiles: Sequence[float], axes: Optional[Sequence[int]] = None) -> np.ndarray:
    """`SamplePercentilesCalculator`: xarray `DataArray.quantile(q, dim=axes)` → `np.nanquantile(..., method="linear")`.

    `q` is handed to numpy as a float64 *array* (as xarray does); with numpy 2 a python-float q would be cast to
    the fp32 array dtype and shift the virtual index by whole elements at n ≈ 1.7e7 (Appendix A.3 (ii)).
    Returns float64 of shape (len(percentiles),) + kept dims (keepdims=True layout for broadcasting).
    """
    q = np.asarray([p / 100.0 for p in percentiles], dtype=np.float64)
    return np.nanquantile(x, q, axis=_reduce_axes(x.ndim, axes), method="linear", keepdims=True)


def scale_range(
    x: np.ndarray,
    min_percentile: float = 0.0,
    max_percentile: float = 100.0,
    axes: Optional[Sequence[int]] = None,
    eps: float = 1e-6,
    arithmetic: str = "float64",
) -> np.ndarray:
    """bioimageio `scale_range`: (x - lower) / (upper - lower + eps), no clipping (Appendix A.2).

    arithmetic="float64": the quantiles are shaped float64 arrays, so the elementwise arithmetic promotes to
    fp64 and the result is cast back to fp32 by the trailing EnsureDtype.  arithmetic="float32": the variant in
    which 0-d fp64 quantiles do not promote an fp32 tensor under numpy 1.26 value-based casting (Appendix A.3 (i)).
    The two differ by <= 1 ulp(fp32) of the operands — the 1e-5 tolerance north_star allows.
    """
    x = np.asarray(x, dtype=np.float32)
    lower, upper = sample_percentiles(x, (min_percentile, max_percentile), axes)
    if arithmetic == "float64":
        return ((x.astype(np.float64) - lower) / (upper - lower + eps)).astype(np.float32)
    lo32 = lower.astype(np.float32)
    den32 = (upper - lower + eps).astype(np.float32)
    return ((x - lo32) / den32).astype(np.float32)


def mean_var(x: np.ndarray, axes: Optional[Sequence[int]] = None) -> Tuple[np.ndarray, np.ndarray]:
    """`MeanVarStdCalculator`: mean = x.mean(axes); c = x - mean; var = dot(c, c)/n (ddof 0), in the tensor dtype."""
    ax = _reduce_axes(x.ndim, axes)
    x = np.asarray(x, dtype=np.float32)
    mean = x.mean(axis=ax, keepdims=True)
    c = x - mean
    n = np.prod([x.shape[a] for a in ax])
    var = (c * c).sum(axis=ax, keepdims=True) / n
    return mean, var


def zero_mean_unit_variance(x: np.ndarray, axes: Optional[Sequence[int]] = None, eps: float = 1e-6) -> np.ndarray:
    """bioimageio `zero_mean_unit_variance`: (x - mean) / (sqrt(var) + eps) in fp32 (Appendix A.2)."""
    x = np.asarray(x, dtype=np.float32)
    mean, var = mean_var(x, axes)
    return ((x - mean) / (np.sqrt(var) + np.float32(eps))).astype(np.float32)


def fixed_zero_mean_unit_variance(x: np.ndarray, mean: float, std: float, eps: float = 1e-6) -> np.ndarray:
    x = np.asarray(x, dtype=np.float32)
    return ((x - np.float32(mean)) / (np.float32(std) + np.float32(eps))).astype(np.float32)


def scale_mean_variance(x: np.ndarray, ref: np.ndarray, axes: Optional[Sequence[int]] = None, eps: float = 1e-6) -> np.ndarray:
    """bioimageio `scale_mean_variance` (Appendix A.2): (x - mean) / (std + eps) * (ref_std + eps) + ref_mean, the
    statistics of x and of the reference tensor taken over `axes` (population variance, like zero_mean_unit_variance)."""
    x = np.asarray(x, dtype=np.float32)
    ref = np.asarray(ref, dtype=np.float32)
    mean, var = mean_var(x, axes)
    rmean, rvar = mean_var(ref, axes)
    std, rstd = np.sqrt(var) + np.float32(eps), np.sqrt(rvar) + np.float32(eps)
    return ((x - mean) / std * rstd + rmean).astype(np.float32)


def clip(x: np.ndarray, lo: float, hi: float) -> np.ndarray:
    """bioimageio `clip` → np.clip."""
    return np.clip(np.asarray(x, dtype=np.float32), np.float32(lo), np.float32(hi))


def scale_linear(x: np.ndarray, gain: float, offset: float) -> np.ndarray:
    """bioimageio `scale_linear`: x * gain + offset in fp32."""
    x = np.asarray(x, dtype=np.float32)
    return (x * np.float32(gain) + np.float32(offset)).astype(np.float32)


def sigmoid(x: np.ndarray) -> np.ndarray:
    """bioimageio `sigmoid`: 1 / (1 + exp(-x)) in fp32."""
    x = np.asarray(x, dtype=np.float32)
    with np.errstate(over="ignore"):
        return (np.float32(1.0) / (np.float32(1.0) + np.exp(-x))).astype(np.float32)


def binarize(x: np.ndarray, threshold: float) -> np.ndarray:
    """bioimageio `binarize`: x > threshold (strict) → bool; the trailing EnsureDtype casts it (uint8 here)."""
    return (np.asarray(x, dtype=np.float32) > np.float32(threshold)).astype(np.uint8)


# --------------------------------------------------------------------------------------------------
# bioimageio.core blocking (third-party) — predict_sample_with_blocking,
# ref: apps/model-runner/runtime_deployment.py:277-278
# --------------------------------------------------------------------------------------------------


def split_axis_into_blocks(S: int, B: int, hl: int, hr: int) -> List[Tuple[int, int]]:
    """Inner slices of one axis: block length B includes the halo; the last inner slice is shifted back."""
    inner = B - hl - hr
    if inner <= 0:
        raise ValueError(f"halo ({hl},{hr}) leaves no inner region in block {B}")
    if S < inner:
        raise ValueError(f"sample axis {S} shorter than the inner block {inner}")
    return [(min(p, S - inner), min(p + inner, S)) for p in range(0, S, inner)]


def split_into_blocks(x: np.ndarray, block: Sequence[int], halo: Sequence[int]):
    """Yield (inner_slices, block_array) over the trailing len(block) axes of x (leading axes kept whole).

    Outer slice = inner ± halo clipped to the sample; the clipped part is reflect-padded
    (`np.pad(mode="reflect")`) so every block has exactly the requested shape.  Blocks come in C order.
    """
    nd = len(block)
    lead = x.ndim - nd
    shape = x.shape[lead:]
    per_axis = [split_axis_into_blocks(shape[a], block[a], halo[a], halo[a]) for a in range(nd)]
    for inner in itertools.product(*per_axis):
        sl, pad = [], []
        for a, (i0, i1) in enumerate(inner):
            o0, o1 = i0 - halo[a], i1 + halo[a]
            c0, c1 = max(o0, 0), min(o1, shape[a])
            sl.append(slice(c0, c1))
            pad.append((c0 - o0, o1 - c1))
        data = x[(slice(None),) * lead + tuple(sl)]
        if any(p != (0, 0) for p in pad):
            data = np.pad(data, [(0, 0)] * lead + pad, mode="reflect")
        yield inner, data


def blocks_tensor(x: np.ndarray, block: Sequence[int], halo: Sequence[int]) -> np.ndarray:
    """All blocks of an (N, C, *spatial) array stacked as (N*T, C, *block) in (n, block-C-order) order."""
    out = []
    for n in range(x.shape[0]):
        for _, data in split_into_blocks(x[n], block, halo):
            out.append(data)
    return np.stack(out, axis=0)


def from_blocks(tiles: np.ndarray, sample_shape: Sequence[int], block: Sequence[int], halo: Sequence[int]) -> np.ndarray:
    """`Sample.from_blocks`: out = full(NaN); out[inner_slice] = block[local inner] in block order (last writer wins).

    tiles: (N*T, C, *block); sample_shape: (N, C, *spatial).
    """
    nd = len(block)
    N, C = sample_shape[:2]
    spatial = sample_shape[2:]
    out = np.full(tuple(sample_shape), np.nan, dtype=tiles.dtype)
    per_axis = [split_axis_into_blocks(spatial[a], block[a], halo[a], halo[a]) for a in range(nd)]
    t = 0
    for n in range(N):
        for inner in itertools.product(*per_axis):
            dst = tuple(slice(i0, i1) for i0, i1 in inner)
            src = tuple(slice(halo[a], block[a] - halo[a]) for a in range(nd))
            out[(n, slice(None)) + dst] = tiles[(t, slice(None)) + src]
            t += 1
    return out


# --------------------------------------------------------------------------------------------------
# fibsem-mito-analysis: in-tree numpy pipeline
# --------------------------------------------------------------------------------------------------


def fibsem_normalize(image_np: np.ndarray) -> np.ndarray:
    """ref: apps/fibsem-mito-analysis/analysis_deployment.py:241-242 (verbatim arithmetic).

    `p1`/`p99` are numpy float64 scalars.  They are applied as fp64 here (what numpy >= 2 does for a float64
    scalar; numpy 1.26 would have stayed in fp32 — same <= 1 ulp difference as in scale_range).
    """
    image_np = np.asarray(image_np, dtype=np.float32)
    p1, p99 = np.percentile(image_np, [1, 99])
    return np.clip((image_np.astype(np.float64) - p1) / (p99 - p1 + 1e-6), 0, 1).astype(np.float32)


def fibsem_window(tile_size: int) -> np.ndarray:
    """ref: analysis_deployment.py:125-127."""
    yy = np.linspace(-1, 1, tile_size)
    xx = np.linspace(-1, 1, tile_size)
    return np.outer(np.exp(-2 * yy**2), np.exp(-2 * xx**2)).astype(np.float32)


def fibsem_tile_coords(shape: Sequence[int], tile_size: int, overlap: int) -> List[Tuple[int, ...]]:
    """ref: analysis_deployment.py:121,149-153 (generalised to n-D: product of range(0, S, stride))."""
    stride = tile_size - overlap
    return list(itertools.product(*[range(0, s, stride) for s in shape]))


def fibsem_tiles(image_norm: np.ndarray, tile_size: int, overlap: int) -> np.ndarray:
    """The padded tiles `process_tile` submits — ref: analysis_deployment.py:132-138.  (T, *tile) fp32."""
    nd = image_norm.ndim
    tiles = []
    for origin in fibsem_tile_coords(image_norm.shape, tile_size, overlap):
        sl = tuple(slice(o, min(o + tile_size, s)) for o, s in zip(origin, image_norm.shape))
        tile = image_norm[sl].copy()
        if any(d < tile_size for d in tile.shape):
            tile = np.pad(tile, [(0, tile_size - d) for d in tile.shape], mode="reflect")
        tiles.append(tile)
    return np.stack(tiles, axis=0) if tiles else np.zeros((0,) + (tile_size,) * nd, np.float32)


def fibsem_blend(probs: np.ndarray, shape: Sequence[int], tile_size: int, overlap: int) -> np.ndarray:
    """Gaussian-blended accumulation — ref: analysis_deployment.py:125-130,140-156, tiles in tile_coords order.

    probs: (T, *tile) fp32 network outputs for the tiles of `fibsem_tiles`.  n-D window = fp32 cast of the
    fp64 outer product of the 1-D exp(-2 t^2) tables (2-D: exactly `fibsem_window`).
    """
    nd = len(shape)
    t1 = np.exp(-2 * np.linspace(-1, 1, tile_size) ** 2)
    win = t1
    for _ in range(nd - 1):
        win = np.multiply.outer(win, t1)
    weight_win = win.astype(np.float32)
    prob_acc = np.zeros(tuple(shape), dtype=np.float64)
    weight_acc = np.zeros(tuple(shape), dtype=np.float64)
    for t, origin in enumerate(fibsem_tile_coords(shape, tile_size, overlap)):
        sl = tuple(slice(o, min(o + tile_size, s)) for o, s in zip(origin, shape))
        ext = tuple(slice(0, s.stop - s.start) for s in sl)
        prob = probs[t].astype(np.float32)[ext]
        w = weight_win[ext]
        prob_acc[sl] += prob * w
        weight_acc[sl] += w
    out = np.zeros(tuple(shape), dtype=np.float64)
    np.divide(prob_acc, weight_acc, out=out, where=weight_acc > 0)
    return out.astype(np.float32)


def label(mask: np.ndarray, connectivity: Optional[int] = None) -> np.ndarray:
    """`skimage.measure.label(mask)` (ref: analysis_deployment.py:175) via scipy with the same structuring element.

    connectivity defaults to mask.ndim (8-conn / 26-conn); numbering is raster order of first pixel (Appendix C).
    """
    mask = np.asarray(mask) != 0
    c = mask.ndim if connectivity is None else connectivity
    lab, _ = ndi.label(mask, structure=ndi.generate_binary_structure(mask.ndim, c))
    return lab.astype(np.int32)


def remove_small_objects(mask: np.ndarray, min_size: int) -> np.ndarray:
    """`skimage.morphology.remove_small_objects(mask, min_size)` (ref: analysis_deployment.py:166), restated: components
    under connectivity 1, sizes by bincount, those with fewer than min_size pixels removed."""
    mask = np.asarray(mask) != 0
    lab, _ = ndi.label(mask, structure=ndi.generate_binary_structure(mask.ndim, 1))
    sizes = np.bincount(lab.ravel())
    too_small = sizes < min_size
    out = mask.copy()
    out[too_small[lab]] = False
    return out


def disk(radius: int) -> np.ndarray:
    """`skimage.morphology.disk(radius)`: (X^2 + Y^2) <= radius^2 on a (2r+1)^2 grid."""
    L = np.arange(-radius, radius + 1)
    X, Y = np.meshgrid(L, L)
    return (X**2 + Y**2) <= radius**2


def binary_closing_disk(mask: np.ndarray, radius: int = 4) -> np.ndarray:
    """ref: analysis_deployment.py:169 — ndi.binary_closing(binary, structure=morphology.disk(4))."""
    return ndi.binary_closing(np.asarray(mask) != 0, structure=disk(radius))


def distance_transform_edt(mask: np.ndarray) -> np.ndarray:
    """ref: analysis_deployment.py:170 — ndi.distance_transform_edt(closed), fp64."""
    return ndi.distance_transform_edt(np.asarray(mask) != 0)


def _exclude_border(labels: np.ndarray, width: int) -> np.ndarray:
    out = labels.copy()
    if width > 0:
        for ax in range(out.ndim):
            sl = [slice(None)] * out.ndim
            sl[ax] = slice(None, width)
            out[tuple(sl)] = 0
            sl[ax] = slice(-width, None)
            out[tuple(sl)] = 0
    return out


def _peak_mask(image: np.ndarray, footprint: np.ndarray, threshold: float, mask: np.ndarray) -> np.ndarray:
    """skimage.feature.peak._get_peak_mask: pixels that equal the maximum over the footprint (edges replicated) and lie
    above the threshold; an image in which every masked pixel is such a maximum has no peaks except pixels a binary
    opening removes."""
    if footprint.size == 1 or image.size == 1:
        return image > threshold
    out = image == ndi.maximum_filter(image, footprint=footprint, mode="nearest")
    if np.all(out[mask]):
        out[:] = False
        out[np.logical_xor(mask, ndi.binary_opening(mask))] = True
    out &= image > threshold
    return out


def _ensure_spacing(coord: np.ndarray, spacing: float) -> np.ndarray:
    """skimage._shared.coord.ensure_spacing with p_norm = inf, restated as the greedy pass it performs: walk the points in
    the given order; a point that is still in play removes every later point at Chebyshev distance < spacing.  (skimage
    walks batches of growing size through a KD-tree; removed points never remove others, so the batches do not change
    the outcome.)"""
    from scipy.spatial import cKDTree

    if len(coord) == 0:
        return coord
    near = cKDTree(coord).query_ball_point(coord, r=spacing, p=np.inf)
    rejected = np.zeros(len(coord), dtype=bool)
    for i, cand in enumerate(near):
        if rejected[i]:
            continue
        for c in cand:
            if c != i and np.abs(coord[c] - coord[i]).max() < spacing:
                rejected[c] = True
    return coord[~rejected]


def peak_local_max(image: np.ndarray, min_distance: int, labels: np.ndarray) -> np.ndarray:
    """`skimage.feature.peak_local_max(image, min_distance=d, labels=labels)` with every other argument at its default
    (ref: analysis_deployment.py:171; third-party, PARITY UNPINNED — restated from skimage/feature/peak.py):
    exclude_border = d on every axis, threshold = image.min(), footprint = ones((2d+1,)*ndim); per label, the bounding box
    of the border-stripped label is cut out, pixels outside the label are set to the lowest float, peaks are pixels equal
    to the window maximum (window clipped to the box) above the threshold, sorted by value (descending, stable, so raster
    order among equals) and thinned to Chebyshev distance >= d.  Returns (n, ndim) coordinates."""
    image = np.asarray(image)
    threshold = image.min()
    footprint = np.ones((2 * min_distance + 1,) * image.ndim, dtype=bool)
    stripped = _exclude_border(labels.astype(int, casting="safe"), min_distance)
    found = []
    for idx, roi in enumerate(ndi.find_objects(stripped)):
        if roi is None:
            continue
        label_mask = labels[roi] == idx + 1
        obj = image[roi].copy()
        obj[np.logical_not(label_mask)] = np.finfo(image.dtype).min if np.issubdtype(image.dtype, np.floating) else np.iinfo(image.dtype).min
        mask = _peak_mask(obj, footprint, threshold, label_mask)
        coord = np.nonzero(mask)
        order = np.argsort(-obj[coord], kind="stable")
        coord = _ensure_spacing(np.transpose(coord)[order], min_distance)
        for ax, s in enumerate(roi):
            coord[:, ax] += s.start
        found.append(coord)
    return np.vstack(found) if found else np.empty((0, image.ndim), dtype=int)


_WS_LIB = None


def _watershed_lib():
    """oracle/watershed.c compiled with gcc into oracle/_build/ (built by __graft_entry__.build(), or here on first use)."""
    global _WS_LIB
    if _WS_LIB is None:
        import ctypes
        import os
        import subprocess

        here = os.path.dirname(os.path.abspath(__file__))
        so = os.path.join(here, "_build", "liboracle_watershed.so")
        src = os.path.join(here, "watershed.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            os.makedirs(os.path.dirname(so), exist_ok=True)
            subprocess.run(["gcc", "-O2", "-shared", "-fPIC", "-o", so, src], check=True)
        lib = ctypes.CDLL(so)
        lib.oracle_watershed.restype = ctypes.c_int
        lib.oracle_watershed.argtypes = [ctypes.c_void_p] * 4 + [ctypes.c_int64, ctypes.c_int64, ctypes.c_int]
        _WS_LIB = lib
    return _WS_LIB


def watershed(image: np.ndarray, markers: np.ndarray, mask: np.ndarray, ties: str = "heap") -> np.ndarray:
    """`skimage.segmentation.watershed(image, markers, mask=mask)` for 2-D inputs (ref: analysis_deployment.py:175); the
    flood itself is oracle/watershed.c.  ties="heap": skimage's binary heap decides between markers that tie in value;
    ties="fifo": first pushed, first popped (what the CUDA flood implements)."""
    image = np.ascontiguousarray(image, dtype=np.float64)
    markers = np.ascontiguousarray(markers, dtype=np.int32)
    mask = np.ascontiguousarray(np.asarray(mask) != 0, dtype=np.uint8)
    assert image.ndim == 2 and markers.shape == image.shape == mask.shape
    out = np.empty_like(markers)
    rc = _watershed_lib().oracle_watershed(image.ctypes.data, markers.ctypes.data, mask.ctypes.data, out.ctypes.data,
                                           image.shape[0], image.shape[1], {"heap": 0, "fifo": 1}[ties])
    if rc != 0:
        raise MemoryError("oracle_watershed")
    return out


def prob_to_instances(prob: np.ndarray, ties: str = "heap", parts: Optional[dict] = None) -> np.ndarray:
    """ref: analysis_deployment.py:160-176 `_prob_to_instances`, line by line."""
    binary = remove_small_objects(prob > 0.5, 300)
    if not binary.any():
        return np.zeros(prob.shape, dtype=np.int32)
    closed = binary_closing_disk(binary, 4)
    dist = distance_transform_edt(closed)
    coords = peak_local_max(dist, 8, closed)
    m = np.zeros(closed.shape, dtype=bool)
    if len(coords):
        m[tuple(coords.T)] = True
    markers = label(m)
    if parts is not None:
        parts.update(closed=closed, dist=dist, coords=coords, markers=markers)
    return watershed(-dist, markers, closed, ties).astype(np.int32)


def regionprops_table(labels: np.ndarray, pixel_size_nm: float = 5.0) -> Dict[str, list]:
    """The property loop of ref: analysis_deployment.py:256-271, with skimage's definitions (Appendix C):

    area = pixel count; centroid = mean coordinate; inertia tensor eigenvalues l1 >= l2 from the central second
    moments / area; axis_major/minor_length = 4 sqrt(l); eccentricity = sqrt(1 - l2/l1) (0 if l1 == 0).
    """
    pixel_um = pixel_size_nm / 1000.0
    out = {k: [] for k in ("label", "area_um2", "aspect_ratio", "eccentricity", "centroid_y", "centroid_x")}
    n = int(labels.max()) if labels.size else 0
    objs = ndi.find_objects(labels, max_label=n)
    for l in range(1, n + 1):
        if objs[l - 1] is None:
            continue
        sl = objs[l - 1]   # skimage works on the bounding-box crop as well; the coordinates are put back in the image frame
        ys, xs = np.nonzero(labels[sl] == l)
        ys, xs = ys + sl[0].start, xs + sl[1].start
        area = float(ys.size)
        cy, cx = ys.mean(), xs.mean()
        dy, dx = ys - cy, xs - cx
        mu20, mu02, mu11 = (dy * dy).sum() / area, (dx * dx).sum() / area, (dy * dx).sum() / area
        # skimage inertia_tensor for 2-D: [[mu02, -mu11], [-mu11, mu20]] (in its x/y naming) — same eigenvalues
        tr, det = mu20 + mu02, mu20 * mu02 - mu11 * mu11
        disc = np.sqrt(max(tr * tr / 4.0 - det, 0.0))
        l1, l2 = tr / 2.0 + disc, max(tr / 2.0 - disc, 0.0)
        major, minor = 4.0 * np.sqrt(l1), 4.0 * np.sqrt(l2)
        ecc = 0.0 if l1 == 0 else float(np.sqrt(max(1.0 - l2 / l1, 0.0)))
        out["label"].append(l)
        out["area_um2"].append(area * pixel_um**2)
        out["aspect_ratio"].append(float(major / (minor + 1e-6)))
        out["eccentricity"].append(ecc)
        out["centroid_y"].append(float(cy))
        out["centroid_x"].append(float(cx))
    return out


# --------------------------------------------------------------------------------------------------
# cellpose.metrics.average_precision (third-party, cellpose 4.0.7)
# call site ref: apps/cellpose-finetuning/main.py:2000-2019
# --------------------------------------------------------------------------------------------------


def label_overlap(x: np.ndarray, y: np.ndarray) -> np.ndarray:
    """`_label_overlap`: overlap[i, j] = #pixels with x == i and y == j, shape (1+x.max(), 1+y.max())."""
    x = x.ravel().astype(np.int64)
    y = y.ravel().astype(np.int64)
    nx, ny = 1 + int(x.max()), 1 + int(y.max())
    return np.bincount(x * ny + y, minlength=nx * ny).reshape(nx, ny).astype(np.uint64)


def intersection_over_union(masks_true: np.ndarray, masks_pred: np.ndarray) -> np.ndarray:
    """`_intersection_over_union` (Appendix B.1): fp64 iou incl. background row/col, NaN → 0."""
    overlap = label_overlap(masks_true, masks_pred).astype(np.float64)
    n_pixels_pred = overlap.sum(axis=0, keepdims=True)
    n_pixels_true = overlap.sum(axis=1, keepdims=True)
    with np.errstate(invalid="ignore", divide="ignore"):
        iou = overlap / (n_pixels_pred + n_pixels_true - overlap)
    iou[np.isnan(iou)] = 0.0
    return iou


def true_positive(iou: np.ndarray, th: float) -> int:
    """`_true_positive` (Appendix B.2): Hungarian assignment on -(iou >= th) - iou / (2 n_min)."""
    n_min = min(iou.shape[0], iou.shape[1])
    costs = -(iou >= th).astype(float) - iou / (2 * n_min)
    true_ind, pred_ind = linear_sum_assignment(costs)
    match_ok = iou[true_ind, pred_ind] >= th
    return int(match_ok.sum())


def average_precision(masks_true, masks_pred, threshold=(0.5, 0.75, 0.9)):
    """`cellpose.metrics.average_precision` (Appendix B): returns (ap, tp, fp, fn), each (n_images, n_thr) float32."""
    not_list = False
    if not isinstance(masks_true, list):
        masks_true, masks_pred, not_list = [masks_true], [masks_pred], True
    if not isinstance(threshold, (list, tuple, np.ndarray)):
        threshold = [threshold]
    n, m = len(masks_true), len(threshold)
    ap = np.zeros((n, m), np.float32)
    tp = np.zeros((n, m), np.float32)
    fp = np.zeros((n, m), np.float32)
    fn = np.zeros((n, m), np.float32)
    n_true = np.array([int(np.max(t)) for t in masks_true])
    n_pred = np.array([int(np.max(p)) for p in masks_pred])
    for i in range(n):
        if n_pred[i] > 0:
            iou = intersection_over_union(masks_true[i], masks_pred[i])[1:, 1:]
            for k, th in enumerate(threshold):
                tp[i, k] = true_positive(iou, th) if iou.size else 0
        fp[i] = n_pred[i] - tp[i]
        fn[i] = n_true[i] - tp[i]
        with np.errstate(invalid="ignore", divide="ignore"):
            ap[i] = tp[i] / (tp[i] + fp[i] + fn[i])
    if not_list:
        ap, tp, fp, fn = ap[0], tp[0], fp[0], fn[0]
    return ap, tp, fp, fn


def mean_ap(ap: np.ndarray) -> np.ndarray:
    """ref: apps/cellpose-finetuning/main.py:2008 — np.nanmean over images."""
    with np.errstate(invalid="ignore"):
        return np.nanmean(ap, axis=0)


def binary_metrics(pred: np.ndarray, target: np.ndarray) -> Dict[str, float]:
    """ref: apps/cellpose-finetuning/main.py:1225-1270 — adaptive threshold + TP/FP/FN/TN + ratios."""

    def thr(a):
        return 0.5 if (float(a.min()) >= 0.0 and float(a.max()) <= 1.0) else 0.0

    pr = (pred > thr(pred)).ravel()
    gt = (target > thr(target)).ravel()
    tp = int(np.sum(pr & gt))
    fp = int(np.sum(pr & ~gt))
    fn = int(np.sum(~pr & gt))
    tn = int(np.sum(~pr & ~gt))

    def sd(a, b):
        return float(a / b) if b != 0 else 0.0

    precision, recall = sd(tp, tp + fp), sd(tp, tp + fn)
    return {
        "tp": tp, "fp": fp, "fn": fn, "tn": tn,
        "precision": precision, "recall": recall,
        "f1": sd(2 * precision * recall, precision + recall),
        "pixel_accuracy": sd(tp + tn, tp + tn + fp + fn),
        "iou": sd(tp, tp + fp + fn),
    }


# --------------------------------------------------------------------------------------------------
# synthetic inputs — shared by tests and bench so both arms see identical data
# --------------------------------------------------------------------------------------------------


def gaussian_blobs(shape: Sequence[int], n: int, sigma=(4.0, 10.0), seed: int = 0) -> np.ndarray:
    """Sum of n isotropic Gaussian blobs in [0, ~1], float32; separable per-blob evaluation in a local window."""
    rng = np.random.default_rng(seed)
    out = np.zeros(tuple(shape), np.float32)
    nd = len(shape)
    for _ in range(n):
        c = [rng.uniform(0, s) for s in shape]
        sg = rng.uniform(*sigma)
        r = int(np.ceil(4 * sg))
        sl, axes = [], []
        for a in range(nd):
            lo, hi = max(int(c[a]) - r, 0), min(int(c[a]) + r + 1, shape[a])
            sl.append(slice(lo, hi))
            axes.append(np.exp(-0.5 * ((np.arange(lo, hi) - c[a]) / sg) ** 2).astype(np.float32))
        blob = axes[0]
        for a in range(1, nd):
            blob = np.multiply.outer(blob, axes[a])
        out[tuple(sl)] += blob
    return out


def synthetic_image(shape: Sequence[int], n_blobs: int, seed: int, microscope: bool = True,
                    with_truth: bool = False):
    """C2-style frame: blobs*1000 + 100 + shot-like noise + 0.1 % hot pixels at 65535 (exercises the percentiles).
    with_truth=True also returns the noise-free blob field (ground truth for fitting a read-out / making GT labels)."""
    rng = np.random.default_rng(seed + 7919)
    clean = gaussian_blobs(shape, n_blobs, seed=seed)
    if not microscope:
        img = (clean + 0.1 * rng.standard_normal(clean.shape).astype(np.float32)).astype(np.float32)
    else:
        img = clean * 1000.0 + 100.0
        img = img + np.sqrt(img) * rng.standard_normal(img.shape).astype(np.float32)
        hot = rng.random(img.shape) < 0.001
        img[hot] = 65535.0
        img = img.astype(np.float32)
    return (img, clean) if with_truth else img
