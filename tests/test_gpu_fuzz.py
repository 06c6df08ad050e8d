"""Seeded random geometries through the kernels against the oracle — the cases nobody thought of writing down.
Sizes are small so the CPU side stays in seconds; every draw is reproducible from its index."""
import os

import numpy as np
import pytest
import torch

from bioengine_worker_b200 import geometry as G
from oracle import oracle as O

pytestmark = pytest.mark.gpu
_X = int(os.environ.get("BE_FUZZ", "1"))   # BE_FUZZ=10 runs ten times as many draws


def K():
    from bioengine_worker_b200 import kernels

    return kernels


@pytest.mark.parametrize("case", range(24 * _X))
def test_fuzz_fibsem_gather_blend_label(case, dev):
    rng = np.random.default_rng(1000 + case)
    nd = 2 if case % 3 else 3
    B = int(rng.choice([16, 32, 48, 64] if nd == 2 else [8, 16, 24]))
    ov = int(rng.choice([0, 4, B // 4, B // 2 - 4 if B >= 16 else 0]))
    ov = max(0, min(ov, B // 2))
    lo, hi = (B + 1, 5 * B) if nd == 2 else (B + 1, 3 * B)
    spatial = tuple(int(rng.integers(lo, hi)) for _ in range(nd))
    if rng.random() < 0.5:   # aligned widths take the TMA kernels, the rest the general ones
        spatial = spatial[:-1] + (max(B + 4, spatial[-1] // 4 * 4),)
    C = int(rng.integers(1, 3))
    img = rng.standard_normal((1, C) + spatial).astype(np.float32)
    axes = [G.fibsem_axis(s, B, ov) for s in spatial]
    if nd == 2:
        axes = [G.identity_axis(1)] + axes
    daxes = [K().DeviceAxis(a, dev) for a in axes]
    tiles = K().tile_gather(torch.from_numpy(img).to(dev), daxes)
    want_tiles = np.stack([O.fibsem_tiles(img[0, c], B, ov) for c in range(C)], axis=1)
    assert np.array_equal(tiles.cpu().numpy(), want_tiles), (case, spatial, B, ov)
    logits = (tiles * 3.0 - 0.5).contiguous()
    w1 = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
    wins = (w1,) * 3 if nd == 3 else (torch.ones(1, dtype=torch.float64, device=dev), w1, w1)
    thr = float(rng.choice([0.5, 0.35, 0.8]))
    _, mask = K().stitch_blend(logits, (1, C) + spatial, daxes, wins, epilogue=1, want_out=False, want_mask=True, thr=thr)
    ln = logits.cpu().numpy()
    want = np.stack([O.fibsem_blend(ln[:, c], spatial, B, ov) for c in range(C)])
    want_mask = O.binarize(O.sigmoid(want), thr)
    assert np.array_equal(mask.cpu().numpy()[0], want_mask), (case, spatial, B, ov, thr)
    conn = int(rng.integers(1, nd + 1))
    lab, n = K().ccl(mask[0], conn)   # channels as a batch of masks
    for c in range(C):
        wl = O.label(want_mask[c], conn)
        assert np.array_equal(lab[c].cpu().numpy(), wl), (case, c, conn)
        assert int(n[c]) == int(wl.max())


@pytest.mark.parametrize("case", range(12 * _X))
def test_fuzz_bioimageio_gather_crop(case, dev):
    rng = np.random.default_rng(2000 + case)
    nd = 2 if case % 2 else 3
    block = tuple(int(rng.choice([16, 24, 32, 64] if nd == 2 else [8, 16, 32])) for _ in range(nd))
    halo = tuple(int(rng.integers(0, b // 4 + 1)) for b in block)
    spatial = tuple(int(rng.integers(b, 4 * b)) for b in block)
    N, C = int(rng.integers(1, 3)), int(rng.integers(1, 3))
    x = rng.standard_normal((N, C) + spatial).astype(np.float32)
    axes = [G.bioimageio_axis(s, b, h) for s, b, h in zip(spatial, block, halo)]
    if nd == 2:
        axes = [G.identity_axis(1)] + axes
    daxes = [K().DeviceAxis(a, dev) for a in axes]
    tiles = K().tile_gather(torch.from_numpy(x).to(dev), daxes)
    assert np.array_equal(tiles.cpu().numpy(), O.blocks_tensor(x, block, halo)), (case, spatial, block, halo)
    out, _ = K().stitch_crop(tiles, (N, C) + spatial, daxes)
    assert np.array_equal(out.cpu().numpy(), x), (case, spatial, block, halo)   # identity network: exact round trip


@pytest.mark.parametrize("case", range(10 * _X))
def test_fuzz_quantiles(case, dev):
    rng = np.random.default_rng(3000 + case)
    n_seg = int(rng.integers(1, 4))
    seg_len = int(rng.choice([257, 4099, (1 << 18) + int(rng.integers(0, 9)), (1 << 19) + 3, 300000]))
    kind = case % 5
    x = rng.standard_normal((n_seg, seg_len)).astype(np.float32)
    if kind == 1:
        x = np.round(x * 3)
    elif kind == 2:
        x[rng.random(x.shape) < 0.3] = np.nan
    elif kind == 3:
        x = np.abs(x) ** 6 * 1e4
    elif kind == 4:
        x[rng.random(x.shape) < 0.6] = 0.0
    q = np.sort(rng.random(2))
    if case % 4 == 0:
        q = np.array([0.0, 1.0])
    xd = torch.from_numpy(x).to(dev)
    got, _ = K().quantile(xd, n_seg, torch.from_numpy(q).to(dev), 1e-6)
    import warnings
    with warnings.catch_warnings(), np.errstate(all="ignore"):
        warnings.simplefilter("ignore")
        want = np.nanquantile(x, q, axis=1, method="linear").T
    assert np.array_equal(got.cpu().numpy(), want, equal_nan=True), (case, n_seg, seg_len, kind, q)


@pytest.mark.parametrize("case", range(24 * _X))
def test_fuzz_peaks_and_flood(case, dev):
    """Random small masks (blobs, bars, noise, borders) through peak_local_max and the marker flood, random integer
    landscapes through the flood alone — against the oracle's restatement of skimage (first-in-first-out tie order)."""
    import scipy.ndimage as ndi

    rng = np.random.default_rng(7000 + case)
    H, W = int(rng.integers(20, 140)), int(rng.integers(20, 160))
    kind = case % 4
    if kind == 0:     # smooth blobs
        mask = ndi.gaussian_filter(rng.standard_normal((H, W)), rng.uniform(2, 6)) > rng.uniform(-0.02, 0.05)
    elif kind == 1:   # bars and boxes: plateaus of the distance, tied seeds
        mask = np.zeros((H, W), dtype=bool)
        for _ in range(int(rng.integers(2, 8))):
            y, x = int(rng.integers(0, H - 4)), int(rng.integers(0, W - 4))
            mask[y:y + int(rng.integers(3, 30)), x:x + int(rng.integers(3, 60))] = True
    elif kind == 2:   # dense noise, then closed: thin and ragged
        mask = ndi.binary_closing(rng.random((H, W)) < rng.uniform(0.2, 0.6), iterations=1)
    else:             # everything, or nearly
        mask = rng.random((H, W)) < 0.995
    d = int(rng.choice([1, 2, 3, 8, 8, 8]))
    dist = O.distance_transform_edt(mask)
    want = O.peak_local_max(dist, d, mask)
    m = torch.from_numpy(mask.astype(np.uint8)).to(dev)
    d2 = K().distance_transform_edt_sq(m[None])[0]
    peaks = K().peak_local_max(d2, m, d).cpu().numpy()
    want_img = np.zeros((H, W), dtype=np.uint8)
    if len(want):
        want_img[tuple(want.T)] = 1
    assert np.array_equal(peaks, want_img), (case, kind, d)
    markers = O.label(want_img)
    flood = O.watershed(-dist, markers, mask, ties="fifo")
    got = K().watershed(d2, torch.from_numpy(markers).to(dev), m, negate=True).cpu().numpy()
    assert np.array_equal(got, flood), (case, kind, d)
    # an arbitrary landscape with arbitrary seeds (some of them several pixels large)
    h = rng.integers(0, int(rng.choice([2, 5, 50, 5000])), size=(H, W)).astype(np.int32)
    mk = np.zeros((H, W), dtype=np.int32)
    ys, xs = np.nonzero(mask)
    if len(ys):
        for l in range(1, int(rng.integers(1, 12)) + 1):
            j = int(rng.integers(0, len(ys)))
            y, x = ys[j], xs[j]
            blob = np.zeros((H, W), dtype=bool)
            blob[y:y + int(rng.integers(1, 3)), x:x + int(rng.integers(1, 4))] = True
            mk[blob & mask & (mk == 0)] = l
    flood = O.watershed(h.astype(np.float64), mk, mask, ties="fifo")
    got = K().watershed(torch.from_numpy(h).to(dev), torch.from_numpy(mk).to(dev), m).cpu().numpy()
    assert np.array_equal(got, flood), (case, kind, "landscape")
