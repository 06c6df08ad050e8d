"""Pins the oracle: (1) hand-derived known answers, written out literally here and independent of the oracle's code;
(2) the committed vectors of tests/golden (made by tests/golden/make_golden.py) must keep reproducing.  CPU only."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from oracle import paths as P

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

MASK_6x6 = np.array([
    [1, 0, 0, 1, 1, 0],
    [0, 1, 0, 0, 1, 0],
    [0, 0, 0, 0, 0, 0],
    [1, 1, 0, 0, 0, 1],
    [0, 1, 0, 0, 1, 0],
    [0, 0, 0, 1, 0, 0],
], dtype=np.uint8)
# 4-connectivity: numbered in raster order of each component's first pixel
LABELS_6x6_C1 = np.array([
    [1, 0, 0, 2, 2, 0],
    [0, 3, 0, 0, 2, 0],
    [0, 0, 0, 0, 0, 0],
    [4, 4, 0, 0, 0, 5],
    [0, 4, 0, 0, 6, 0],
    [0, 0, 0, 7, 0, 0],
], dtype=np.int32)
# 8-connectivity: the diagonal pairs (0,0)-(1,1) and (3,5)-(4,4)-(5,3) merge
LABELS_6x6_C2 = np.array([
    [1, 0, 0, 2, 2, 0],
    [0, 1, 0, 0, 2, 0],
    [0, 0, 0, 0, 0, 0],
    [3, 3, 0, 0, 0, 4],
    [0, 3, 0, 0, 4, 0],
    [0, 0, 0, 4, 0, 0],
], dtype=np.int32)


def test_label_known_answer_6x6():
    assert np.array_equal(O.label(MASK_6x6, 1), LABELS_6x6_C1)
    assert np.array_equal(O.label(MASK_6x6, 2), LABELS_6x6_C2)
    assert np.array_equal(O.label(MASK_6x6), LABELS_6x6_C2)  # skimage default: connectivity = ndim


def test_quantile_known_answers():
    x = np.array([4.0, 1.0, 3.0, 2.0, np.nan, 5.0], np.float32)  # sorted valid: 1 2 3 4 5
    q = O.sample_percentiles(x, (0.0, 25.0, 50.0, 90.0, 100.0)).ravel()
    # virtual index q*(n-1): 0, 1, 2, 3.6, 4  →  1, 2, 3, 4 + 0.6, 5
    np.testing.assert_allclose(q, [1.0, 2.0, 3.0, 4.6, 5.0], rtol=0, atol=1e-15)
    y = O.scale_range(np.array([0.0, 5.0, 10.0], np.float32), 0.0, 100.0, eps=0.0)
    assert np.array_equal(y, np.array([0.0, 0.5, 1.0], np.float32))


def test_zmuv_and_pointwise_known_answers():
    x = np.array([[1.0, 3.0], [5.0, 7.0]], np.float32)  # mean 4, population var 5
    y = O.zero_mean_unit_variance(x, eps=0.0)
    np.testing.assert_allclose(y, (x - 4.0) / np.sqrt(5.0), rtol=1e-6)
    assert np.array_equal(O.binarize(np.array([0.5, 0.50001, 0.4], np.float32), 0.5), [0, 1, 0])  # strict >
    assert O.sigmoid(np.array([0.0], np.float32))[0] == np.float32(0.5)
    assert np.array_equal(O.clip(np.array([-1.0, 0.3, 9.0], np.float32), 0, 1), np.array([0, 0.3, 1], np.float32))


def test_blocking_known_answer_1d():
    # S=10, block 6 incl. halo 1 → inner 4; inner slices (0,4) (4,8) (6,10): the last one is shifted back
    assert O.split_axis_into_blocks(10, 6, 1, 1) == [(0, 4), (4, 8), (6, 10)]
    x = np.arange(10, dtype=np.float32)[None]
    blocks = [d[0].tolist() for _, d in O.split_into_blocks(x, (6,), (1,))]
    assert blocks == [[1, 0, 1, 2, 3, 4],        # left halo reflected (edge sample not repeated)
                      [3, 4, 5, 6, 7, 8],
                      [5, 6, 7, 8, 9, 8]]        # right halo reflected
    tiles = np.array(blocks, np.float32)[:, None, :] * 10
    out = O.from_blocks(tiles, (1, 1, 10), (6,), (1,))[0, 0]
    assert out.tolist() == [0, 10, 20, 30, 40, 50, 60, 70, 80, 90]


def test_fibsem_tiling_known_answer_1d():
    # S=10, tile 6, overlap 2 → stride 4 → origins 0, 4, 8; the last tile has 2 valid samples and is reflect-padded
    x = np.arange(10, dtype=np.float32)
    assert O.fibsem_tile_coords((10,), 6, 2) == [(0,), (4,), (8,)]
    assert O.fibsem_tiles(x, 6, 2).tolist() == [[0, 1, 2, 3, 4, 5], [4, 5, 6, 7, 8, 9], [8, 9, 8, 9, 8, 9]]
    # blending constant tiles returns the constant; blending tile index shows the weights
    ones = np.ones((3, 6), np.float32)
    assert np.array_equal(O.fibsem_blend(ones * 7, (10,), 6, 2), np.full(10, 7, np.float32))
    w = np.exp(-2 * np.linspace(-1, 1, 6) ** 2).astype(np.float32).astype(np.float64)
    idx = np.stack([np.full(6, k, np.float32) for k in range(3)])
    got = O.fibsem_blend(idx, (10,), 6, 2)
    want4 = (0 * w[4] + 1 * w[0]) / (w[4] + w[0])  # output 4: tile 0 offset 4 and tile 1 offset 0
    assert got[3] == 0 and got[6] == 1 and abs(got[4] - want4) < 1e-6


def test_average_precision_known_answers():
    t = np.zeros((4, 8), np.int32)
    p = np.zeros((4, 8), np.int32)
    t[0:2, 0:4] = 1   # 8 px
    p[0:2, 0:3] = 1   # 6 px inside → IoU 6/8 = 0.75
    t[2:4, 4:8] = 2   # 8 px
    p[3:4, 4:8] = 2   # 4 px inside → IoU 0.5
    p[0, 7] = 3       # false positive
    ap, tp, fp, fn = O.average_precision([t], [p], threshold=[0.5, 0.75, 0.9])
    assert tp.tolist() == [[2, 1, 0]] and fp.tolist() == [[1, 2, 3]] and fn.tolist() == [[0, 1, 2]]
    np.testing.assert_allclose(ap, [[2 / 3, 1 / 4, 0.0]], rtol=1e-7)
    # IoU exactly 0.5 twice for one true object: the assignment can use only one of them
    t2 = np.zeros((2, 8), np.int32)
    p2 = np.zeros((2, 8), np.int32)
    t2[:, 0:8] = 1
    p2[:, 0:4], p2[:, 4:8] = 1, 2
    assert O.average_precision([t2], [p2], threshold=[0.5])[1].tolist() == [[1]]


def test_regionprops_known_answer():
    lab = np.zeros((6, 8), np.int32)
    lab[1:3, 2:6] = 1  # 2 x 4 rectangle: centroid (1.5, 3.5); variances (2^2-1)/12 and (4^2-1)/12
    r = O.regionprops_table(lab, pixel_size_nm=1000.0)
    assert r["label"] == [1] and r["area_um2"] == [8.0] and r["centroid_y"] == [1.5] and r["centroid_x"] == [3.5]
    l1, l2 = 15 / 12, 3 / 12
    np.testing.assert_allclose(r["eccentricity"], [np.sqrt(1 - l2 / l1)], rtol=1e-12)
    np.testing.assert_allclose(r["aspect_ratio"], [4 * np.sqrt(l1) / (4 * np.sqrt(l2) + 1e-6)], rtol=1e-12)


def test_committed_vectors_reproduce():
    z = np.load(os.path.join(G, "pre_tiles.npz"))
    img = z["img"]
    assert np.array_equal(O.scale_range(img, 1.0, 99.8, axes=(1, 2, 3)), z["scale_range"])
    assert np.array_equal(O.fibsem_normalize(img[0, 0]), z["fibsem_norm"])
    assert np.array_equal(P.crop_pre(img, (48, 48), (8, 8)), z["crop_tiles"])
    assert np.array_equal(P.c2_pre(img, 48, 16), z["fibsem_tiles"])
    np.testing.assert_allclose(O.zero_mean_unit_variance(img, axes=(1, 2, 3)), z["zmuv"], rtol=1e-6, atol=1e-6)
    s = np.load(os.path.join(G, "stitch_post.npz"))
    prob, mask, labels = P.c2_post(s["blend_logits"], (96, 120), 48, 16, 2)
    assert np.array_equal(mask, s["blend_mask"]) and np.array_equal(labels, s["blend_labels"])
    np.testing.assert_allclose(prob, s["blend_prob"], rtol=1e-6)
    _, cmask, clabels = P.crop_post(s["crop_logits"], (1, 1, 96, 120), (48, 48), (8, 8))
    assert np.array_equal(cmask, s["crop_mask"]) and np.array_equal(clabels, s["crop_labels"])
    l = np.load(os.path.join(G, "labels.npz"))
    for key, (m, c) in {"l2_c1": ("m2", 1), "l2_c2": ("m2", 2), "l3_c1": ("m3", 1), "l3_c2": ("m3", 2), "l3_c3": ("m3", 3)}.items():
        assert np.array_equal(O.label(l[m], c), l[key])
    mt = np.load(os.path.join(G, "metric.npz"))
    assert np.array_equal(O.label_overlap(mt["true"], mt["pred"]), mt["overlap"])
    ap, tp, fp, fn = O.average_precision([mt["true"]], [mt["pred"]], threshold=[0.5, 0.75, 0.9])
    assert np.array_equal(tp, mt["tp"]) and np.array_equal(ap, mt["ap"], equal_nan=True)
    z = np.load(os.path.join(G, "instances.npz"))
    for name in ("blobs", "twins"):
        parts = {}
        assert np.array_equal(O.prob_to_instances(z[f"{name}_prob"], ties="fifo", parts=parts), z[f"{name}_fifo"])
        assert np.array_equal(O.prob_to_instances(z[f"{name}_prob"], ties="heap"), z[f"{name}_heap"])
        assert np.array_equal(parts["coords"], z[f"{name}_peaks"]) and np.array_equal(parts["markers"], z[f"{name}_markers"])
        assert np.array_equal(parts["closed"], z[f"{name}_closed"])
    assert np.array_equal(z["blobs_fifo"], z["blobs_heap"])            # no tied seeds there
    assert 0 < (z["twins_fifo"] != z["twins_heap"]).sum() < 200         # tied seeds: the two orders part on a few boundary pixels
