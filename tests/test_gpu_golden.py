"""The CUDA path against the committed golden vectors (tests/golden/*.npz) and the hand-derived known answers."""
import os

import numpy as np
import pytest
import torch

from bioengine_worker_b200 import geometry as G
from tests.test_oracle_golden import LABELS_6x6_C1, LABELS_6x6_C2, MASK_6x6

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def K():
    from bioengine_worker_b200 import kernels

    return kernels


def test_known_answer_6x6_labels(dev):
    m = torch.from_numpy(MASK_6x6[None]).to(dev)
    assert np.array_equal(K().ccl(m, 1)[0].cpu().numpy()[0], LABELS_6x6_C1)
    lab, n = K().ccl(m, 2)
    assert np.array_equal(lab.cpu().numpy()[0], LABELS_6x6_C2) and n.item() == 4


def test_known_answer_quantiles(dev):
    x = torch.tensor([4.0, 1.0, 3.0, 2.0, float("nan"), 5.0], device=dev)
    for qs, want in (((0.0, 0.25), (1.0, 2.0)), ((0.5, 0.9), (3.0, 4.6)), ((1.0,), (5.0,))):
        got, _ = K().quantile(x, 1, torch.tensor(qs, dtype=torch.float64, device=dev))
        np.testing.assert_allclose(got.cpu().numpy()[0], want, rtol=0, atol=1e-15)


def test_golden_pre_and_tiles(dev):
    z = np.load(os.path.join(GOLD, "pre_tiles.npz"))
    x = torch.from_numpy(z["img"]).to(dev)
    q = torch.tensor([0.01, 0.998], dtype=torch.float64, device=dev)
    quant, norm = K().quantile(x, 1, q, 1e-6)
    assert np.array_equal(quant.cpu().numpy()[0], z["quantiles"][[0, 2]])
    assert np.array_equal(K().normalize(x, norm).cpu().numpy(), z["scale_range"])
    ax = [K().DeviceAxis(a, dev) for a in (G.identity_axis(1), G.bioimageio_axis(96, 48, 8), G.bioimageio_axis(120, 48, 8))]
    assert np.array_equal(K().tile_gather(x, ax, norm=norm, norm_mode=1).cpu().numpy(), z["crop_tiles"])
    fx = [K().DeviceAxis(a, dev) for a in (G.identity_axis(1), G.fibsem_axis(96, 48, 16), G.fibsem_axis(120, 48, 16))]
    assert np.array_equal(K().tile_gather(x, fx, norm=norm, norm_mode=1).cpu().numpy(), z["fibsem_tiles"])
    q2 = torch.tensor([0.01, 0.99], dtype=torch.float64, device=dev)
    _, n2 = K().quantile(x, 1, q2, 1e-6)
    assert np.array_equal(K().normalize(x, n2, clip=(0.0, 1.0)).cpu().numpy()[0, 0], z["fibsem_norm"])
    _, nz = K().meanvar(x, 1, 1e-6)
    np.testing.assert_allclose(K().normalize(x, nz).cpu().numpy(), z["zmuv"], rtol=1e-5, atol=1e-5)


def test_golden_stitch_post_labels(dev):
    s = np.load(os.path.join(GOLD, "stitch_post.npz"))
    fx = [K().DeviceAxis(a, dev) for a in (G.identity_axis(1), G.fibsem_axis(96, 48, 16), G.fibsem_axis(120, 48, 16))]
    w = torch.from_numpy(G.gauss_window_1d(48)).to(dev)
    one = torch.ones(1, dtype=torch.float64, device=dev)
    prob, mask = K().stitch_blend(torch.from_numpy(s["blend_logits"]).to(dev), (1, 1, 96, 120), fx, (one, w, w), 1, 0.5,
                                  want_mask=True)
    assert np.array_equal(prob.cpu().numpy()[0, 0], s["blend_prob"])
    assert np.array_equal(mask.cpu().numpy()[0, 0], s["blend_mask"])
    # mask-only call (band-test fast path) gives the same mask
    _, mask2 = K().stitch_blend(torch.from_numpy(s["blend_logits"]).to(dev), (1, 1, 96, 120), fx, (one, w, w), 1, 0.5,
                                want_out=False, want_mask=True)
    assert torch.equal(mask, mask2)
    assert np.array_equal(K().ccl(mask[0], 2)[0].cpu().numpy()[0], s["blend_labels"])
    ax = [K().DeviceAxis(a, dev) for a in (G.identity_axis(1), G.bioimageio_axis(96, 48, 8), G.bioimageio_axis(120, 48, 8))]
    cprob, cmask = K().stitch_crop(torch.from_numpy(s["crop_logits"]).to(dev), (1, 1, 96, 120), ax, 1, 0.5, want_mask=True)
    assert np.array_equal(cprob.cpu().numpy(), s["crop_prob"]) and np.array_equal(cmask.cpu().numpy(), s["crop_mask"])
    assert np.array_equal(K().ccl(cmask[:, 0], None)[0].cpu().numpy(), s["crop_labels"])


def test_golden_labels_all_connectivities(dev):
    l = np.load(os.path.join(GOLD, "labels.npz"))
    for key, (m, c) in {"l2_c1": ("m2", 1), "l2_c2": ("m2", 2), "l3_c1": ("m3", 1), "l3_c2": ("m3", 2), "l3_c3": ("m3", 3)}.items():
        got = K().ccl(torch.from_numpy(l[m][None].astype(np.uint8)).to(dev), c)[0].cpu().numpy()[0]
        assert np.array_equal(got, l[key]), key


def test_golden_metric(dev):
    from bioengine_worker_b200.metrics import average_precision

    mt = np.load(os.path.join(GOLD, "metric.npz"))
    ap, tp, fp, fn = average_precision([mt["true"]], [mt["pred"]], threshold=[0.5, 0.75, 0.9], device=dev)
    assert np.array_equal(tp, mt["tp"]) and np.array_equal(fp, mt["fp"]) and np.array_equal(fn, mt["fn"])
    assert np.array_equal(ap, mt["ap"], equal_nan=True)


def test_golden_instances(dev):
    """Committed vectors of the fibsem post-processing: closing, squared distance, peaks, markers and the flood."""
    z = np.load(os.path.join(GOLD, "instances.npz"))
    for name in ("blobs", "twins"):
        mask = torch.from_numpy((z[f"{name}_prob"] > 0.5).astype(np.uint8)).to(dev)
        kept = K().remove_small_objects(mask[None], 300)
        closed = K().binary_closing(kept, K().disk_half_widths(4))
        assert np.array_equal(closed.cpu().numpy()[0].astype(bool), z[f"{name}_closed"]), name
        d2 = K().distance_transform_edt_sq(closed)
        assert np.array_equal(d2.cpu().numpy()[0], z[f"{name}_dist2"]), name
        peaks = K().peak_local_max(d2[0], closed[0], 8).cpu().numpy()
        want = np.zeros(peaks.shape, np.uint8)
        want[tuple(z[f"{name}_peaks"].T)] = 1
        assert np.array_equal(peaks, want), name
        got, n = K().instances_from_mask(mask)
        assert np.array_equal(got.cpu().numpy(), z[f"{name}_fifo"]) and int(n.item()) == z[f"{name}_markers"].max(), name
