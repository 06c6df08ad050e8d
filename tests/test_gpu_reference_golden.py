"""The CUDA path against fixtures written by the REFERENCE's own functions (tests/golden/ref_*.npz; see
tests/golden/make_reference_golden.py).  Tiles, masks and counts bit-exact; blended probabilities within 1e-5 relative
(north_star's tolerance for floats) — and in practice bit-exact, which is asserted too."""
import os

import numpy as np
import pytest
import torch

from bioengine_worker_b200 import geometry as G
from oracle import oracle as O
from tests.test_reference_golden import N_INFER_CASES, infer_case

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def K():
    from bioengine_worker_b200 import kernels

    return kernels


@pytest.mark.parametrize("ci", range(N_INFER_CASES))
def test_gather_and_blend_equal_reference_infer_tiled(dev, ci):
    """ref: analysis_deployment.py:108-156 through be_tile_gather_f32 and be_stitch_blend_f32."""
    z = np.load(os.path.join(GOLD, "ref_infer_tiled.npz"))
    (H, W, tile, ov), img, tiles, probs, out = infer_case(z, ci)
    fx = [K().DeviceAxis(a, dev) for a in (G.identity_axis(1), G.fibsem_axis(H, tile, ov), G.fibsem_axis(W, tile, ov))]
    x = torch.from_numpy(img[None, None]).to(dev)
    got_tiles = K().tile_gather(x, fx).cpu().numpy()
    assert np.array_equal(got_tiles[:, 0], tiles)
    w = torch.from_numpy(G.gauss_window_1d(tile)).to(dev)
    one = torch.ones(1, dtype=torch.float64, device=dev)
    p = torch.from_numpy(probs[:, None]).to(dev)
    got, mask = K().stitch_blend(p, (1, 1, H, W), fx, (one, w, w), 0, 0.5, want_mask=True)
    got = got.cpu().numpy()[0, 0]
    np.testing.assert_allclose(got, out, rtol=1e-5, atol=0)
    assert np.array_equal(got, out)
    assert np.array_equal(mask.cpu().numpy()[0, 0], (out > np.float32(0.5)).astype(np.uint8))
    # the mask-only call decides most pixels without the exact quotient; same mask
    _, mask2 = K().stitch_blend(p, (1, 1, H, W), fx, (one, w, w), 0, 0.5, want_out=False, want_mask=True)
    assert torch.equal(mask, mask2)
    # with the reference's fp32 window table (itself bit-exact) the fp32 filter kernel decides; same mask
    table = K().blend_window((one, w, w))
    assert np.array_equal(table.cpu().numpy(), O.fibsem_window(tile))
    _, mask3 = K().stitch_blend(p, (1, 1, H, W), fx, (one, w, w), 0, 0.5, want_out=False, want_mask=True, window=table)
    assert torch.equal(mask, mask3)


@pytest.mark.parametrize("name", ["u8", "f32", "flat", "hot"])
def test_normalize_equals_reference_analyze(dev, name):
    """ref: analysis_deployment.py:241-242 through be_quantile_f32 + be_normalize_f32 (clip 0..1)."""
    z = np.load(os.path.join(GOLD, "ref_analyze_norm.npz"))
    x = torch.from_numpy(z[f"{name}_image"].astype(np.float32)).to(dev)
    q = torch.tensor([0.01, 0.99], dtype=torch.float64, device=dev)
    _, norm = K().quantile(x, 1, q, 1e-6)
    got = K().normalize(x, norm, clip=(0.0, 1.0)).cpu().numpy()
    np.testing.assert_allclose(got, z[f"{name}_norm"], rtol=1e-5, atol=0)
    assert np.array_equal(got, z[f"{name}_norm"])


@pytest.mark.parametrize("name", ["prob_prob", "logit_prob", "prob_label", "edges", "empty"])
def test_binary_metrics_equal_reference(dev, name):
    """ref: apps/cellpose-finetuning/main.py:1225-1270 through be_binary_confusion_f32."""
    from bioengine_worker_b200.metrics import binary_metrics

    z = np.load(os.path.join(GOLD, "ref_binary_metrics.npz"))
    m = binary_metrics(z[f"{name}_pred"], z[f"{name}_target"], device=dev)
    assert [m["tp"], m["fp"], m["fn"], m["tn"]] == z[f"{name}_counts"].tolist()
    got = np.array([m["pixel_accuracy"], m["precision"], m["recall"], m["f1"], m["iou"]], np.float64)
    assert np.array_equal(got, z[f"{name}_ratios"])
