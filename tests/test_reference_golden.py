"""The oracle against fixtures written by the REFERENCE's own functions (tests/golden/ref_*.npz, made by
tests/golden/make_reference_golden.py from /root/reference/apps/fibsem-mito-analysis/analysis_deployment.py and
/root/reference/apps/cellpose-finetuning/main.py, imported unmodified under two stub modules).  CPU only; the GPU
counterpart is tests/test_gpu_reference_golden.py.  Everything here is bit-exact."""
import os

import numpy as np
import pytest

from oracle import oracle as O

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
N_INFER_CASES = 6


def ref_probs(tiles, ramp):
    """The stand-in network of make_reference_golden.py: probs[k] = fl32(tiles[k] * ramp + fl32(0.01) * fl32((k+1) % 7))."""
    k = (np.arange(1, len(tiles) + 1) % 7).astype(np.float32)
    return (tiles * ramp[None] + (np.float32(0.01) * k)[:, None, None]).astype(np.float32)


def infer_case(z, ci):
    H, W, tile, ov = (int(v) for v in z[f"c{ci}_geom"])
    return (H, W, tile, ov), z[f"c{ci}_img"], z[f"c{ci}_tiles"], ref_probs(z[f"c{ci}_tiles"], z[f"c{ci}_ramp"]), z[f"c{ci}_out"]


@pytest.mark.parametrize("ci", range(N_INFER_CASES))
def test_oracle_tiles_and_blend_equal_reference_infer_tiled(ci):
    """ref: analysis_deployment.py:108-156 — the tiles the reference hands to the network and its blended result."""
    z = np.load(os.path.join(G, "ref_infer_tiled.npz"))
    (H, W, tile, ov), img, tiles, probs, out = infer_case(z, ci)
    assert np.array_equal(O.fibsem_tiles(img, tile, ov), tiles)
    got = O.fibsem_blend(probs, (H, W), tile, ov)
    assert got.dtype == np.float32 and np.array_equal(got, out)


@pytest.mark.parametrize("name", ["u8", "f32", "flat", "hot"])
def test_oracle_normalize_equals_reference_analyze(name):
    """ref: analysis_deployment.py:236-242 — np.array(list, float32), p1/p99, clip to [0, 1], float32."""
    z = np.load(os.path.join(G, "ref_analyze_norm.npz"))
    assert np.array_equal(O.fibsem_normalize(z[f"{name}_image"].astype(np.float32)), z[f"{name}_norm"])


@pytest.mark.parametrize("name", ["prob_prob", "logit_prob", "prob_label", "edges", "empty"])
def test_oracle_binary_metrics_equal_reference(name):
    """ref: apps/cellpose-finetuning/main.py:1225-1270 — adaptive thresholds, the four counts, the five ratios."""
    z = np.load(os.path.join(G, "ref_binary_metrics.npz"))
    m = O.binary_metrics(z[f"{name}_pred"], z[f"{name}_target"])
    assert [m["tp"], m["fp"], m["fn"], m["tn"]] == z[f"{name}_counts"].tolist()
    got = np.array([m["pixel_accuracy"], m["precision"], m["recall"], m["f1"], m["iou"]], np.float64)
    assert np.array_equal(got, z[f"{name}_ratios"])
