"""Writes tests/golden/ref_*.npz from the REFERENCE's own code, imported unmodified from /root/reference.

Run from the repo root, in the build container (the GPU box has no /root/reference; it only reads the fixtures):
    python tests/golden/make_reference_golden.py

The two reference files on this path import `ray.serve` and `hypha_rpc.utils.schema` at module top; neither package
is installed here and neither takes part in the arithmetic, so two stub modules are put into sys.modules
(`serve.deployment` wraps the class in an object exposing `.func_or_class`, the way Ray does; `schema_method` is the
identity).  Everything below calls the reference's functions, not the oracle's:

  ref_infer_tiled.npz   `MitoAnalysisDeployment._infer_tiled`   apps/fibsem-mito-analysis/analysis_deployment.py:108-156
                        with `_infer_full` (the network round trip, :84-106) replaced by a fixed function that records
                        the tile it is handed and returns `tile * ramp + bias_k` (position- and call-dependent, so the
                        blend weights and the tile order matter)
  ref_analyze_norm.npz  `MitoAnalysisDeployment.analyze`        analysis_deployment.py:192-286, with
                        `_prob_to_instances` replaced by a recorder (scikit-image is not installed) and the network
                        replaced by the identity: the recorded `prob` is the normalised image of :241-242
  ref_binary_metrics.npz `_binary_threshold`, `_compute_binary_metrics`  apps/cellpose-finetuning/main.py:1225-1270
                        on the three value-range cases of the adaptive threshold

numpy here is 2.x: `p1`/`p99` are float64 scalars and promote the normalisation to fp64 (NEP 50); the reference pins no
numpy version for this app (analysis_deployment.py has no runtime_env)."""
import asyncio
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.dirname(os.path.abspath(__file__))


def install_stubs():
    class _Deployment:
        def __init__(self, cls):
            self.func_or_class = cls

    def deployment(*args, **kwargs):
        if len(args) == 1 and not kwargs and isinstance(args[0], type):
            return _Deployment(args[0])
        return lambda cls: _Deployment(cls)

    ray = types.ModuleType("ray")
    serve = types.ModuleType("ray.serve")
    serve.deployment = deployment
    serve.multiplexed = lambda *a, **k: (a[0] if a and callable(a[0]) else (lambda f: f))
    serve.get_multiplexed_model_id = lambda: ""
    ray.serve = serve
    hy = types.ModuleType("hypha_rpc")
    hyu = types.ModuleType("hypha_rpc.utils")
    hys = types.ModuleType("hypha_rpc.utils.schema")
    hys.schema_method = lambda f=None, **k: (f if f is not None else (lambda g: g))
    hy.utils, hyu.schema = hyu, hys
    for name, mod in (("ray", ray), ("ray.serve", serve), ("hypha_rpc", hy), ("hypha_rpc.utils", hyu),
                      ("hypha_rpc.utils.schema", hys)):
        sys.modules.setdefault(name, mod)


def load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def image(shape, seed):
    """Smooth structure + noise in [0, 1] (what `_infer_tiled` receives after the normalise lines)."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:shape[0], 0:shape[1]]
    f = 0.5 + 0.3 * np.sin(y / 7.0 + seed) * np.cos(x / 11.0) + 0.15 * rng.standard_normal(shape)
    return np.clip(f, 0, 1).astype(np.float32)


# (H, W, tile_size, overlap): interior overlaps; last tile narrower than the overlap (multi-bounce reflect padding);
# a one-pixel last tile (np.pad reflect on a single sample); image smaller than one tile; exact multiple of the stride
INFER_CASES = [
    (300, 410, 128, 32),
    (200, 104, 64, 16),
    (97, 193, 64, 16),
    (40, 50, 64, 16),
    (192, 96, 64, 16),
    (150, 70, 32, 0),
]


def ref_probs(tiles, ramp):
    """The stand-in network's outputs for the recorded tiles (tests rebuild them with the same expression)."""
    k = (np.arange(1, len(tiles) + 1) % 7).astype(np.float32)
    return (tiles * ramp[None] + (np.float32(0.01) * k)[:, None, None]).astype(np.float32)


def make_infer_tiled(mito_cls, out):
    for ci, (H, W, tile, ov) in enumerate(INFER_CASES):
        inst = mito_cls()
        img = image((H, W), 100 + ci)
        rng = np.random.default_rng(500 + ci)
        ramp = (0.5 + rng.random((tile, tile))).astype(np.float32)
        seen, given = [], []

        async def fake_infer_full(t, ramp=ramp, seen=seen, given=given):
            seen.append(np.array(t, dtype=np.float32, copy=True))
            p = (t * ramp + np.float32(0.01) * np.float32(len(seen) % 7)).astype(np.float32)
            given.append(p)
            return p

        inst._infer_full = fake_infer_full
        res = asyncio.run(inst._infer_tiled(img, tile_size=tile, overlap=ov))
        out[f"c{ci}_geom"] = np.array([H, W, tile, ov], np.int64)
        out[f"c{ci}_img"] = img
        out[f"c{ci}_tiles"] = np.stack(seen)
        out[f"c{ci}_ramp"] = ramp  # probs[k] = fl32(tiles[k] * ramp + fl32(0.01) * fl32((k + 1) % 7)), see ref_probs()
        assert all(np.array_equal(g, ref_probs(np.stack(seen), ramp)[k]) for k, g in enumerate(given))
        out[f"c{ci}_out"] = res
        assert res.dtype == np.float32 and res.shape == (H, W)


def make_analyze_norm(mito_cls, out):
    rng = np.random.default_rng(77)
    cases = {
        "u8": rng.integers(0, 256, (60, 90)).astype(np.uint8),
        "f32": (rng.standard_normal((48, 64)) * 300 + 1000).astype(np.float32),
        "flat": np.full((16, 16), 7.0, np.float32),  # p99 == p1: the 1e-6 in the denominator decides
        "hot": np.where(rng.random((80, 80)) < 0.02, 65535.0, rng.random((80, 80)) * 100).astype(np.float32),
    }
    for name, arr in cases.items():
        inst = mito_cls()
        got = {}

        async def identity(img_norm):
            return img_norm

        async def identity_tiled(img_norm, tile_size=512, overlap=64):
            return img_norm

        def record(prob, got=got):
            got["prob"] = np.array(prob, copy=True)
            return np.zeros(prob.shape, np.int32)

        inst._infer_full, inst._infer_tiled, inst._prob_to_instances = identity, identity_tiled, record
        res = asyncio.run(inst.analyze(arr.tolist(), pixel_size_nm=5.0, tile_size=512, overlap=64))
        assert res["n_mitochondria"] == 0 and res["image_shape"] == list(arr.shape)
        out[f"{name}_image"] = arr
        out[f"{name}_norm"] = got["prob"]
        assert got["prob"].dtype == np.float32


def make_binary_metrics(main_mod, out):
    import torch

    rng = np.random.default_rng(9)
    shape = (3, 40, 56)
    cases = {
        # both inside [0, 1] → thresholds 0.5 / 0.5
        "prob_prob": (rng.random(shape).astype(np.float32), (rng.random(shape) < 0.3).astype(np.float32)),
        # logits (range outside [0,1]) vs {0,1} target → 0.0 / 0.5
        "logit_prob": ((rng.standard_normal(shape) * 2).astype(np.float32), (rng.random(shape) < 0.4).astype(np.float32)),
        # probabilities vs integer instance labels (max > 1) → 0.5 / 0.0
        "prob_label": (rng.random(shape).astype(np.float32), rng.integers(0, 5, shape).astype(np.float32)),
        # values exactly on the range limits and on the thresholds
        "edges": (np.array([[0.0, 0.5, 1.0, 0.50001]], np.float32), np.array([[0.0, 1.0, 0.5, 1.0]], np.float32)),
        # nothing predicted: every ratio goes through _safe_div's zero branch
        "empty": (np.zeros((2, 8, 8), np.float32), np.zeros((2, 8, 8), np.float32)),
    }
    for name, (pred, target) in cases.items():
        p, t = torch.from_numpy(pred), torch.from_numpy(target)
        pr, gt = main_mod._binary_threshold(p, t)
        m = main_mod._compute_binary_metrics(p, t)
        md = m.model_dump() if hasattr(m, "model_dump") else dict(m)
        out[f"{name}_pred"], out[f"{name}_target"] = pred, target
        out[f"{name}_pr"], out[f"{name}_gt"] = pr.numpy(), gt.numpy()
        out[f"{name}_ratios"] = np.array(
            [md["pixel_accuracy"], md["precision"], md["recall"], md["f1"], md["iou"]], np.float64)
        out[f"{name}_counts"] = np.array(
            [int((pr & gt).sum()), int((pr & ~gt).sum()), int((~pr & gt).sum()), int((~pr & ~gt).sum())], np.int64)


def main():
    install_stubs()
    mito = load(os.path.join(REF, "apps/fibsem-mito-analysis/analysis_deployment.py"), "ref_fibsem_analysis")
    mito_cls = mito.MitoAnalysisDeployment.func_or_class
    a, b = {}, {}
    make_infer_tiled(mito_cls, a)
    np.savez_compressed(os.path.join(OUT, "ref_infer_tiled.npz"), **a)
    make_analyze_norm(mito_cls, b)
    np.savez_compressed(os.path.join(OUT, "ref_analyze_norm.npz"), **b)
    try:
        cp = load(os.path.join(REF, "apps/cellpose-finetuning/main.py"), "ref_cellpose_main")
    except Exception as e:  # its module top may want more than the two stubs
        print("cellpose-finetuning/main.py did not import:", repr(e))
        raise
    c = {}
    make_binary_metrics(cp, c)
    np.savez_compressed(os.path.join(OUT, "ref_binary_metrics.npz"), **c)
    for f in ("ref_infer_tiled.npz", "ref_analyze_norm.npz", "ref_binary_metrics.npz"):
        print(f, os.path.getsize(os.path.join(OUT, f)), "bytes")


if __name__ == "__main__":
    main()
