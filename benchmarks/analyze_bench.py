#!/usr/bin/env python
"""`MitoAnalysisDeployment.analyze` end to end on one 2048^2 image (host array in → labels + properties out), with the
oracle's CPU restatement of the same call (ref: apps/fibsem-mito-analysis/analysis_deployment.py:236-286; the network's
outputs are taken from the GPU run, so the forward is on neither clock for the CPU side).  Diagnostic."""
import asyncio
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bioengine_worker_b200 import models  # noqa: E402
from bioengine_worker_b200.analysis_deployment import MitoAnalysisDeployment  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    tile, ov = 512, 64
    dev = torch.device("cuda:0")
    models.strict_fp32()
    rng = np.random.default_rng(0)
    img = (O.gaussian_blobs((S, S), S * S // 3000, sigma=(6.0, 16.0), seed=2) * 170 + 40 + 6 * rng.standard_normal((S, S))).clip(0, 255).astype(np.uint8)
    net = models.build_model("UNet2d", seed=2, final_sigmoid=True).to(dev)
    # a stand-in network that segments the blobs: smooth, threshold — the post-processing then has real objects to split
    level = float(np.percentile(O.fibsem_normalize(img.astype(np.float32)), 72))
    seg = lambda t: torch.sigmoid((torch.nn.functional.avg_pool2d(t, 5, 1, 2) - level) * 30.0)  # noqa: E731
    out = {}
    for name, model in (("unet2d_random_init", net), ("smooth_and_threshold", seg)):
        dep = MitoAnalysisDeployment(model=model, device=dev)
        asyncio.run(dep.analyze(img, tile_size=tile, overlap=ov, as_numpy=True))   # warm-up: pipeline + workspaces
        torch.cuda.synchronize()
        times = []
        for _ in range(5):
            t0 = time.perf_counter()
            res = asyncio.run(dep.analyze(img, tile_size=tile, overlap=ov, as_numpy=True))
            times.append((time.perf_counter() - t0) * 1e3)
        out[name] = {"ms_median": float(np.median(times)), "n_mitochondria": res["n_mitochondria"]}
    # the CPU side of the second model: same tiles, same network outputs, numpy / scipy / the C flood on one core
    norm = O.fibsem_normalize(img.astype(np.float32))
    t0 = time.perf_counter()
    norm = O.fibsem_normalize(img.astype(np.float32))
    tiles = O.fibsem_tiles(norm, tile, ov)
    t_pre = time.perf_counter() - t0
    with torch.no_grad():
        probs = seg(torch.from_numpy(tiles[:, None]).to(dev)).cpu().numpy()[:, 0]
    t0 = time.perf_counter()
    prob = O.fibsem_blend(probs, (S, S), tile, ov)
    t1 = time.perf_counter()
    labels = O.prob_to_instances(prob, ties="fifo")
    t2 = time.perf_counter()
    props = O.regionprops_table(labels, 5.0)
    t3 = time.perf_counter()
    out["cpu_oracle_one_core"] = {"normalise+tiles_ms": t_pre * 1e3, "blend_ms": (t1 - t0) * 1e3, "instances_ms": (t2 - t1) * 1e3,
                                  "regionprops_ms": (t3 - t2) * 1e3, "n_mitochondria": int(labels.max()),
                                  "labels_equal_gpu": bool(np.array_equal(labels, res["labels"])),
                                  "n_properties": len(props["label"])}
    out["config"] = f"{S}x{S} uint8 image, tile {tile} overlap {ov}, analyze(as_numpy=True), forward included on the GPU side only"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
