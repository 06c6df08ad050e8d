#!/usr/bin/env python
"""Time of the fibsem post-processing (`_prob_to_instances`, ref: analysis_deployment.py:160-176) on one 2048^2 mask of
touching blobs: device stages by CUDA events, the oracle (scipy + the C flood) on one host core beside it.  Diagnostic."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bioengine_worker_b200 import kernels as K  # noqa: E402
from oracle import oracle as O  # noqa: E402


def ev(fn, n=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    S = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    dev = torch.device("cuda:0")
    img = O.gaussian_blobs((S, S), S * S // 2500, sigma=(6.0, 16.0), seed=1)
    prob = np.clip(img * 1.3, 0, 1).astype(np.float32)
    mask = torch.from_numpy((prob > 0.5).astype(np.uint8)).to(dev)
    kept = K.remove_small_objects(mask[None], 300)
    closed = K.binary_closing(kept, K.disk_half_widths(4))
    d2 = K.distance_transform_edt_sq(closed)
    peaks = K.peak_local_max(d2[0], closed[0], 8)
    markers, n_seeds = K.ccl(peaks[None], 2)
    out = K.watershed(d2[0], markers[0], closed[0], negate=True)
    t = {
        "remove_small_objects": ev(lambda: K.remove_small_objects(mask[None], 300)),
        "binary_closing": ev(lambda: K.binary_closing(kept, K.disk_half_widths(4))),
        "edt_sq": ev(lambda: K.distance_transform_edt_sq(closed)),
        "peak_local_max": ev(lambda: K.peak_local_max(d2[0], closed[0], 8)),
        "label_peaks": ev(lambda: K.ccl(peaks[None], 2)),
        "watershed(regions+flood)": ev(lambda: K.watershed(d2[0], markers[0], closed[0], negate=True)),
        "all": ev(lambda: K.instances_from_mask(mask)),
    }
    torch.cuda.synchronize()
    K.watershed(d2[0], markers[0], closed[0], negate=True)
    torch.cuda.synchronize()
    st = K._default_ws._buf[:64].view(torch.int64).cpu().numpy()
    t0 = time.perf_counter()
    parts = {}
    want = O.prob_to_instances(prob, ties="fifo", parts=parts)
    cpu = time.perf_counter() - t0
    regions = int(K.ccl(closed, 1)[1].item())
    print(json.dumps({"size": S, "foreground": float(closed.float().mean().item()), "regions": regions,
                      "seeds": int(n_seeds.item()), "ms": {k: round(v, 3) for k, v in t.items()},
                      "cpu_oracle_ms_one_core": round(cpu * 1e3, 1), "flood_stats[steps,entries,max_queue,max_steps]": None if st is None else st[1:5].tolist(), "equal_to_oracle": bool(np.array_equal(out.cpu().numpy(), want))}))


if __name__ == "__main__":
    main()
