#!/usr/bin/env python
"""The kernel path of bench.py (quantiles → gather → mask-only blend → labels, forward excluded, stand-in logits) over a
range of image sizes: where the path is launch / latency bound and where it streams.  Diagnostic; CUDA events, L2 flushed
between steps."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for S in (512, 1024, 2048, 4096, 8192, 16384):
        g = torch.Generator(device=dev).manual_seed(S)
        # speckle-free synthetic content made on the device: smooth field + noise, a few hot pixels
        coarse = torch.randn((1, 1, S // 32, S // 32), device=dev, generator=g)
        x = torch.nn.functional.interpolate(coarse, size=(S, S), mode="bilinear", align_corners=False)
        x = (x * 300 + 600 + 20 * torch.randn((1, 1, S, S), device=dev, generator=g)).contiguous()
        x.view(-1)[:: 1009] = 65535.0
        pipe = TiledPipeline((1, 1, S, S), PipelineSpec(block=(256, 256), halo=(32, 32), convention="fibsem", overlap=(64, 64),
                                                        preprocess="scale_range", connectivity=2, return_mask=False), device=dev)
        tiles = pipe.preprocess_and_tile(x)
        logits = ((tiles - 0.62) * 14.0).contiguous()
        res = pipe.stitch_and_label(logits)
        n_comp = int(res["n_labels"].item())
        steps = 20 if S <= 8192 else 8
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        for i in range(steps + 3):
            flush.fill_(i & 255)
            if i >= 3:
                ev[i - 3][0].record()
            pipe.preprocess_and_tile(x)
            pipe.stitch_and_label(logits)
            if i >= 3:
                ev[i - 3][1].record()
        torch.cuda.synchronize()
        ms = float(np.median([a.elapsed_time(b) for a, b in ev]))
        # the four C-ABI calls on their own (same flush, same events)
        from bioengine_worker_b200 import kernels as K

        q = torch.tensor([0.01, 0.998], dtype=torch.float64, device=dev)
        _, norm = K.quantile(x, 1, q, 1e-6, ws=pipe.ws)
        _, _, bits = K.stitch_blend(logits, (1, 1, S, S), pipe.axes, pipe.windows, 1, 0.5, want_out=False, want_bits=True,
                                    window=pipe.window_table)
        calls = {"quantiles": lambda: K.quantile(x, 1, q, 1e-6, ws=pipe.ws),
                 "gather": lambda: K.tile_gather(x, pipe.axes, norm=norm, norm_mode=1, out=pipe.tiles),
                 "blend": lambda: K.stitch_blend(logits, (1, 1, S, S), pipe.axes, pipe.windows, 1, 0.5, want_out=False,
                                                 want_bits=True, window=pipe.window_table),
                 "labels": lambda: K.ccl_bits(bits, (1, S, S), 2, ws=pipe.ws)}
        stage = {}
        for name, fn in calls.items():
            ts = []
            for i in range(8):
                flush.fill_(i)
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                fn()
                b.record()
                torch.cuda.synchronize()
                if i >= 2:
                    ts.append(a.elapsed_time(b))
            stage[name] = round(float(np.median(ts)), 4)
        E = pipe.n_tiles * 256 * 256 / (S * S)
        alg = (14 + 8 * E) * S * S
        rows.append({"size": S, "tiles": pipe.n_tiles, "ms": round(ms, 4), "Gpx_s": round(S * S / ms / 1e6, 1),
                     "alg_TB_s": round(alg / ms / 1e9, 3), "components": n_comp,
                     "foreground": round(float(res["labels"].gt(0).float().mean().item()), 3), "stage_ms": stage})
        del pipe, tiles, logits, res, x
        torch.cuda.empty_cache()
    print(json.dumps({"path": "scale_range p1/p99.8 -> 256-px tiles overlap 64 -> blend + sigmoid + binarize (mask only) -> 8-conn labels",
                      "rows": rows}))


if __name__ == "__main__":
    main()
