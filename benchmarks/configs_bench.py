#!/usr/bin/env python
"""Timings for the BASELINE.json configurations that are not the bench.py line (C3 volume, C4 screening, C5 stream).
Diagnostic only — bench.py is the contract.  Usage:

    python benchmarks/configs_bench.py c3|c5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 benchmarks/configs_bench.py c4

Prints one JSON line per configuration (rank 0).  The UNet forward is plain torch and is timed separately where it is
part of the loop (C4, C5); kernel times are CUDA events on the launching stream.
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bioengine_worker_b200 import geometry as G  # noqa: E402
from bioengine_worker_b200 import kernels as K  # noqa: E402
from bioengine_worker_b200 import models  # noqa: E402
from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline  # noqa: E402


def ev_time(fn, n=5, warm=2):
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


def c3(dev):
    """512^3 volume, fibsem p1/p99 + clip, 128^3 tiles / overlap 32 (216 tiles), blend, > 0.5, 26-connected labels."""
    S, B, ov = 512, 128, 32
    g = torch.Generator(device=dev).manual_seed(3)
    # smooth random field (blob-like structures a few tens of voxels across) + pixel noise
    coarse = torch.randn((1, 1, 24, 24, 24), device=dev, generator=g)
    x = torch.nn.functional.interpolate(coarse, size=(S, S, S), mode="trilinear", align_corners=False)
    x = (x * 40 + 120 + 4 * torch.randn(x.shape, device=dev, generator=g)).contiguous()
    spec = PipelineSpec(block=(B,) * 3, halo=(16,) * 3, convention="fibsem", overlap=(ov,) * 3, preprocess="fibsem",
                        sigmoid=True, connectivity=3, forward_batch=8)
    pipe = TiledPipeline((1, 1, S, S, S), spec, device=dev)
    tiles = pipe.preprocess_and_tile(x)
    # stand-in logits with blob-like structure (the 3-D UNet forward is outside the measured kernels)
    logits = ((tiles - 0.78) * 12.0).contiguous()  # foreground = brightest ~15 % of the volume
    t_pre = ev_time(lambda: pipe.preprocess_and_tile(x))
    res = pipe.stitch_and_label(logits)
    t_post = ev_time(lambda: pipe.stitch_and_label(logits))
    _, _, bits = K.stitch_blend(logits, (1, 1, S, S, S), pipe.axes, pipe.windows, 1, 0.5, want_out=False, want_bits=True)
    t_blend = ev_time(lambda: K.stitch_blend(logits, (1, 1, S, S, S), pipe.axes, pipe.windows, 1, 0.5, want_out=False, want_bits=True))
    t_ccl = ev_time(lambda: K.ccl_bits(bits, (1, S, S, S), 3, ws=pipe.ws))
    q = torch.tensor([0.01, 0.99], dtype=torch.float64, device=dev)
    _, norm = K.quantile(x, 1, q, 1e-6, ws=pipe.ws)
    t_quant = ev_time(lambda: K.quantile(x, 1, q, 1e-6, ws=pipe.ws))
    t_gather = ev_time(lambda: K.tile_gather(x, pipe.axes, norm=norm, norm_mode=1, clip=(0.0, 1.0), out=pipe.tiles))
    vox = S**3
    E = pipe.n_tiles * B**3 / vox
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
    stages = {}
    for name, ms, bpv in (("quantile_select", t_quant, 4.0), ("normalise+tile_gather", t_gather, 4 + 4 * E),
                          ("blend_stitch+sigmoid+binarize", t_blend, 4 * E + 1), ("ccl", t_ccl, 5.0)):
        gbs = bpv * vox / (ms * 1e-3) / 1e9
        stages[name] = {"ms": round(ms, 4), "alg_bytes_per_vox": round(bpv, 3), "alg_GBps": round(gbs, 1),
                        "frac_of_peak": round(gbs / peak, 4)}
    alg = (14 + 8 * E) * vox
    total = t_pre + t_post
    return {"config": "C3 512^3, 128^3 tiles overlap 32 (216 tiles), fibsem norm, blend, sigmoid, binarize, CC 26-conn",
            "pre_ms": t_pre, "post_ms": t_post, "ms_per_volume": total, "stages": stages,
            "Gvox_per_s": vox / (total * 1e-3) / 1e9,
            "path_roofline": {"alg_bytes_per_vox": round(14 + 8 * E, 3), "achieved_GBps": round(alg / (total * 1e-3) / 1e9, 1),
                              "peak": peak, "frac": round(alg / (total * 1e-3) / 1e9 / peak, 4),
                              "target_ms_at_0.6": round(alg / (0.6 * peak * 1e9) * 1e3, 3)},
            "n_components": int(res["n_labels"].item()),
            "foreground": float(res["mask"].float().mean().item()), "data": "synthetic, logits = affine stand-in for the 3-D forward"}


def c5(dev, frames=200):
    """2048^2 frame stream: H2D (pinned) → scale_range → whole-frame forward → sigmoid/binarize → CC → label stats → D2H of
    (count, areas, centroids); per-frame latency."""
    from oracle import oracle as O

    H = W = 2048
    models.strict_fp32()
    net = models.build_model("UNet2d", seed=5).to(dev)
    base, clean = O.synthetic_image((1, 1, H, W), 800, seed=5, with_truth=True)
    host = torch.from_numpy(np.clip(base, 0, 65535).astype(np.uint16)).pin_memory()   # frames arrive as uint16
    spec = PipelineSpec(block=(512, 512), halo=(32, 32), convention="bioimageio", preprocess="scale_range", sigmoid=True,
                        connectivity=2, forward_batch=25)
    pipe = TiledPipeline((1, 1, H, W), spec, device=dev)
    models.calibrate_head(net, pipe.preprocess_and_tile(pipe.upload(host))[::5].contiguous())
    cap = 1 << 18
    lat, lat_nofwd = [], []
    for i in range(frames + 10):
        t0 = time.perf_counter()
        x = pipe.upload(host)   # 8 MB of uint16 over PCIe, float32 on the device
        tiles = pipe.preprocess_and_tile(x)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        logits = pipe.forward(net, tiles)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        res = pipe.stitch_and_label(logits)
        sums, bbox, status = K.label_stats(res["labels"][0], cap)
        n = int(res["n_labels"].item())
        stats = sums[0, : min(n + 1, cap), :4].cpu()
        t3 = time.perf_counter()
        if i >= 10:
            lat.append((t3 - t0) * 1e3)
            lat_nofwd.append((t1 - t0 + t3 - t2) * 1e3)
    lat, nf = np.array(lat), np.array(lat_nofwd)
    return {"config": "C5 2048^2 uint16 frame stream (pinned host), upload + ensure_dtype + scale_range + 512/32 blocks + UNet2d + "
                      "sigmoid/binarize + CC + label stats + download of the stats",
            "frames": frames, "h2d_bytes_per_frame": H * W * 2, "p50_ms": float(np.percentile(lat, 50)), "p99_ms": float(np.percentile(lat, 99)),
            "p50_ms_without_forward": float(np.percentile(nf, 50)), "p99_ms_without_forward": float(np.percentile(nf, 99)),
            "n_components_last": n}


def c4(dev, rank, world, n_images=256, n_models=8, size=1024):
    """Screening: n_models random-init UNets x n_images synthetic size^2 images with GT labels, sharded by image index."""
    import torch.distributed as dist

    from bioengine_worker_b200.screening import screen_models
    from oracle import oracle as O

    models.strict_fp32()
    rng = np.random.default_rng(4)
    # a small pool of distinct images reused across the index range keeps host-side generation out of the picture
    pool = []
    for i in range(8):
        img, clean = O.synthetic_image((size, size), 200, seed=4000 + i, microscope=False, with_truth=True)
        pool.append((img, O.label(clean > 0.5, 2)))
    images = [pool[i % 8][0] for i in range(n_images)]
    gts = [pool[i % 8][1] for i in range(n_images)]
    nets = []
    probe = torch.from_numpy(O.zero_mean_unit_variance(images[0][None, None], axes=(1, 2, 3))).to(dev)
    for m in range(n_models):
        net = models.build_model("UNet2d", seed=40 + m, depth=3, initial_features=8 if m % 2 else 16).to(dev)
        nets.append(models.calibrate_head(net, probe))
    level = float((0.5 - images[0].mean()) / images[0].std())  # the GT level in normalised units
    # one candidate that really segments (smooth + threshold): it should rank first
    nets[n_models // 2] = lambda t: (torch.nn.functional.avg_pool2d(t, 5, 1, 2) - level) * 8.0
    screen_models(nets[:1], images[: 2 * world], gts[: 2 * world], device=dev, batch=2, rank=rank, world=world)  # warm-up
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = screen_models(nets, images, gts, device=dev, batch=16, rank=rank, world=world)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    return {"config": f"C4 screening: {n_models} models x {n_images} images {size}^2, mAP@[0.5,0.75,0.9], {world} GPU(s)",
            "seconds": dt, "Mpx_per_s_model_images": n_models * n_images * size * size / dt / 1e6,
            "mean_ap": np.round(res["mean_ap"], 4).tolist(), "ranking": res["ranking"].tolist()}


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "c3"
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    out = {"c3": lambda: c3(dev), "c5": lambda: c5(dev), "c4": lambda: c4(dev, rank, world)}[which]()
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
