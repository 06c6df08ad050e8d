"""World-size-2 gloo run of the screening merge: each rank fills only its image shard, the all-reduce must rebuild the
single-process table and every rank must compute the same AP / ranking (CPU only)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_case(n_models=3, n_images=7, seed=0):
    rng = np.random.default_rng(seed)
    gts, preds = [], [[] for _ in range(n_models)]
    for i in range(n_images):
        a = O.gaussian_blobs((96, 120), 10 + i, seed=seed + i)
        gts.append(O.label(a > 0.5, 2))
        for m in range(n_models):
            b = a + (0.1 + 0.15 * m) * rng.standard_normal(a.shape).astype(np.float32)
            preds[m].append(O.label(b > 0.5, 2))
    return gts, preds


def _full_table(gts, preds, thr):
    n_models, n_images = len(preds), len(gts)
    t = np.zeros((n_models, n_images, len(thr) + 2), np.int64)
    for m in range(n_models):
        _, tp, _, _ = O.average_precision(gts, preds[m], threshold=list(thr))
        t[m, :, :len(thr)] = tp.astype(np.int64)
        t[m, :, len(thr)] = [g.max() for g in gts]
        t[m, :, len(thr) + 1] = [p.max() for p in preds[m]]
    return t


def _worker(rank, world, port, full, thr, out):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from bioengine_worker_b200.screening import merge_and_rank, shard_indices

    local = np.zeros_like(full)
    mine = shard_indices(full.shape[1], rank, world)
    local[:, mine] = full[:, mine]
    res = merge_and_rank(torch.from_numpy(local), thr)
    out[rank] = (res["table"], res["ap"], res["mean_ap"], res["ranking"])
    dist.destroy_process_group()


def test_two_rank_merge_equals_single_process():
    thr = (0.5, 0.75, 0.9)
    gts, preds = _make_case()
    full = _full_table(gts, preds, thr)
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(2, _free_port(), full, thr, out), nprocs=2, join=True)
    assert sorted(out.keys()) == [0, 1]
    for rank in (0, 1):
        table, ap, mean_ap, ranking = out[rank]
        assert np.array_equal(table, full)
        for m in range(len(preds)):
            want_ap, _, _, _ = O.average_precision(gts, preds[m], threshold=list(thr))
            assert np.array_equal(ap[m], want_ap, equal_nan=True)
            assert np.array_equal(mean_ap[m], O.mean_ap(want_ap), equal_nan=True)
        assert ranking.tolist() == np.argsort(-mean_ap[:, 0], kind="stable").tolist()


def test_shards_partition_the_images():
    from bioengine_worker_b200.screening import shard_indices

    for world in (1, 2, 4, 8):
        seen = sorted(i for r in range(world) for i in shard_indices(1024, r, world))
        assert seen == list(range(1024))
