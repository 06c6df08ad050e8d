"""Label-overlap histogram, threshold matching and label statistics against the oracle (all integer-exact)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def K():
    from bioengine_worker_b200 import kernels

    return kernels


def dense_overlap(table, i, nt, npred):
    keys = table["keys"][i].cpu().numpy().astype(np.uint64)
    cnt = table["counts"][i].cpu().numpy().astype(np.uint32)
    out = np.zeros((nt, npred), np.uint64)
    used = keys != np.uint64(0xFFFFFFFFFFFFFFFF)
    t = (keys[used] >> np.uint64(32)).astype(np.int64)
    p = (keys[used] & np.uint64(0xFFFFFFFF)).astype(np.int64)
    out[t, p] = cnt[used]
    return out


def make_pair(rng, shape=(256, 320), n_blobs=40, seed=0):
    a = O.gaussian_blobs(shape, n_blobs, seed=seed)
    b = O.gaussian_blobs(shape, n_blobs, seed=seed) + 0.15 * rng.standard_normal(shape).astype(np.float32)
    return O.label(a > 0.5, 2), O.label(b > 0.45, 2)


def test_overlap_table_equals_dense_histogram(dev, rng):
    pairs = [make_pair(rng, seed=s) for s in range(3)]
    t = torch.from_numpy(np.stack([p[0] for p in pairs])).to(dev)
    p = torch.from_numpy(np.stack([p[1] for p in pairs])).to(dev)
    tab = K().label_overlap(t, p, capacity=1 << 14, cap_true=512, cap_pred=4096)
    assert tab["status"].cpu().tolist() == [0, 0, 0]
    for i, (lt, lp) in enumerate(pairs):
        want = O.label_overlap(lt, lp)
        got = dense_overlap(tab, i, want.shape[0], want.shape[1])
        assert np.array_equal(got[1:, 1:], want[1:, 1:])     # the table holds the pairs with both labels > 0 ...
        assert not got[0].any() and not got[:, 0].any()      # ... the background row / column lives in the areas
        assert np.array_equal(tab["area_true"][i].cpu().numpy()[: want.shape[0]], want.sum(1).astype(np.int32))
        assert np.array_equal(tab["area_pred"][i].cpu().numpy()[: want.shape[1]], want.sum(0).astype(np.int32))
        assert tab["max_lbl"][i].cpu().tolist() == [lt.max(), lp.max()]


@pytest.mark.parametrize("shape", [(256, 320), (131, 257), (7, 1000), (300, 5)])
def test_overlap_with_device_bounds_ragged_shapes_and_flat_input(dev, rng, shape):
    """Per-image label bounds on the device cut the tables to what each image needs (the rest stays uninitialised and
    must never be read); rows that are not multiples of 4 / 128; the same pixels handed over without row structure."""
    from bioengine_worker_b200.metrics import counts_to_tp

    pairs = [make_pair(rng, shape=shape, n_blobs=12 + 9 * s, seed=10 + s) for s in range(3)]
    t = torch.from_numpy(np.stack([p[0] for p in pairs])).to(dev)
    p = torch.from_numpy(np.stack([p[1] for p in pairs])).to(dev)
    bounds = torch.tensor([[int(a.max()), int(b.max())] for a, b in pairs], dtype=torch.int32, device=dev)
    thr = torch.tensor([0.5, 0.75, 0.9], dtype=torch.float64, device=dev)
    want_tp = O.average_precision([a for a, _ in pairs], [b for _, b in pairs])[1]
    for tt, pp, bb in ((t, p, bounds), (t, p, None), (t.reshape(3, -1), p.reshape(3, -1), bounds)):
        tab = K().label_overlap(tt, pp, capacity=1 << 15, cap_true=2048, cap_pred=8192, bounds=bb)
        assert tab["status"].cpu().tolist() == [0, 0, 0]
        c = K().match_counts(tab, thr).cpu().numpy()
        assert np.array_equal(counts_to_tp(c, [0.5, 0.75, 0.9]), want_tp.astype(np.int64))
        for i, (lt, lp) in enumerate(pairs):
            want = O.label_overlap(lt, lp)
            assert np.array_equal(tab["area_true"][i].cpu().numpy()[: want.shape[0]], want.sum(1).astype(np.int32))
            assert np.array_equal(tab["area_pred"][i].cpu().numpy()[: want.shape[1]], want.sum(0).astype(np.int32))
            assert tab["max_lbl"][i].cpu().tolist() == [lt.max(), lp.max()]
    # a bound that is too small is reported, not overrun
    low = bounds.clone()
    low[1, 1] = 1
    st = K().label_overlap(t, p, capacity=1 << 15, cap_true=2048, cap_pred=8192, bounds=low)["status"].cpu().tolist()
    assert st[1] == 2 and st[0] == 0 and st[2] == 0


def test_overlap_crowded_tiles_bypass_the_shared_table(dev, rng):
    """Every pixel its own label on both sides: a 128 x 128 tile holds 16 K distinct pairs, far more than the CTA's
    shared-memory table, so most pairs take the direct path to the global table."""
    H, W = 200, 260
    lt = (np.arange(H * W, dtype=np.int32).reshape(H, W) % 5000) + 1
    lp = rng.permutation(H * W).astype(np.int32).reshape(H, W) % 7000 + 1
    lp[::3] = 0
    t, p = torch.from_numpy(lt[None]).to(dev), torch.from_numpy(lp[None]).to(dev)
    tab = K().label_overlap(t, p, capacity=1 << 18, cap_true=5001, cap_pred=7001)
    assert tab["status"].item() == 0
    want = O.label_overlap(lt, lp)
    got = dense_overlap(tab, 0, want.shape[0], want.shape[1])
    assert np.array_equal(got[1:, 1:], want[1:, 1:])
    assert np.array_equal(tab["area_true"][0].cpu().numpy(), want.sum(1).astype(np.int32))
    assert np.array_equal(tab["area_pred"][0].cpu().numpy(), want.sum(0).astype(np.int32))


def test_overlap_table_overflow_and_range_are_reported(dev, rng):
    lt, lp = make_pair(rng)
    t = torch.from_numpy(lt[None]).to(dev)
    p = torch.from_numpy(lp[None]).to(dev)
    assert K().label_overlap(t, p, capacity=8, cap_true=512, cap_pred=4096)["status"].item() == 1
    assert K().label_overlap(t, p, capacity=1 << 14, cap_true=4, cap_pred=4096)["status"].item() == 2


def gpu_tp(lt, lp, thr, dev):
    t = torch.from_numpy(np.ascontiguousarray(lt[None].astype(np.int32))).to(dev)
    p = torch.from_numpy(np.ascontiguousarray(lp[None].astype(np.int32))).to(dev)
    tab = K().label_overlap(t, p, capacity=1 << 14, cap_true=1024, cap_pred=4096)
    assert tab["status"].item() == 0
    c = K().match_counts(tab, torch.tensor(thr, dtype=torch.float64, device=dev)).cpu().numpy()[0]
    return c[:, 0] - c[:, 1], c


def test_true_positives_equal_hungarian(dev, rng):
    thr = [0.5, 0.75, 0.9]
    for s in range(4):
        lt, lp = make_pair(rng, seed=10 + s)
        tp, _ = gpu_tp(lt, lp, thr, dev)
        iou = O.intersection_over_union(lt, lp)[1:, 1:]
        assert tp.tolist() == [O.true_positive(iou, th) for th in thr]


def test_iou_exactly_one_half_tie(dev):
    """Appendix B.3: one true object split into two preds of exactly half its area — two candidates at IoU 0.5,
    Hungarian matches one."""
    lt = np.zeros((8, 16), np.int32)
    lp = np.zeros((8, 16), np.int32)
    lt[2:4, 2:10] = 1
    lp[2:4, 2:6] = 1
    lp[2:4, 6:10] = 2
    lt[6:8, 0:4] = 2
    lt[6:8, 4:8] = 3  # and the mirrored case: two trues inside one pred
    lp[6:8, 0:8] = 3
    thr = [0.5, 0.75]
    tp, raw = gpu_tp(lt, lp, thr, dev)
    iou = O.intersection_over_union(lt, lp)[1:, 1:]
    assert [O.true_positive(iou, th) for th in thr] == [2, 0]
    assert raw[0].tolist() == [4, 2]  # four candidate pairs, two labels with a second partner
    assert tp.tolist() == [2, 0]


def test_label_stats_exact(dev, rng):
    lab = O.label(O.gaussian_blobs((300, 410), 60, seed=4) > 0.4, 2)
    n = int(lab.max())
    sums, bbox, status = K().label_stats(torch.from_numpy(lab[None]).to(dev), cap=n + 1)
    assert status.item() == 0
    sums, bbox = sums.cpu().numpy()[0], bbox.cpu().numpy()[0]
    for l in range(1, n + 1):
        ys, xs = np.nonzero(lab == l)
        want = [ys.size, 0, ys.sum(), xs.sum(), 0, (ys * ys).sum(), (xs * xs).sum(), 0, 0, (ys * xs).sum()]
        assert sums[l].tolist() == want
        assert bbox[l].tolist() == [0, ys.min(), xs.min(), 0, ys.max(), xs.max()]


def test_label_stats_3d(dev):
    lab = O.label(O.gaussian_blobs((40, 50, 60), 30, sigma=(2.0, 5.0), seed=5) > 0.4, 3)
    n = int(lab.max())
    sums, bbox, _ = K().label_stats(torch.from_numpy(lab[None]).to(dev), cap=n + 1)
    sums = sums.cpu().numpy()[0]
    for l in range(1, n + 1):
        zs, ys, xs = np.nonzero(lab == l)
        want = [zs.size, zs.sum(), ys.sum(), xs.sum(), (zs * zs).sum(), (ys * ys).sum(), (xs * xs).sum(),
                (zs * ys).sum(), (zs * xs).sum(), (ys * xs).sum()]
        assert sums[l].tolist() == want
