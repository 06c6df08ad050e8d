"""Mask clean-up and distance transform (first part of the fibsem `_prob_to_instances`) against scipy: all bit-exact."""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def K():
    from bioengine_worker_b200 import kernels

    return kernels


def masks(rng, shape, kind):
    if kind == "blobs":
        return np.stack([O.gaussian_blobs(shape, 25, seed=s) + 0.15 * rng.standard_normal(shape) > 0.45 for s in (1, 2)])
    if kind == "noise":
        return rng.random((2,) + shape) < 0.5
    if kind == "sparse":
        return rng.random((2,) + shape) < 0.03
    if kind == "border":
        m = np.zeros((2,) + shape, bool)
        m[0, :7, :], m[0, :, -5:], m[1, 3:-3, 3:-3] = True, True, True
        return m
    raise ValueError(kind)


@pytest.mark.parametrize("shape", [(64, 64), (97, 130), (300, 257), (5, 33)])
@pytest.mark.parametrize("kind", ["blobs", "noise", "sparse", "border"])
def test_remove_small_objects_closing_edt(shape, kind, dev, rng):
    m = masks(rng, shape, kind)
    md = torch.from_numpy(m.astype(np.uint8)).to(dev)
    for min_size in (1, 5, 300):
        got = K().remove_small_objects(md, min_size).cpu().numpy()
        want = np.stack([O.remove_small_objects(x, min_size) for x in m])
        assert np.array_equal(got.astype(bool), want), (kind, min_size)
    hw = K().disk_half_widths(4)
    assert hw.tolist() == [0, 2, 3, 3, 4, 3, 3, 2, 0] and np.array_equal(O.disk(4).sum(1), 2 * hw + 1)
    closed = K().binary_closing(md, hw)
    want_closed = np.stack([O.binary_closing_disk(x, 4) for x in m])
    assert np.array_equal(closed.cpu().numpy().astype(bool), want_closed), kind
    for radius in (1, 2, 7):
        got = K().binary_closing(md, K().disk_half_widths(radius)).cpu().numpy().astype(bool)
        assert np.array_equal(got, np.stack([O.binary_closing_disk(x, radius) for x in m])), (kind, radius)
    for src, name in ((md, "mask"), (closed, "closed")):
        a = src.cpu().numpy()
        if not (a == 0).any(axis=(1, 2)).all():
            continue  # no zero pixel: undefined
        dist = K().distance_transform_edt(src).cpu().numpy()
        want = np.stack([O.distance_transform_edt(x) for x in a])
        assert dist.dtype == np.float64 and np.array_equal(dist, want), (kind, name)


def test_prob_to_instances_front_half_full_size(dev):
    """threshold → remove_small_objects(300) → closing(disk 4) → EDT on a 2048^2 probability-like map, vs scipy."""
    rng = np.random.default_rng(5)
    prob = O.gaussian_blobs((2048, 2048), 900, seed=9) + 0.12 * rng.standard_normal((2048, 2048)).astype(np.float32)
    binary = prob > 0.5
    md = torch.from_numpy(binary.astype(np.uint8)).to(dev)[None]
    kept = K().remove_small_objects(md, 300)
    want_kept = O.remove_small_objects(binary, 300)
    assert np.array_equal(kept.cpu().numpy()[0].astype(bool), want_kept)
    closed = K().binary_closing(kept, K().disk_half_widths(4))
    want_closed = O.binary_closing_disk(want_kept, 4)
    assert np.array_equal(closed.cpu().numpy()[0].astype(bool), want_closed)
    dist = K().distance_transform_edt(closed).cpu().numpy()[0]
    assert np.array_equal(dist, O.distance_transform_edt(want_closed))


def test_morph_errors(dev):
    from bioengine_worker_b200._lib import BeError

    m = torch.zeros((1, 8, 8), dtype=torch.uint8, device=dev)
    with pytest.raises(BeError):
        K().binary_closing(m, np.array([0, 1, 2], np.int32))  # not symmetric
    with pytest.raises(BeError):
        K().binary_closing(m, np.full(41, 1, np.int32))       # radius 20 > 15
