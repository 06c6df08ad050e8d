"""Seeds + marker flood (second part) against the oracle's restatement of skimage's peak_local_max and
watershed (oracle/oracle.py, oracle/watershed.c; third-party algorithm, parity unpinned).  Bit-exact labels."""
import numpy as np
import pytest
import scipy.ndimage as ndi

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import oracle as O  # noqa: E402


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def _K():
    from bioengine_worker_b200 import kernels as K

    return K


def tubes_and_blobs(shape, seed):
    """Straight and diagonal tubes (ridges of equal distance), squares (flat tops), touching discs, things on the border."""
    rng = np.random.default_rng(seed)
    H, W = shape
    m = np.zeros(shape, dtype=bool)
    yy, xx = np.mgrid[:H, :W]
    for _ in range(6):
        y, x, r = rng.integers(20, H - 20), rng.integers(20, W - 20), rng.integers(6, 18)
        m |= (yy - y) ** 2 + (xx - x) ** 2 <= r * r
        m |= (yy - y) ** 2 + (xx - x - int(1.6 * r)) ** 2 <= r * r          # a touching twin of the same size: tied seeds
    for _ in range(4):
        y, x, h, w = rng.integers(0, H - 12), rng.integers(0, W - 60), rng.integers(5, 12), rng.integers(30, 120)
        m[y:y + h, x:x + w] = True                                           # horizontal tube (may touch the border)
    for _ in range(3):
        y, x, h, w = rng.integers(0, H - 60), rng.integers(0, W - 12), rng.integers(30, 100), rng.integers(5, 12)
        m[y:y + h, x:x + w] = True
    for _ in range(3):
        y, x, s = rng.integers(0, H - 40), rng.integers(0, W - 40), rng.integers(9, 30)
        m[y:y + s, x:x + s] = True
    d = np.abs((yy - xx) - rng.integers(-W // 3, H // 3)) <= 4
    m |= d & (yy > H // 4) & (yy < 3 * H // 4)
    return m


@pytest.mark.parametrize("shape,seed", [((200, 260), 0), ((333, 257), 1), ((96, 512), 2), ((70, 90), 3)])
def test_peak_local_max_matches_oracle(dev, shape, seed):
    K = _K()
    mask = tubes_and_blobs(shape, seed)
    dist = O.distance_transform_edt(mask)
    want = O.peak_local_max(dist, 8, mask)
    m = torch.from_numpy(mask.astype(np.uint8)).to(dev)
    d2 = K.distance_transform_edt_sq(m[None])[0]
    assert np.array_equal(d2.cpu().numpy(), np.rint(dist**2).astype(np.int32))
    peaks = K.peak_local_max(d2, m, 8).cpu().numpy()
    want_img = np.zeros(shape, dtype=np.uint8)
    if len(want):
        want_img[tuple(want.T)] = 1
    assert len(want) > 0
    assert np.array_equal(peaks, want_img)


@pytest.mark.parametrize("d", [1, 3, 8])
def test_peak_local_max_on_noise_and_plateaus(dev, d):
    """Not a distance map: integer noise with many equal values (the greedy thinning decides) inside an irregular label."""
    K = _K()
    rng = np.random.default_rng(d)
    img = rng.integers(0, 6, size=(150, 170)).astype(np.int32)
    mask = ndi.binary_dilation(rng.random((150, 170)) < 0.02, iterations=6)
    img[~mask] = 0
    want = O.peak_local_max(img.astype(np.float64), d, mask)
    peaks = K.peak_local_max(torch.from_numpy(img).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev), d).cpu().numpy()
    want_img = np.zeros(img.shape, dtype=np.uint8)
    want_img[tuple(want.T)] = 1
    assert want_img.sum() > 10
    assert np.array_equal(peaks, want_img)


def test_peak_local_max_trivial_image_and_empty(dev):
    """Every masked pixel is a window maximum (one-pixel-wide lines, distance 1 everywhere): skimage returns only what a
    binary opening removes; nothing at all for an empty or border-only mask."""
    K = _K()
    mask = np.zeros((64, 80), dtype=bool)
    mask[20, 10:70] = True
    mask[10:50, 40] = True
    mask[30:33, 20:23] = True    # a 3x3 block survives the opening (its centre) → not isolated
    dist = O.distance_transform_edt(mask)
    want = O.peak_local_max(dist, 8, mask)
    m = torch.from_numpy(mask.astype(np.uint8)).to(dev)
    peaks = K.peak_local_max(K.distance_transform_edt_sq(m[None])[0], m, 8).cpu().numpy()
    want_img = np.zeros(mask.shape, dtype=np.uint8)
    want_img[tuple(want.T)] = 1
    assert np.array_equal(peaks, want_img)
    for mk in (np.zeros((64, 80), dtype=bool), np.pad(np.zeros((48, 64), dtype=bool), 8, constant_values=True)):
        m = torch.from_numpy(mk.astype(np.uint8)).to(dev)
        assert len(O.peak_local_max(O.distance_transform_edt(mk), 8, mk)) == 0
        assert int(K.peak_local_max(K.distance_transform_edt_sq(m[None])[0], m, 8).sum().item()) == 0


@pytest.mark.parametrize("seed,levels", [(0, 4), (1, 40), (2, 100000)])
def test_watershed_random_heights_match_oracle(dev, seed, levels):
    """Random integer heights (few levels: huge plateaus; many: hardly a tie) on a random mask with random seeds: every
    flood runs down into lower basins all the time, which is the part of the order that is hard to get right."""
    K = _K()
    rng = np.random.default_rng(seed)
    shape = (120, 150)
    h = rng.integers(0, levels, size=shape).astype(np.int32)
    mask = ndi.binary_dilation(rng.random(shape) < 0.03, iterations=5)
    mk = np.zeros(shape, dtype=np.int32)
    ys, xs = np.nonzero(mask)
    pick = rng.choice(len(ys), size=60, replace=False)
    mk[ys[pick], xs[pick]] = np.arange(1, 61)
    want = O.watershed(h.astype(np.float64), mk, mask, ties="fifo")
    got = K.watershed(torch.from_numpy(h).to(dev), torch.from_numpy(mk).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev))
    assert np.array_equal(got.cpu().numpy(), want)
    if levels == 100000:   # seeds hardly ever tie: the binary heap gives the same flood
        assert (O.watershed(h.astype(np.float64), mk, mask, ties="heap") != want).mean() < 1e-3


@pytest.mark.parametrize("levels", [3, 1000])
def test_watershed_long_queues(dev, levels):
    """One open region of 1.2e5 pixels with noise heights: the queue of waiting pixels grows to tens of thousands of
    entries and shrinks again (many entries per thread in every pass over the queue)."""
    K = _K()
    rng = np.random.default_rng(levels)
    shape = (300, 400)
    h = rng.integers(0, levels, size=shape).astype(np.int32)
    mask = np.ones(shape, dtype=bool)
    mk = np.zeros(shape, dtype=np.int32)
    for l, (y, x) in enumerate(zip(rng.integers(0, 300, 7), rng.integers(0, 400, 7)), start=1):
        mk[y, x] = l
    want = O.watershed(h.astype(np.float64), mk, mask, ties="fifo")
    got = K.watershed(torch.from_numpy(h).to(dev), torch.from_numpy(mk).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev))
    assert np.array_equal(got.cpu().numpy(), want)
    st = K._default_ws._buf[:64].view(torch.int64).cpu().numpy()
    assert st[3] > 3072          # longest queue


def test_watershed_multi_pixel_seeds_and_unreached_regions(dev):
    K = _K()
    rng = np.random.default_rng(5)
    shape = (90, 130)
    h = (rng.integers(0, 9, size=shape) + np.add.outer(np.arange(90) % 17, np.arange(130) % 23)).astype(np.int32)
    mask = np.ones(shape, dtype=bool)
    mask[:, 64] = False          # two regions; the right one gets no seed and must stay 0
    mk = np.zeros(shape, dtype=np.int32)
    mk[10:14, 10:20] = 3
    mk[60, 5:40] = 1
    mk[80:85, 50:55] = 2
    want = O.watershed(h.astype(np.float64), mk, mask, ties="fifo")
    got = K.watershed(torch.from_numpy(h).to(dev), torch.from_numpy(mk).to(dev), torch.from_numpy(mask.astype(np.uint8)).to(dev)).cpu().numpy()
    assert np.array_equal(got, want)
    assert (got[:, 65:] == 0).all() and (got[:, :64] > 0).all()


@pytest.mark.parametrize("shape,seed", [((256, 320), 0), ((512, 512), 1), ((700, 333), 2)])
def test_instances_match_the_reference_post_processing(dev, shape, seed):
    """`_prob_to_instances` end to end on blob images whose blobs touch: threshold → remove_small_objects(300) →
    closing(disk 4) → EDT → peak_local_max(8) → label → watershed."""
    K = _K()
    img = O.gaussian_blobs(shape, shape[0] * shape[1] // 2500, sigma=(6.0, 16.0), seed=seed)
    prob = np.clip(img * 1.3, 0, 1).astype(np.float32)
    parts = {}
    want = O.prob_to_instances(prob, ties="fifo", parts=parts)
    mask = torch.from_numpy((prob > 0.5).astype(np.uint8)).to(dev)
    got, n_seeds = K.instances_from_mask(mask)
    assert int(n_seeds.item()) == len(parts["coords"]) == want.max()
    assert want.max() > ndi.label(parts["closed"])[1]          # some blobs were split
    assert np.array_equal(got.cpu().numpy(), want)
    heap = O.prob_to_instances(prob, ties="heap")
    diff = int((heap != want).sum())
    assert diff <= 0.001 * want.size, diff                     # complete ties between seeds: a few boundary pixels at most


def test_instances_of_tubes_and_tied_twins(dev):
    """Shapes built to produce seeds of exactly equal height inside one region (twin discs, long straight tubes): the
    first-in-first-out order is what the kernel promises; where skimage's heap order would differ is counted, not hidden."""
    K = _K()
    mask = tubes_and_blobs((400, 500), 7)
    want = O.prob_to_instances(mask.astype(np.float32), ties="fifo")
    got, _ = K.instances_from_mask(torch.from_numpy(mask.astype(np.uint8)).to(dev))
    assert np.array_equal(got.cpu().numpy(), want)
    heap = O.prob_to_instances(mask.astype(np.float32), ties="heap")
    assert ((heap > 0) == (want > 0)).all()
    assert (heap != want).mean() < 0.01


def test_instances_of_an_empty_mask(dev):
    K = _K()
    got, n = K.instances_from_mask(torch.zeros((128, 160), dtype=torch.uint8, device=dev))
    assert int(n.item()) == 0 and int(got.abs().sum().item()) == 0
    small = torch.zeros((128, 160), dtype=torch.uint8, device=dev)
    small[10:20, 10:20] = 1       # 100 px < 300: removed
    got, n = K.instances_from_mask(small)
    assert int(n.item()) == 0 and int(got.abs().sum().item()) == 0
