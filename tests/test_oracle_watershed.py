"""The oracle's restatement of skimage's peak_local_max / watershed (third-party, not installable here: parity unpinned)
held to hand-derived known answers, and its two tie orders compared."""
import numpy as np
import scipy.ndimage as ndi

from oracle import oracle as O


def test_flood_on_a_line_known_answers():
    # two seeds of equal height at the ends of a ridge: they leave in raster order, the fronts meet in the middle and the
    # pixel both reach at the same height goes to the front that was pushed first
    h = np.array([[0, 1, 2, 3, 2, 1, 0]], dtype=np.float64)
    mk = np.array([[1, 0, 0, 0, 0, 0, 2]], dtype=np.int32)
    mask = np.ones_like(mk, dtype=bool)
    assert O.watershed(h, mk, mask, "fifo").tolist() == [[1, 1, 1, 1, 2, 2, 2]]
    # a flood that crosses a crest runs down into the basin behind it before anything higher is touched
    h = np.array([[0, 5, 1, 1, 5, 3]], dtype=np.float64)
    mk = np.array([[1, 0, 0, 0, 0, 2]], dtype=np.int32)
    for ties in ("fifo", "heap"):
        assert O.watershed(h, mk, np.ones_like(mk, dtype=bool), ties).tolist() == [[1, 1, 1, 1, 2, 2]]
    # the mask stops the flood; unreached pixels stay 0
    mask = np.array([[1, 1, 0, 1, 1, 1]], dtype=bool)
    assert O.watershed(h, mk, mask, "fifo").tolist() == [[1, 1, 0, 2, 2, 2]]


def test_flood_neighbour_order_up_left_right_down():
    # flat image, one seed in the middle of a plus-shaped mask and a second seed that competes for the corner pixels:
    # on a plateau pixels leave in push order, and a pixel pushes up, left, right, down in that order
    h = np.zeros((3, 3))
    mk = np.zeros((3, 3), dtype=np.int32)
    mk[1, 1] = 1
    out = O.watershed(h, mk, np.ones((3, 3), dtype=bool), "fifo")
    assert (out == 1).all()
    mk[0, 0] = 2          # raster-first seed: pushes (0,1) and (1,0) before seed 1 moves
    out = O.watershed(h, mk, np.ones((3, 3), dtype=bool), "fifo")
    assert out.tolist() == [[2, 2, 2], [2, 1, 1], [2, 1, 1]]


def test_heap_and_fifo_agree_without_tied_seeds():
    rng = np.random.default_rng(0)
    for _ in range(5):
        h = rng.random((60, 70))
        mask = ndi.binary_dilation(rng.random((60, 70)) < 0.05, iterations=4)
        mk = np.zeros((60, 70), dtype=np.int32)
        ys, xs = np.nonzero(mask)
        pick = rng.choice(len(ys), 25, replace=False)
        mk[ys[pick], xs[pick]] = np.arange(1, 26)
        assert np.array_equal(O.watershed(h, mk, mask, "heap"), O.watershed(h, mk, mask, "fifo"))


def test_peak_local_max_known_answers():
    # a 41x60 bar: distance ridge of height 5 along its middle row; border stripping removes nothing (bar is inside);
    # the window maximum keeps the ridge pixels 5 away from the bar's ends, the thinning keeps every 8th in raster order
    mask = np.zeros((41, 60), dtype=bool)
    mask[16:25, 10:50] = True
    dist = O.distance_transform_edt(mask)
    c = O.peak_local_max(dist, 8, mask)
    assert (c[:, 0] == 20).all()
    assert c[:, 1].tolist() == [14, 22, 30, 38]          # ridge x = 14..45 at height 5, greedy from the left: 14, 22, 30, 38
    # the same bar pushed against the left border: columns < 8 are stripped from the label before the bounding box is
    # taken, so the box — and with it the candidates — starts at column 8
    mask = np.zeros((41, 60), dtype=bool)
    mask[16:25, 0:30] = True
    c = O.peak_local_max(O.distance_transform_edt(mask), 8, mask)
    assert (c[:, 0] == 20).all() and c[:, 1].min() == 8
    # nothing but border → no seeds; a lone disc → its centre
    assert len(O.peak_local_max(np.ones((20, 20)), 8, np.ones((20, 20), dtype=bool))) == 0
    yy, xx = np.mgrid[:50, :50]
    disc = (yy - 25) ** 2 + (xx - 24) ** 2 <= 100
    assert O.peak_local_max(O.distance_transform_edt(disc), 8, disc).tolist() == [[25, 24]]


def test_prob_to_instances_splits_touching_discs():
    yy, xx = np.mgrid[:80, :120]
    prob = (((yy - 40) ** 2 + (xx - 40) ** 2 <= 18**2) | ((yy - 40) ** 2 + (xx - 72) ** 2 <= 17**2)).astype(np.float32)
    out = O.prob_to_instances(prob)
    assert out.max() == 2 and ndi.label(prob > 0.5)[1] == 1
    assert out[40, 30] == 1 and out[40, 85] == 2
    assert ((out > 0) == O.binary_closing_disk(prob > 0.5, 4)).all()
    assert O.prob_to_instances(np.zeros((30, 30), np.float32)).max() == 0
