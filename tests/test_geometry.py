"""Host tile tables (product code) against the oracle's np.pad / slicing restatement of the reference."""
import numpy as np
import pytest

from bioengine_worker_b200 import geometry as G
from oracle import oracle as O


@pytest.mark.parametrize("n,lo,hi", [(7, -9, 20), (2, -5, 9), (1, -3, 4), (64, -200, 300)])
def test_reflect_index_is_numpy_reflect(n, lo, hi):
    base = np.arange(100, 100 + n)
    padded = np.pad(base, (-lo, hi - n), mode="reflect")
    idx = G.reflect_index(np.arange(lo, hi), 0, n)
    assert np.array_equal(base[idx], padded)


@pytest.mark.parametrize("S,B,halo", [(4096, 256, 32), (512, 128, 16), (200, 64, 8), (48, 48, 0), (100, 64, 16), (97, 33, 5)])
def test_bioimageio_axis_matches_oracle_blocks(S, B, halo):
    x = np.arange(S, dtype=np.float32)[None] * 3 + 1  # (1, S): one leading axis kept whole
    ax = G.bioimageio_axis(S, B, halo)
    blocks = [d[0] for _, d in O.split_into_blocks(x, (B,), (halo,))]
    assert ax.T == len(blocks)
    got = x[0][ax.src].reshape(ax.T, B)
    assert np.array_equal(got, np.stack(blocks))
    # reassembly: last writer wins
    tiles = (np.arange(ax.T * B, dtype=np.float32)).reshape(ax.T, 1, B)
    want = O.from_blocks(tiles, (1, 1, S), (B,), (halo,))[0, 0]
    assert np.array_equal(tiles.reshape(-1)[ax.sel], want)
    assert not np.isnan(want).any()


@pytest.mark.parametrize("S,B,ov", [(4096, 256, 64), (700, 512, 64), (513, 512, 64), (449, 64, 16), (1000, 128, 32)])
def test_fibsem_axis_matches_numpy_pad(S, B, ov):
    x = (np.arange(S, dtype=np.float32) * 2 + 5)
    ax = G.fibsem_axis(S, B, ov)
    tiles = O.fibsem_tiles(x, B, ov)
    assert ax.T == tiles.shape[0] == len(range(0, S, B - ov))
    assert np.array_equal(x[ax.src].reshape(ax.T, B), tiles)


def test_window_is_reference_window():
    w = G.gauss_window_1d(256)
    assert np.array_equal(np.multiply.outer(w, w).astype(np.float32), O.fibsem_window(256))


def test_too_small_sample_is_an_error():
    with pytest.raises(ValueError):
        G.bioimageio_axis(100, 256, 32)
    with pytest.raises(ValueError):
        G.bioimageio_axis(512, 64, 32)
    with pytest.raises(ValueError):
        G.fibsem_axis(512, 64, 64)
