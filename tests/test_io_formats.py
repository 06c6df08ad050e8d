"""Wire formats of the entry path: file content → ndarray as `imageio.v3.imread` would return it
(ref: apps/model-runner/entry_deployment.py:29,1196-1263).  CPU only."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

from bioengine_worker_b200 import io as bio
from bioengine_worker_b200.entry_deployment import EntryDeployment


def _enc(ext, arr):
    ok, buf = cv2.imencode(ext, arr)
    assert ok
    return buf.tobytes()


def test_png_tiff_npy_round_trip(tmp_path):
    rng = np.random.default_rng(0)
    g8 = rng.integers(0, 256, (37, 53), dtype=np.uint8)
    g16 = rng.integers(0, 65536, (37, 53), dtype=np.uint16)
    rgb = rng.integers(0, 256, (20, 31, 3), dtype=np.uint8)
    f32 = rng.random((19, 23)).astype(np.float32)
    assert np.array_equal(bio.decode_image(_enc(".png", g8), ".png"), g8)
    assert np.array_equal(bio.decode_image(_enc(".png", g16), ".PNG"), g16)
    assert np.array_equal(bio.decode_image(_enc(".png", rgb[:, :, ::-1]), ".png"), rgb)       # file holds RGB, array is RGB
    assert np.array_equal(bio.decode_image(_enc(".tiff", g16), ".tiff"), g16)
    got = bio.decode_image(_enc(".tif", f32), ".tif")
    assert got.dtype == np.float32 and np.array_equal(got, f32)
    jpg = bio.decode_image(_enc(".jpg", g8), ".jpg")
    assert jpg.shape == g8.shape and jpg.dtype == np.uint8
    p = tmp_path / "a.npy"
    np.save(p, f32)
    assert np.array_equal(bio.load_image(str(p)), f32)
    p2 = tmp_path / "b.tif"
    p2.write_bytes(_enc(".tif", g16))
    assert np.array_equal(EntryDeployment._load_image_from_source("file://" + str(p2)), g16)


def test_multipage_tiff(tmp_path):
    rng = np.random.default_rng(1)
    pages = [rng.integers(0, 65536, (16, 24), dtype=np.uint16) for _ in range(3)]
    p = tmp_path / "stack.tif"
    assert cv2.imwritemulti(str(p), pages)
    got = bio.load_image(str(p))
    assert got.shape == (3, 16, 24) and all(np.array_equal(got[i], pages[i]) for i in range(3))


def test_errors_match_the_reference(tmp_path):
    with pytest.raises(ValueError, match="Unsupported file extension"):
        bio.load_image(str(tmp_path / "x.bmp"))
    with pytest.raises(FileNotFoundError):
        bio.load_image(str(tmp_path / "missing.png"))
    bad = tmp_path / "bad.png"
    bad.write_bytes(b"not an image")
    with pytest.raises(ValueError, match="Failed to parse image"):
        bio.load_image(str(bad))
