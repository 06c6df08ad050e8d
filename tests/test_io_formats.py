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


def test_fetch_image_over_http(tmp_path):
    """Direct URL sources (ref: entry_deployment.py:1217-1258) against a local HTTP server: .npy and .png contents, the
    query string ignored for the extension, 404 → FileNotFoundError, unsupported extension / unreadable content →
    ValueError, nobody listening → FileNotFoundError."""
    import functools
    import http.server
    import threading

    import cv2

    from bioengine_worker_b200.io import fetch_image

    rng = np.random.default_rng(0)
    arr = rng.normal(size=(5, 7)).astype(np.float32)
    img = rng.integers(0, 65535, size=(9, 11), dtype=np.uint16)
    np.save(tmp_path / "a.npy", arr)
    cv2.imwrite(str(tmp_path / "b.png"), img)
    (tmp_path / "broken.png").write_bytes(b"not a png")
    (tmp_path / "c.txt").write_text("x")

    class Quiet(http.server.SimpleHTTPRequestHandler):
        def log_message(self, *a):
            pass

    srv = http.server.ThreadingHTTPServer(("127.0.0.1", 0), functools.partial(Quiet, directory=str(tmp_path)))
    port = srv.server_address[1]
    t = threading.Thread(target=srv.serve_forever, daemon=True)
    t.start()
    try:
        base = f"http://127.0.0.1:{port}"
        assert np.array_equal(fetch_image(f"{base}/a.npy"), arr)
        assert np.array_equal(fetch_image(f"{base}/b.png?X-Amz-Signature=abc"), img)
        with pytest.raises(FileNotFoundError, match="does not exist or has expired"):
            fetch_image(f"{base}/missing.npy", retries=1)
        with pytest.raises(ValueError, match="Unsupported file extension"):
            fetch_image(f"{base}/c.txt")
        with pytest.raises(ValueError, match="Failed to parse image"):
            fetch_image(f"{base}/broken.png")
    finally:
        srv.shutdown()
        srv.server_close()
    with pytest.raises(FileNotFoundError, match="Failed to download"):
        fetch_image(f"http://127.0.0.1:{port}/a.npy", timeout=2.0, retries=1)   # the server is gone
