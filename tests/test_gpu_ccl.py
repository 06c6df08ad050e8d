"""Connected-component labelling: labels AND numbering bit-exact against scipy.ndimage.label (the oracle)."""
import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def K():
    from bioengine_worker_b200 import kernels

    return kernels


def run(mask_np, conn, dev):
    m = torch.from_numpy(np.ascontiguousarray(mask_np.astype(np.uint8))).to(dev)
    labels, n = K().ccl(m, conn)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), n.cpu().numpy()


@pytest.mark.parametrize("conn", [1, 2])
@pytest.mark.parametrize("shape", [(6, 6), (1, 1), (64, 64), (65, 63), (200, 333), (512, 512), (1, 300), (300, 1)])
@pytest.mark.parametrize("density", [0.1, 0.45, 0.6, 0.9])
def test_ccl_2d_random(shape, conn, density, dev, rng):
    mask = rng.random((2,) + shape) < density
    got, n = run(mask, conn, dev)
    for i in range(2):
        want = O.label(mask[i], conn)
        assert np.array_equal(got[i], want)
        assert n[i] == want.max()


def test_ccl_2d_structured(dev):
    H, W = 300, 400
    yy, xx = np.mgrid[:H, :W]
    cases = {
        "empty": np.zeros((H, W), bool),
        "full": np.ones((H, W), bool),
        "checker": (yy + xx) % 2 == 0,          # one component under 8-conn, H*W/2 under 4-conn
        "hstripes": yy % 2 == 0,
        "vstripes": xx % 2 == 0,
        "spiral": ((np.hypot(yy - H / 2, xx - W / 2) + 6 * np.arctan2(yy - H / 2, xx - W / 2)) % 12) < 5,
        "diag": (yy - xx) % 7 == 0,
        "anti": (yy + xx) % 7 == 0,
    }
    for name, m in cases.items():
        for conn in (1, 2):
            got, n = run(m[None], conn, dev)
            want = O.label(m, conn)
            assert np.array_equal(got[0], want), (name, conn)
            assert n[0] == want.max()


def test_ccl_blobs_like_the_pipeline(dev):
    img = O.gaussian_blobs((1024, 1024), 400, seed=11)
    for thr in (0.2, 0.5):
        m = img > thr
        got, n = run(m[None], 2, dev)
        assert np.array_equal(got[0], O.label(m, 2))


@pytest.mark.parametrize("conn", [1, 2, 3])
@pytest.mark.parametrize("shape", [(5, 6, 7), (8, 8, 64), (9, 17, 65), (40, 50, 70), (3, 100, 130)])
def test_ccl_3d_random(shape, conn, dev, rng):
    for density in (0.08, 0.3, 0.55):
        mask = rng.random((1,) + shape) < density
        got, n = run(mask, conn, dev)
        want = O.label(mask[0], conn)
        assert np.array_equal(got[0], want), (shape, conn, density)
        assert n[0] == want.max()


def test_ccl_3d_blobs(dev):
    vol = O.gaussian_blobs((96, 128, 160), 150, sigma=(3.0, 7.0), seed=3)
    m = vol > 0.4
    got, n = run(m[None], 3, dev)
    assert np.array_equal(got[0], O.label(m, 3))


def test_ccl_full_size_2d_properties(dev):
    """4096^2 at ~40 % density (~265 k components): full comparison with the oracle."""
    rng = np.random.default_rng(2)
    m = rng.random((4096, 4096)) < 0.4
    got, n = run(m[None], 2, dev)
    want = O.label(m, 2)
    assert n[0] == want.max()
    assert np.array_equal(got[0], want)


def test_ccl_idempotent_on_own_output(dev, rng):
    m = rng.random((1, 500, 700)) < 0.5
    got, _ = run(m, 2, dev)
    again, _ = run(got > 0, 2, dev)
    assert np.array_equal(got, again)


def pack_bits(mask_np):
    """(..., X) bool → (..., ceil(X/64)) int64 words, lowest bit = first pixel of the word (the kernels' layout)."""
    X = mask_np.shape[-1]
    W = (X + 63) // 64
    padded = np.zeros(mask_np.shape[:-1] + (W * 64,), np.uint8)
    padded[..., :X] = mask_np
    by = np.packbits(padded.reshape(mask_np.shape[:-1] + (W, 64)), axis=-1, bitorder="little")
    return np.ascontiguousarray(by).view(np.uint64).reshape(mask_np.shape[:-1] + (W,)).astype(np.int64)


@pytest.mark.parametrize("shape,conn", [((2, 200, 333), 2), ((1, 64, 64), 1), ((3, 129, 640), 2), ((1, 700, 1), 2),
                                        ((2, 9, 40, 130), 3), ((1, 16, 16, 256), 1)])
def test_ccl_from_bit_packed_mask_and_from_logits(shape, conn, dev, rng):
    """The three front ends of the labelling kernel (uint8 mask, bit-packed mask, fp32 values with the threshold fused)
    must give the same labels: scipy's, numbering included."""
    from bioengine_worker_b200 import _lib

    x = (rng.standard_normal(shape) * 1.5 - 0.6).astype(np.float32)
    x[..., ::7, :] += 1.0
    for epi, mask in ((_lib.BE_EPI_SIGMOID, O.binarize(O.sigmoid(x), 0.5)), (_lib.BE_EPI_NONE, O.binarize(x, 0.25))):
        thr = 0.5 if epi == _lib.BE_EPI_SIGMOID else 0.25
        want = np.stack([O.label(m, conn) for m in mask])
        got_u8, n_u8 = run(mask, conn, dev)
        assert np.array_equal(got_u8, want)
        bits = torch.from_numpy(pack_bits(mask.astype(bool))).to(dev)
        # garbage past the end of ragged rows must be ignored
        if shape[-1] % 64:
            bits[..., -1] |= torch.tensor(-1, dtype=torch.int64, device=dev) << (shape[-1] % 64)
        from bioengine_worker_b200.kernels import Workspace
        fresh = Workspace()   # not the workspace the uint8 call above left its packed mask in
        fresh.get(1 << 26, dev).fill_(0xA5)
        got_b, n_b = K().ccl_bits(bits.contiguous(), shape, conn, ws=fresh)
        assert np.array_equal(got_b.cpu().numpy(), want) and np.array_equal(n_b.cpu().numpy(), n_u8)
        got_f, n_f = K().ccl_logits(torch.from_numpy(x).to(dev), epi, thr, conn)
        assert np.array_equal(got_f.cpu().numpy(), want) and np.array_equal(n_f.cpu().numpy(), n_u8)


def test_ccl_logits_values_planted_on_the_threshold(dev):
    """sigmoid(x) > 0.5 decided inside the labelling kernel: values within a few ulp of logit(0.5) = 0 and exactly on it
    must fall on the side numpy's float32 sigmoid puts them."""
    from bioengine_worker_b200 import _lib

    base = np.float32(0.0)
    vals = [base]
    for _ in range(40):
        vals.append(np.nextafter(vals[-1], np.float32(1)))
    vals = np.array(vals + [-v for v in vals] + [1e-4, -1e-4, 9.9e-5, -9.9e-5, 1.01e-4], np.float32)
    x = np.tile(vals, 50)[: 64 * 64].reshape(1, 64, 64).copy()
    mask = O.binarize(O.sigmoid(x), 0.5)
    got, _ = K().ccl_logits(torch.from_numpy(x).to(dev), _lib.BE_EPI_SIGMOID, 0.5, 2)
    assert np.array_equal(got.cpu().numpy()[0], O.label(mask[0], 2))
