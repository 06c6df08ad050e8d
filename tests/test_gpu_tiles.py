"""Tile gather and both stitchers against the oracle: bit-exact tiles, bit-exact crop stitch, blend within 1e-5
(observed: bit-exact — same fp64 operation order)."""
import numpy as np
import pytest
import torch

from bioengine_worker_b200 import geometry as G
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def K():
    from bioengine_worker_b200 import kernels

    return kernels


def dev_axes(axes, dev):
    return [K().DeviceAxis(a, dev) for a in axes]


@pytest.mark.parametrize("shape,B,halo", [((2, 3, 200, 333), 64, 8), ((1, 1, 512, 512), 256, 32), ((1, 2, 97, 130), 33, 5)])
def test_gather_2d_bioimageio_bit_exact(shape, B, halo, dev, rng):
    x = rng.standard_normal(shape).astype(np.float32)
    axes = dev_axes([G.identity_axis(1), G.bioimageio_axis(shape[2], B, halo), G.bioimageio_axis(shape[3], B, halo)], dev)
    tiles = K().tile_gather(torch.from_numpy(x).to(dev), axes).cpu().numpy()
    assert np.array_equal(tiles, O.blocks_tensor(x, (B, B), (halo, halo)))


def test_gather_3d_bioimageio_bit_exact(dev, rng):
    shape, B, halo = (1, 2, 70, 90, 100), 32, 4
    x = rng.standard_normal(shape).astype(np.float32)
    axes = dev_axes([G.bioimageio_axis(s, B, halo) for s in shape[2:]], dev)
    tiles = K().tile_gather(torch.from_numpy(x).to(dev), axes).cpu().numpy()
    assert np.array_equal(tiles, O.blocks_tensor(x, (B,) * 3, (halo,) * 3))


def test_gather_fused_normalise_equals_normalise_then_gather(dev):
    img = O.synthetic_image((2, 1, 300, 420), 80, seed=5)
    xd = torch.from_numpy(img).to(dev)
    q = torch.tensor([0.01, 0.998], dtype=torch.float64, device=dev)
    _, norm = K().quantile(xd, 2, q, 1e-6)
    axes = dev_axes([G.identity_axis(1), G.bioimageio_axis(300, 128, 16), G.bioimageio_axis(420, 128, 16)], dev)
    fused = K().tile_gather(xd, axes, norm=norm, norm_mode=1).cpu().numpy()
    want = O.blocks_tensor(O.scale_range(img, 1.0, 99.8, axes=(1, 2, 3)), (128, 128), (16, 16))
    assert np.array_equal(fused, want)


@pytest.mark.parametrize("H,W,B,ov", [(700, 900, 256, 64), (513, 512, 512, 64), (300, 449, 64, 16)])
def test_fibsem_gather_and_blend_2d(H, W, B, ov, dev, rng):
    img = rng.random((H, W)).astype(np.float32)
    axes = dev_axes([G.identity_axis(1), G.fibsem_axis(H, B, ov), G.fibsem_axis(W, B, ov)], dev)
    tiles = K().tile_gather(torch.from_numpy(img).to(dev)[None, None], axes)
    want_tiles = O.fibsem_tiles(img, B, ov)
    assert np.array_equal(tiles.cpu().numpy()[:, 0], want_tiles)
    # a fake "network": a fixed nonlinear map per tile, so overlapping tiles disagree and blending matters
    probs = torch.sin(tiles * 3.0) * (1.0 + 0.01 * torch.arange(tiles.shape[0], device=dev).view(-1, 1, 1, 1))
    probs = probs.contiguous()
    w1 = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
    one = torch.ones(1, dtype=torch.float64, device=dev)
    out, mask = K().stitch_blend(probs, (1, 1, H, W), axes, (one, w1, w1), want_mask=True, thr=0.5)
    want = O.fibsem_blend(probs.cpu().numpy()[:, 0], (H, W), B, ov)
    np.testing.assert_allclose(out.cpu().numpy()[0, 0], want, rtol=1e-5, atol=0)
    assert np.array_equal(out.cpu().numpy()[0, 0], want)
    assert np.array_equal(mask.cpu().numpy()[0, 0], O.binarize(want, 0.5))


def test_fibsem_blend_3d(dev, rng):
    shape, B, ov = (70, 81, 90), 32, 8
    vol = rng.random(shape).astype(np.float32)
    axes = dev_axes([G.fibsem_axis(s, B, ov) for s in shape], dev)
    tiles = K().tile_gather(torch.from_numpy(vol).to(dev)[None, None], axes)
    assert np.array_equal(tiles.cpu().numpy()[:, 0], O.fibsem_tiles(vol, B, ov))
    probs = (tiles * tiles).contiguous()
    w1 = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
    out, _ = K().stitch_blend(probs, (1, 1) + shape, axes, (w1, w1, w1))
    want = O.fibsem_blend(probs.cpu().numpy()[:, 0], shape, B, ov)
    assert np.array_equal(out.cpu().numpy()[0, 0], want)


@pytest.mark.parametrize("shape,B,halo", [((2, 2, 200, 333), 64, 8), ((1, 1, 512, 512), 256, 32)])
def test_crop_stitch_round_trip_and_last_writer(shape, B, halo, dev, rng):
    x = rng.standard_normal(shape).astype(np.float32)
    axes = dev_axes([G.identity_axis(1), G.bioimageio_axis(shape[2], B, halo), G.bioimageio_axis(shape[3], B, halo)], dev)
    tiles = K().tile_gather(torch.from_numpy(x).to(dev), axes)
    out, _ = K().stitch_crop(tiles, shape, axes)
    assert np.array_equal(out.cpu().numpy(), x)  # gather → crop is the identity
    # distinct per-tile content shows which block wins in the shifted-last-block overlap
    t2 = torch.from_numpy(rng.standard_normal(tuple(tiles.shape)).astype(np.float32)).to(dev)
    out2, mask2 = K().stitch_crop(t2, shape, axes, want_mask=True, thr=0.25)
    want = O.from_blocks(t2.cpu().numpy(), shape, (B, B), (halo, halo))
    assert np.array_equal(out2.cpu().numpy(), want)
    assert np.array_equal(mask2.cpu().numpy(), O.binarize(want, 0.25))
    # fused sigmoid epilogue == stitch then sigmoid op
    out3, mask3 = K().stitch_crop(t2, shape, axes, epilogue=1, want_mask=True, thr=0.5)
    assert np.array_equal(out3.cpu().numpy(), K().sigmoid(out2).cpu().numpy())
    assert np.array_equal(mask3.cpu().numpy(), O.binarize(O.sigmoid(want), 0.5))


def test_crop_stitch_3d(dev, rng):
    shape, B, halo = (1, 1, 70, 90, 100), 32, 4
    axes = dev_axes([G.bioimageio_axis(s, B, halo) for s in shape[2:]], dev)
    T = axes[0].ax.T * axes[1].ax.T * axes[2].ax.T
    t = torch.from_numpy(rng.standard_normal((T, 1, B, B, B)).astype(np.float32)).to(dev)
    out, _ = K().stitch_crop(t, shape, axes)
    assert np.array_equal(out.cpu().numpy(), O.from_blocks(t.cpu().numpy(), shape, (B,) * 3, (halo,) * 3))


def test_c2_geometry_full_size_round_trip(dev):
    """BASELINE config 2: 4096^2, 256-px blocks incl. 32-px halo → 484 tiles; gather→crop must return the image."""
    g = torch.Generator(device=dev).manual_seed(2)
    x = torch.randn((1, 1, 4096, 4096), device=dev, generator=g)
    axes = dev_axes([G.identity_axis(1), G.bioimageio_axis(4096, 256, 32), G.bioimageio_axis(4096, 256, 32)], dev)
    tiles = K().tile_gather(x, axes)
    assert tuple(tiles.shape) == (484, 1, 256, 256)
    out, _ = K().stitch_crop(tiles, (1, 1, 4096, 4096), axes)
    assert torch.equal(out, x)
    # blend of tiles cut from one image returns that image up to rounding (weights cancel)
    faxes = dev_axes([G.identity_axis(1), G.fibsem_axis(4096, 256, 64), G.fibsem_axis(4096, 256, 64)], dev)
    ft = K().tile_gather(x, faxes)
    w1 = torch.from_numpy(G.gauss_window_1d(256)).to(dev)
    one = torch.ones(1, dtype=torch.float64, device=dev)
    bl, _ = K().stitch_blend(ft, (1, 1, 4096, 4096), faxes, (one, w1, w1))
    torch.testing.assert_close(bl, x, rtol=1e-5, atol=1e-6)


# ---- the TMA-fed kernels (aligned geometry) next to the general ones: every path of both ----

@pytest.mark.parametrize("shape,block,halo", [
    ((2, 3, 96, 128), (64, 64), (8, 8)),        # batch + channels, interior and border tiles in one launch
    ((1, 1, 300, 64), (200, 16), (4, 2)),       # 16-px-wide tiles: more than 64 rows per shared-memory stage
    ((1, 2, 24, 40, 64), (8, 16, 32), (1, 2, 4)),  # 3-D, different block per axis
    ((1, 1, 64, 1024), (32, 512), (4, 16)),     # 512-px-wide tiles (two 4-px-unit sweeps per row group)
])
def test_gather_tma_paths_bit_exact(shape, block, halo, dev, rng):
    x = rng.standard_normal(shape).astype(np.float32)
    nd = len(block)
    axes = [G.bioimageio_axis(s, b, h) for s, b, h in zip(shape[2:], block, halo)]
    if nd == 2:
        axes = [G.identity_axis(1)] + axes
    tiles = K().tile_gather(torch.from_numpy(x).to(dev), dev_axes(axes, dev)).cpu().numpy()
    assert np.array_equal(tiles, O.blocks_tensor(x, block, halo))
    # the fused normalisation through the same kernel: per-sample constants, clip
    norm = torch.tensor([[0.1, 1.7]] * shape[0], dtype=torch.float64, device=dev)
    fused = K().tile_gather(torch.from_numpy(x).to(dev), dev_axes(axes, dev), norm=norm, norm_mode=2, clip=(0.0, 1.0))
    want = np.clip((x.astype(np.float64) - 0.1) / 1.7, 0.0, 1.0).astype(np.float32)
    assert np.array_equal(fused.cpu().numpy(), O.blocks_tensor(want, block, halo))


@pytest.mark.parametrize("spatial,B,ov,N,C", [
    ((128, 256), 64, 16, 2, 2),      # rows kernel, several (n, c) planes
    ((40, 52, 96), 32, 8, 1, 1),     # rows kernel, 3-D: a row's kz tiles are consecutive steps
    ((40, 52, 96), 32, 8, 1, 2),
    ((70, 81, 90), 32, 8, 1, 1),     # X not a multiple of 4: the general kernel
])
@pytest.mark.parametrize("thr", [0.5, 0.3])
def test_blend_kernels_all_epilogues(spatial, B, ov, N, C, thr, dev, rng):
    nd = len(spatial)
    axes = [G.fibsem_axis(s, B, ov) for s in spatial]
    if nd == 2:
        axes = [G.identity_axis(1)] + axes
    T = int(np.prod([a.T for a in axes]))
    logits = (rng.standard_normal((N * T, C) + (B,) * nd) * 2).astype(np.float32)
    w1 = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
    wins = (w1,) * 3 if nd == 3 else (torch.ones(1, dtype=torch.float64, device=dev), w1, w1)
    ld = torch.from_numpy(logits).to(dev)
    daxes = dev_axes(axes, dev)
    shape = (N, C) + tuple(spatial)
    want = np.stack([np.stack([O.fibsem_blend(logits[n * T:(n + 1) * T, c], spatial, B, ov) for c in range(C)])
                     for n in range(N)])
    out, mask = K().stitch_blend(ld, shape, daxes, wins, want_mask=True, thr=thr)             # raw blend + threshold
    assert np.array_equal(out.cpu().numpy(), want)
    assert np.array_equal(mask.cpu().numpy(), O.binarize(want, thr))
    prob, pmask = K().stitch_blend(ld, shape, daxes, wins, epilogue=1, want_mask=True, thr=thr)   # sigmoid, both outputs
    assert np.array_equal(prob.cpu().numpy(), O.sigmoid(want))
    assert np.array_equal(pmask.cpu().numpy(), O.binarize(O.sigmoid(want), thr))
    _, only = K().stitch_blend(ld, shape, daxes, wins, epilogue=1, want_out=False, want_mask=True, thr=thr)  # band test
    assert np.array_equal(only.cpu().numpy(), O.binarize(O.sigmoid(want), thr))
    # the mask alone, bit-packed, for both epilogues (2-D: the fp32 interval filter decides most pixels)
    from tests.test_gpu_ccl import pack_bits
    table = K().blend_window(wins) if nd == 2 else None   # 2-D + table: the fp32 interval filter kernel
    for epi, m in ((1, O.binarize(O.sigmoid(want), thr)), (0, O.binarize(want, thr))):
        for tab in (None, table):
            _, none, bits = K().stitch_blend(ld, shape, daxes, wins, epilogue=epi, want_out=False, want_bits=True, thr=thr,
                                             window=tab)
            assert none is None and np.array_equal(bits.cpu().numpy(), pack_bits(m.astype(bool)))
            _, both, bits2 = K().stitch_blend(ld, shape, daxes, wins, epilogue=epi, want_out=False, want_mask=True,
                                              want_bits=True, thr=thr, window=tab)
            assert np.array_equal(both.cpu().numpy(), m) and torch.equal(bits, bits2)
    _, m3, bits3 = K().stitch_blend(ld, shape, daxes, wins, epilogue=1, want_mask=True, want_bits=True, thr=thr)  # with the prob map
    assert np.array_equal(bits3.cpu().numpy(), pack_bits(m3.cpu().numpy().astype(bool)))


def test_blend_band_test_on_values_at_the_threshold(dev):
    """Logits planted exactly at, one ulp below and one ulp above logit(thr) — where the division-free band test must
    hand over to the exact quotient and numpy's fp32 sigmoid."""
    B, ov, H, W = 64, 16, 96, 128
    axes = [G.identity_axis(1), G.fibsem_axis(H, B, ov), G.fibsem_axis(W, B, ov)]
    T = axes[1].T * axes[2].T
    for thr in (0.5, 0.3, 0.9):
        L = np.float32(np.log(thr / (1 - thr)))
        vals = np.array([L, np.nextafter(L, np.float32(-9)), np.nextafter(L, np.float32(9)), L + np.float32(5e-5),
                         L - np.float32(5e-5), 0.0, -0.0], np.float32)
        logits = np.tile(vals, (T * B * B + len(vals) - 1) // len(vals))[:T * B * B].reshape(T, 1, B, B).copy()
        w1 = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
        wins = (torch.ones(1, dtype=torch.float64, device=dev), w1, w1)
        want = O.fibsem_blend(logits[:, 0], (H, W), B, ov)
        table = K().blend_window(wins)
        assert np.array_equal(table.cpu().numpy(), O.fibsem_window(B))   # the reference's weight_win, bit for bit
        for tab in (None, table):
            _, only = K().stitch_blend(torch.from_numpy(logits).to(dev), (1, 1, H, W), dev_axes(axes, dev), wins, epilogue=1,
                                       want_out=False, want_mask=True, thr=thr, window=tab)
            assert np.array_equal(only.cpu().numpy()[0, 0], O.binarize(O.sigmoid(want), thr)), thr
        # no sigmoid: blended values planted on the threshold itself (every tile holds the same value, so the blend
        # returns it up to rounding of the quotient)
        t32 = np.float32(thr)
        vals2 = np.array([t32, np.nextafter(t32, np.float32(-9)), np.nextafter(t32, np.float32(9)), t32 * np.float32(1.000001),
                          t32 * np.float32(0.999999), 0.0, 1.0], np.float32)
        img = np.tile(vals2, (H * W + len(vals2) - 1) // len(vals2))[:H * W].reshape(H, W)
        plain = O.fibsem_tiles(img, B, ov)[:, None]
        want2 = O.fibsem_blend(plain[:, 0], (H, W), B, ov)
        for tab in (None, table):
            _, only2 = K().stitch_blend(torch.from_numpy(plain).to(dev), (1, 1, H, W), dev_axes(axes, dev), wins, epilogue=0,
                                        want_out=False, want_mask=True, thr=thr, window=tab)
            assert np.array_equal(only2.cpu().numpy()[0, 0], O.binarize(want2, thr)), thr


def test_kernels_write_nothing_outside_their_outputs(dev, rng):
    """Guard bands around the output buffers (compute-sanitizer is not available on the GPU pool): shapes whose rows are
    not a whole number of row groups / stages, several planes, both blend kernels and both gather kernels."""
    from bioengine_worker_b200 import _lib
    kern = K()
    guard = 4096
    stream = kern._stream()
    for spatial, B, ov, C in (((52, 96), 32, 8, 2), ((40, 52, 96), 32, 8, 2), ((50, 90), 32, 8, 1), ((77, 1024 + 64), 256, 64, 1)):
        nd = len(spatial)
        axes = [G.fibsem_axis(s, B, ov) for s in spatial]
        if nd == 2:
            axes = [G.identity_axis(1)] + axes
        az, ay, ax = dev_axes(axes, dev)
        T = az.ax.T * ay.ax.T * ax.ax.T
        Z, Y, X = az.ax.S, ay.ax.S, ax.ax.S
        n_px = C * Z * Y * X
        img = torch.from_numpy(rng.standard_normal((1, C, Z, Y, X)).astype(np.float32)).to(dev)
        # gather into a guarded tile buffer
        n_t = T * C * az.ax.B * ay.ax.B * ax.ax.B
        tbuf = torch.full((n_t + 2 * guard,), 7.5, dtype=torch.float32, device=dev)
        _lib.call("be_tile_gather_f32", img.data_ptr(), tbuf[guard:].data_ptr(), 1, C, Z, Y, X, az.ax.T, ay.ax.T, ax.ax.T,
                  az.ax.B, ay.ax.B, ax.ax.B, az.src.data_ptr(), ay.src.data_ptr(), ax.src.data_ptr(), 0, 0, 0, 0.0, 0.0, stream)
        assert bool((tbuf[:guard] == 7.5).all()) and bool((tbuf[guard + n_t:] == 7.5).all()), (spatial, "gather")
        assert not bool((tbuf[guard:guard + n_t] == 7.5).any())
        # blend into guarded outputs, every epilogue combination
        w1 = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
        one = torch.ones(1, dtype=torch.float64, device=dev)
        wz = w1 if nd == 3 else one
        wtab = kern.blend_window((wz, w1, w1)) if nd == 2 else None
        for want_out, epi, use_table in ((True, 0, False), (True, 1, True), (False, 1, False), (False, 1, True), (False, 0, True)):
            obuf = torch.full((n_px + 2 * guard,), -3.25, dtype=torch.float32, device=dev)
            mbuf = torch.full((n_px + 2 * guard,), 9, dtype=torch.uint8, device=dev)
            n_words = C * Z * Y * ((X + 63) // 64)
            bbuf = torch.full((n_words + 2 * guard,), 0x5555, dtype=torch.int64, device=dev)
            _lib.call("be_stitch_blend_f32", tbuf[guard:].data_ptr(), obuf[guard:].data_ptr() if want_out else 0,
                      mbuf[guard:].data_ptr(), bbuf[guard:].data_ptr(), 1, C, Z, Y, X, az.ax.T, ay.ax.T, ax.ax.T, az.ax.B, ay.ax.B, ax.ax.B,
                      az.ax.stride, ay.ax.stride, ax.ax.stride, wz.data_ptr(), w1.data_ptr(), w1.data_ptr(),
                      wtab.data_ptr() if (nd == 2 and use_table) else 0, epi, 0.5, stream)
            assert bool((mbuf[:guard] == 9).all()) and bool((mbuf[guard + n_px:] == 9).all()), (spatial, epi, "mask")
            assert bool((mbuf[guard:guard + n_px] <= 1).all())
            assert bool((bbuf[:guard] == 0x5555).all()) and bool((bbuf[guard + n_words:] == 0x5555).all()), (spatial, epi, "bits")
            from tests.test_gpu_ccl import pack_bits
            want_bits = pack_bits(mbuf[guard:guard + n_px].cpu().numpy().reshape(C * Z * Y, X).astype(bool))
            assert np.array_equal(bbuf[guard:guard + n_words].cpu().numpy().reshape(C * Z * Y, -1), want_bits), (spatial, epi)
            assert bool((obuf[:guard] == -3.25).all()) and bool((obuf[guard + n_px:] == -3.25).all()), (spatial, epi, "out")
            if want_out:
                assert not bool((obuf[guard:guard + n_px] == -3.25).any())


@pytest.mark.parametrize("shape,block,halo", [((2, 1, 100, 200), (48, 64), (8, 8)), ((1, 2, 70, 81), (32, 32), (4, 4)),
                                              ((1, 1, 20, 33, 130), (8, 16, 64), (1, 2, 4))])
def test_crop_stitch_packed_mask(shape, block, halo, dev, rng):
    """The bit-packed mask of the crop stitcher (rows that end inside a word, several planes, 3-D) against the byte mask."""
    from tests.test_gpu_ccl import pack_bits

    nd = len(block)
    axes = [G.bioimageio_axis(s, b, h) for s, b, h in zip(shape[2:], block, halo)]
    if nd == 2:
        axes = [G.identity_axis(1)] + axes
    T = int(np.prod([a.T for a in axes]))
    logits = (rng.standard_normal((shape[0] * T, shape[1]) + tuple(block)) * 2).astype(np.float32)
    ld = torch.from_numpy(logits).to(dev)
    want = O.binarize(O.sigmoid(O.from_blocks(logits, shape, block, halo)), 0.5)
    out, mask, bits = K().stitch_crop(ld, shape, dev_axes(axes, dev), 1, 0.5, want_out=True, want_mask=True, want_bits=True)
    assert np.array_equal(mask.cpu().numpy(), want)
    assert np.array_equal(bits.cpu().numpy(), pack_bits(want.astype(bool)))
    _, none, only = K().stitch_crop(ld, shape, dev_axes(axes, dev), 1, 0.5, want_out=False, want_bits=True)
    assert none is None and torch.equal(only, bits)
    labels, n = K().ccl_bits(only.reshape(-1), (shape[0] * shape[1],) + tuple(shape[2:]), None)
    want_l = np.stack([O.label(m) for m in want.reshape((-1,) + tuple(shape[2:]))])
    assert np.array_equal(labels.cpu().numpy(), want_l)
