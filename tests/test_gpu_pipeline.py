"""Whole path on the device against the whole-path oracle.  Downstream integer results are compared on identical
logits (the GPU forward's output is handed to the CPU oracle), as SURVEY.md Appendix C prescribes."""
import numpy as np
import pytest
import torch

from oracle import oracle as O
from oracle import paths as P

pytestmark = pytest.mark.gpu


def make(shape, spec_kwargs, dev, seed=1, arch="UNet2d", **model_kwargs):
    from bioengine_worker_b200 import models
    from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline

    models.strict_fp32()
    pipe = TiledPipeline(shape, PipelineSpec(**spec_kwargs), device=dev)
    model = models.build_model(arch, seed=seed, **model_kwargs).to(dev)
    return pipe, model


def test_c2_shape_path_fibsem_blend(dev):
    H, W, B, ov = 700, 900, 256, 64
    img = O.synthetic_image((1, 1, H, W), 120, seed=2)
    pipe, model = make((1, 1, H, W), dict(block=(B, B), halo=(32, 32), convention="fibsem", overlap=(ov, ov),
                                          preprocess="scale_range", return_prob=True, connectivity=2), dev)
    x = torch.from_numpy(img).to(dev)
    tiles = pipe.preprocess_and_tile(x)
    assert np.array_equal(tiles.cpu().numpy(), P.c2_pre(img, B, ov))
    from bioengine_worker_b200 import models

    models.calibrate_head(model, tiles[::3].contiguous())
    logits = pipe.forward(model, tiles)
    res = pipe.stitch_and_label(logits)
    prob, mask, labels = P.c2_post(logits.cpu().numpy(), (H, W), B, ov, 2)
    np.testing.assert_allclose(res["prob"].cpu().numpy()[0, 0], prob, rtol=1e-5)
    assert np.array_equal(res["mask"].cpu().numpy()[0, 0], mask)
    assert np.array_equal(res["labels"].cpu().numpy()[0, 0], labels)
    assert res["n_labels"].item() == labels.max() > 0
    # the public one-call API gives the same answer from a host array
    res2 = pipe.run(img, model)
    assert torch.equal(res2["labels"], res["labels"])


def test_bioimageio_crop_path_batch(dev):
    shape, B, halo = (2, 1, 300, 420), 128, 16
    img = O.synthetic_image(shape, 90, seed=7)
    pipe, model = make(shape, dict(block=(B, B), halo=(halo, halo), convention="bioimageio", preprocess="scale_range",
                                   return_prob=True), dev)
    x = torch.from_numpy(img).to(dev)
    tiles = pipe.preprocess_and_tile(x)
    assert np.array_equal(tiles.cpu().numpy(), P.crop_pre(img, (B, B), (halo, halo)))
    from bioengine_worker_b200 import models

    models.calibrate_head(model, tiles[::5].contiguous())
    logits = pipe.forward(model, tiles)
    res = pipe.stitch_and_label(logits)
    assert res["n_labels"].min().item() > 0
    prob, mask, labels = P.crop_post(logits.cpu().numpy(), shape, (B, B), (halo, halo))
    assert np.array_equal(res["prob"].cpu().numpy(), prob)
    assert np.array_equal(res["mask"].cpu().numpy(), mask)
    assert np.array_equal(res["labels"].cpu().numpy()[:, 0], labels)


def test_c1_zmuv_binarize_untiled(dev):
    """BASELINE config 1 shape: 512^2, zero_mean_unit_variance, binarize, CC — one block covering the image."""
    shape = (1, 1, 512, 512)
    img = O.synthetic_image(shape, 60, seed=1, microscope=False)
    pipe, model = make(shape, dict(block=(512, 512), halo=(0, 0), convention="bioimageio",
                                   preprocess="zero_mean_unit_variance", sigmoid=False, return_prob=True), dev)
    x = torch.from_numpy(img).to(dev)
    tiles = pipe.preprocess_and_tile(x)
    np.testing.assert_allclose(tiles.cpu().numpy(), O.zero_mean_unit_variance(img, axes=(1, 2, 3)), rtol=1e-5, atol=1e-5)
    logits = pipe.forward(model, tiles)
    res = pipe.stitch_and_label(logits)
    raw = logits.cpu().numpy()
    mask = O.binarize(raw, 0.5)
    assert np.array_equal(res["mask"].cpu().numpy(), mask)
    assert np.array_equal(res["labels"].cpu().numpy()[0, 0], O.label(mask[0, 0]))


def test_c3_shape_path_3d(dev):
    """BASELINE config 3 in miniature: fibsem p1/p99 + clip, 3-D tiles with overlap, blend, > 0.5, 26-connected labels."""
    D, B, ov = (80, 96, 112), 32, 8
    vol = O.synthetic_image((1, 1) + D, 60, seed=3)
    pipe, model = make((1, 1) + D, dict(block=(B,) * 3, halo=(4,) * 3, convention="fibsem", overlap=(ov,) * 3,
                                        preprocess="fibsem", sigmoid=True, return_prob=True, forward_batch=16), dev,
                       arch="UNet3d", depth=3, initial_features=8)
    x = torch.from_numpy(vol).to(dev)
    tiles = pipe.preprocess_and_tile(x)
    want_tiles = O.fibsem_tiles(O.fibsem_normalize(vol[0, 0]), B, ov)
    assert np.array_equal(tiles.cpu().numpy()[:, 0], want_tiles)
    from bioengine_worker_b200 import models

    models.calibrate_head(model, tiles[::7].contiguous())
    logits = pipe.forward(model, tiles)
    res = pipe.stitch_and_label(logits)
    assert res["n_labels"].item() > 0
    prob, mask, labels = P.c2_post(logits.cpu().numpy(), D, B, ov, 3)
    assert np.array_equal(res["mask"].cpu().numpy()[0, 0], mask)
    assert np.array_equal(res["labels"].cpu().numpy()[0, 0], labels)


def test_run_stream_equals_run(dev):
    """Throughput mode (three streams, double buffering) returns exactly what the one-image call returns."""
    shape = (1, 1, 200, 260)
    imgs = [O.synthetic_image(shape, 40, seed=20 + i) for i in range(5)]
    pipe, model = make(shape, dict(block=(64, 64), halo=(8, 8), convention="fibsem", overlap=(16, 16),
                                   preprocess="scale_range", connectivity=2), dev)
    from bioengine_worker_b200 import models

    models.calibrate_head(model, pipe.preprocess_and_tile(torch.from_numpy(imgs[0]).to(dev))[::3].contiguous())
    want = [pipe.run(im, model) for im in imgs]
    want = [(w["labels"].cpu().numpy().copy(), int(w["n_labels"].item())) for w in want]
    host_in = [torch.from_numpy(im).pin_memory() for im in imgs]
    host_l = [torch.empty((1, 1) + shape[2:], dtype=torch.int32).pin_memory() for _ in imgs]
    host_n = [torch.empty((1,), dtype=torch.int32).pin_memory() for _ in imgs]
    pipe.run_stream(host_in, model, host_l, host_n).synchronize()
    torch.cuda.synchronize()
    for (wl, wn), gl, gn in zip(want, host_l, host_n):
        assert wn == int(gn.item()) > 0
        assert np.array_equal(wl, gl.numpy())


def test_c3_full_size_properties(dev):
    """BASELINE config 3 at full size (512^3, 128^3 tiles, 16-voxel halo → overlap 32, 216 tiles): size-independent
    properties instead of an oracle run — blending tiles cut from one volume returns that volume, the crop stitch of the
    bioimageio blocks is the identity, and 26-connected labelling is idempotent and agrees with the oracle on a slab."""
    from bioengine_worker_b200 import geometry as G
    from bioengine_worker_b200 import kernels as K

    S, B, halo = 512, 128, 16
    g = torch.Generator(device=dev).manual_seed(3)
    x = torch.rand((1, 1, S, S, S), device=dev, generator=g)
    fax = [K.DeviceAxis(G.fibsem_axis(S, B, 2 * halo), dev) for _ in range(3)]
    tiles = K.tile_gather(x, fax)
    assert tuple(tiles.shape) == (216, 1, B, B, B)
    w = torch.from_numpy(G.gauss_window_1d(B)).to(dev)
    out, mask = K.stitch_blend(tiles, (1, 1, S, S, S), fax, (w, w, w), want_mask=True, thr=0.7)
    torch.testing.assert_close(out, x, rtol=1e-5, atol=1e-6)
    del tiles
    bax = [K.DeviceAxis(G.bioimageio_axis(S, B, halo), dev) for _ in range(3)]
    btiles = K.tile_gather(x, bax)
    back, _ = K.stitch_crop(btiles, (1, 1, S, S, S), bax)
    assert torch.equal(back, x)
    del btiles, back, out
    labels, n = K.ccl(mask[0], 3)
    assert int(n.item()) > 0 and int(labels.max().item()) == int(n.item())
    again, n2 = K.ccl((labels > 0).to(torch.uint8), 3)
    assert torch.equal(labels, again) and int(n2.item()) == int(n.item())
    # the oracle on a slab: labels restricted to z < 40 must induce the same partition as labelling the slab alone
    slab = mask[0, 0, :40].cpu().numpy()
    got = K.ccl(mask[0, :, :40].contiguous(), 3)[0].cpu().numpy()[0]
    assert np.array_equal(got, O.label(slab, 3))


def test_c3_full_size_labels_equal_scipy(dev):
    """BASELINE config 3 at full size: the 512^3 mask the device pipeline produces, labelled on the device (26- and
    6-connected, from the packed mask) and by scipy on the identical mask: equal, numbering included."""
    import scipy.ndimage as ndi

    from bioengine_worker_b200 import kernels as K

    S = 512
    g = torch.Generator(device=dev).manual_seed(33)
    coarse = torch.randn((1, 1, 20, 20, 20), device=dev, generator=g)
    x = torch.nn.functional.interpolate(coarse, size=(S, S, S), mode="trilinear", align_corners=False)[0, 0]
    x = x + 0.35 * torch.randn(x.shape, device=dev, generator=g)     # blobs with ragged, speckled borders
    mask = (x > 0.9).to(torch.uint8).contiguous()
    del x
    m_np = mask.cpu().numpy()
    from tests.test_gpu_ccl import pack_bits

    bits = torch.from_numpy(pack_bits(m_np.astype(bool))).to(dev)
    for conn in (3, 1):
        want, n_want = ndi.label(m_np, structure=ndi.generate_binary_structure(3, conn))
        got, n = K.ccl_bits(bits.reshape(-1), (1, S, S, S), conn)
        assert int(n.item()) == n_want
        assert np.array_equal(got.cpu().numpy()[0], want.astype(np.int32))
        del want
    got8, n8 = K.ccl(mask[None], 3)
    assert torch.equal(got8, K.ccl_bits(bits.reshape(-1), (1, S, S, S), 3)[0])


def test_c5_frame_full_size_labels_and_stats_equal_oracle(dev):
    """BASELINE config 5 at full size: one 2048^2 uint16 frame through upload → ensure_dtype → scale_range → 512-px blocks
    with a 32-px halo → forward → sigmoid → binarize → 8-connected labels → per-label statistics, against the oracle on the
    same logits (labels bit-exact, areas and centroid sums exact integers)."""
    from bioengine_worker_b200 import kernels as K
    from bioengine_worker_b200 import models
    from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline

    H = W = 2048
    models.strict_fp32()
    frame = np.clip(O.synthetic_image((1, 1, H, W), 800, seed=5), 0, 65535).astype(np.uint16)
    spec = PipelineSpec(block=(512, 512), halo=(32, 32), convention="bioimageio", preprocess="scale_range", sigmoid=True,
                        connectivity=2, forward_batch=25)
    pipe = TiledPipeline((1, 1, H, W), spec, device=dev)
    net = models.build_model("UNet2d", seed=5, depth=2, initial_features=8).to(dev)
    tiles = pipe.preprocess_and_tile(pipe.upload(frame))
    models.calibrate_head(net, tiles[::5].contiguous())
    want_tiles = P.crop_pre(frame.astype(np.float32), (512, 512), (32, 32))
    assert np.array_equal(tiles.cpu().numpy(), want_tiles)
    logits = pipe.forward(net, tiles)
    res = pipe.stitch_and_label(logits)
    _, mask, labels = P.crop_post(logits.cpu().numpy(), (1, 1, H, W), (512, 512), (32, 32), 2)
    assert 0.02 < mask.mean() < 0.98
    assert np.array_equal(res["mask"].cpu().numpy(), mask)
    assert np.array_equal(res["labels"].cpu().numpy()[:, 0], labels)
    n = int(res["n_labels"].item())
    assert n == labels.max()
    sums, bbox, status = K.label_stats(res["labels"][0], n + 1)
    assert int(status.item()) == 0
    s = sums.cpu().numpy()[0]
    lab = labels[0]
    ys, xs = np.nonzero(lab)
    l = lab[ys, xs]
    assert np.array_equal(s[:, 0], np.bincount(l, minlength=n + 1))                      # areas
    assert np.array_equal(s[:, 2], np.bincount(l, weights=ys, minlength=n + 1).astype(np.int64))   # sum of y
    assert np.array_equal(s[:, 3], np.bincount(l, weights=xs, minlength=n + 1).astype(np.int64))   # sum of x


def test_c2_path_on_an_image_four_times_the_benchmark_size(dev):
    """8192 x 8192 (67 M pixels, 1849 tiles of 256 px, 485 MB of tiles): quantiles, tiles, the mask-only blend, the packed
    mask and the labels against the oracle on identical stand-in logits — index arithmetic beyond 2^26 elements per image
    and 2^28 tile elements, ragged last tiles (8192 = 42 * 192 + 128)."""
    from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline

    S, B, ov = 8192, 256, 64
    img = O.synthetic_image((1, 1, S, S), 3000, seed=8)
    pipe = TiledPipeline((1, 1, S, S), PipelineSpec(block=(B, B), halo=(32, 32), convention="fibsem", overlap=(ov, ov),
                                                    preprocess="scale_range", connectivity=2), device=dev)
    tiles = pipe.preprocess_and_tile(torch.from_numpy(img).to(dev))
    want_tiles = P.c2_pre(img, B, ov)
    assert tiles.shape == want_tiles.shape == (43 * 43, 1, B, B)
    assert np.array_equal(tiles.cpu().numpy(), want_tiles)
    logits = ((tiles - 0.15) * 9.0).contiguous()       # stand-in for the forward: blobs above the threshold
    res = pipe.stitch_and_label(logits)
    _, mask, labels = P.c2_post(logits.cpu().numpy(), (S, S), B, ov, 2)
    assert 0.01 < mask.mean() < 0.6
    assert np.array_equal(res["mask"].cpu().numpy()[0, 0], mask)
    assert np.array_equal(res["labels"].cpu().numpy()[0, 0], labels)
    assert int(res["n_labels"].item()) == labels.max() > 1000
