"""The reference-facing classes: RuntimeDeployment.predict (RDF-driven) and MitoAnalysisDeployment.analyze, driven the
way the reference's own local-dev mains drive them (`RuntimeDeployment.func_or_class()`, asyncio.run), against the
whole-path oracle on identical logits."""
import asyncio

import numpy as np
import pytest
import torch
import yaml

from oracle import oracle as O
from oracle import paths as P

pytestmark = pytest.mark.gpu


def write_model(tmp_path, pre, post, halo=16, out_dtype="float32", seed=3, arch="UNet2d", arch_kwargs=None, calibrate_on=None):
    from bioengine_worker_b200 import models

    arch_kwargs = arch_kwargs or {"depth": 2, "initial_features": 8}
    net = models.build_model(arch, seed=seed, **arch_kwargs)
    if calibrate_on is not None:
        models.calibrate_head(net, torch.from_numpy(calibrate_on))
    torch.save(net.state_dict(), tmp_path / "weights.pt")
    rdf = {
        "format_version": "0.5.3", "type": "model", "name": "test-unet",
        "inputs": [{"id": "raw", "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["c0"]},
                                           {"type": "space", "id": "y", "size": {"min": 64, "step": 32}},
                                           {"type": "space", "id": "x", "size": {"min": 64, "step": 32}}],
                    "preprocessing": pre}],
        "outputs": [{"id": "mask", "data": {"type": out_dtype},
                     "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["p"]},
                              {"type": "space", "id": "y", "halo": halo}, {"type": "space", "id": "x", "halo": halo}],
                     "postprocessing": post}],
        "weights": {"pytorch_state_dict": {"source": "weights.pt",
                                           "architecture": {"callable": arch, "kwargs": arch_kwargs}}},
    }
    (tmp_path / "rdf.yaml").write_text(yaml.safe_dump(rdf))
    return str(tmp_path / "rdf.yaml"), net


def test_predict_with_blocking_scale_range_sigmoid_binarize(tmp_path, dev):
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    img = O.synthetic_image((300, 420), 60, seed=9)
    norm = O.scale_range(img[None, None], 1.0, 99.8, axes=(1, 2, 3))
    pre = [{"id": "ensure_dtype", "kwargs": {"dtype": "float32"}},
           {"id": "scale_range", "kwargs": {"min_percentile": 1.0, "max_percentile": 99.8, "axes": ["y", "x"]}}]
    post = [{"id": "sigmoid"}, {"id": "binarize", "kwargs": {"threshold": 0.5}}]
    rdf_path, net = write_model(tmp_path, pre, post, halo=16, out_dtype="uint8", calibrate_on=norm[:, :, :128, :128])
    rt = RuntimeDeployment.func_or_class()
    # default_blocksize_parameter n=2 → block = 64 + 2*32 = 128 (halo 16 inside)
    out = asyncio.run(rt.predict(rdf_path, img, default_blocksize_parameter=2))
    assert list(out) == ["mask"] and out["mask"].shape == (1, 1, 300, 420) and out["mask"].dtype == np.uint8
    tiles = P.crop_pre(img[None, None], (128, 128), (16, 16))
    with torch.no_grad():
        logits = net.to(dev)(torch.from_numpy(tiles).to(dev)).cpu().numpy()
    _, mask, _ = P.crop_post(logits, (1, 1, 300, 420), (128, 128), (16, 16))
    assert 0.02 < mask.mean() < 0.98
    assert np.array_equal(out["mask"], mask)
    # second call hits the pipeline cache and agrees
    assert np.array_equal(asyncio.run(rt.predict(rdf_path, {"raw": img[None, None]}, default_blocksize_parameter=2))["mask"], mask)


def test_predict_without_blocking_zmuv_float_output(tmp_path, dev):
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    img = O.synthetic_image((128, 160), 12, seed=4, microscope=False)
    pre = [{"id": "zero_mean_unit_variance", "kwargs": {"axes": ["y", "x"]}}]
    post = [{"id": "sigmoid"}]
    rdf_path, net = write_model(tmp_path, pre, post)
    rt = RuntimeDeployment.func_or_class()
    out = asyncio.run(rt.predict(rdf_path, img))
    xn = O.zero_mean_unit_variance(img[None, None], axes=(1, 2, 3))
    with torch.no_grad():
        want = O.sigmoid(net.to(dev)(torch.from_numpy(xn).to(dev)).cpu().numpy())
    assert out["mask"].dtype == np.float32 and out["mask"].shape == want.shape
    np.testing.assert_allclose(out["mask"], want, rtol=1e-4, atol=1e-6)  # forward sees inputs that differ by <= 1e-5 rel


def test_predict_errors_follow_the_reference(tmp_path):
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    rt = RuntimeDeployment.func_or_class()
    with pytest.raises(FileNotFoundError):
        asyncio.run(rt.predict(str(tmp_path / "missing.yaml"), np.zeros((64, 64), np.float32)))
    rdf_path, _ = write_model(tmp_path, [], [])
    with pytest.raises(ValueError):
        asyncio.run(rt.predict(rdf_path, np.zeros((64, 64), np.float32), device="cpu"))
    with pytest.raises(ValueError):
        asyncio.run(rt.predict(rdf_path, np.zeros((1, 1, 1, 64, 64), np.float32)))
    report = asyncio.run(rt.test(rdf_path))
    assert report["status"] == "passed" and report["details"][0]["name"] == "mask"


def test_analyze_matches_fibsem_oracle(dev):
    from bioengine_worker_b200 import models
    from bioengine_worker_b200.analysis_deployment import MitoAnalysisDeployment

    models.strict_fp32()
    H, W, tile, ov = 300, 410, 128, 32
    rng = np.random.default_rng(0)
    img = (O.gaussian_blobs((H, W), 40, seed=6) * 200 + 20 + 5 * rng.standard_normal((H, W))).clip(0, 255).astype(np.float32)
    net = models.build_model("UNet2d", seed=2, depth=2, initial_features=8, final_sigmoid=True).to(dev)
    norm = O.fibsem_normalize(img)
    tiles = O.fibsem_tiles(norm, tile, ov)[:, None]
    # centre the pre-sigmoid head output so the probabilities straddle 0.5
    net.final_sigmoid = False
    models.calibrate_head(net, torch.from_numpy(tiles[::3]).to(dev))
    net.final_sigmoid = True
    dep = MitoAnalysisDeployment(model=net, device=dev)
    res = asyncio.run(dep.analyze(img.tolist(), pixel_size_nm=5.0, tile_size=tile, overlap=ov, cleanup=False))
    with torch.no_grad():
        probs = net(torch.from_numpy(tiles).to(dev)).cpu().numpy()[:, 0]
    prob = O.fibsem_blend(probs, (H, W), tile, ov)
    labels = O.label(prob > 0.5)
    assert res["n_mitochondria"] == labels.max() > 0
    assert np.array_equal(np.asarray(res["labels"], dtype=np.int32), labels)
    want = O.regionprops_table(labels, 5.0)
    assert res["properties"]["label"] == want["label"]
    for k in ("area_um2", "centroid_y", "centroid_x"):
        np.testing.assert_allclose(res["properties"][k], want[k], rtol=1e-12)
    # aspect_ratio = major / (minor + 1e-6) is ill-conditioned for one-pixel-thin objects (minor is cancellation noise in
    # any floating-point evaluation); compare tightly where the minor axis is resolved, loosely elsewhere
    ar_got, ar_want = np.asarray(res["properties"]["aspect_ratio"]), np.asarray(want["aspect_ratio"])
    thin = ar_want > 1e3
    np.testing.assert_allclose(ar_got[~thin], ar_want[~thin], rtol=1e-6)
    assert (ar_got[thin] > 1e3).all()
    np.testing.assert_allclose(res["properties"]["eccentricity"], want["eccentricity"], rtol=1e-6, atol=1e-6)
    assert res["image_shape"] == [H, W] and set(res) >= {"labels", "properties", "n_mitochondria", "pixel_size_nm", "model"}
    # the default: the reference's `_prob_to_instances` (size filter, closing, distance peaks, watershed)
    res3 = asyncio.run(dep.analyze(img.tolist(), tile_size=tile, overlap=ov, as_numpy=True))   # cleanup=True is the default
    want3 = O.prob_to_instances(prob, ties="fifo")
    assert np.array_equal(res3["labels"], want3) and res3["n_mitochondria"] == want3.max()
    assert res3["properties"]["label"] == O.regionprops_table(want3, 5.0)["label"]
    # small image → single untiled call (ref: analysis_deployment.py:247-248)
    small = img[:100, :120]
    res2 = asyncio.run(dep.analyze(small.tolist(), tile_size=128, overlap=32, as_numpy=True, cleanup=False))
    with torch.no_grad():
        p2 = net(torch.from_numpy(O.fibsem_normalize(small)[None, None]).to(dev)).cpu().numpy()[0, 0]
    assert np.array_equal(res2["labels"], O.label(p2 > 0.5))
    with pytest.raises(ValueError):
        asyncio.run(dep.analyze(np.zeros((2, 3, 4)).tolist()))


def test_entry_deployment_infer_mirrors_the_reference_call(tmp_path, dev):
    from bioengine_worker_b200.entry_deployment import EntryDeployment
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    pkg = tmp_path / "models" / "tiny-unet"
    pkg.mkdir(parents=True)
    rdf_path, _ = write_model(pkg, [{"id": "zero_mean_unit_variance", "kwargs": {"axes": ["y", "x"]}}], [{"id": "sigmoid"}])
    entry = EntryDeployment(RuntimeDeployment.func_or_class(), models_dir=str(tmp_path / "models"))
    img = O.synthetic_image((96, 128), 8, seed=1, microscope=False)
    direct = asyncio.run(entry.infer("tiny-unet", img))
    assert list(direct) == ["mask"] and direct["mask"].shape == (1, 1, 96, 128)
    np.save(tmp_path / "in.npy", img)
    via_file = asyncio.run(entry.infer("tiny-unet", str(tmp_path / "in.npy"), return_download_url=True))
    assert via_file["mask"].startswith("file://")
    assert np.array_equal(np.load(via_file["mask"][7:]), direct["mask"])
    with pytest.raises(FileNotFoundError):
        asyncio.run(entry.infer("no-such-model", img))
    with pytest.raises(ValueError):      # the extension is checked before anything is fetched
        asyncio.run(entry.infer("tiny-unet", "https://example.org/x.bmp"))
    with pytest.raises(FileNotFoundError):   # nothing listens there: the source cannot be downloaded
        asyncio.run(entry.infer("tiny-unet", "http://127.0.0.1:9/x.npy"))


def test_npy_wire_path_round_trip(tmp_path, dev):
    from bioengine_worker_b200 import io as bio

    rng = np.random.default_rng(0)
    staging = bio.PinnedStaging()
    for arr in (rng.random((3, 40, 50)).astype(np.float32), rng.integers(0, 65535, (64, 64)).astype(np.uint16),
                rng.integers(0, 255, (5, 7)).astype(np.uint8)):
        p = tmp_path / "a.npy"
        np.save(p, arr)
        t = bio.load_npy_to_device(str(p), dev, staging)
        assert t.dtype == torch.float32 and np.array_equal(t.cpu().numpy(), arr.astype(np.float32))
    lab = torch.arange(24, dtype=torch.int32, device=dev).reshape(4, 6)
    bio.save_npy_from_device(str(tmp_path / "out.npy"), lab, staging)
    assert np.array_equal(np.load(tmp_path / "out.npy"), lab.cpu().numpy())


def test_scale_mean_variance_op_and_in_predict(tmp_path, dev):
    """bioimageio `scale_mean_variance`: the output is brought to the mean / std of a reference tensor
    — in zoo models the raw input, whose statistics are taken before pre-processing touches it."""
    from bioengine_worker_b200 import kernels as K
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    rng = np.random.default_rng(11)
    x = (rng.standard_normal((2, 3, 40, 50)) * 3 + 1).astype(np.float32)
    ref = (rng.standard_normal((2, 3, 40, 50)) * 7 - 4).astype(np.float32)
    for axes, n_seg in (((1, 2, 3), 2), ((2, 3), 6)):
        ref_norm = K.meanvar(torch.from_numpy(ref).to(dev), n_seg, 1e-6)[1]
        got = K.scale_mean_variance(torch.from_numpy(x).to(dev), ref_norm, n_seg, 1e-6).cpu().numpy()
        np.testing.assert_allclose(got, O.scale_mean_variance(x, ref, axes=axes), rtol=1e-5, atol=2e-5)

    img = O.synthetic_image((128, 160), 12, seed=6, microscope=False) * 50 + 200   # a raw image far from unit scale
    pre = [{"id": "zero_mean_unit_variance", "kwargs": {"axes": ["y", "x"]}}]
    post = [{"id": "scale_mean_variance", "kwargs": {"reference_tensor": "raw", "axes": ["y", "x"]}}]
    rdf_path, net = write_model(tmp_path, pre, post)
    out = asyncio.run(RuntimeDeployment.func_or_class().predict(rdf_path, img))["mask"]
    xn = O.zero_mean_unit_variance(img[None, None], axes=(1, 2, 3))
    with torch.no_grad():
        y = net.to(dev)(torch.from_numpy(xn).to(dev)).cpu().numpy()
    want = O.scale_mean_variance(y, img[None, None], axes=(1, 2, 3))
    assert out.shape == want.shape and out.dtype == np.float32
    np.testing.assert_allclose(out, want, rtol=1e-4, atol=1e-3)   # forward sees inputs that differ by <= 1e-5 rel
    assert abs(float(out.mean()) - float(img.mean())) < 1e-2 * abs(float(img.mean()))
    assert abs(float(out.std()) - float(img.std())) < 1e-2 * float(img.std())


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_pipeline_on_second_gpu_while_first_is_current():
    """The library launches on the current device: the front end must make the tensors' device current (and use that
    device's stream) even when the caller's current device is another one."""
    from bioengine_worker_b200 import kernels as K
    from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline

    torch.cuda.set_device(0)
    d1 = torch.device("cuda:1")
    img = O.synthetic_image((1, 1, 256, 320), 30, seed=1)
    spec = PipelineSpec(block=(64, 64), halo=(8, 8), convention="bioimageio", preprocess="scale_range")
    pipe = TiledPipeline(img.shape, spec, device=d1)
    res = pipe.run(img, lambda t: (t * 6.0 - 2.0).contiguous())
    torch.cuda.synchronize(d1)
    assert torch.cuda.current_device() == 0
    want_tiles = O.blocks_tensor(O.scale_range(img, 1.0, 99.8), (64, 64), (8, 8))
    want_mask = O.binarize(O.sigmoid(O.from_blocks(want_tiles * np.float32(6.0) - np.float32(2.0), img.shape, (64, 64), (8, 8))), 0.5)
    assert np.array_equal(res["mask"].cpu().numpy(), want_mask)
    assert np.array_equal(res["labels"].cpu().numpy()[0, 0], O.label(want_mask[0, 0], 2))
    with pytest.raises(ValueError):
        K.normalize(torch.zeros(8, device="cuda:0"), torch.zeros((1, 2), dtype=torch.float64, device=d1))


def test_predict_fixed_zmuv_clip_scale_linear_and_integer_input(tmp_path, dev):
    """`fixed_zero_mean_unit_variance` (scalar and per-channel), `clip`, `scale_linear` in one pre-processing chain, on a
    uint8 input that is uploaded as uint8 and converted on the device (bioimageio's leading ensure_dtype)."""
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    rng = np.random.default_rng(12)
    img8 = rng.integers(0, 256, (1, 1, 96, 128)).astype(np.uint8)
    pre = [{"id": "ensure_dtype", "kwargs": {"dtype": "float32"}},
           {"id": "fixed_zero_mean_unit_variance", "kwargs": {"mean": 103.5, "std": 57.25}},
           {"id": "clip", "kwargs": {"min": -1.5, "max": 1.25}},
           {"id": "scale_linear", "kwargs": {"gain": 0.5, "offset": 0.125}}]
    rdf_path, net = write_model(tmp_path, pre, [{"id": "sigmoid"}])
    rt = RuntimeDeployment.func_or_class()
    out = asyncio.run(rt.predict(rdf_path, img8))
    x = img8.astype(np.float32)
    xn = O.scale_linear(O.clip(O.fixed_zero_mean_unit_variance(x, 103.5, 57.25), -1.5, 1.25), 0.5, 0.125)
    with torch.no_grad():
        want = O.sigmoid(net.to(dev)(torch.from_numpy(xn).to(dev)).cpu().numpy())
    np.testing.assert_allclose(out["mask"], want, rtol=1e-4, atol=1e-6)
    # the fused form (a single normaliser in the chain goes through the gather): fp64 arithmetic, <= 1e-5 from the fp32 reading
    sub = tmp_path / "fused"
    sub.mkdir()
    rdf2, net2 = write_model(sub, [{"id": "fixed_zero_mean_unit_variance", "kwargs": {"mean": [103.5], "std": [57.25], "axis": "channel"}}], [])
    out2 = asyncio.run(rt.predict(rdf2, img8))
    with torch.no_grad():
        want2 = net2.to(dev)(torch.from_numpy(O.fixed_zero_mean_unit_variance(x, 103.5, 57.25)).to(dev)).cpu().numpy()
    np.testing.assert_allclose(out2["mask"], want2, rtol=1e-4, atol=1e-5)


def test_scale_mean_variance_stays_on_the_device(dev):
    from bioengine_worker_b200 import kernels as K

    rng = np.random.default_rng(5)
    x = (rng.standard_normal((2, 3, 40, 56)) * 7 + 3).astype(np.float32)
    ref = (rng.standard_normal((2, 3, 40, 56)) * 0.5 - 2).astype(np.float32)
    xd, rd = torch.from_numpy(x).to(dev), torch.from_numpy(ref).to(dev)
    ref_norm = K.meanvar(rd, 6, 1e-6)[1]
    got = K.scale_mean_variance(xd, ref_norm, 6, 1e-6).cpu().numpy()
    want = O.scale_mean_variance(x, ref, axes=(2, 3))
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-5)


def test_predict_validates_the_sample_against_the_rdf(tmp_path, dev):
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    rdf_path, _ = write_model(tmp_path, [{"id": "zero_mean_unit_variance", "kwargs": {"axes": ["y", "x"]}}], [])
    rt = RuntimeDeployment.func_or_class()
    ok = np.zeros((1, 1, 96, 128), np.float32)       # 64 + k * 32 on both axes
    assert asyncio.run(rt.predict(rdf_path, ok))["mask"].shape == (1, 1, 96, 128)
    for bad in (np.zeros((1, 2, 96, 128), np.float32),    # two channels, one declared
                np.zeros((1, 1, 100, 128), np.float32),   # 100 is not 64 + k * 32
                np.zeros((1, 1, 32, 128), np.float32)):   # below the minimum
        with pytest.raises(ValueError):
            asyncio.run(rt.predict(rdf_path, bad))
    with pytest.raises(ValueError):
        asyncio.run(rt.predict(rdf_path, {"raw": ok, "extra": ok}))
    # with blocking any size at least as large as the inner block goes
    assert asyncio.run(rt.predict(rdf_path, np.zeros((1, 1, 100, 131), np.float32), default_blocksize_parameter=1))["mask"].shape == (1, 1, 100, 131)


_TWO_IN_TWO_OUT = '''
import torch
import torch.nn as nn


class TwoHeads(nn.Module):
    """Two input tensors (1 and 2 channels), two outputs (a logit map and a 3-channel feature map); receptive field 5."""

    def __init__(self, width=6):
        super().__init__()
        self.a = nn.Conv2d(1, width, 3, padding=1)
        self.b = nn.Conv2d(2, width, 3, padding=1)
        self.mix = nn.Conv2d(2 * width, width, 3, padding=1)
        self.logit = nn.Conv2d(width, 1, 1)
        self.feat = nn.Conv2d(width, 3, 1)

    def forward(self, raw, aux):
        h = torch.relu(self.mix(torch.cat([torch.relu(self.a(raw)), torch.relu(self.b(aux))], dim=1)))
        return self.logit(h) * 40.0, self.feat(h)
'''


def write_two_input_package(root, with_tests_from=None):
    """A model package that ships its own architecture file, two inputs with different pre-processing, two outputs with
    different post-processing and (optionally) test tensors."""
    import importlib.util

    (root / "net.py").write_text(_TWO_IN_TWO_OUT)
    spec = importlib.util.spec_from_file_location("two_heads_for_test", root / "net.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(11)
    net = mod.TwoHeads(width=6).eval()
    with torch.no_grad():    # centre the logit head on a probe so the mask is neither empty nor full
        g = torch.Generator().manual_seed(12)
        probe = net(torch.randn((1, 1, 64, 64), generator=g), torch.rand((1, 2, 64, 64), generator=g) * 1.5 - 0.75)[0]
        net.logit.bias.sub_(probe.median() / 40.0)
    torch.save(net.state_dict(), root / "weights.pt")
    sp = lambda halo=None: [{"type": "space", "id": a, **({"halo": halo} if halo is not None else {"size": {"min": 32, "step": 16}})} for a in "yx"]
    rdf = {
        "format_version": "0.5.3", "type": "model", "name": "two-heads",
        "inputs": [
            {"id": "raw", "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["em"]}] + sp(),
             "preprocessing": [{"id": "zero_mean_unit_variance", "kwargs": {"axes": ["y", "x"]}}]},
            {"id": "aux", "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["u", "v"]}] + sp(),
             "preprocessing": [{"id": "scale_linear", "kwargs": {"gain": 0.5, "offset": -1.0}},
                               {"id": "clip", "kwargs": {"min": -0.75, "max": 0.75}}]},
        ],
        "outputs": [
            {"id": "mask", "data": {"type": "uint8"},
             "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["fg"]}] + sp(8),
             "postprocessing": [{"id": "sigmoid"}, {"id": "binarize", "kwargs": {"threshold": 0.5}}]},
            {"id": "features", "data": {"type": "float32"},
             "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["f0", "f1", "f2"]}] + sp(8),
             "postprocessing": [{"id": "scale_linear", "kwargs": {"gain": 2.0, "offset": 0.25}}]},
        ],
        "weights": {"pytorch_state_dict": {"source": "weights.pt",
                                           "architecture": {"source": "net.py", "callable": "TwoHeads", "kwargs": {"width": 6}}}},
    }
    if with_tests_from is not None:
        for t, arr in zip(rdf["inputs"] + rdf["outputs"], with_tests_from):
            np.save(root / f"test_{t['id']}.npy", arr)
            t["test_tensor"] = {"source": f"test_{t['id']}.npy"}
    (root / "rdf.yaml").write_text(yaml.safe_dump(rdf))
    return str(root / "rdf.yaml"), net


def _two_heads_oracle(net, raw, aux, dev, block=None, halo=None):
    """The reference order on the CPU: whole-sample pre-processing per input → (halo blocks) → network → (inner-region
    reassembly) → post-processing per output."""
    a = O.zero_mean_unit_variance(raw, axes=(2, 3))
    b = O.clip(O.scale_linear(aux, 0.5, -1.0), -0.75, 0.75)
    if block is not None:
        a, b = O.blocks_tensor(a, block, halo), O.blocks_tensor(b, block, halo)
    with torch.no_grad():
        lg, ft = net.to(dev)(torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev))
    lg, ft = lg.cpu().numpy(), ft.cpu().numpy()
    if block is not None:
        lg = O.from_blocks(lg, (raw.shape[0], 1) + raw.shape[2:], block, halo)
        ft = O.from_blocks(ft, (raw.shape[0], 3) + raw.shape[2:], block, halo)
    return O.binarize(O.sigmoid(lg), 0.5), O.scale_linear(ft, 2.0, 0.25)


def test_predict_two_inputs_two_outputs_with_the_packages_own_architecture(tmp_path, dev):
    from bioengine_worker_b200 import models
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    models.strict_fp32()
    rng = np.random.default_rng(4)
    raw = (O.gaussian_blobs((144, 176), 25, seed=4)[None, None] * 90 + 10 + rng.standard_normal((1, 1, 144, 176))).astype(np.float32)
    aux = rng.uniform(-1, 4, size=(1, 2, 144, 176)).astype(np.float32)
    rdf_path, net = write_two_input_package(tmp_path)
    rt = RuntimeDeployment.func_or_class()
    # whole sample (144 = 32 + 7*16, 176 = 32 + 9*16)
    out = asyncio.run(rt.predict(rdf_path, {"raw": raw, "aux": aux}))
    mask, feat = _two_heads_oracle(net, raw, aux, dev)
    assert list(out) == ["mask", "features"] and out["mask"].dtype == np.uint8 and out["features"].shape == (1, 3, 144, 176)
    assert 0.05 < mask.mean() < 0.95
    assert np.array_equal(out["mask"], mask)
    np.testing.assert_allclose(out["features"], feat, rtol=1e-5, atol=1e-6)
    # blocks of 32 + 2*16 = 64 with the outputs' 8-px halo: the network's receptive field (5) fits inside the halo
    outb = asyncio.run(rt.predict(rdf_path, {"aux": aux, "raw": raw}, default_blocksize_parameter=2))
    maskb, featb = _two_heads_oracle(net, raw, aux, dev, (64, 64), (8, 8))
    assert np.array_equal(outb["mask"], maskb)
    np.testing.assert_allclose(outb["features"], featb, rtol=1e-5, atol=1e-6)
    # what create_sample_for_model rejects
    for bad in (raw, {"raw": raw}, {"raw": raw, "aux": aux, "more": aux}, {"raw": raw, "aux": aux[:, :, :128]},
                {"raw": raw, "aux": aux[:, :1]}):
        with pytest.raises(ValueError):
            asyncio.run(rt.predict(rdf_path, bad))


def test_model_test_uses_the_packaged_test_tensors(tmp_path, dev):
    """`RuntimeDeployment.test` = bioimageio.core.test_model: run the packaged test inputs, compare with the packaged test
    outputs; a package whose outputs do not reproduce fails."""
    from bioengine_worker_b200 import models
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    models.strict_fp32()
    rng = np.random.default_rng(5)
    raw = rng.normal(50, 20, size=(1, 1, 64, 80)).astype(np.float32)
    aux = rng.uniform(-1, 4, size=(1, 2, 64, 80)).astype(np.float32)
    good, bad = tmp_path / "good", tmp_path / "bad"
    good.mkdir(), bad.mkdir()
    _, net = write_two_input_package(good)
    mask, feat = _two_heads_oracle(net, raw, aux, dev)
    rdf_good, _ = write_two_input_package(good, with_tests_from=(raw, aux, mask, feat))
    rdf_bad, _ = write_two_input_package(bad, with_tests_from=(raw, aux, mask, feat + 0.01))
    rt = RuntimeDeployment.func_or_class()
    rep = asyncio.run(rt.test(rdf_good))
    assert rep["status"] == "passed" and rep["tested_against"] == "packaged test tensors"
    assert [d["status"] for d in rep["details"]] == ["passed", "passed"]
    rep = asyncio.run(rt.test(rdf_bad))
    assert rep["status"] == "failed" and [d["status"] for d in rep["details"]] == ["passed", "failed"]
    with pytest.raises(ValueError):
        asyncio.run(rt.test(rdf_good, additional_requirements="numpy"))


_RESAMPLERS = '''
import torch
import torch.nn as nn
import torch.nn.functional as F


class Up2(nn.Module):
    """3x3 convolution, then nearest-neighbour upsampling by 2: output twice the input size."""
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(1, 2, 3, padding=1)

    def forward(self, x):
        return F.interpolate(self.conv(x), scale_factor=2, mode="nearest")


class Down2(nn.Module):
    """3x3 convolution, then 2x2 average pooling: output half the input size."""
    def __init__(self):
        super().__init__()
        self.conv = nn.Conv2d(1, 1, 3, padding=1)

    def forward(self, x):
        return F.avg_pool2d(self.conv(x), 2)
'''


@pytest.mark.parametrize("name,pixel,halo_out,factor", [("Up2", 0.5, 4, 2.0), ("Down2", 2.0, 2, 0.5)])
def test_predict_output_of_another_size(tmp_path, dev, name, pixel, halo_out, factor):
    """Outputs whose size is the input's times a factor (`SizeReference` + axis `scale`): the blocks, their halo and the
    reassembly are in output pixels on the way out."""
    import importlib.util

    from bioengine_worker_b200 import models
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    models.strict_fp32()
    (tmp_path / "net.py").write_text(_RESAMPLERS)
    spec = importlib.util.spec_from_file_location("resamplers_for_test", tmp_path / "net.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(21)
    net = getattr(mod, name)().eval()
    torch.save(net.state_dict(), tmp_path / "weights.pt")
    rdf = {
        "format_version": "0.5.3", "type": "model", "name": name,
        "inputs": [{"id": "raw", "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["c"]}]
                    + [{"type": "space", "id": a, "size": {"min": 32, "step": 16}} for a in "yx"],
                    "preprocessing": [{"id": "zero_mean_unit_variance", "kwargs": {"axes": ["y", "x"]}}]}],
        "outputs": [{"id": "out", "data": {"type": "float32"},
                     "axes": [{"type": "batch"}, {"type": "channel", "channel_names": [f"o{i}" for i in range(2 if name == "Up2" else 1)]}]
                     + [{"type": "space", "id": a, "scale": pixel, "halo": halo_out, "size": {"tensor_id": "raw", "axis_id": a}} for a in "yx"]}],
        "weights": {"pytorch_state_dict": {"source": "weights.pt", "architecture": {"source": "net.py", "callable": name}}},
    }
    (tmp_path / "rdf.yaml").write_text(yaml.safe_dump(rdf))
    rng = np.random.default_rng(8)
    x = rng.normal(10, 3, size=(1, 1, 144, 176)).astype(np.float32)
    pre = O.zero_mean_unit_variance(x, axes=(2, 3))
    rt = RuntimeDeployment.func_or_class()
    out = asyncio.run(rt.predict(str(tmp_path / "rdf.yaml"), x))["out"]
    with torch.no_grad():
        want = net.to(dev)(torch.from_numpy(pre).to(dev)).cpu().numpy()
    assert out.shape == want.shape == (1, want.shape[1], int(144 * factor), int(176 * factor))
    np.testing.assert_allclose(out, want, rtol=1e-5, atol=1e-6)
    # blocks of 32 + 2*16 = 64 input pixels; the halo is declared in output pixels
    outb = asyncio.run(rt.predict(str(tmp_path / "rdf.yaml"), x, default_blocksize_parameter=2))["out"]
    halo_in = int(halo_out / factor)
    tiles = O.blocks_tensor(pre, (64, 64), (halo_in, halo_in))
    with torch.no_grad():
        t_out = net(torch.from_numpy(tiles).to(dev)).cpu().numpy()
    bo = int(64 * factor)
    wantb = O.from_blocks(t_out, want.shape, (bo, bo), (halo_out, halo_out))
    np.testing.assert_allclose(outb, wantb, rtol=1e-5, atol=1e-6)
    # away from the image border (reflect-padded halo there, zero padding in the whole-sample run) the halo covers the
    # network's receptive field and the two runs agree
    np.testing.assert_allclose(outb[..., 8:-8, 8:-8], want[..., 8:-8, 8:-8], rtol=1e-4, atol=1e-5)
    if name == "Down2":   # 150 is 32 + k*16 for no k, and with blocking 150 * 0.5 is fine but 151 is not whole
        with pytest.raises(ValueError):
            asyncio.run(rt.predict(str(tmp_path / "rdf.yaml"), x[:, :, :143], default_blocksize_parameter=2))


def test_per_channel_scale_linear_and_binarize(dev):
    """bioimageio `scale_linear` / `binarize` with one value per channel (`axis: channel`)."""
    from bioengine_worker_b200.rdf import Op
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    rt = RuntimeDeployment.func_or_class()
    rng = np.random.default_rng(3)
    x = rng.normal(0, 2, size=(2, 3, 37, 41)).astype(np.float32)
    gain, offset, thr = [0.5, 2.0, -1.25], [1.0, -3.0, 0.125], [0.1, -2.0, 1.5]
    xt = torch.from_numpy(x).to(dev)
    y = rt._run_ops_unfused(xt.clone(), [Op("scale_linear", {"gain": gain, "offset": offset, "axis": "channel"})], 2, 3)
    want = x * np.asarray(gain, np.float32)[None, :, None, None] + np.asarray(offset, np.float32)[None, :, None, None]
    assert np.array_equal(y.cpu().numpy(), want.astype(np.float32))
    y = rt._run_ops_unfused(xt.clone(), [Op("scale_linear", {"gain": 3.0, "offset": offset})], 2, 3)
    assert np.array_equal(y.cpu().numpy(), (x * np.float32(3.0) + np.asarray(offset, np.float32)[None, :, None, None]).astype(np.float32))
    x[0, 1, 0, :3] = [-2.0, np.nextafter(np.float32(-2.0), np.float32(0)), np.nextafter(np.float32(-2.0), np.float32(-3))]   # on / beside a threshold
    b = rt._run_ops_unfused(torch.from_numpy(x).to(dev), [Op("binarize", {"threshold": thr, "axis": "channel"})], 2, 3)
    assert np.array_equal(b.cpu().numpy(), (x > np.asarray(thr, np.float32)[None, :, None, None]).astype(np.float32))
    with pytest.raises(ValueError):
        rt._run_ops_unfused(xt.clone(), [Op("scale_linear", {"gain": [1.0, 2.0]})], 2, 3)
    with pytest.raises(ValueError):
        rt._run_ops_unfused(xt.clone(), [Op("binarize", {"threshold": thr, "axis": "x"})], 2, 3)


def test_screening_app_equals_single_process_screening(tmp_path, dev):
    """The screening app (entry → 3 shards on a worker object) against `screen_models` in one process and the oracle's AP:
    same integer table, same AP, same ranking; images given as arrays and as .npy sources."""
    from bioengine_worker_b200 import models
    from bioengine_worker_b200.screening import screen_models
    from bioengine_worker_b200.screening_deployment import ScreeningEntry, ScreeningWorker

    models.strict_fp32()
    n, size = 7, 160
    imgs, gts = [], []
    for i in range(n):
        img, clean = O.synthetic_image((size, size), 14 + i, seed=500 + i, with_truth=True)
        imgs.append(img)
        gts.append(O.label(clean > 0.5, 2))
    probe = O.zero_mean_unit_variance(imgs[0][None, None], axes=(1, 2, 3))
    rdfs = []
    for m in range(2):
        d = tmp_path / f"model{m}"
        d.mkdir()
        rdf_path, _ = write_model(d, [], [], seed=60 + m, calibrate_on=probe if m == 0 else -probe)
        rdfs.append(rdf_path)
    worker = ScreeningWorker.func_or_class(batch=3)
    entry = ScreeningEntry.func_or_class(worker, n_shards=3)
    sources = []
    for i, img in enumerate(imgs):     # every other image travels as a file
        if i % 2:
            np.save(tmp_path / f"img{i}.npy", img)
            sources.append(str(tmp_path / f"img{i}.npy"))
        else:
            sources.append(img)
    res = asyncio.run(entry.screen(rdfs, sources, gts, [0.5, 0.75]))
    nets = [worker._net(p, dev) for p in rdfs]
    one = screen_models(nets, imgs, gts, thresholds=(0.5, 0.75), device=dev, batch=4)
    assert res["n_shards"] == 3
    assert np.array_equal(res["table"], one["table"])
    assert np.array_equal(res["ap"], one["ap"], equal_nan=True) and res["ranking_index"] == one["ranking"].tolist()
    x = torch.from_numpy(np.stack([O.zero_mean_unit_variance(im[None, None], axes=(1, 2, 3))[0] for im in imgs])).to(dev)
    with torch.no_grad():
        logits = nets[0](x).cpu().numpy()
    preds = [O.label(O.binarize(O.sigmoid(l), 0.5)[0], 2) for l in logits]
    ap, tp, _, _ = O.average_precision(gts, preds, threshold=[0.5, 0.75])
    assert np.array_equal(res["table"][0, :, :2], tp.astype(np.int64))
    assert np.array_equal(res["ap"][0], ap, equal_nan=True)
    assert sum(p.max() for p in preds) > 0


def test_statistics_over_the_batch_axis(dev):
    """`axes` naming the batch axis (0.5) / `mode: per_dataset` (0.4): one set of percentiles or mean / std for all samples
    of the call, per channel or jointly, applied to every sample."""
    from bioengine_worker_b200 import kernels as K
    from bioengine_worker_b200.rdf import Op
    from bioengine_worker_b200.runtime_deployment import RuntimeDeployment

    rng = np.random.default_rng(12)
    x = (rng.normal(0, 1, size=(3, 2, 40, 56)) * np.array([1.0, 5.0])[None, :, None, None]
         + np.array([0.0, 10.0, -4.0])[:, None, None, None]).astype(np.float32)
    xt = torch.from_numpy(x).to(dev)
    cases = [
        (Op("zero_mean_unit_variance", {"axes": ["batch", "y", "x"]}), O.zero_mean_unit_variance(x, axes=(0, 2, 3))),
        (Op("zero_mean_unit_variance", {"axes": ["b", "c", "y", "x"]}), O.zero_mean_unit_variance(x, axes=(0, 1, 2, 3))),
        (Op("zero_mean_unit_variance", {"mode": "per_dataset", "axes": ["c", "y", "x"]}), O.zero_mean_unit_variance(x, axes=(0, 1, 2, 3))),
        (Op("scale_range", {"axes": ["batch", "y", "x"], "min_percentile": 2.0, "max_percentile": 99.0}),
         O.scale_range(x, 2.0, 99.0, axes=(0, 2, 3))),
        (Op("scale_range", {"axes": ["batch", "channel", "y", "x"], "min_percentile": 1.0, "max_percentile": 99.8}),
         O.scale_range(x, 1.0, 99.8, axes=(0, 1, 2, 3))),
    ]
    for op, want in cases:
        norm, mode = RuntimeDeployment._norm_params(xt, op, 3, 2)
        got = K.normalize(xt.clone(), norm).cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6, err_msg=str(op.kwargs))
        assert norm.shape[0] == (3 if mode == 2 else 6)
