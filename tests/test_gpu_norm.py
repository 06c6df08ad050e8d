"""CUDA statistics / elementwise kernels against the oracle (bit-exact order statistics, <= 1e-5 rel floats)."""
import warnings

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # north_star: "within 1e-5 relative for normalised floats"


def K():
    from bioengine_worker_b200 import kernels

    return kernels


def _quant(x_np, q, n_seg, dev, eps=1e-6):
    x = torch.from_numpy(x_np).to(dev)
    qd = torch.tensor(q, dtype=torch.float64, device=dev)
    quant, norm = K().quantile(x, n_seg, qd, eps)
    torch.cuda.synchronize()
    return quant.cpu().numpy(), (None if norm is None else norm.cpu().numpy())


@pytest.mark.parametrize("n", [1, 2, 3, 5, 31, 1000, 4099, 1 << 16, (1 << 20) + 7])
@pytest.mark.parametrize("q", [(0.01, 0.998), (0.0, 1.0), (0.5,), (0.25, 0.75)])
def test_quantile_is_numpy_nanquantile_bit_exact(n, q, dev, rng):
    x = (rng.standard_normal(n) * 100 + 50).astype(np.float32)
    got, _ = _quant(x, q, 1, dev)
    want = np.nanquantile(x, np.asarray(q, np.float64), method="linear")
    assert np.array_equal(got[0], want)


def test_quantile_segments_ties_negatives_and_nans(dev, rng):
    n_seg, seg_len = 7, 50_001
    x = rng.integers(-40, 40, size=(n_seg, seg_len)).astype(np.float32)  # heavy ties, both signs, +-0
    x[1] *= 1e-30
    x[2] = 7.0  # constant segment
    x[3, ::3] = np.nan
    x[4, :] = np.nan  # all-NaN segment → NaN
    x[5, 5] = np.inf
    x[5, 6] = -np.inf
    x[6] = np.abs(x[6]) * 0.0 - 0.0
    q = (0.01, 0.99)
    got, norm = _quant(x, q, n_seg, dev)
    with warnings.catch_warnings(), np.errstate(all="ignore"):
        warnings.simplefilter("ignore")  # all-NaN slice
        want = np.nanquantile(x, np.asarray(q, np.float64), axis=1, method="linear").T
    assert np.array_equal(got, want, equal_nan=True)
    ok = ~np.isnan(want[:, 0])
    assert np.array_equal(norm[ok, 0], want[ok, 0])
    assert np.array_equal(norm[ok, 1], want[ok, 1] - want[ok, 0] + 1e-6)


@pytest.mark.parametrize("seg_len", [50_001, (1 << 18) + 77])   # whole-segment radix select / sampled path
@pytest.mark.parametrize("q", [(0.0, 1.0), (0.01, 0.99), (0.04, 0.97), (0.5,)])
def test_quantile_ranks_that_land_on_infinities(seg_len, q, dev, rng):
    """scale_range defaults to percentiles 0 / 100, so a single inf pixel puts a rank on it.  numpy's lerp then gives
    inf - inf = NaN or inf * 0 = NaN (or inf when only one neighbour is infinite and gamma > 0); the device lerp performs
    the same operations on the same two order statistics."""
    x = (rng.standard_normal((4, seg_len)) * 10).astype(np.float32)
    x[0, 7] = np.inf                               # one +inf: only q = 1 touches it
    x[1, rng.random(seg_len) < 0.05] = np.inf      # 5 % +inf
    x[1, rng.random(seg_len) < 0.03] = -np.inf     # 3 % -inf
    x[2, rng.random(seg_len) < 0.6] = -np.inf      # median on -inf
    x[3, :] = np.inf
    x[3, ::3] = np.nan
    got, _ = _quant(x, q, 4, dev)
    with warnings.catch_warnings(), np.errstate(all="ignore"):
        warnings.simplefilter("ignore")
        want = np.nanquantile(x, np.asarray(q, np.float64), axis=1, method="linear").T
    assert np.array_equal(got, want, equal_nan=True)


def test_quantile_c2_image_full_size(dev):
    """BASELINE config 2 input: 4096^2, p1.0 / p99.8 with hot pixels; numpy on the same data is the checker."""
    img = O.synthetic_image((2048, 2048), 1000, seed=2)
    img = np.tile(img, (2, 2))  # 4096^2 without paying 4x generation time
    got, _ = _quant(img, (0.01, 0.998), 1, dev)
    want = np.nanquantile(img, np.asarray((0.01, 0.998)), method="linear")
    assert np.array_equal(got[0], want)


def test_percentile_fibsem_call_site(dev, rng):
    img = (rng.random((700, 900)) * 255).astype(np.float32)
    got, _ = _quant(img, (0.01, 0.99), 1, dev)
    p1, p99 = np.percentile(img, [1, 99])
    assert got[0, 0] == p1 and got[0, 1] == p99


@pytest.mark.parametrize("shape,n_seg", [((1, 1, 512, 512), 1), ((3, 2, 129, 77), 6), ((1, 1, 5), 1)])
def test_meanvar_and_zmuv(shape, n_seg, dev, rng):
    x = (rng.standard_normal(shape) * 37 + 1000).astype(np.float32)
    xd = torch.from_numpy(x).to(dev)
    stats, norm = K().meanvar(xd, n_seg, 1e-6)
    y = K().normalize(xd, norm)
    torch.cuda.synchronize()
    x64 = x.reshape(n_seg, -1).astype(np.float64)
    np.testing.assert_allclose(stats[:, 0].cpu().numpy(), x64.mean(1), rtol=1e-12)
    np.testing.assert_allclose(stats[:, 1].cpu().numpy(), x64.var(1), rtol=1e-9)
    axes = tuple(range(2, x.ndim)) if n_seg > 1 else None
    want = O.zero_mean_unit_variance(x, axes=axes)
    np.testing.assert_allclose(y.cpu().numpy(), want, rtol=RTOL, atol=RTOL)
    # bitwise reproducible statistics
    stats2, _ = K().meanvar(xd, n_seg, 1e-6)
    assert torch.equal(stats, stats2)


def test_scale_range_apply_matches_both_reference_arithmetics(dev):
    img = O.synthetic_image((1, 1, 600, 1000), 200, seed=3)
    xd = torch.from_numpy(img).to(dev)
    q = torch.tensor([0.01, 0.998], dtype=torch.float64, device=dev)
    _, norm = K().quantile(xd, 1, q, 1e-6)
    y = K().normalize(xd, norm).cpu().numpy()
    want64 = O.scale_range(img, 1.0, 99.8, axes=None, arithmetic="float64")
    want32 = O.scale_range(img, 1.0, 99.8, axes=None, arithmetic="float32")
    assert np.array_equal(y, want64)  # same fp64 operations → same bits
    np.testing.assert_allclose(y, want32, rtol=RTOL, atol=RTOL)


def test_fibsem_normalise_clip(dev, rng):
    img = (rng.random((513, 300)) * 255).astype(np.float32)
    xd = torch.from_numpy(img).to(dev)
    q = torch.tensor([0.01, 0.99], dtype=torch.float64, device=dev)
    _, norm = K().quantile(xd, 1, q, 1e-6)
    y = K().normalize(xd, norm, clip=(0.0, 1.0)).cpu().numpy()
    assert np.array_equal(y, O.fibsem_normalize(img))


def test_elementwise_ops_bit_exact(dev, rng):
    x = (rng.standard_normal(100_003) * 4).astype(np.float32)
    x[:64] = np.linspace(-3e-7, 3e-7, 64, dtype=np.float32)  # straddles the sigmoid(x) == 0.5 rounding boundary
    x[64:70] = [np.nan, np.inf, -np.inf, 0.0, -0.0, 88.9]
    x[70:74] = [-88.0, -104.0, 17.0, -17.0]
    xd = torch.from_numpy(x).to(dev)
    k = K()
    with np.errstate(all="ignore"):
        assert np.array_equal(k.clip(xd, -1.5, 2.25).cpu().numpy(), O.clip(x, -1.5, 2.25), equal_nan=True)
        assert np.array_equal(k.scale_linear(xd, 1.7, -0.3).cpu().numpy(), O.scale_linear(x, 1.7, -0.3), equal_nan=True)
        s = k.sigmoid(xd).cpu().numpy()
        want = O.sigmoid(x)
        np.testing.assert_allclose(s, want, rtol=RTOL, equal_nan=True)
        # the device exp restates numpy's SIMD float32 exp operation for operation, so the values agree bit for bit
        assert np.array_equal(s, want, equal_nan=True)
        # the binarised mask of the sigmoid must be bit-exact
        assert np.array_equal(k.binarize(k.sigmoid(xd), 0.5).cpu().numpy(), O.binarize(want, 0.5))
        assert np.array_equal(k.binarize(xd, 0.5).cpu().numpy(), O.binarize(x, 0.5))


def test_unaligned_views(dev, rng):
    x = (rng.standard_normal(10_007)).astype(np.float32)
    xd = torch.from_numpy(x).to(dev)[3:]  # 12-byte offset: exercises the scalar head/tail paths
    got, _ = K().quantile(xd, 1, torch.tensor([0.3, 0.9], dtype=torch.float64, device=dev))
    assert np.array_equal(got.cpu().numpy()[0], np.nanquantile(x[3:], np.array([0.3, 0.9])))
    assert np.array_equal(K().binarize(xd, 0.1).cpu().numpy(), O.binarize(x[3:], 0.1))


def test_errors_are_loud(dev):
    from bioengine_worker_b200._lib import BeError

    x = torch.zeros(16, dtype=torch.float32, device=dev)
    with pytest.raises(BeError):
        K().quantile(x, 1, torch.tensor([0.1, 0.2, 0.3], dtype=torch.float64, device=dev))
    with pytest.raises(ValueError):
        K().quantile(x.cpu(), 1, torch.tensor([0.1], dtype=torch.float64, device=dev))


# ---- sampled-pivot path (segments >= 2^18 elements) ----

def _sample_positions(seg_len, seg=0, m=2048):
    """The positions qsel_sample_kernel reads (csrc/be_norm.cu: hash32 / qsel_sample_kernel), restated in numpy so a test
    can plant an unrepresentative sample and force the full-segment fallback."""
    i = np.arange(m, dtype=np.uint64)  # m = kSample
    x = (i * np.uint64(0x9E3779B9) + np.uint64(seg)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    x = (x * np.uint64(0x7FEB352D)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(15)
    x = (x * np.uint64(0x846CA68B)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    stride = np.uint64(seg_len // m)
    return (i * stride + x % stride).astype(np.int64)


@pytest.mark.parametrize("q", [(0.01, 0.998), (0.0, 1.0), (0.5,), (0.25, 0.75), (1e-7, 0.9999999)])
def test_sampled_quantile_distributions(q, dev, rng):
    n = (1 << 20) + 12345
    cases = {
        "normal": (rng.standard_normal(n) * 100 + 50).astype(np.float32),
        "uint8 image": rng.integers(0, 256, n).astype(np.float32),                 # 256 distinct values: all ties
        "half zeros": np.where(rng.random(n) < 0.5, 0.0, rng.random(n)).astype(np.float32),
        "constant": np.full(n, 3.25, np.float32),
        "two values": np.where(rng.random(n) < 0.3, -1.0, 2.0).astype(np.float32),
        "signed zeros": np.where(rng.random(n) < 0.5, -0.0, 0.0).astype(np.float32),
        "heavy tail": (rng.standard_cauchy(n) * 1e3).astype(np.float32),
        "sorted": np.sort(rng.standard_normal(n)).astype(np.float32),
        # ~160 distinct values: the list between the pivots is thousands of copies of one or two values, so the level-0
        # bins cannot be gathered into the leader CTA and the cluster-wide levels run
        "rounded normal": np.round(rng.standard_normal(n) * 20).astype(np.float32),
    }
    for name, x in cases.items():
        got, _ = _quant(x, q, 1, dev)
        want = np.nanquantile(x, np.asarray(q, np.float64), method="linear")
        assert np.array_equal(got[0], want), name


def test_sampled_quantile_nans_infs_and_segments(dev, rng):
    n_seg, seg_len = 5, (1 << 18) + 77
    x = (rng.standard_normal((n_seg, seg_len)) * 10).astype(np.float32)
    x[0, ::2] = np.nan                     # half NaN
    x[1, :] = np.nan                       # all NaN → NaN
    x[2, :1000] = np.inf
    x[2, 1000:3000] = -np.inf
    x[3] = np.round(x[3])                  # ties
    q = (0.01, 0.99)
    got, norm = _quant(x, q, n_seg, dev)
    with warnings.catch_warnings(), np.errstate(all="ignore"):
        warnings.simplefilter("ignore")
        want = np.nanquantile(x, np.asarray(q, np.float64), axis=1, method="linear").T
    assert np.array_equal(got, want, equal_nan=True)
    ok = ~np.isnan(want[:, 0])
    assert np.array_equal(norm[ok, 0], want[ok, 0])
    assert np.array_equal(norm[ok, 1], want[ok, 1] - want[ok, 0] + 1e-6)


def test_sampled_quantile_falls_back_when_the_sample_lies(dev, rng):
    """Plant tiny values exactly where the sampler looks: its brackets then miss the true quantiles completely and the
    selection must fall back to the whole segment — still exact."""
    n = 1 << 19
    x = (rng.random(n) * 1000 + 5000).astype(np.float32)
    x[_sample_positions(n)] = np.linspace(0.0, 1.0, 2048, dtype=np.float32)
    for q in ((0.5, 0.9), (0.01, 0.998)):
        got, _ = _quant(x, q, 1, dev)
        assert np.array_equal(got[0], np.nanquantile(x, np.asarray(q, np.float64), method="linear"))


def test_sampled_quantile_many_segments_one_call(dev, rng):
    n_seg, seg_len = 24, 1 << 18
    x = rng.gamma(2.0, 50.0, size=(n_seg, seg_len)).astype(np.float32)
    got, _ = _quant(x, (0.01, 0.998), n_seg, dev)
    want = np.nanquantile(x, np.array([0.01, 0.998]), axis=1, method="linear").T
    assert np.array_equal(got, want)


def test_sampled_quantile_long_tied_lists_stream_every_level(dev, rng):
    """8 M elements rounded to ~300 values: a bracket holds ~10^5 equal elements, more than a CTA's shared-memory list, so
    the selection streams its slice on every level (the path a fallback to a whole segment takes as well)."""
    n = 1 << 23
    x = np.round(rng.standard_normal(n) * 40).astype(np.float32)
    for q in ((0.01, 0.998), (0.3, 0.7)):
        got, _ = _quant(x, q, 1, dev)
        assert np.array_equal(got[0], np.nanquantile(x, np.asarray(q, np.float64), method="linear"))


def test_sampled_quantile_ranks_on_and_next_to_pivot_ties(dev, rng):
    """Ranks that fall inside the run of elements equal to a pivot, and one element either side of it."""
    n = 1 << 19
    base = rng.random(n).astype(np.float32)
    for frac in (0.2, 0.5, 0.8):
        x = base.copy()
        x[rng.random(n) < 0.3] = np.float32(frac)      # 30 % of the data is one value sitting at quantile ~frac
        xs = np.sort(x)
        lo, hi = np.searchsorted(xs, np.float32(frac), "left"), np.searchsorted(xs, np.float32(frac), "right")
        for k in (lo - 1, lo, lo + 1, (lo + hi) // 2, hi - 2, hi - 1, hi):
            q = (k / (n - 1), min(1.0, (k + 0.5) / (n - 1)))
            got, _ = _quant(x, q, 1, dev)
            assert np.array_equal(got[0], np.nanquantile(x, np.asarray(q, np.float64), method="linear")), (frac, k)


def test_selection_with_the_portable_cluster_size(dev):
    """The selection kernel takes clusters of 16 CTAs where the device grants them; BE_B200_CLUSTER_MAX=8 (read once per
    process, hence the subprocess) forces the portable size a smaller part would fall back to."""
    import os, subprocess, sys, textwrap

    code = textwrap.dedent('''
        import numpy as np, torch
        from bioengine_worker_b200 import kernels as K
        rng = np.random.default_rng(7)
        dev = torch.device("cuda:0")
        for x in ((rng.standard_normal((2, (1 << 19) + 5)) * 50).astype(np.float32),
                  np.round(rng.standard_normal((1, 1 << 20)) * 20).astype(np.float32)):
            q = np.array([0.01, 0.998])
            got, _ = K.quantile(torch.from_numpy(x).to(dev), x.shape[0], torch.from_numpy(q).to(dev), 1e-6)
            want = np.nanquantile(x, q, axis=1, method="linear").T
            assert np.array_equal(got.cpu().numpy(), want)
        print("ok")
    ''')
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, BE_B200_CLUSTER_MAX="8", PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", ""))
    out = subprocess.run([sys.executable, "-c", code], env=env, cwd=root, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


@pytest.mark.parametrize("dtype", ["uint8", "int8", "uint16", "int16", "int32", "uint32", "float16", "float64"])
@pytest.mark.parametrize("n", [1, 7, 4096, 100_003])
def test_cast_to_float32_is_numpy_astype(dtype, n, dev, rng):
    """`ensure_dtype(float32)` on the device: every value of the small types, rounding of the wide ones."""
    dt = np.dtype(dtype)
    if dt.kind in "ui":
        info = np.iinfo(dt)
        x = rng.integers(info.min, info.max, size=n, endpoint=True, dtype=dt)
        x[: min(n, 4)] = [info.min, info.max, info.max - 1, 0][: min(n, 4)]
    elif dt == np.float16:
        x = rng.standard_normal(n).astype(dt)
        x[: min(n, 3)] = [np.inf, -0.0, 6.1e-5][: min(n, 3)]
    else:
        x = rng.standard_normal(n) * 10.0 ** rng.integers(-40, 40, n)     # beyond the float32 range too
    for off in (0, 1):   # an unaligned start as well
        t = torch.from_numpy(x).to(dev)[off:]
        if t.numel() == 0:
            continue
        with np.errstate(over="ignore"):
            want = x[off:].astype(np.float32)
        got = K().to_float32(t.contiguous() if off == 0 else t.clone()).cpu().numpy()
        assert np.array_equal(got, want, equal_nan=True), (dtype, n, off)


def test_pipeline_takes_uint16_frames(dev):
    """A uint16 frame uploaded as such gives exactly what its float32 copy gives (C5's frames are uint16)."""
    from bioengine_worker_b200.pipeline import PipelineSpec, TiledPipeline

    img16 = np.clip(O.synthetic_image((1, 1, 200, 264), 30, seed=9), 0, 65535).astype(np.uint16)
    spec = PipelineSpec(block=(64, 64), halo=(8, 8), convention="bioimageio", preprocess="scale_range")
    pipe = TiledPipeline(img16.shape, spec, device=dev)
    fwd = lambda t: (t * 6.0 - 2.0).contiguous()
    a = pipe.run(img16, fwd)
    b = pipe.run(img16.astype(np.float32), fwd)
    assert torch.equal(a["labels"], b["labels"]) and torch.equal(a["mask"], b["mask"])
    host_l = [torch.empty((1, 1, 200, 264), dtype=torch.int32).pin_memory() for _ in range(3)]
    host_n = [torch.empty((1,), dtype=torch.int32).pin_memory() for _ in range(3)]
    frames = [torch.from_numpy(img16).pin_memory() for _ in range(3)]
    pipe.run_stream(frames, fwd, host_l, host_n).synchronize()
    for l in host_l:
        assert np.array_equal(l.numpy(), a["labels"].cpu().numpy())


def test_quantile_of_a_volume_sized_segment(dev):
    """2^26 + 12345 elements in one segment (the 8192-element sample path; C3's 512^3 volume is twice this): exact order
    statistics against numpy, with hot voxels and a constant stretch in the data."""
    from bioengine_worker_b200 import kernels as K

    rng = np.random.default_rng(26)
    n = (1 << 26) + 12345
    x = rng.normal(120.0, 40.0, size=n).astype(np.float32)
    x[rng.integers(0, n, 50000)] = 65535.0
    x[1000:200000] = 77.0
    q = np.array([0.01, 0.5, 0.99, 0.998])
    want = np.quantile(x.astype(np.float64), q)
    xt = torch.from_numpy(x).to(dev)
    for qa in (q[[0, 2]], q[[1, 3]]):
        got, _ = K.quantile(xt, 1, torch.tensor(qa, dtype=torch.float64, device=dev), 1e-6)
        assert np.array_equal(got.cpu().numpy()[0], np.quantile(x.astype(np.float64), qa))
    assert want[3] > want[2] > want[1] > want[0]
