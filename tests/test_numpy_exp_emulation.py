"""The float32 exp the CUDA sigmoid uses (csrc/be_common.cuh: be_numpy_expf) is a restatement of numpy's own SIMD
float32 exp.  This test keeps the restatement honest on the CPU: the same operation sequence, written in numpy,
must reproduce np.exp bit for bit (it does on AVX2+FMA3 / AVX512F hosts, which is every x86 box this runs on)."""
import numpy as np
import pytest

f32 = np.float32


def fma(a, b, c):
    # exact a*b in fp64 (24+24 significant bits), one rounding of the sum: an fp32 fused multiply-add
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


def numpy_expf_restated(x):
    x = x.astype(f32)
    magic = f32(12582912.0)
    k = (x * f32(1.442695040888963407359924681001892137)).astype(f32)
    k = ((k + magic).astype(f32) - magic).astype(f32)
    r = fma(k, f32(-6.93145752e-1), x)
    r = fma(k, f32(-1.42860677e-6), r)
    num = fma(f32(5.082762527590693718096e-04), r, f32(6.757896990527504603057e-03))
    for c in (5.114512081637298353406e-02, 2.473615434895520810817e-01, 7.257664613233124478488e-01,
              9.999999999980870924916e-01):
        num = fma(num, r, f32(c))
    den = fma(f32(2.159509375685829852307e-02), r, f32(-2.742335390411667452936e-01))
    den = fma(den, r, f32(1.0))
    return np.ldexp((num / den).astype(f32), k.astype(np.int32)).astype(f32)


def _simd_exp_host():
    try:
        from numpy._core._multiarray_umath import __cpu_features__ as feat
    except Exception:  # pragma: no cover
        return False
    return bool(feat.get("AVX512F") or (feat.get("AVX2") and feat.get("FMA3")))


@pytest.mark.skipif(not _simd_exp_host(), reason="numpy uses libm expf on this host, not its SIMD loop")
@pytest.mark.parametrize("scale", [1e-7, 1e-3, 1.0, 4.0, 20.0])
def test_restated_exp_is_numpy_exp(scale):
    rng = np.random.default_rng(int(scale * 1e7) % 1000)
    x = (rng.standard_normal(500_000) * scale).astype(f32)
    assert np.array_equal(np.exp(x), numpy_expf_restated(x))
