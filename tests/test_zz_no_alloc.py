"""be_quantile_f32_noalloc — the quantile entry point for hosts that cannot allocate device memory (its scratch is a
fixed array inside the library instead of a caller workspace).

Called through the C-ABI exactly as such a host would: raw pointers only, no workspace.  Its results must be the
ones be_quantile_f32 gives (bit for bit: same kernels, same data), also on a second call, which finds the shared
scratch array still holding what the first call left there.  Collected last on purpose.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(x_np, n_seg, q, dev, eps=1e-6):
    from bioengine_worker_b200 import _lib

    x = torch.from_numpy(x_np).to(dev)
    qd = torch.tensor(q, dtype=torch.float64, device=dev)
    seg_len = x.numel() // n_seg
    stream = torch.cuda.current_stream().cuda_stream
    out = {}
    for name in ("be_quantile_f32", "be_quantile_f32_noalloc"):
        quant = torch.full((n_seg, len(q)), -1.0, dtype=torch.float64, device=dev)
        norm = torch.full((n_seg, 2), -1.0, dtype=torch.float64, device=dev)
        args = [x.data_ptr(), n_seg, seg_len, qd.data_ptr(), len(q), eps, quant.data_ptr(), norm.data_ptr()]
        if name == "be_quantile_f32":
            ws = torch.empty(_lib.query("be_quantile_workspace_bytes", n_seg, seg_len), dtype=torch.uint8, device=dev)
            args.append(ws.data_ptr())
        _lib.call(name, *args, stream)
        torch.cuda.synchronize()
        out[name] = (quant.cpu().numpy(), norm.cpu().numpy())
    return out["be_quantile_f32"], out["be_quantile_f32_noalloc"]


def test_noalloc_matches_workspace_entry_point_twice_in_a_row(dev):
    rng = np.random.default_rng(7)
    first = (rng.standard_normal((3, 200_003)) * 50 + 10).astype(np.float32)
    first[1, ::5] = np.nan
    second = rng.integers(-5, 5, size=(8, 65_536)).astype(np.float32)  # other shape, heavy ties, the maximum 8 segments
    for x, n_seg in ((first, 3), (second, 8)):
        (q_ws, n_ws), (q_na, n_na) = _both(x, n_seg, (0.01, 0.998), dev)
        assert np.array_equal(q_ws, q_na, equal_nan=True)
        assert np.array_equal(n_ws, n_na, equal_nan=True)
        want = np.nanquantile(x, np.array([0.01, 0.998]), axis=1).T
        assert np.array_equal(q_na, want)


def test_noalloc_ranks_on_infinities(dev):
    """Ranks that land in the key buckets of -inf / +inf (those buckets also span NaN encodings)."""
    import warnings

    rng = np.random.default_rng(11)
    x = (rng.standard_normal((3, 70_001)) * 10).astype(np.float32)
    x[0, 3] = np.inf
    x[1, rng.random(70_001) < 0.05] = np.inf
    x[1, rng.random(70_001) < 0.05] = -np.inf
    x[2, :] = -np.inf
    for q in ((0.0, 1.0), (0.02, 0.97)):
        (q_ws, _), (q_na, _) = _both(x, 3, q, dev)
        with warnings.catch_warnings(), np.errstate(all="ignore"):
            warnings.simplefilter("ignore")
            want = np.nanquantile(x, np.array(q), axis=1).T
        assert np.array_equal(q_na, want, equal_nan=True)
        assert np.array_equal(q_ws, q_na, equal_nan=True)


def test_noalloc_rejects_more_segments_than_its_scratch_holds(dev):
    from bioengine_worker_b200 import _lib

    x = torch.zeros((9, 128), dtype=torch.float32, device=dev)
    q = torch.tensor([0.5], dtype=torch.float64, device=dev)
    quant = torch.empty((9, 1), dtype=torch.float64, device=dev)
    with pytest.raises(_lib.BeError) as e:
        _lib.call("be_quantile_f32_noalloc", x.data_ptr(), 9, 128, q.data_ptr(), 1, 1e-6, quant.data_ptr(), None,
                  torch.cuda.current_stream().cuda_stream)
    assert e.value.code == _lib.BE_ERR_TOO_LARGE and "8 segments" in str(e.value)
