"""The C-ABI library loads and exports every symbol include/be_b200.h declares (no compute calls: CPU-safe)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "be_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(be_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = declared_symbols()
    for must in ("be_version", "be_quantile_f32", "be_tile_gather_f32", "be_stitch_blend_f32", "be_ccl_u8_i32",
                 "be_label_overlap_i32", "be_match_counts", "be_label_stats_i32"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from bioengine_worker_b200 import _lib

    assert os.path.exists(_lib.LIB_PATH), "libbe_b200.so not built: run `python __graft_entry__.py`"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [s for s in declared_symbols() if not hasattr(lib, s)]
    assert not missing, f"declared in be_b200.h but not exported: {missing}"


def test_bindings_cover_the_header():
    from bioengine_worker_b200 import _lib

    assert sorted(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()
    assert lib.be_version() == 1000
    assert lib.be_last_error_string() == b""


def test_workspace_queries_are_host_only():
    from bioengine_worker_b200 import _lib

    assert _lib.query("be_ccl_workspace_bytes", 1, 1, 4096, 4096) >= 4 * 4096 * 4096
    assert _lib.query("be_quantile_workspace_bytes", 1, 1 << 24) > 0
    assert _lib.query("be_meanvar_workspace_bytes", 4, 1 << 20) > 0
    assert _lib.query("be_match_workspace_bytes", 2, 512, 4096, 3) > 0


def test_invalid_arguments_fail_loudly_without_a_gpu():
    from bioengine_worker_b200 import _lib

    with pytest.raises(_lib.BeError) as e:
        _lib.call("be_binarize_f32_u8", None, None, 16, 0.5, None)
    assert "null pointer" in str(e.value)
    with pytest.raises(_lib.BeError):
        _lib.call("be_ccl_u8_i32", 8, 8, 1, 1, 0, 0, 2, 8, 8, None)  # empty image
