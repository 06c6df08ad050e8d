"""B200-native tiled bioimage inference path (drop-in for the model-runner hot path of aicell-lab/bioengine-worker).

Layout: csrc/ (sm_100a CUDA kernels + C-ABI → libbe_b200.so), _lib (ctypes binding), kernels (torch-tensor front end),
geometry (host tile tables), pipeline / runtime_deployment / analysis_deployment / metrics (reference-facing API).
"""
from . import _lib, geometry  # noqa: F401

__all__ = ["_lib", "geometry"]
