"""Host-side logic of the model-runner mirror that needs no GPU: the RDF reader (0.5 and 0.4 spellings), the sample /
input-name rules of `create_sample_for_model`, axis handling, and loading an architecture from a package's own file."""
import numpy as np
import pytest
import yaml

from bioengine_worker_b200 import models
from bioengine_worker_b200.rdf import load_model_spec
from bioengine_worker_b200.runtime_deployment import RuntimeDeployment


def _write(tmp_path, rdf):
    (tmp_path / "rdf.yaml").write_text(yaml.safe_dump(rdf))
    return str(tmp_path / "rdf.yaml")


def test_rdf_05_fields(tmp_path):
    rdf = {
        "format_version": "0.5.3", "type": "model", "name": "m",
        "inputs": [{"id": "raw", "test_tensor": {"source": "in.npy"},
                    "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["a", "b"]},
                             {"type": "space", "id": "y", "size": {"min": 64, "step": 16}},
                             {"type": "space", "id": "x", "size": 128}],
                    "preprocessing": [{"id": "ensure_dtype", "kwargs": {"dtype": "float32"}},
                                      {"id": "scale_range", "kwargs": {"min_percentile": 1, "max_percentile": 99.8, "axes": ["y", "x"]}}]}],
        "outputs": [{"id": "out", "data": {"type": "uint8"}, "test_tensor": {"source": "out.npy"},
                     "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["p"]},
                              {"type": "space", "id": "y", "halo": 8}, {"type": "space", "id": "x", "halo": 12}],
                     "postprocessing": [{"id": "sigmoid"}, {"id": "binarize", "kwargs": {"threshold": 0.3}}]}],
        "weights": {"pytorch_state_dict": {"source": "w.pt", "architecture": {"source": "net.py", "callable": "Net", "kwargs": {"k": 3}}}},
    }
    spec = load_model_spec(_write(tmp_path, rdf))
    i, o = spec.inputs[0], spec.outputs[0]
    assert (i.id, i.axes, i.n_channels, i.test_tensor) == ("raw", "bcyx", 2, "in.npy")
    assert i.size_min == {"y": 64, "x": 128} and i.size_step == {"y": 16, "x": 0}
    assert [op.name for op in i.ops] == ["ensure_dtype", "scale_range"] and i.ops[1].kwargs["max_percentile"] == 99.8
    assert (o.id, o.dtype, o.halo, o.test_tensor) == ("out", "uint8", {"y": 8, "x": 12}, "out.npy")
    assert (spec.architecture, spec.arch_source, spec.arch_kwargs, spec.weights_source) == ("Net", "net.py", {"k": 3}, "w.pt")


def test_rdf_04_spellings(tmp_path):
    rdf = {
        "format_version": "0.4.10", "type": "model", "name": "old",
        "inputs": [{"name": "input0", "axes": "bcyx", "data_type": "float32", "shape": {"min": [1, 1, 32, 32], "step": [0, 0, 16, 16]},
                    "preprocessing": [{"name": "zero_mean_unit_variance", "kwargs": {"axes": "yx"}}]}],
        "outputs": [{"name": "output0", "axes": "bcyx", "data_type": "float32", "halo": [0, 0, 4, 4]}],
        "test_inputs": ["ti.npy"], "test_outputs": ["to.npy"],
        "weights": {"pytorch_state_dict": {"source": "w.pt", "architecture": "unet.py:UNet", "kwargs": {"depth": 2}}},
    }
    spec = load_model_spec(_write(tmp_path, rdf))
    assert spec.inputs[0].size_min == {"y": 32, "x": 32} and spec.inputs[0].size_step == {"y": 16, "x": 16}
    assert spec.outputs[0].halo == {"y": 4, "x": 4}
    assert (spec.architecture, spec.arch_source, spec.arch_kwargs) == ("UNet", "unet.py", {"depth": 2})
    assert spec.inputs[0].test_tensor == "ti.npy" and spec.outputs[0].test_tensor == "to.npy"
    assert spec.inputs[0].ops[0].name == "zero_mean_unit_variance"


def test_rdf_errors(tmp_path):
    with pytest.raises(FileNotFoundError):
        load_model_spec(str(tmp_path / "missing.yaml"))
    with pytest.raises(ValueError):
        load_model_spec(_write(tmp_path, {"type": "dataset"}))
    with pytest.raises(ValueError):   # no pytorch_state_dict entry
        load_model_spec(_write(tmp_path, {"inputs": [], "outputs": [], "weights": {"onnx": {"source": "m.onnx"}}}))


def test_sample_arrays_follow_create_sample_for_model(tmp_path):
    two = {"inputs": [{"id": "a", "axes": "bcyx"}, {"id": "b", "axes": "bcyx"}], "outputs": [{"id": "o", "axes": "bcyx"}],
           "weights": {"pytorch_state_dict": {"architecture": {"callable": "UNet2d"}}}}
    one = {"inputs": [{"id": "a", "axes": "bcyx"}], "outputs": [{"id": "o", "axes": "bcyx"}],
           "weights": {"pytorch_state_dict": {"architecture": {"callable": "UNet2d"}}}}
    s2, s1 = load_model_spec(_write(tmp_path, two)), load_model_spec(_write(tmp_path, one))
    x = np.zeros((1, 1, 8, 8), np.float32)
    f = RuntimeDeployment._sample_arrays
    assert list(f(x, s1)) == ["a"] and list(f({"a": x}, s1)) == ["a"] and list(f({"whatever": x}, s1)) == ["a"]
    assert list(f({"b": x, "a": x}, s2)) == ["a", "b"]
    for bad, spec in ((x, s2), ({"a": x}, s2), ({"a": x, "b": x, "c": x}, s2), ({"a": x, "z": x}, s1)):
        with pytest.raises(ValueError):
            f(bad, spec)


def test_to_ncsp_axis_handling():
    f = RuntimeDeployment._to_ncsp
    assert f(np.zeros((5, 7), np.float64), "bcyx").shape == (1, 1, 5, 7)
    assert f(np.zeros((5, 7), np.float64), "bcyx").dtype == np.float64        # converted on the device
    assert f(np.zeros((2, 5, 7, 3), np.uint8), "byxc").shape == (2, 3, 5, 7) and f(np.zeros((2, 5, 7, 3), np.uint8), "byxc").dtype == np.uint8
    assert f(np.zeros((4, 5, 6), np.int64), "bczyx").shape == (1, 1, 4, 5, 6)
    with pytest.raises(ValueError):
        f(np.zeros((1, 1, 1, 5, 7)), "bcyx")


def test_architecture_from_the_package(tmp_path):
    (tmp_path / "net.py").write_text("import torch.nn as nn\nclass Tiny(nn.Module):\n    def __init__(self, c=2):\n        super().__init__()\n        self.conv = nn.Conv2d(1, c, 1)\n    def forward(self, x):\n        return self.conv(x)\n")
    net = models.build_from_source(tmp_path, "net.py", "Tiny", c=3)
    assert net.conv.out_channels == 3 and not net.training and all(not p.requires_grad for p in net.parameters())
    with pytest.raises(ValueError):
        models.build_from_source(tmp_path, "net.py", "Missing")
    with pytest.raises(FileNotFoundError):
        models.build_from_source(tmp_path, "../outside.py", "Tiny")
    with pytest.raises(FileNotFoundError):
        models.build_from_source(tmp_path, "nope.py", "Tiny")


def test_output_ratio_from_size_reference(tmp_path):
    from fractions import Fraction

    rdf = {"inputs": [{"id": "raw", "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["c"]},
                                              {"type": "space", "id": "y", "scale": 2.0, "size": {"min": 32, "step": 16}},
                                              {"type": "space", "id": "x", "size": {"min": 32, "step": 16}}]}],
           "outputs": [{"id": "o", "axes": [{"type": "batch"}, {"type": "channel", "channel_names": ["c"]},
                                             {"type": "space", "id": "y", "scale": 1.0, "size": {"tensor_id": "raw", "axis_id": "y"}},
                                             {"type": "space", "id": "x", "scale": 4.0, "size": {"tensor_id": "raw", "axis_id": "x"}}]}],
           "weights": {"pytorch_state_dict": {"architecture": {"callable": "UNet2d"}}}}
    spec = load_model_spec(_write(tmp_path, rdf))
    assert RuntimeDeployment._output_ratio(spec, spec.outputs[0], ["y", "x"]) == [Fraction(2), Fraction(1, 4)]
    rdf["outputs"][0]["axes"][2]["size"]["offset"] = -4
    with pytest.raises(ValueError):
        spec = load_model_spec(_write(tmp_path, rdf))
        RuntimeDeployment._output_ratio(spec, spec.outputs[0], ["y", "x"])
    old = {"inputs": [{"name": "i", "axes": "bcyx", "shape": {"min": [1, 1, 32, 32], "step": [0, 0, 16, 16]}}],
           "outputs": [{"name": "o", "axes": "bcyx", "shape": {"reference_tensor": "i", "scale": [1, 1, 2, 0.5], "offset": [0, 0, 0, 0]}}],
           "weights": {"pytorch_state_dict": {"architecture": "UNet2d"}}}
    spec = load_model_spec(_write(tmp_path, old))
    assert RuntimeDeployment._output_ratio(spec, spec.outputs[0], ["y", "x"]) == [Fraction(2), Fraction(1, 2)]


def test_host_binding_is_a_no_op_without_nvml(monkeypatch):
    """`bind_host_to_gpu` never raises: without NVML (this container) or with the switch set it reports why it did nothing."""
    import os

    from bioengine_worker_b200.hostbind import bind_host_to_gpu

    before = os.sched_getaffinity(0)
    info = bind_host_to_gpu(0)
    assert isinstance(info["bound"], bool) and (info["bound"] or "reason" in info)
    monkeypatch.setenv("BE_B200_NO_NUMA_BIND", "1")
    assert bind_host_to_gpu(0)["reason"].startswith("disabled")
    assert os.sched_getaffinity(0) == before or info["bound"]
