"""Minimal bioimage.io resource-description (rdf.yaml) reader: only what the tiled inference path needs.

Understands the parts of model RDF 0.5 (and the 0.4 spellings of the same things) that drive
`predict_sample_with[out]_blocking`: input/output ids and axes, pre-/post-processing op lists,
parameterised input sizes (`size: {min, step}`), output halo, packaged test tensors, and the `pytorch_state_dict` weights
entry with its architecture — a callable of this package (`UNet2d`, `UNet3d`) or a class in a source file of the model
package.  Output sizes follow an input's through `SizeReference` and the axes' `scale` (0.4: `shape.reference_tensor` /
`scale`); statistics over the batch axis / `mode: per_dataset` use the samples of the call (row f2 of SURVEY §8f).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from pathlib import Path
from typing import Any, Dict, List, Optional, Tuple

import yaml


@dataclass
class Op:
    name: str
    kwargs: Dict[str, Any] = field(default_factory=dict)


@dataclass
class TensorSpec:
    id: str
    axes: str                                   # e.g. "bcyx" / "bczyx"
    dtype: str = "float32"
    ops: List[Op] = field(default_factory=list)
    size_min: Dict[str, int] = field(default_factory=dict)   # spatial axis id -> min
    size_step: Dict[str, int] = field(default_factory=dict)  # spatial axis id -> step (0 = fixed)
    halo: Dict[str, int] = field(default_factory=dict)
    n_channels: Optional[int] = None            # from `channel_names` / a fixed channel size, if declared
    test_tensor: Optional[str] = None           # packaged example (`test_tensor.source`, or 0.4 `test_inputs/outputs`)
    scale: Dict[str, float] = field(default_factory=dict)     # spatial axis id -> physical size of one pixel (default 1)
    size_ref: Dict[str, Tuple[str, str, int]] = field(default_factory=dict)   # axis -> (tensor id, axis id, offset)
    ref_scale: Dict[str, float] = field(default_factory=dict)  # 0.4 outputs: size = reference size * scale + 2 * offset


@dataclass
class ModelSpec:
    root: Path
    name: str
    inputs: List[TensorSpec]
    outputs: List[TensorSpec]
    architecture: str
    arch_kwargs: Dict[str, Any]
    weights_source: Optional[str]
    seed: int = 0
    arch_source: Optional[str] = None           # python file inside the package that defines `architecture`


def _ops(entries) -> List[Op]:
    out = []
    for e in entries or []:
        name = e.get("id", e.get("name"))
        if name is None:
            raise ValueError(f"processing entry without id/name: {e}")
        out.append(Op(str(name), dict(e.get("kwargs") or {})))
    return out


def _axis_letter(ax: Dict[str, Any]) -> str:
    t = ax.get("type")
    if t == "batch":
        return "b"
    if t == "channel":
        return "c"
    if t == "space":
        return str(ax.get("id", "x"))
    if t == "index":
        return "i"
    raise ValueError(f"unsupported axis type {t!r}")


def _tensor(entry: Dict[str, Any], is_output: bool) -> TensorSpec:
    tid = str(entry.get("id", entry.get("name", "output0" if is_output else "input0")))
    key = "postprocessing" if is_output else "preprocessing"
    axes = entry.get("axes")
    spec = TensorSpec(id=tid, axes="", ops=_ops(entry.get(key)))
    if isinstance(axes, str):  # RDF 0.4: "bcyx", shape / halo lists
        spec.axes = axes
        spec.dtype = str(entry.get("data_type", "float32"))
        halo = entry.get("halo")
        if halo:
            spec.halo = {a: int(h) for a, h in zip(axes, halo) if a in "zyx"}
        shape = entry.get("shape")
        if isinstance(shape, dict) and "min" in shape:
            spec.size_min = {a: int(m) for a, m in zip(axes, shape["min"]) if a in "zyx"}
            spec.size_step = {a: int(s) for a, s in zip(axes, shape.get("step", [0] * len(axes))) if a in "zyx"}
        if isinstance(shape, dict) and "reference_tensor" in shape:   # implicit output shape
            ref = str(shape["reference_tensor"])
            for a, sc, off in zip(axes, shape.get("scale", [1.0] * len(axes)), shape.get("offset", [0] * len(axes))):
                if a in "zyx":
                    spec.ref_scale[a] = float(sc)
                    spec.size_ref[a] = (ref, a, int(round(2 * float(off))))
        return spec
    letters = []
    for ax in axes or []:
        l = _axis_letter(ax)
        letters.append(l)
        if l == "c":
            names = ax.get("channel_names")
            if isinstance(names, (list, tuple)) and names:
                spec.n_channels = len(names)
            elif isinstance(ax.get("size"), int):
                spec.n_channels = int(ax["size"])
        if l in "zyx":
            size = ax.get("size")
            if isinstance(size, dict) and "min" in size:
                spec.size_min[l], spec.size_step[l] = int(size["min"]), int(size.get("step", 0))
            elif isinstance(size, int):
                spec.size_min[l], spec.size_step[l] = int(size), 0
            elif isinstance(size, dict) and "tensor_id" in size:   # SizeReference
                spec.size_ref[l] = (str(size["tensor_id"]), str(size.get("axis_id", l)), int(size.get("offset", 0)))
            if "scale" in ax:
                spec.scale[l] = float(ax["scale"])
            if "halo" in ax:
                spec.halo[l] = int(ax["halo"])
    spec.axes = "".join(letters)
    tt = entry.get("test_tensor")
    if isinstance(tt, dict) and tt.get("source"):
        spec.test_tensor = str(tt["source"])
    data = entry.get("data")
    if isinstance(data, dict) and "type" in data:
        spec.dtype = str(data["type"])
    return spec


def load_model_spec(rdf_path: str) -> ModelSpec:
    path = Path(rdf_path)
    if not path.exists():
        raise FileNotFoundError(f"RDF not found: {rdf_path}")
    rdf = yaml.safe_load(path.read_text())
    if not isinstance(rdf, dict) or "inputs" not in rdf or "outputs" not in rdf:
        raise ValueError(f"{rdf_path} is not a model RDF (needs inputs/outputs)")
    weights = (rdf.get("weights") or {}).get("pytorch_state_dict")
    if not weights:
        raise ValueError("only `pytorch_state_dict` weights are supported by the B200 runtime")
    arch = weights.get("architecture") or {}
    source = None
    if isinstance(arch, dict):      # 0.5: {callable, kwargs} with `source: file.py` (ArchitectureFromFile) or without
        callable_ = arch.get("callable")
        source = arch.get("source")
    else:                           # 0.4: "file.py:ClassName" or a bare class name
        callable_ = str(arch).split(":")[-1]
        if ":" in str(arch):
            source = str(arch).rsplit(":", 1)[0]
    kwargs = dict((arch.get("kwargs") if isinstance(arch, dict) else None) or weights.get("kwargs") or {})
    inputs = [_tensor(e, False) for e in rdf["inputs"]]
    outputs = [_tensor(e, True) for e in rdf["outputs"]]
    for specs, key in ((inputs, "test_inputs"), (outputs, "test_outputs")):   # 0.4 keeps the examples at top level
        for t, src in zip(specs, rdf.get(key) or []):
            t.test_tensor = t.test_tensor or str(src)
    return ModelSpec(
        root=path.parent, name=str(rdf.get("name", path.parent.name)), inputs=inputs, outputs=outputs,
        architecture=str(callable_), arch_kwargs=kwargs, weights_source=weights.get("source"),
        seed=int((rdf.get("config") or {}).get("b200", {}).get("seed", 0)),
        arch_source=str(source) if source else None,
    )
