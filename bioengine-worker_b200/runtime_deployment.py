"""`RuntimeDeployment` — B200 replacement for ref: apps/model-runner/runtime_deployment.py.

Same class name, same `predict(...)` / `test(...)` signatures, same `{output_id: ndarray}` result, same error
conventions (FileNotFoundError for a missing RDF, ValueError for bad inputs, CUDA OOM re-raised as a plain RuntimeError
after `torch.cuda.empty_cache()`, ref: runtime_deployment.py:253-254,296-310).  What changed is underneath: instead of
bioimageio.core running numpy pre-/post-processing on the CPU and shipping every block to the GPU and back, the sample
is uploaded once and `TiledPipeline` runs statistics, normalise + halo-block gather, the torch forward, inner-region
stitching and post-processing on the device.

`ray` is optional: with Ray Serve installed the class is a deployment with the reference's options; without it the
decorators are no-ops and `RuntimeDeployment.func_or_class()` still constructs the plain object (the reference's own
local-dev pattern, ref: runtime_deployment.py:328).
"""
from __future__ import annotations

import hashlib
import json
import logging
import os
from collections import OrderedDict
from fractions import Fraction
from pathlib import Path
from typing import Dict, List, Literal, Optional, Union

import numpy as np
import torch

from . import _lib, kernels as K, models
from .pipeline import PipelineSpec, TiledPipeline
from .rdf import ModelSpec, Op, load_model_spec

logger = logging.getLogger("ray.serve")

from ._serve_compat import deployment

_deployment = deployment(
    ray_actor_options={"num_cpus": 1, "num_gpus": 1, "memory": 12 * 1024 * 1024 * 1024},
    max_ongoing_requests=1,
    autoscaling_config={"min_replicas": 1, "initial_replicas": 1, "max_replicas": 2,
                        "target_num_ongoing_requests_per_replica": 0.8, "metrics_interval_s": 2.0,
                        "look_back_period_s": 10.0, "downscale_delay_s": 300, "upscale_delay_s": 0.0},
    health_check_period_s=30.0, health_check_timeout_s=30.0, graceful_shutdown_timeout_s=120.0,
    graceful_shutdown_wait_loop_s=2.0,
)


_NORMALISERS = ("scale_range", "zero_mean_unit_variance", "fixed_zero_mean_unit_variance")


class _LoadedModel:
    """A model description with its weights on the device (what bioimageio calls a loaded prediction pipeline)."""

    def __init__(self, spec: ModelSpec, device: torch.device, blocksize_parameter: Optional[int]):
        self.spec, self.device, self.n = spec, device, blocksize_parameter
        if spec.arch_source:     # the package ships its own architecture file
            self.net = models.build_from_source(spec.root, spec.arch_source, spec.architecture, **spec.arch_kwargs)
        else:
            self.net = models.build_model(spec.architecture, seed=spec.seed, **spec.arch_kwargs)
        if spec.weights_source:
            src = spec.root / spec.weights_source
            if not src.exists():
                raise FileNotFoundError(f"weights not found: {src}")
            self.net.load_state_dict(torch.load(src, map_location="cpu"))
        self.net = self.net.to(device).eval()
        self.pipelines: Dict[tuple, TiledPipeline] = {}


@_deployment
class RuntimeDeployment:
    """Internal deployment for running model inference on the B200 tiled path."""

    def __init__(self):
        self._kwargs_cache = {}
        self._pipelines: "OrderedDict[str, _LoadedModel]" = OrderedDict()
        self._cache_size = int(os.environ.get("PIPELINE_CACHE_SIZE", 10))
        models.strict_fp32()

    # === model test ===

    async def test(self, rdf_path: str, additional_requirements: Optional[List[str]] = None) -> dict:
        """ref: runtime_deployment.py:86-99 (`bioimageio.core.test_model`): run the model on its packaged test inputs and
        compare with its packaged test outputs (absolute 1.5e-4 — bioimageio's `decimal=4` — plus 1e-3 relative).  A
        package without test tensors is run once on a zero sample of its minimum shape (status "passed" = it ran)."""
        if additional_requirements is not None and not isinstance(additional_requirements, list):
            raise ValueError("additional_requirements must be a list of strings.")
        spec = load_model_spec(rdf_path)
        have_tests = all(t.test_tensor and (spec.root / t.test_tensor).exists() for t in spec.inputs + spec.outputs)
        if not have_tests:
            arrays = {}
            for inp in spec.inputs:
                spatial = [max(inp.size_min.get(a, 64), 16) for a in inp.axes if a in "zyx"]
                arrays[inp.id] = np.zeros([1, inp.n_channels or spec.arch_kwargs.get("in_channels", 1)] + spatial, np.float32)
            out = await self.predict(rdf_path, arrays)
            return {"status": "passed", "name": spec.name, "tested_against": "zero sample (no packaged test tensors)",
                    "details": [{"name": k, "shape": list(v.shape), "dtype": str(v.dtype)} for k, v in out.items()]}
        arrays = {inp.id: np.load(spec.root / inp.test_tensor) for inp in spec.inputs}
        out = await self.predict(rdf_path, arrays)
        details, ok = [], True
        for o in spec.outputs:
            want, got = np.load(spec.root / o.test_tensor), out[o.id]
            same_shape = tuple(want.shape) == tuple(got.shape)
            close = same_shape and bool(np.allclose(got.astype(np.float64), want.astype(np.float64), rtol=1e-3, atol=1.5e-4))
            ok &= close
            details.append({"name": o.id, "shape": list(got.shape), "dtype": str(got.dtype), "status": "passed" if close else "failed",
                            "max_abs_diff": float(np.abs(got.astype(np.float64) - want).max()) if same_shape else None})
        return {"status": "passed" if ok else "failed", "name": spec.name, "tested_against": "packaged test tensors",
                "details": details}

    # === prediction ===

    def _set_prediction_kwargs(self, rdf_path, weights_format, device, default_blocksize_parameter,
                               latest_remote_modified=None) -> str:
        """Cache key of a loaded pipeline — same recipe as ref: runtime_deployment.py:160-185."""
        pipeline_kwargs = {
            "rdf_path": rdf_path,
            "latest_remote_modified": latest_remote_modified,
            "create_kwargs": {"weights_format": weights_format, "device": device,
                              "default_blocksize_parameter": default_blocksize_parameter},
        }
        cache_key = hashlib.md5(json.dumps(pipeline_kwargs, sort_keys=True).encode()).hexdigest()
        self._kwargs_cache[cache_key] = pipeline_kwargs
        return cache_key

    async def _create_prediction_pipeline(self, cache_key: str) -> _LoadedModel:
        if cache_key in self._pipelines:
            self._pipelines.move_to_end(cache_key)
            self._kwargs_cache.pop(cache_key, None)
            return self._pipelines[cache_key]
        pipeline_kwargs = self._kwargs_cache.get(cache_key)
        if not pipeline_kwargs:
            raise ValueError(f"No pipeline kwargs found for cache key: {cache_key}")
        try:
            ck = pipeline_kwargs["create_kwargs"]
            if ck["weights_format"] not in (None, "pytorch_state_dict"):
                raise ValueError(f"weights_format {ck['weights_format']!r} is not supported (pytorch_state_dict only)")
            if ck["device"] == "cpu":
                raise ValueError("device='cpu' is not available in the B200 runtime (no CPU fallback)")
            spec = load_model_spec(pipeline_kwargs["rdf_path"])
            loaded = _LoadedModel(spec, torch.device("cuda", torch.cuda.current_device()),
                                  ck["default_blocksize_parameter"])
            self._pipelines[cache_key] = loaded
            while len(self._pipelines) > self._cache_size:  # LRU like serve.multiplexed
                self._pipelines.popitem(last=False)
            return loaded
        finally:
            self._kwargs_cache.pop(cache_key, None)

    @staticmethod
    def _to_ncsp(x: np.ndarray, axes: str) -> np.ndarray:
        """Bring an input array to (N, C, *spatial), guessing the axes the way `create_sample_for_model` does when the
        array has fewer dimensions than the model declares (missing leading batch / channel axes are added)."""
        want = len(axes)
        if x.ndim > want:
            raise ValueError(f"input has {x.ndim} dimensions, model expects at most {want} ({axes})")
        while x.ndim < want:
            x = x[None]
        order = [axes.index(a) for a in "bc" if a in axes] + [axes.index(a) for a in "zyx" if a in axes]
        # element types the device converts (`ensure_dtype(float32)` there) stay as they are: fewer bytes over PCIe
        dtype = x.dtype if x.dtype.name in K.NUMPY_DEVICE_DTYPES else np.float32
        return np.ascontiguousarray(np.transpose(x, order), dtype=dtype)

    def _run_ops_unfused(self, x: torch.Tensor, ops: List[Op], n: int, c: int,
                         ref_norms: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
        for op in ops:
            k = op.kwargs
            if op.name == "ensure_dtype":
                continue
            if op.name == "scale_mean_variance":
                # statistics of the reference tensor were taken before anything touched it (see _reference_norms)
                if ref_norms is None or id(op) not in ref_norms:
                    raise ValueError(f"scale_mean_variance: reference tensor {k.get('reference_tensor')!r} is not an input")
                n_seg = self._segments_of(op, n, c)
                if ref_norms[id(op)].shape[0] != n_seg:
                    raise ValueError("scale_mean_variance: the reference tensor has a different number of channels")
                x = K.scale_mean_variance(x, ref_norms[id(op)], n_seg, float(k.get("eps", 1e-6)), out=x)
                continue
            if op.name in _NORMALISERS:
                x = K.normalize(x, self._norm_params(x, op, n, c)[0], out=x)
            elif op.name == "clip":
                x = K.clip(x, float(k["min"]), float(k["max"]), out=x)
            elif op.name == "scale_linear":
                gain, offset = k.get("gain", 1.0), k.get("offset", 0.0)
                if isinstance(gain, (list, tuple)) or isinstance(offset, (list, tuple)):   # one value per channel
                    self._check_per_channel(op, (gain, offset), c)
                    x = K.scale_linear_per_channel(x, gain, offset, out=x)
                else:
                    x = K.scale_linear(x, float(gain), float(offset), out=x)
            elif op.name == "sigmoid":
                x = K.sigmoid(x, out=x)
            elif op.name == "binarize":
                thr = k["threshold"]
                if isinstance(thr, (list, tuple)):   # x > t[c]  ⇔  x - t[c] > 0 (a float subtraction keeps the sign exactly)
                    self._check_per_channel(op, (thr,), c)
                    x = K.to_float32(K.binarize(K.scale_linear_per_channel(x, 1.0, [-float(t) for t in thr]), 0.0))
                else:
                    x = K.to_float32(K.binarize(x, float(thr)))
            else:
                raise ValueError(f"unsupported processing op {op.name!r}")
        return x

    @staticmethod
    def _check_per_channel(op: Op, values, c: int) -> None:
        axis = op.kwargs.get("axis", "channel")
        if axis not in ("c", "channel"):
            raise ValueError(f"{op.name}: parameters along axis {axis!r} are not supported (channel only)")
        for v in values:
            if isinstance(v, (list, tuple)) and len(v) != c:
                raise ValueError(f"{op.name}: {len(v)} values for {c} channel(s)")

    @staticmethod
    def _segments_of(op: Op, n: int, c: int) -> int:
        """How many independent segments the statistics of `op` have: `axes` lists the REDUCED axes (Appendix A.2)."""
        axes = op.kwargs.get("axes")
        if axes is not None and any(a in ("b", "batch") for a in axes):
            raise ValueError("statistics over the batch axis are not supported")
        joint_c = axes is None or any(a in ("c", "channel") for a in axes)
        return n if joint_c else n * c

    def _reference_norms(self, ops: List[Op], tensors: Dict[str, torch.Tensor]) -> Dict[int, torch.Tensor]:
        """{id(op): {ref_mean, ref_std + eps}} for every scale_mean_variance op whose reference tensor is in `tensors`
        (bioimageio computes input-tensor measures on the sample as given, before any pre-processing)."""
        out = {}
        for op in ops:
            if op.name != "scale_mean_variance":
                continue
            ref = tensors.get(str(op.kwargs.get("reference_tensor")))
            if ref is None:
                continue
            n_seg = self._segments_of(op, ref.shape[0], ref.shape[1])
            out[id(op)] = K.meanvar(ref, n_seg, float(op.kwargs.get("eps", 1e-6)))[1]
        return out

    @staticmethod
    def _norm_params(x: torch.Tensor, op: Op, n: int, c: int):
        """→ (norm [n_seg, 2] fp64 on device, norm_mode for the gather).  `axes` lists the REDUCED axes (Appendix A.2)."""
        k = op.kwargs
        axes = k.get("axes")
        joint_c = axes is None or any(a in ("c", "channel") for a in axes)
        # statistics that include the batch axis (`axes` naming it, or 0.4's `mode: per_dataset`, whose "dataset" inside one
        # predict call is the samples of that call): one set of numbers for all samples, applied to each of them
        over_batch = (axes is not None and any(a in ("b", "batch") for a in axes)) or k.get("mode") == "per_dataset"
        n_seg, mode = (n, 2) if joint_c else (n * c, 1)
        eps = float(k.get("eps", 1e-6))
        if op.name in ("scale_range", "zero_mean_unit_variance"):
            xs, n_stat = x, n_seg
            if over_batch:
                xs, n_stat = (x, 1) if joint_c else (x.transpose(0, 1).contiguous(), c)
            if op.name == "scale_range":
                q = torch.tensor([float(k.get("min_percentile", 0.0)) / 100.0, float(k.get("max_percentile", 100.0)) / 100.0],
                                 dtype=torch.float64, device=x.device)
                norm = K.quantile(xs, n_stat, q, eps)[1]
            else:
                norm = K.meanvar(xs, n_stat, eps)[1]
            return (norm.repeat(n, 1).contiguous() if over_batch else norm), mode
        mean, std = k["mean"], k["std"]
        if isinstance(mean, (list, tuple)):
            vals = [[float(m), float(s) + eps] for m, s in zip(mean, std)] * n
            return torch.tensor(vals, dtype=torch.float64, device=x.device), 1
        return torch.tensor([[float(mean), float(std) + eps]] * n, dtype=torch.float64, device=x.device), 2

    @staticmethod
    def _validate_sample(x: np.ndarray, inp, blocking: bool) -> None:
        """What `create_sample_for_model` checks against the model description (ref: runtime_deployment.py:270-274): the
        channel count the RDF declares, and — without blocking — that every parameterised spatial axis has a size the
        model accepts (min + k * step)."""
        sp_axes = [a for a in inp.axes if a in "zyx"]
        if inp.n_channels is not None and x.shape[1] != inp.n_channels:
            raise ValueError(f"input {inp.id!r} has {x.shape[1]} channel(s), the model expects {inp.n_channels}")
        for a, size in zip(sp_axes, x.shape[2:]):
            lo, step = inp.size_min.get(a), inp.size_step.get(a, 0)
            if lo is None:
                continue
            if size < lo and not blocking:
                raise ValueError(f"input {inp.id!r}: axis {a} has size {size}, the model needs at least {lo}")
            if not blocking and step > 0 and (size - lo) % step:
                raise ValueError(f"input {inp.id!r}: axis {a} has size {size}, the model accepts {lo} + k*{step}")
            if not blocking and step == 0 and size != lo:
                raise ValueError(f"input {inp.id!r}: axis {a} has size {size}, the model needs exactly {lo}")

    def _predict_sample(self, loaded: _LoadedModel, arrays: Dict[str, np.ndarray], blocking: bool) -> Dict[str, np.ndarray]:
        """One sample = one array per declared input tensor (ref: `create_sample_for_model`, runtime_deployment.py:270-274)
        → `{output_id: ndarray}` for every declared output.  All inputs share the batch size and the spatial shape (they
        are tiled with one geometry); every input has its own pre-processing, every output its own post-processing."""
        spec = loaded.spec
        inp0 = spec.inputs[0]
        xs = []
        for inp in spec.inputs:
            x_host = self._to_ncsp(np.asarray(arrays[inp.id]), inp.axes)
            if x_host.size == 0:
                raise ValueError("empty input")
            self._validate_sample(x_host, inp, blocking)
            xs.append(K.to_float32(torch.from_numpy(x_host).to(loaded.device, non_blocking=True)))
        x = xs[0]
        N = x.shape[0]
        spatial = tuple(x.shape[2:])
        for inp, xi in zip(spec.inputs[1:], xs[1:]):
            if xi.shape[0] != N or tuple(xi.shape[2:]) != spatial:
                raise ValueError(f"input {inp.id!r} has batch/spatial shape {(xi.shape[0],) + tuple(xi.shape[2:])}, "
                                 f"input {inp0.id!r} has {(N,) + spatial}: inputs are tiled together and must agree")
        sp_axes = [a for a in inp0.axes if a in "zyx"]
        ratios = [self._output_ratio(spec, o, sp_axes) for o in spec.outputs]      # output pixels per input pixel, per axis
        halos = []
        for o, r in zip(spec.outputs, ratios):   # an output's halo is given in its own pixels: bring it to input pixels
            h = [Fraction(o.halo.get(a, 0)) / ra for a, ra in zip(sp_axes, r)]
            if any(v.denominator != 1 for v in h):
                raise ValueError(f"output {o.id!r}: halo {o.halo} is not a whole number of input pixels at scale {[float(x) for x in r]}")
            halos.append(tuple(int(v) for v in h))
        if len(set(halos)) != 1:
            raise ValueError(f"outputs declare different halos {halos}; one block geometry serves all outputs")
        if blocking:
            block = tuple(inp0.size_min.get(a, s) + loaded.n * inp0.size_step.get(a, 0) for a, s in zip(sp_axes, spatial))
            halo = halos[0]
            if any(s < b - 2 * h for s, b, h in zip(spatial, block, halo)):
                raise ValueError(f"sample {spatial} is smaller than the block {block} minus halo")
        else:
            block, halo = spatial, (0,) * len(spatial)

        all_ops = [o for t in spec.inputs + spec.outputs for o in t.ops]
        ref_norms = self._reference_norms(all_ops, {str(i.id): xi for i, xi in zip(spec.inputs, xs)})  # before normalising
        key = (tuple(x.shape), block, halo)
        pipe = loaded.pipelines.get(key)
        if pipe is None:
            pipe = TiledPipeline(tuple(x.shape), PipelineSpec(block=block, halo=halo, convention="bioimageio",
                                                              preprocess="none", label=False), device=loaded.device)
            loaded.pipelines[key] = pipe
        # whole-sample pre-processing first (statistics are per sample, never per block), fused into the gather
        tiles = []
        for j, (inp, xi) in enumerate(zip(spec.inputs, xs)):
            pre = [o for o in inp.ops if o.name != "ensure_dtype"]
            out_buf = pipe.tiles if j == 0 and xi.shape[1] == x.shape[1] else None
            if len(pre) == 1 and pre[0].name in _NORMALISERS:
                norm, mode = self._norm_params(xi, pre[0], N, xi.shape[1])
                tiles.append(K.tile_gather(xi, pipe.axes, norm=norm, norm_mode=mode, out=out_buf))
            else:
                tiles.append(K.tile_gather(self._run_ops_unfused(xi, pre, N, xi.shape[1], ref_norms), pipe.axes, out=out_buf))
        if len(tiles) == 1:
            raw = pipe.forward(loaded.net, tiles[0])
        else:
            fb = max(1, int(pipe.spec.forward_batch))
            raw = self._gather_outputs([loaded.net(*[t[i:i + fb] for t in tiles]) for i in range(0, tiles[0].shape[0], fb)])
        logits_by_output = self._outputs_in_order(raw, spec)
        results = {}
        for o, lg, r in zip(spec.outputs, logits_by_output, ratios):
            axes, out_spatial = self._output_axes(pipe, r, spatial, block, halo, tuple(lg.shape[2:]), o.id)
            results[str(o.id)] = self._finish_output(axes, lg.contiguous(), o, N, out_spatial, ref_norms)
        return results

    @staticmethod
    def _output_ratio(spec: ModelSpec, outp, sp_axes) -> List[Fraction]:
        """Output pixels per input pixel along every spatial axis: bioimageio's `SizeReference` gives
        size_out = size_ref * scale_ref / scale_out + offset (0.5; `scale` = physical size of a pixel) or
        size_out = size_ref * scale + 2 * offset (0.4 implicit output shape).  Offsets other than 0 (valid convolutions that
        shrink the output) are not supported: such models declare a halo instead."""
        by_id = {t.id: t for t in spec.inputs}
        out = []
        for a in sp_axes:
            ref = outp.size_ref.get(a)
            r = Fraction(1)
            if ref is not None:
                tensor_id, axis_id, offset = ref
                if offset != 0:
                    raise ValueError(f"output {outp.id!r}: size offset {offset} on axis {a} is not supported")
                if a in outp.ref_scale:
                    r = Fraction(outp.ref_scale[a]).limit_denominator(64)
                else:
                    src = by_id.get(tensor_id)
                    if src is None:
                        raise ValueError(f"output {outp.id!r} takes its size from {tensor_id!r}, which is not an input")
                    r = (Fraction(src.scale.get(axis_id, 1.0)).limit_denominator(64)
                         / Fraction(outp.scale.get(a, 1.0)).limit_denominator(64))
            elif a in outp.scale:
                r = 1 / Fraction(outp.scale[a]).limit_denominator(64)
            if r <= 0:
                raise ValueError(f"output {outp.id!r}: scale {float(r)} on axis {a}")
            out.append(r)
        return out

    @staticmethod
    def _output_axes(pipe: TiledPipeline, ratio, spatial, block, halo, tile_spatial, out_id):
        """The stitcher's per-axis tables in OUTPUT pixels.  At scale 1 they are the gather's; otherwise the same blocks
        scaled: block, halo, inner region and sample size times the ratio (all must stay whole numbers), cached per output."""
        if all(r == 1 for r in ratio):
            if tuple(tile_spatial) != tuple(block):
                raise ValueError(f"output {out_id!r}: the network returned tiles of {tuple(tile_spatial)} for blocks of {tuple(block)}")
            return pipe.axes, tuple(spatial)
        cache = pipe.__dict__.setdefault("_scaled_axes", {})
        key = (out_id, tuple(ratio))
        if key not in cache:
            from . import geometry as G

            dims = []
            for s_, b_, h_, r in zip(spatial, block, halo, ratio):
                vals = [Fraction(s_) * r, Fraction(b_) * r, Fraction(h_) * r]
                if any(v.denominator != 1 for v in vals):
                    raise ValueError(f"output {out_id!r}: sample {s_} / block {b_} / halo {h_} do not scale by {float(r)} to whole pixels")
                dims.append(tuple(int(v) for v in vals))
            axes = [G.bioimageio_axis(s_, b_, h_) for s_, b_, h_ in dims]
            if len(axes) == 2:
                axes = [G.identity_axis(1)] + axes
            cache[key] = ([K.DeviceAxis(a, pipe.device) for a in axes], tuple(d[0] for d in dims), tuple(d[1] for d in dims))
        axes, out_spatial, out_block = cache[key]
        if tuple(tile_spatial) != out_block:
            raise ValueError(f"output {out_id!r}: the network returned tiles of {tuple(tile_spatial)}, the description implies {out_block}")
        return axes, out_spatial

    @staticmethod
    def _gather_outputs(chunks):
        """Per-chunk network results (tensor, tuple / list or dict of tensors) → the same structure over all tiles."""
        first = chunks[0]
        if isinstance(first, torch.Tensor):
            return torch.cat(chunks)
        if isinstance(first, dict):
            return {k: torch.cat([c[k] for c in chunks]) for k in first}
        return [torch.cat([c[j] for c in chunks]) for j in range(len(first))]

    @staticmethod
    def _outputs_in_order(raw, spec: ModelSpec) -> List[torch.Tensor]:
        if isinstance(raw, torch.Tensor):
            raw = [raw]
        if isinstance(raw, dict):
            missing = [o.id for o in spec.outputs if o.id not in raw]
            if missing:
                raise ValueError(f"the network returned no tensor named {missing}")
            raw = [raw[o.id] for o in spec.outputs]
        if len(raw) != len(spec.outputs):
            raise ValueError(f"the network returned {len(raw)} tensor(s), the model declares {len(spec.outputs)} output(s)")
        return list(raw)

    def _finish_output(self, axes, logits: torch.Tensor, outp, N: int, spatial, ref_norms) -> np.ndarray:
        post = [o for o in outp.ops if o.name != "ensure_dtype"]
        out_shape = (N, logits.shape[1]) + tuple(spatial)
        names = [o.name for o in post]
        if names in (["sigmoid"], ["binarize"], ["sigmoid", "binarize"], []):
            epi = _lib.BE_EPI_SIGMOID if "sigmoid" in names else _lib.BE_EPI_NONE
            if "binarize" in names:
                thr = float(post[-1].kwargs["threshold"])
                _, mask = K.stitch_crop(logits, out_shape, axes, epi, thr, want_out=False, want_mask=True)
                result = mask
            else:
                result, _ = K.stitch_crop(logits, out_shape, axes, epi)
        else:
            stitched, _ = K.stitch_crop(logits, out_shape, axes)
            result = self._run_ops_unfused(stitched, post, N, logits.shape[1], ref_norms)
        want_dtype = getattr(torch, outp.dtype, torch.float32)
        if result.dtype == want_dtype:
            out = result
        elif want_dtype == torch.float32:
            out = K.to_float32(result.contiguous())     # trailing ensure_dtype(float32), on the device
        else:
            out = result.to(want_dtype)
        arr = out.cpu().numpy()
        # back to the model's declared output axis order
        have = [a for a in "bc" if a in outp.axes] + [a for a in "zyx" if a in outp.axes]
        if "".join(have) != outp.axes and len(have) == arr.ndim:
            arr = np.transpose(arr, [have.index(a) for a in outp.axes])
        return arr

    @staticmethod
    def _sample_arrays(inputs, spec: ModelSpec) -> Dict[str, np.ndarray]:
        """`create_sample_for_model`'s input handling: one array for a one-input model, else `{input_id: array}` naming
        every declared input and nothing else."""
        ids = [i.id for i in spec.inputs]
        if not isinstance(inputs, dict):
            if len(ids) != 1:
                raise ValueError(f"the model declares {len(ids)} input tensors {ids}; pass a dict with one array each")
            return {ids[0]: inputs}
        if len(ids) == 1 and len(inputs) == 1 and ids[0] not in inputs:
            return {ids[0]: next(iter(inputs.values()))}   # a single unnamed-by-id array for a single input
        unknown = [k for k in inputs if k not in ids]
        missing = [k for k in ids if k not in inputs]
        if missing:
            raise ValueError(f"inputs must contain tensor(s) {missing}")
        if unknown:
            raise ValueError(f"unknown input tensor(s) {unknown}; the model declares {ids}")
        return {k: inputs[k] for k in ids}

    async def predict(
        self,
        rdf_path: str,
        inputs: Union[np.ndarray, Dict[str, np.ndarray]],
        weights_format: Optional[str] = None,
        device: Literal["cuda", "cpu"] = None,
        default_blocksize_parameter: Optional[int] = None,
        sample_id: str = "sample",
        latest_remote_modified: Optional[float] = None,
    ) -> Dict[str, np.ndarray]:
        """Run inference on model using the B200 tiled pipeline (signature of ref: runtime_deployment.py:234-243)."""
        try:
            if not Path(rdf_path).exists():
                raise FileNotFoundError(f"RDF not found: {rdf_path}")
            cache_key = self._set_prediction_kwargs(rdf_path=rdf_path, weights_format=weights_format, device=device,
                                                    default_blocksize_parameter=default_blocksize_parameter,
                                                    latest_remote_modified=latest_remote_modified)
            loaded = await self._create_prediction_pipeline(cache_key)
            arrays = self._sample_arrays(inputs, loaded.spec)
            if any(np.asarray(v).size == 0 for v in arrays.values()):
                raise ValueError("empty input")
            logger.info(f"Running inference for sample '{sample_id}'")
            return self._predict_sample(loaded, arrays, blocking=bool(default_blocksize_parameter))
        except Exception as e:
            torch.cuda.empty_cache()
            err_type = type(e).__name__
            if err_type in ("OutOfMemoryError", "CUDAOutOfMemoryError") or ("out of memory" in str(e).lower()):
                oom_msg = f"CUDA out of memory during inference: {str(e)}"
                logger.error(oom_msg)
                raise RuntimeError(oom_msg) from None
            logger.error(f"Prediction failed: {str(e)}")
            raise
