"""The app directories a BioEngine worker would deploy: every `file:Class` of a manifest resolves the way the builder
resolves it (exec the file, take the class — ref: bioengine/apps/builder.py:1210,1453-1508), the first class has the
public methods the reference's schema lists, and later deployments are constructor kwargs of the first one."""
import inspect
import runpy
from pathlib import Path

import pytest
import yaml

APPS = Path(__file__).resolve().parents[1] / "apps"


@pytest.mark.parametrize("app,methods", [
    ("model-runner-b200", {"EntryDeployment": ["infer"], "RuntimeDeployment": ["predict", "test"]}),
    ("fibsem-mito-analysis-b200", {"MitoAnalysisDeployment": ["analyze"]}),
    ("model-screening-b200", {"ScreeningEntry": ["screen"], "ScreeningWorker": ["score_shard"]}),
])
def test_manifest_deployments_resolve(app, methods):
    man = yaml.safe_load((APPS / app / "manifest.yaml").read_text())
    assert man["type"] == "ray-serve" and man["deployments"]
    classes = []
    for entry in man["deployments"]:
        stem, cls_name = entry.split(":")
        ns = runpy.run_path(str(APPS / app / f"{stem}.py"))
        cls = ns[cls_name]
        cls = getattr(cls, "func_or_class", cls)      # serve.deployment wraps the class; without Ray it is the class
        classes.append((stem, cls))
        for m in methods[cls_name]:
            assert inspect.iscoroutinefunction(getattr(cls, m)), (cls_name, m)
    first = classes[0][1]
    # ref: bioengine/apps/builder.py:1453-1465 — the entry class must carry schema methods
    schema = [m for m in dir(first) if callable(getattr(first, m)) and hasattr(getattr(first, m), "__schema__")]
    assert set(methods[first.__name__]) <= set(schema), schema
    # ref: builder.py:892-975,1482-1508 — later deployments are injected through __init__ parameters named after the file
    # stem and annotated DeploymentHandle (the annotation object itself is compared); every other parameter has a default
    from bioengine_worker_b200._serve_compat import DeploymentHandle

    params = inspect.signature(first.__init__).parameters
    for stem, _ in classes[1:]:
        assert stem in params, f"{first.__name__}.__init__ must take the {stem} handle"
        assert params[stem].annotation is DeploymentHandle
    injected = {stem for stem, _ in classes[1:]}
    for name, prm in params.items():
        if name not in injected and name != "self":
            assert prm.default is not inspect.Parameter.empty, f"{name} would have to be given in the manifest"
    # lifecycle hooks the builder awaits when present (builder.py:669-890)
    for hook in ("async_init", "test_deployment", "check_health"):
        assert inspect.iscoroutinefunction(getattr(first, hook)), hook


def test_entry_calls_the_runtime_through_a_deployment_handle(tmp_path, monkeypatch):
    """ref: entry_deployment.py:1969-1977 — `await self.runtime_deployment.predict.remote(...)`: a Ray DeploymentHandle has
    no awaitable `predict(...)`, only `predict.remote(...)`."""
    import asyncio

    import numpy as np

    from bioengine_worker_b200.entry_deployment import EntryDeployment

    calls = []

    class _Method:
        async def remote(self, **kw):
            calls.append(kw)
            return {"mask": np.ones((1, 1, 4, 4), np.uint8)}

    class _Handle:
        predict = _Method()

    (tmp_path / "m1").mkdir()
    (tmp_path / "m1" / "rdf.yaml").write_text("name: m1\n")
    cls = getattr(EntryDeployment, "func_or_class", EntryDeployment)
    entry = cls(runtime_deployment=_Handle(), models_dir=str(tmp_path))
    monkeypatch.setattr(entry, "_check_runtime_available", lambda: None)
    asyncio.run(entry.async_init())
    out = asyncio.run(entry.infer("m1", np.zeros((4, 4), np.float32), default_blocksize_parameter=2))
    assert out["mask"].shape == (1, 1, 4, 4)
    assert calls[0]["rdf_path"].endswith("m1/rdf.yaml") and calls[0]["default_blocksize_parameter"] == 2
    url = asyncio.run(entry.infer("m1", np.zeros((4, 4), np.float32), return_download_url=True))["mask"]
    assert url.startswith("file://") and np.load(url[7:]).shape == (1, 1, 4, 4)
    with pytest.raises(FileNotFoundError):
        asyncio.run(entry.infer("missing", np.zeros((4, 4), np.float32)))


def test_runtime_predict_signature_matches_the_reference():
    ns = runpy.run_path(str(APPS / "model-runner-b200" / "runtime_deployment.py"))
    cls = getattr(ns["RuntimeDeployment"], "func_or_class", ns["RuntimeDeployment"])
    names = list(inspect.signature(cls.predict).parameters)
    # ref: apps/model-runner/runtime_deployment.py:234-243
    assert names == ["self", "rdf_path", "inputs", "weights_format", "device", "default_blocksize_parameter", "sample_id",
                     "latest_remote_modified"]


def test_screening_entry_fans_out_and_merges_integer_tables():
    """The entry asks every shard (as `handle.score_shard.remote(...)`), sums the int64 tables and ranks: a fake worker that
    answers from a known table shows the merge is exact and the ranking follows the mean AP at the first threshold."""
    import asyncio

    import numpy as np

    from bioengine_worker_b200.screening_deployment import ScreeningEntry, rank_from_table

    n_models, n_images, thr = 3, 7, [0.5, 0.75]
    rng = np.random.default_rng(0)
    n_true = rng.integers(5, 40, size=n_images)
    full = np.zeros((n_models, n_images, 4), dtype=np.int64)
    for m in range(n_models):
        n_pred = n_true + rng.integers(-3, 4, size=n_images)
        tp50 = np.minimum(n_true, n_pred) - rng.integers(0, 4, size=n_images) - 2 * m     # model 0 is the best
        full[m, :, 0], full[m, :, 1] = np.maximum(tp50, 0), np.maximum(tp50 - 3, 0)
        full[m, :, 2], full[m, :, 3] = n_true, n_pred
    asked = []

    class Method:
        def __init__(self, fn):
            self.remote = fn

    class FakeHandle:   # like a DeploymentHandle: methods are only reachable as `.remote`
        def __init__(self):
            async def score_shard(**kw):
                asked.append((kw["shard"], kw["n_shards"]))
                nm, ni, nt = len(kw["rdf_paths"]), len(kw["images"]), len(kw["thresholds"])
                t = np.zeros((nm, ni, nt + 2), dtype=np.int64)
                if (nm, ni, nt + 2) == full.shape:
                    t[:, kw["shard"]::kw["n_shards"]] = full[:, kw["shard"]::kw["n_shards"]]
                return t
            self.score_shard = Method(score_shard)

    entry = ScreeningEntry(FakeHandle(), n_shards=4)
    res = asyncio.run(entry.screen([f"m{i}/rdf.yaml" for i in range(n_models)], [np.zeros((4, 4))] * n_images,
                                   [np.zeros((4, 4), np.int32)] * n_images, thr))
    assert sorted(asked) == [(s, 4) for s in range(4)]
    assert np.array_equal(res["table"], full)
    want = rank_from_table(full, thr)
    assert np.array_equal(res["ap"], want["ap"]) and res["ranking_index"] == want["ranking"].tolist()
    assert res["ranking"][0] == "m0/rdf.yaml"
    tp, nt, npred = full[0, :, 0], full[0, :, 2], full[0, :, 3]
    np.testing.assert_allclose(res["ap"][0, :, 0], (tp / (nt + npred - tp)).astype(np.float32), rtol=1e-6)
    # fewer images than shards: one shard per image
    asked.clear()
    asyncio.run(entry.screen(["m0/rdf.yaml"], [np.zeros((4, 4))] * 2, [np.zeros((4, 4), np.int32)] * 2))
    assert sorted(asked) == [(0, 2), (1, 2)]
    import pytest as _pytest
    with _pytest.raises(ValueError):
        asyncio.run(entry.screen(["m0/rdf.yaml"], [np.zeros((4, 4))] * 2, [np.zeros((4, 4), np.int32)]))
