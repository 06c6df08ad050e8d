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
    params = inspect.signature(first.__init__).parameters
    for stem, _ in classes[1:]:
        assert stem in params, f"{first.__name__}.__init__ must take the {stem} handle"


def test_runtime_predict_signature_matches_the_reference():
    ns = runpy.run_path(str(APPS / "model-runner-b200" / "runtime_deployment.py"))
    cls = getattr(ns["RuntimeDeployment"], "func_or_class", ns["RuntimeDeployment"])
    names = list(inspect.signature(cls.predict).parameters)
    # ref: apps/model-runner/runtime_deployment.py:234-243
    assert names == ["self", "rdf_path", "inputs", "weights_format", "device", "default_blocksize_parameter", "sample_id",
                     "latest_remote_modified"]
