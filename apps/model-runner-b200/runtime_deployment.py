"""`runtime_deployment:RuntimeDeployment` of the model-runner-b200 app.

The builder `exec`s this file (ref: bioengine/apps/builder.py:1210); the class comes from the library package, already
wrapped in `serve.deployment(...)` with the reference's options when Ray Serve is importable
(ref: apps/model-runner/runtime_deployment.py:31-55).  Ray's `runtime_env` for it needs only torch + numpy + pyyaml and
the built `libbe_b200.so` on the replica's node.
"""
from bioengine_worker_b200.runtime_deployment import RuntimeDeployment  # noqa: F401
