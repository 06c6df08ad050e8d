"""What the deployment classes need from Ray Serve and hypha-rpc, with stand-ins when those packages are absent.

The reference's app builder (ref: bioengine/apps/builder.py) `exec`s a deployment file, expects a
`serve.deployment(...)`-wrapped class (`.func_or_class` is the plain class), collects the `__schema__` attribute that
`hypha_rpc.utils.schema.schema_method` puts on public methods (:1453-1465) and checks that the parameters through which
later deployments are injected are annotated `ray.serve.handle.DeploymentHandle` (:1482-1508).  With Ray / hypha-rpc
installed the real objects are used; without them (the build image, plain-Python use, the tests) the stand-ins below keep
the same shape so the classes import and run unchanged.
"""
from __future__ import annotations

import inspect

try:  # pragma: no cover - ray is not installed in the build image
    from ray import serve
    from ray.serve.handle import DeploymentHandle

    HAVE_RAY = True

    def deployment(**options):
        return serve.deployment(**options)

except Exception:  # ray absent: the decorated class stays a plain class that also answers `.func_or_class`
    HAVE_RAY = False

    class DeploymentHandle:  # noqa: D401 - annotation stand-in
        """Stand-in for ray.serve.handle.DeploymentHandle (annotation only)."""

    def deployment(**options):
        def wrap(cls):
            cls.func_or_class = cls
            cls._deployment_options = options
            return cls

        return wrap


try:  # pragma: no cover - hypha-rpc is not installed in the build image
    from hypha_rpc.utils.schema import schema_method
except Exception:

    def schema_method(func=None, **_kwargs):
        """Stand-in: marks the method the way hypha-rpc does (a `__schema__` dict with name, description, parameters)."""

        def mark(f):
            params = [p for p in inspect.signature(f).parameters if p != "self"]
            f.__schema__ = {"name": f.__name__, "description": (f.__doc__ or "").strip(),
                            "parameters": {"type": "object", "properties": {p: {} for p in params}}}
            return f

        return mark(func) if func is not None else mark


async def call_handle(handle, method: str, **kwargs):
    """`await handle.method.remote(**kwargs)` for a Ray DeploymentHandle (ref: entry_deployment.py:1969-1977),
    `await handle.method(**kwargs)` for a plain object (local use, tests)."""
    m = getattr(handle, method)
    if hasattr(m, "remote"):
        return await m.remote(**kwargs)
    return await m(**kwargs)
