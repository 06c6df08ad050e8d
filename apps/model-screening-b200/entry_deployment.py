"""`entry_deployment:ScreeningEntry` of the model-screening-b200 app (composition as ref: apps/model-runner/manifest.yaml:18-20)."""
from bioengine_worker_b200.screening_deployment import ScreeningEntry  # noqa: F401
