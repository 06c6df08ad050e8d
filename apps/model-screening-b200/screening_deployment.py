"""`screening_deployment:ScreeningWorker` of the model-screening-b200 app: one replica per GPU."""
from bioengine_worker_b200.screening_deployment import ScreeningWorker  # noqa: F401
