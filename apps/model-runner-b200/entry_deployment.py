"""`entry_deployment:EntryDeployment` of the model-runner-b200 app: the data-path subset of the reference's entry class
(ref: apps/model-runner/entry_deployment.py:1869-1981).  The reference's own EntryDeployment (model cache, S3 up/download,
search, validation) can be used unchanged instead — it only calls `runtime_deployment.predict.remote(...)`."""
from bioengine_worker_b200.entry_deployment import EntryDeployment  # noqa: F401
