"""`analysis_deployment:MitoAnalysisDeployment` of the fibsem-mito-analysis-b200 app
(ref: apps/fibsem-mito-analysis/analysis_deployment.py:41-47,192-286)."""
from bioengine_worker_b200.analysis_deployment import MitoAnalysisDeployment  # noqa: F401
