"""b200fit -- B200-native fit engine behind Orange3-Spark's clustering / classification widgets.

Drop-in surface: module objects `b200fit.clustering`,
`b200fit.classification`, `b200fit.feature`, `b200fit.evaluation` whose classes satisfy what the reference's
widget bases need from `pyspark.ml.*` (orangecontrib/spark/base/spark_ml_estimator.py:19-25,
orangecontrib/spark/utils/ml_api_utils.py:7-72).  All arithmetic runs in libb200fit.so
(CUDA, sm_100a); there is no CPU fallback.
"""
from . import classification, clustering, evaluation, feature  # noqa: F401
from .frame import ColumnarFrame, DeviceMatrix  # noqa: F401

__version__ = "0.1.0"
