"""CPU oracle for the b200fit hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

This package is a plain numpy / fp64 restatement of the arithmetic that the
reference (Gjerman/Orange3-Spark) reaches when its widgets call

    VectorAssembler(inputCols=..., outputCol='features').transform(df)
        reference: orangecontrib/spark/widgets/ml/spark_ml_dataset.py:575-576
    method_instance.fit(self.in_df, params=paramMap)
        reference: orangecontrib/spark/base/spark_ml_estimator.py:22
    model_instance.transform(self.in_df)
        reference: orangecontrib/spark/widgets/ml/spark_ml_model.py:53

The reference itself contains none of that arithmetic: it lives in a
third-party dependency that is absent from /root/reference and is NOT pinned
(setup.py:66-76 lists only Orange3, pandas, py4j, sqlparse; README.md:41-42
says "you need to have Apache Spark installed").  The dependency is Apache
Spark MLlib (org.apache.spark.ml.clustering.KMeans,
org.apache.spark.ml.classification.LogisticRegression,
org.apache.spark.ml.feature.{VectorAssembler,StandardScaler}) plus scalanlp
Breeze (breeze.optimize.{LBFGS,OWLQN,StrongWolfeLineSearch,
BacktrackingLineSearch}).  What is restated here is their published algorithm
as summarised in SURVEY.md Appendix A (A.1 .. A.6).

PARITY UNPINNED.  The reference ships no test, golden vector or fixture for
this path and neither it nor Spark can run in this
environment (no JVM, no pyspark).  The only pins are three known-answer values
recalled from upstream pyspark doctests, reproduced in
tests/test_oracle_kat.py, and scikit-learn cross-checks at the optimum.

Who may import this package: tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product (orange3-spark_b200/) never
imports it and has no CPU fallback.
"""

from .assemble import assemble, standard_scaler_fit, standard_scaler_transform, summarize  # noqa: F401
from .kmeans import (  # noqa: F401
    fast_squared_distance,
    find_closest_scalar,
    kmeans_assign,
    kmeans_lloyd,
    kmeans_step,
)
from .logreg import (  # noqa: F401
    lr_binary_lossgrad,
    lr_fit,
    lr_multinomial_lossgrad,
    lr_predict,
)
from .optim import LBFGS, OWLQN  # noqa: F401
from .evaluate import area_under_pr, area_under_roc, multiclass_metrics, silhouette_sqeuclid  # noqa: F401
