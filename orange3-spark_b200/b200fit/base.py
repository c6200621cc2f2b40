"""Estimator / Transformer / Model protocol consumed by the reference widgets.

  OWSparkEstimator.apply:  method_instance.fit(self.in_df, params=paramMap)
      reference: orangecontrib/spark/base/spark_ml_estimator.py:19-25
  OWSparkTransformer.apply: method_instance.transform(self.in_df, params=paramMap)
      reference: orangecontrib/spark/base/spark_ml_transformer.py:129-139
  OWSparkMLMOdel.transform: model_instance.transform(self.in_df)
      reference: orangecontrib/spark/widgets/ml/spark_ml_model.py:49-54
  output channel typed `pyspark.ml.Model`    spark_ml_estimator.py:15

Calls are synchronous (the widget calls them on the Qt thread), exceptions
propagate as plain Python exceptions, models are immutable after fit.
"""
from .param import Params


class Transformer(Params):
    def _transform(self, dataset):  # pragma: no cover - abstract
        raise NotImplementedError

    def transform(self, dataset, params=None):
        if params:
            return self.copy(params)._transform(dataset)
        return self._transform(dataset)


class Estimator(Params):
    def _fit(self, dataset):  # pragma: no cover - abstract
        raise NotImplementedError

    def fit(self, dataset, params=None):
        if isinstance(params, (list, tuple)):
            return [self.fit(dataset, p) for p in params]
        if params:
            return self.copy(params)._fit(dataset)
        return self._fit(dataset)


class Model(Transformer):
    """Fitted model.  Registered as a virtual subclass of pyspark.ml.Model when pyspark is
    importable so Orange's dynamic channel check isinstance(obj, pyspark.ml.Model) passes."""


def _register_with_pyspark():
    try:  # pyspark is absent in this environment; this branch is untested here
        import pyspark.ml as _pm

        for ours, theirs in ((Model, "Model"), (Estimator, "Estimator"), (Transformer, "Transformer")):
            cls = getattr(_pm, theirs, None)
            if cls is not None and hasattr(cls, "register"):
                cls.register(ours)
    except Exception:
        pass


_register_with_pyspark()
