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
from abc import ABCMeta, abstractmethod

from .param import Params


# The bases are ABCs, as in pyspark: the reference's discovery lists every class of the module with a `fit` /
# `transform` that is NOT abstract (orangecontrib/spark/utils/ml_api_utils.py:53-72, inspect.isabstract), and the
# bases are imported into clustering / classification / feature.
class Transformer(Params, metaclass=ABCMeta):
    @abstractmethod
    def _transform(self, dataset):
        raise NotImplementedError

    def transform(self, dataset, params=None):
        if params:
            return self.copy(params)._transform(dataset)
        return self._transform(dataset)


class Estimator(Params, metaclass=ABCMeta):
    @abstractmethod
    def _fit(self, dataset):
        raise NotImplementedError

    def fit(self, dataset, params=None):
        if isinstance(params, (list, tuple)):
            return [self.fit(dataset, p) for p in params]
        if params:
            return self.copy(params)._fit(dataset)
        return self._fit(dataset)


class Model(Transformer, metaclass=ABCMeta):
    """Fitted model (abstract: `_transform` comes from the concrete model).  Registered as a virtual subclass of
    pyspark.ml.Model when pyspark is importable so Orange's dynamic channel check isinstance(obj, pyspark.ml.Model)
    passes."""

    def transform(self, dataset, params=None):
        from .frame import return_like

        return return_like(dataset, super().transform(dataset, params))


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
