"""Headless stand-in for the reference's widget plumbing (no Qt, no Orange, no pyspark).

It performs, with plain Python objects, the same steps the GUI performs around the hot path:
  * list estimator classes of a module          reference: orangecontrib/spark/utils/ml_api_utils.py:70-72
  * read ctor defaults + .params docs           reference: orangecontrib/spark/utils/ml_api_utils.py:7-50
  * show every default as TEXT, coerce the text the user leaves/edits back to a typed value
                                                reference: orangecontrib/spark/utils/gui_utils.py:78-94
  * key the values by a freshly made Param(instance, name, '')
                                                reference: orangecontrib/spark/base/spark_ml_transformer.py:117-123
  * call instance.fit(df, params=paramMap)      reference: orangecontrib/spark/base/spark_ml_estimator.py:19-25
Written from the behaviour described in SURVEY.md 3.3 / 8(b); not a copy of the GUI code.
"""
import inspect
from collections import OrderedDict


class ForeignParam:
    """Plays the role of pyspark.ml.param.Param(parent, name, doc): a DIFFERENT class from
    b200fit.param.Param, so the estimator must resolve it by name."""

    def __init__(self, parent, name, doc):
        self.parent = getattr(parent, "uid", str(parent))
        self.name = name
        self.doc = doc

    def __hash__(self):
        return hash((self.parent, self.name))

    def __eq__(self, other):
        return isinstance(other, ForeignParam) and (self.parent, self.name) == (other.parent, other.name)


def list_estimators(module):
    return {n: c for n, c in inspect.getmembers(module, inspect.isclass)
            if "fit" in dir(c) and not inspect.isabstract(c) and not n.startswith("Java")}


def list_transformers(module):
    return {n: c for n, c in inspect.getmembers(module, inspect.isclass)
            if "transform" in dir(c) and not inspect.isabstract(c) and not n.endswith("Model")
            and not n.startswith("Java")}


def describe(cls):
    """(qualified name, doc head, OrderedDict name -> [position, default, doc])."""
    name = str(cls).split("'")[1]
    doc = str(inspect.getdoc(cls)).split(">>>")[0].strip()
    table = OrderedDict()
    for pname, p in inspect.signature(cls).parameters.items():
        table[pname] = [len(table), p.default]
    for p in cls().params:
        if p.name in table:
            table[p.name].append(p.doc)
    return name, doc, table


def text_to_value(text):
    """What a text box yields after the GUI's coercion rules."""
    val = str(text).strip()
    if val in ("None", ""):
        return None
    if val in ("True", "False"):
        return val == "True"
    try:
        as_float = float(val)
    except ValueError:
        return val
    try:
        as_int = int(val)
    except ValueError:
        return as_float
    return as_float if "." in val else as_int


def press_apply(cls, in_df, edits=None, objects=None):
    """Simulates 'Apply' on an Estimator widget: every ctor kwarg appears as the text of its
    default; `edits` maps names to the text a user typed.  `objects` maps names to values that
    cannot be typed into a text box (the `initialModel` extension) and are put in the map as-is."""
    _, _, table = describe(cls)
    inst = cls()
    param_map = {}
    for pname, meta in table.items():
        text = str(meta[1]) if not edits or pname not in edits else str(edits[pname])
        param_map[ForeignParam(inst, pname, "")] = text_to_value(text)
    for pname, obj in (objects or {}).items():
        param_map[ForeignParam(inst, pname, "")] = obj
    return inst.fit(in_df, params=param_map)


def list_evaluators(module):
    """reference: orangecontrib/spark/utils/ml_api_utils.py:59-62"""
    return {n: c for n, c in inspect.getmembers(module, inspect.isclass)
            if "evaluate" in dir(c) and not inspect.isabstract(c)
            and "java_model" not in inspect.getfullargspec(c).args and not n.startswith("Java") and n != "Evaluator"}


def press_apply_evaluator(cls, in_df, edits=None):
    """Simulates 'Apply' on the Evaluation widget (reference: orangecontrib/spark/widgets/ml/spark_ml_evaluation.py:42-55):
    metric names come out of the metricName doc text, one call per metric with metricName overridden in the map."""
    _, _, table = describe(cls)
    metric_names = table["metricName"][2].split("(")[-1].replace(")", "").split("|")
    inst = cls()
    param_map = {}
    for pname, meta in table.items():
        text = str(meta[1]) if not edits or pname not in edits else str(edits[pname])
        param_map[ForeignParam(inst, pname, "")] = text_to_value(text)
    values = {}
    for metric in metric_names:
        param_map[ForeignParam(inst, "metricName", "")] = metric
        values[metric] = inst.apply(in_df, param_map)
    return values


class _FakeSparkSession:
    def createDataFrame(self, pdf):
        return FakeSparkDataFrame(pdf)


class FakeSparkDataFrame:
    """The slice of pyspark.sql.DataFrame the bridge touches: `.columns`, `.toPandas()` (Spark's columnar collect)
    and `.sparkSession.createDataFrame(pandas_frame)` for the way back.  Counts its collects."""

    def __init__(self, pdf):
        self._pdf = pdf
        self.sparkSession = _FakeSparkSession()
        self.collected = 0

    @property
    def columns(self):
        return list(self._pdf.columns)

    def toPandas(self):
        self.collected += 1
        return self._pdf.copy()
