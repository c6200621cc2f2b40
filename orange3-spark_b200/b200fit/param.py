"""pyspark.ml-shaped parameter plumbing (drop-in boundary).

What the reference needs from an estimator class (all evidence is reference code):
  * no-arg constructible, every ctor kwarg has a default, `inspect.signature(cls)`
    gives (name -> default) in display order   orangecontrib/spark/utils/ml_api_utils.py:24-28
  * instance has `.params`: objects with `.name` and `.doc`      ml_api_utils.py:38-42
  * `fit(dataset, params=paramMap)` where the keys are FRESHLY SYNTHESISED
    `pyspark.ml.param.Param(instance, name, '')` objects, so they must be resolved by
    name, never by identity              orangecontrib/spark/base/spark_ml_transformer.py:117-123
  * values arrive as None / bool / int / float / str after the GUI's string coercion
    (ints for float params, None for empty boxes)   orangecontrib/spark/utils/gui_utils.py:78-94
"""
import itertools

_uid_counter = itertools.count()


class Param:
    """Same attribute surface as pyspark.ml.param.Param: parent (uid), name, doc."""

    def __init__(self, parent, name, doc, typeConverter=None):
        self.parent = parent.uid if hasattr(parent, "uid") else str(parent)
        self.name = str(name)
        self.doc = str(doc)
        self.typeConverter = typeConverter or (lambda v: v)

    def __repr__(self):
        return "Param(parent=%r, name=%r, doc=%r)" % (self.parent, self.name, self.doc)

    def __str__(self):
        return self.parent + "__" + self.name

    def __hash__(self):
        return hash(str(self))

    def __eq__(self, other):
        return isinstance(other, Param) and self.parent == other.parent and self.name == other.name


class TypeConverters:
    """Accept what the widget's string coercion produces: ints for float params, etc."""

    @staticmethod
    def toInt(v):
        if isinstance(v, bool):
            raise TypeError("expected an int, got bool")
        if isinstance(v, float) and v != int(v):
            raise TypeError("expected an int, got %r" % v)
        return int(v)

    @staticmethod
    def toFloat(v):
        if isinstance(v, bool):
            raise TypeError("expected a float, got bool")
        return float(v)

    @staticmethod
    def toString(v):
        return str(v)

    @staticmethod
    def toBoolean(v):
        if isinstance(v, bool):
            return v
        if isinstance(v, str) and v in ("True", "False", "true", "false"):
            return v.lower() == "true"
        raise TypeError("expected a bool, got %r" % (v,))

    @staticmethod
    def toListString(v):
        if isinstance(v, str):
            return [s.strip() for s in v.strip("[]() ").split(",") if s.strip()]
        return [str(s) for s in v]

    @staticmethod
    def identity(v):
        return v


class Params:
    """Holds declared Params, their defaults and explicitly-set values."""

    _declared = ()  # tuple of (name, doc, converter) filled by subclasses

    def __init__(self):
        self.uid = "%s_%012x" % (type(self).__name__, next(_uid_counter))
        self._params = []
        self._defaultParamMap = {}
        self._paramMap = {}
        for name, doc, conv in self._declared:
            p = Param(self, name, doc, conv)
            self._params.append(p)
            setattr(self, name, p)

    @property
    def params(self):
        return list(self._params)

    def _param_by_name(self, name):
        for p in self._params:
            if p.name == name:
                return p
        raise ValueError("%s has no param named %r" % (type(self).__name__, name))

    def _resolve(self, key):
        """Param object of ours, a foreign Param-like object (anything with .name), or a str."""
        if isinstance(key, str):
            return self._param_by_name(key)
        name = getattr(key, "name", None)
        if name is None:
            raise TypeError("cannot resolve param key %r" % (key,))
        return self._param_by_name(name)

    def _setDefault(self, **kw):
        for k, v in kw.items():
            self._defaultParamMap[k] = v
        return self

    def _set(self, **kw):
        for k, v in kw.items():
            p = self._param_by_name(k)
            if v is None:  # the widget sends None for an empty box: leave unset -> default applies
                self._paramMap.pop(k, None)
                continue
            try:
                self._paramMap[k] = p.typeConverter(v)
            except (TypeError, ValueError) as e:
                raise TypeError("Invalid param value given for param %r: %s" % (k, e))
        return self

    def isSet(self, param):
        return self._resolve(param).name in self._paramMap

    def hasDefault(self, param):
        return self._resolve(param).name in self._defaultParamMap

    def isDefined(self, param):
        return self.isSet(param) or self.hasDefault(param)

    def hasParam(self, name):
        return any(p.name == name for p in self._params)

    def getParam(self, name):
        return self._param_by_name(name)

    def getOrDefault(self, param):
        name = self._resolve(param).name
        if name in self._paramMap:
            return self._paramMap[name]
        return self._defaultParamMap.get(name)

    def extractParamMap(self, extra=None):
        m = {p: self.getOrDefault(p) for p in self._params if self.isDefined(p)}
        if extra:
            for k, v in extra.items():
                m[self._resolve(k)] = v
        return m

    def explainParam(self, param):
        p = self._resolve(param)
        return "%s: %s (default: %r, current: %r)" % (p.name, p.doc, self._defaultParamMap.get(p.name),
                                                      self._paramMap.get(p.name, "undefined"))

    def explainParams(self):
        return "\n".join(self.explainParam(p) for p in self._params)

    def copy(self, extra=None):
        """New instance (fresh uid kept equal, as pyspark does) with `extra` applied by NAME."""
        import copy as _copy

        that = _copy.copy(self)
        that._paramMap = dict(self._paramMap)
        that._defaultParamMap = dict(self._defaultParamMap)
        if extra:
            for k, v in extra.items():
                that._set(**{self._resolve(k).name: v})
        return that


def make_accessors(cls):
    """Adds pyspark-style setX / getX for every declared param."""
    for name, _doc, _conv in cls._declared:
        cap = name[0].upper() + name[1:]

        def getter(self, _n=name):
            return self.getOrDefault(_n)

        def setter(self, value, _n=name):
            return self._set(**{_n: value})

        if not hasattr(cls, "get" + cap):
            setattr(cls, "get" + cap, getter)
        if not hasattr(cls, "set" + cap):
            setattr(cls, "set" + cap, setter)
    return cls
