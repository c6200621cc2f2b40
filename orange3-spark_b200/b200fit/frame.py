"""A minimal columnar stand-in for the pyspark.sql.DataFrame the widgets pass around.

The reference only ever uses: `df.columns` (spark_ml_dataset.py:414-430,
spark_ml_transformer.py:104-105), `df[col]` / `withColumn` (spark_ml_dataset.py:577-578),
`.cache()` (spark_ml_transformer.py:134-135) and hands the object to fit / transform.
Columns here are numpy arrays, torch tensors (host or device) or a DeviceMatrix (the
'features' vector column packed by VectorAssembler).
"""
import numpy as np


class DeviceMatrix:
    """Row-major fp32 feature tile X[n, ld] resident in HBM; ld % 4 == 0 so every row is a
    legal 16-byte / TMA row.  `row0` is the global index of the first row (row shards)."""

    def __init__(self, data, d, row0=0):
        assert data.dim() == 2 and data.is_cuda and data.stride(1) == 1
        self.data = data
        self.n = int(data.shape[0])
        self.ld = int(data.stride(0)) if self.n > 1 else int(data.shape[1])
        self.d = int(d)
        self.row0 = int(row0)

    def __len__(self):
        return self.n

    def numpy(self):
        return self.data[:, : self.d].cpu().numpy()


class _Col:
    """`df['name']` handle supporting .cast('double') as used at spark_ml_dataset.py:578."""

    def __init__(self, frame, name):
        self.frame, self.name = frame, name

    def cast(self, kind):
        v = self.frame._cols[self.name]
        if kind not in ("double", "float", "int", "integer", "long", "bigint"):
            raise ValueError("unsupported cast %r" % kind)
        return ("cast", v, kind)


class ColumnarFrame:
    def __init__(self, columns=None):
        self._cols = dict(columns or {})
        n = {len(v) for v in self._cols.values()}
        if len(n) > 1:
            raise ValueError("ragged columns: lengths %s" % sorted(n))

    @classmethod
    def from_pandas(cls, pdf):
        cols = {}
        for name in pdf.columns:
            v = pdf[name].to_numpy()
            if v.dtype == object and len(v) and hasattr(v[0], "toArray"):  # pyspark Vector cells
                v = np.stack([x.toArray() for x in v])
            cols[str(name)] = v
        return cls(cols)

    @property
    def columns(self):
        return list(self._cols.keys())

    def __getitem__(self, name):
        if name not in self._cols:
            raise KeyError("no such column %r; have %s" % (name, self.columns))
        return _Col(self, name)

    def col(self, name):
        if name not in self._cols:
            raise KeyError("no such column %r; have %s" % (name, self.columns))
        return self._cols[name]

    def withColumn(self, name, value):
        cols = dict(self._cols)
        if isinstance(value, tuple) and len(value) == 3 and value[0] == "cast":
            _, v, kind = value
            if kind in ("double", "float"):
                value = _cast_float(v, np.float64 if kind == "double" else np.float32)
            else:
                value = _cast_float(v, np.int64 if kind in ("long", "bigint") else np.int32)
        elif isinstance(value, _Col):
            value = value.frame._cols[value.name]
        cols[name] = value
        return ColumnarFrame(cols)

    def select(self, *names):
        return ColumnarFrame({n: self.col(n) for n in names})

    def count(self):
        for v in self._cols.values():
            return len(v)
        return 0

    def cache(self):
        return self

    def toPandas(self):
        import pandas as pd

        out = {}
        for k, v in self._cols.items():
            if isinstance(v, DeviceMatrix):
                out[k] = list(v.numpy())
            elif hasattr(v, "cpu"):
                out[k] = v.cpu().numpy()
            else:
                out[k] = v
        return pd.DataFrame(out)


def _cast_float(v, dt):
    if hasattr(v, "cpu"):
        import torch

        tmap = {np.float64: torch.float64, np.float32: torch.float32, np.int64: torch.int64, np.int32: torch.int32}
        return v.to(tmap[dt])
    return np.asarray(v).astype(dt)


def as_frame(dataset):
    """ColumnarFrame as-is; dict of columns; pandas DataFrame; or (UNTESTED here: no pyspark)
    a pyspark DataFrame, collected through toPandas()."""
    if isinstance(dataset, ColumnarFrame):
        return dataset
    if isinstance(dataset, dict):
        return ColumnarFrame(dataset)
    try:
        import pandas as pd

        if isinstance(dataset, pd.DataFrame):
            return ColumnarFrame.from_pandas(dataset)
    except ImportError:  # pragma: no cover
        pass
    if hasattr(dataset, "toPandas"):
        return ColumnarFrame.from_pandas(dataset.toPandas())
    raise TypeError("cannot interpret %r as a dataset" % type(dataset))
