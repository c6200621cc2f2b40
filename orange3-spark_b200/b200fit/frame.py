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
    legal 16-byte / TMA row.  `row0` is the global index of the first row (row shards).

    `col_scale` (optional, fp32[d] on the device) makes this a VIEW of data * col_scale -- what
    StandardScalerModel.transform returns when it only scales (withStd, no centering): consumers
    that can fold the factor into their kernels (the KMeans filter pipeline) read `data` and
    `col_scale`; everybody else calls materialized(), which writes the scaled copy once."""

    def __init__(self, data, d, row0=0, col_scale=None):
        assert data.dim() == 2 and data.is_cuda and data.stride(1) == 1
        self.data = data
        self.n = int(data.shape[0])
        self.ld = int(data.stride(0)) if self.n > 1 else int(data.shape[1])
        self.d = int(d)
        self.row0 = int(row0)
        self.col_scale = col_scale
        self._copy = None

    def __len__(self):
        return self.n

    def materialized(self):
        """A DeviceMatrix without a pending column scale (self when there is none)."""
        if self.col_scale is None:
            return self
        if self._copy is None:
            import torch

            from . import _native as N

            Y = torch.empty_like(self.data)
            N.check(N.lib().b2f_scale_columns(N.ptr(self.data), N.ptr(Y), self.n, self.d, self.ld,
                                              N.ptr(self.col_scale), None, N.stream_ptr()), "b2f_scale_columns")
            self._copy = DeviceMatrix(Y, self.d, self.row0)
        return self._copy

    def rows(self, idx):
        """Rows idx (device int64 tensor) as fp32 [m, d], scaled the way the kernels see them."""
        r = self.data[idx, : self.d]
        return r if self.col_scale is None else r * self.col_scale[None, :]

    def numpy(self):
        return self.materialized().data[:, : self.d].cpu().numpy()


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
    def __init__(self, columns=None, origin=None):
        self._cols = dict(columns or {})
        # (pandas frame the columns were taken from, names taken): lets Model.transform hand the caller's own
        # frame type back with only the NEW columns converted (return_like)
        self._origin = origin
        n = {len(v) for v in self._cols.values()}
        if len(n) > 1:
            raise ValueError("ragged columns: lengths %s" % sorted(n))

    @classmethod
    def from_pandas(cls, pdf):
        cols = {}
        for name in pdf.columns:
            v = pdf[name].to_numpy()
            if v.dtype == object and len(v):
                if hasattr(v[0], "toArray"):  # pyspark.ml.linalg Vector cells (dense or sparse)
                    v = np.stack([x.toArray() for x in v])
                elif isinstance(v[0], (np.ndarray, list, tuple)):  # Arrow list<double> / numpy-array cells
                    v = np.stack([np.asarray(x, dtype=np.float64) for x in v])
            cols[str(name)] = v
        return cls(cols, origin=(pdf, {k: id(v) for k, v in cols.items()}))

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
        return ColumnarFrame(cols, origin=self._origin)

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


def _host_column(v):
    """A frame column as something a pandas DataFrame can hold: 1-D -> numpy, 2-D (vector column) -> list of rows."""
    if isinstance(v, DeviceMatrix):
        return list(v.numpy())
    if hasattr(v, "cpu"):
        v = v.cpu().numpy()
    v = np.asarray(v)
    return list(v) if v.ndim == 2 else v


def return_like(dataset, out):
    """Hand a transform result back in the CALLER's frame type (the reference's Model widget
    does `self.out_df = model.transform(self.in_df)`, orangecontrib/spark/widgets/ml/spark_ml_model.py:53-54, and
    sends it on as a DataFrame):
      ColumnarFrame / dict in           -> ColumnarFrame out (device-resident columns stay on the device);
      pandas DataFrame in               -> pandas DataFrame = the caller's frame + the new columns;
      pyspark-like DataFrame in         -> `session.createDataFrame(pandas frame)` of the same session
        (anything with toPandas() and a sparkSession / sql_ctx exposing createDataFrame; vector-valued new
        columns become pyspark.ml.linalg dense vectors when pyspark is importable, lists otherwise).
    A result that holds a DeviceMatrix (VectorAssembler / StandardScaler output: tens of GB at the BASELINE
    sizes) is never copied back: it stays a ColumnarFrame handle, which offers what the widgets use
    (.columns, [name], withColumn, cache)."""
    if isinstance(dataset, (ColumnarFrame, dict)) or not isinstance(out, ColumnarFrame):
        return out
    if any(isinstance(v, DeviceMatrix) for v in out._cols.values()):
        return out
    if out._origin is None:
        return out
    pdf, taken = out._origin
    res = pdf.copy(deep=False)
    for c in out.columns:  # only columns that are new, or were replaced, since the frame was taken from `pdf`
        if taken.get(c) != id(out._cols[c]):
            res[c] = _host_column(out._cols[c])
    session = getattr(dataset, "sparkSession", None) or getattr(dataset, "sql_ctx", None)
    if hasattr(dataset, "toPandas") and session is not None and hasattr(session, "createDataFrame"):
        try:
            from pyspark.ml.linalg import Vectors  # absent here; the list form below is what gets tested

            for c in res.columns:
                if len(res) and isinstance(res[c].iloc[0], np.ndarray):
                    res[c] = [Vectors.dense(x) for x in res[c]]
        except ImportError:
            pass
        return session.createDataFrame(res)
    return res


def as_frame(dataset):
    """ColumnarFrame as-is; dict of columns; pandas DataFrame; or a pyspark-like DataFrame (anything with
    toPandas(): with Arrow enabled that is Spark's columnar collect), see return_like for the way back."""
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
