"""b200fit.feature -- VectorAssembler and StandardScaler with the pyspark.ml.feature surface.

VectorAssembler is what the Dataset Builder widget runs:
    VA = VectorAssembler(inputCols=attributes, outputCol='features'); VA.transform(in_df)
    reference: orangecontrib/spark/widgets/ml/spark_ml_dataset.py:575-576   
Here it is the columnar loader K1: ordered numeric columns -> device-resident row-major
fp32 tiles (b2f_pack_columns), host columns being streamed in row chunks over two CUDA
streams so the H2D copies overlap the packing kernel.
"""
import numpy as np

from . import _native as N
from .base import Estimator, Model, Transformer
from .dist import Comm
from .frame import ColumnarFrame, DeviceMatrix, as_frame
from .param import TypeConverters as TC
from .param import make_accessors

_NP2CODE = {np.dtype("float32"): N.F32, np.dtype("float64"): N.F64, np.dtype("int32"): N.I32,
            np.dtype("int64"): N.I64, np.dtype("uint8"): N.U8, np.dtype("bool"): N.U8}


def _torch():
    import torch

    return torch


def _as_host_or_device_1d(col):
    """Returns (tensor, dtype_code).  numpy -> torch view (no copy where possible)."""
    torch = _torch()
    if isinstance(col, torch.Tensor):
        t = col
        if t.dtype == torch.bool:
            t = t.view(torch.uint8)
        code = {torch.float32: N.F32, torch.float64: N.F64, torch.int32: N.I32, torch.int64: N.I64,
                torch.uint8: N.U8}.get(t.dtype)
        if code is None:
            t, code = t.to(torch.float64), N.F64
        return t.contiguous(), code
    a = np.asarray(col)
    if a.dtype == object:
        if any(v is None for v in a.ravel()):
            raise ValueError("VectorAssembler: null value in input column (handleInvalid='error')")
        a = a.astype(np.float64)
    if a.dtype not in _NP2CODE:
        a = a.astype(np.float64)
    code = _NP2CODE[a.dtype]
    if a.dtype == np.bool_:
        a = a.view(np.uint8)
    return torch.from_numpy(np.ascontiguousarray(a)), code


def pack_columns(cols, chunk_rows=1 << 22, handle_invalid="error", row0=0):
    """K1: list of 1-D columns (numpy / torch, host or device) -> DeviceMatrix."""
    N.require_cuda()
    torch = _torch()
    L = N.lib()
    flat = []
    for c in cols:
        if isinstance(c, DeviceMatrix):
            c = c.materialized()
            for j in range(c.d):
                flat.append((c.data[:, j].contiguous(), N.F32))
            continue
        if getattr(c, "ndim", 1) == 2:  # vector-valued column: flatten its slots
            arr = c if isinstance(c, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(c))
            for j in range(arr.shape[1]):
                flat.append(_as_host_or_device_1d(arr[:, j]))
            continue
        flat.append(_as_host_or_device_1d(c))
    d = len(flat)
    if d == 0:
        raise ValueError("VectorAssembler: no input columns")
    n = int(flat[0][0].shape[0])
    for t, _ in flat:
        if int(t.shape[0]) != n:
            raise ValueError("VectorAssembler: ragged columns")
    ld = (d + 3) // 4 * 4
    dev = torch.device("cuda", torch.cuda.current_device())
    X = torch.empty((n, ld), dtype=torch.float32, device=dev)
    bad = torch.zeros(1, dtype=torch.int32, device=dev)
    codes = torch.tensor([c for _, c in flat], dtype=torch.int32, device=dev)
    all_dev = all(t.is_cuda for t, _ in flat)
    if n == 0:
        return DeviceMatrix(X, d, row0)
    if all_dev:
        ptrs = torch.tensor([t.data_ptr() for t, _ in flat], dtype=torch.int64, device=dev)
        N.check(L.b2f_pack_columns(N.ptr(ptrs), N.ptr(codes), d, n, N.ptr(X), ld, N.ptr(bad), N.stream_ptr()),
                "b2f_pack_columns")
    else:
        # stream host columns chunk by chunk: copy stream fills staging buffer (i & 1) while the
        # compute stream packs the other one
        esz = [t.element_size() for t, _ in flat]
        chunk = max(1, min(chunk_rows, n))
        stage = [[torch.empty(chunk, dtype=t.dtype, device=dev) for t, _ in flat] for _ in range(2)]
        ptrs = [torch.tensor([s.data_ptr() for s in stage[b]], dtype=torch.int64, device=dev) for b in range(2)]
        main = torch.cuda.current_stream()
        copy_s = torch.cuda.Stream()
        copied = [torch.cuda.Event() for _ in range(2)]
        packed = [torch.cuda.Event() for _ in range(2)]
        copy_s.wait_stream(main)
        i = 0
        for a in range(0, n, chunk):
            b = min(n, a + chunk)
            buf = i & 1
            with torch.cuda.stream(copy_s):
                if i >= 2:
                    copy_s.wait_event(packed[buf])
                for (t, _), s in zip(flat, stage[buf]):
                    src = t[a:b]
                    s[: b - a].copy_(src, non_blocking=True)
                copied[buf].record(copy_s)
            main.wait_event(copied[buf])
            N.check(L.b2f_pack_columns(N.ptr(ptrs[buf]), N.ptr(codes), d, b - a, N.ptr(X[a:]), ld, N.ptr(bad),
                                       N.stream_ptr()), "b2f_pack_columns")
            packed[buf].record(main)
            i += 1
        del esz
    if handle_invalid == "error" and int(bad.item()) != 0:
        raise ValueError("VectorAssembler: encountered NaN while assembling a row with handleInvalid='error'; "
                         "remove the NaNs or use handleInvalid='keep'")
    return DeviceMatrix(X, d, row0)


def to_device_matrix(col, allow_scaled=False):
    """Accepts whatever sits in a features column and returns a DeviceMatrix.  A scaled VIEW (DeviceMatrix with a
    pending col_scale) is handed through only to callers that fold the factor into their kernels."""
    torch = _torch()
    if isinstance(col, DeviceMatrix):
        return col if allow_scaled else col.materialized()
    if isinstance(col, torch.Tensor) and col.is_cuda and col.dim() == 2 and col.dtype == torch.float32 \
            and col.stride(1) == 1 and col.shape[1] % 4 == 0 and col.data_ptr() % 16 == 0 \
            and (col.shape[0] <= 1 or col.stride(0) % 4 == 0):
        return DeviceMatrix(col, col.shape[1])
    if getattr(col, "ndim", None) == 2 or (isinstance(col, torch.Tensor) and col.dim() == 2):
        return pack_columns([col], handle_invalid="keep")
    raise TypeError("features column must be a DeviceMatrix or a 2-D array, got %r" % type(col))


def to_device_vector(col, n_expected=None):
    """1-D label / weight column -> fp32 device tensor."""
    torch = _torch()
    if isinstance(col, torch.Tensor):
        t = col.to(device="cuda", dtype=torch.float32)
    else:
        t = torch.from_numpy(np.ascontiguousarray(np.asarray(col, dtype=np.float32))).cuda()
    if n_expected is not None and t.shape[0] != n_expected:
        raise ValueError("column length %d != number of rows %d" % (t.shape[0], n_expected))
    return t.contiguous()


def column_stats(X, w=None, comm=None):
    """Summarizer (A.3 pass 0 / A.6): weighted mean, UNBIASED weighted std and W, reduced over ranks."""
    torch = _torch()
    comm = comm or Comm()
    L = N.lib()
    d = X.d
    out = torch.zeros(2 * d + 2, dtype=torch.float64, device=X.data.device)
    shift = torch.zeros(d, dtype=torch.float32, device=X.data.device)
    if comm.rank == 0 and X.n > 0:
        shift.copy_(X.data[0, :d])
    comm.broadcast(shift, 0)
    N.check(L.b2f_colstats(N.ptr(X.data), X.n, d, X.ld, N.ptr(w), N.ptr(shift), N.ptr(out), N.stream_ptr()),
            "b2f_colstats")
    comm.all_reduce_sum(out)
    o = out.cpu().numpy()
    sh = shift.cpu().numpy().astype(np.float64)
    s1, s2, W, W2 = o[:d], o[d:2 * d], o[2 * d], o[2 * d + 1]
    mean_s = s1 / W
    denom = W - W2 / W
    var = (s2 - W * mean_s * mean_s) / denom if denom > 0 else np.zeros(d)
    return mean_s + sh, np.sqrt(np.maximum(var, 0.0)), W


@make_accessors
class VectorAssembler(Transformer):
    """A feature transformer that merges multiple columns into a vector column.

    B200-native: the output column is a device-resident row-major fp32 matrix.
    """

    _declared = (
        ("inputCols", "input column names.", TC.toListString),
        ("outputCol", "output column name.", TC.toString),
        ("handleInvalid", "How to handle invalid data (NaN). Options are 'error' (throw an error) or 'keep' "
                          "(return the NaN in the output).", TC.toString),
    )

    def __init__(self, inputCols=None, outputCol=None, handleInvalid="error"):
        super().__init__()
        self._setDefault(handleInvalid="error")
        self._set(inputCols=inputCols, outputCol=outputCol, handleInvalid=handleInvalid)

    def setParams(self, inputCols=None, outputCol=None, handleInvalid="error"):
        return self._set(inputCols=inputCols, outputCol=outputCol, handleInvalid=handleInvalid)

    def _transform(self, dataset):
        frame = as_frame(dataset)
        names = self.getOrDefault("inputCols")
        if not names:
            raise ValueError("VectorAssembler: inputCols is not set")
        out = self.getOrDefault("outputCol") or (self.uid + "__output")
        hi = self.getOrDefault("handleInvalid")
        if hi not in ("error", "keep"):
            raise ValueError("VectorAssembler: handleInvalid must be 'error' or 'keep' ('skip' is not supported)")
        X = pack_columns([frame.col(c) for c in names], handle_invalid=hi)
        return frame.withColumn(out, X)


@make_accessors
class StandardScaler(Estimator):
    """Standardizes features by removing the mean and scaling to unit variance using column
    summary statistics on the samples in the training set (unbiased sample std)."""

    _declared = (
        ("withMean", "Center data with mean", TC.toBoolean),
        ("withStd", "Scale to unit standard deviation", TC.toBoolean),
        ("inputCol", "input column name.", TC.toString),
        ("outputCol", "output column name.", TC.toString),
    )

    def __init__(self, withMean=False, withStd=True, inputCol=None, outputCol=None):
        super().__init__()
        self._setDefault(withMean=False, withStd=True)
        self._set(withMean=withMean, withStd=withStd, inputCol=inputCol, outputCol=outputCol)

    def _fit(self, dataset):
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(self.getOrDefault("inputCol")))
        mean, std, _ = column_stats(X)
        m = StandardScalerModel(mean, std)
        m._set(withMean=self.getOrDefault("withMean"), withStd=self.getOrDefault("withStd"),
               inputCol=self.getOrDefault("inputCol"), outputCol=self.getOrDefault("outputCol"))
        return m


@make_accessors
class StandardScalerModel(Model):
    """Model fitted by StandardScaler (A.6): x * (1/std) (0 where std == 0), optional centering."""

    _declared = StandardScaler._declared

    def __init__(self, mean=None, std=None):
        super().__init__()
        self._setDefault(withMean=False, withStd=True)
        self.mean = np.asarray(mean, dtype=np.float64) if mean is not None else None
        self.std = np.asarray(std, dtype=np.float64) if std is not None else None

    def _transform(self, dataset):
        torch = _torch()
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(self.getOrDefault("inputCol")))
        out = self.getOrDefault("outputCol") or (self.uid + "__output")
        scale = np.ones(X.ld)
        shift = np.zeros(X.ld)
        if self.getOrDefault("withStd"):
            scale[: X.d] = np.where(self.std != 0, 1.0 / np.where(self.std != 0, self.std, 1.0), 0.0)
        if self.getOrDefault("withMean"):
            shift[: X.d] = self.mean
        sc = torch.from_numpy(scale[: X.d].astype(np.float32)).to(X.data.device)
        if not self.getOrDefault("withMean"):
            # pure scaling: hand back a VIEW (raw rows + per-column factor).  KMeans folds the factor into its
            # operand conversion (b2f_kmeans_step_scaled), so cfg 5 never writes a second 200M x 64 matrix;
            # any other consumer materialises the copy on first use (DeviceMatrix.materialized).
            return frame.withColumn(out, DeviceMatrix(X.data, X.d, X.row0, col_scale=sc))
        sh = torch.from_numpy(shift[: X.d].astype(np.float32)).to(X.data.device)
        Y = torch.empty_like(X.data)
        N.check(N.lib().b2f_scale_columns(N.ptr(X.data), N.ptr(Y), X.n, X.d, X.ld, N.ptr(sc), N.ptr(sh),
                                          N.stream_ptr()), "b2f_scale_columns")
        torch.cuda.current_stream().synchronize()  # sc / sh are temporaries of this call
        return frame.withColumn(out, DeviceMatrix(Y, X.d, X.row0))
