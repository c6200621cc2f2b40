"""Oracle (test infrastructure): CPU mirror of the device data generator
(orange3-spark_b200/csrc/synth.cu): Philox4x32-10 keyed on
(seed, global row, column) and an 8-term Irwin-Hall normal.  Integer arithmetic plus one
exact int->fp32 conversion and IEEE fp32 mul/add, so the values are bit-identical to the GPU's.
No reference counterpart (the reference ships no data generator)."""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
IH_SCALE = np.float32(float.fromhex("0x1.3988e2p-16"))
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint32) for v in (c0, c1, c2, c3))
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & _MASK).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & _MASK).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32((int(k0) + int(W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def _ih_normal(words):
    s = np.zeros(words[0].shape, dtype=np.int64)
    for w in words:
        s += (w & np.uint32(0xFFFF)).astype(np.int64) + (w >> np.uint32(16)).astype(np.int64)
    return (s - 262140).astype(np.float32) * IH_SCALE


def znorm(seed, rows, cols, stream):
    rows = np.asarray(rows, dtype=np.uint64)
    r = philox4x32_10((rows & _MASK).astype(np.uint32), (rows >> np.uint64(32)).astype(np.uint32),
                      np.asarray(cols, dtype=np.uint32), np.uint32(stream), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return _ih_normal(r)


def urand32(seed, rows, col, stream):
    rows = np.asarray(rows, dtype=np.uint64)
    return philox4x32_10((rows & _MASK).astype(np.uint32), (rows >> np.uint64(32)).astype(np.uint32),
                         np.uint32(col), np.uint32(stream), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)[0]


def mixture(n, row0, d, seed, n_components, center_scale, col_scale=None):
    """Mirror of b2f_synth_mixture: X[n, d] fp32."""
    rows = np.arange(row0, row0 + n, dtype=np.uint64)
    comp = (urand32(seed, rows, 0, 1) % np.uint32(n_components)).astype(np.uint64)
    cols = np.arange(d, dtype=np.uint32)
    z = znorm(seed, rows[:, None], cols[None, :], 0)
    mu = np.float32(center_scale) * znorm(seed, comp[:, None], cols[None, :], 2)
    X = (mu + z).astype(np.float32)
    if col_scale is not None:
        X = X * np.asarray(col_scale, dtype=np.float32)[None, :]
    return X


def mixture_centers(d, seed, n_components, center_scale):
    comp = np.arange(n_components, dtype=np.uint64)
    return np.float32(center_scale) * znorm(seed, comp[:, None], np.arange(d, dtype=np.uint32)[None, :], 2)


def logistic_features(n, row0, d, seed, col_scale=None):
    """Mirror of the X part of b2f_synth_logistic (labels depend on device expf; take them from the GPU)."""
    rows = np.arange(row0, row0 + n, dtype=np.uint64)
    z = znorm(seed, rows[:, None], np.arange(d, dtype=np.uint32)[None, :], 0)
    if col_scale is not None:
        z = z * np.asarray(col_scale, dtype=np.float32)[None, :]
    return z.astype(np.float32)
