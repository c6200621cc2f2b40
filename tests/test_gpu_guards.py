"""Out-of-bounds check of our own (compute-sanitizer is not available on the GPU pool): every buffer a kernel is
handed is carved out of one arena with poisoned guard bands on both sides; after the call the bands must be intact and
the results must still match the oracle.  This catches out-of-bounds WRITES (a stray read leaves no trace in a band;
reads are covered only as far as a wrong value would show in the result).  Ragged shapes on purpose (row counts that are not multiples of the 128-row
tile, the smallest and an odd number of centre tiles)."""
import numpy as np
import pytest
import torch

import oracle
from oracle import synth

pytestmark = pytest.mark.gpu
SEED = 20240611
GUARD = 4096
POISON = 0xA5


class Arena:
    """Carves 1024-byte aligned buffers out of one allocation, each followed (and the first preceded) by a guard band."""

    def __init__(self, total_bytes, dev):
        self.buf = torch.full((total_bytes,), POISON, dtype=torch.uint8, device=dev)
        base = self.buf.data_ptr()
        self.off = (-base) % 1024 + GUARD
        self.guards = [(self.off - GUARD, self.off)]

    def take(self, nbytes, dtype, shape, zero=True):
        a = self.off
        b = a + nbytes
        t = self.buf[a:b].view(dtype).view(*shape)
        if zero:
            t.zero_()
        nxt = (b + GUARD + 1023) // 1024 * 1024
        self.guards.append((b, nxt))
        self.off = nxt
        assert self.off <= self.buf.numel(), "arena too small"
        return t

    def check(self, what):
        torch.cuda.synchronize()
        for a, b in self.guards:
            band = self.buf[a:b]
            bad = (band != POISON).nonzero()
            assert bad.numel() == 0, "%s wrote outside its buffers: guard band [%d, %d) touched at offset %d" % (
                what, a, b, a + int(bad[0]))


def _imports():
    from b200fit import _native as N
    import gpu_util as G

    return N, G


@pytest.mark.parametrize("n,d,k,algo_name", [(4099, 128, 256, "KM_TCH"), (1000, 64, 384, "KM_TCH"), (129, 64, 128, "KM_TCH"),
                                             (4099, 128, 256, "KM_TCB"), (3001, 128, 1024, "KM_TCH"),
                                             (777, 48, 10, "KM_GENERIC")])
def test_kmeans_step_stays_inside_its_buffers(n, d, k, algo_name):
    N, G = _imports()
    L = N.lib()
    algo = getattr(N, algo_name)
    dev = G.dev()
    ld = (d + 3) // 4 * 4
    Xh = synth.mixture(n, 0, d, SEED, max(k, 2), 2.0)
    C = Xh[np.random.default_rng(n).choice(n, size=k, replace=False)].astype(np.float64)
    prep_b, acc_b = int(L.b2f_kmeans_prep_bytes(k, d)), int(L.b2f_kmeans_acc_bytes(k, d))
    ws_b = int(L.b2f_kmeans_ws_bytes(n, k, d))
    ar = Arena(n * ld * 4 + k * d * 8 + prep_b + acc_b + ws_b + n * 4 + 16 * (GUARD + 1024), dev)
    X = ar.take(n * ld * 4, torch.float32, (n, ld))
    X[:, :d] = torch.from_numpy(Xh).to(dev)
    cen = ar.take(k * d * 8, torch.float64, (k, d))
    cen.copy_(torch.from_numpy(C))
    prep = ar.take(prep_b, torch.uint8, (prep_b,))
    acc = ar.take(acc_b, torch.uint8, (acc_b,))
    ws = ar.take(ws_b, torch.uint8, (ws_b,))
    labels = ar.take(n * 4, torch.int32, (n,))
    s = G.stream()
    N.check(L.b2f_kmeans_prepare(N.ptr(cen), k, d, N.ptr(prep), s), "prepare")
    ar.check("b2f_kmeans_prepare")
    N.check(L.b2f_kmeans_acc_zero(N.ptr(acc), k, d, s), "acc_zero")
    if algo in (N.KM_TCH, N.KM_TCB):
        N.check(L.b2f_kmeans_step_ws(N.ptr(X), n, d, ld, None, N.ptr(prep), k, N.ptr(acc), N.ptr(labels), N.ptr(ws),
                                     ws_b, algo, s), "step_ws")
    else:
        N.check(L.b2f_kmeans_step(N.ptr(X), n, d, ld, None, N.ptr(prep), k, N.ptr(acc), N.ptr(labels), algo, s), "step")
    ar.check("b2f_kmeans_step (%s)" % algo_name)
    N.check(L.b2f_kmeans_fold(N.ptr(acc), k, d, s), "fold")
    ar.check("b2f_kmeans_fold")
    plen = int(L.b2f_kmeans_payload_len(k, d))
    p = acc[: plen * 8].view(torch.float64).cpu().numpy()
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    got = labels.cpu().numpy()
    assert not ((got != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    assert p[k * d + k + 1] == n
    if (got == ref_lab).all():
        np.testing.assert_array_equal(p[k * d: k * d + k], np.bincount(ref_lab, minlength=k))
    assert p[k * d + k] == pytest.approx(float(ref_d.sum()), rel=1e-5, abs=1e-3)


@pytest.mark.parametrize("n,d,k", [(1001, 3, 4), (4099, 130, 7)])
def test_scaler_and_silhouette_stay_inside_their_buffers(n, d, k):
    N, G = _imports()
    L = N.lib()
    dev = G.dev()
    ld = (d + 3) // 4 * 4
    rng = np.random.default_rng(d)
    Xh = (rng.normal(size=(n, d)) * 2 + 1).astype(np.float32)
    lab_h = rng.integers(0, k, size=n).astype(np.int32)
    sc_b = int(L.b2f_silhouette_scratch_bytes(k, d))
    ar = Arena(2 * n * ld * 4 + 3 * d * 8 + n * 4 + k * (d + 2) * 8 + sc_b + 16 * (GUARD + 1024), dev)
    X = ar.take(n * ld * 4, torch.float32, (n, ld))
    X[:, :d] = torch.from_numpy(Xh).to(dev)
    Y = ar.take(n * ld * 4, torch.float32, (n, ld))
    scale = ar.take(d * 4, torch.float32, (d,))
    scale.copy_(torch.from_numpy(rng.uniform(0.5, 2.0, size=d).astype(np.float32)))
    stats_cols = ar.take((2 * d + 2) * 8, torch.float64, (2 * d + 2,))
    lab = ar.take(n * 4, torch.int32, (n,))
    lab.copy_(torch.from_numpy(lab_h))
    stats = ar.take(k * (d + 2) * 8, torch.float64, (k * (d + 2),))
    scratch = ar.take(sc_b, torch.uint8, (sc_b,))
    out = ar.take(16, torch.float64, (2,))
    s = G.stream()
    N.check(L.b2f_colstats(N.ptr(X), n, d, ld, None, None, N.ptr(stats_cols), s), "colstats")
    ar.check("b2f_colstats")
    N.check(L.b2f_scale_columns(N.ptr(X), N.ptr(Y), n, d, ld, N.ptr(scale), None, s), "scale")
    ar.check("b2f_scale_columns")
    N.check(L.b2f_silhouette_stats(N.ptr(X), n, d, ld, N.ptr(lab), None, k, N.ptr(stats), None, s), "sil stats")
    N.check(L.b2f_silhouette_score(N.ptr(X), n, d, ld, N.ptr(lab), None, k, N.ptr(stats), N.ptr(scratch), N.ptr(out), s),
            "sil score")
    ar.check("b2f_silhouette_*")
    sc = stats_cols.cpu().numpy()
    np.testing.assert_allclose(sc[:d], Xh.astype(np.float64).sum(0), rtol=1e-10, atol=1e-8)
    assert sc[2 * d] == n
    np.testing.assert_allclose(Y.cpu().numpy()[:, :d], Xh * scale.cpu().numpy()[None, :], rtol=1e-6)
    tot, cnt = out.cpu().numpy()
    assert tot / cnt == pytest.approx(oracle.silhouette_sqeuclid(Xh.astype(np.float64), lab_h), abs=1e-5)


@pytest.mark.parametrize("n,d,K", [(1001, 3, 2), (4099, 300, 2), (3001, 128, 5), (2500, 512, 20), (1200, 257, 8)])
def test_lr_passes_stay_inside_their_buffers(n, d, K):
    """Binary pass, CUDA-core multinomial pass and the tcgen05 multinomial pass (d % 128 == 0, ld == d)."""
    N, G = _imports()
    L = N.lib()
    dev = G.dev()
    ld = (d + 3) // 4 * 4
    KM = 1 if K == 2 else K
    rng = np.random.default_rng(n + d + K)
    Xh = rng.normal(size=(n, d)).astype(np.float32)
    yh = rng.integers(0, K, size=n).astype(np.float32)
    B = rng.normal(size=(KM, d)) * 0.3 / np.sqrt(d)
    b0 = rng.normal(size=KM) * 0.1
    sc_b = int(L.b2f_lr_scratch_bytes(d, K))
    ar = Arena(n * ld * 4 + n * 4 + KM * d * 8 + KM * 8 + (KM * d + KM + 2) * 8 + sc_b + 16 * (GUARD + 1024), dev)
    X = ar.take(n * ld * 4, torch.float32, (n, ld))
    X[:, :d] = torch.from_numpy(Xh).to(dev)
    y = ar.take(n * 4, torch.float32, (n,))
    y.copy_(torch.from_numpy(yh))
    beta = ar.take(KM * d * 8, torch.float64, (KM, d))
    beta.copy_(torch.from_numpy(B))
    icpt = ar.take(KM * 8, torch.float64, (KM,))
    icpt.copy_(torch.from_numpy(b0))
    out = ar.take((KM * d + KM + 2) * 8, torch.float64, (KM * d + KM + 2,))
    scratch = ar.take(sc_b, torch.uint8, (sc_b,), zero=False)
    s = G.stream()
    for rep in range(2):  # twice: the scratch keeps the first call's contents
        out.zero_()
        if K == 2:
            N.check(L.b2f_lr_binary_lossgrad(N.ptr(X), N.ptr(y), None, n, d, ld, N.ptr(beta), N.ptr(icpt), N.ptr(out),
                                             N.ptr(scratch), s), "lr_binary")
        else:
            N.check(L.b2f_lr_multinomial_lossgrad(N.ptr(X), N.ptr(y), None, n, d, ld, K, N.ptr(beta), N.ptr(icpt),
                                                  N.ptr(out), N.ptr(scratch), s), "lr_multinomial")
        ar.check("LR pass (K = %d, d = %d, call %d)" % (K, d, rep))
    p = out.cpu().numpy()
    W = p[KM * d + KM + 1]
    assert W == n
    if K == 2:
        rl, rg, _ = oracle.lr_binary_lossgrad(Xh, yh, None, B[0], float(b0[0]), np.ones(d))
        rg = rg[None, :]
    else:
        rl, rg, _ = oracle.lr_multinomial_lossgrad(Xh, yh, None, B, b0, np.ones(d))
    assert p[KM * d + KM] / W == pytest.approx(rl, rel=1e-5)
    scale = np.abs(rg).max() + 1e-30
    np.testing.assert_allclose(p[: KM * d].reshape(KM, d) / W, rg, rtol=2e-5, atol=2e-5 * scale)
