"""GPU parity tests for the tcgen05 / TMA / TMEM Lloyd kernel (B2F_KM_TC3, 3xTF32) through the
C ABI: labels, sums, counts and cost vs the CPU oracle, per-iteration centres, and agreement
with the CUDA-core kernel.  Same tolerances as test_gpu_kmeans.py."""
import numpy as np
import pytest
import torch

import oracle
from oracle import synth

pytestmark = pytest.mark.gpu
SEED = 20240611


def _imports():
    from b200fit import _native as N
    import gpu_util as G

    return N, G


TC_SHAPES = [(128, 32, 128), (1, 32, 128), (1000, 64, 128), (5000, 128, 256), (4099, 128, 1024), (3000, 96, 384),
             (20000, 128, 128)]


@pytest.mark.parametrize("n,d,k", TC_SHAPES)
def test_tc3_step_matches_oracle(n, d, k):
    N, G = _imports()
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    rng = np.random.default_rng(n + d + k)
    if n >= k:
        C = Xh[rng.choice(n, size=k, replace=False)].astype(np.float64)
    else:
        C = synth.mixture_centers(d, SEED, k, 2.0).astype(np.float64)
    X, _, ld = G.upload_matrix(Xh)
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TC3)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    bad = (labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)
    assert not bad.any(), "%d label mismatches outside the 1e-6 gap band" % bad.sum()
    _, rcost, rsums, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    assert nrows == n and counts.sum() == n
    if (labels == ref_lab).all():
        np.testing.assert_array_equal(counts, rcounts)
        scale = np.abs(Xh).astype(np.float64).sum(axis=0).max() / k + 1.0
        np.testing.assert_allclose(sums, rsums, rtol=1e-5, atol=1e-6 * scale)
    # rows that coincide with their centre have distance ~0 (n <= k): absolute slack of fp32 cancellation there only
    assert cost == pytest.approx(rcost, rel=1e-5, abs=0.0 if n > k else 1e-3 * n)


def test_tc3_agrees_with_generic_kernel_and_predict():
    N, G = _imports()
    L = N.lib()
    n, d, k = 30000, 128, 512
    Xh = synth.mixture(n, 0, d, SEED + 1, k, 2.0)
    C = synth.mixture_centers(d, SEED + 1, k, 2.0).astype(np.float64) + 0.05
    X, _, ld = G.upload_matrix(Xh)
    lt, st, ct, costt, _ = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TC3)
    lg, sg, cg, costg, _ = G.kmeans_step_raw(X, n, d, ld, C, N.KM_GENERIC)
    assert (lt == lg).mean() > 0.9999
    assert costt == pytest.approx(costg, rel=1e-5)
    cen, prep, acc = G.kmeans_state(C, k, d)
    lab = torch.empty(n, dtype=torch.int32, device=G.dev())
    N.check(L.b2f_kmeans_predict(N.ptr(X), n, d, ld, N.ptr(prep), k, N.ptr(lab), None, N.KM_TC3, G.stream()), "predict")
    assert np.array_equal(lab.cpu().numpy(), lt)


def test_tc3_fit_per_iteration_parity():
    """BASELINE config 2 shape at reduced N (d = 128, k = 1024, fixed initialModel, tol = 0)."""
    N, G = _imports()
    from b200fit.clustering import CudaLloydEngine
    from b200fit.dist import Comm
    from b200fit.feature import pack_columns

    n, d, k, iters = 60000, 128, 1024, 5
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C0 = Xh[np.random.default_rng(2).choice(n, size=k, replace=False)].astype(np.float64)
    ref = oracle.kmeans_lloyd(Xh.astype(np.float64), C0, max_iter=iters, tol=0.0)
    eng = CudaLloydEngine(pack_columns([Xh]), k, algo=N.KM_AUTO)
    eng.set_centers(C0)
    comm = Comm()
    for it in range(iters):
        payload = eng.step()
        comm.all_reduce_sum(payload)
        _, cost = eng.finalize(payload, 0.0)
        sc = np.abs(ref["centers_hist"][it]).max()
        np.testing.assert_allclose(eng.get_centers(), ref["centers_hist"][it], rtol=1e-5, atol=1e-5 * sc)
        assert cost == pytest.approx(ref["cost_hist"][it], rel=1e-5)
        np.testing.assert_array_equal(eng.get_counts(payload), ref["counts_hist"][it])


def test_tc1_filter_pass_is_sane():
    N, G = _imports()
    n, d, k = 8000, 128, 256
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C = synth.mixture_centers(d, SEED, k, 2.0).astype(np.float64) + 0.02
    X, _, ld = G.upload_matrix(Xh)
    lt, _, counts, cost, _ = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TC1)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    clear = (ref_second - ref_d) > 0.05 * ref_second
    assert (lt[clear] == ref_lab[clear]).all() and counts.sum() == n
    assert cost == pytest.approx(float(ref_d.sum()), rel=5e-3)


def test_tc_rejects_unsupported_shapes():
    N, G = _imports()
    Xh = np.zeros((100, 48), dtype=np.float32)
    X, _, ld = G.upload_matrix(Xh)
    with pytest.raises(N.B200FitError):
        G.kmeans_step_raw(X, 100, 48, ld, np.zeros((128, 48)), N.KM_TC3)
    with pytest.raises(N.B200FitError):
        G.kmeans_step_raw(X[:, :32].contiguous(), 100, 32, 32, np.zeros((100, 32)), N.KM_TC3)


# (k / 128 odd: the two MMA issuers, the two epilogue warps per quarter and the three accumulators then fall out of
# step with the row tiles; 45000 rows = 352 row tiles: several per CTA)
TCB_SHAPES = [(128, 64, 128), (1000, 64, 128), (5000, 128, 256), (4099, 128, 1024), (20000, 128, 128),
              (30000, 64, 256), (3000, 64, 384), (45000, 128, 640), (40000, 64, 128),
              # k > 1024: row counts through global atomics instead of the per-CTA shared-memory counters; at d = 64
              # more centre tiles than stages (no resident tiles)
              (9000, 64, 2048), (6000, 128, 1536)]


FILTERS = ["TCB", "TCH"]  # bf16x3 filter (three MMAs per product) / fp16 filter (one MMA, scaled operands)


@pytest.mark.parametrize("filt", FILTERS)
@pytest.mark.parametrize("n,d,k", TCB_SHAPES)
def test_tcb_step_matches_oracle(n, d, k, filt):
    """Low-precision filter pass + 3xTF32 re-check of uncertain rows: same bar as the 3xTF32 kernel."""
    N, G = _imports()
    algo = getattr(N, "KM_" + filt)
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    rng = np.random.default_rng(n + d + k)
    C = Xh[rng.choice(n, size=k, replace=False)].astype(np.float64) if n >= k else \
        synth.mixture_centers(d, SEED, k, 2.0).astype(np.float64)
    X, _, ld = G.upload_matrix(Xh)
    info = {}
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, algo, info=info)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    bad = (labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)
    assert not bad.any(), "%d label mismatches outside the 1e-6 gap band" % bad.sum()
    _, rcost, rsums, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    assert nrows == n and counts.sum() == n
    assert info["rechecked"] == info["listed"] and info["uncertain"] <= n
    assert (info["pairs"] == 0) == (filt == "TCB") or info["uncertain"] == 0
    if (labels == ref_lab).all():
        np.testing.assert_array_equal(counts, rcounts)
        scale = np.abs(Xh).astype(np.float64).sum(axis=0).max() / k + 1.0
        np.testing.assert_allclose(sums, rsums, rtol=1e-5, atol=1e-6 * scale)
    # fp16 filter: the cost comes from sum ||x||^2 and the accumulated sums, a difference of large terms whose fp32
    # roundings only show when the true cost is ~0 (rows sitting on their centres: the n == k case).
    # bf16x3 filter: each certain row carries its score error (|err| <= tau ~ 0.1 here, zero mean).
    if n > k and filt == "TCH":
        assert cost == pytest.approx(rcost, rel=1e-5)
    else:
        assert cost == pytest.approx(rcost, rel=1e-5, abs=(0.3 * np.sqrt(n) if n > k else 2e-2 * n))


@pytest.mark.parametrize("filt", FILTERS)
def test_tcb_uncertain_rows_are_rechecked_and_overflow_is_visible(filt):
    """Two centres almost on top of each other make most of their rows uncertain."""
    N, G = _imports()
    algo = getattr(N, "KM_" + filt)
    n, d, k = 20000, 128, 128
    Xh = synth.mixture(n, 0, d, SEED, 4, 2.0)
    C = synth.mixture_centers(d, SEED, k, 2.0).astype(np.float64)
    # near-duplicate of component 0's centre.  (Index 6, not 5: the fp16 filter's epilogue folds columns {0, 1} mod 4
    # and {2, 3} mod 4 of a centre tile in separate chains and can only bound the THIRD best score -- the condition
    # for its pair check -- when the two best sit in different chains; same-chain pairs go to the 3xTF32 re-check.)
    C[6] = C[0] + 1e-3
    X, _, ld = G.upload_matrix(Xh)
    info = {}
    big = int(N.lib().b2f_kmeans_ws_bytes(16 * n, k, d))  # room for every row on the re-check list
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, algo, ws_bytes=big, info=info)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert info["uncertain"] > n // 8 and nrows == n
    assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    if filt == "TCH":  # rows torn between the two near-duplicates have exactly two candidates: the pair check
        assert info["pairs"] > n // 8
    # a workspace with room for 128 rows only: the excess is dropped and the row count shows it
    small = 256 + 5 * 1024 + 128 * (28 + 4 * d) + 4096
    info2 = {}
    _, _, _, _, nrows2 = G.kmeans_step_raw(X, n, d, ld, C, algo, ws_bytes=small, info=info2)
    assert info2["cap"] == 128
    assert info2["uncertain"] == info["uncertain"] and info2["rechecked"] < info2["listed"]
    assert nrows2 == n - (info2["listed"] - info2["rechecked"])


@pytest.mark.parametrize("scale", [1e-6, 1.0, 3e5])
def test_tch_operand_scaling_and_fp16_overflow_rows(scale):
    """The fp16 filter scales its operands by a power of two taken from the centres; rows with entries far above
    the centres' range overflow fp16 and must come back through the re-check with the right label."""
    N, G = _imports()
    n, d, k = 6000, 128, 128
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C = synth.mixture_centers(d, SEED, k, 2.0).astype(np.float64) * scale
    Xh = (Xh * np.float32(scale)).astype(np.float32)
    Xh[17, 5] = np.float32(1e4 * scale)   # beyond 65504 / S for these centres
    Xh[4000, :] *= np.float32(500.0)
    Xh[123, :] = 0.0
    X, _, ld = G.upload_matrix(Xh)
    info = {}
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TCH, info=info)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert nrows == n and counts.sum() == n and 2 <= info["uncertain"] < n // 4
    assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    _, rcost, rsums, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    assert cost == pytest.approx(rcost, rel=1e-5)
    if (labels == ref_lab).all():
        np.testing.assert_array_equal(counts, rcounts)


@pytest.mark.parametrize("filt", FILTERS)
def test_tcb_fit_per_iteration_parity_and_fallback(filt):
    N, G = _imports()
    from b200fit.clustering import CudaLloydEngine, lloyd
    from b200fit.dist import Comm
    from b200fit.feature import pack_columns

    n, d, k, iters = 60000, 128, 1024, 5
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C0 = Xh[np.random.default_rng(2).choice(n, size=k, replace=False)].astype(np.float64)
    ref = oracle.kmeans_lloyd(Xh.astype(np.float64), C0, max_iter=iters, tol=0.0)
    eng = CudaLloydEngine(pack_columns([Xh]), k, algo=getattr(N, "KM_" + filt))
    eng.set_centers(C0)
    it, costs, conv, payload = lloyd(eng, Comm(), iters, 0.0)
    sc = np.abs(ref["centers"]).max()
    np.testing.assert_allclose(eng.get_centers(), ref["centers"], rtol=1e-5, atol=1e-5 * sc)
    np.testing.assert_allclose(costs, ref["cost_hist"], rtol=1e-5)
    np.testing.assert_array_equal(eng.get_counts(payload), ref["counts_hist"][-1])
    assert eng.fallbacks == 0


@pytest.mark.parametrize("filt", FILTERS)
def test_filter_pipeline_is_repeatable_across_many_runs(filt):
    """The filter kernel is a web of mbarrier hand-offs between twenty warps; a timing-dependent race would show as
    run-to-run variation.  Twelve runs of the same step (about 16 row tiles per CTA, an odd number of centre tiles)
    must give identical labels and counts, equal to the oracle's."""
    N, G = _imports()
    n, d, k = 300000, 128, 384
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C = Xh[np.random.default_rng(5).choice(n, size=k, replace=False)].astype(np.float64)
    X, _, ld = G.upload_matrix(Xh)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    first = None
    for run in range(12):
        labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, getattr(N, "KM_" + filt))
        assert nrows == n and counts.sum() == n
        if first is None:
            first = (labels.copy(), counts.copy(), cost)
            assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
        else:
            assert np.array_equal(labels, first[0]), "labels changed between runs (run %d)" % run
            assert np.array_equal(counts, first[1]), "counts changed between runs (run %d)" % run
            # labels and counts are bit-identical; the cost is not: the fp16 path forms it from the fp32 replica sums,
            # whose rounding depends on the order in which the reductions land (observed: 1e-9 relative)
            assert cost == pytest.approx(first[2], rel=1e-7)
