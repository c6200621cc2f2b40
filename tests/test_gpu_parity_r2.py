"""GPU parity tests added in round 2 (all through the C ABI or the public classes):

  * the exact fallback (3xTF32 WITH its certainty test + fp64 tail) at step level and driven through a whole fit;
  * the pair check of the fp16 filter (rows with exactly two candidate centres decided by two fp64 distances);
  * the certainty pipeline on the predict path (labels, distances, computeCost);
  * the StandardScaler fold (b2f_kmeans_step_scaled: no scaled copy) against the materialised copy and the oracle;
  * FULL-SIZE accumulation checks: the fp32 replica sums of the Lloyd step at 20M x 128 / k = 1024 and 40M x 64 /
    k = 256, and the multinomial gradient at 20M x 512 (its accumulator lives in tensor memory for a CTA's whole
    life), each against an fp64 re-accumulation on the device from the kernel's own labels / iterate;
  * constant-label LogisticRegression, string `thresholds`, the DataFrame return path.
"""
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


def _near_duplicate_setup(n=20000, d=128, k=128):
    Xh = synth.mixture(n, 0, d, SEED, 4, 2.0)
    C = synth.mixture_centers(d, SEED, k, 2.0).astype(np.float64)
    C[6] = C[0] + 1e-3  # rows of component 0 are torn between centres 0 and 6
    return Xh, C


def test_tc3_with_workspace_certifies_its_labels():
    """B2F_KM_TC3 through b2f_kmeans_step_ws: rows inside the 3xTF32 band go to the fp64 kernel (header [2] > 0),
    exact duplicates tie to the lower index, nothing is dropped."""
    N, G = _imports()
    n, d, k = 20000, 128, 128
    Xh, C = _near_duplicate_setup(n, d, k)
    C[7] = C[1]  # an exact duplicate: every row of component 1 is an exact tie
    X, _, ld = G.upload_matrix(Xh)
    info = {}
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TC3, use_ws=True, info=info)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert nrows == n and counts.sum() == n
    assert info["exact"] > n // 8 and info["cap"] >= n and info["listed"] == 0 and info["pairs"] == 0
    assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    assert counts[7] == 0 and not (labels == 7).any()  # findClosest: the lower index wins exact ties
    _, rcost, rsums, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    assert cost == pytest.approx(rcost, rel=1e-5)


@pytest.mark.parametrize("d", [64, 128])
def test_one_centre_tile_many_row_tiles_per_cta(d):
    """k = 128 (ONE centre tile per row tile) with eight row tiles per CTA, through the 3xTF32 kernel WITH its
    certainty test: nothing but the row-norm hand-off then keeps its split warps from running two row tiles ahead of
    the epilogue, and that hand-off once had no back-pressure (the kernel timed out in a barrier wait from ~4 row
    tiles per CTA on).  Also through the filter pipeline, whose re-check pass is the same kernel."""
    N, G = _imports()
    k = 128
    n = 148 * 128 * 8 + 77
    Xh = synth.mixture(n, 0, d, SEED + 21, 32, 2.0)
    C = Xh[np.random.default_rng(5).choice(n, size=k, replace=False)].astype(np.float64)
    C[64:] = C[:64] + 1e-4  # every row is torn between a centre and its shifted copy: all of them get re-checked
    X, _, ld = G.upload_matrix(Xh)
    X64 = Xh.astype(np.float64)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(X64, C, return_second=True)
    _, rcost, rsums, rcounts = oracle.kmeans_step(X64, C)
    big = int(N.lib().b2f_kmeans_ws_bytes(6 * n, k, d))  # lists with room for every row
    for algo, use_ws in ((N.KM_TC3, True), (N.KM_TCB, False), (N.KM_TCH, False)):
        info = {}
        labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, algo, use_ws=use_ws, info=info,
                                                              ws_bytes=big)
        assert nrows == n and counts.sum() == n
        assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
        assert cost == pytest.approx(rcost, rel=1e-5)
        if algo == N.KM_TCB:  # no pair check in the bf16x3 filter: every row goes through the re-check kernel
            assert info["rechecked"] > n // 2


@pytest.mark.parametrize("d,k", [(32, 128), (96, 256), (64, 384), (128, 640), (64, 1024), (128, 256), (64, 256)])
def test_six_row_tiles_per_cta_across_shapes(d, k):
    """The barrier rings of the tensor-core kernels only wrap with several row tiles per CTA, which the small parity
    shapes (one or two tiles per CTA) never reach: one step of every tensor-core path at ~114 000 rows, odd and even
    numbers of centre tiles, against the oracle."""
    N, G = _imports()
    n = 148 * 128 * 6 + 5
    Xh = synth.mixture(n, 0, d, SEED + 31 + d + k, 24, 2.0)
    C = Xh[np.random.default_rng(d + k).choice(n, size=k, replace=False)].astype(np.float64)
    X, _, ld = G.upload_matrix(Xh)
    X64 = Xh.astype(np.float64)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(X64, C, return_second=True)
    _, rcost, rsums, rcounts = oracle.kmeans_step(X64, C)
    big = int(N.lib().b2f_kmeans_ws_bytes(6 * n, k, d))
    runs = [(N.KM_TC3, False), (N.KM_TC3, True)]
    if d in (64, 128):
        runs += [(N.KM_TCB, False), (N.KM_TCH, False)]
    sc = np.abs(rsums).max()
    for algo, use_ws in runs:
        labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, algo, use_ws=use_ws, ws_bytes=big)
        assert nrows == n and counts.sum() == n
        bad = (labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)
        if algo == N.KM_TC3 and not use_ws:  # no certainty test: 3xTF32 labels, a wider tie band
            bad = (labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second, rel=1e-5)
        assert not bad.any()
        assert cost == pytest.approx(rcost, rel=1e-5)
        if not (labels != ref_lab).any():
            np.testing.assert_array_equal(counts, rcounts)
            np.testing.assert_allclose(sums, rsums, rtol=1e-5, atol=1e-5 * sc)


@pytest.mark.parametrize("filt", ["TCB", "TCH"])
def test_fit_driven_through_the_exact_fallback_matches_oracle(filt):
    """A workspace far too small for the uncertain rows: every pass of the fit overflows its lists, lloyd() redoes
    it with exact=True, and centres / costs / counts must still be the oracle's."""
    N, G = _imports()
    from b200fit.clustering import CudaLloydEngine, lloyd
    from b200fit.dist import Comm
    from b200fit.feature import pack_columns
    from oracle.kmeans import kmeans_update

    n, d, k, iters = 60000, 128, 256, 4
    Xh = synth.mixture(n, 0, d, SEED, 8, 2.0)  # 8 components, 256 centres: most rows have close competitors
    C0 = Xh[np.random.default_rng(3).choice(n, size=k, replace=False)].astype(np.float64)
    small = 256 + 5 * 1024 + 128 * (28 + 4 * d) + 4096 + 4 * n  # lists of 128 rows; as a flat list: every row fits
    eng = CudaLloydEngine(pack_columns([Xh]), k, algo=getattr(N, "KM_" + filt), ws_bytes=small)
    eng.set_centers(C0)
    X64 = Xh.astype(np.float64)
    # Checked iteration by iteration, each against the oracle's step FROM THE ENGINE'S OWN CENTRES: this data is
    # chosen to have near ties everywhere, so a whole trajectory is chaotic -- centres that differ in the last bits
    # (the fp32 replica sums are added in a different order every run) flip one torn row in the next pass and the
    # flip grows from there (measured: 1 run in ~30 ends 0.05 away from the oracle's fourth iterate, costs 7e-7
    # apart).  One pass from given centres has no such freedom: its labels are certified.
    for i in range(iters):
        C = eng.get_centers()
        ref_lab, rcost, rsums, rcounts = oracle.kmeans_step(X64, C)
        refC, _ = kmeans_update(C, rsums, rcounts, 0.0)
        it, costs, conv, payload = lloyd(eng, Comm(), 1, 0.0)
        assert it == 1 and eng.fallbacks == i + 1, "every pass was expected to overflow the 128-row lists"
        sc = np.abs(refC).max()
        np.testing.assert_allclose(eng.get_centers(), refC, rtol=1e-5, atol=1e-5 * sc)
        np.testing.assert_allclose(costs, [rcost], rtol=1e-5)
        np.testing.assert_array_equal(eng.get_counts(payload), rcounts)
        _, ref_d, ref_second = oracle.kmeans_assign(X64, C, return_second=True)
        lab = eng.labels_buf.cpu().numpy()
        assert not ((lab != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()


def test_pair_check_decides_two_candidate_rows_exactly():
    """fp16 filter, rows between two close centres: they must leave through the pair list (not the 3xTF32 re-check)
    with the fp64 label, and accumulate like any other row."""
    N, G = _imports()
    n, d, k = 50000, 128, 256
    Xh = synth.mixture(n, 0, d, SEED + 3, 16, 2.0)
    mu = synth.mixture_centers(d, SEED + 3, 16, 2.0).astype(np.float64)
    rng = np.random.default_rng(8)
    C = synth.mixture_centers(d, SEED + 9, k, 2.0).astype(np.float64)  # far-away fillers
    # Two centres per component, 0.5 apart: a wide band of torn rows, everything else far.  Where the two sit
    # matters to the filter: its epilogue folds columns {0, 1} mod 4 and {2, 3} mod 4 of a centre tile in separate
    # chains, and two epilogue warps take alternate tiles.  Half of the pairs share a tile (different chains), the
    # other half straddle two tiles, so both the in-tile and the cross-warp merge of the (best, second, third)
    # triples are exercised.
    for j in range(16):
        delta = rng.normal(size=d)
        delta *= 0.25 / np.linalg.norm(delta)
        a, b = (4 * j, 4 * j + 2) if j < 8 else (4 * j + 1, 128 + 4 * j + 3)
        C[a], C[b] = mu[j] + delta, mu[j] - delta
    X, _, ld = G.upload_matrix(Xh)
    info = {}
    big = int(N.lib().b2f_kmeans_ws_bytes(8 * n, k, d))  # lists normally hold n / 5 rows; here most rows are torn
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TCH, ws_bytes=big, info=info)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert nrows == n and counts.sum() == n
    assert info["pairs"] > n // 4, info
    assert info["listed"] < info["pairs"] // 4, info  # two close centres in one chain of one tile: the re-check list
    assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    _, rcost, rsums, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    if (labels == ref_lab).all():
        np.testing.assert_array_equal(counts, rcounts)
        scale = np.abs(Xh).astype(np.float64).sum(axis=0).max() / k + 1.0
        np.testing.assert_allclose(sums, rsums, rtol=1e-5, atol=1e-6 * scale)
    assert cost == pytest.approx(rcost, rel=1e-5)


def test_predict_runs_the_certainty_pipeline_and_reports_distances():
    """KMeansModel.transform / computeCost at the cfg-2 shape: labels inside the 1e-6 band of the oracle's, distances
    those of the labelled centre, and no CUDA-core all-centres pass behind them."""
    N, G = _imports()
    from b200fit.clustering import KMeansModel, predict_labels
    from b200fit.feature import pack_columns
    from b200fit.frame import ColumnarFrame

    n, d, k = 300000, 128, 1024
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C = Xh[np.random.default_rng(11).choice(n, size=k, replace=False)].astype(np.float64)
    C[100] = C[3] + 1e-4
    X = pack_columns([Xh])
    info = {}
    lab, dist = predict_labels(X, C, want_dist=True, info=info)
    lab, dist = lab.cpu().numpy(), dist.cpu().numpy().astype(np.float64)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert info["dropped_first_pass"] == 0
    assert not ((lab != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    np.testing.assert_allclose(dist, ref_d, rtol=2e-5, atol=1e-4)
    model = KMeansModel(C)
    frame = ColumnarFrame({"features": X})
    assert model.computeCost(frame) == pytest.approx(float(ref_d.sum()), rel=1e-5)
    out = model.transform(frame)
    assert np.array_equal(out.col("prediction").cpu().numpy(), lab)
    # header of a direct call: [3] is the list capacity, nothing dropped
    L = N.lib()
    cen, prep, acc = G.kmeans_state(C, k, d)
    wsb = int(L.b2f_kmeans_ws_bytes(n, k, d))
    ws = torch.zeros(wsb, dtype=torch.uint8, device=G.dev())
    l2 = torch.empty(n, dtype=torch.int32, device=G.dev())
    N.check(L.b2f_kmeans_predict_ws(N.ptr(X.data), n, d, X.ld, None, N.ptr(prep), k, N.ptr(l2), None, N.ptr(ws), wsb,
                                    N.KM_AUTO, G.stream()), "predict_ws")
    h = ws[:20].view(torch.int32).cpu().numpy()
    assert h[3] >= n // 5 - 128 and h[0] <= h[3] and h[2] <= h[3] and h[4] > 0
    assert np.array_equal(l2.cpu().numpy(), lab)


@pytest.mark.parametrize("n,d,k", [(40000, 64, 256), (30000, 128, 384)])
def test_scaler_fold_equals_the_materialised_copy(n, d, k):
    """b2f_kmeans_step_scaled on raw rows + per-column factor vs b2f_kmeans_step_ws on the scaled copy: same labels,
    counts and (to fp32 rounding of the row sums) sums and cost; both equal to the oracle on the scaled data."""
    N, G = _imports()
    L = N.lib()
    rng = np.random.default_rng(d + k)
    col = np.exp(rng.uniform(np.log(0.1), np.log(10.0), size=d)).astype(np.float32)
    Xraw = (synth.mixture(n, 0, d, SEED + 5, k, 2.0) * col[None, :]).astype(np.float32)
    sc = (1.0 / Xraw.astype(np.float64).std(axis=0, ddof=1)).astype(np.float32)
    X, _, ld = G.upload_matrix(Xraw)
    scd = torch.from_numpy(sc).to(G.dev())
    Y = torch.empty_like(X)
    N.check(L.b2f_scale_columns(N.ptr(X), N.ptr(Y), n, d, ld, N.ptr(scd), None, G.stream()), "scale")
    Yh = Y.cpu().numpy()[:, :d]
    C = Yh[rng.choice(n, size=k, replace=False)].astype(np.float64)
    i1, i2 = {}, {}
    lab_f, sums_f, cnt_f, cost_f, rows_f = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TCH, col_scale=scd, info=i1)
    lab_m, sums_m, cnt_m, cost_m, rows_m = G.kmeans_step_raw(Y, n, d, ld, C, N.KM_TCH, info=i2)
    assert rows_f == rows_m == n
    assert np.array_equal(lab_f, lab_m) and np.array_equal(cnt_f, cnt_m)
    assert i1["uncertain"] == i2["uncertain"] and i1["pairs"] == i2["pairs"]
    scale = np.abs(Yh).astype(np.float64).sum(axis=0).max() / k + 1.0
    np.testing.assert_allclose(sums_f, sums_m, rtol=2e-6, atol=1e-6 * scale)
    assert cost_f == pytest.approx(cost_m, rel=1e-6)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Yh.astype(np.float64), C, return_second=True)
    assert not ((lab_f != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    _, rcost, rsums, rcounts = oracle.kmeans_step(Yh.astype(np.float64), C)
    assert cost_f == pytest.approx(rcost, rel=1e-5)
    if (lab_f == ref_lab).all():
        np.testing.assert_array_equal(cnt_f, rcounts)
        np.testing.assert_allclose(sums_f, rsums, rtol=1e-5, atol=1e-6 * scale)


def test_scaled_view_is_not_materialised_by_the_kmeans_pipeline():
    """StandardScalerModel.transform hands back a view; KMeans.fit / transform / computeCost consume it through the
    fused pass and never write the scaled matrix (cfg 5: 51 GB at full size)."""
    N, G = _imports()
    from b200fit.clustering import KMeans
    from b200fit.feature import StandardScaler, pack_columns
    from b200fit.frame import ColumnarFrame

    n, d, k = 40000, 64, 256
    rng = np.random.default_rng(4)
    col = np.exp(rng.uniform(np.log(0.1), np.log(10.0), size=d)).astype(np.float32)
    Xraw = (synth.mixture(n, 0, d, SEED + 5, k, 2.0) * col[None, :]).astype(np.float32)
    frame = ColumnarFrame({"raw": pack_columns([Xraw])})
    sm = StandardScaler(inputCol="raw", outputCol="features").fit(frame)
    scaled = sm.transform(frame)
    view = scaled.col("features")
    assert view.col_scale is not None and view.data.data_ptr() == frame.col("raw").data.data_ptr()
    mean, std = oracle.standard_scaler_fit(Xraw.astype(np.float64))
    Ys = oracle.standard_scaler_transform(Xraw.astype(np.float64), mean, std).astype(np.float32).astype(np.float64)
    C0 = Ys[rng.choice(n, size=k, replace=False)]
    model = KMeans(k=k, maxIter=6, tol=0.0, initialModel=C0).fit(scaled)
    assert view._copy is None, "the fit wrote a scaled copy"
    ref = oracle.kmeans_lloyd(Ys, C0, max_iter=6, tol=0.0)
    sc = np.abs(ref["centers"]).max()
    np.testing.assert_allclose(np.stack(model.clusterCenters()), ref["centers"], rtol=1e-5, atol=1e-5 * sc)
    np.testing.assert_allclose(model.summary.costHistory, ref["cost_hist"], rtol=1e-5)
    assert model.summary.clusterSizes == [int(c) for c in ref["counts_hist"][-1]]
    cost = model.computeCost(scaled)
    lab = model.transform(scaled).col("prediction").cpu().numpy()
    assert view._copy is None, "transform / computeCost wrote a scaled copy"
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Ys, ref["centers"], return_second=True)
    assert not ((lab != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    assert cost == pytest.approx(float(ref_d.sum()), rel=1e-5)
    # random / k-means|| initialisation read rows through the view as well
    m2 = KMeans(k=k, maxIter=2, seed=7, initMode="random").fit(scaled)
    assert view._copy is None and np.isfinite(m2.summary.trainingCost)


@pytest.mark.parametrize("n,d,k", [(5000, 128, 256), (30000, 64, 256), (4099, 128, 1024)])
def test_fp16_filter_with_per_row_key_scales_matches_oracle(n, d, k):
    """Operands prepared WITHOUT the data's largest row norm (plain b2f_kmeans_prepare, what predict on unseen data
    uses): the epilogue scales every score into its row's own key window.  Same bar as the global-scale default."""
    N, G = _imports()
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C = Xh[np.random.default_rng(n + d + k).choice(n, size=k, replace=False)].astype(np.float64)
    X, _, ld = G.upload_matrix(Xh)
    info = {}
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TCH, info=info, global_scale=False)
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert nrows == n and counts.sum() == n
    assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    _, rcost, rsums, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    assert cost == pytest.approx(rcost, rel=1e-5)
    # the two modes quantise differently but certify alike: about the same number of uncertain rows
    info_g = {}
    lab_g = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TCH, info=info_g)[0]
    assert np.array_equal(lab_g, labels) or not ((lab_g != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    assert abs(info_g["uncertain"] - info["uncertain"]) <= 0.1 * max(info["uncertain"], 50)


def test_rows_longer_than_the_declared_norm_are_rechecked_not_trusted():
    """Global key scale built for rows up to HALF the data's largest norm: longer rows may leave the 22-bit key
    window, so the filter must hand every one of them to the re-check -- and the labels must still be the oracle's."""
    N, G = _imports()
    n, d, k = 20000, 128, 256
    Xh = synth.mixture(n, 0, d, SEED + 4, k, 2.0)
    Xh[::7] *= np.float32(3.0)  # a seventh of the rows three times as long
    C = synth.mixture_centers(d, SEED + 4, k, 2.0).astype(np.float64)
    X, _, ld = G.upload_matrix(Xh)
    norms = np.sqrt((Xh.astype(np.float64) ** 2).sum(1))
    limit = 0.5 * norms.max()
    info = {}
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, N.KM_TCH, info=info, xmax_override=limit,
                                                          ws_bytes=int(N.lib().b2f_kmeans_ws_bytes(8 * n, k, d)))
    ref_lab, ref_d, ref_second = oracle.kmeans_assign(Xh.astype(np.float64), C, return_second=True)
    assert nrows == n and counts.sum() == n
    assert info["listed"] >= int((norms > limit * 1.001).sum())  # every long row went to the 3xTF32 re-check
    assert not ((labels != ref_lab) & ~G.allowed_label_mismatch(ref_d, ref_second)).any()
    _, rcost, _, rcounts = oracle.kmeans_step(Xh.astype(np.float64), C)
    assert cost == pytest.approx(rcost, rel=1e-5)


def _fp64_sums_from_labels(X, labels, k, d, col_scale=None, chunk=2_000_000):
    """Test-side reference: sums[k, d] and counts re-accumulated in fp64 on the device from the labels the step
    produced (torch.index_add_ in double; independent of every kernel of the library)."""
    sums = torch.zeros((k, d), dtype=torch.float64, device=X.device)
    n = X.shape[0]
    for a in range(0, n, chunk):
        b = min(n, a + chunk)
        rows = X[a:b, :d]
        if col_scale is not None:
            rows = rows * col_scale[None, :]  # fp32 product: the value the kernels see
        sums.index_add_(0, labels[a:b].long(), rows.double())
    counts = torch.bincount(labels.long(), minlength=k).double()
    return sums, counts


@pytest.mark.parametrize("n,d,k,comps,scaled", [(20_000_000, 128, 1024, 1024, False), (40_000_000, 64, 256, 256, True)])
def test_full_size_accumulation_matches_fp64_reaccumulation(n, d, k, comps, scaled):
    """The fp32 replica sums at the size they are used at: ~20 000 (cfg 2) / ~150 000 (cfg 5 shape) rows added per
    centre before the fp64 fold.  Centres from the step's sums must be within 1e-5 relative of an fp64
    re-accumulation from the same labels; the error actually found is printed."""
    N, G = _imports()
    L = N.lib()
    free, _ = torch.cuda.mem_get_info()
    if free < n * d * 4 * 2.2:
        pytest.skip("not enough free device memory for the full-size check")
    X = torch.empty((n, d), dtype=torch.float32, device=G.dev())
    N.check(L.b2f_synth_mixture(N.ptr(X), n, 0, d, d, SEED, comps, 2.0, None, G.stream()), "synth")
    col_scale = None
    if scaled:
        col_scale = torch.exp(torch.linspace(-2.0, 2.0, d, device=G.dev())).float().contiguous()
    idx = torch.from_numpy(np.random.default_rng(1).choice(n, size=k, replace=False)).to(G.dev())
    C0 = X[idx, :]
    if scaled:
        C0 = C0 * col_scale[None, :]
    C = C0.double().cpu().numpy()
    cen, prep, acc = G.kmeans_state(C, k, d)
    wsb = int(L.b2f_kmeans_ws_bytes(n, k, d))
    ws = torch.zeros(wsb, dtype=torch.uint8, device=G.dev())
    labels = torch.full((n,), -1, dtype=torch.int32, device=G.dev())
    if scaled:
        N.check(L.b2f_kmeans_step_scaled(N.ptr(X), n, d, d, N.ptr(col_scale), N.ptr(prep), k, N.ptr(acc),
                                         N.ptr(labels), N.ptr(ws), wsb, G.stream()), "step_scaled")
    else:
        N.check(L.b2f_kmeans_step_ws(N.ptr(X), n, d, d, None, N.ptr(prep), k, N.ptr(acc), N.ptr(labels), N.ptr(ws), wsb,
                                     N.KM_AUTO, G.stream()), "step_ws")
    N.check(L.b2f_kmeans_fold(N.ptr(acc), k, d, G.stream()), "fold")
    plen = int(L.b2f_kmeans_payload_len(k, d))
    p = acc[: plen * 8].view(torch.float64)
    kd = k * d
    assert float(p[kd + k + 1].item()) == n, "rows dropped"
    assert int(labels.min().item()) >= 0 and int(labels.max().item()) < k
    ref_sums, ref_counts = _fp64_sums_from_labels(X, labels, k, d, col_scale)
    assert torch.equal(p[kd:kd + k], ref_counts)
    got_c = (p[:kd].view(k, d) / ref_counts.clamp(min=1)[:, None]).cpu().numpy()
    ref_c = (ref_sums / ref_counts.clamp(min=1)[:, None]).cpu().numpy()
    live = ref_counts.cpu().numpy() > 0
    scale = np.abs(ref_c[live]).max()
    err = np.abs(got_c[live] - ref_c[live]).max() / scale
    rel = (np.abs(got_c[live] - ref_c[live]) / np.maximum(np.abs(ref_c[live]), 1e-3 * scale)).max()
    print("\nfull-size accumulation %dx%d k=%d: max |centre error| / max |centre| = %.3g (the bar: 1e-5), max "
          "elementwise relative error with a floor of 1e-3 of the scale = %.3g (entries near zero; informational), "
          "rows per centre ~%d" % (n, d, k, err, rel, n // k))
    # the criterion of every other centre comparison in this suite: |got - ref| <= 1e-5 |ref| + 1e-5 * scale
    np.testing.assert_allclose(got_c[live], ref_c[live], rtol=1e-5, atol=1e-5 * scale)
    assert err < 1e-5
    # cost: sum over rows of the fp64 distance to the labelled centre, from a 2M-row sample scaled up is not exact;
    # use the identity cost = sum ||x||^2 - 2 sum_j c_j . S_j + sum_j n_j ||c_j||^2 in fp64 on the reference sums
    xn = torch.zeros((), dtype=torch.float64, device=G.dev())
    for a in range(0, n, 4_000_000):
        rows = X[a:a + 4_000_000]
        if scaled:
            rows = rows * col_scale[None, :]
        xn += (rows.double() ** 2).sum()
    Cd = torch.from_numpy(C).to(G.dev())
    ref_cost = float((xn - 2.0 * (Cd * ref_sums).sum() + (ref_counts * (Cd * Cd).sum(1)).sum()).item())
    assert float(p[kd + k].item()) == pytest.approx(ref_cost, rel=1e-5)


def test_full_size_labels_of_the_filter_pipeline_equal_the_3xtf32_pipeline():
    """20M x 128, k = 1024: the fp16 filter pipeline (filter, pair check, re-check, fp64 tail) and the independent
    3xTF32 pipeline (3xTF32 scores with their own certainty test, fp64 tail) both claim fp64-exact labels, by
    different arithmetic; at a size the oracle cannot reach they must give the SAME label for every row, the same
    counts, and costs equal to 1e-5 (the 3xTF32 kernel forms its cost per row in fp32)."""
    N, G = _imports()
    L = N.lib()
    n, d, k = 20_000_000, 128, 1024
    free, _ = torch.cuda.mem_get_info()
    if free < n * d * 4 * 1.8:
        pytest.skip("not enough free device memory for the full-size check")
    X = torch.empty((n, d), dtype=torch.float32, device=G.dev())
    N.check(L.b2f_synth_mixture(N.ptr(X), n, 0, d, d, SEED, k, 2.0, None, G.stream()), "synth")
    idx = torch.from_numpy(np.random.default_rng(4).choice(n, size=k, replace=False)).to(G.dev())
    C = X[idx, :].double().cpu().numpy()
    C[17] = C[900] + 1e-4  # a near-duplicate pair: a band of torn rows on top of the natural ones
    out = {}
    for name, algo in (("filter", N.KM_TCH), ("tc3", N.KM_TC3)):
        info = {}
        labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, d, C, algo, use_ws=True, info=info)
        assert nrows == n and counts.sum() == n, name
        out[name] = (labels, counts, cost, info)
    lf, lt = out["filter"][0], out["tc3"][0]
    diff = int((lf != lt).sum())
    print("\nfull-size label agreement: %d of %d rows differ; filter: %s; 3xTF32: %s" % (diff, n, out["filter"][3],
                                                                                       out["tc3"][3]))
    assert diff == 0
    assert np.array_equal(out["filter"][1], out["tc3"][1])
    assert out["filter"][2] == pytest.approx(out["tc3"][2], rel=1e-5)
    assert out["filter"][3]["pairs"] > n // 200 and out["tc3"][3]["exact"] > 0


def test_full_size_multinomial_gradient_matches_fp64_reference():
    """cfg-4 size (20M x 512, 20 classes): the tensor-core pass keeps G^T in an fp32 tensor-memory accumulator for
    the CTA's whole life (~135 000 rows) and forms the products from bf16 pieces.  Reference: fp64 softmax and fp64
    GEMM on the device (torch), chunked.  The bar is north_star's 1e-5 relative to the largest gradient entry."""
    N, G = _imports()
    L = N.lib()
    n, d, K = 20_000_000, 512, 20
    free, _ = torch.cuda.mem_get_info()
    if free < n * d * 4 * 1.3:
        pytest.skip("not enough free device memory for the full-size check")
    rng = np.random.default_rng(SEED + 2)
    scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d)).astype(np.float32)
    w_true = (rng.standard_normal((K, d)) / np.sqrt(d)).astype(np.float32)
    X = torch.empty((n, d), dtype=torch.float32, device=G.dev())
    y = torch.empty(n, dtype=torch.float32, device=G.dev())
    sc_d, w_d = torch.from_numpy(scale).to(G.dev()), torch.from_numpy(w_true).to(G.dev())
    b_d = torch.zeros(K, dtype=torch.float32, device=G.dev())
    N.check(L.b2f_synth_logistic(N.ptr(X), N.ptr(y), n, 0, d, d, SEED + 2, N.ptr(sc_d), N.ptr(w_d), N.ptr(b_d), K,
                                 G.stream()), "synth")
    beta = rng.standard_normal((K, d)) * 0.05 / scale[None, :]
    icpt = rng.standard_normal(K) * 0.1
    loss, grad, gb, W = G.lr_lossgrad_raw(X, y, None, n, d, d, K, beta, icpt)
    Bd = torch.from_numpy(beta).to(G.dev())
    bd = torch.from_numpy(icpt).to(G.dev())
    Gref = torch.zeros((K, d), dtype=torch.float64, device=G.dev())
    gbref = torch.zeros(K, dtype=torch.float64, device=G.dev())
    lref = torch.zeros((), dtype=torch.float64, device=G.dev())
    chunk = 500_000
    for a in range(0, n, chunk):
        Xc = X[a:a + chunk].double()
        yc = y[a:a + chunk].long()
        M = Xc @ Bd.T + bd[None, :]
        lse = torch.logsumexp(M, dim=1)
        lref += (lse - M.gather(1, yc[:, None])[:, 0]).sum()
        P = torch.exp(M - lse[:, None])
        P.scatter_add_(1, yc[:, None], -torch.ones_like(lse)[:, None])
        Gref += P.T @ Xc
        gbref += P.sum(0)
    Gref, gbref, lref = (Gref / n).cpu().numpy(), (gbref / n).cpu().numpy(), float(lref.item()) / n
    gscale = np.abs(Gref).max()
    err = np.abs(grad - Gref).max() / gscale
    print("\nfull-size multinomial gradient: max |g - g_ref| / max |g_ref| = %.3g (max |g_ref| = %.3g), intercept part "
          "%.3g, loss rel %.3g" % (err, gscale, np.abs(gb - gbref).max() / np.abs(gbref).max(), abs(loss - lref) / lref))
    assert W == n
    assert loss == pytest.approx(lref, rel=1e-5)
    assert err < 1e-5
    np.testing.assert_allclose(gb, gbref, rtol=1e-5, atol=1e-5 * np.abs(gbref).max())


def test_constant_label_short_circuits_like_mllib():
    N, G = _imports()
    from b200fit.classification import LogisticRegression
    from b200fit.feature import pack_columns
    from b200fit.frame import ColumnarFrame

    rng = np.random.default_rng(0)
    Xh = rng.normal(size=(500, 8)).astype(np.float32)
    for const, want in ((1.0, np.inf), (0.0, -np.inf)):
        yh = np.full(500, const)
        m = LogisticRegression(maxIter=10).fit(ColumnarFrame({"features": pack_columns([Xh]), "label": yh}))
        ref = oracle.lr_fit(Xh, yh.astype(np.int64), max_iter=10)
        assert m.intercept == want == ref["intercept"]
        assert not m.coefficients.any() and not ref["coefficients"].any()
        assert m.summary.objectiveHistory == [] == ref["objectiveHistory"] and m.summary.totalIterations == 0
        pred = m.transform(ColumnarFrame({"features": pack_columns([Xh])})).col("prediction").cpu().numpy()
        assert (pred == const).all()
    # without an intercept the same data trains normally (and stays finite)
    m = LogisticRegression(maxIter=5, fitIntercept=False, regParam=0.1).fit(
        ColumnarFrame({"features": pack_columns([Xh]), "label": np.ones(500)}))
    assert np.isfinite(m.coefficients).all() and m.intercept == 0.0


def test_thresholds_arrive_as_the_string_the_widget_hands_over():
    N, G = _imports()
    from b200fit.classification import LogisticRegressionModel
    from b200fit.feature import pack_columns
    from b200fit.frame import ColumnarFrame

    rng = np.random.default_rng(1)
    Xh = rng.normal(size=(300, 4)).astype(np.float32)
    coef = rng.normal(size=(3, 4))
    model = LogisticRegressionModel(coef, np.zeros(3), 3)
    frame = ColumnarFrame({"features": pack_columns([Xh])})
    _, prob, _ = oracle.lr_predict(Xh, coef, np.zeros(3))
    for th in ("[0.2, 0.5, 0.3]", "0.2,0.5,0.3", [0.2, 0.5, 0.3]):
        model._set(thresholds=th)
        pred = model.transform(frame).col("prediction").cpu().numpy()
        assert np.array_equal(pred, np.argmax(prob / np.array([0.2, 0.5, 0.3])[None, :], axis=1))
    model._set(thresholds="None")
    assert np.array_equal(model.transform(frame).col("prediction").cpu().numpy(), np.argmax(prob, axis=1))
    model._set(thresholds="[0.2, x]")
    with pytest.raises(ValueError):
        model.transform(frame)


def test_model_transform_hands_back_the_callers_frame_type():
    """pandas in -> pandas out; a pyspark-shaped frame in (toPandas + sparkSession.createDataFrame) -> the same kind
    out, with `prediction` appended and the caller's columns untouched."""
    N, G = _imports()
    import pandas as pd

    from b200fit.clustering import KMeans
    from fake_widget import FakeSparkDataFrame

    rng = np.random.default_rng(2)
    Xh = np.concatenate([rng.normal(size=(200, 3)), rng.normal(size=(200, 3)) + 8.0]).astype(np.float64)
    pdf = pd.DataFrame({"id": np.arange(400), "features": list(Xh)})
    model = KMeans(k=2, maxIter=5, initialModel=Xh[[0, 399]]).fit(pdf)
    out = model.transform(pdf)
    assert isinstance(out, pd.DataFrame) and list(out.columns) == ["id", "features", "prediction"]
    assert (out["prediction"].to_numpy()[:200] == 0).all() and (out["prediction"].to_numpy()[200:] == 1).all()
    sdf = FakeSparkDataFrame(pdf)
    out2 = model.transform(sdf)
    assert isinstance(out2, FakeSparkDataFrame) and out2.columns == ["id", "features", "prediction"]
    assert sdf.collected == 1 and out2.toPandas()["prediction"].tolist() == out["prediction"].tolist()
