"""N > 1 on real GPUs (SURVEY.md 8(e), section 4 "1-GPU == N-GPU"): two NCCL ranks, contiguous row shards, one
all-reduce per unit.  A 2-rank fit must give the counts of the 1-rank fit exactly and its centres / objective history
to 1e-7 (the sums are added in a different order, so bit equality is not on offer); an error condition in ONE shard
must raise on every rank instead of leaving the other blocked in a collective; a rank with an empty shard must
stay in step.  Skipped when fewer than two GPUs are visible."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 20240611


def _fit_all(torch_dev, rank, world, Xh, yb, ym, C0, empty_rank=None):
    """KMeans + binary LR + multinomial LR on this rank's shard; returns plain numpy results."""
    from b200fit.classification import LogisticRegression
    from b200fit.clustering import KMeans
    from b200fit.dist import shard_bounds
    from b200fit.feature import pack_columns
    from b200fit.frame import ColumnarFrame

    n = len(Xh)
    if empty_rank is None:
        lo, hi = shard_bounds(n, rank, world)
    else:
        lo, hi = (n, n) if rank == empty_rank else (0, n)
    X = pack_columns([Xh[lo:hi]], row0=lo) if hi > lo else None
    if X is None:
        from b200fit.frame import DeviceMatrix

        X = DeviceMatrix(torch.empty((0, Xh.shape[1]), dtype=torch.float32, device=torch_dev), Xh.shape[1], lo)
    km = KMeans(k=len(C0), maxIter=5, tol=0.0, initialModel=C0).fit(ColumnarFrame({"features": X}))
    out = {"centers": np.stack(km.clusterCenters()), "sizes": np.array(km.summary.clusterSizes),
           "costs": np.array(km.summary.costHistory)}
    lrb = LogisticRegression(maxIter=12, regParam=0.01).fit(ColumnarFrame({"features": X, "label": yb[lo:hi]}))
    out["b_coef"], out["b_icpt"] = lrb.coefficients.copy(), lrb.intercept
    out["b_hist"] = np.array(lrb.summary.objectiveHistory)
    lrm = LogisticRegression(maxIter=8, regParam=0.01, elasticNetParam=0.5, family="multinomial").fit(
        ColumnarFrame({"features": X, "label": ym[lo:hi]}))
    out["m_coef"], out["m_icpt"] = lrm.coefficientMatrix.copy(), lrm.interceptVector.copy()
    out["m_hist"] = np.array(lrm.summary.objectiveHistory)
    return out


def _data():
    from oracle import synth  # the shared Philox generator (test infrastructure)

    n, d, k = 60000, 128, 1024
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    C0 = Xh[np.random.default_rng(2).choice(n, size=k, replace=False)].astype(np.float64)
    rng = np.random.default_rng(3)
    wt = rng.normal(size=d) / np.sqrt(d)
    yb = (rng.uniform(size=n) < 1.0 / (1.0 + np.exp(-(Xh @ wt)))).astype(np.float64)
    ym = np.argmax(Xh @ rng.normal(size=(d, 5)) / np.sqrt(d) + rng.gumbel(size=(n, 5)), axis=1).astype(np.float64)
    return Xh, yb, ym, C0


def _worker(rank, world, port, out):
    for p in (ROOT, os.path.join(ROOT, "orange3-spark_b200"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        from b200fit.classification import LogisticRegression
        from b200fit.dist import shard_bounds
        from b200fit.feature import pack_columns
        from b200fit.frame import ColumnarFrame

        Xh, yb, ym, C0 = _data()
        res = {"sharded": _fit_all(dev, rank, world, Xh, yb, ym, C0)}
        res["empty"] = _fit_all(dev, rank, world, Xh, yb, ym, C0, empty_rank=1)
        # one bad label in rank 1's shard: BOTH ranks must raise (and then still be able to talk to each other)
        lo, hi = shard_bounds(len(Xh), rank, world)
        ybad = yb[lo:hi].copy()
        if rank == 1:
            ybad[5] = -1.0
        try:
            LogisticRegression(maxIter=2).fit(ColumnarFrame({"features": pack_columns([Xh[lo:hi]]), "label": ybad}))
            res["bad_label"] = "no error"
        except ValueError as e:
            res["bad_label"] = "ValueError: %s" % e
        t = torch.ones(1, device=dev)
        dist.all_reduce(t)
        res["alive_after_error"] = float(t.item())
        out.put((rank, res))
    except Exception as e:
        import traceback

        out.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


def test_two_rank_fits_equal_the_one_rank_fits():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    for p in (ROOT, os.path.join(ROOT, "orange3-spark_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    torch.cuda.set_device(0)
    Xh, yb, ym, C0 = _data()
    one = _fit_all(torch.device("cuda", 0), 0, 1, Xh, yb, ym, C0)
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29600 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(out.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for rank in (0, 1):
        assert isinstance(got[rank], dict), got[rank]
        assert got[rank]["bad_label"].startswith("ValueError"), got[rank]["bad_label"]
        assert got[rank]["alive_after_error"] == 2.0
        for tag in ("sharded", "empty"):
            two = got[rank][tag]
            assert np.array_equal(two["sizes"], one["sizes"]), "%s: cluster sizes differ" % tag
            sc = np.abs(one["centers"]).max()
            np.testing.assert_allclose(two["centers"], one["centers"], rtol=1e-7, atol=1e-7 * sc)
            np.testing.assert_allclose(two["costs"], one["costs"], rtol=1e-7)
            for key in ("b_hist", "m_hist"):
                assert len(two[key]) == len(one[key]), "%s %s: different number of optimizer states" % (tag, key)
                np.testing.assert_allclose(two[key], one[key], rtol=1e-7)
            np.testing.assert_allclose(two["b_coef"], one["b_coef"], rtol=1e-5, atol=1e-7 * np.abs(one["b_coef"]).max())
            np.testing.assert_allclose(two["m_coef"], one["m_coef"], rtol=1e-5, atol=1e-6 * np.abs(one["m_coef"]).max())
    # ranks of one run hold identical models (no root, no broadcast: every rank ran the same finalize)
    np.testing.assert_array_equal(got[0]["sharded"]["centers"], got[1]["sharded"]["centers"])
    np.testing.assert_array_equal(got[0]["sharded"]["b_coef"], got[1]["sharded"]["b_coef"])
