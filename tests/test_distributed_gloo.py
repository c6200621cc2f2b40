"""N > 1 host logic on CPU: world_size-2 gloo process group, row shards per rank, ONE sum
all-reduce per unit, identical finalize on every rank.  The CUDA engines are
replaced by oracle-backed stand-ins with the same interface (the oracle is allowed in tests/);
what is under test is b200fit.dist, the Lloyd loop and the LR objective/optimizer plumbing."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class OracleLloydEngine:
    """Same interface as b200fit.clustering.CudaLloydEngine, numpy arithmetic from oracle/."""

    def __init__(self, Xshard, k):
        self.X, self.k, self.d = Xshard.astype(np.float64), k, Xshard.shape[1]

    def set_centers(self, C):
        self.C = np.array(C, dtype=np.float64)

    def step(self):
        import oracle

        _, cost, sums, counts = oracle.kmeans_step(self.X, self.C)
        return torch.from_numpy(np.concatenate([sums.ravel(), counts, [cost, float(len(self.X))]]))

    def finalize(self, payload, tol):
        from oracle.kmeans import kmeans_update

        p = payload.numpy()
        kd = self.k * self.d
        self.C, conv = kmeans_update(self.C, p[:kd].reshape(self.k, self.d), p[kd:kd + self.k], tol)
        return conv, float(p[kd + self.k])

    def get_centers(self):
        return self.C.copy()


class OracleLogRegEngine:
    def __init__(self, X, y, K):
        self.X, self.y, self.K, self.d = X, y, K, X.shape[1]

    def lossgrad(self, beta, icpt):
        import oracle

        n = len(self.X)
        one = np.ones(self.d)
        if self.K == 1:
            loss, g, gb = oracle.lr_binary_lossgrad(self.X, self.y, None, beta.ravel(), float(icpt[0]), one)
            g, gb = g[None, :], np.array([gb])
        else:
            loss, g, gb = oracle.lr_multinomial_lossgrad(self.X, self.y, None, beta.reshape(self.K, self.d), icpt, one)
        return torch.from_numpy(np.concatenate([g.ravel() * n, gb * n, [loss * n, float(n)]]))


def _worker(rank, world, port, out):
    for p in (ROOT, os.path.join(ROOT, "orange3-spark_b200"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle
        from b200fit.classification import make_objective
        from b200fit.clustering import lloyd
        from b200fit.dist import Comm, shard_bounds
        from b200fit.optim import LBFGS

        comm = Comm()
        assert comm.world == world and comm.rank == rank
        rng = np.random.default_rng(0)  # same global data on every rank
        n, d, k = 3001, 6, 5
        X = rng.normal(size=(n, d)) + 4.0 * rng.integers(0, 3, size=(n, 1))
        C0 = X[:k].copy()
        lo, hi = shard_bounds(n, rank, world)
        eng = OracleLloydEngine(X[lo:hi], k)
        eng.set_centers(C0)
        it, costs, conv, payload = lloyd(eng, comm, 15, 1e-6)
        ref = oracle.kmeans_lloyd(X, C0, max_iter=15, tol=1e-6)
        assert it == ref["numIter"] and conv == ref["converged"]
        np.testing.assert_allclose(eng.get_centers(), ref["centers"], rtol=1e-12, atol=1e-12)
        np.testing.assert_allclose(costs, ref["cost_hist"], rtol=1e-12)
        assert float(payload[-1]) == n  # row counts summed over shards

        y = (rng.uniform(size=n) < 1 / (1 + np.exp(-(X @ rng.normal(size=d) * 0.2)))).astype(np.float64)
        std = X.std(axis=0, ddof=1)
        inv = 1.0 / std
        f = make_objective(OracleLogRegEngine(X[lo:hi], y[lo:hi], 1), comm, 1, d, inv, np.full(d, 0.01), True,
                           np.ones(d))
        trail = LBFGS(30, 10, 1e-9).run(f, np.zeros(d + 1))
        ref_lr = oracle.lr_fit(X, y.astype(np.int64), None, reg_param=0.01, max_iter=30, tol=1e-9, return_states=True)
        x0 = np.zeros(d + 1)
        x0[d] = ref_lr["states"][0]["x"][d]
        trail = LBFGS(30, 10, 1e-9).run(f, x0)
        assert len(trail) == len(ref_lr["states"])
        np.testing.assert_allclose(trail[-1].x, ref_lr["states"][-1]["x"], rtol=1e-8, atol=1e-10)
        # an EMPTY shard on one rank: it must still join every collective of the Lloyd loop
        lo2, hi2 = (0, n) if rank == 0 else (n, n)
        eng2 = OracleLloydEngine(X[lo2:hi2], k)
        eng2.set_centers(C0)
        it2, costs2, conv2, _ = lloyd(eng2, comm, 15, 1e-6)
        assert it2 == ref["numIter"] and conv2 == ref["converged"]
        np.testing.assert_allclose(eng2.get_centers(), ref["centers"], rtol=1e-12, atol=1e-12)
        # an error condition seen by ONE rank is raised by all of them (no rank left waiting in a collective)
        assert comm.any_rank(rank == 1, device="cpu") is True and comm.any_rank(False, device="cpu") is False
        t = torch.tensor([float(rank)])
        comm.all_reduce_max(t)
        assert t.item() == world - 1
        b = torch.tensor([123.0 if rank == 0 else 0.0])
        comm.broadcast(b, 0)
        assert b.item() == 123.0
        out.put((rank, "ok"))
    except Exception as e:  # surface the failure to the parent
        import traceback

        out.put((rank, "FAIL: %s\n%s" % (e, traceback.format_exc())))
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo_matches_single_process_oracle():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = [out.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res


def test_shard_bounds_partition_rows_exactly():
    from b200fit.dist import shard_bounds

    for n in (0, 1, 7, 100, 100_000_001):
        for world in (1, 2, 3, 8):
            b = [shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
