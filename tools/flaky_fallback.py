"""Dev tool: repeat the fit of tests/test_gpu_parity_r2.py::test_fit_driven_through_the_exact_fallback (every pass
overflows its lists and is redone exactly) on freshly allocated, deliberately dirtied buffers and report how often the
centres / costs differ from the first run's."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "orange3-spark_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
from b200fit import _native as N  # noqa: E402
from b200fit.clustering import CudaLloydEngine, lloyd  # noqa: E402
from b200fit.dist import Comm  # noqa: E402
from b200fit.feature import pack_columns  # noqa: E402
from oracle import synth  # noqa: E402  (dev tool: data generator only)

n, d, k, iters = 60000, 128, 256, 4
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
Xh = synth.mixture(n, 0, d, int(os.environ.get("SEED", "20240611")), 8, 2.0)
C0 = Xh[np.random.default_rng(3).choice(n, size=k, replace=False)].astype(np.float64)
import oracle  # noqa: E402
ref = oracle.kmeans_lloyd(Xh.astype(np.float64), C0, max_iter=iters, tol=0.0)
small = 256 + 5 * 1024 + 128 * (28 + 4 * d) + 4096 + 4 * n
for algo in ("KM_TCB", "KM_TCH"):
    first, bad = None, 0
    for r in range(reps):
        torch.cuda.empty_cache()
        junk = torch.randint(-2**31, 2**31 - 1, (64 << 20,), dtype=torch.int32, device="cuda")  # dirty 256 MB
        del junk
        eng = CudaLloydEngine(pack_columns([Xh]), k, algo=getattr(N, algo), ws_bytes=small)
        eng.set_centers(C0)
        it, costs, conv, payload = lloyd(eng, Comm(), iters, 0.0)
        C = eng.get_centers()
        cnt = eng.get_counts(payload)
        if first is None:
            first = (C, np.array(costs), cnt)
            print("  first run vs oracle: centres %.3g, costs rel %.3g" %
                  (np.abs(C - ref["centers"]).max(), np.abs(np.array(costs) / np.array(ref["cost_hist"]) - 1).max()))
            continue
        dd = np.abs(C - first[0]).max()
        if dd > 1e-6 or eng.fallbacks != iters:
            bad += 1
            if bad <= 5:
                print("  rep %d: centres differ by %.3g; costs rel diff %s; counts differ at %d; fallbacks %d" %
                      (r, dd, np.abs(np.array(costs) / first[1] - 1), (cnt != first[2]).sum(), eng.fallbacks))
    print("%s: %d of %d repeats differ" % (algo, bad, reps - 1))
