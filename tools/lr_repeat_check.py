"""Dev tool: run the multinomial loss+gradient pass several times on the same inputs; report run-to-run differences
(a deterministic kernel gives none) and the error against the CPU oracle."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "orange3-spark_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import oracle
from b200fit import _native as N
import gpu_util as G

for n, d, K in ((60000, 128, 5), (30000, 128, 5), (60000, 256, 5), (60000, 384, 7), (60000, 512, 20), (300000, 128, 5)):
    rng = np.random.default_rng(n + d)
    Xh = (rng.normal(size=(n, d)) * np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d))).astype(np.float32)
    yh = rng.integers(0, K, size=n).astype(np.float32)
    B = rng.normal(size=(K, d)) * 0.05
    b0 = rng.normal(size=K) * 0.1
    X, _, ld = G.upload_matrix(Xh)
    y = torch.from_numpy(yh).to(G.dev())
    outs = []
    for rep in range(4):
        loss, g, gb, W = G.lr_lossgrad_raw(X, y, None, n, d, ld, K, B, b0)
        outs.append((loss, g.copy(), gb.copy()))
    rl, rg, rgb = oracle.lr_multinomial_lossgrad(Xh, yh, None, B, b0, np.ones(d))
    sc = np.abs(rg).max()
    rr = max(np.abs(o[1] - outs[0][1]).max() for o in outs[1:]) / sc
    print("n=%d d=%d K=%d: run-to-run max |dg|/max|g| = %.3g, loss diff %.3g; vs oracle: grad %.3g, icpt %.3g, loss %.3g" % (
        n, d, K, rr, max(abs(o[0] - outs[0][0]) for o in outs[1:]) / abs(rl),
        np.abs(outs[0][1] - rg).max() / sc, np.abs(outs[0][2] - rgb).max() / np.abs(rgb).max(), abs(outs[0][0] - rl) / abs(rl)))
