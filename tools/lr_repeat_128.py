"""Dev tool: the d = 128 multinomial determinism probe alone (see tools/lr_repeat_check.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "orange3-spark_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

from b200fit import _native as N
import gpu_util as G

n, d, K = 60000, 128, 5
rng = np.random.default_rng(n + d)
Xh = (rng.normal(size=(n, d)) * np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d))).astype(np.float32)
yh = rng.integers(0, K, size=n).astype(np.float32)
B = rng.normal(size=(K, d)) * 0.05
b0 = rng.normal(size=K) * 0.1
X, _, ld = G.upload_matrix(Xh)
y = torch.from_numpy(yh).to(G.dev())
outs = [G.lr_lossgrad_raw(X, y, None, n, d, ld, K, B, b0) for _ in range(6)]
sc = np.abs(outs[0][1]).max()
print(os.environ.get("B2F_LIB_PATH", "product"), "run-to-run max |dg|/max|g| = %.3g  loss diffs %s" % (
    max(np.abs(o[1] - outs[0][1]).max() for o in outs[1:]) / sc, ["%.2e" % abs(o[0] - outs[0][0]) for o in outs[1:]]))
