"""Dev tool: time one LR loss+gradient pass (binary or multinomial) for a shape on one GPU."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orange3-spark_b200"))
import numpy as np
import torch

from b200fit import _native as N
from b200fit.classification import CudaLogRegEngine
from b200fit.frame import DeviceMatrix

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=4_000_000)
ap.add_argument("--d", type=int, default=512)
ap.add_argument("--K", type=int, default=20)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
L = N.lib()
dev = torch.device("cuda", 0)
rng = np.random.default_rng(0)
X = torch.empty((a.n, a.d), dtype=torch.float32, device=dev)
y = torch.empty(a.n, dtype=torch.float32, device=dev)
KM = 1 if a.K == 2 else a.K
wt = torch.from_numpy((rng.standard_normal((KM, a.d)) / np.sqrt(a.d)).astype(np.float32)).to(dev)
bt = torch.zeros(KM, dtype=torch.float32, device=dev)
N.check(L.b2f_synth_logistic(N.ptr(X), N.ptr(y), a.n, 0, a.d, a.d, 7, None, N.ptr(wt), N.ptr(bt), a.K,
                             N.stream_ptr()), "synth")
torch.cuda.synchronize()
eng = CudaLogRegEngine(DeviceMatrix(X, a.d), y, None, a.K)
eng.beta.copy_(torch.from_numpy(rng.standard_normal(KM * a.d) * 0.05).to(dev))
ts = []
for i in range(a.reps + 2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.lossgrad_device()
    e1.record()
    torch.cuda.synchronize()
    if i >= 2:
        ts.append(e0.elapsed_time(e1))
ms = sum(ts) / len(ts)
print("LR K=%d n=%d d=%d: %.3f ms  %.0f GB/s  (%.2f ms per 20M rows)" %
      (a.K, a.n, a.d, ms, a.n * (4.0 * a.d + 4) / ms / 1e6, ms * 2e7 / a.n))
