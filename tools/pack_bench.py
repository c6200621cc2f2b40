"""Dev tool: time b2f_pack_columns (K1) on device-resident fp32 columns; prints ms and GB/s (read + write)."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orange3-spark_b200"))
from b200fit import _native as N  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=20_000_000)
ap.add_argument("--d", type=int, default=128)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
L = N.lib()
dev = torch.device("cuda:0")
n, d = a.n, a.d
cols = torch.empty((d, n), dtype=torch.float32, device=dev)
for j in range(d):
    cols[j].copy_(torch.arange(n, device=dev, dtype=torch.float32).mul_(1e-6).add_(float(j)))
X = torch.empty((n, d), dtype=torch.float32, device=dev)
ptrs = torch.tensor([cols[j].data_ptr() for j in range(d)], dtype=torch.int64, device=dev)
codes = torch.zeros(d, dtype=torch.int32, device=dev)
bad = torch.zeros(1, dtype=torch.int32, device=dev)


def once():
    N.check(L.b2f_pack_columns(N.ptr(ptrs), N.ptr(codes), d, n, N.ptr(X), d, N.ptr(bad), N.stream_ptr()), "pack")


for _ in range(3):
    once()
torch.cuda.synchronize()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(a.reps):
    once()
t1.record()
torch.cuda.synchronize()
ms = t0.elapsed_time(t1) / a.reps
ok = torch.equal(X[12345 % n], cols[:, 12345 % n]) and torch.equal(X[n - 1], cols[:, n - 1]) and int(bad.item()) == 0
print("pack n=%d d=%d: %.3f ms  %.0f GB/s  ok=%s" % (n, d, ms, n * d * 8.0 / ms / 1e6, ok))
