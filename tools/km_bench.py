"""Dev tool: time b2f_kmeans_step alone (CUDA events) for a given shape / algo on one GPU."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orange3-spark_b200"))
import numpy as np
import torch

from b200fit import _native as N

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=20_000_000)
ap.add_argument("--d", type=int, default=128)
ap.add_argument("--k", type=int, default=1024)
ap.add_argument("--algo", type=int, default=2)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--per-row", action="store_true", help="plain b2f_kmeans_prepare (per-row key scales) instead of "
                                                         "b2f_kmeans_prepare_x with the data's largest row norm")
ap.add_argument("--predict", action="store_true", help="labels only (no accumulation)")
ap.add_argument("--predict-ws", action="store_true", help="labels only through the certainty pipeline (filter pass "
                                                            "without accumulation)")
ap.add_argument("--lloyd", type=int, default=0, help="run this many real Lloyd iterations first (centres move)")
a = ap.parse_args()
L = N.lib()
dev = torch.device("cuda", 0)
X = torch.empty((a.n, a.d), dtype=torch.float32, device=dev)
N.check(L.b2f_synth_mixture(N.ptr(X), a.n, 0, a.d, a.d, 20240611, a.k, 2.0, None, N.stream_ptr()), "synth")
idx = torch.from_numpy(np.random.default_rng(0).choice(a.n, a.k, replace=False)).to(dev)
C = X[idx].double().contiguous()
prep = torch.empty(L.b2f_kmeans_prep_bytes(a.k, a.d), dtype=torch.uint8, device=dev)
acc = torch.empty(L.b2f_kmeans_acc_bytes(a.k, a.d), dtype=torch.uint8, device=dev)
xm = torch.zeros(1, dtype=torch.float32, device=dev)
if not a.per_row:
    N.check(L.b2f_row_norm_max(N.ptr(X), a.n, a.d, a.d, None, N.ptr(xm), N.stream_ptr()), "row_norm_max")
XMAX = float(xm.item())


def prepare():
    N.check(L.b2f_kmeans_prepare_x(N.ptr(C), a.k, a.d, XMAX, N.ptr(prep), N.stream_ptr()), "prep")


prepare()
lab = torch.empty(a.n, dtype=torch.int32, device=dev)
wsb = int(L.b2f_kmeans_ws_bytes(a.n, a.k, a.d))
ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
status = torch.zeros(2, dtype=torch.int32, device=dev)
status_f64 = torch.zeros(2, dtype=torch.float64, device=dev)
for i in range(a.lloyd):
    N.check(L.b2f_kmeans_acc_zero(N.ptr(acc), a.k, a.d, N.stream_ptr()), "zero")
    N.check(L.b2f_kmeans_step_ws(N.ptr(X), a.n, a.d, a.d, None, N.ptr(prep), a.k, N.ptr(acc), N.ptr(lab),
                                 N.ptr(ws), wsb, 0, N.stream_ptr()), "step_ws")
    N.check(L.b2f_kmeans_fold(N.ptr(acc), a.k, a.d, N.stream_ptr()), "fold")
    N.check(L.b2f_kmeans_finalize(N.ptr(acc), a.k, a.d, 0.0, N.ptr(C), N.ptr(status), N.ptr(status_f64),
                                  N.stream_ptr()), "finalize")
    prepare()
ts = []
for i in range(a.reps + 2):
    N.check(L.b2f_kmeans_acc_zero(N.ptr(acc), a.k, a.d, N.stream_ptr()), "zero")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if a.predict_ws:
        N.check(L.b2f_kmeans_predict_ws(N.ptr(X), a.n, a.d, a.d, None, N.ptr(prep), a.k, N.ptr(lab), None, N.ptr(ws),
                                        wsb, a.algo, N.stream_ptr()), "predict_ws")
    elif a.predict:
        N.check(L.b2f_kmeans_predict(N.ptr(X), a.n, a.d, a.d, N.ptr(prep), a.k, N.ptr(lab), None, a.algo,
                                     N.stream_ptr()), "predict")
    elif a.algo in (4, 5):
        N.check(L.b2f_kmeans_step_ws(N.ptr(X), a.n, a.d, a.d, None, N.ptr(prep), a.k, N.ptr(acc), N.ptr(lab),
                                     N.ptr(ws), wsb, a.algo, N.stream_ptr()), "step_ws")
    else:
        N.check(L.b2f_kmeans_step(N.ptr(X), a.n, a.d, a.d, None, N.ptr(prep), a.k, N.ptr(acc), None, a.algo,
                                  N.stream_ptr()), "step")
    e1.record()
    torch.cuda.synchronize()
    if i >= 2:
        ts.append(e0.elapsed_time(e1))
ms = sum(ts) / len(ts)
fl = 2.0 * a.n * a.k * a.d
if a.algo in (4, 5):
    print("uncertain rows: %s of %d" % (ws[:12].view(torch.int32).cpu().tolist(), a.n))
print("algo %d n=%d d=%d k=%d: %.3f ms  %.1f TFLOP/s algorithmic  (%.1f ms per 100M rows)" %
      (a.algo, a.n, a.d, a.k, ms, fl / ms / 1e9, ms * 1e8 / a.n))
