#!/usr/bin/env python
"""bench.py -- BASELINE.json metric: KMeans Lloyd iterations/s on 100M x 128 fp32, k = 1024
(configs[1]) plus LR loss+gradient passes/s on 50M x 256 (configs[2]), rows sharded over N B200s.

  python bench.py --gpus N --steps K --warmup W          # this repo's CUDA path (one JSON line)
  python bench.py --impl reference ...                   # CPU restatement on the host cores

A "step" is one full Lloyd iteration: assignment + accumulation kernel, fold, all-reduce of
[sums|counts|cost], centre update and operand re-preparation.  Inputs are synthetic
(Philox, generated on device, seed 20240611) and far larger than L2 (51.2 GB vs 126 MB).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "orange3-spark_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

SEED = 20240611
KM = dict(n=100_000_000, d=128, k=1024, comps=1024, center_scale=2.0)
LR = dict(n=50_000_000, d=256)
LRM = dict(n=20_000_000, d=512, K=20)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=0, help="DEV ONLY: override the row count (marks the line reduced)")
    ap.add_argument("--lr-rows", type=int, default=0, help="DEV ONLY: override the LR row count")
    ap.add_argument("--lrm-rows", type=int, default=0, help="DEV ONLY: override the multinomial LR row count")
    ap.add_argument("--algo", type=int, default=0, help="B2F_KM_* code (0 = auto)")
    ap.add_argument("--no-e2e", action="store_true", help="DEV ONLY: skip the host-buffer end-to-end leg")
    ap.add_argument("--no-lr", action="store_true", help="DEV ONLY: skip the LR leg")
    ap.add_argument("--no-cpu", action="store_true", help="DEV ONLY: skip the cpu_baseline leg")
    return ap.parse_args()


TRAFFIC_FILE = os.path.join("profiles", "r2_traffic.json")


def measured_traffic(kernel, rows):
    """(DRAM bytes per launch, where the figure comes from).  ncu cannot run inside a timed bench, so the figure is
    the one of the committed `ncu --set full` capture of this kernel, quoted only when this launch has the captured
    size -- and labelled as such in the JSON line (`traffic_source`), never presented as measured in this run."""
    try:
        with open(os.path.join(ROOT, TRAFFIC_FILE)) as f:
            e = json.load(f)[kernel]
        if e["rows"] == rows:
            return e["traffic"], "committed ncu capture %s (%s), not measured in this run" % (TRAFFIC_FILE, e.get("source", kernel))
    except Exception:
        pass
    return None, None


def with_traffic(roofline, kernel, rows):
    roofline["traffic"], roofline["traffic_source"] = measured_traffic(kernel, rows)
    return roofline


def cpu_kmeans_sample(n_total, steps, warmup=1, n_s=400_000):
    """THE CPU routine of both CPU arms (cpu_baseline and --impl reference): the oracle's Lloyd iteration
    (assignment + accumulation + centre update) on rows [0, n_s) of the SAME Philox data the GPU arm generates on
    the device (oracle/synth.py mirrors csrc/synth.cu bit for bit), fixed seeded initial centres, all host threads
    through numpy/OpenBLAS.  Returns (seconds per iteration on the sample, description)."""
    import numpy as np

    import oracle
    from oracle import synth

    n_s = min(n_s, n_total)
    X = synth.mixture(n_s, 0, KM["d"], SEED, KM["comps"], KM["center_scale"]).astype(np.float64)
    C = X[np.random.default_rng(SEED).choice(n_s, KM["k"], replace=False)].copy()
    for _ in range(warmup):
        oracle.kmeans_step(X[:50_000], C)
    t0 = time.perf_counter()
    for _ in range(steps):
        _, _, sums, counts = oracle.kmeans_step(X, C)
        C, _ = oracle.kmeans.kmeans_update(C, sums, counts, 0.0)
    dt = (time.perf_counter() - t0) / steps
    what = ("rows [0, %d) of the %d Philox rows (seed %d, same generator as the device), %d Lloyd iterations "
            "(assign + accumulate + update), linear extrapolation in N; numpy/OpenBLAS fp64 CPU restatement of MLlib "
            "(Spark local[N]: unavailable on this box)" % (n_s, n_total, SEED, steps))
    return dt, n_s, what


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return dict(hbm=p["hbm_gbs"], tensor=p["bf16_tflops_sustained"], tensor_burst=p["bf16_tflops"], src="measured")
    except Exception:
        return dict(hbm=6650.0, tensor=1400.0, tensor_burst=1590.0, src="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([s.strip() for s in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for nm, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def reference_arm(args):
    """CPU restatement (oracle, numpy/OpenBLAS on all host threads) of one Lloyd iteration on a
    bounded sample of the same workload, extrapolated linearly in N (every unit is linear in N)."""
    import numpy as np

    import oracle

    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    n_total = args.rows or KM["n"]
    dt, n_s, sample = cpu_kmeans_sample(n_total, args.steps, warmup=max(1, args.warmup))
    value = 1.0 / (dt * n_total / n_s)
    line = {"impl": "reference", "metric": "kmeans_lloyd_iter_per_s", "value": value, "unit": "iter/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 * n_total / n_s,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(n_total, args.gpus),
            "cpu_baseline": {"value": value, "unit": "iter/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(n_total, gpus):
    return {"workload": "KMeans k=1024 maxIter=20 on synthetic %dx128 fp32, fixed initialModel (BASELINE configs[1])"
                        % n_total, "rows": n_total, "d": KM["d"], "k": KM["k"], "parallelism": "rows/%d" % gpus,
            "l2": "inputs (%.1f GB per GPU) far exceed the 126 MB L2; no flush needed"
                  % (n_total * KM["d"] * 4 / gpus / 1e9)}


def b200_arm(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from b200fit import _native as N
    from b200fit._kmeans_init import init_random
    from b200fit.classification import CudaLogRegEngine
    from b200fit.clustering import CudaLloydEngine, KMeans
    from b200fit.dist import Comm, shard_bounds
    from b200fit.feature import VectorAssembler
    from b200fit.frame import ColumnarFrame, DeviceMatrix

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    from b200fit.dist import bind_to_gpu_numa_node

    numa = bind_to_gpu_numa_node(local) if world > 1 else None  # before any pinned allocation
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    comm = Comm()
    L = N.lib()
    pk = peaks()
    n_total = args.rows or KM["n"]
    d, k = KM["d"], KM["k"]
    lo, hi = shard_bounds(n_total, rank, world)
    n = hi - lo

    def sync_all():
        comm.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        comm.all_reduce_max(t)
        return float(t.item())

    # ---------------- KMeans: device-resident inputs ----------------
    Xt = torch.empty((n, d), dtype=torch.float32, device=dev)
    N.check(L.b2f_synth_mixture(N.ptr(Xt), n, lo, d, d, SEED, KM["comps"], KM["center_scale"], None, N.stream_ptr()),
            "b2f_synth_mixture")
    X = DeviceMatrix(Xt, d, lo)
    C0 = init_random(X, n_total, k, SEED, comm)  # 1024 seeded distinct data rows
    eng = CudaLloydEngine(X, k, algo=args.algo)
    eng.set_centers(C0)

    def iteration(ev=None):
        payload = eng.step(events=ev)
        comm.all_reduce_sum(payload)
        eng.finalize_async(payload, 0.0)

    for _ in range(args.warmup):
        iteration()
    sync_all()
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = eng.launches
    with ClockSampler(local) as clk:
        sync_all()
        t0.record()
        for i in range(args.steps):
            iteration(kev[i])
        t1.record()
        sync_all()
    ms_step = max_over_ranks(t0.elapsed_time(t1)) / args.steps
    kern_ms = sum(a.elapsed_time(b) for a, b in kev) / args.steps
    launches = eng.launches - launches0
    cost_after = float(eng.status_f64.cpu()[1].item())
    rows_after = eng.rows_counted(eng.payload)
    hdr = eng.ws[:20].view(torch.int32).cpu().numpy() if eng.use_ws else [0] * 5
    uncertain, pair_rows = int(hdr[0]), min(int(hdr[4]), int(hdr[3]))
    algo_name = ("fp16 filter + candidate check + 3xTF32 re-check + fp64 tail (B2F_KM_TCH)" if eng.use_ws and args.algo in (0, 5) else "bf16x3 filter + 3xTF32 re-check (B2F_KM_TCB)" if eng.use_ws else "algo %d" % args.algo)
    value = 1e3 / ms_step
    flops = 2.0 * n * k * d
    achieved = flops / (kern_ms * 1e-3) / 1e12
    line = {"metric": "kmeans_lloyd_iter_per_s", "value": value, "unit": "iter/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None,
            "dtype": ("f16 filter + tf32x3 re-check + f64 tail -> f32 sums (folded in f64)"
                      if eng.use_ws and args.algo in (0, 5) else
                      "bf16x3 filter + tf32x3 re-check -> f32 (sums/cost folded in f64)") if eng.use_ws else
                     "tf32x3->f32 (sums/cost folded in f64)",
            "data": "synthetic", "config": workload_config(n_total, world), "gpu_launches": launches,
            "clocks": clk.summary(),
            "roofline": with_traffic({"bound": "tensor", "achieved": achieved, "peak": pk["tensor"], "unit": "TFLOP/s",
                                      "frac": achieved / pk["tensor"],
                                      "kernel": "b2f_kmeans_step_ws: " + algo_name,
                                      "kernel_ms": kern_ms,
                                      "peak_source": pk["src"] + " bf16 sustained (no TF32 figure measured)",
                                      "algorithmic_flops_per_launch": flops},
                                     "km_tcb_kernel" if eng.use_ws else "none", n),
            "check": {"cost_after_last_timed_iter": cost_after, "rows_accumulated_last_iter_this_rank": rows_after,
                      "rows_this_rank": n, "rechecked_rows_last_iter_this_rank": uncertain,
                      "pair_checked_rows_last_iter_this_rank": pair_rows}}
    if args.rows:
        line["reduced"] = True
    if numa:
        line["numa_binding_rank0"] = numa

    # ---------------- end to end through the public API with HOST buffers ----------------
    if not args.no_e2e:
        line["e2e"] = e2e_leg(args, torch, Xt, C0, n, d, k, comm, ColumnarFrame, VectorAssembler, KMeans,
                              max_over_ranks, sync_all)
    del eng, X, Xt
    torch.cuda.empty_cache()

    # ---------------- the other legs (a leg that fails is reported as such; the headline line is still printed) ----
    def leg(name, fn, *a):
        try:
            line[name] = fn(*a)
        except Exception as e:  # noqa: BLE001
            line[name] = {"error": "%s: %s" % (type(e).__name__, str(e).splitlines()[0] if str(e) else "")}
        torch.cuda.empty_cache()

    if not args.no_lr:
        # LR binary (configs[2]) / multinomial (configs[3]): one fused loss+gradient pass per step
        leg("lr", lr_leg, args, torch, N, L, comm, rank, world, dev, pk, CudaLogRegEngine, DeviceMatrix, shard_bounds,
            max_over_ranks, sync_all)
        leg("lr_multinomial", lrm_leg, args, torch, N, L, comm, rank, world, dev, pk, CudaLogRegEngine, DeviceMatrix,
            shard_bounds, max_over_ranks, sync_all)
        leg("pipeline_cfg5", cfg5_leg, args, torch, N, L, comm, rank, world, dev, pk, DeviceMatrix, ColumnarFrame,
            shard_bounds, max_over_ranks, sync_all)
        leg("loader", loader_leg, args, torch, N, L, dev, pk, rank, world, max_over_ranks, sync_all)
        leg("cfg1", cfg1_leg, args, torch, comm, rank, max_over_ranks, sync_all)
    if rank == 0 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(n_total)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def e2e_leg(args, torch, Xt, C0, n, d, k, comm, ColumnarFrame, VectorAssembler, KMeans, max_over_ranks, sync_all):
    """Same metric through the reference-facing call chain with HOST columns in pinned memory:
    VectorAssembler.transform (H2D + pack) -> KMeans(maxIter=M, initialModel).fit -> centres to host."""
    import psutil

    iters = 20
    n_full = n
    avail = psutil.virtual_memory().available
    if n * d * 4 * 1.25 > avail:  # never drive the box out of host memory: shrink and SAY so
        n = int(avail * 0.6 / (d * 4))
    cols = {}
    for j in range(d):
        h = torch.empty(n, dtype=torch.float32, pin_memory=True)
        h.copy_(Xt[:n, j])
        cols["f%d" % j] = h
    torch.cuda.synchronize()
    df = ColumnarFrame(cols)
    va = VectorAssembler(inputCols=["f%d" % j for j in range(d)], outputCol="features", handleInvalid="keep")
    km = KMeans(k=k, maxIter=iters, tol=0.0, initialModel=C0)
    km._algo = args.algo
    sync_all()
    t0 = time.perf_counter()
    model = km.fit(va.transform(df))
    centers = model.clusterCenters()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=Xt.device)
    comm.all_reduce_max(t)
    dt = float(t.item())
    assert len(centers) == k and model.summary.numIter == iters
    hist = model.summary.costHistory
    # full-size invariants: the Lloyd cost never increases, every row is in exactly one cluster
    monotone = all(b <= a * (1.0 + 1e-7) for a, b in zip(hist, hist[1:]))
    sizes = torch.tensor([float(sum(model.summary.clusterSizes))], dtype=torch.float64, device=Xt.device)
    rows = torch.tensor([float(n)], dtype=torch.float64, device=Xt.device)
    comm.all_reduce_sum(rows)
    scale = n / float(n_full)  # 1.0 unless host RAM forced a smaller e2e sample
    return {"value": iters / dt * scale, "unit": "iter/s", "rows_used": n, "rows_full": n_full,
            "host_ram_limited": n != n_full, "h2d_bytes_per_step": n * d * 4 / iters,
            "d2h_bytes_per_step": (k * d * 8 + 24 * iters) / iters, "iterations": iters, "seconds": dt,
            "check": {"cost_non_increasing_over_the_fit": bool(monotone), "cost_first": hist[0], "cost_last": hist[-1],
                      "cluster_sizes_sum_to_rows": bool(float(sizes.item()) == float(rows.item()))},
            "what": "VectorAssembler.transform(host pinned columns) + KMeans.fit(maxIter=20, tol=0) + clusterCenters(); "
                    "the one H2D of X is amortised over the fit's 20 iterations"}


def lr_leg(args, torch, N, L, comm, rank, world, dev, pk, CudaLogRegEngine, DeviceMatrix, shard_bounds,
           max_over_ranks, sync_all):
    import numpy as np

    n_total = args.lr_rows or (LR["n"] if not args.rows else max(args.rows // 2, 1))
    d = LR["d"]
    lo, hi = shard_bounds(n_total, rank, world)
    n = hi - lo
    rng = np.random.default_rng(SEED)
    scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d)).astype(np.float32)
    w_true = (rng.standard_normal(d) / np.sqrt(d)).astype(np.float32)
    Xt = torch.empty((n, d), dtype=torch.float32, device=dev)
    y = torch.empty(n, dtype=torch.float32, device=dev)
    sc_d, w_d = torch.from_numpy(scale).to(dev), torch.from_numpy(w_true).to(dev)  # keep alive across the call
    b_d = torch.tensor([0.3], dtype=torch.float32, device=dev)
    N.check(L.b2f_synth_logistic(N.ptr(Xt), N.ptr(y), n, lo, d, d, SEED + 1, N.ptr(sc_d), N.ptr(w_d), N.ptr(b_d), 2,
                                 N.stream_ptr()), "b2f_synth_logistic")
    torch.cuda.synchronize()
    eng = CudaLogRegEngine(DeviceMatrix(Xt, d, lo), y, None, 2)
    beta = (rng.standard_normal(d) * 0.05 / scale)
    eng.beta.copy_(torch.from_numpy(beta).to(dev))

    def one_pass():
        comm.all_reduce_sum(eng.lossgrad_device())

    for _ in range(args.warmup):
        one_pass()
    sync_all()
    steps = max(args.steps, 5)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        one_pass()
    t1.record()
    sync_all()
    ms = max_over_ranks(t0.elapsed_time(t1)) / steps
    p = eng.out.cpu().numpy()
    bytes_pass = n * (4.0 * d + 4.0)
    gbs = bytes_pass / (ms * 1e-3) / 1e9
    e2e = None if args.no_e2e else lr_e2e(args, torch, Xt, y, n, d, comm, max_over_ranks, sync_all)
    return {"metric": "lr_binary_lossgrad_pass_per_s", "value": 1e3 / ms, "unit": "pass/s", "ms_per_step": ms,
            "config": {"workload": "binary LogisticRegression loss+gradient pass on synthetic %dx256 fp32 "
                                   "(BASELINE configs[2])" % n_total, "rows": n_total, "d": d},
            "roofline": with_traffic({"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                                      "frac": gbs / pk["hbm"], "kernel": "lr_binary_kernel",
                                      "algorithmic_bytes_per_launch": bytes_pass, "peak_source": pk["src"]},
                                     "lr_binary_kernel", n),
            "e2e": e2e,
            "cpu_baseline": cpu_lr_baseline(n_total, d, 1, SEED + 1, SEED) if (rank == 0 and not args.no_cpu) else None,
            "check": {"loss": float(p[d + 1] / p[d + 2]), "weight_sum": float(p[d + 2])}}


def lr_e2e(args, torch, Xt, y, n, d, comm, max_over_ranks, sync_all):
    """Binary LR end to end through the public API with HOST columns: VectorAssembler.transform (H2D + pack) ->
    LogisticRegression(regParam=0.01, maxIter=10).fit -> coefficients on the host.  Optimizer, per-pass host<->device
    traffic and the summarizer pass are all inside the timed region; the figure is loss+gradient passes per second."""
    import psutil

    from b200fit.classification import LogisticRegression
    from b200fit.feature import VectorAssembler
    from b200fit.frame import ColumnarFrame

    n_full = n
    try:
        torch._C._host_emptyCache()  # pinned blocks of the KMeans end-to-end leg (other sizes: not reusable here)
    except Exception:
        pass
    avail = psutil.virtual_memory().available
    if n * d * 4 * 1.25 > avail:
        n = int(avail * 0.6 / (d * 4))
    cols = {}
    for j in range(d):
        h = torch.empty(n, dtype=torch.float32, pin_memory=True)
        h.copy_(Xt[:n, j])
        cols["f%d" % j] = h
    lab = torch.empty(n, dtype=torch.float32, pin_memory=True)
    lab.copy_(y[:n])
    cols["label"] = lab
    torch.cuda.synchronize()
    df = ColumnarFrame(cols)
    va = VectorAssembler(inputCols=["f%d" % j for j in range(d)], outputCol="features", handleInvalid="keep")
    lr = LogisticRegression(regParam=0.01, maxIter=10, tol=0.0)
    sync_all()
    t0 = time.perf_counter()
    model = lr.fit(va.transform(df))
    coef = model.coefficients
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device=Xt.device)
    comm.all_reduce_max(t)
    dt = float(t.item())
    passes = int(model.summary.lossGradPasses)
    assert len(coef) == d and passes >= 1
    scale = n / float(n_full)
    return {"value": passes / dt * scale, "unit": "pass/s", "passes": passes, "iterations": model.summary.totalIterations,
            "seconds": dt, "rows_used": n, "rows_full": n_full, "host_ram_limited": n != n_full,
            "h2d_bytes_per_step": (n * (d + 1) * 4 + passes * (d + 1) * 8) / passes,
            "d2h_bytes_per_step": (d + 3) * 8,
            "what": "VectorAssembler.transform(host pinned columns) + LogisticRegression(regParam=0.01, maxIter=10).fit: "
                    "one H2D of X and y, the summarizer pass, then per pass beta up / payload down and the host L-BFGS step"}


def cfg1_leg(args, torch, comm, rank, max_over_ranks, sync_all):
    """BASELINE configs[0]: 1M x 16, k = 8 through the Dataset Builder -> widget path, timed as a user runs it:
    host columns -> VectorAssembler.transform -> KMeans.fit(df, params=<the widget's param map>) -> model.transform,
    the parameter map built from TEXT exactly as the GUI builds it (tests/fake_widget.py replays
    orangecontrib/spark/base/spark_ml_transformer.py:117-123).  The CPU oracle runs the same fit at FULL size beside it
    (this is the one config the reference's own CPU path can run)."""
    import numpy as np

    from oracle import synth
    import oracle

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import fake_widget as fw

    from b200fit.clustering import KMeans
    from b200fit.feature import VectorAssembler
    from b200fit.frame import ColumnarFrame

    n, d, k, iters = 1_000_000, 16, 8, 20
    if comm.world > 1:
        return None  # a 64 MB problem for ONE device (the fit would otherwise join this job's collectives)
    Xh = synth.mixture(n, 0, d, SEED + 10, k, 3.0)
    C0 = Xh[np.random.default_rng(SEED).choice(n, k, replace=False)].astype(np.float64)
    cols = {"f%d" % j: torch.from_numpy(np.ascontiguousarray(Xh[:, j])).pin_memory() for j in range(d)}
    va = VectorAssembler(inputCols=list(cols), outputCol="features")

    def once():
        t0 = time.perf_counter()
        df = va.transform(ColumnarFrame(cols))
        model = fw.press_apply(KMeans, df, edits={"k": str(k), "maxIter": str(iters), "tol": "1e-4"},
                               objects={"initialModel": C0})
        lab = model.transform(df).col("prediction").cpu()
        torch.cuda.synchronize()
        return time.perf_counter() - t0, model, lab

    once()
    dts = []
    for _ in range(3):
        dt, model, lab = once()
        dts.append(dt)
    dt = sorted(dts)[1]
    t0 = time.perf_counter()
    ref = oracle.kmeans_lloyd(Xh.astype(np.float64), C0, max_iter=iters, tol=1e-4)
    cpu_dt = time.perf_counter() - t0
    it = model.summary.numIter
    same = it == ref["numIter"] and np.allclose(np.stack(model.clusterCenters()), ref["centers"], rtol=1e-5, atol=1e-5)
    return {"metric": "kmeans_lloyd_iter_per_s", "value": it / dt, "unit": "iter/s", "seconds": dt, "iterations": it,
            "config": {"workload": "KMeans k=8 maxIter=20 tol=1e-4 on synthetic 1000000x16 through VectorAssembler -> "
                                   "widget param map -> fit -> transform, host columns (BASELINE configs[0])",
                       "rows": n, "d": d, "k": k},
            "note": "launch- and host-latency bound (6.4e7 bytes, 2.6e8 flop per iteration): a parity config, not a "
                    "roofline target",
            "cpu_baseline": {"value": ref["numIter"] / cpu_dt, "unit": "iter/s", "cores": os.cpu_count() or 1,
                             "kind": "port", "sample": "FULL size: the same fit by the oracle (numpy/OpenBLAS), %d "
                                                       "iterations in %.2f s" % (ref["numIter"], cpu_dt)},
            "check": {"centres_match_oracle_1e-5": bool(same), "iterations_gpu": it, "iterations_oracle": ref["numIter"],
                      "label_histogram": np.bincount(lab.numpy().astype(np.int64), minlength=k).tolist()}}


def loader_leg(args, torch, N, L, dev, pk, rank, world, max_over_ranks, sync_all):
    """K1 alone: b2f_pack_columns on device-resident fp32 columns (the streaming pass behind VectorAssembler),
    n x 128 with n chosen so columns + output fit beside nothing else (51 GB at 50M rows per GPU)."""
    n_total = 50_000_000 if not args.rows else max(args.rows // 2, 1)
    n = n_total // world
    d = 128
    free, _ = torch.cuda.mem_get_info()
    if free < n * d * 8 * 1.1:
        n = int(free / (d * 8 * 1.2))
    cols = torch.empty((d, n), dtype=torch.float32, device=dev)
    for j in range(d):
        cols[j].copy_(torch.arange(n, device=dev, dtype=torch.float32).mul_(1e-6).add_(float(j)))
    X = torch.empty((n, d), dtype=torch.float32, device=dev)
    ptrs = torch.tensor([cols[j].data_ptr() for j in range(d)], dtype=torch.int64, device=dev)
    codes = torch.zeros(d, dtype=torch.int32, device=dev)
    bad = torch.zeros(1, dtype=torch.int32, device=dev)

    def once():
        N.check(L.b2f_pack_columns(N.ptr(ptrs), N.ptr(codes), d, n, N.ptr(X), d, N.ptr(bad), N.stream_ptr()),
                "b2f_pack_columns")

    for _ in range(max(args.warmup, 1)):
        once()
    sync_all()
    steps = max(args.steps, 5)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        once()
    t1.record()
    sync_all()
    ms = max_over_ranks(t0.elapsed_time(t1)) / steps
    ok = bool(int(bad.item()) == 0 and float(X[12345 % n, 7].item()) == float(cols[7, 12345 % n].item()))
    bytes_pass = n * d * 8.0
    gbs = bytes_pass / (ms * 1e-3) / 1e9
    return {"metric": "loader_rows_per_s", "value": n * world / (ms * 1e-3), "unit": "rows/s", "ms_per_step": ms,
            "config": {"workload": "b2f_pack_columns (VectorAssembler, K1): 128 device fp32 columns -> row-major fp32, "
                                   "%d rows per GPU" % n, "rows_per_gpu": n, "d": d},
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s", "frac": gbs / pk["hbm"],
                         "traffic": None, "traffic_source": None, "kernel": "pack_columns_kernel",
                         "algorithmic_bytes_per_launch": bytes_pass, "peak_source": pk["src"]},
            "check": {"spot_value_ok": ok}}


def cfg5_leg(args, torch, N, L, comm, rank, world, dev, pk, DeviceMatrix, ColumnarFrame, shard_bounds, max_over_ranks,
             sync_all):
    """BASELINE configs[4]: StandardScaler -> KMeans k=256 on 200M x 64 (row-sharded): the scaler's two HBM passes
    (column statistics, x / sigma) timed once each, then Lloyd iterations on the scaled matrix."""
    import numpy as np

    from b200fit._kmeans_init import init_random
    from b200fit.clustering import CudaLloydEngine
    from b200fit.feature import StandardScaler

    n_total = 200_000_000 if not args.rows else max(args.rows, 1)
    d, k = 64, 256
    lo, hi = shard_bounds(n_total, rank, world)
    n = hi - lo
    rng = np.random.default_rng(SEED + 5)
    col_scale = np.exp(rng.uniform(np.log(0.1), np.log(10.0), size=d)).astype(np.float32)
    cs_d = torch.from_numpy(col_scale).to(dev)  # keep alive across the call
    Xt = torch.empty((n, d), dtype=torch.float32, device=dev)
    N.check(L.b2f_synth_mixture(N.ptr(Xt), n, lo, d, d, SEED + 5, k, 2.0, N.ptr(cs_d), N.stream_ptr()),
            "b2f_synth_mixture")
    torch.cuda.synchronize()
    frame = ColumnarFrame({"raw": DeviceMatrix(Xt, d, lo)})
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    sync_all()
    ev[0].record()
    sm = StandardScaler(inputCol="raw", outputCol="features").fit(frame)
    ev[1].record()
    scaled = sm.transform(frame)
    ev[2].record()
    sync_all()
    fit_ms, tr_ms = max_over_ranks(ev[0].elapsed_time(ev[1])), max_over_ranks(ev[1].elapsed_time(ev[2]))
    del Xt, frame
    torch.cuda.empty_cache()
    X = scaled.col("features")  # a VIEW: raw rows + per-column factor, folded into the Lloyd pass (no scaled copy)
    folded = getattr(X, "col_scale", None) is not None
    C0 = init_random(X, n_total, k, SEED + 5, comm)
    eng = CudaLloydEngine(X, k)
    eng.set_centers(C0)

    def iteration():
        payload = eng.step()
        comm.all_reduce_sum(payload)
        eng.finalize_async(payload, 0.0)

    for _ in range(args.warmup):
        iteration()
    sync_all()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(args.steps):
        iteration()
    t1.record()
    sync_all()
    ms = max_over_ranks(t0.elapsed_time(t1)) / args.steps
    flops = 2.0 * n * k * d
    achieved = flops / (ms * 1e-3) / 1e12
    rows_after = eng.rows_counted(eng.payload)
    return {"metric": "kmeans_lloyd_iter_per_s", "value": 1e3 / ms, "unit": "iter/s", "ms_per_step": ms,
            "config": {"workload": "StandardScaler -> KMeans k=256 on synthetic %dx64 fp32 (BASELINE configs[4])"
                                   % n_total, "rows": n_total, "d": d, "k": k},
            "scaler": {"fit_ms": fit_ms, "transform_ms": tr_ms,
                       "fit_gbs": n * d * 4.0 / (fit_ms * 1e-3) / 1e9,
                       "transform_gbs": None if folded else n * d * 8.0 / (tr_ms * 1e-3) / 1e9,
                       "folded_into_lloyd_pass": folded,
                       "note": "fit = column statistics pass (+ host sync); transform returns a view (raw rows + "
                               "per-column factor) that b2f_kmeans_step_scaled consumes: no scaled copy is written"
                               if folded else "fit = column statistics pass (+ host sync), transform = b2f_scale_columns"},
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": pk["tensor"], "unit": "TFLOP/s",
                         "frac": achieved / pk["tensor"], "traffic": None, "traffic_source": None,
                         "kernel": "b2f_kmeans_step_scaled (whole step)" if folded else "b2f_kmeans_step_ws (whole step)",
                         "hbm_floor_frac": (n * d * 4.0 / (ms * 1e-3) / 1e9) / pk["hbm"],
                         "algorithmic_flops_per_launch": flops, "peak_source": pk["src"] + " bf16 sustained"},
            "check": {"rows_accumulated_last_iter": rows_after, "std_first_cols": [float(v) for v in sm.std[:3]],
                      "col_scale_first_cols": [float(v) for v in col_scale[:3]]}}


def lrm_leg(args, torch, N, L, comm, rank, world, dev, pk, CudaLogRegEngine, DeviceMatrix, shard_bounds,
            max_over_ranks, sync_all):
    """BASELINE configs[3]: multinomial LR (20 classes) loss+gradient pass on 20M x 512."""
    import numpy as np

    n_total = args.lrm_rows or (LRM["n"] if not args.rows else max(args.rows // 4, 1))
    d, K = LRM["d"], LRM["K"]
    lo, hi = shard_bounds(n_total, rank, world)
    n = hi - lo
    rng = np.random.default_rng(SEED + 2)
    scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d)).astype(np.float32)
    w_true = (rng.standard_normal((K, d)) / np.sqrt(d)).astype(np.float32)
    Xt = torch.empty((n, d), dtype=torch.float32, device=dev)
    y = torch.empty(n, dtype=torch.float32, device=dev)
    sc_d, w_d = torch.from_numpy(scale).to(dev), torch.from_numpy(w_true).to(dev)  # keep alive across the call
    b_d = torch.zeros(K, dtype=torch.float32, device=dev)
    N.check(L.b2f_synth_logistic(N.ptr(Xt), N.ptr(y), n, lo, d, d, SEED + 2, N.ptr(sc_d), N.ptr(w_d), N.ptr(b_d), K,
                                 N.stream_ptr()), "b2f_synth_logistic")
    torch.cuda.synchronize()
    eng = CudaLogRegEngine(DeviceMatrix(Xt, d, lo), y, None, K)
    eng.beta.copy_(torch.from_numpy((rng.standard_normal((K, d)) * 0.05 / scale[None, :]).ravel()).to(dev))

    def one_pass():
        comm.all_reduce_sum(eng.lossgrad_device())

    for _ in range(args.warmup):
        one_pass()
    sync_all()
    steps = max(args.steps, 5)
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(steps):
        one_pass()
    t1.record()
    sync_all()
    ms = max_over_ranks(t0.elapsed_time(t1)) / steps
    p = eng.out.cpu().numpy()
    bytes_pass = n * (4.0 * d + 4.0)
    gbs = bytes_pass / (ms * 1e-3) / 1e9
    return {"metric": "lr_multinomial_lossgrad_pass_per_s", "value": 1e3 / ms, "unit": "pass/s", "ms_per_step": ms,
            "config": {"workload": "multinomial LogisticRegression (20 classes) loss+gradient pass on synthetic "
                                   "%dx512 fp32 (BASELINE configs[3])" % n_total, "rows": n_total, "d": d, "K": K},
            "roofline": with_traffic({"bound": "hbm", "achieved": gbs, "peak": pk["hbm"], "unit": "GB/s",
                                      "frac": gbs / pk["hbm"],
                                      "kernel": "lr_mn_tc_kernel (tcgen05: 3xTF32 margins GEMM + bf16 two-piece gradient GEMM)",
                                      "algorithmic_bytes_per_launch": bytes_pass, "peak_source": pk["src"]},
                                     "lr_mn_tc_kernel", n),
            "cpu_baseline": cpu_lr_baseline(n_total, d, K, SEED + 2, SEED + 2) if (rank == 0 and not args.no_cpu) else None,
            "check": {"loss": float(p[K * d + K] / p[K * d + K + 1]), "weight_sum": float(p[K * d + K + 1])}}


def cpu_baseline(n_total):
    """Oracle ('port') timed on the host cores on a bounded sample (about 10-30 s of CPU work)."""
    dt, n_s, what = cpu_kmeans_sample(n_total, 3)
    return {"value": 1.0 / (dt * n_total / n_s), "unit": "iter/s", "cores": os.cpu_count() or 1, "kind": "port",
            "sample": what}


def cpu_lr_baseline(n_total, d, K, seed, rng_seed):
    """Oracle loss+gradient pass (numpy/OpenBLAS, all host threads) on rows [0, n_s) of the leg's Philox features;
    the labels of the sample are drawn on the host (a pass costs the same whatever they are)."""
    import numpy as np

    import oracle
    from oracle import synth

    n_s = min(n_total, 400_000 if K == 1 else 200_000)
    rng = np.random.default_rng(rng_seed)  # the leg draws its column scales from this stream first
    scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d)).astype(np.float32)
    X = synth.logistic_features(n_s, 0, d, seed, scale)
    reps = 3
    if K == 1:
        y = (rng.uniform(size=n_s) < 0.5).astype(np.float64)
        beta = rng.standard_normal(d) * 0.05 / scale
        oracle.lr_binary_lossgrad(X[:20_000], y[:20_000], None, beta, 0.1, np.ones(d))
        t0 = time.perf_counter()
        for _ in range(reps):
            oracle.lr_binary_lossgrad(X, y, None, beta, 0.1, np.ones(d))
    else:
        y = rng.integers(0, K, size=n_s).astype(np.float64)
        B = rng.standard_normal((K, d)) * 0.05 / scale[None, :]
        oracle.lr_multinomial_lossgrad(X[:20_000], y[:20_000], None, B, np.zeros(K), np.ones(d))
        t0 = time.perf_counter()
        for _ in range(reps):
            oracle.lr_multinomial_lossgrad(X, y, None, B, np.zeros(K), np.ones(d))
    dt = (time.perf_counter() - t0) / reps
    return {"value": 1.0 / (dt * n_total / n_s), "unit": "pass/s", "cores": os.cpu_count() or 1, "kind": "port",
            "sample": "rows [0, %d) of the %d Philox rows, %d oracle loss+gradient passes, linear extrapolation in N; "
                      "numpy/OpenBLAS fp64 CPU restatement of MLlib's aggregator (Spark local[N]: unavailable on this box)"
                      % (n_s, n_total, reps)}


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        reference_arm(a)
    else:
        b200_arm(a)
