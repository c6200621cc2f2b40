"""Helpers for the -m gpu parity tests: thin ctypes calls straight through the C ABI."""
import numpy as np
import torch

from b200fit import _native as N


def dev():
    return torch.device("cuda", 0)


def stream():
    return N.stream_ptr()


def upload_matrix(Xh):
    """Host [n, d] float32 -> (device tensor [n, ld] zero padded, d, ld)."""
    Xh = np.ascontiguousarray(Xh, dtype=np.float32)
    n, d = Xh.shape
    ld = (d + 3) // 4 * 4
    X = torch.zeros((n, ld), dtype=torch.float32, device=dev())
    X[:, :d] = torch.from_numpy(Xh).to(dev())
    return X, d, ld


def row_norm_max(X, n, d, ld, col_scale=None):
    """max ||x|| over the rows (of x * col_scale), as b2f_row_norm_max reports it (fp32, rounded up)."""
    out = torch.zeros(1, dtype=torch.float32, device=dev())
    N.check(N.lib().b2f_row_norm_max(N.ptr(X), n, d, ld, N.ptr(col_scale), N.ptr(out), stream()), "row_norm_max")
    return float(out.item())


def kmeans_state(C, k, d, xmax=0.0):
    """xmax > 0: operands prepared with a global key scale for rows up to that norm (b2f_kmeans_prepare_x)."""
    L = N.lib()
    cen = torch.from_numpy(np.ascontiguousarray(C, dtype=np.float64)).to(dev())
    prep = torch.empty(L.b2f_kmeans_prep_bytes(k, d), dtype=torch.uint8, device=dev())
    acc = torch.empty(L.b2f_kmeans_acc_bytes(k, d), dtype=torch.uint8, device=dev())
    N.check(L.b2f_kmeans_prepare_x(N.ptr(cen), k, d, float(xmax), N.ptr(prep), stream()), "prepare")
    N.check(L.b2f_kmeans_acc_zero(N.ptr(acc), k, d, stream()), "acc_zero")
    return cen, prep, acc


def kmeans_step_raw(X, n, d, ld, C, algo, w=None, want_labels=True, ws_bytes=None, info=None, use_ws=False,
                    col_scale=None, global_scale=True, xmax_override=None):
    """One b2f_kmeans_step (or _step_ws for B2F_KM_TCB / B2F_KM_TCH, or any algo with use_ws; _step_scaled with a
    column scale) + fold.  Returns labels, sums[k,d], counts[k], cost, nrows (host).
    global_scale: prepare the operands with the data's largest row norm (what the fit does); False = the per-row key
    scales of a plain b2f_kmeans_prepare (what predict on unseen data uses)."""
    L = N.lib()
    k = C.shape[0]
    xmax = row_norm_max(X, n, d, ld, col_scale) if (global_scale and n > 0) else 0.0
    if xmax_override is not None:
        xmax = xmax_override
    cen, prep, acc = kmeans_state(C, k, d, xmax)
    labels = torch.full((n,), -1, dtype=torch.int32, device=dev()) if want_labels else None
    if algo in (N.KM_TCB, N.KM_TCH) or use_ws or col_scale is not None:
        wsb = int(L.b2f_kmeans_ws_bytes(n, k, d)) if ws_bytes is None else ws_bytes
        ws = torch.zeros(wsb, dtype=torch.uint8, device=dev())
        if col_scale is not None:
            N.check(L.b2f_kmeans_step_scaled(N.ptr(X), n, d, ld, N.ptr(col_scale), N.ptr(prep), k, N.ptr(acc),
                                             N.ptr(labels), N.ptr(ws), wsb, stream()), "step_scaled")
        else:
            N.check(L.b2f_kmeans_step_ws(N.ptr(X), n, d, ld, N.ptr(w), N.ptr(prep), k, N.ptr(acc), N.ptr(labels),
                                         N.ptr(ws), wsb, algo, stream()), "step_ws")
        if info is not None:
            # workspace header (include/b200fit.h): [0] rows listed for the 3xTF32 re-check, [1] rows re-checked,
            # [2] rows listed for the fp64 kernel, [3] capacity of each list, [4] rows offered to the candidate check (the pair check of earlier versions)
            hdr = ws[:20].view(torch.int32).cpu().numpy()
            info["listed"], info["rechecked"], info["exact"], info["cap"] = (int(v) for v in hdr[:4])
            info["pairs"] = min(int(hdr[4]), int(hdr[3]))
            info["uncertain"] = info["listed"] + info["pairs"]
    else:
        N.check(L.b2f_kmeans_step(N.ptr(X), n, d, ld, N.ptr(w), N.ptr(prep), k, N.ptr(acc), N.ptr(labels), algo,
                                  stream()), "step")
    N.check(L.b2f_kmeans_fold(N.ptr(acc), k, d, stream()), "fold")
    plen = int(L.b2f_kmeans_payload_len(k, d))
    p = acc[: plen * 8].view(torch.float64).cpu().numpy()
    kd = k * d
    return (labels.cpu().numpy() if want_labels else None, p[:kd].reshape(k, d).copy(), p[kd:kd + k].copy(),
            float(p[kd + k]), float(p[kd + k + 1]))


def allowed_label_mismatch(dist_best, dist_second, rel=1e-6):
    """Rows whose oracle top-2 distance gap is below `rel` relative may disagree (north_star)."""
    return (dist_second - dist_best) <= rel * np.maximum(dist_second, 1e-300)


def lr_lossgrad_raw(X, y, w, n, d, ld, n_classes, beta, icpt):
    L = N.lib()
    K = 1 if n_classes == 2 else n_classes
    out = torch.zeros(K * d + K + 2, dtype=torch.float64, device=dev())
    scratch = torch.empty(L.b2f_lr_scratch_bytes(d, n_classes), dtype=torch.uint8, device=dev())
    b = torch.from_numpy(np.ascontiguousarray(beta, dtype=np.float64).ravel()).to(dev())
    ic = torch.from_numpy(np.ascontiguousarray(np.atleast_1d(icpt), dtype=np.float64)).to(dev())
    if K == 1:
        N.check(L.b2f_lr_binary_lossgrad(N.ptr(X), N.ptr(y), N.ptr(w), n, d, ld, N.ptr(b), N.ptr(ic), N.ptr(out),
                                         N.ptr(scratch), stream()), "lr_binary")
    else:
        N.check(L.b2f_lr_multinomial_lossgrad(N.ptr(X), N.ptr(y), N.ptr(w), n, d, ld, n_classes, N.ptr(b), N.ptr(ic),
                                              N.ptr(out), N.ptr(scratch), stream()), "lr_multinomial")
    p = out.cpu().numpy()
    W = p[K * d + K + 1]
    return p[K * d + K] / W, p[: K * d].reshape(K, d) / W, p[K * d: K * d + K] / W, W
