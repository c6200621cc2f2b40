"""b2f_lr_binary_lossgrad_noalloc: the loss+gradient pass for callers that cannot give the
library any scratch memory (hosts that own only their input/output buffers, or that must not
call cudaMalloc).  It keeps its per-CTA partial sums in an array built into the library.

The check: through the C ABI, its output must equal the regular entry point's (which uses
caller-provided scratch) on the same inputs, twice in a row -- the second call runs on top of
whatever the first one left in the shared array -- and, so that the comparison is not just one
code path against its twin, the CPU oracle's loss and gradient for the same rows.  The file name sorts last so that a failure here
cannot hide the result of any other test in a stop-at-first-failure run."""
import numpy as np
import pytest
import torch

import oracle

pytestmark = pytest.mark.gpu


def _both(N, G, X, y, w, n, d, ld, beta, icpt):
    L = N.lib()
    dev = G.dev()
    b = torch.from_numpy(np.ascontiguousarray(beta, dtype=np.float64)).to(dev)
    ic = torch.tensor([icpt], dtype=torch.float64, device=dev)
    ref = torch.zeros(d + 3, dtype=torch.float64, device=dev)
    scratch = torch.empty(L.b2f_lr_scratch_bytes(d, 2), dtype=torch.uint8, device=dev)
    N.check(L.b2f_lr_binary_lossgrad(N.ptr(X), N.ptr(y), N.ptr(w), n, d, ld, N.ptr(b), N.ptr(ic), N.ptr(ref),
                                     N.ptr(scratch), G.stream()), "b2f_lr_binary_lossgrad")
    got = torch.zeros(d + 3, dtype=torch.float64, device=dev)
    N.check(L.b2f_lr_binary_lossgrad_noalloc(N.ptr(X), N.ptr(y), N.ptr(w), n, d, ld, N.ptr(b), N.ptr(ic), N.ptr(got),
                                             G.stream()), "b2f_lr_binary_lossgrad_noalloc")
    torch.cuda.synchronize()
    return got.cpu().numpy(), ref.cpu().numpy()


def test_noalloc_entry_matches_regular_entry_on_two_calls_in_a_row():
    from b200fit import _native as N
    import gpu_util as G

    rng = np.random.default_rng(123)
    # second call uses a SMALLER problem, so stale partials from the first one sit right behind its slices
    for n, d in ((20011, 300), (4099, 37)):
        Xh = rng.normal(size=(n, d)).astype(np.float32)
        yh = (rng.uniform(size=n) < 0.4).astype(np.float32)
        wh = rng.uniform(0.5, 2.0, size=n).astype(np.float32)
        X, _, ld = G.upload_matrix(Xh)
        y = torch.from_numpy(yh).to(G.dev())
        w = torch.from_numpy(wh).to(G.dev())
        beta = rng.normal(size=d) * 0.1
        got, ref = _both(N, G, X, y, w, n, d, ld, beta, 0.25)
        # same kernels, same launch shape, same deterministic fold: the tolerance of the parity tests
        # (1e-5 relative) is far looser than needed
        scale = np.abs(ref).max()
        np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-5 * scale)
        W = float(wh.astype(np.float64).sum())
        assert got[d + 2] == pytest.approx(W, rel=1e-12)
        # and against the oracle (numpy fp64 restatement of the MLlib aggregator), at the parity tests' tolerance
        rl, rg, rgb = oracle.lr_binary_lossgrad(Xh, yh, wh.astype(np.float64), beta, 0.25, np.ones(d))
        assert got[d + 1] / W == pytest.approx(rl, rel=1e-5)
        np.testing.assert_allclose(got[:d] / W, rg, rtol=1e-5, atol=1e-5 * np.abs(rg).max())
        assert got[d] / W == pytest.approx(rgb, rel=1e-5, abs=1e-5 * np.abs(rg).max())


def test_noalloc_entry_rejects_rows_wider_than_its_scratch():
    from b200fit import _native as N
    import gpu_util as G

    L = N.lib()
    n, d = 64, 516  # padded width 516 > 512
    X = torch.zeros((n, d), dtype=torch.float32, device=G.dev())
    y = torch.zeros(n, dtype=torch.float32, device=G.dev())
    b = torch.zeros(d, dtype=torch.float64, device=G.dev())
    ic = torch.zeros(1, dtype=torch.float64, device=G.dev())
    out = torch.zeros(d + 3, dtype=torch.float64, device=G.dev())
    rc = L.b2f_lr_binary_lossgrad_noalloc(N.ptr(X), N.ptr(y), None, n, d, d, N.ptr(b), N.ptr(ic), N.ptr(out),
                                          G.stream())
    assert rc == -1
    assert b"scratch limit" in L.b2f_last_error()
