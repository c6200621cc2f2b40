"""Oracle (test infrastructure): Breeze L-BFGS / OWL-QN restated in numpy fp64.

Restates SURVEY.md Appendix A.4 / A.5 (breeze/optimize/{FirstOrderMinimizer,
LBFGS,StrongWolfe,OWLQN,BacktrackingLineSearch}.scala -- scalanlp Breeze, an
un-vendored, un-pinned dependency of Spark, itself un-pinned by the reference:
setup.py:66-76).  Reached from the reference through
`method_instance.fit(...)` orangecontrib/spark/base/spark_ml_estimator.py:22.

L-BFGS + Strong Wolfe is pinned by the pyspark doctest trajectory KAT
(tests/test_oracle_kat.py); OWL-QN + backtracking is MEDIUM CONFIDENCE and
un-probed (only its optimum is cross-checked).
"""
import math

import numpy as np


class FirstOrderException(Exception):
    pass


class _Cached:
    """CachedDiffFunction: remembers the last evaluated point only, so one data
    pass happens per *distinct* trial point."""

    def __init__(self, f):
        self.f = f
        self.last_x = None
        self.last = None
        self.evals = 0

    def __call__(self, x):
        if self.last_x is not None and x.shape == self.last_x.shape and np.array_equal(x, self.last_x):
            return self.last
        v, g = self.f(x)
        self.evals += 1
        self.last_x = x.copy()
        self.last = (float(v), np.asarray(g, dtype=np.float64).copy())
        return self.last


def _strong_wolfe(phi, init, max_zoom=10, max_ls=10):
    """StrongWolfeLineSearch.minimize.  phi(t) -> (value, directional derivative)."""
    c1, c2 = 1e-4, 0.9
    t = init
    low = (0.0,) + phi(0.0)  # (t, fval, dd)
    fval, dd = low[1], low[2]
    if dd > 0:
        raise FirstOrderException("Line search invoked with non-descent direction: %r" % dd)

    def interp(l, r):
        d1 = l[2] + r[2] - 3.0 * (l[1] - r[1]) / (l[0] - r[0])
        disc = d1 * d1 - l[2] * r[2]
        d2 = math.sqrt(disc) if disc >= 0 else float("nan")
        mult = r[0] - l[0]
        tt = r[0] - mult * (r[2] + d2 - d1) / (r[2] - l[2] + 2.0 * d2)
        lb = l[0] + 0.1 * (r[0] - l[0])
        ub = l[0] + 0.9 * (r[0] - l[0])
        if tt < lb:
            return lb
        if tt > ub:
            return ub
        return tt

    def zoom(lo, hi):
        for _ in range(max_zoom):
            tt = interp(hi, lo) if lo[0] > hi[0] else interp(lo, hi)
            c = (tt,) + phi(tt)
            if c[1] > fval + c1 * c[0] * dd or c[1] >= lo[1]:
                hi = c
            else:
                if abs(c[2]) <= c2 * abs(dd):
                    return c[0]
                if c[2] * (hi[0] - lo[0]) >= 0:
                    hi = lo
                lo = c
        raise FirstOrderException("Line search zoom failed")

    for i in range(max_ls):
        c = (t,) + phi(t)
        if math.isinf(c[1]) or math.isnan(c[1]):
            t /= 2.0
        else:
            if c[1] > fval + c1 * t * dd or (c[1] >= low[1] and i > 0):
                return zoom(low, c)
            if abs(c[2]) <= c2 * abs(dd):
                return c[0]
            if c[2] >= 0:
                return zoom(c, low)
            low = c
            t *= 1.5
    raise FirstOrderException("Line search failed")


def _backtracking(phi, initfval, init, shrink, max_iter=20, grow=2.1, c_armijo=1e-4, c_wolfe=0.9,
                  min_alpha=1e-10, max_alpha=1e10):
    """BacktrackingLineSearch.iterations(...).last.alpha, including the upstream
    quirk that the first state carries `initfval` (value at alpha=0 handed in by
    the caller) together with the derivative at alpha=init."""
    f0, df0 = phi(0.0)
    fderiv = phi(init)[1]
    alpha, fval = init, initfval
    it = 0
    while it < max_iter:
        if fval > f0 + alpha * df0 * c_armijo:
            mult = shrink
        elif fderiv < c_wolfe * df0:
            mult = grow
        elif fderiv > -c_wolfe * df0:
            mult = shrink
        else:
            mult = 1.0
        if mult == 1.0:
            return alpha
        new_alpha = alpha * mult
        if it >= max_iter:
            raise FirstOrderException("LineSearchFailed")
        if new_alpha < min_alpha:
            raise FirstOrderException("StepSizeUnderflow")
        if new_alpha > max_alpha:
            raise FirstOrderException("StepSizeOverflow")
        fval, fderiv = phi(new_alpha)
        alpha = new_alpha
        it += 1
    return alpha


class LBFGS:
    """breeze.optimize.LBFGS(maxIter, m, tolerance) with defaultConvergenceCheck
    (maxIter || function values converged over a 10-value window seeded with +inf
    || gradient converged || search failed)."""

    def __init__(self, max_iter=100, m=10, tol=1e-6, fval_memory=10):
        self.max_iter, self.m, self.tol, self.fval_memory = max_iter, m, tol, fval_memory

    # hooks overridden by OWLQN ------------------------------------------------
    def _adjust(self, x, g, v):
        return v, g

    def _take_step(self, st, d, alpha):
        return st["x"] + d * alpha

    def _direction(self, st):
        return self._two_loop(st["hist"], st["grad"])

    def _step_size(self, st, f, d):
        x = st["x"]

        def phi(t):
            v, g = f(x + d * t)
            return v, float(g @ d)

        init = 1.0 / float(np.linalg.norm(d)) if st["iter"] == 0 else 1.0
        alpha = _strong_wolfe(phi, init)
        if alpha * float(np.linalg.norm(st["grad"])) < 1e-10:
            raise FirstOrderException("StepSizeUnderflow")
        return alpha

    # -------------------------------------------------------------------------
    def _two_loop(self, hist, grad):
        S, Y = hist
        n = len(S)
        if n > 0:
            sy = float(S[0] @ Y[0])
            yy = float(Y[0] @ Y[0])
            if sy < 0 or math.isnan(sy):
                raise FirstOrderException("NaNHistory")
            diag = sy / yy
        else:
            diag = 1.0
        q = grad.copy()
        a = np.zeros(n)
        rho = np.zeros(n)
        for i in range(n):
            rho[i] = float(S[i] @ Y[i])
            a[i] = float(S[i] @ q) / rho[i]
            if math.isnan(a[i]):
                raise FirstOrderException("NaNHistory")
            q = q - a[i] * Y[i]
        q = q * diag
        for i in range(n - 1, -1, -1):
            beta = float(Y[i] @ q) / rho[i]
            q = q + (a[i] - beta) * S[i]
        return -q

    def _converged(self, st):
        if st["iter"] >= self.max_iter and self.max_iter >= 0:
            return "max iterations"
        fv = st["fvals"]
        if len(fv) >= 2 and abs(st["adj_value"] - max(fv)) <= self.tol * st["init_adj"]:
            return "function values converged"
        if float(np.linalg.norm(st["adj_grad"])) <= max(self.tol * st["adj_value"], 1e-8):
            return "gradient converged"
        if st.get("search_failed"):
            return "search failed"
        return None

    def iterations(self, f, x0):
        """Yields every optimizer state (including the initial one) up to and
        including the first state for which the convergence check fires."""
        fc = _Cached(f)
        x = np.asarray(x0, dtype=np.float64).copy()
        v, g = fc(x)
        av, ag = self._adjust(x, g, v)
        st = dict(x=x, value=v, grad=g, adj_value=av, adj_grad=ag, iter=0, init_adj=av,
                  hist=([], []), fvals=[float("inf")], evals=fc.evals)
        failed_once = False
        while True:
            yield st
            if self._converged(st) is not None:
                return
            try:
                d = self._direction(st)
                alpha = self._step_size(st, fc, d)
                xn = self._take_step(st, d, alpha)
                v, g = fc(xn)
                av, ag = self._adjust(xn, g, v)
                S, Y = st["hist"]
                S = [xn - st["x"]] + S
                Y = [g - st["grad"]] + Y
                fvals = (st["fvals"] + [v])[-self.fval_memory:]
                failed_once = False
                st = dict(x=xn, value=v, grad=g, adj_value=av, adj_grad=ag, iter=st["iter"] + 1,
                          init_adj=st["init_adj"], hist=(S[: self.m], Y[: self.m]), fvals=fvals,
                          evals=fc.evals)
            except FirstOrderException:
                if failed_once:
                    st = dict(st, search_failed=True)
                else:
                    failed_once = True
                    st = dict(st, hist=([], []))

    def minimize(self, f, x0):
        states = list(self.iterations(f, x0))
        return states[-1], states


class OWLQN(LBFGS):
    """breeze.optimize.OWLQN(maxIter, m, l1reg, tolerance): `l1` is the per-index
    L1 weight vector (0 for intercepts)."""

    def __init__(self, max_iter, m, l1, tol):
        super().__init__(max_iter, m, tol)
        self.l1 = np.asarray(l1, dtype=np.float64)

    def _adjust(self, x, g, v):
        l1 = self.l1
        adjv = v + float(np.abs(l1 * x).sum())
        dp, dm = g + l1, g - l1
        at_zero = np.where(dm > 0, dm, np.where(dp < 0, dp, 0.0))
        adj = np.where(x == 0.0, at_zero, g + np.sign(x) * l1)
        adj = np.where(l1 == 0.0, g, adj)
        return adjv, adj

    def _direction(self, st):
        d = self._two_loop(st["hist"], st["adj_grad"])
        return np.where(d * st["adj_grad"] < 0, d, 0.0)

    def _take_step(self, st, d, alpha):
        stepped = st["x"] + d * alpha
        orth = np.where(st["x"] != 0, np.sign(st["x"]), np.sign(-st["adj_grad"]))
        return stepped * (np.sign(stepped) == np.sign(orth))

    def _step_size(self, st, f, d):
        def phi(t):
            xn = self._take_step(st, d, t)
            v, g = f(xn)
            av, ag = self._adjust(xn, g, v)
            return av, float(ag @ d)

        it = st["iter"]
        init = 0.5 / float(np.linalg.norm(st["grad"])) if it < 1 else 1.0
        return _backtracking(phi, st["value"], init, shrink=0.1 if it < 1 else 0.5)
