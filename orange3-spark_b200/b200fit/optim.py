"""Host-side quasi-Newton drivers for LogisticRegression: L-BFGS with a strong-Wolfe line
search and OWL-QN with a backtracking search, following the Breeze optimizers that MLlib's
LogisticRegression.train drives.  Reached from the reference
through fit(): orangecontrib/spark/base/spark_ml_estimator.py:22.

The vectors are tiny (d+1 or K(d+1) doubles); every objective evaluation is one fused
GPU pass plus one all-reduce, and one evaluation happens per DISTINCT trial point
(last-point cache, as Breeze's CachedDiffFunction does).
"""
import math
from dataclasses import dataclass, field

import numpy as np


class LineSearchError(Exception):
    pass


@dataclass
class OptState:
    x: np.ndarray
    value: float
    grad: np.ndarray
    adj_value: float
    adj_grad: np.ndarray
    iter: int
    init_adj_value: float
    mem_s: list = field(default_factory=list)   # newest first
    mem_y: list = field(default_factory=list)
    fvals: list = field(default_factory=lambda: [math.inf])
    search_failed: bool = False
    n_evals: int = 0


class _LastPointCache:
    def __init__(self, fn):
        self.fn, self.key, self.val, self.count = fn, None, None, 0

    def __call__(self, x):
        if self.key is None or self.key.shape != x.shape or not np.array_equal(self.key, x):
            v, g = self.fn(x)
            self.count += 1
            self.key, self.val = x.copy(), (float(v), np.array(g, dtype=np.float64))
        return self.val


class Bracket:
    __slots__ = ("t", "f", "df")

    def __init__(self, t, f, df):
        self.t, self.f, self.df = t, f, df


class StrongWolfe:
    """Bracketing phase with 1.5x expansion, zoom phase with safeguarded cubic interpolation."""

    C1, C2 = 1e-4, 0.9

    def __init__(self, max_zoom=10, max_search=10, slack=0.0):
        """`slack`: absolute resolution of the objective values.  Every comparison BETWEEN OBJECTIVE VALUES is
        relaxed by it, so that once the true decrease of a step falls below the rounding noise of the fp32 passes the
        slope tests (whose inputs are accurate to ~1e-9) decide alone.  0 reproduces Breeze exactly."""
        self.max_zoom, self.max_search, self.slack = max_zoom, max_search, slack

    @staticmethod
    def _cubic(a, b):
        d1 = a.df + b.df - 3.0 * (a.f - b.f) / (a.t - b.t)
        rad = d1 * d1 - a.df * b.df
        d2 = math.sqrt(rad) if rad >= 0.0 else math.nan
        t = b.t - (b.t - a.t) * (b.df + d2 - d1) / (b.df - a.df + 2.0 * d2)
        lo, hi = a.t + 0.1 * (b.t - a.t), a.t + 0.9 * (b.t - a.t)
        if t < lo:
            return lo
        if t > hi:
            return hi
        return t

    def search(self, phi, t0):
        start = Bracket(0.0, *phi(0.0))
        f0, g0 = start.f, start.df
        eps = self.slack
        if g0 > 0:
            raise LineSearchError("not a descent direction (slope %g)" % g0)

        def zoom(lo, hi):
            for _ in range(self.max_zoom):
                t = self._cubic(hi, lo) if lo.t > hi.t else self._cubic(lo, hi)
                c = Bracket(t, *phi(t))
                if c.f > f0 + self.C1 * c.t * g0 + eps or c.f >= lo.f + eps:
                    hi = c
                    continue
                if abs(c.df) <= self.C2 * abs(g0):
                    return c.t
                if c.df * (hi.t - lo.t) >= 0:
                    hi = lo
                lo = c
            raise LineSearchError("zoom did not converge")

        t, prev = t0, start
        for i in range(self.max_search):
            c = Bracket(t, *phi(t))
            if math.isinf(c.f) or math.isnan(c.f):
                t *= 0.5
                continue
            if c.f > f0 + self.C1 * t * g0 + eps or (i > 0 and c.f >= prev.f + eps):
                return zoom(prev, c)
            if abs(c.df) <= self.C2 * abs(g0):
                return c.t
            if c.df >= 0:
                return zoom(c, prev)
            prev = c
            t *= 1.5
        raise LineSearchError("bracketing did not converge")


class LBFGS:
    def __init__(self, max_iter=100, memory=10, tol=1e-6, window=10, resolution=0.0):
        """`resolution`: relative resolution of the objective (0 = exact arithmetic, Breeze's behaviour).
        The GPU passes evaluate margins in fp32, so objective differences below ~1e-8 relative are noise
        (measured: ~5e-10 between neighbouring points at n = 5000, tools/lr_noise_check.py), while the gradient is
        good to ~3e-9 absolute.  Two consequences: the line searches compare objective values with a slack of
        resolution * |f| (so the last, superlinear steps are accepted on their slopes, as they are in fp64), and when
        a line search still fails while the last accepted step improved the objective by less than the resolution the
        run stops at once instead of burning ~20 more passes on Breeze's reset-and-retry."""
        self.max_iter, self.memory, self.tol, self.window = max_iter, memory, tol, window
        self.resolution = resolution

    # ---- pieces OWL-QN overrides ------------------------------------------------------
    def adjust(self, x, grad, value):
        return value, grad

    def step_to(self, st, direction, alpha):
        return st.x + direction * alpha

    def direction(self, st):
        return self.two_loop(st, st.grad)

    def step_size(self, st, fn, direction):
        def phi(a):
            v, g = fn(st.x + direction * a)
            return v, float(np.dot(g, direction))

        t0 = 1.0 / float(np.linalg.norm(direction)) if st.iter == 0 else 1.0
        alpha = StrongWolfe(slack=self.resolution * abs(st.value)).search(phi, t0)
        if alpha * float(np.linalg.norm(st.grad)) < 1e-10:
            raise LineSearchError("step size underflow")
        return alpha

    # -----------------------------------------------------------------------------------
    def two_loop(self, st, grad):
        S, Y = st.mem_s, st.mem_y
        m = len(S)
        gamma = 1.0
        if m:
            sy, yy = float(np.dot(S[0], Y[0])), float(np.dot(Y[0], Y[0]))
            if sy < 0 or math.isnan(sy):
                raise LineSearchError("curvature history is not positive")
            gamma = sy / yy
        q = grad.copy()
        alphas, rhos = [0.0] * m, [0.0] * m
        for i in range(m):
            rhos[i] = float(np.dot(S[i], Y[i]))
            alphas[i] = float(np.dot(S[i], q)) / rhos[i]
            if math.isnan(alphas[i]):
                raise LineSearchError("NaN in history")
            q = q - alphas[i] * Y[i]
        q = q * gamma
        for i in reversed(range(m)):
            b = float(np.dot(Y[i], q)) / rhos[i]
            q = q + (alphas[i] - b) * S[i]
        return -q

    def stop_reason(self, st):
        if 0 <= self.max_iter <= st.iter:
            return "maxIter"
        if len(st.fvals) >= 2 and abs(st.adj_value - max(st.fvals)) <= self.tol * st.init_adj_value:
            return "objective converged"
        if float(np.linalg.norm(st.adj_grad)) <= max(self.tol * st.adj_value, 1e-8):
            return "gradient converged"
        if st.search_failed:
            return "line search failed"
        return None

    def run(self, objective, x0, callback=None):
        """Returns the list of states, initial one included, up to the first converged state."""
        fn = _LastPointCache(objective)
        x = np.array(x0, dtype=np.float64)
        v, g = fn(x)
        av, ag = self.adjust(x, g, v)
        st = OptState(x, v, g, av, ag, 0, av, n_evals=fn.count)
        trail = []
        retried = False
        while True:
            trail.append(st)
            if callback:
                callback(st)
            if self.stop_reason(st) is not None:
                return trail
            try:
                p = self.direction(st)
                a = self.step_size(st, fn, p)
                xn = self.step_to(st, p, a)
                v, g = fn(xn)
                av, ag = self.adjust(xn, g, v)
                st = OptState(xn, v, g, av, ag, st.iter + 1, st.init_adj_value,
                              ([xn - st.x] + st.mem_s)[: self.memory], ([g - st.grad] + st.mem_y)[: self.memory],
                              (st.fvals + [v])[-self.window:], False, fn.count)
                retried = False
            except LineSearchError:
                at_resolution = (self.resolution > 0.0 and len(trail) >= 2 and
                                 abs(trail[-2].adj_value - st.adj_value) <= self.resolution * abs(st.adj_value))
                if retried or at_resolution:
                    st = OptState(st.x, st.value, st.grad, st.adj_value, st.adj_grad, st.iter, st.init_adj_value,
                                  st.mem_s, st.mem_y, st.fvals, True, fn.count)
                else:
                    retried = True
                    st = OptState(st.x, st.value, st.grad, st.adj_value, st.adj_grad, st.iter, st.init_adj_value,
                                  [], [], st.fvals, False, fn.count)


class OWLQN(LBFGS):
    """Orthant-wise L-BFGS; `l1` holds the per-coordinate L1 weight (0 = unpenalised)."""

    def __init__(self, max_iter, memory, l1, tol, resolution=0.0):
        super().__init__(max_iter, memory, tol, resolution=resolution)
        self.l1 = np.asarray(l1, dtype=np.float64)

    def adjust(self, x, grad, value):
        l1 = self.l1
        total = value + float(np.sum(np.abs(l1 * x)))
        right, left = grad + l1, grad - l1
        pg_zero = np.where(left > 0, left, np.where(right < 0, right, 0.0))
        pg = np.where(x == 0.0, pg_zero, grad + np.sign(x) * l1)
        return total, np.where(l1 == 0.0, grad, pg)

    def direction(self, st):
        p = self.two_loop(st, st.adj_grad)
        return np.where(p * st.adj_grad < 0, p, 0.0)

    def step_to(self, st, direction, alpha):
        moved = st.x + direction * alpha
        orthant = np.where(st.x != 0, np.sign(st.x), np.sign(-st.adj_grad))
        return moved * (np.sign(moved) == np.sign(orthant))

    def step_size(self, st, fn, direction):
        def phi(a):
            xn = self.step_to(st, direction, a)
            v, g = fn(xn)
            av, ag = self.adjust(xn, g, v)
            return av, float(np.dot(ag, direction))

        first = st.iter < 1
        t0 = 0.5 / float(np.linalg.norm(st.grad)) if first else 1.0
        return backtrack(phi, st.value, t0, 0.1 if first else 0.5, slack=self.resolution * abs(st.adj_value))


def backtrack(phi, f_at_zero_hint, t0, shrink, grow=2.1, c_armijo=1e-4, c_wolfe=0.9, max_iter=20, t_min=1e-10,
              t_max=1e10, slack=0.0):
    """Breeze's BacktrackingLineSearch, including its first-trial behaviour: the first test
    pairs the caller-supplied value at 0 with the slope measured at t0."""
    f0, g0 = phi(0.0)
    f_t, g_t = f_at_zero_hint, phi(t0)[1]
    t = t0
    for k in range(max_iter):
        if f_t > f0 + t * g0 * c_armijo + slack:
            factor = shrink
        elif g_t < c_wolfe * g0:
            factor = grow
        elif g_t > -c_wolfe * g0:
            factor = shrink
        else:
            return t
        t_new = t * factor
        if t_new < t_min:
            raise LineSearchError("step size underflow")
        if t_new > t_max:
            raise LineSearchError("step size overflow")
        f_t, g_t = phi(t_new)
        t = t_new
    return t
