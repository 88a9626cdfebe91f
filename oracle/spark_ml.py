"""Driver-side restatement of Spark ML 2.4.4 LinearRegression and Breeze 0.13.2 optimisers.

TEST INFRASTRUCTURE ONLY (checker + timed CPU baseline).  PARITY UNPINNED: the reference
(App:120-126 calls `new LinearRegression().setMaxIter(40).setRegParam(1).setElasticNetParam(1)`
and `lr.fit(df)`) delegates all arithmetic to jars that are absent from /root/reference
(org.apache.spark:spark-mllib_2.11:2.4.4, org.scalanlp:breeze_2.11:0.13.2, pom.xml:13-32).
What follows restates those libraries' published algorithms:

  * ml.optim.WeightedLeastSquares.fit               -> wls_fit_from_moments
  * ml.optim.NormalEquationSolver (Cholesky / QN)   -> cholesky_solve / quasi_newton_solve
  * ml.regression.LinearRegression.train (l-bfgs)   -> lbfgs_path_fit
  * breeze.optimize.{FirstOrderMinimizer,LBFGS,OWLQN,BacktrackingLineSearch,
    StrongWolfeLineSearch,CachedDiffFunction}       -> class LBFGS / OWLQN below
  * mllib.evaluation.RegressionMetrics              -> regression_metrics

Vectors are small (d or d+1 entries) so plain numpy / Python loops are fine here; the O(n) data
passes live in dq4ml_oracle.c.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

from . import passes


class FirstOrderException(Exception):
    pass


# ----------------------------------------------------------------------------- Breeze line searches

def backtracking_line_search(ff, init, initfval, shrink_step, grow_step=2.1, c_armijo=1e-4, c_wolfe=0.9,
                             max_iterations=20, min_alpha=1e-10, max_alpha=1e10):
    """breeze.optimize.BacktrackingLineSearch.iterations(...).last.alpha

    Quirk kept on purpose: the first trial state carries `initfval` (the constructor argument, OWLQN passes
    state.value) as its function value, and only the derivative is taken from f(init).
    """
    f0, df0 = ff(0.0)
    initfderiv = ff(init)[1]
    alpha, fval, fderiv = init, initfval, initfderiv
    it = 0
    while True:
        if fval > f0 + alpha * df0 * c_armijo:
            mult = shrink_step
        elif fderiv < c_wolfe * df0:
            mult = grow_step
        elif fderiv > -c_wolfe * df0:
            mult = shrink_step
        else:
            mult = 1.0
        if mult == 1.0:
            return alpha
        new_alpha = alpha * mult
        if new_alpha < min_alpha:
            raise FirstOrderException("StepSizeUnderflow")
        if new_alpha > max_alpha:
            raise FirstOrderException("StepSizeOverflow")
        prev_alpha = alpha
        fval, fderiv = ff(new_alpha)
        alpha = new_alpha
        it += 1
        # `.takeWhile(!done && iter < maxIterations).last`: the state with iter == maxIterations is evaluated
        # but never emitted, so the caller gets the alpha before it.
        if it >= max_iterations:
            return prev_alpha


def strong_wolfe_line_search(ff, init, max_zoom_iter=10, max_line_search_iter=10, bound=None):
    """breeze.optimize.StrongWolfeLineSearch.minimize (Nocedal & Wright p.58-60); with `bound`, .minimizeWithBound
    (the variant LBFGSB uses: the step never grows past the bound and is returned as soon as it is reached with
    sufficient decrease)."""
    c1, c2 = 1e-4, 0.9
    if bound is not None and not init <= bound:
        raise ValueError("requirement failed: init value should <= bound")  # IllegalArgumentException in Breeze

    def phi(t):
        v, dd = ff(t)
        return (t, dd, v)  # Bracket(t, dd, fval)

    def interp(l, r):
        lt, ldd, lf = l
        rt, rdd, rf = r
        d1 = ldd + rdd - 3 * (lf - rf) / (lt - rt)
        d2 = math.sqrt(d1 * d1 - ldd * rdd) if d1 * d1 - ldd * rdd >= 0 else float("nan")
        multiplier = rt - lt
        t = rt - multiplier * (rdd + d2 - d1) / (rdd - ldd + 2 * d2)
        lbound = lt + 0.1 * (rt - lt)
        ubound = lt + 0.9 * (rt - lt)
        if t < lbound:
            return lbound
        if t > ubound:
            return ubound
        return t

    t = init
    low = phi(0.0)
    fval, dd = low[2], low[1]
    if dd > 0:
        raise FirstOrderException("Line search invoked with non-descent direction: %r" % dd)

    def zoom(linit, rinit):
        lo, hi = linit, rinit
        for _ in range(max_zoom_iter):
            tt = interp(hi, lo) if lo[0] > hi[0] else interp(lo, hi)
            c = phi(tt)
            if c[2] > fval + c1 * c[0] * dd or c[2] >= lo[2]:
                hi = c
            else:
                if abs(c[1]) <= c2 * abs(dd):
                    return c[0]
                if c[1] * (hi[0] - lo[0]) >= 0:
                    hi = lo
                lo = c
        raise FirstOrderException("Line search zoom failed")

    for i in range(max_line_search_iter):
        c = phi(t)
        if math.isinf(c[2]) or math.isnan(c[2]):
            t /= 2.0
        else:
            if (c[2] > fval + c1 * t * dd) or (c[2] >= low[2] and i > 0):
                return zoom(low, c)
            if abs(c[1]) <= c2 * abs(dd):
                return c[0]
            if c[1] >= 0:
                return zoom(c, low)
            low = c
            if bound is None:
                t *= 1.5
            elif t == bound:
                return bound  # sufficient decrease at the bound, curvature condition not met
            else:
                t = min(t * 1.5, bound)
    raise FirstOrderException("Line search failed")


# ----------------------------------------------------------------------------- Breeze FirstOrderMinimizer family

class CachedDiffFunction:
    """breeze.optimize.CachedDiffFunction: remembers only the last evaluated point (copied AFTER the call,
    so in-place edits made by the wrapped function are part of the remembered key)."""

    def __init__(self, f):
        self.f = f
        self.last = None
        self.evals = 0

    def __call__(self, x):
        if self.last is None or not np.array_equal(x, self.last[0]):
            v, g = self.f(x)
            self.evals += 1
            self.last = (x.copy(), v, g)
        return self.last[1], self.last[2]


@dataclass
class State:
    x: np.ndarray
    value: float
    grad: np.ndarray
    adjusted_value: float
    adjusted_gradient: np.ndarray
    iter: int
    initial_adj_val: float
    history: tuple  # (mem_step list, mem_grad_delta list), newest first
    fvals: list = field(default_factory=lambda: [float("inf")])
    search_failed: bool = False


class LBFGS:
    """breeze.optimize.LBFGS(maxIter, m, tolerance) with FirstOrderMinimizer.defaultConvergenceCheck."""

    FVAL_MEMORY = 20  # defaultConvergenceCheck(..., fvalMemory = 20)

    def __init__(self, max_iter, m, tol):
        self.max_iter, self.m, self.tol = max_iter, m, tol

    # -- hooks overridden by OWLQN / LBFGSB
    def adjust(self, x, grad, value):
        return value, grad

    def initial_history(self, x):
        return ([], [])

    def update_history(self, state, x, grad):
        mem_step, mem_gd = state.history
        return (([x - state.x] + mem_step)[: self.m], ([grad - state.grad] + mem_gd)[: self.m])

    def take_step(self, state, direction, step):
        return state.x + direction * step

    def choose_descent_direction(self, state, grad=None):
        g = state.grad if grad is None else grad
        mem_step, mem_gd = state.history
        hl = len(mem_step)
        if hl > 0:
            sy = float(mem_step[0] @ mem_gd[0])
            yy = float(mem_gd[0] @ mem_gd[0])
            if sy < 0 or math.isnan(sy):
                raise FirstOrderException("NaNHistory")
            diag = sy / yy
        else:
            diag = 1.0
        d = g.copy()
        a = [0.0] * hl
        rho = [0.0] * hl
        for i in range(hl):
            rho[i] = float(mem_step[i] @ mem_gd[i])
            a[i] = float(mem_step[i] @ d) / rho[i]
            if math.isnan(a[i]):
                raise FirstOrderException("NaNHistory")
            d += -a[i] * mem_gd[i]
        d *= diag
        for i in range(hl - 1, -1, -1):
            beta = float(mem_gd[i] @ d) / rho[i]
            d += (a[i] - beta) * mem_step[i]
        d *= -1.0
        return d

    def determine_step_size(self, state, f, direction):
        x = state.x

        def ff(alpha):
            v, g = f(x + direction * alpha)
            return v, float(g @ direction)

        init = 1.0 / float(np.linalg.norm(direction)) if state.iter == 0 else 1.0
        alpha = strong_wolfe_line_search(ff, init, 10, 10)
        if alpha * float(np.linalg.norm(state.grad)) < 1e-10:
            raise FirstOrderException("StepSizeUnderflow")
        return alpha

    # -- FirstOrderMinimizer
    def _converged(self, s):
        if s.iter >= self.max_iter and self.max_iter >= 0:
            return "MaxIterations"
        if len(s.fvals) >= 2 and abs(s.adjusted_value - max(s.fvals)) <= self.tol * s.initial_adj_val:
            return "FunctionValuesConverged"
        if float(np.linalg.norm(s.adjusted_gradient)) <= max(self.tol * s.adjusted_value, 1e-8):
            return "GradientConverged"
        if s.search_failed:
            return "SearchFailed"
        return None

    def iterations(self, f, init):
        """Yields every optimiser state including the initial one and the one on which a convergence test fires."""
        x = init
        value, grad = f(x)
        adj_v, adj_g = self.adjust(x, grad, value)
        state = State(x, value, grad, adj_v, adj_g, 0, adj_v, self.initial_history(x))
        failed_once = False
        while True:
            yield state
            self.reason = self._converged(state)
            if self.reason is not None:
                return
            try:
                direction = self.choose_descent_direction(state)
                step = self.determine_step_size(state, f, direction)
                x = self.take_step(state, direction, step)
                value, grad = f(x)
                adj_v, adj_g = self.adjust(x, grad, value)
                history = self.update_history(state, x, grad)
                fvals = (state.fvals + [value])[-self.FVAL_MEMORY:]
                failed_once = False
                state = State(x, value, grad, adj_v, adj_g, state.iter + 1, state.initial_adj_val, history, fvals)
            except FirstOrderException:
                if not failed_once:
                    failed_once = True
                    state = State(state.x, state.value, state.grad, state.adjusted_value, state.adjusted_gradient,
                                  state.iter, state.initial_adj_val, self.initial_history(state.x), state.fvals,
                                  state.search_failed)
                else:
                    state = State(state.x, state.value, state.grad, state.adjusted_value, state.adjusted_gradient,
                                  state.iter, state.initial_adj_val, state.history, state.fvals, True)


class OWLQN(LBFGS):
    """breeze.optimize.OWLQN(maxIter, m, l1reg: Int => Double, tolerance)."""

    def __init__(self, max_iter, m, l1reg, tol):
        super().__init__(max_iter, m, tol)
        self.l1reg = l1reg

    def adjust(self, x, grad, value):
        adj_value = value
        res = grad.copy()
        for i in range(x.size):
            l1 = self.l1reg(i)
            assert l1 >= 0.0
            if l1 == 0.0:
                continue
            xv, v = float(x[i]), float(grad[i])
            adj_value += abs(l1 * xv)
            if xv == 0.0:
                dp, dm = v + l1, v - l1
                res[i] = dm if dm > 0 else (dp if dp < 0 else 0.0)
            else:
                res[i] = v + math.copysign(1.0, xv) * l1
        return adj_value, res

    def choose_descent_direction(self, state, grad=None):
        d = super().choose_descent_direction(state, state.adjusted_gradient)
        return np.where(d * state.adjusted_gradient < 0, d, 0.0)

    def take_step(self, state, direction, step):
        stepped = state.x + direction * step
        orthant = np.where(state.x != 0, np.sign(state.x), np.sign(-state.adjusted_gradient))
        return stepped * (np.sign(stepped) == np.sign(orthant))

    def determine_step_size(self, state, f, direction):
        it = state.iter

        def ff(alpha):
            new_x = self.take_step(state, direction, alpha)
            v, new_g = f(new_x)
            adjv, adjgrad = self.adjust(new_x, new_g, v)
            return adjv, float(adjgrad @ direction)

        shrink = 0.1 if it < 1 else 0.5
        init = 0.5 / float(np.linalg.norm(state.grad)) if it < 1 else 1.0
        return backtracking_line_search(ff, init, state.value, shrink)


class LBFGSB(LBFGS):
    """breeze.optimize.LBFGSB(lowerBounds, upperBounds, maxIter, m, tolerance, maxZoomIter = 64, maxLineSearchIter = 64):
    Byrd, Lu, Nocedal & Zhu's L-BFGS-B as Breeze 0.13.2 states it -- generalized Cauchy point (algorithm CP), subspace
    minimisation by the primal method, compact representation B = theta I - W M W^T with W = [Y, theta S], strong-Wolfe
    line search capped at the largest feasible step.  LinearRegression uses it for loss = huber, where only the scale
    sigma has a bound (sigma >= Double.MinPositiveValue).

    One deliberate difference: the subspace step is shortened by Byrd et al.'s alpha* (largest alpha <= 1 that stays
    inside the bounds).  Breeze 0.13.2's findAlpha, as this restatement recalls it, takes max(alpha, min(., 1)) starting
    from alpha = 1 and therefore never shortens the step; an infeasible subspace minimum then makes the largest feasible
    line-search step smaller than the initial step 1 and StrongWolfeLineSearch.minimizeWithBound fails its
    `require(init <= bound)` -- the fit dies with an IllegalArgumentException.  Wherever Breeze's version succeeds
    (alpha* = 1) the two are the same algorithm; alpha_cuts counts the steps where they are not."""

    PROJ_GRADIENT_EPS = 1e-5
    EPS = 2.2e-16

    def __init__(self, lower, upper, max_iter, m, tol, max_zoom_iter=64, max_line_search_iter=64):
        super().__init__(max_iter, m, tol)
        self.lower = np.asarray(lower, dtype=np.float64)
        self.upper = np.asarray(upper, dtype=np.float64)
        self.max_zoom_iter, self.max_ls_iter = max_zoom_iter, max_line_search_iter
        self.alpha_cuts = 0

    # history = (theta, W, M, Y, S) with Y, S of shape (n, k), newest column last
    def initial_history(self, x):
        n = x.size
        return (1.0, np.zeros((n, 1)), np.zeros((1, 1)), np.zeros((n, 0)), np.zeros((n, 0)))

    def update_history(self, state, x, grad):
        theta, w_mat, m_mat, y_hist, s_hist = state.history
        new_s, new_y = x - state.x, grad - state.grad
        if y_hist.shape[1] < self.m:
            y_new = np.concatenate([y_hist, new_y[:, None]], axis=1)
            s_new = np.concatenate([s_hist, new_s[:, None]], axis=1)
        else:
            y_new = np.concatenate([y_hist[:, 1:], new_y[:, None]], axis=1)
            s_new = np.concatenate([s_hist[:, 1:], new_s[:, None]], axis=1)
        curvature = abs(float(new_s @ new_y))
        if self.EPS * float(new_y @ new_y) < curvature:
            theta_new = float(new_y @ new_y) / float(new_y @ new_s)
            w_new = np.concatenate([y_new, theta_new * s_new], axis=1)
            a = s_new.T @ y_new
            low = np.tril(a, -1)
            dmat = -np.diag(np.diag(a))
            mm = np.block([[dmat, low.T], [low, theta_new * (s_new.T @ s_new)]])
            return (theta_new, w_new, np.linalg.inv(mm), y_new, s_new)
        return (theta, w_mat, m_mat, y_hist, s_hist)

    def _cauchy_point(self, history, x, g):
        theta, w_mat, m_mat = history[0], history[1], history[2]
        n = x.size
        d = np.zeros(n)
        t = np.full(n, np.finfo(np.float64).max)
        for i in range(n):
            gi = g[i]
            if gi != 0.0:
                ti = (x[i] - self.upper[i]) / gi if gi < 0 else (x[i] - self.lower[i]) / gi
                d[i] = 0.0 if ti == 0.0 else -gi
                t[i] = ti
        x_cauchy = x.copy()
        p = w_mat.T @ d
        c = np.zeros(m_mat.shape[0])
        f_deriv = float(g @ d)
        f_second = -theta * f_deriv - float(p @ (m_mat @ p))
        dt_min = -(f_deriv / f_second)
        old_t = 0.0
        order = sorted(range(n), key=lambda i: t[i])  # stable, like sortWith on equal keys? (ties are measure-zero here)
        i = next((k for k, idx in enumerate(order) if t[idx] != 0.0), n)
        if i < n:
            b = order[i]
            min_t = t[b]
            delta_t = min_t - old_t
            while delta_t <= dt_min and i < n:
                x_cauchy[b] = self.upper[b] if d[b] > 0 else self.lower[b]
                zb = x_cauchy[b] - x[b]
                c = c + p * delta_t
                wb = w_mat[b, :].copy()
                f_deriv += delta_t * f_second + g[b] * g[b] + theta * g[b] * zb - g[b] * float(wb @ (m_mat @ c))
                f_second += -theta * g[b] * g[b] - 2.0 * (g[b] * float(wb @ (m_mat @ p))) - g[b] * g[b] * float(wb @ (m_mat @ wb))
                p = p + wb * g[b]
                d[b] = 0.0
                dt_min = -f_deriv / f_second
                old_t = min_t
                i += 1
                if i < n:
                    b = order[i]
                    min_t = t[b]
                    delta_t = min_t - old_t
        dt_min = max(dt_min, 0.0)
        old_t += dt_min
        for k in range(i, n):
            idx = order[k]
            x_cauchy[idx] = x[idx] + old_t * d[idx]
        c = c + p * dt_min
        return x_cauchy, c

    def _subspace_min(self, history, x_cauchy, x, c, g):
        theta, w_mat, m_mat = history[0], history[1], history[2]
        inv_theta = 1.0 / theta
        free = [i for i in range(x.size) if x_cauchy[i] != self.upper[i] and x_cauchy[i] != self.lower[i]]
        wz = w_mat[free, :].T  # (2k, free)
        full_r = g + (x_cauchy - x) * theta - w_mat @ (m_mat @ c)
        rc = full_r[free]
        v = m_mat @ (wz @ rc)
        n_mat = np.eye(wz.shape[0]) - m_mat @ ((wz @ wz.T) * inv_theta)
        v = np.linalg.solve(n_mat, v)
        wzv = (wz.T @ v) * (inv_theta * inv_theta)
        du = -(rc * inv_theta + wzv)
        # findAlpha, as Byrd et al. define it: the largest alpha <= 1 that keeps x_cauchy + alpha du inside the bounds
        alpha_star = 1.0
        with np.errstate(over="ignore"):  # overflow to inf against the +-Double.MaxValue bounds is harmless under min
            for k, idx in enumerate(free):
                if du[k] > 0:
                    alpha_star = min(alpha_star, (self.upper[idx] - x_cauchy[idx]) / du[k])
                elif du[k] < 0:
                    alpha_star = min(alpha_star, (self.lower[idx] - x_cauchy[idx]) / du[k])
        self.alpha_cuts += int(alpha_star < 1.0)
        out = x_cauchy.copy()
        for k, idx in enumerate(free):
            out[idx] += alpha_star * du[k]
        return out

    def choose_descent_direction(self, state, grad=None):
        x, g = state.x, state.grad
        with np.errstate(over="ignore"):  # (x - Double.MaxValue) / g overflows to +-inf, as it does on the JVM
            x_cauchy, c = self._cauchy_point(state.history, x, g)
        if state.iter == 0:
            return x_cauchy - x
        return self._subspace_min(state.history, x_cauchy, x, c, g) - x

    def take_step(self, state, direction, step):
        return np.minimum(np.maximum(state.x + direction * step, self.lower), self.upper)

    def determine_step_size(self, state, f, direction):
        def ff(alpha):
            v, g = f(self.take_step(state, direction, alpha))
            return v, float(g @ direction)

        min_bound = float("inf")
        with np.errstate(over="ignore"):
            for i in range(direction.size):
                di = direction[i]
                if di != 0.0:
                    bnd = self.lower[i] if di < 0.0 else self.upper[i]
                    step_bound = (bnd - state.x[i]) / di
                    assert step_bound >= 0.0
                    min_bound = min(min_bound, step_bound)
        return strong_wolfe_line_search(ff, 1.0, self.max_zoom_iter, self.max_ls_iter, bound=min_bound)

    def _converged(self, s):
        proj = np.minimum(np.maximum(s.x - s.grad, self.lower), self.upper) - s.x
        if abs(float(np.max(np.abs(proj)))) < self.PROJ_GRADIENT_EPS:
            return "ProjectedStepConverged"
        return super()._converged(s)


def run_optimizer(opt, cost_fun, x0):
    f = CachedDiffFunction(cost_fun)
    hist = []
    state = None
    for state in opt.iterations(f, np.array(x0, dtype=np.float64)):
        hist.append(state.adjusted_value)
    return state, hist, f.evals


# ----------------------------------------------------------------------------- packed Cholesky (LAPACK dppsv 'U')

class SingularMatrixException(Exception):
    pass


def cholesky_solve(aa_packed, ab):
    """mllib.linalg.CholeskyDecomposition.solve -> LAPACK dppsv('U'): packed upper, column-major."""
    k = ab.size
    A = np.array(aa_packed, dtype=np.float64)
    b = np.array(ab, dtype=np.float64)
    if k > 16:  # the real thing: LAPACK dppsv('U') on the packed matrix, exactly what Spark calls
        try:
            from scipy.linalg import lapack
            x, info = lapack.dppsv(k, A, b, lower=0)
            if info > 0:
                raise SingularMatrixException(f"leading minor of order {info} is not positive definite")
            if info == 0:
                return np.asarray(x, dtype=np.float64).reshape(-1)
        except ImportError:
            pass

    def idx(i, j):  # i <= j
        return i + j * (j + 1) // 2

    for j in range(k):  # dpptrf upper: U^T U
        s = A[idx(j, j)]
        for i in range(j):
            t = A[idx(i, j)]
            for l in range(i):
                t -= A[idx(l, i)] * A[idx(l, j)]
            t /= A[idx(i, i)]
            A[idx(i, j)] = t
            s -= t * t
        if not (s > 0.0):
            raise SingularMatrixException(f"leading minor of order {j + 1} is not positive definite")
        A[idx(j, j)] = math.sqrt(s)
    for i in range(k):  # U^T y = b
        t = b[i]
        for l in range(i):
            t -= A[idx(l, i)] * b[l]
        b[i] = t / A[idx(i, i)]
    for i in range(k - 1, -1, -1):  # U x = y
        t = b[i]
        for l in range(i + 1, k):
            t -= A[idx(i, l)] * b[l]
        b[i] = t / A[idx(i, i)]
    return b


def _spmv_upper(k, ap, x):
    """BLAS dspmv 'U' y = A x for packed upper column-major A."""
    y = np.zeros(k)
    p = 0
    for j in range(k):
        for i in range(j + 1):
            a = ap[p]
            p += 1
            y[i] += a * x[j]
            if i != j:
                y[j] += a * x[i]
    return y


def quasi_newton_solve(b_bar, bb_bar, ab, aa, a_bar, fit_intercept, max_iter, tol, l1_reg_func):
    """ml.optim.QuasiNewtonSolver.solve with NormalEquationCostFun (intercept slot overwritten in place)."""
    nf = a_bar.size
    k = nf + 1 if fit_intercept else nf
    x0 = np.zeros(k)
    if fit_intercept:
        x0[k - 1] = b_bar

    def cost(coef):  # coef aliases the optimiser's vector, as Vectors.fromBreeze does
        if fit_intercept:
            dot = 0.0
            for j in range(nf):
                dot += coef[j] * a_bar[j]
            coef[nf] = b_bar - dot
        aax = _spmv_upper(k, aa, coef)
        loss = 0.5 * bb_bar - float(coef @ ab) + 0.5 * float(coef @ aax)
        return loss, aax - ab

    opt = OWLQN(max_iter, 10, l1_reg_func, tol) if l1_reg_func is not None else LBFGS(max_iter, 10, tol)
    state, hist, _ = run_optimizer(opt, cost, x0)
    return state.x.copy(), hist


# ----------------------------------------------------------------------------- WeightedLeastSquares.fit

@dataclass
class LRParams:
    max_iter: int = 100
    reg_param: float = 0.0
    elastic_net_param: float = 0.0
    tol: float = 1e-6
    fit_intercept: bool = True
    standardization: bool = True
    solver: str = "auto"  # auto | normal | l-bfgs
    loss: str = "squaredError"  # squaredError | huber
    epsilon: float = 1.35


@dataclass
class LRModel:
    coefficients: np.ndarray
    intercept: float
    objective_history: list
    total_iterations: int
    solver_used: str
    data_passes: int = 0
    scale: float = 1.0  # LinearRegressionModel.scale: the huber loss's sigma (1.0 for squaredError)


def wls_fit_from_moments(mom, d, p: LRParams, solver_type="auto"):
    """ml.optim.WeightedLeastSquares.fit after the treeAggregate (standardizeLabel = true as LinearRegression passes)."""
    count, w_sum, ww_sum, b_sum, bb_sum = mom[:5]
    a_sum = np.array(mom[5:5 + d])
    ab_sum = np.array(mom[5 + d:5 + 2 * d])
    aa_sum = np.array(mom[5 + 2 * d:])
    tri_k = d * (d + 1) // 2
    if count <= 0 or w_sum <= 0:
        raise ValueError("Training dataset is empty / sum of weights is not positive")  # Aggregator.validate
    fit_intercept = p.fit_intercept
    raw_b_bar = b_sum / w_sum
    raw_b_std = math.sqrt(max(bb_sum / w_sum - raw_b_bar * raw_b_bar, 0.0))
    b_std = abs(raw_b_bar) if raw_b_std == 0.0 else raw_b_std
    if raw_b_std == 0:
        if fit_intercept or raw_b_bar == 0.0:
            return LRModel(np.zeros(d), raw_b_bar, [0.0], 1, "trivial")
        if p.reg_param > 0.0:
            raise ValueError("The standard deviation of the label is zero. Model cannot be regularized with standardization=true")
    b_bar = raw_b_bar / b_std
    bb_bar = (bb_sum / w_sum) / (b_std * b_std)
    a_std = np.zeros(d)
    i, j = 0, 2
    while i < tri_k:
        l = j - 2
        aw = a_sum[l] / w_sum
        a_std[l] = math.sqrt(max(aa_sum[i] / w_sum - aw * aw, 0.0))
        i += j
        j += 1
    inv_w = 1.0 / w_sum  # BLAS.scal(1.0 / wSum, ...)
    a_bar = a_sum * inv_w
    ab_bar = ab_sum * inv_w
    aa_bar = aa_sum * inv_w
    for q in range(d):
        if a_std[q] == 0.0:
            a_bar[q] = 0.0
            ab_bar[q] = 0.0
        else:
            a_bar[q] /= a_std[q]
            ab_bar[q] /= (a_std[q] * b_std)
    pidx = 0
    for jj in range(d):
        for ii in range(jj + 1):
            if a_std[jj] == 0.0 or a_std[ii] == 0.0:
                aa_bar[pidx] = 0.0
            else:
                aa_bar[pidx] /= (a_std[ii] * a_std[jj])
            pidx += 1
    eff_reg = p.reg_param / b_std
    eff_l1 = p.elastic_net_param * eff_reg
    eff_l2 = (1.0 - p.elastic_net_param) * eff_reg
    i, j = 0, 2
    while i < tri_k:
        lam = eff_l2
        if not p.standardization:
            std = a_std[j - 2]
            lam = lam / (std * std) if std != 0.0 else 0.0
        aa_bar[i] += lam
        i += j
        j += 1

    def get_ata():
        return np.concatenate([aa_bar, a_bar, [1.0]]) if fit_intercept else aa_bar.copy()

    def get_atb():
        return np.concatenate([ab_bar, [b_bar]]) if fit_intercept else ab_bar.copy()

    use_qn = (solver_type == "auto" and p.elastic_net_param != 0.0 and p.reg_param != 0.0) or solver_type == "quasi-newton"
    l1_func = None
    if use_qn and eff_l1 != 0.0:
        def l1_func(index):
            if fit_intercept and index == d:
                return 0.0
            if p.standardization:
                return eff_l1
            return eff_l1 / a_std[index] if a_std[index] != 0.0 else 0.0
    if use_qn:
        sol, hist = quasi_newton_solve(b_bar, bb_bar, get_atb(), get_ata(), a_bar, fit_intercept, p.max_iter, p.tol, l1_func)
        used = "wls-owlqn" if l1_func is not None else "wls-lbfgs"
    else:
        try:
            sol = cholesky_solve(get_ata(), get_atb())
            hist = [0.0]
            used = "wls-cholesky"
        except SingularMatrixException:
            if solver_type != "auto":
                raise
            sol, hist = quasi_newton_solve(b_bar, bb_bar, get_atb(), get_ata(), a_bar, fit_intercept, p.max_iter, p.tol, None)
            used = "wls-lbfgs-fallback"
    if fit_intercept:
        coef = sol[:-1].copy()
        intercept = sol[-1] * b_std
    else:
        coef = sol.copy()
        intercept = 0.0
    for q in range(d):
        coef[q] *= (b_std / a_std[q]) if a_std[q] != 0.0 else 0.0
    return LRModel(coef, float(intercept), list(hist), len(hist), used, 1)


# ----------------------------------------------------------------------------- LinearRegression.train, l-bfgs branch

# ---- instance weights (LinearRegression.setWeightCol): numpy restatements, small inputs only --------------------------
# The lab never sets weightCol and Spark's sources are not under /root/reference (Maven dependency spark-mllib_2.11
# 2.4.4): these follow the published algorithm of ml.optim.WeightedLeastSquares.Aggregator.add,
# mllib.stat.MultivariateOnlineSummarizer.add(sample, weight) / .mean / .variance and
# ml.optim.aggregator.LeastSquaresAggregator.add.  PARITY UNPINNED, like the rest of the oracle.

def _dq_keep_mask(dq, n):
    if dq is None:
        return np.ones(n, dtype=bool)
    g = np.asarray(dq[0], dtype=np.int32)
    pr = np.asarray(dq[1], dtype=np.float64)
    with np.errstate(invalid="ignore"):
        return ~(pr < 20.0) & ~(pr <= 0.0) & ~((g < 14) & (pr > 90.0))


def _weighted_table(features, label, weight, dq):
    y = np.asarray(label, dtype=np.float64)
    w = np.asarray(weight, dtype=np.float64)
    if np.any(~(w >= 0.0)):
        raise ValueError("instance weight must be >= 0")  # require(w >= 0.0) in Aggregator.add
    keep = _dq_keep_mask(dq, y.size)
    x = np.stack([np.asarray(f, dtype=np.float64)[keep] for f in features], axis=1) if len(features) else np.zeros((int(keep.sum()), 0))
    return x, y[keep], w[keep]


def wls_moments_weighted(features, label, weight, dq=None):
    """WeightedLeastSquares.Aggregator.add over every kept row (zero-weight rows still count), public moment layout."""
    x, y, w = _weighted_table(features, label, weight, dq)
    d = x.shape[1]
    aa = (x * w[:, None]).T @ x
    packed = [aa[i, j] for j in range(d) for i in range(j + 1)]
    return np.array([float(y.size), w.sum(), (w * w).sum(), (w * y).sum(), (w * y * y).sum()]
                    + list((x * w[:, None]).sum(axis=0)) + list((x * (w * y)[:, None]).sum(axis=0)) + packed)


def summarizer_weighted(features, label, weight, dq=None):
    """MultivariateOnlineSummarizer with weights: rows with w == 0 are skipped; variance = sum w (v - mean)^2 /
    (W - sum w^2 / W) when that denominator is positive, else 0."""
    x, y, w = _weighted_table(features, label, weight, dq)
    on = w != 0.0
    x, y, w = x[on], y[on], w[on]
    cols = np.concatenate([x, y[:, None]], axis=1)
    w_sum, ww_sum = w.sum(), (w * w).sum()
    mean = (cols * w[:, None]).sum(axis=0) / w_sum if w_sum > 0 else np.zeros(cols.shape[1])
    denom = w_sum - ww_sum / w_sum if w_sum > 0 else 0.0
    m2n = (w[:, None] * (cols - mean) ** 2).sum(axis=0)
    var = np.maximum(m2n / denom, 0.0) if denom > 0 else np.zeros(cols.shape[1])
    return {"count": float(w.size), "mean": mean, "variance": var, "weight_sum": w_sum}


def lsq_loss_grad_weighted(features, label, weight, coef, fstd, fmean, ystd, ymean, fit_intercept=True, dq=None):
    """LeastSquaresAggregator.add: diff = x.(c/sigma) - y/sigma_y + offset; lossSum += w diff^2/2; gradSum_j += w diff x_j/sigma_j."""
    x, y, w = _weighted_table(features, label, weight, dq)
    fstd = np.asarray(fstd, dtype=np.float64)
    eff = np.where(fstd != 0.0, np.asarray(coef, dtype=np.float64) / np.where(fstd != 0.0, fstd, 1.0), 0.0)
    offset = (ymean / ystd - float(eff @ np.asarray(fmean))) if fit_intercept else 0.0
    diff = x @ eff - y / ystd + offset
    grad = np.where(fstd != 0.0, ((w * diff)[:, None] * x).sum(axis=0) / np.where(fstd != 0.0, fstd, 1.0), 0.0)
    return 0.5 * float((w * diff * diff).sum()), float(w.sum()), grad


def lbfgs_path_fit(features, label, p: LRParams, dq=None, threads=1, weight=None):
    d = len(features)
    if weight is None:
        summ = passes.summarizer(features, label, dq=dq, threads=threads)
    else:
        summ = summarizer_weighted(features, label, weight, dq=dq)
    n_pass = 1
    y_mean = float(summ["mean"][d])
    raw_y_std = math.sqrt(float(summ["variance"][d]))
    if raw_y_std == 0.0:
        if p.fit_intercept or y_mean == 0.0:
            return LRModel(np.zeros(d), y_mean, [0.0], 1, "trivial", n_pass)
        if p.reg_param != 0.0:
            raise ValueError("The standard deviation of the label is zero. Model cannot be regularized.")
    y_std = raw_y_std if raw_y_std > 0 else abs(y_mean)
    f_mean = summ["mean"][:d].copy()
    f_std = np.sqrt(summ["variance"][:d])
    eff_reg = p.reg_param / y_std
    eff_l1 = p.elastic_net_param * eff_reg
    eff_l2 = (1.0 - p.elastic_net_param) * eff_reg
    evals = [0]

    def cost(c):
        if weight is None:
            loss_sum, w_sum, grad_sum = passes.lsq_loss_grad(features, label, c, f_std, f_mean, y_std, y_mean,
                                                             p.fit_intercept, dq=dq, threads=threads)
        else:
            loss_sum, w_sum, grad_sum = lsq_loss_grad_weighted(features, label, weight, c, f_std, f_mean, y_std, y_mean,
                                                               p.fit_intercept, dq=dq)
        evals[0] += 1
        loss = loss_sum / w_sum
        grad = grad_sum / w_sum
        if eff_l2 != 0.0:  # L2Regularization over the d feature coefficients
            s = 0.0
            for j in range(d):
                if p.standardization:
                    s += c[j] * c[j]
                    grad[j] += eff_l2 * c[j]
                elif f_std[j] != 0.0:
                    temp = c[j] / (f_std[j] * f_std[j])
                    s += c[j] * temp
                    grad[j] += eff_l2 * temp
            loss += 0.5 * s * eff_l2
        return loss, grad

    if p.elastic_net_param == 0.0 or eff_reg == 0.0:
        opt = LBFGS(p.max_iter, 10, p.tol)
        used = "lbfgs"
    else:
        def l1(index):
            if p.standardization:
                return eff_l1
            return eff_l1 / f_std[index] if f_std[index] != 0.0 else 0.0
        opt = OWLQN(p.max_iter, 10, l1, p.tol)
        used = "owlqn"
    state, hist, n_eval = run_optimizer(opt, cost, np.zeros(d))
    raw = state.x.copy()
    for j in range(d):
        raw[j] *= (y_std / f_std[j]) if f_std[j] != 0.0 else 0.0
    intercept = (y_mean - float(raw @ f_mean)) if p.fit_intercept else 0.0
    return LRModel(raw, float(intercept), list(hist), len(hist), used, n_pass + n_eval)


def huber_loss_grad(x, y, w, params, fstd, fit_intercept, epsilon):
    """ml.optim.aggregator.HuberAggregator.add over every row with w != 0: returns (lossSum, weightSum, gradientSum)
    with gradientSum laid out like the parameters [coefficients (d), intercept (if fitted), sigma].  The label is NOT
    standardised; features enter as value / featuresStd (columns with std 0 are skipped)."""
    d = x.shape[1]
    params = np.asarray(params, dtype=np.float64)
    sigma = params[-1]
    intercept = params[d] if fit_intercept else 0.0
    on = w != 0.0
    x, y, w = x[on], y[on], w[on]
    ok = fstd != 0.0
    xs = np.where(ok, x / np.where(ok, fstd, 1.0), 0.0)
    margin = xs @ params[:d] + intercept
    lin = y - margin
    inner = np.abs(lin) <= sigma * epsilon
    with np.errstate(over="ignore", invalid="ignore"):  # a line-search trial may put sigma next to its lower bound
        lds = lin / sigma
        lds2 = lin * lin / sigma
    loss = np.where(inner, 0.5 * w * (sigma + lds2),
                    0.5 * w * (sigma + 2.0 * epsilon * np.abs(lin) - sigma * epsilon * epsilon))
    sign = np.where(lin >= 0, -1.0, 1.0)
    factor = np.where(inner, -1.0 * w * lds, w * sign * epsilon)  # multiplies value / std (and 1 for the intercept)
    grad = np.zeros(params.size)
    grad[:d] = (factor[:, None] * xs).sum(axis=0)
    if fit_intercept:
        grad[d] = factor.sum()
    grad[-1] = np.where(inner, 0.5 * w * (1.0 - lds * lds), 0.5 * w * (1.0 - epsilon * epsilon)).sum()
    return float(loss.sum()), float(w.sum()), grad


def huber_path_fit(features, label, p: LRParams, dq=None, threads=1, weight=None):
    """LinearRegression.train with loss = huber (Spark 2.3+): summarizer for the feature scales, HuberAggregator per
    evaluation, L2 regularisation only (regParam NOT divided by yStd), Breeze LBFGSB with sigma > 0, start at all ones."""
    d = len(features)
    if p.elastic_net_param != 0.0:
        raise ValueError("requirement failed: LinearRegression with huber loss only supports L2 regularization")
    if p.solver == "normal":
        raise ValueError("requirement failed: LinearRegression with huber loss doesn't support normal solver")
    if not p.epsilon > 1.0:
        raise ValueError("epsilon must be > 1.0")
    if weight is None:
        summ = passes.summarizer(features, label, dq=dq, threads=threads)
    else:
        summ = summarizer_weighted(features, label, weight, dq=dq)
    y_mean = float(summ["mean"][d])
    raw_y_std = math.sqrt(float(summ["variance"][d]))
    if raw_y_std == 0.0:
        if p.fit_intercept or y_mean == 0.0:
            return LRModel(np.zeros(d), y_mean, [0.0], 1, "trivial", 1)
        if p.reg_param != 0.0:
            raise ValueError("The standard deviation of the label is zero. Model cannot be regularized.")
    f_std = np.sqrt(summ["variance"][:d])
    eff_l2 = p.reg_param  # (1 - elasticNetParam) * regParam with elasticNetParam == 0
    y = np.asarray(label, dtype=np.float64)
    w_all = np.ones(y.size) if weight is None else np.asarray(weight, dtype=np.float64)
    x, yk, wk = _weighted_table(features, label, w_all, dq)
    dim = d + 2 if p.fit_intercept else d + 1

    def cost(c):
        loss_sum, w_sum, grad_sum = huber_loss_grad(x, yk, wk, c, f_std, p.fit_intercept, p.epsilon)
        loss = loss_sum / w_sum
        grad = grad_sum / w_sum
        if eff_l2 != 0.0:
            ssum = 0.0
            for j in range(d):
                if p.standardization:
                    ssum += c[j] * c[j]
                    grad[j] += eff_l2 * c[j]
                elif f_std[j] != 0.0:
                    temp = c[j] / (f_std[j] * f_std[j])
                    ssum += c[j] * temp
                    grad[j] += eff_l2 * temp
            loss += 0.5 * ssum * eff_l2
        return loss, grad

    lower = np.full(dim, -np.finfo(np.float64).max)
    lower[dim - 1] = 5e-324  # Double.MinPositiveValue
    upper = np.full(dim, np.finfo(np.float64).max)
    opt = LBFGSB(lower, upper, p.max_iter, 10, p.tol)
    state, hist, n_eval = run_optimizer(opt, cost, np.ones(dim))
    par = state.x.copy()
    raw = par[:d].copy()
    for j in range(d):
        raw[j] *= (1.0 / f_std[j]) if f_std[j] != 0.0 else 0.0
    intercept = float(par[d]) if p.fit_intercept else 0.0
    model = LRModel(raw, intercept, list(hist), len(hist), "lbfgsb", 1 + n_eval, scale=float(par[-1]))
    model.alpha_cuts = opt.alpha_cuts  # steps on which Breeze 0.13.2 (as recalled) would have thrown instead
    return model


def linear_regression_fit(features, label, p: LRParams, dq=None, threads=1, weight=None):
    """ml.regression.LinearRegression.train: solver selection (auto & d <= 4096 -> WLS, else l-bfgs)."""
    d = len(features)
    if p.loss == "huber":
        return huber_path_fit(features, label, p, dq=dq, threads=threads, weight=weight)
    if (p.solver == "auto" and d <= 4096) or p.solver == "normal":
        if weight is None:
            mom = passes.wls_moments(features, label, dq=dq, threads=threads)
        else:
            mom = wls_moments_weighted(features, label, weight, dq=dq)
        return wls_fit_from_moments(mom, d, p, "auto")
    return lbfgs_path_fit(features, label, p, dq=dq, threads=threads, weight=weight)


def regression_metrics(metrics_out, fit_intercept=True):
    """mllib.evaluation.RegressionMetrics: RMSE, r2, MAE, MSE from [n, SSerr, sum_y, sum_yy, sum_abs_err]."""
    n, ss_err, sy, syy, sabs = [float(v) for v in metrics_out]
    mse = ss_err / n
    ss_tot = (syy - sy * sy / n) if fit_intercept else syy
    return {"rmse": math.sqrt(mse), "mse": mse, "r2": 1.0 - ss_err / ss_tot, "mae": sabs / n}
