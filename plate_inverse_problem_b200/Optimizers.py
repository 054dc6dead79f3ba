"""Optimiser loops over the fused CUDA evaluations: the drivers of
`/root/reference/source/jax_plate/Optimizers.py` (gradient descent `:231-254`, coordinate
descent `:257-367`, trust-region Newton `:153-228` with its sub-problem solver `:70-122`),
plus the Gauss-Newton / Levenberg-Marquardt loop the reference only sketches (`:139-144`).

The reference obtains derivatives with `jax.value_and_grad(f)` / `jax.jacobian(jax.grad(f))`;
jax is not part of this image, so an objective here is a callable that carries its own
derivatives -- exactly what `Problem.getLossFunction` returns:

    f(x) -> float                      f.value_and_grad(x) -> (float, grad[P])
    f.gauss_newton(x) -> (f, g, H_gn)  (optional: the model `2 r^T J / n, 2 J^T J / n`)
    f.hessian(x) -> H[P, P]            (optional: exact Hessian for method="newt")

Every iteration is ONE batched native call (`jpb_eval_loss_grad` / `jpb_eval_jacobian`):
assembly, factorisation, both solves and the contraction stay on the device; the host sees
`coef[F, K]` going in and `ct_coef[F, K]` coming back (170 KB at F = 400) and applies the
chain rule through the material transform.  Iterates, step rules, stopping tests and the
`optResult` record follow the reference so that histories can be compared step by step
(tests/test_host.py does that against the reference's own loops on analytic objectives).
"""
from __future__ import annotations

from collections import namedtuple

import numpy as np

optResult = namedtuple("optResult", ["x", "f", "f_history", "x_history", "grad_history", "niter", "status"])


def _value_and_grad(f):
    if hasattr(f, "value_and_grad"):
        return f.value_and_grad
    raise TypeError("the objective must carry .value_and_grad(x) (e.g. Problem.getLossFunction(...)); "
                    "there is no automatic differentiation here")


class FixedParameterFunction:
    """`Optimizers.py:9-56`: `function` with some of its parameters pinned to constants."""

    def __init__(self, function, param_size: int, fixed_indices, fixed_values):
        if isinstance(fixed_indices, (int, np.integer)):
            fixed_indices, fixed_values = (int(fixed_indices),), (float(fixed_values),)
        assert len(fixed_indices) == len(fixed_values)
        self.func = function
        self.array = np.zeros(param_size)
        self.array[list(fixed_indices)] = fixed_values
        self.free_idx = np.array([i for i in range(param_size) if i not in set(fixed_indices)])

    def full(self, params):
        x = self.array.copy()
        x[self.free_idx] = params
        return x

    def __call__(self, params, *args):
        return self.func(self.full(params), *args)

    def value_and_grad(self, params):
        v, g = _value_and_grad(self.func)(self.full(params))
        return v, np.asarray(g)[self.free_idx]


# ---------------------------------------------------------------- trust-region sub-problem
def _step(B, g, lam):
    sd = np.linalg.solve(B + lam * np.eye(B.shape[0]), -g)
    return sd, float(np.linalg.norm(sd))


def solve_trust_region_model(B, g, delta, rtol=1e-6, max_iter=100):
    """min_p g.p + p.B.p / 2 subject to |p| <= delta (`Optimizers.py:70-122`): the Newton step
    when it is interior and B is positive semi-definite, else the boundary solution
    p(lam) = -(B + lam I)^-1 g with lam found by bracketing (doubling) and bisection on
    |p(lam)| = delta.  Returns (p, lam, predicted decrease)."""
    B, g = np.asarray(B, dtype=np.float64), np.asarray(g, dtype=np.float64)
    eig = np.linalg.eigvalsh(B)
    sd, norm = _step(B, g, 0.0)
    if norm <= delta and eig.min() >= 0:
        return sd, 0.0, float(-(g @ sd + 0.5 * sd @ B @ sd))
    lo = max(0.0, float((-eig).max()))
    hi = lo + 1.0
    sd, norm = _step(B, g, hi)
    for _ in range(max_iter):
        if norm <= delta:
            break
        lo, hi = hi, 2.0 * hi
        sd, norm = _step(B, g, hi)
    assert norm <= delta, "Failed to find upper bound for lambda"
    lam = hi
    for _ in range(2 * max_iter):
        if norm <= delta and delta - norm <= delta * rtol:
            break
        lam = 0.5 * (lo + hi)
        sd, norm = _step(B, g, lam)
        if norm < delta:
            hi = lam
        else:
            lo = lam
    if norm > delta:
        lam = hi
        sd, norm = _step(B, g, lam)
    pred = float(-(g @ sd + 0.5 * sd @ B @ sd))
    assert pred >= 0, "Predicted improvement for quadratic model is negative"
    return sd, lam, pred


def _model(f, method):
    """x -> (f, g, B) for the trust-region / LM loops."""
    vg = _value_and_grad(f)
    if method in ("gn", "lvmq"):
        if not hasattr(f, "gauss_newton"):
            raise TypeError('method "gn" needs an objective with .gauss_newton(x) (Problem.getLossFunction)')
        return f.gauss_newton
    if method == "newt":
        if hasattr(f, "hessian"):
            def update(x):
                v, g = vg(x)
                return v, g, np.asarray(f.hessian(x))
            return update

        def update_fd(x):        # exact-Hessian stand-in: central differences of the adjoint gradient
            v, g = vg(x)
            n = x.size
            H = np.zeros((n, n))
            for j in range(n):
                h = 1e-6 * max(abs(x[j]), 1e-8)
                e = np.zeros(n)
                e[j] = h
                H[:, j] = (np.asarray(vg(x + e)[1]) - np.asarray(vg(x - e)[1])) / (2 * h)
            return v, g, 0.5 * (H + H.T)
        return update_fd
    raise NotImplementedError(f"Method <<{method}>> not implemented")


def optimize_trust_region(f, x_0, N_steps=10, delta_max=1.0, delta=None, eta=0.15, method="newt",
                          steps_to_stall=10):
    """`Optimizers.py:153-228`.  method: "newt" (Hessian from `f.hessian`, else finite
    differences of the gradient) or "gn" (Gauss-Newton model, one Jacobian evaluation)."""
    if delta is None:
        delta = delta_max / 10.0
    if eta > 0.25:
        raise ValueError(f"eta should be below 0.25; got {eta:f}")
    if eta < 0:
        raise ValueError(f"eta should be positive; got {eta:f}")
    update_model = _model(f, method)
    f_hist, x_hist, g_hist = [], [], []
    refresh, status, idle = True, "Running", 0
    x = np.array(x_0, dtype=np.float64)
    k = 0
    for k in range(N_steps):
        if refresh:
            cur_f, g, B = update_model(x)
        try:
            sd, lam, predicted = solve_trust_region_model(B, g, delta)
        except AssertionError as e:
            status = str(e)
            break
        ratio = (cur_f - f(x + sd)) / predicted
        if ratio < 0.25:
            delta /= 4.0
        elif ratio >= 0.75 and lam > 0.0:      # the step was on the boundary
            delta = min(2.0 * delta, delta_max)
        if ratio >= eta:
            x = x + sd
            refresh, idle = True, 0
        else:
            refresh, idle = False, idle + 1
        f_hist.append(cur_f)
        x_hist.append(x.copy())
        g_hist.append(g)
        if cur_f < 1e-16:
            status = "Converged"
            break
        if idle >= steps_to_stall:
            status = "Stalled"
            break
    return optResult(x, cur_f, f_hist, x_hist, g_hist, k, status)


def optimize_lm(f, x_0, N_steps=20, mu=1e-3, mu_up=10.0, mu_down=0.2, f_min=1e-16, gtol=0.0):
    """Levenberg-Marquardt on the Gauss-Newton model (`Optimizers.py:139-144` is the reference's
    commented stub of that model): step p = -(H_gn + mu diag(H_gn))^-1 g, accepted when the loss
    decreases (mu shrinks), otherwise mu grows and the model is reused."""
    model = _model(f, "gn")
    x = np.array(x_0, dtype=np.float64)
    f_hist, x_hist, g_hist = [], [], []
    status = "Running"
    cur_f, g, H = model(x)
    k = 0
    for k in range(N_steps):
        f_hist.append(cur_f)
        x_hist.append(x.copy())
        g_hist.append(g)
        if cur_f <= f_min or np.linalg.norm(g) <= gtol:
            status = "Converged"
            break
        D = np.diag(np.maximum(np.diag(H), 1e-300))
        p = np.linalg.solve(H + mu * D, -g)
        new_f = f(x + p)
        if np.isfinite(new_f) and new_f < cur_f:
            x = x + p
            mu = max(mu * mu_down, 1e-12)
            cur_f, g, H = model(x)
        else:
            mu *= mu_up
            if mu > 1e12:
                status = "Stalled"
                break
    return optResult(x, cur_f, f_hist, x_hist, g_hist, k, status)


def optimize_gd(f, x_0, N_steps=100, h=0.01, f_min=1e-8):
    """`Optimizers.py:231-254`: x <- x - h * grad."""
    vg = _value_and_grad(f)
    x = np.array(x_0, dtype=np.float64)
    x_hist, f_hist, g_hist = [], [], []
    status = "Running"
    k = 0
    for k in range(N_steps):
        cur_f, g = vg(x)
        x_hist.append(x.copy())
        f_hist.append(cur_f)
        g_hist.append(g)
        if cur_f <= f_min:
            status = "Converged"
            break
        x = x - h * np.asarray(g)
    return optResult(x, cur_f, f_hist, x_hist, g_hist, k, status)


def optimize_cd(f, x_0, N_steps=100, h=0.01, f_min=1e-8):
    """`Optimizers.py:257-286`: cyclic coordinate descent, one gradient component per move."""
    vg = _value_and_grad(f)
    x = np.array(x_0, dtype=np.float64)
    n = x.size
    assert n >= 2
    x_hist, f_hist, g_hist = [], [], []
    status = "Running"
    k = 0
    for k in range(N_steps):
        for i in range(n):
            cur_f, g = vg(x)
            gi = np.zeros(n)
            gi[i] = np.asarray(g)[i]
            x_hist.append(x.copy())
            f_hist.append(cur_f)
            g_hist.append(gi)
            if cur_f <= f_min:       # the reference leaves only the inner loop here, too
                status = "Converged"
                break
            x = x - h * gi
    return optResult(x, cur_f, f_hist, x_hist, g_hist, k, status)


def optimize_cd_mem2(f, x_0, N_steps=100, h=0.01, f_min=1e-8):
    """`Optimizers.py:325-367`: coordinate descent with a step size per coordinate that is cut
    by 5 (and the move redone from the previous point) whenever the move increased the loss."""
    vg = _value_and_grad(f)
    x = np.array(x_0, dtype=np.float64)
    n = x.size
    assert n >= 2
    hs = np.full(n, float(h))
    x_hist, f_hist, g_hist = [], [], []
    status = "Running"
    k = 0
    for k in range(N_steps):
        for i in range(n):
            cur_f, g = vg(x)
            gi = np.zeros(n)
            gi[i] = np.asarray(g)[i]
            x_hist.append(x.copy())
            f_hist.append(cur_f)
            g_hist.append(gi)
            if cur_f <= f_min:
                status = "Converged"
                break
            x = x - hs[i] * gi
            if f(x) > f_hist[-1]:
                hs[i] /= 5
                x = x_hist[-1] - hs[i] * gi
    return optResult(x, cur_f, f_hist, x_hist, g_hist, k, status)
