"""Restatement of ``torchdiffeq 0.2.3`` ``odeint`` in plain PyTorch (CPU oracle).

TEST INFRASTRUCTURE.  torchdiffeq is the third-party ODE solver the reference
calls at 19 sites (e.g. NLBAC_Unicycle_RL_training/Unicycle_RL_training/
sac_cbf_clf/model.py:252, sac_cbf_clf/sac_cbf_clf.py:453); it is pinned to
0.2.3 by the reference README.md:33 only, is not vendored under
/root/reference and cannot be installed here (no network).  This file restates
the published algorithm of rtqichen/torchdiffeq (``_impl/odeint.py``,
``fixed_grid.py``, ``rk_common.py``, ``dopri5.py``, ``interp.py``, ``misc.py``)
as summarised in SURVEY.md Appendix A.  Parity of the solver is therefore
"unpinned upstream"; it is anchored in tests/test_oracle_solver.py by analytic
known answers and, for the dopri5 tableau / error weights (up to torchdiffeq's
2/3 scale) / dense output / initial step, by scipy's independent RK45.

Differentiable by ordinary autograd through the tensor ops (discretise-then-
optimise), exactly like plain ``odeint`` (the reference never uses
``odeint_adjoint``).
"""
import math

import torch

SOLVER_NAMES = ("euler", "midpoint", "rk4", "dopri5")


# --------------------------------------------------------------------------- #
# fixed-grid solvers (fixed_grid.py / solvers.py: FixedGridODESolver)
# --------------------------------------------------------------------------- #
def _euler_step(func, t0, dt, t1, y):
    return dt * func(t0, y)


def _midpoint_step(func, t0, dt, t1, y):
    half = 0.5 * dt
    y_mid = y + func(t0, y) * half
    return dt * func(t0 + half, y_mid)


def _rk4_alt_step(func, t0, dt, t1, y):
    """rk4 in torchdiffeq is the 3/8 rule (rk_common.rk4_alt_step_func)."""
    third = 1.0 / 3.0
    k1 = func(t0, y)
    k2 = func(t0 + dt * third, y + dt * k1 * third)
    k3 = func(t0 + dt * (2.0 * third), y + dt * (k2 - k1 * third))
    k4 = func(t1, y + dt * (k1 - k2 + k3))
    return (k1 + 3 * (k2 + k3) + k4) * dt * 0.125


_FIXED = {"euler": _euler_step, "midpoint": _midpoint_step, "rk4": _rk4_alt_step}


def _fixed_grid(t, step_size):
    """FixedGridODESolver._grid_constructor_from_step_size."""
    if step_size is None:
        return t
    start, end = t[0], t[-1]
    niters = int(torch.ceil((end - start) / step_size + 1).item())
    grid = torch.arange(0, niters, dtype=t.dtype, device=t.device) * step_size + start
    grid[-1] = t[-1]
    return grid


def _integrate_fixed(func, y0, t, method, step_size):
    step = _FIXED[method]
    grid = _fixed_grid(t, step_size)
    sol = [y0]
    j = 1
    y = y0
    for t0, t1 in zip(grid[:-1], grid[1:]):
        dt = t1 - t0
        dy = step(func, t0, dt, t1, y)
        y1 = y + dy
        while j < len(t) and t1 >= t[j]:
            if t[j] == t0:
                sol.append(y)
            elif t[j] == t1:
                sol.append(y1)
            else:
                slope = (t[j] - t0) / (t1 - t0)
                sol.append(y + slope * (y1 - y))
            j += 1
        y = y1
    return torch.stack(sol, 0)


# --------------------------------------------------------------------------- #
# dopri5 (dopri5.py / rk_common.py: RKAdaptiveStepsizeODESolver)
# --------------------------------------------------------------------------- #
DP_ALPHA = [1 / 5, 3 / 10, 4 / 5, 8 / 9, 1.0, 1.0]
DP_BETA = [
    [1 / 5],
    [3 / 40, 9 / 40],
    [44 / 45, -56 / 15, 32 / 9],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
    [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84],
]
DP_C_SOL = [35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84, 0]
DP_C_ERR = [
    35 / 384 - 1951 / 21600,
    0,
    500 / 1113 - 22642 / 50085,
    125 / 192 - 451 / 720,
    -2187 / 6784 - -12231 / 42400,
    11 / 84 - 649 / 6300,
    -1.0 / 60.0,
]
DP_C_MID = [
    6025192743 / 30085553152 / 2,
    0,
    51252292925 / 65400821598 / 2,
    -2691868925 / 45128329728 / 2,
    187940372067 / 1594534317056 / 2,
    -1776094331 / 19743644256 / 2,
    11237099 / 235043384 / 2,
]


def _rms_norm(x):
    return x.pow(2).mean().sqrt()


def _select_initial_step(func, t0, y0, order, rtol, atol, norm, f0):
    """misc._select_initial_step; returns a float64 python-side scalar tensor.

    The step size is treated as a constant of the backward pass (SURVEY.md
    App. A.3), so it is computed on detached values.
    """
    with torch.no_grad():
        dtype = y0.dtype
        scale = atol + torch.abs(y0) * rtol
        d0 = norm(y0 / scale).abs()
        d1 = norm(f0 / scale).abs()
        if d0 < 1e-5 or d1 < 1e-5:
            h0 = torch.tensor(1e-6, dtype=dtype)
        else:
            h0 = 0.01 * d0 / d1
        h0 = h0.abs()
        y1 = y0 + h0 * f0
        f1 = func(t0 + h0, y1)
        d2 = torch.abs(norm((f1 - f0) / scale) / h0)
        if d1 <= 1e-15 and d2 <= 1e-15:
            h1 = torch.max(torch.tensor(1e-6, dtype=dtype), h0 * 1e-3)
        else:
            h1 = (0.01 / max(d1, d2)) ** (1.0 / float(order + 1))
        h1 = h1.abs()
        return torch.min(100 * h0, h1).to(torch.float64)


def _optimal_step_size(last_step, error_ratio, safety=0.9, ifactor=10.0, dfactor=0.2, order=5):
    """misc._optimal_step_size (under no_grad upstream)."""
    with torch.no_grad():
        if error_ratio == 0:
            return last_step * ifactor
        if error_ratio < 1:
            dfactor = torch.ones((), dtype=last_step.dtype)
        error_ratio = error_ratio.type_as(last_step)
        exponent = torch.tensor(order, dtype=last_step.dtype).reciprocal()
        factor = torch.min(torch.tensor(ifactor, dtype=last_step.dtype),
                           torch.max(safety / error_ratio ** exponent,
                                     torch.as_tensor(dfactor, dtype=last_step.dtype)))
        return last_step * factor


def _runge_kutta_step(func, y0, f0, t0, dt, t1):
    """rk_common._runge_kutta_step for the dopri5 tableau (k kept as a list)."""
    t0 = t0.to(y0.dtype)
    dt = dt.to(y0.dtype)
    t1 = t1.to(y0.dtype)
    k = [f0]
    for i, (alpha_i, beta_i) in enumerate(zip(DP_ALPHA, DP_BETA)):
        ti = t1 if alpha_i == 1.0 else t0 + alpha_i * dt
        acc = 0
        for b, kj in zip(beta_i, k):
            if b != 0:
                acc = acc + kj * (b * dt)
        yi = y0 + acc
        k.append(func(ti, yi))
    y1 = yi  # c_sol == beta[-1] ++ [0]  (Dormand–Prince)
    f1 = k[-1]
    err = 0
    for c, kj in zip(DP_C_ERR, k):
        if c != 0:
            err = err + kj * (c * dt)
    return y1, f1, err, k


def _interp_fit(y0, y1, y_mid, f0, f1, dt):
    """interp._interp_fit: 4th-order polynomial through the step."""
    a = 2 * dt * (f1 - f0) - 8 * (y1 + y0) + 16 * y_mid
    b = dt * (5 * f0 - 3 * f1) + 18 * y0 + 14 * y1 - 32 * y_mid
    c = dt * (f1 - 4 * f0) - 11 * y0 - 5 * y1 + 16 * y_mid
    d = dt * f0
    e = y0
    return [e, d, c, b, a]


def _interp_evaluate(coeffs, t0, t1, t):
    x = ((t - t0) / (t1 - t0)).to(coeffs[0].dtype)
    total = coeffs[0] + x * coeffs[1]
    x_power = x
    for coeff in coeffs[2:]:
        x_power = x_power * x
        total = total + x_power * coeff
    return total


def _integrate_dopri5(func, y0, t, rtol, atol, stats=None, max_num_steps=2 ** 31 - 1, replay_dt=None):
    """`replay_dt` (test hook): a recorded list of accepted step sizes; every step is then taken with the
    recorded size and accepted, which separates arithmetic parity from step-size selection (the error
    estimate is a difference of nearly equal fp32 numbers, so two fp32 implementations legitimately pick
    step sizes that differ in the 3rd digit)."""
    norm = _rms_norm
    t = t.to(torch.float64)
    f0 = func(t[0].to(y0.dtype), y0)
    dt = _select_initial_step(func, t[0].to(y0.dtype), y0.detach(), 4, rtol, atol, norm, f0.detach())
    y, f, t_prev, t_cur = y0, f0, t[0], t[0]
    interp = [y0] * 5
    sol = [y0]
    n_acc = n_rej = 0
    dts = []
    if replay_dt is not None:
        replay_dt = [torch.tensor(float(v), dtype=torch.float64) for v in replay_dt]
        dt = replay_dt[0]
    for i in range(1, len(t)):
        next_t = t[i]
        n_steps = 0
        while next_t > t_cur:
            assert n_steps < max_num_steps
            t1 = t_cur + dt
            assert t1 > t_cur, "underflow in dt"
            y1, f1, err, k = _runge_kutta_step(func, y, f, t_cur, dt, t1)
            with torch.no_grad():
                tol = atol + rtol * torch.max(y.abs(), y1.abs())
                ratio = norm(err / tol).abs()
            accept = bool(ratio <= 1) or replay_dt is not None
            if accept:
                dts.append(float(dt))
                dt32 = dt.to(y.dtype)
                acc = 0
                for c, kj in zip(DP_C_MID, k):
                    if c != 0:
                        acc = acc + kj * (c * dt32)
                y_mid = y + acc
                interp = _interp_fit(y, y1, y_mid, f, f1, dt32)
                t_prev, t_cur, y, f = t_cur, t1, y1, f1
                n_acc += 1
            else:
                n_rej += 1
            if replay_dt is None:
                dt = _optimal_step_size(dt, ratio)
            elif n_acc < len(replay_dt):
                dt = replay_dt[n_acc]
            n_steps += 1
        sol.append(_interp_evaluate(interp, t_prev, t_cur, next_t))
    if stats is not None:
        stats["n_accepted"] = n_acc
        stats["n_rejected"] = n_rej
        stats["nfe"] = 2 + 6 * (n_acc + n_rej)
        stats["step_dt"] = dts
    return torch.stack(sol, 0)


# --------------------------------------------------------------------------- #
# entry (odeint.py / misc._check_inputs)
# --------------------------------------------------------------------------- #
def odeint(func, y0, t, *, rtol=1e-7, atol=1e-9, method=None, options=None, event_fn=None, _stats=None):
    """``torchdiffeq.odeint`` as the reference calls it.

    Returns ``[len(t), *y0.shape]`` with ``solution[0] == y0``.
    """
    if event_fn is not None:
        raise NotImplementedError("event handling is not on the NLBAC path")
    if not torch.is_tensor(y0):
        raise NotImplementedError("tuple state is not on the NLBAC path")
    if not torch.is_tensor(t) or t.ndim != 1:
        raise ValueError("t must be one dimensional")
    if method is None:
        method = "dopri5"
    if method not in SOLVER_NAMES:
        raise ValueError('Invalid method "{}". Must be one of {}'.format(
            method, '{"' + '", "'.join(SOLVER_NAMES) + '"}.'))
    options = dict(options or {})
    if t.device != y0.device:
        t = t.to(y0.device)
    d = t[1:] - t[:-1]
    if not (bool((d > 0).all()) or bool((d < 0).all())):
        raise AssertionError("t must be strictly increasing or decreasing")
    if bool((d < 0).all()):
        # reversed time: integrate dy/ds = -f(-s, y) over s = -t (misc._check_inputs)
        base = func
        func = lambda s, y: -base(-s, y)  # noqa: E731
        t = -t
    user = func
    func = lambda tt, y: user(tt.to(y.dtype), y)  # noqa: E731  (_PerturbFunc casts t)
    if method == "dopri5":
        return _integrate_dopri5(func, y0, t, rtol, atol, stats=_stats, replay_dt=options.get("_replay_dt"))
    return _integrate_fixed(func, y0, t, method, options.get("step_size"))
