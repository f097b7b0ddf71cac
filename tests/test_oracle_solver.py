"""Anchor the torchdiffeq restatement (oracle/torchdiffeq_restated.py) with analytic known answers.

torchdiffeq itself is absent so these are the pins: matrix exponential,
convergence orders, the 3/8-rule closed form, step_size gridding, error behaviour.
"""
import math

import pytest
import torch

from oracle.torchdiffeq_restated import odeint, _fixed_grid


class Lin(torch.nn.Module):
    def __init__(self, A):
        super().__init__()
        self.A = torch.nn.Parameter(A)

    def forward(self, t, y):
        return y @ self.A.T


def _A():
    return torch.tensor([[-0.3, 1.1, 0.0], [-1.1, -0.2, 0.4], [0.2, 0.0, -0.5]], dtype=torch.float64)


def test_solution_starts_at_y0_and_shape():
    f = Lin(_A())
    y0 = torch.randn(5, 3, dtype=torch.float64)
    t = torch.linspace(0, 1, 7, dtype=torch.float64)
    for m in ("euler", "midpoint", "rk4", "dopri5"):
        sol = odeint(f, y0, t, method=m)
        assert sol.shape == (7, 5, 3)
        assert torch.equal(sol[0], y0)


@pytest.mark.parametrize("method,order", [("euler", 1), ("midpoint", 2), ("rk4", 4)])
def test_fixed_order_of_convergence(method, order):
    A = _A()
    f = Lin(A)
    y0 = torch.tensor([[1.0, -0.5, 0.25]], dtype=torch.float64)
    exact = y0 @ torch.linalg.matrix_exp(A).T
    errs = []
    for n in (8, 16, 32):
        t = torch.linspace(0, 1, n + 1, dtype=torch.float64)
        errs.append((odeint(f, y0, t, method=method)[-1] - exact).abs().max().item())
    slope1 = math.log2(errs[0] / errs[1])
    slope2 = math.log2(errs[1] / errs[2])
    assert abs(slope1 - order) < 0.35 and abs(slope2 - order) < 0.25, (errs, slope1, slope2)


def test_rk4_is_three_eighths_rule_closed_form():
    # for y' = a y one 3/8-rule step multiplies by 1 + z + z^2/2 + z^3/6 + z^4/24 (any 4-stage order-4 RK),
    # so distinguish it from classic RK4 with a NONLINEAR scalar problem evaluated by hand.
    f = lambda t, y: y * y + t  # noqa: E731
    y0 = torch.tensor([0.7], dtype=torch.float64)
    h = 0.3
    t = torch.tensor([0.0, h], dtype=torch.float64)
    got = odeint(f, y0, t, method="rk4")[-1].item()
    y = 0.7
    k1 = y * y + 0.0
    k2 = (y + h * k1 / 3) ** 2 + h / 3
    k3 = (y + h * (k2 - k1 / 3)) ** 2 + 2 * h / 3
    k4 = (y + h * (k1 - k2 + k3)) ** 2 + h
    want = y + h * (k1 + 3 * (k2 + k3) + k4) / 8
    assert abs(got - want) < 1e-14
    # classic RK4 gives a different number
    c1 = y * y
    c2 = (y + h * c1 / 2) ** 2 + h / 2
    c3 = (y + h * c2 / 2) ** 2 + h / 2
    c4 = (y + h * c3) ** 2 + h
    classic = y + h * (c1 + 2 * c2 + 2 * c3 + c4) / 6
    assert abs(got - classic) > 1e-6


def test_step_size_grid_and_interpolation():
    t = torch.tensor([0.0, 0.05, 0.1])
    g = _fixed_grid(t, 0.02)
    assert torch.allclose(g, torch.tensor([0.0, 0.02, 0.04, 0.06, 0.08, 0.1]))
    f = Lin(_A().float())
    y0 = torch.randn(4, 3)
    sol = odeint(f, y0, t, method="euler", options={"step_size": 0.02})
    # manual: 0.05 lies inside step [0.04, 0.06] → linear interpolation
    y = y0
    ys = [y0]
    for _ in range(5):
        y = y + 0.02 * f(None, y)
        ys.append(y)
    mid = ys[2] + (0.05 - 0.04) / 0.02 * (ys[3] - ys[2])
    assert torch.allclose(sol[1], mid, atol=1e-6)
    assert torch.allclose(sol[2], ys[5], atol=1e-6)


def test_dopri5_accuracy_and_stats():
    A = _A()
    f = Lin(A)
    y0 = torch.randn(6, 3, dtype=torch.float64)
    t = torch.tensor([0.0, 0.7, 2.0], dtype=torch.float64)
    stats = {}
    sol = odeint(f, y0, t, method="dopri5", rtol=1e-8, atol=1e-10, _stats=stats)
    for i in (1, 2):
        exact = y0 @ torch.linalg.matrix_exp(A * t[i]).T
        assert (sol[i] - exact).abs().max() < 1e-6
    assert stats["n_accepted"] >= 2 and stats["nfe"] == 2 + 6 * (stats["n_accepted"] + stats["n_rejected"])


def test_dopri5_order_five():
    A = _A()
    f = Lin(A)
    y0 = torch.tensor([[1.0, -0.5, 0.25]], dtype=torch.float64)
    from oracle.torchdiffeq_restated import _runge_kutta_step
    errs = []
    for h in (0.4, 0.2):
        y1, _, _, _ = _runge_kutta_step(f, y0, f(None, y0), torch.tensor(0.0, dtype=torch.float64),
                                        torch.tensor(h, dtype=torch.float64), torch.tensor(h, dtype=torch.float64))
        errs.append((y1 - y0 @ torch.linalg.matrix_exp(A * h).T).abs().max().item())
    assert 5.5 < math.log2(errs[0] / errs[1]) < 6.5  # local error O(h^6)


def test_reversed_time_and_errors():
    torch.manual_seed(0)
    f = Lin(_A())
    y0 = torch.randn(2, 3, dtype=torch.float64)
    fwd = odeint(f, y0, torch.tensor([0.0, 0.5], dtype=torch.float64), method="rk4")[-1]
    back = odeint(f, fwd, torch.tensor([0.5, 0.0], dtype=torch.float64), method="rk4")[-1]
    assert (back - y0).abs().max() < 5e-3   # one rk4 step of 0.5 each way
    with pytest.raises(ValueError):
        odeint(f, y0, torch.tensor([0.0, 1.0]), method="rk45")
    with pytest.raises(AssertionError):
        odeint(f, y0, torch.tensor([0.0, 1.0, 0.5]), method="euler")


def test_gradcheck_fixed_and_adaptive():
    A = _A()
    f = Lin(A)
    y0 = torch.randn(2, 3, dtype=torch.float64, requires_grad=True)
    t = torch.tensor([0.0, 0.1, 0.3], dtype=torch.float64)
    for m in ("euler", "rk4"):
        assert torch.autograd.gradcheck(lambda y: odeint(f, y, t, method=m), (y0,), atol=1e-7)


def test_dopri5_step_arithmetic_matches_scipy_rk45():
    """An INDEPENDENT implementation of the published Dormand-Prince 5(4) pair (scipy.integrate.RK45; Hairer et al.
    / Shampine's continuous extension) pins what it can of the dopri5 restatement, with the NLBAC control-affine
    network as the right-hand side (float64):
      * the tableau: y1 and f(y1) of one step from the same (y0, f0, h) agree to round-off;
      * the error weights: torchdiffeq's c_error is scipy's E scaled by -2/3 (torchdiffeq subtracts a differently
        normalised 4th-order solution: -1/60 on k7 where Hairer's pair has 1/40) -- all 7 entries proportional;
      * the dense output: the quartic through (y0, y1, y_mid, f0, f1) with torchdiffeq's DPS_C_MID is the same
        polynomial as scipy's (5 conditions fix a quartic), compared at 5 interior points;
      * _select_initial_step (Hairer's rule, order 4 -> exponent 1/5): identical first step size.
    NOT pinned by this (torchdiffeq-specific, restated from its published source): the 2/3 scale itself, the step
    controller's no-shrink-after-accept rule, and not clipping the last step."""
    import numpy as np
    from scipy.integrate import RK45
    from scipy.integrate._ivp.rk import rk_step
    from oracle import nlbac_ref
    from oracle import torchdiffeq_restated as R
    torch.manual_seed(3)
    m = nlbac_ref.AffineNODE(3, 3, 6).double()
    for p in m.parameters():
        p.data.mul_(1.5)
    B = 16
    y0 = torch.randn(B, 5, dtype=torch.float64)

    def fun(t, y):
        with torch.no_grad():
            return m(torch.tensor(t), torch.from_numpy(np.ascontiguousarray(y)).reshape(B, 5)).reshape(-1).numpy()

    t0, h = 0.0, 0.13
    with torch.no_grad():
        f0 = m(torch.tensor(t0), y0)
        y1, f1, err, k = R._runge_kutta_step(m, y0, f0, torch.tensor(t0, dtype=torch.float64),
                                             torch.tensor(h, dtype=torch.float64), torch.tensor(t0 + h, dtype=torch.float64))
    K = np.empty((7, B * 5))
    y1_s, f1_s = rk_step(fun, t0, y0.reshape(-1).numpy(), f0.reshape(-1).numpy(), h, RK45.A, RK45.B, RK45.C, K)
    assert np.abs(y1.reshape(-1).numpy() - y1_s).max() < 1e-13
    assert np.abs(f1.reshape(-1).numpy() - f1_s).max() < 1e-12
    for i in range(7):
        assert np.abs(k[i].reshape(-1).numpy() - K[i]).max() < 1e-12
    ratios = [c / e for c, e in zip(R.DP_C_ERR, RK45.E) if e != 0]
    assert len(ratios) == 6 and all(abs(r + 2.0 / 3.0) < 1e-12 for r in ratios), ratios
    assert R.DP_C_ERR[1] == 0 and RK45.E[1] == 0
    err_s = K.T.dot(RK45.E) * h
    assert np.abs(err.reshape(-1).numpy() + (2.0 / 3.0) * err_s).max() < 1e-15 + 1e-9 * np.abs(err_s).max()
    # dense output
    y_mid = y0 + sum(kj * (c * h) for c, kj in zip(R.DP_C_MID, k))
    coeffs = R._interp_fit(y0, y1, y_mid, f0, f1, torch.tensor(h, dtype=torch.float64))
    Q = K.T.dot(RK45.P)
    for theta in (0.1, 0.3, 0.5, 0.77, 0.95):
        want = y0.reshape(-1).numpy() + h * theta * Q.dot(np.cumprod(np.full(4, theta)) / theta)
        got = R._interp_evaluate(coeffs, torch.tensor(t0, dtype=torch.float64), torch.tensor(t0 + h, dtype=torch.float64),
                                 torch.tensor(t0 + theta * h, dtype=torch.float64)).reshape(-1).numpy()
        assert np.abs(got - want).max() < 1e-9 * max(1.0, np.abs(want).max()), theta
    # initial step
    rtol, atol = 1e-6, 1e-8
    st = {}
    with torch.no_grad():
        R.odeint(m, y0, torch.tensor([0.0, 3.0], dtype=torch.float64), method="dopri5", rtol=rtol, atol=atol, _stats=st)
    s = RK45(fun, 0.0, y0.reshape(-1).numpy(), 3.0, rtol=rtol, atol=atol)
    assert abs(st["step_dt"][0] - s.h_abs) <= 1e-12 * s.h_abs, (st["step_dt"][0], s.h_abs)
