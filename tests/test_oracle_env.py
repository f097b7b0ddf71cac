"""oracle/envs.py (numpy restatement of the reference's UnicycleEnv) against trajectories produced by the reference's
own class (tests/golden/make_golden_env.py -> env_U.npz): goal bonus, hazard violations, plain motion."""
import os
import numpy as np
import pytest

from oracle import envs as oenv

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_U.npz"))
NROLL = 4


def _stack(key):
    return np.stack([G["%s_%d" % (key, i)] for i in range(NROLL)], axis=1)      # [T, R, ...]


def test_oracle_env_matches_reference_rollouts():
    act = _stack("action")
    T = act.shape[0]
    state, last, step = oenv.reset(NROLL)
    np.testing.assert_allclose(oenv.get_obs(state), np.stack([G["obs0_%d" % i] for i in range(NROLL)]), rtol=1e-13)
    want = {k: _stack(k) for k in ("obs", "reward", "constraint", "center", "next_center", "done", "goal_met", "nviol", "cost", "state")}
    for t in range(T):
        obs, r, c, cen, ncen, done, met, nv, cost = oenv.step(state, last, step, act[t])
        # float64 with the same operation order: agreement to a few ulp (numpy's small matmul may fuse differently)
        for name, got in (("obs", obs), ("reward", r), ("constraint", c), ("center", cen), ("next_center", ncen),
                          ("cost", cost), ("state", state)):
            np.testing.assert_allclose(got, want[name][t], rtol=1e-11, atol=1e-12, err_msg="%s at step %d" % (name, t))
        assert (done == want["done"][t]).all() and (met == want["goal_met"][t]).all(), t
        assert (nv == want["nviol"][t]).all(), t
    assert want["goal_met"].any() and want["nviol"].sum() > 100      # the goldens exercise both branches


def test_oracle_env_time_limit():
    state, last, step = oenv.reset(2)
    step[:] = oenv.MAX_STEPS - 2
    a = np.zeros((2, 2))
    assert not oenv.step(state, last, step, a)[5].any()
    assert oenv.step(state, last, step, a)[5].all()               # episode_step == max_episode_steps


GC = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_C.npz"))


def _stack_c(key):
    return np.stack([GC["%s_%d" % (key, i)] for i in range(NROLL)], axis=1)


def test_oracle_cars_env_matches_reference_rollouts():
    state0 = np.stack([GC["state0_%d" % i] for i in range(NROLL)])
    state, t, step = oenv.cars_reset(state0[:, 1] - 3.0)            # the offset the reference drew at reset
    np.testing.assert_allclose(state, state0, rtol=0, atol=1e-15)
    np.testing.assert_allclose(oenv.cars_obs(state), np.stack([GC["obs0_%d" % i] for i in range(NROLL)]), rtol=1e-15)
    act = _stack_c("action")
    want = {k: _stack_c(k) for k in ("obs", "reward", "constraint", "prev", "next", "done", "reached", "nviol", "cost", "state")}
    for i in range(act.shape[0]):
        obs, r, c, prev, nxt, done, reached, nv, cost = oenv.cars_step(state, t, step, act[i])
        for name, got in (("obs", obs), ("reward", r), ("constraint", c), ("prev", prev), ("next", nxt), ("cost", cost), ("state", state)):
            np.testing.assert_allclose(got, want[name][i], rtol=1e-12, atol=1e-12, err_msg="%s at step %d" % (name, i))
        assert (done == want["done"][i]).all() and (reached == want["reached"][i]).all() and (nv == want["nviol"][i]).all(), i
    assert want["reached"].sum() > 20 and want["nviol"].sum() > 100 and want["done"][-1].all()


GP = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_P.npz"))
NROLL_P = 6


def _stack_p(key):
    return np.stack([GP["%s_%d" % (key, i)] for i in range(NROLL_P)], axis=1)


def test_oracle_pvtol_env_matches_reference_rollouts():
    state, step = oenv.pvtol_reset(NROLL_P)
    obs_prev = np.stack([GP["obs0_%d" % i] for i in range(NROLL_P)])
    np.testing.assert_allclose(oenv.pvtol_obs(state), obs_prev, rtol=1e-13)
    act = _stack_p("action")
    want = {k: _stack_p(k) for k in ("obs", "reward", "constraint", "done", "goal_met", "nviol", "cost", "state")}
    for i in range(act.shape[0]):
        obs, r, c, pre, done, met, nv, cost = oenv.pvtol_step(state, step, act[i])
        np.testing.assert_allclose(pre, obs_prev, rtol=1e-11, atol=1e-12)        # lya_pre_term = previous observation
        obs_prev = want["obs"][i]
        for name, got in (("obs", obs), ("reward", r), ("constraint", c), ("cost", cost), ("state", state)):
            np.testing.assert_allclose(got, want[name][i], rtol=1e-11, atol=1e-11, err_msg="%s at step %d" % (name, i))
        assert (done == want["done"][i]).all() and (met == want["goal_met"][i]).all() and (nv == want["nviol"][i]).all(), i
    assert want["goal_met"].any() and (want["nviol"].sum(axis=(0, 1)) > 50).all()   # every violation kind occurs
