"""VecUnicycleEnv (nlbac_env_unicycle_step) against the reference's own UnicycleEnv trajectories
(tests/golden/env_U.npz) and against the numpy oracle on a large random batch."""
import os
import numpy as np
import pytest
import torch

from oracle import envs as oenv

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_U.npz"))
NROLL = 4


def _stack(key):
    return np.stack([G["%s_%d" % (key, i)] for i in range(NROLL)], axis=1)


def test_vec_env_replays_reference_rollouts():
    import nlbac_b200
    env = nlbac_b200.VecUnicycleEnv(NROLL)
    obs0 = env.reset()
    np.testing.assert_allclose(obs0.cpu().numpy(), np.stack([G["obs0_%d" % i] for i in range(NROLL)]), rtol=1e-13)
    act = torch.from_numpy(_stack("action")).cuda()
    want = {k: _stack(k) for k in ("obs", "reward", "constraint", "center", "next_center", "done", "goal_met", "nviol", "cost", "state")}
    for t in range(act.shape[0]):
        obs, r, c, cen, ncen, done, info = env.step(act[t])
        got = {"obs": obs, "reward": r, "constraint": c, "center": cen, "next_center": ncen, "cost": info["safety_cost"],
               "state": env.state}
        for k, v in got.items():     # float64 on both sides; device cos / sin / exp differ from libm by <= 2 ulp
            np.testing.assert_allclose(v.cpu().numpy(), want[k][t], rtol=1e-11, atol=1e-12, err_msg="%s at step %d" % (k, t))
        assert (done.cpu().numpy() == want["done"][t]).all(), t
        assert (info["goal_met"].cpu().numpy() == want["goal_met"][t]).all(), t
        assert (info["num_safety_violation"].cpu().numpy() == want["nviol"][t]).all(), t


def test_vec_env_large_batch_vs_oracle_and_masked_reset():
    import nlbac_b200
    E = 200_003                                     # ragged over the grid
    rng = np.random.RandomState(7)
    env = nlbac_b200.VecUnicycleEnv(E)
    # scatter the states over the arena so that hazards, the goal disc and the time limit are all hit
    st = np.stack([rng.uniform(-3, 3, E), rng.uniform(-3, 3, E), rng.uniform(-np.pi, np.pi, E)], axis=1)
    env.state.copy_(torch.from_numpy(st))
    env.episode_step.copy_(torch.from_numpy(rng.randint(0, 1201, E).astype(np.int32)))
    last = rng.uniform(0, 7, E)
    env.last_goal_dist.copy_(torch.from_numpy(last))
    ostate, olast, ostep = st.copy(), last.copy(), env.episode_step.cpu().numpy().copy()
    for _ in range(3):
        a = rng.uniform([-3.5, -12.0], [3.5, 12.0], size=(E, 2))
        obs, r, c, cen, ncen, done, info = env.step(torch.from_numpy(a).cuda())
        w = oenv.step(ostate, olast, ostep, a)
        inside_edge = np.abs(np.sqrt(((w[4][:, None, :] - oenv.HAZARDS[None]) ** 2).sum(2)) - oenv.HAZ_R).min(1) < 1e-9
        goal_edge = np.abs(w[2] - oenv.GOAL_SIZE) < 1e-9
        ok = ~(inside_edge | goal_edge)             # decisions within 1e-9 of a boundary may legitimately differ
        for got, want in ((obs, w[0]), (c, w[2]), (cen, w[3]), (ncen, w[4]), (env.state, ostate)):
            np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-11, atol=1e-12)
        np.testing.assert_allclose(r.cpu().numpy()[ok], w[1][ok], rtol=1e-10, atol=1e-10)
        assert (done.cpu().numpy()[ok] == w[5][ok]).all() and (info["goal_met"].cpu().numpy()[ok] == w[6][ok]).all()
        assert (info["num_safety_violation"].cpu().numpy()[ok] == w[7][ok]).all()
        np.testing.assert_allclose(info["safety_cost"].cpu().numpy()[ok], w[8][ok], rtol=1e-9, atol=1e-12)
        assert w[5].any() and w[7].sum() > 0 and w[6].any()
    # masked reset: only the finished environments start over
    mask = done
    before = env.state.clone()
    obs = env.reset(mask)
    m = mask.cpu().numpy()
    assert (env.episode_step.cpu().numpy()[m] == 0).all()
    np.testing.assert_array_equal(env.state.cpu().numpy()[~m], before.cpu().numpy()[~m])
    np.testing.assert_array_equal(env.state.cpu().numpy()[m], np.tile([-2.5, -2.5, 0.0], (m.sum(), 1)))
    np.testing.assert_allclose(obs.cpu().numpy(), oenv.get_obs(env.state.cpu().numpy()), rtol=1e-11, atol=1e-12)


def test_vec_env_argument_errors():
    import nlbac_b200
    env = nlbac_b200.VecUnicycleEnv(3)
    with pytest.raises(ValueError):
        env.step(torch.zeros(2, 2, device="cuda"))
    with pytest.raises(RuntimeError):
        nlbac_b200.VecUnicycleEnv(3, device="cpu")


GC = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_C.npz"))


def _stack_c(key):
    return np.stack([GC["%s_%d" % (key, i)] for i in range(NROLL)], axis=1)


def test_vec_cars_env_replays_reference_rollouts():
    import nlbac_b200
    env = nlbac_b200.VecSimulatedCarsEnv(NROLL)
    state0 = np.stack([GC["state0_%d" % i] for i in range(NROLL)])
    obs0 = env.reset(noise=torch.from_numpy(state0[:, 1] - 3.0))
    np.testing.assert_allclose(env.state.cpu().numpy(), state0, rtol=0, atol=1e-15)
    np.testing.assert_allclose(obs0.cpu().numpy(), np.stack([GC["obs0_%d" % i] for i in range(NROLL)]), rtol=1e-15)
    act = torch.from_numpy(_stack_c("action")).cuda()
    want = {k: _stack_c(k) for k in ("obs", "reward", "constraint", "prev", "next", "done", "reached", "nviol", "cost", "state")}
    for i in range(act.shape[0]):
        obs, r, c, prev, nxt, done, info = env.step(act[i].reshape(-1, 1))
        got = {"obs": obs, "reward": r, "constraint": c, "prev": prev, "next": nxt, "cost": info["safety_cost"], "state": env.state}
        for k, v in got.items():
            np.testing.assert_allclose(v.cpu().numpy(), want[k][i], rtol=1e-11, atol=1e-12, err_msg="%s at step %d" % (k, i))
        assert (done.cpu().numpy() == want["done"][i]).all(), i
        assert (info["reached"].cpu().numpy() == want["reached"][i]).all(), i
        assert (info["num_safety_violation"].cpu().numpy() == want["nviol"][i]).all(), i
        assert not info["goal_met"].any()


def test_vec_cars_env_large_batch_vs_oracle_and_masked_reset():
    import nlbac_b200
    E = 150_001
    rng = np.random.RandomState(11)
    env = nlbac_b200.VecSimulatedCarsEnv(E)
    noise = rng.normal(0, 0.5, E)
    env.reset(noise=torch.from_numpy(noise))
    ostate, ot, ostep = oenv.cars_reset(noise)
    # spread the gaps so that braking branches, the reward band and both violations occur
    ostate[:, ::2] += rng.uniform(-6, 6, (E, 5))
    ot += rng.uniform(0, 6, E)
    ostep += rng.randint(0, 300, E).astype(np.int32)
    env.state.copy_(torch.from_numpy(ostate)); env.t.copy_(torch.from_numpy(ot)); env.episode_step.copy_(torch.from_numpy(ostep))
    for _ in range(3):
        a = rng.uniform(-3, 3, E)
        obs, r, c, prev, nxt, done, info = env.step(torch.from_numpy(a).cuda())
        w = oenv.cars_step(ostate, ot, ostep, a)
        for got, want in ((obs, w[0]), (c, w[2]), (prev, w[3]), (nxt, w[4]), (env.state, ostate), (env.t, ot)):
            np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-11, atol=1e-11)
        d34, d45 = ostate[:, 4] - ostate[:, 6], ostate[:, 6] - ostate[:, 8]
        ok = (np.abs(np.abs(d34 - 9.5) - 0.5) > 1e-9) & (np.abs(d34 - 2.5) > 1e-9) & (np.abs(d45 - 2.5) > 1e-9)
        np.testing.assert_allclose(r.cpu().numpy()[ok], w[1][ok], rtol=1e-11, atol=1e-12)
        assert (done.cpu().numpy() == w[5]).all() and (info["reached"].cpu().numpy()[ok] == w[6][ok]).all()
        assert (info["num_safety_violation"].cpu().numpy()[ok] == w[7][ok]).all()
        np.testing.assert_allclose(info["safety_cost"].cpu().numpy()[ok], w[8][ok], rtol=1e-9, atol=1e-11)
        assert w[5].any() and w[6].any() and (w[7] == 2).any()
    before = env.state.clone()
    n2 = rng.normal(0, 0.5, E)
    env.reset(done, noise=torch.from_numpy(n2))
    m = done.cpu().numpy()
    assert (env.episode_step.cpu().numpy()[m] == 0).all() and (env.t.cpu().numpy()[m] == 0).all()
    np.testing.assert_array_equal(env.state.cpu().numpy()[~m], before.cpu().numpy()[~m])
    np.testing.assert_allclose(env.state.cpu().numpy()[m], oenv.cars_reset(n2[m])[0], rtol=0, atol=1e-15)


GP = np.load(os.path.join(os.path.dirname(__file__), "golden", "env_P.npz"))
NROLL_P = 6
P_KINDS = (("obstacles", "safety_cost_obstacles"), ("safety_operator", "safety_cost_operator_val"),
           ("y_max", "safety_cost_y_max_val"), ("y_min", "safety_cost_y_min_val"))


def _stack_p(key):
    return np.stack([GP["%s_%d" % (key, i)] for i in range(NROLL_P)], axis=1)


def test_vec_pvtol_env_replays_reference_rollouts():
    import nlbac_b200
    env = nlbac_b200.VecPvtolEnv(NROLL_P)
    obs_prev = env.reset().cpu().numpy()
    np.testing.assert_allclose(obs_prev, np.stack([GP["obs0_%d" % i] for i in range(NROLL_P)]), rtol=1e-13)
    act = torch.from_numpy(_stack_p("action")).cuda()
    want = {k: _stack_p(k) for k in ("obs", "reward", "constraint", "done", "goal_met", "nviol", "cost", "state")}
    for i in range(act.shape[0]):
        obs, r, c, pre, obs2, done, info = env.step(act[i])
        assert obs2 is obs
        np.testing.assert_allclose(pre.cpu().numpy(), obs_prev, rtol=1e-11, atol=1e-12)
        obs_prev = want["obs"][i]
        cost = torch.stack([info[ck] for _, ck in P_KINDS], dim=1)
        nv = torch.stack([info["num_safety_violation_" + k] for k, _ in P_KINDS], dim=1)
        for k, v in (("obs", obs), ("reward", r), ("constraint", c), ("cost", cost), ("state", env.state)):
            np.testing.assert_allclose(v.cpu().numpy(), want[k][i], rtol=1e-11, atol=1e-11, err_msg="%s at step %d" % (k, i))
        assert (done.cpu().numpy() == want["done"][i]).all() and (info["goal_met"].cpu().numpy() == want["goal_met"][i]).all(), i
        assert (nv.cpu().numpy() == want["nviol"][i]).all(), i


def test_vec_pvtol_env_large_batch_vs_oracle_and_masked_reset():
    import nlbac_b200
    E = 120_007
    rng = np.random.RandomState(5)
    env = nlbac_b200.VecPvtolEnv(E)
    st = np.stack([rng.uniform(-7, 7, E), rng.uniform(-120, 120, E), rng.uniform(-np.pi, np.pi, E), rng.uniform(-200, 200, E),
                   rng.uniform(-3, 3, E), rng.uniform(0, 3, E), rng.uniform(-7, 7, E)], axis=1)
    st[: E // 2, 1] = rng.uniform(-6, 6, E // 2)          # half of them inside the arena (hazards, goal disc)
    st[: E // 2, 3] = rng.uniform(-3, 3, E // 2)
    st[:, 6] = st[:, 0] + rng.uniform(-0.5, 0.5, E)
    ostep = rng.randint(0, 2001, E).astype(np.int32)
    env.state.copy_(torch.from_numpy(st)); env.episode_step.copy_(torch.from_numpy(ostep))
    ostate = st.copy()
    for _ in range(3):
        a = rng.uniform([-3.5, -15.0], [3.5, 15.0], size=(E, 2))
        obs, r, c, pre, _, done, info = env.step(torch.from_numpy(a).cuda())
        w = oenv.pvtol_step(ostate, ostep, a)
        for got, want in ((obs, w[0]), (c, w[2]), (pre, w[3]), (env.state, ostate)):
            np.testing.assert_allclose(got.cpu().numpy(), want, rtol=1e-10, atol=1e-10)
        d = np.sqrt(((ostate[:, None, :2] - oenv.P_HAZARDS[None]) ** 2).sum(2))
        dop = np.abs(ostate[:, 0] - ostate[:, 6])
        ok = (np.abs(d - 0.25).min(1) > 1e-9) & (np.abs(w[2] - 3.5) > 1e-9) & (np.abs(dop - 1.0) > 1e-9) & \
             (np.abs(np.abs(ostate[:, 1]) - 100.0) > 1e-9)
        cost = torch.stack([info[ck] for _, ck in P_KINDS], dim=1).cpu().numpy()
        nv = torch.stack([info["num_safety_violation_" + k] for k, _ in P_KINDS], dim=1).cpu().numpy()
        np.testing.assert_allclose(r.cpu().numpy()[ok], w[1][ok], rtol=1e-10, atol=1e-10)
        assert (done.cpu().numpy()[ok] == w[4][ok]).all() and (info["goal_met"].cpu().numpy()[ok] == w[5][ok]).all()
        assert (nv[ok] == w[6][ok]).all()
        np.testing.assert_allclose(cost[ok], w[7][ok], rtol=1e-9, atol=1e-9)
        assert w[5].any() and (w[6].sum(0) > 0).all() and w[4].any()
    before = env.state.clone()
    obs = env.reset(done)
    m = done.cpu().numpy()
    assert (env.episode_step.cpu().numpy()[m] == 0).all()
    np.testing.assert_array_equal(env.state.cpu().numpy()[~m], before.cpu().numpy()[~m])
    np.testing.assert_array_equal(env.state.cpu().numpy()[m], np.tile([-4.5, -4.5, 0, 0, 0, 1.0, -4.5], (m.sum(), 1)))
    np.testing.assert_allclose(obs.cpu().numpy(), oenv.pvtol_obs(env.state.cpu().numpy()), rtol=1e-10, atol=1e-12)
