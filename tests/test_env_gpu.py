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
