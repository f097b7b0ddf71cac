"""CPU restatement of the reference's UnicycleEnv for E environments at once (TEST INFRASTRUCTURE, see
oracle/__init__.py): numpy float64, one environment per row.  Follows U envs/unicycle_env.py line by line:
_step :92-155, goal_met :157-164, reset :166-184, get_obs :263-279, _get_dynamics :281-300, obs_compass :302-318,
_goal_dist_center :320-321.  Pinned by tests/golden/env_U.npz, which the reference's own class produced
(tests/golden/make_golden_env.py)."""
import numpy as np

DT, L_P, DES_V = 0.02, 0.03, 2.5
GOAL = np.array([2.5, 2.5])
GOAL_SIZE, REWARD_GOAL, MAX_STEPS = 0.3, 500, 1200
HAZ_R, COST_COEF = 0.5, 1.0
HAZARDS = np.array([[0., 0.], [0., 1.], [0., -1.], [-1., 1.], [-1., -1.], [1., -1.], [1., 1.]]) * 1.5


def get_obs(state):
    c, s = np.cos(state[:, 2]), np.sin(state[:, 2])
    rel = GOAL - state[:, :2]
    dist = np.sqrt(np.sum(rel * rel, axis=1))
    v0 = rel[:, 0] * c + rel[:, 1] * s          # vec @ R, R = [[c, -s], [s, c]]
    v1 = rel[:, 0] * (-s) + rel[:, 1] * c
    n = np.sqrt(v0 * v0 + v1 * v1) + 0.001
    return np.stack([state[:, 0], state[:, 1], c, s, v0 / n, v1 / n, np.exp(-dist)], axis=1)


def reset(E):
    state = np.tile(np.array([-2.5, -2.5, 0.0]), (E, 1))
    nc = np.array([-2.47, -2.5])
    last = np.full(E, np.linalg.norm(GOAL - nc))
    return state, last, np.zeros(E, dtype=np.int32)


def step(state, last_goal_dist, episode_step, action):
    """In-place like the reference; returns (obs, reward, constraint, center, next_center, done, goal_met, nviol, cost)."""
    c, s = np.cos(state[:, 2]), np.sin(state[:, 2])
    center = np.stack([state[:, 0] + L_P * c, state[:, 1] + L_P * s], axis=1)
    a0, a1 = action[:, 0], action[:, 1]
    state[:, 0] += DT * (c * a0)
    state[:, 1] += DT * (s * a0)
    state[:, 2] += DT * a1
    c, s = np.cos(state[:, 2]), np.sin(state[:, 2])
    k = DT * 0.1
    state[:, 0] -= (k * c) * c
    state[:, 1] -= (k * s) * c
    nc = np.stack([state[:, 0] + L_P * c, state[:, 1] + L_P * s], axis=1)
    episode_step += 1
    d = GOAL - nc
    dist = np.sqrt(np.sum(d * d, axis=1))
    reward = -np.square(a0 - DES_V) * 0.1 + (last_goal_dist - dist) * 30
    last_goal_dist[:] = dist
    met = dist <= GOAL_SIZE
    reward = reward + np.where(met, REWARD_GOAL, 0.0)
    done = met | (episode_step >= MAX_STEPS)
    d2 = np.sum((nc[:, None, :] - HAZARDS[None]) ** 2, axis=2)         # [E, 7]
    inside = d2 < HAZ_R ** 2
    nviol = inside.sum(axis=1).astype(np.int32)
    cost = np.where(inside, (HAZ_R - np.sqrt(d2)) / HAZ_R * COST_COEF, 0.0).sum(axis=1)
    return get_obs(state), reward, dist.copy(), center, nc, done, met, nviol, cost
