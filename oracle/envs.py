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


# ------------------------------------------------------------------------------------------------
# SimulatedCarsEnv (C envs/simulated_cars_env.py: step :42-146, _get_reward :148-155, reset :158-177, _get_obs :197-213)
# ------------------------------------------------------------------------------------------------
def cars_obs(state):
    o = state.copy()
    o[:, ::2] /= 100.0
    o[:, 1::2] /= 30.0
    return o


def cars_reset(noise):
    E = noise.shape[0]
    state = np.zeros((E, 10))
    state[:, ::2] = [42.0, 34.0, 26.0, 18.0, 10.0]
    state[:, 1::2] = (3.0 + noise)[:, None]
    state[:, 7] = 3.0
    return state, np.zeros(E), np.zeros(E, dtype=np.int32)


def cars_step(state, t, episode_step, action):
    """action [E]; in-place; returns (obs, reward, constraint, prev, next, done, reached, nviol, cost)."""
    pos, vels = state[:, ::2].copy(), state[:, 1::2].copy()
    vd = 3.0 * np.ones_like(vels)
    vd[:, 0] -= 4 * np.sin(t)
    acc = 4.0 * (vd - vels)
    d01, d12, d24 = pos[:, 0] - pos[:, 1], pos[:, 1] - pos[:, 2], pos[:, 2] - pos[:, 4]
    acc[:, 1] += -20.0 * d01 * (d01 < 6.5)
    acc[:, 2] += -20.0 * d12 * (d12 < 6.5)
    acc[:, 3] = 0.0
    acc[:, 4] += -20.0 * d24 * (d24 < 13.0)
    acc *= 1.1
    prev = state[:, 4:8].copy()
    f = np.zeros_like(state)
    f[:, ::2] = vels
    f[:, 1::2] = acc
    f[:, 7] = 0.0
    g = np.zeros_like(state)
    g[:, 7] = 1.0
    state += 0.02 * (f + g * action[:, None])
    t += 0.02
    episode_step += 1
    nxt = state[:, 4:8].copy()
    d34, d45 = state[:, 4] - state[:, 6], state[:, 6] - state[:, 8]
    reward = -0.5 * np.abs(action ** 2) / 300
    reached = np.abs(d34 - 9.5) < 0.5
    reward = reward + np.where(reached, 2.0, 0.0)
    done = episode_step >= 300
    v1, v2 = d34 < 2.5, d45 < 2.5
    nviol = v1.astype(np.int32) + v2.astype(np.int32)
    cost = np.where(v1, np.abs(d34 - 2.5), 0.0) + np.where(v2, np.abs(d45 - 2.5), 0.0)
    return cars_obs(state), reward, np.abs(d34 - 9.5), prev, nxt, done, reached, nviol, cost


# ------------------------------------------------------------------------------------------------
# PvtolEnv (P envs/pvtol_env.py: _step :84-216, goal_met :218-224, reset :226-246, _get_dynamics :248-277,
# get_obs :361-377, obs_compass :379-405)
# ------------------------------------------------------------------------------------------------
P_GOAL = np.array([4.5, 4.5])
P_HAZARDS = np.array([[-2.5, -2.5], [-2.5, 2.5], [0.0, -3.5], [0.0, 3.5], [-4.5, 0.0]])


def pvtol_obs(state):
    c, s = np.cos(state[:, 2]), np.sin(state[:, 2])
    rel = P_GOAL - state[:, :2]
    dist = np.sqrt(np.sum(rel * rel, axis=1))
    v0 = rel[:, 0] * c + rel[:, 1] * s
    v1 = rel[:, 0] * (-s) + rel[:, 1] * c
    n = np.sqrt(v0 * v0 + v1 * v1) + 0.001
    return np.stack([state[:, 0], state[:, 1], c, s, state[:, 3], state[:, 4], state[:, 5], state[:, 6], v0 / n, v1 / n,
                     np.exp(-dist)], axis=1)


def pvtol_reset(E):
    return np.tile(np.array([-4.5, -4.5, 0.0, 0.0, 0.0, 1.0, -4.5]), (E, 1)), np.zeros(E, dtype=np.int32)


def pvtol_step(state, episode_step, action):
    """returns (obs, reward, constraint, lya_pre_term, done, goal_met, nviol [E,4], cost [E,4])."""
    pre = pvtol_obs(state)
    th, thrust = state[:, 2].copy(), state[:, 5].copy()
    f = np.zeros((state.shape[0], 6))
    f[:, 0], f[:, 1] = state[:, 3], state[:, 4]
    f[:, 3] = -np.sin(th) * thrust
    f[:, 4] = np.cos(th) * thrust - 1.0
    gu = np.zeros_like(f)
    gu[:, 2], gu[:, 5] = action[:, 1], action[:, 0]
    state[:, :6] += 0.02 * (f + gu)
    state[:, 6] = state[:, 6] + 0.7 * (state[:, 0] - state[:, 6])
    episode_step += 1
    rel = P_GOAL - state[:, :2]
    dist = np.sqrt(np.sum(rel * rel, axis=1))
    met = dist <= 3.5
    reward = -1e-3 * dist + np.where(met, 1500.0, 0.0)
    done = met | (episode_step >= 2000)
    E = state.shape[0]
    nviol, cost = np.zeros((E, 4), dtype=np.int32), np.zeros((E, 4))
    d2 = np.sum((state[:, None, :2] - P_HAZARDS[None]) ** 2, axis=2)
    inside = d2 < 0.25 ** 2
    nviol[:, 0] = inside.sum(1)
    cost[:, 0] = np.where(inside, (0.25 - np.sqrt(d2)) / 0.25, 0.0).sum(1)
    dop = state[:, 0] - state[:, 6]
    far = np.abs(dop) >= 1.0
    nviol[:, 1] = far
    cost[:, 1] = np.where(far, np.where(dop <= -1.0, -dop - 1.0, dop - 1.0), 0.0)
    hi, lo = state[:, 1] > 100.0, state[:, 1] < -100.0
    nviol[:, 2], nviol[:, 3] = hi, lo
    cost[:, 2] = np.where(hi, state[:, 1] - 100.0, 0.0)
    cost[:, 3] = np.where(lo, -100.0 - state[:, 1], 0.0)
    return pvtol_obs(state), reward, dist, pre, done, met, nviol, cost
