"""Golden trajectories of the reference's OWN UnicycleEnv (U envs/unicycle_env.py), run in this container from
/root/reference with a stand-in `gym` module (gym is not installed; the class only needs gym.Env and spaces.Box as
containers).  Writes tests/golden/env_U.npz: for R rollouts of T steps, the actions fed and everything step() returned.
Rollouts are steered so that hazard violations, the goal bonus and plain motion all occur."""
import os, sys, types
import numpy as np

REF = "/root/reference/NLBAC_Unicycle_RL_training/Unicycle_RL_training"
gym = types.ModuleType("gym")
class _Env: pass
class _Box:
    def __init__(self, low=None, high=None, shape=None):
        self.low, self.high, self.shape = low, high, shape
spaces = types.ModuleType("gym.spaces"); spaces.Box = _Box
gym.Env, gym.spaces = _Env, spaces
sys.modules["gym"], sys.modules["gym.spaces"] = gym, spaces
sys.path.insert(0, REF)
from envs.unicycle_env import UnicycleEnv   # noqa: E402

def rollout(seed, T, mode):
    env = UnicycleEnv(seed)
    rng = np.random.RandomState(1000 + seed)
    obs0 = env.reset()
    rec = {k: [] for k in ("action", "obs", "reward", "constraint", "center", "next_center", "done", "goal_met", "nviol", "cost", "state")}
    for t in range(T):
        if mode == "goal":      # head for the goal with a proportional steering law + noise
            rel = env.goal_pos - env.state[:2]
            err = np.arctan2(rel[1], rel[0]) - env.state[2]
            err = (err + np.pi) % (2 * np.pi) - np.pi
            a = np.array([3.0 + 0.3 * rng.randn(), 4.0 * err + 0.5 * rng.randn()])
        elif mode == "hazard":  # drive straight at the hazard at (-1.5, -1.5), slowly
            rel = np.array([-1.5, -1.5]) - env.state[:2]
            err = np.arctan2(rel[1], rel[0]) - env.state[2]
            err = (err + np.pi) % (2 * np.pi) - np.pi
            a = np.array([1.0 + 0.1 * rng.randn(), 6.0 * err])
        else:
            a = rng.uniform([-3.5, -12.0], [3.5, 12.0])
        obs, r, c, cen, ncen, done, info = env.step(a)
        rec["action"].append(a); rec["obs"].append(obs); rec["reward"].append(r); rec["constraint"].append(c)
        rec["center"].append(np.array(cen)); rec["next_center"].append(np.array(ncen)); rec["done"].append(bool(done))
        rec["goal_met"].append(bool(info.get("goal_met", False))); rec["nviol"].append(int(info.get("num_safety_violation", 0)))
        rec["cost"].append(float(info.get("safety_cost", 0.0))); rec["state"].append(env.state.copy())
    return obs0, {k: np.array(v) for k, v in rec.items()}

if __name__ == "__main__":
    out = {}
    for i, (seed, mode) in enumerate([(0, "goal"), (1, "hazard"), (2, "random"), (3, "random")]):
        obs0, rec = rollout(seed, 260, mode)
        out["obs0_%d" % i] = obs0
        for k, v in rec.items():
            out["%s_%d" % (k, i)] = v
        print(mode, "goal_met", rec["goal_met"].any(), "violations", int(rec["nviol"].sum()), "done at", int(np.argmax(rec["done"])) if rec["done"].any() else None)
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "env_U.npz"), **out)


def rollout_cars(seed, T, mode):
    REFC = "/root/reference/NLBAC_SimulatedCarsFollowing_RL_training/Simulated_Car_Following_RL_training"
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_cars_env", os.path.join(REFC, "envs", "simulated_cars_env.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    env = mod.SimulatedCarsEnv(seed)
    rng = np.random.RandomState(2000 + seed)
    obs0 = env.reset()
    state0 = env.state.copy()
    keys = ("action", "obs", "reward", "constraint", "prev", "next", "done", "reached", "nviol", "cost", "state")
    rec = {k: [] for k in keys}
    for t in range(T):
        if mode == "brake":     a = np.array([-3.0 + 0.2 * rng.randn()])      # falls back onto car 5
        elif mode == "rush":    a = np.array([3.0 - 0.2 * rng.rand()])         # closes in on car 3
        elif mode == "follow":  a = np.array([np.clip(2.0 * ((env.state[4] - env.state[6]) - 9.5) + 1.5 * (env.state[5] - env.state[7]), -3, 3)])
        else:                   a = rng.uniform(-3.0, 3.0, size=1)
        obs, r, c, prev, nxt, done, info = env.step(a)
        rec["action"].append(a[0]); rec["obs"].append(obs); rec["reward"].append(float(r)); rec["constraint"].append(float(c))
        rec["prev"].append(np.array(prev)); rec["next"].append(np.array(nxt)); rec["done"].append(bool(done))
        rec["reached"].append(int(info["reached"])); rec["nviol"].append(int(info["num_safety_violation"]))
        rec["cost"].append(float(info["safety_cost"])); rec["state"].append(env.state.copy())
    return obs0, state0, {k: np.array(v) for k, v in rec.items()}


if __name__ == "__main__":
    out = {}
    for i, (seed, mode) in enumerate([(0, "brake"), (1, "rush"), (2, "follow"), (3, "random")]):
        obs0, state0, rec = rollout_cars(seed, 320, mode)
        out["obs0_%d" % i], out["state0_%d" % i] = obs0, state0
        for k, v in rec.items():
            out["%s_%d" % (k, i)] = v
        print("cars", mode, "reached", int(rec["reached"].sum()), "violations", int(rec["nviol"].sum()), "done at", int(np.argmax(rec["done"])))
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "env_C.npz"), **out)


def rollout_pvtol(seed, T, mode):
    REFP = "/root/reference/NLBAC_pvtol_RL_training/Pvtol_RL_training"
    import importlib.util
    u = types.ModuleType("envs.utils"); u.to_pixel = lambda m, shift=0: 1.5 * 37.795 * m + shift
    pk = types.ModuleType("Pvtol_RL_training"); pe = types.ModuleType("Pvtol_RL_training.envs"); pu = types.ModuleType("Pvtol_RL_training.envs.utils")
    pu.to_pixel = u.to_pixel
    for name, m in (("Pvtol_RL_training", pk), ("Pvtol_RL_training.envs", pe), ("Pvtol_RL_training.envs.utils", pu)):
        sys.modules.setdefault(name, m)
    spec = importlib.util.spec_from_file_location("ref_pvtol_env", os.path.join(REFP, "envs", "pvtol_env.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    env = mod.PvtolEnv(seed)
    rng = np.random.RandomState(3000 + seed)
    obs0 = env.reset()
    keys = ("action", "obs", "reward", "constraint", "pre", "done", "goal_met", "nviol", "cost", "state")
    rec = {k: [] for k in keys}
    kinds = (("obstacles", "safety_cost_obstacles"), ("safety_operator", "safety_cost_operator_val"),
             ("y_max", "safety_cost_y_max_val"), ("y_min", "safety_cost_y_min_val"))
    for t in range(T):
        s = env.state
        if mode == "goal":       # PD attitude + thrust law flying up and right towards the goal disc
            th_des = np.clip(-0.25 * (4.5 - s[0]) + 0.35 * s[3], -0.6, 0.6)
            a = np.array([np.clip(4.0 * (1.25 + 0.4 * (4.5 - s[1]) - 0.8 * s[4] - s[5]), -3.5, 3.5), np.clip(8.0 * (th_des - s[2]), -15, 15)])
        elif mode == "hazard":   # hover-climb straight up the x = -4.5 line into the hazard at (-4.5, 0)
            a = np.array([np.clip(4.0 * (1.0 + 0.3 * (0.0 - s[1]) - 0.9 * s[4] - s[5]), -3.5, 3.5), np.clip(-8.0 * s[2], -15, 15)])
        elif mode == "sideways": # tilt to -90 degrees, then full thrust rate: outruns the safety operator (|x - x_op| >= 1)
            a = np.array([3.5, -15.0 if s[2] > -1.5 else 0.0])
        elif mode == "climb":    # upright, full thrust rate: y > y_max
            a = np.array([3.5, 0.0])
        elif mode == "fall":     # thrust goes negative: y < y_min
            a = np.array([-3.5, 0.0])
        else:
            a = rng.uniform([-3.5, -15.0], [3.5, 15.0])
        obs, r, c, pre, obs2, done, info = env.step(a)
        assert np.array_equal(obs, obs2)
        rec["action"].append(a); rec["obs"].append(obs); rec["reward"].append(float(r)); rec["constraint"].append(float(c))
        rec["pre"].append(np.array(pre)); rec["done"].append(bool(done)); rec["goal_met"].append(bool(info.get("goal_met", False)))
        rec["nviol"].append([int(info.get("num_safety_violation_" + k, 0)) for k, _ in kinds])
        rec["cost"].append([float(info.get(ck, 0.0)) for _, ck in kinds]); rec["state"].append(env.state.copy())
    return obs0, {k: np.array(v) for k, v in rec.items()}


if __name__ == "__main__":
    out = {}
    for i, (seed, mode) in enumerate([(0, "goal"), (1, "hazard"), (2, "sideways"), (3, "climb"), (4, "fall"), (5, "random")]):
        obs0, rec = rollout_pvtol(seed, 520, mode)
        assert np.array_equal(rec["pre"][1:], rec["obs"][:-1]) and np.array_equal(rec["pre"][0], obs0)
        del rec["pre"]                     # lya_pre_term is the previous observation: checked here, not stored
        out["obs0_%d" % i] = obs0
        for k, v in rec.items():
            out["%s_%d" % (k, i)] = v
        print("pvtol", mode, "goal_met", bool(rec["goal_met"].any()), "violations", rec["nviol"].sum(0).tolist(), "final", np.round(rec["state"][-1], 2).tolist())
    np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "env_P.npz"), **out)
