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
