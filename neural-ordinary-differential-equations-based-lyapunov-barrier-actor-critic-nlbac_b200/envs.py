"""Vectorised environments on the device.

``VecUnicycleEnv(num_envs)`` is E independent copies of the reference's ``UnicycleEnv`` (U envs/unicycle_env.py)
advanced by ONE kernel launch per step (``nlbac_env_unicycle_step``): same constants, same float64 arithmetic, same
return tuple as ``UnicycleEnv.step`` -- ``(obs, reward, constraint, center_pos, next_center_pos, done, info)`` -- with
a leading environment axis on every element.  ``info`` carries the reference's optional dict keys as dense tensors
(``num_safety_violation`` / ``safety_cost`` are 0 where the reference would omit the key, ``goal_met`` is a bool
tensor).  ``reset(mask)`` re-initialises the selected environments (all when ``mask`` is None), which is what the
reference's training loop does at the top of each episode (U main.py:52).  There is no CPU path.
"""
import numpy as np
import torch

from . import _lib


def _p(t):
    return 0 if t is None else t.data_ptr()


class VecUnicycleEnv:
    dynamics_mode = "Unicycle"

    def __init__(self, num_envs, device="cuda"):
        self.num_envs = int(num_envs)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("VecUnicycleEnv runs on a CUDA device only (no CPU fallback)")
        # the reference's constants (U envs/unicycle_env.py:21-47), kept as attributes under the same names
        self.dt = 0.02
        self.max_episode_steps = 1200
        self.reward_goal = 500
        self.goal_size = 0.3
        self.goal_pos = np.array([2.5, 2.5])
        self.hazards_radius = 0.5
        self.hazards_locations = np.array([[0., 0.], [0., 1.], [0., -1.], [-1., 1.], [-1., -1.], [1., -1.], [1., 1.]]) * 1.5
        self.bds = np.array([[-3., -3.], [3., 3.]])
        self.action_low = np.array([-3.5, -12.0])
        self.action_high = np.array([3.5, 12.0])
        E, dev, f64 = self.num_envs, self.device, torch.float64
        self.state = torch.empty(E, 3, dtype=f64, device=dev)
        self.last_goal_dist = torch.empty(E, dtype=f64, device=dev)
        self.episode_step = torch.empty(E, dtype=torch.int32, device=dev)
        self.reset()

    def reset(self, mask=None):
        """Re-initialise the environments selected by ``mask`` (bool [E]; None = all); returns obs [E, 7]."""
        E = self.num_envs
        obs = torch.empty(E, 7, dtype=torch.float64, device=self.device)
        m = None
        if mask is not None:
            m = mask.to(device=self.device, dtype=torch.uint8).contiguous()
            if m.shape != (E,):
                raise ValueError("mask must have shape [num_envs]")
            obs.copy_(self.get_obs())
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().nlbac_env_unicycle_reset(self.state.data_ptr(), self.last_goal_dist.data_ptr(),
                                                          self.episode_step.data_ptr(), _p(m), E, obs.data_ptr(),
                                                          _lib.raw_stream()), "env_unicycle_reset")
        return obs

    def get_obs(self):
        """[x, y, cos th, sin th, compass_x, compass_y, exp(-dist)] of the current states (U :263-279)."""
        s = self.state
        goal = torch.as_tensor(self.goal_pos, dtype=s.dtype, device=s.device)
        c, sn = torch.cos(s[:, 2]), torch.sin(s[:, 2])
        rel = goal - s[:, :2]
        v0 = rel[:, 0] * c + rel[:, 1] * sn
        v1 = rel[:, 0] * (-sn) + rel[:, 1] * c
        n = torch.sqrt(v0 * v0 + v1 * v1) + 0.001
        return torch.stack([s[:, 0], s[:, 1], c, sn, v0 / n, v1 / n, torch.exp(-rel.norm(dim=1))], dim=1)

    def step(self, action):
        """One ``UnicycleEnv.step`` for every environment; ``action`` [E, 2] (any float dtype, on the device)."""
        E, dev, f64 = self.num_envs, self.device, torch.float64
        a = torch.as_tensor(action, device=dev).to(f64).contiguous()
        if a.shape != (E, 2):
            raise ValueError("action must have shape [num_envs, 2], got %s" % (tuple(a.shape),))
        obs = torch.empty(E, 7, dtype=f64, device=dev)
        reward, constraint = torch.empty(E, dtype=f64, device=dev), torch.empty(E, dtype=f64, device=dev)
        center, next_center = torch.empty(E, 2, dtype=f64, device=dev), torch.empty(E, 2, dtype=f64, device=dev)
        cost = torch.empty(E, dtype=f64, device=dev)
        done = torch.empty(E, dtype=torch.uint8, device=dev)
        met = torch.empty(E, dtype=torch.uint8, device=dev)
        nviol = torch.empty(E, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().nlbac_env_unicycle_step(
                self.state.data_ptr(), self.last_goal_dist.data_ptr(), self.episode_step.data_ptr(), a.data_ptr(), E,
                obs.data_ptr(), reward.data_ptr(), constraint.data_ptr(), center.data_ptr(), next_center.data_ptr(),
                done.data_ptr(), met.data_ptr(), nviol.data_ptr(), cost.data_ptr(), _lib.raw_stream()),
                "env_unicycle_step")
        info = {"goal_met": met.bool(), "num_safety_violation": nviol, "safety_cost": cost}
        return obs, reward, constraint, center, next_center, done.bool(), info


class VecSimulatedCarsEnv:
    """E copies of the reference's ``SimulatedCarsEnv`` (C envs/simulated_cars_env.py); ``step(action[E, 1] or [E])``
    returns ``(obs, reward, constraint, previous_positions, next_positions, done, info)`` like the reference, with
    ``info = {reached, goal_met (all False), num_safety_violation, safety_cost}`` as tensors.  ``reset(mask, noise)``:
    ``noise`` [E] is the per-environment initial-velocity offset (the reference draws ``np.random.normal(0, 0.5)``
    once per reset); sampled on the device when omitted."""
    dynamics_mode = "SimulatedCars"

    def __init__(self, num_envs, device="cuda"):
        self.num_envs = int(num_envs)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("VecSimulatedCarsEnv runs on a CUDA device only (no CPU fallback)")
        self.max_episode_steps, self.dt = 300, 0.02
        self.kp, self.k_brake = 4.0, 20.0
        self.should_keep, self.should_keep_thre, self.reward_goal = 9.5, 0.5, 2.0
        E, dev = self.num_envs, self.device
        self.state = torch.empty(E, 10, dtype=torch.float64, device=dev)
        self.t = torch.empty(E, dtype=torch.float64, device=dev)
        self.episode_step = torch.empty(E, dtype=torch.int32, device=dev)
        self.reset()

    def _get_obs(self):
        o = self.state.clone()
        o[:, ::2] /= 100.0
        o[:, 1::2] /= 30.0
        return o

    def reset(self, mask=None, noise=None):
        E, dev = self.num_envs, self.device
        if noise is None:
            noise = torch.randn(E, dtype=torch.float64, device=dev) * 0.5
        noise = torch.as_tensor(noise, device=dev).to(torch.float64).contiguous()
        if noise.shape != (E,):
            raise ValueError("noise must have shape [num_envs]")
        obs = torch.empty(E, 10, dtype=torch.float64, device=dev)
        m = None
        if mask is not None:
            m = mask.to(device=dev, dtype=torch.uint8).contiguous()
            if m.shape != (E,):
                raise ValueError("mask must have shape [num_envs]")
            obs.copy_(self._get_obs())
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().nlbac_env_cars_reset(self.state.data_ptr(), self.t.data_ptr(), self.episode_step.data_ptr(),
                                                      noise.data_ptr(), _p(m), E, obs.data_ptr(), _lib.raw_stream()),
                       "env_cars_reset")
        return obs

    def step(self, action):
        E, dev, f64 = self.num_envs, self.device, torch.float64
        a = torch.as_tensor(action, device=dev).to(f64).reshape(-1).contiguous()
        if a.shape != (E,):
            raise ValueError("action must have shape [num_envs, 1] or [num_envs]")
        obs = torch.empty(E, 10, dtype=f64, device=dev)
        reward, constraint, cost = (torch.empty(E, dtype=f64, device=dev) for _ in range(3))
        prev, nxt = torch.empty(E, 4, dtype=f64, device=dev), torch.empty(E, 4, dtype=f64, device=dev)
        done, reached = torch.empty(E, dtype=torch.uint8, device=dev), torch.empty(E, dtype=torch.uint8, device=dev)
        nviol = torch.empty(E, dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().nlbac_env_cars_step(
                self.state.data_ptr(), self.t.data_ptr(), self.episode_step.data_ptr(), a.data_ptr(), E, obs.data_ptr(),
                reward.data_ptr(), constraint.data_ptr(), prev.data_ptr(), nxt.data_ptr(), done.data_ptr(),
                reached.data_ptr(), nviol.data_ptr(), cost.data_ptr(), _lib.raw_stream()), "env_cars_step")
        info = {"reached": reached.to(torch.int64), "goal_met": torch.zeros(E, dtype=torch.bool, device=dev),
                "num_safety_violation": nviol, "safety_cost": cost}
        return obs, reward, constraint, prev, nxt, done.bool(), info


class VecPvtolEnv:
    """E copies of the reference's ``PvtolEnv`` (P envs/pvtol_env.py); ``step(action[E, 2])`` returns the reference's
    tuple ``(obs, reward, constraint, lya_pre_term, obs, done, info)``; ``info`` holds ``goal_met`` and, per violation
    kind, ``num_safety_violation_{obstacles, safety_operator, y_max, y_min}`` and ``safety_cost_obstacles`` /
    ``safety_cost_operator_val`` / ``safety_cost_y_max_val`` / ``safety_cost_y_min_val`` (0 where the reference omits
    the key)."""
    dynamics_mode = "Pvtol"
    _KINDS = (("obstacles", "safety_cost_obstacles"), ("safety_operator", "safety_cost_operator_val"),
              ("y_max", "safety_cost_y_max_val"), ("y_min", "safety_cost_y_min_val"))

    def __init__(self, num_envs, device="cuda"):
        self.num_envs = int(num_envs)
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("VecPvtolEnv runs on a CUDA device only (no CPU fallback)")
        self.dt, self.max_episode_steps, self.reward_goal, self.goal_size = 0.02, 2000, 1500.0, 3.5
        self.goal_pos = np.array([4.5, 4.5])
        self.hazard_locations = np.array([[-2.5, -2.5], [-2.5, 2.5], [0.0, -3.5], [0.0, 3.5], [-4.5, 0.0]])
        self.hazards_radius, self.operator_dist, self.safety_operator_follow = 0.25, 1.0, 0.7
        self.y_min, self.y_max = -100.0, 100.0
        E, dev = self.num_envs, self.device
        self.state = torch.empty(E, 7, dtype=torch.float64, device=dev)
        self.episode_step = torch.empty(E, dtype=torch.int32, device=dev)
        self.reset()

    def reset(self, mask=None):
        E, dev = self.num_envs, self.device
        obs = torch.empty(E, 11, dtype=torch.float64, device=dev)
        m = None
        if mask is not None:
            m = mask.to(device=dev, dtype=torch.uint8).contiguous()
            if m.shape != (E,):
                raise ValueError("mask must have shape [num_envs]")
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().nlbac_env_pvtol_reset(self.state.data_ptr(), self.episode_step.data_ptr(), _p(m), E,
                                                       obs.data_ptr() if m is None else 0, _lib.raw_stream()),
                       "env_pvtol_reset")
        return obs if m is None else self.get_obs()

    def get_obs(self):
        s = self.state
        goal = torch.as_tensor(self.goal_pos, dtype=s.dtype, device=s.device)
        c, sn = torch.cos(s[:, 2]), torch.sin(s[:, 2])
        rel = goal - s[:, :2]
        v0 = rel[:, 0] * c + rel[:, 1] * sn
        v1 = rel[:, 0] * (-sn) + rel[:, 1] * c
        n = torch.sqrt(v0 * v0 + v1 * v1) + 0.001
        return torch.stack([s[:, 0], s[:, 1], c, sn, s[:, 3], s[:, 4], s[:, 5], s[:, 6], v0 / n, v1 / n,
                            torch.exp(-rel.norm(dim=1))], dim=1)

    def step(self, action):
        E, dev, f64 = self.num_envs, self.device, torch.float64
        a = torch.as_tensor(action, device=dev).to(f64).contiguous()
        if a.shape != (E, 2):
            raise ValueError("action must have shape [num_envs, 2], got %s" % (tuple(a.shape),))
        obs, pre = torch.empty(E, 11, dtype=f64, device=dev), torch.empty(E, 11, dtype=f64, device=dev)
        reward, constraint = torch.empty(E, dtype=f64, device=dev), torch.empty(E, dtype=f64, device=dev)
        done, met = torch.empty(E, dtype=torch.uint8, device=dev), torch.empty(E, dtype=torch.uint8, device=dev)
        nviol = torch.empty(E, 4, dtype=torch.int32, device=dev)
        cost = torch.empty(E, 4, dtype=f64, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().nlbac_env_pvtol_step(
                self.state.data_ptr(), self.episode_step.data_ptr(), a.data_ptr(), E, obs.data_ptr(), reward.data_ptr(),
                constraint.data_ptr(), pre.data_ptr(), done.data_ptr(), met.data_ptr(), nviol.data_ptr(), cost.data_ptr(),
                _lib.raw_stream()), "env_pvtol_step")
        info = {"goal_met": met.bool()}
        for i, (kind, cost_key) in enumerate(self._KINDS):
            info["num_safety_violation_" + kind] = nviol[:, i]
            info[cost_key] = cost[:, i]
        return obs, reward, constraint, pre, obs, done.bool(), info
