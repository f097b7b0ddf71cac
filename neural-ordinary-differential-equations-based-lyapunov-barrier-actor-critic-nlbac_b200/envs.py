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
