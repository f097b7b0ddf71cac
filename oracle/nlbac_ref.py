"""CPU restatement of the reference's NODE models, ``train_step`` and MLPs.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Device-agnostic re-statements
of the classes in the reference's ``sac_cbf_clf/model.py`` (which pin
``device = torch.device("cuda")`` at import, U model.py:11, and so cannot run
on CPU unmodified).  Each is checked against the real reference class, imported
from /root/reference in the build container, by tests/golden/make_golden.py →
tests/test_oracle_golden.py.

Path tags: U = NLBAC_Unicycle_RL_training/Unicycle_RL_training,
C = NLBAC_SimulatedCarsFollowing_RL_training/Simulated_Car_Following_RL_training,
P = NLBAC_pvtol_RL_training/Pvtol_RL_training,
BU/BP = neural_barrier_certificate/... (see SURVEY.md header).
"""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .torchdiffeq_restated import odeint

LOG_SIG_MAX = 2      # U sac_cbf_clf/model.py:7
LOG_SIG_MIN = -20    # U sac_cbf_clf/model.py:8
EPSILON = 1e-6       # U sac_cbf_clf/model.py:9


def _mlp(sizes):
    layers = []
    for i in range(len(sizes) - 1):
        layers.append(nn.Linear(sizes[i], sizes[i + 1]))
        if i < len(sizes) - 2:
            layers.append(nn.ReLU())
    return nn.Sequential(*layers)


class AffineNODE(nn.Module):
    """Control-affine NODE ``x' = f(x) + g(x) u`` — U model.py:177-217, P model.py:179-222.

    ``NeuralODEModel(input_dim, output_dim1, output_dim2)``; hidden 100; f_net has 5
    Linear layers, g_net 4; ``forward(t, s)`` ignores ``t`` and returns zeros for the
    action columns.  State-dict keys match the reference (f_net.{0,2,4,6,8}, g_net.{0,2,4,6}).
    """

    def __init__(self, input_dim, output_dim1, output_dim2, hidden_dim=100):
        super().__init__()
        self.input_dim, self.output_dim1, self.output_dim2 = input_dim, output_dim1, output_dim2
        self.n_s = input_dim
        self.n_u = output_dim2 // output_dim1
        self.f_net = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim, hidden_dim, output_dim1])
        self.g_net = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim, output_dim2])

    def forward(self, t, s):
        n_s, n_u = self.n_s, self.n_u
        x = s[..., :n_s]
        u = s[..., n_s:n_s + n_u]
        f_x = self.f_net(x)
        g_x = self.g_net(x).reshape(-1, n_s, n_u)
        ds = f_x + torch.bmm(g_x, u.reshape(-1, n_u, 1)).squeeze(-1)
        return torch.cat((ds, torch.zeros_like(u)), -1)


class BlackBoxNODE(nn.Module):
    """``x' = net([x, u, t_k])`` — C model.py:179-206 (12→64→64→64→10)."""

    def __init__(self, input_dim, output_dim, hidden_dim=64):
        super().__init__()
        self.input_dim, self.output_dim = input_dim, output_dim
        self.net = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim, output_dim])

    def forward(self, t, s):
        ds = self.net(s)
        pad = torch.zeros_like(s[:, self.output_dim:])
        return torch.cat((ds, pad), -1)


def train_step(model, state, action, next_state, optimizer, loss_func, horizon, solver, time_interval,
               time_batch=None):
    """U model.py:221-260 / P model.py:224-267; with ``time_batch`` C model.py:208-253."""
    assert state.ndim == 2 and action.ndim == 2 and next_state.ndim == 2
    model.train()
    optimizer.zero_grad()
    cols = [state, action] if time_batch is None else [state, action, time_batch]
    model_input_t = torch.cat(cols, dim=-1)
    t_span = torch.tensor([0.0, time_interval]).to(state.device)
    pred = odeint(model, model_input_t, t_span, method=solver, atol=1e-7, rtol=1e-5)[-1]
    pred = pred[:, :state.shape[1]]
    loss = loss_func(pred, next_state)
    loss.backward()
    optimizer.step()
    return loss.item() / horizon


def weights_init_(m):
    """U model.py:14-17."""
    if isinstance(m, nn.Linear):
        torch.nn.init.xavier_uniform_(m.weight, gain=1)
        torch.nn.init.constant_(m.bias, 0)


class QNetwork(nn.Module):
    """U model.py:37-64."""

    def __init__(self, num_inputs, num_actions, hidden_dim):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs + num_actions, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, 1)
        self.linear4 = nn.Linear(num_inputs + num_actions, hidden_dim)
        self.linear5 = nn.Linear(hidden_dim, hidden_dim)
        self.linear6 = nn.Linear(hidden_dim, 1)
        self.apply(weights_init_)

    def forward(self, state, action):
        xu = torch.cat([state, action], 1)
        x1 = self.linear3(F.relu(self.linear2(F.relu(self.linear1(xu)))))
        x2 = self.linear6(F.relu(self.linear5(F.relu(self.linear4(xu)))))
        return x1, x2


class LyaNetwork(nn.Module):
    """U model.py:67-83."""

    def __init__(self, num_inputs, hidden_dim):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, 1)
        self.apply(weights_init_)

    def forward(self, state):
        return self.linear3(F.relu(self.linear2(F.relu(self.linear1(state)))))


class BarrierNetwork(nn.Module):
    """BU model.py:67-86, BP model.py:68-87."""

    def __init__(self, num_inputs, num_actions, hidden_dim):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs + num_actions, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.linear3 = nn.Linear(hidden_dim, 1)
        self.apply(weights_init_)

    def forward(self, state, action):
        xu = torch.cat([state, action], 1)
        return self.linear3(F.relu(self.linear2(F.relu(self.linear1(xu)))))


class GaussianPolicy(nn.Module):
    """U model.py:86-133.  ``sample(state, eps=...)`` takes the N(0,1) draw explicitly
    (SURVEY.md App. D.10); with ``eps=None`` it draws from torch's global RNG like the reference."""

    def __init__(self, num_inputs, num_actions, hidden_dim, action_high=None, action_low=None):
        super().__init__()
        self.linear1 = nn.Linear(num_inputs, hidden_dim)
        self.linear2 = nn.Linear(hidden_dim, hidden_dim)
        self.mean_linear = nn.Linear(hidden_dim, num_actions)
        self.log_std_linear = nn.Linear(hidden_dim, num_actions)
        self.apply(weights_init_)
        if action_high is None:
            self.action_scale = torch.tensor(1.0)
            self.action_bias = torch.tensor(0.0)
        else:
            hi = torch.as_tensor(action_high, dtype=torch.float32)
            lo = torch.as_tensor(action_low, dtype=torch.float32)
            self.action_scale = (hi - lo) / 2.0
            self.action_bias = (hi + lo) / 2.0

    def forward(self, state):
        x = F.relu(self.linear2(F.relu(self.linear1(state))))
        mean = self.mean_linear(x)
        log_std = torch.clamp(self.log_std_linear(x), min=LOG_SIG_MIN, max=LOG_SIG_MAX)
        return mean, log_std

    def sample(self, state, eps=None):
        mean, log_std = self.forward(state)
        std = log_std.exp()
        if eps is None:
            eps = torch.randn_like(mean)
        x_t = mean + std * eps
        y_t = torch.tanh(x_t)
        action = y_t * self.action_scale + self.action_bias
        # Normal(mean, std).log_prob(x_t)
        log_prob = -((x_t - mean) ** 2) / (2 * std ** 2) - log_std - 0.5 * torch.log(torch.tensor(2 * torch.pi))
        log_prob = log_prob - torch.log(self.action_scale * (1 - y_t.pow(2)) + EPSILON)
        log_prob = log_prob.sum(1, keepdim=True)
        mean = torch.tanh(mean) * self.action_scale + self.action_bias
        return action, log_prob, mean
