"""Host-side mirror of the reference's ``sac_cbf_clf/model.py`` for the NODE path.

Same class names, constructor signatures, attribute names and state-dict keys as the reference, so
``node_model.pkl`` checkpoints (U sac_cbf_clf/sac_cbf_clf.py:336-339) load unchanged and
``torch.optim.Adam(model.parameters(), lr=1e-3)`` (U sac_cbf_clf.py:133) sees ordinary nn.Parameters.
The arithmetic of the solve lives in the CUDA kernels behind ``odeint``; nothing here integrates.
"""
import torch
import torch.nn as nn

from .odeint import odeint


def _mlp(sizes):
    layers = []
    for i in range(len(sizes) - 1):
        layers.append(nn.Linear(sizes[i], sizes[i + 1]))   # default nn.Linear init, like the reference NODE
        if i < len(sizes) - 2:
            layers.append(nn.ReLU())
    return nn.Sequential(*layers)


class NeuralODEModel(nn.Module):
    """``NeuralODEModel(input_dim, output_dim1, output_dim2)`` — control-affine ``x' = f(x) + g(x) u``
    (U model.py:177-217, P model.py:179-222; hidden 100) — or ``NeuralODEModel(input_dim, output_dim)`` —
    black-box ``x' = net([x, u, t_k])`` (C model.py:179-206; hidden 64).

    ``forward(t, s)`` returns ``[x', 0...]`` like the reference; it exists for API compatibility
    (eager torch ops on the module's device) and is NOT what ``odeint`` executes.
    """

    def __init__(self, input_dim, output_dim1, output_dim2=None, hidden_dim=None, n_u=None):
        super().__init__()
        self.input_dim = input_dim
        if output_dim2 is not None:
            hidden_dim = hidden_dim or 100
            self.output_dim1, self.output_dim2 = output_dim1, output_dim2
            self.f_net = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim, hidden_dim, output_dim1])
            self.g_net = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim, output_dim2])
        else:
            hidden_dim = hidden_dim or 64
            self.output_dim = output_dim1
            self.net = _mlp([input_dim, hidden_dim, hidden_dim, hidden_dim, output_dim1])
            if n_u is not None:
                self.n_u = n_u

    @property
    def affine(self):
        return hasattr(self, "f_net")

    def forward(self, t, s):
        if self.affine:
            n_s, n_u = self.input_dim, self.output_dim2 // self.output_dim1
            x, u = s[..., :n_s], s[..., n_s:n_s + n_u]
            g = self.g_net(x).reshape(-1, n_s, n_u)
            ds = self.f_net(x) + torch.bmm(g, u.reshape(-1, n_u, 1)).squeeze(-1)
            return torch.cat((ds, torch.zeros_like(u)), -1)
        ds = self.net(s)
        return torch.cat((ds, torch.zeros_like(s[:, self.output_dim:])), -1)


def train_step(model, state, action, next_state, *rest):
    """One NODE model-fit step with the reference's two signatures:

    U/P/BU/BP (U model.py:221-260):
        ``train_step(model, state, action, next_state, optimizer, loss_func, horizon, solver, time_interval)``
    C (C model.py:208-253):
        ``train_step(model, state, action, next_state, time_batch, optimizer, loss_func, horizon, solver, time_interval)``

    zero_grad -> fused solve over ``[0, time_interval]`` -> loss on the state columns -> fused backward
    -> ``optimizer.step()``; returns ``loss.item() / horizon`` like the reference (one host sync).
    """
    if len(rest) == 6:
        time_batch, optimizer, loss_func, horizon, solver, time_interval = rest
        assert state.ndim == 2 and action.ndim == 2 and next_state.ndim == 2 and time_batch.ndim == 2, \
            (state.shape, action.shape, next_state.shape, time_batch.shape)
        cols = (state, action, time_batch)
    elif len(rest) == 5:
        optimizer, loss_func, horizon, solver, time_interval = rest
        assert state.ndim == 2 and action.ndim == 2 and next_state.ndim == 2, \
            (state.shape, action.shape, next_state.shape)
        cols = (state, action)
    else:
        raise TypeError("train_step: expected the reference's 9- or 10-argument signature")
    model.train()
    optimizer.zero_grad()
    model_input_t = torch.cat(cols, dim=-1)
    t_span = [0.0, float(time_interval)]
    pred = odeint(model, model_input_t, t_span, method=solver, atol=1e-7, rtol=1e-5)[-1]
    pred = pred[:, :state.shape[1]]
    loss = loss_func(pred, next_state)
    loss.backward()
    optimizer.step()
    return loss.item() / horizon
