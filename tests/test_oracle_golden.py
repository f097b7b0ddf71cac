"""Pin the oracle restatements (oracle/nlbac_ref.py) to golden vectors produced by the REFERENCE's
own classes (tests/golden/make_golden.py, run in the build container against /root/reference)."""
import os

import pytest
import torch

from oracle import nlbac_ref
from oracle.torchdiffeq_restated import odeint


def _close(a, b, rtol=2e-5, atol=1e-6):
    scale = max(b.abs().max().item(), 1e-12)
    return (a - b).abs().max().item() <= rtol * scale + atol


def _node(tag):
    if tag == "U":
        return nlbac_ref.AffineNODE(3, 3, 6)
    if tag == "P":
        return nlbac_ref.AffineNODE(6, 6, 12)
    return nlbac_ref.BlackBoxNODE(12, 10)


@pytest.mark.parametrize("tag", ["U", "C", "P"])
def test_node_solve_and_grads_match_reference(tag, golden_dir):
    g = torch.load(os.path.join(golden_dir, "node_%s.pt" % tag), weights_only=False)
    model = _node(tag)
    model.load_state_dict(g["state_dict"])  # key layout is part of the boundary
    for name, c in g["cases"].items():
        y0 = g["y0"].clone().requires_grad_(True)
        model.zero_grad()
        sol = odeint(model, y0, c["t"], method=c["method"], atol=1e-7, rtol=1e-5, options=c["options"])
        assert _close(sol.detach(), c["sol"]), name
        (sol * c["cot"]).sum().backward()
        assert _close(y0.grad, c["grad_y0"]), name
        for k, p in model.named_parameters():
            assert _close(p.grad, c["grad_params"][k], rtol=5e-5), (name, k)


@pytest.mark.parametrize("tag", ["U", "C", "P"])
def test_train_step_trace_matches_reference(tag, golden_dir):
    g = torch.load(os.path.join(golden_dir, "node_%s.pt" % tag), weights_only=False)
    model = _node(tag)
    model.load_state_dict(g["state_dict"])
    n_s = g["n_s"]
    y0 = g["y0"]
    x, u = y0[:, :n_s], y0[:, n_s:n_s + (1 if tag == "C" else 2)]
    tk = y0[:, 11:12] if tag == "C" else None
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    tr = g["train"]
    for want in tr["losses"]:
        got = nlbac_ref.train_step(model, x, u, tr["next_state"], opt, torch.nn.MSELoss(), tr["horizon"],
                                   "euler", 0.02, time_batch=tk)
        assert abs(got - want) <= 1e-5 * abs(want) + 1e-9
    for k, v in model.state_dict().items():
        assert _close(v, tr["state_dict_after"][k], rtol=1e-4), k


@pytest.mark.parametrize("tag", ["U", "C", "BP"])
def test_mlps_match_reference(tag, golden_dir):
    g = torch.load(os.path.join(golden_dir, "mlp_%s.pt" % tag), weights_only=False)
    obs, act, H = g["obs"], g["act"], g["hidden"]
    n_o, n_u = obs.shape[1], act.shape[1]
    q = nlbac_ref.QNetwork(n_o, n_u, H)
    q.load_state_dict(g["q"]["state_dict"])
    q1, q2 = q(obs, act)
    assert _close(q1, g["q"]["q1"]) and _close(q2, g["q"]["q2"])
    lya = nlbac_ref.LyaNetwork(g["lya_in"].shape[1], H)
    lya.load_state_dict(g["lya"]["state_dict"])
    assert _close(lya(g["lya_in"]), g["lya"]["v"])
    if "barrier" in g:
        bar = nlbac_ref.BarrierNetwork(n_o, n_u, H)
        bar.load_state_dict(g["barrier"]["state_dict"])
        assert _close(bar(obs, act), g["barrier"]["b"])
    hi = g["action_high"]
    pol = nlbac_ref.GaussianPolicy(n_o, n_u, H, hi, [-h for h in hi])
    pol.load_state_dict(g["policy"]["state_dict"])
    og = obs.clone().requires_grad_()
    a, logp, mean = pol.sample(og, eps=g["eps"])
    assert _close(a, g["policy"]["action"]) and _close(logp, g["policy"]["log_prob"], rtol=1e-4)
    assert _close(mean, g["policy"]["mean"])
    ((a * g["policy"]["ca"]).sum() + (logp * g["policy"]["cl"]).sum()).backward()
    assert _close(og.grad, g["policy"]["grad_obs"], rtol=1e-4)
    for k, p in pol.named_parameters():
        assert _close(p.grad, g["policy"]["grad_params"][k], rtol=1e-4), k
