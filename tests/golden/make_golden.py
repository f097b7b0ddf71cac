"""Generate golden vectors by running the REFERENCE's own Python in the build container.

Run here (no GPU):  ``python tests/golden/make_golden.py``  → tests/golden/*.pt

The reference's ``sac_cbf_clf/model.py`` of each variant is imported by path from
/root/reference with (a) a module named ``torchdiffeq`` that exposes the restated
solver (torchdiffeq 0.2.3 is not installable here) and (b) the
module-global ``device`` patched to CPU (it is hard-coded to "cuda", U model.py:11).
So the NODE right-hand side, ``train_step`` and the MLP classes that produce these
numbers are the reference's, unmodified; only the solver is the restatement.

/root/reference does not exist on the GPU box; the committed .pt files are what
travels.  Nothing under tests/ reads /root/reference at run time except this script.
"""
import importlib.util
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import torchdiffeq_restated  # noqa: E402

REF = "/root/reference"
VARIANTS = {
    "U": "NLBAC_Unicycle_RL_training/Unicycle_RL_training",
    "C": "NLBAC_SimulatedCarsFollowing_RL_training/Simulated_Car_Following_RL_training",
    "P": "NLBAC_pvtol_RL_training/Pvtol_RL_training",
    "BU": "neural_barrier_certificate/neural_barrier_certificate_NLBAC_Unicycle_RL_training/Unicycle_RL_training",
    "BP": "neural_barrier_certificate/neural_barrier_certificate_NLBAC_pvtol_RL_training/Pvtol_RL_training",
}


def import_ref_model(tag):
    """Import <variant>/sac_cbf_clf/model.py as an isolated module."""
    stub = types.ModuleType("torchdiffeq")
    stub.odeint = torchdiffeq_restated.odeint
    sys.modules["torchdiffeq"] = stub
    path = os.path.join(REF, VARIANTS[tag], "sac_cbf_clf", "model.py")
    spec = importlib.util.spec_from_file_location("ref_model_" + tag, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.device = torch.device("cpu")
    return mod


def make_inputs(tag, B, gen):
    """Synthetic batches with the env bounds of SURVEY.md §8(d)."""
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    if tag in ("U", "BU"):
        x = torch.stack([r(B) * 6 - 3, r(B) * 6 - 3, (r(B) * 2 - 1) * torch.pi], 1)
        u = torch.stack([r(B) * 7 - 3.5, r(B) * 24 - 12], 1)
        return x, u, None
    if tag in ("P", "BP"):
        x = torch.stack([r(B) * 14 - 7, r(B) * 12 - 6, (r(B) * 2 - 1) * torch.pi,
                         r(B) * 4 - 2, r(B) * 4 - 2, r(B) * 2], 1)
        u = torch.stack([r(B) * 7 - 3.5, r(B) * 30 - 15], 1)
        return x, u, None
    pos = torch.tensor([42.0, 34, 26, 18, 10]) + (r(B, 5) * 10 - 5)
    vel = 3 + 0.5 * torch.randn(B, 5, generator=gen)
    x = torch.stack([pos, vel], 2).reshape(B, 10)
    u = r(B, 1) * 6 - 3
    tk = r(B, 1) * 6
    return x, u, tk


def build_node(mod, tag):
    torch.manual_seed(0)
    if tag in ("U", "BU"):
        return mod.NeuralODEModel(3, 3, 6)
    if tag in ("P", "BP"):
        return mod.NeuralODEModel(6, 6, 12)
    return mod.NeuralODEModel(12, 10)


def node_case(tag, B=24):
    mod = import_ref_model(tag)
    model = build_node(mod, tag)
    gen = torch.Generator().manual_seed(1234)
    x, u, tk = make_inputs(tag, B, gen)
    cols = [x, u] if tk is None else [x, u, tk]
    y0 = torch.cat(cols, 1)
    n_s = x.shape[1]
    out = {"tag": tag, "state_dict": {k: v.clone() for k, v in model.state_dict().items()},
           "y0": y0.clone(), "n_s": n_s, "cases": {}}
    dt = 0.02
    specs = {
        "euler_1": ("euler", torch.tensor([0.0, dt]), None),
        "midpoint_1": ("midpoint", torch.tensor([0.0, dt]), None),
        "rk4_1": ("rk4", torch.tensor([0.0, dt]), None),
        "rk4_20": ("rk4", dt * torch.arange(21, dtype=torch.float32), None),
        "rk4_sub": ("rk4", torch.tensor([0.0, 0.05, 0.1]), {"step_size": 0.02}),
        "dopri5_1": ("dopri5", torch.tensor([0.0, dt]), None),
        "dopri5_long": ("dopri5", torch.tensor([0.0, 0.5, 1.0]), None),
    }
    for name, (method, t, options) in specs.items():
        y0g = y0.clone().requires_grad_(True)
        model.zero_grad()
        stats = {}
        sol = torchdiffeq_restated.odeint(model, y0g, t, method=method, atol=1e-7, rtol=1e-5,
                                          options=options, _stats=stats)
        cot = torch.randn(sol.shape, generator=gen)
        cot[0] = 0  # sol[0] is y0 itself; keep the test about the solve
        (sol * cot).sum().backward()
        grads = {k: p.grad.clone() for k, p in model.named_parameters()}
        out["cases"][name] = {"method": method, "t": t.clone(), "options": options,
                              "sol": sol.detach().clone(), "cot": cot, "grad_y0": y0g.grad.clone(),
                              "grad_params": grads, "stats": stats}
    # train_step trace through the reference's own function (3 Adam steps, euler like the shipped code)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    nxt = x + 0.02 * torch.randn(x.shape, generator=gen)
    losses = []
    for _ in range(3):
        if tk is None:
            losses.append(mod.train_step(model, x, u, nxt, opt, torch.nn.MSELoss(), B, "euler", dt))
        else:
            losses.append(mod.train_step(model, x, u, nxt, tk, opt, torch.nn.MSELoss(), B, "euler", dt))
    out["train"] = {"next_state": nxt, "losses": losses, "horizon": B,
                    "state_dict_after": {k: v.clone() for k, v in model.state_dict().items()}}
    return out


class _Box:
    def __init__(self, low, high):
        import numpy as np
        self.low, self.high = np.array(low, dtype="float32"), np.array(high, dtype="float32")
        self.shape = self.low.shape


def mlp_case(tag, B=16):
    """Forward values + grads of the reference MLP classes on fixed inputs and injected eps."""
    mod = import_ref_model(tag)
    n_o, n_u, k_lya, hi = {"U": (7, 2, 2, [3.5, 12.0]), "C": (10, 1, 4, [3.0]),
                           "P": (11, 2, 11, [3.5, 15.0]), "BU": (7, 2, 2, [3.5, 12.0]),
                           "BP": (11, 2, 11, [3.5, 15.0])}[tag]
    gen = torch.Generator().manual_seed(99)
    torch.manual_seed(7)
    H = 256
    obs = torch.randn(B, n_o, generator=gen)
    act = torch.randn(B, n_u, generator=gen)
    lya_in = torch.randn(B, k_lya, generator=gen)
    eps = torch.randn(B, n_u, generator=gen)
    out = {"tag": tag, "obs": obs, "act": act, "lya_in": lya_in, "eps": eps, "hidden": H,
           "action_high": hi}
    q = mod.QNetwork(n_o, n_u, H)
    for p in q.parameters():   # biases are zero-initialised upstream; perturb so they are exercised
        if p.ndim == 1:
            p.data.normal_(0, 0.1, generator=gen)
    obs_g, act_g = obs.clone().requires_grad_(), act.clone().requires_grad_()
    q1, q2 = q(obs_g, act_g)
    c1, c2 = torch.randn(B, 1, generator=gen), torch.randn(B, 1, generator=gen)
    ((q1 * c1).sum() + (q2 * c2).sum()).backward()
    out["q"] = {"state_dict": {k: v.clone() for k, v in q.state_dict().items()}, "q1": q1.detach(),
                "q2": q2.detach(), "c1": c1, "c2": c2, "grad_obs": obs_g.grad, "grad_act": act_g.grad,
                "grad_params": {k: p.grad.clone() for k, p in q.named_parameters()}}
    lya = mod.LyaNetwork(k_lya, H)
    for p in lya.parameters():
        if p.ndim == 1:
            p.data.normal_(0, 0.1, generator=gen)
    li = lya_in.clone().requires_grad_()
    v = lya(li)
    cv = torch.randn(B, 1, generator=gen)
    (v * cv).sum().backward()
    out["lya"] = {"state_dict": {k: v_.clone() for k, v_ in lya.state_dict().items()}, "v": v.detach(),
                  "c": cv, "grad_in": li.grad, "grad_params": {k: p.grad.clone() for k, p in lya.named_parameters()}}
    if hasattr(mod, "BarrierNetwork"):
        bar = mod.BarrierNetwork(n_o, n_u, H)
        for p in bar.parameters():
            if p.ndim == 1:
                p.data.normal_(0, 0.1, generator=gen)
        obs_g, act_g = obs.clone().requires_grad_(), act.clone().requires_grad_()
        b = bar(obs_g, act_g)
        cb = torch.randn(B, 1, generator=gen)
        (b * cb).sum().backward()
        out["barrier"] = {"state_dict": {k: v_.clone() for k, v_ in bar.state_dict().items()},
                          "b": b.detach(), "c": cb, "grad_obs": obs_g.grad, "grad_act": act_g.grad,
                          "grad_params": {k: p.grad.clone() for k, p in bar.named_parameters()}}
    pol = mod.GaussianPolicy(n_o, n_u, H, _Box([-h for h in hi], hi))
    for p in pol.parameters():
        if p.ndim == 1:
            p.data.normal_(0, 0.1, generator=gen)
    # inject eps: Normal.rsample draws torch.randn-like noise via _standard_normal → patch it
    import torch.distributions.normal as tdn
    orig = tdn._standard_normal
    tdn._standard_normal = lambda shape, dtype, device: eps.to(dtype)
    try:
        obs_g = obs.clone().requires_grad_()
        a, logp, mean = pol.sample(obs_g)
    finally:
        tdn._standard_normal = orig
    ca, cl = torch.randn(B, n_u, generator=gen), torch.randn(B, 1, generator=gen)
    ((a * ca).sum() + (logp * cl).sum()).backward()
    out["policy"] = {"state_dict": {k: v_.clone() for k, v_ in pol.state_dict().items()},
                     "action": a.detach(), "log_prob": logp.detach(), "mean": mean.detach(),
                     "ca": ca, "cl": cl, "grad_obs": obs_g.grad,
                     "grad_params": {k: p.grad.clone() for k, p in pol.named_parameters()}}
    return out


def main():
    for tag in ("U", "C", "P"):
        torch.save(node_case(tag), os.path.join(HERE, "node_%s.pt" % tag))
        print("node", tag, "ok")
    for tag in ("U", "C", "BP"):
        torch.save(mlp_case(tag), os.path.join(HERE, "mlp_%s.pt" % tag))
        print("mlp", tag, "ok")


if __name__ == "__main__":
    main()
