"""Golden trace of the REFERENCE's own ``SAC_CBF_CLF.update_parameters`` (U sac_cbf_clf/sac_cbf_clf.py:181-319) over 12
consecutive updates, Unicycle variant.  Run here:  ``python tests/golden/make_golden_update.py`` -> update_U.pt

The agent object is created with ``object.__new__`` (its constructor needs gym) and given the attributes the
constructor would have set (U sac_cbf_clf.py:28-134): the reference's own network classes, ``torch.optim.Adam``
optimisers with the reference's learning rates, multipliers, rho.  ``update_parameters`` itself, ``get_cbf_clf_part``,
``get_policy_loss_2``, the backup variants, ``train_step``, ``soft_update`` and ``DynamicsModel.get_state`` run
unmodified.  Two things are injected: the minibatches (a stand-in memory whose ``sample`` returns fixed arrays) and
the Gaussian noise ``Normal.rsample`` draws (``torch.distributions.normal._standard_normal`` pops pre-generated
tensors, recorded in the fixture), because PyTorch's RNG stream is not part of the contract.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import _Box  # noqa: E402
from make_golden_constraints import import_variant, _Env, _Args, _perturb_biases, _sd  # noqa: E402

N_UPDATES, NODE_N = 12, 512


class _Memory:
    """Stand-in for ReplayMemory: ``sample`` returns the next prepared batch (10 fields like U replay_memory.py:21-25)."""

    def __init__(self, batches, position):
        self.batches, self.i, self.position = batches, 0, position

    def sample(self, batch_size):
        b = self.batches[self.i]
        self.i += 1
        assert len(b[0]) == batch_size, (len(b[0]), batch_size)
        return b


def _batch(gen, n):
    r = lambda *s: torch.rand(*s, generator=gen, dtype=torch.float64)  # noqa: E731
    l_p, goal = 0.03, torch.tensor([2.5, 2.5], dtype=torch.float64)

    def obs_of(x, y, th):
        rel = goal - torch.stack([x, y], 1)
        d = rel.norm(dim=1)
        c, s = torch.cos(th), torch.sin(th)
        cx = (c * rel[:, 0] + s * rel[:, 1]) / (d + 1e-3)
        cy = (-s * rel[:, 0] + c * rel[:, 1]) / (d + 1e-3)
        return torch.stack([x, y, c, s, cx, cy, torch.exp(-d)], 1)

    x, y, th = r(n) * 4 - 2, r(n) * 4 - 2, (r(n) * 2 - 1) * np.pi
    act = torch.stack([r(n) * 7 - 3.5, r(n) * 24 - 12], 1)
    x2 = x + 0.02 * act[:, 0] * torch.cos(th)
    y2 = y + 0.02 * act[:, 0] * torch.sin(th)
    th2 = th + 0.02 * act[:, 1]
    obs, nobs = obs_of(x, y, th), obs_of(x2, y2, th2)
    cp = torch.stack([x + l_p * torch.cos(th), y + l_p * torch.sin(th)], 1)
    ncp = torch.stack([x2 + l_p * torch.cos(th2), y2 + l_p * torch.sin(th2)], 1)
    reward = -(goal - torch.stack([x2, y2], 1)).norm(dim=1) * 0.1
    cons = (goal - ncp).norm(dim=1)
    mask = (r(n) > 0.02).double()
    f32 = lambda t: t.float().double().numpy()  # noqa: E731  values exactly representable in float32
    return (f32(obs), f32(act), f32(reward), f32(cons), f32(cp), f32(ncp), f32(nobs), f32(mask), None, None)


def _batch_C(gen, n):
    """SimulatedCars minibatch (C replay layout: obs 10, action 1, reward, constraint, previous_positions 4,
    next_positions 4, next_obs 10, mask, t, next_t)."""
    r = lambda *s: torch.rand(*s, generator=gen, dtype=torch.float64)  # noqa: E731
    pos = torch.tensor([42.0, 36, 31, 26, 21], dtype=torch.float64) + (r(n, 5) * 4 - 2)
    vel = 3 + 0.5 * torch.randn(n, 5, generator=gen, dtype=torch.float64)
    act = r(n, 1) * 6 - 3
    pos2 = pos + 0.02 * vel
    vel2 = vel + 0.02 * torch.randn(n, 5, generator=gen, dtype=torch.float64)
    vel2[:, 3] = vel2[:, 3] + 0.02 * act[:, 0]
    st = torch.stack([pos, vel], 2).reshape(n, 10)
    st2 = torch.stack([pos2, vel2], 2).reshape(n, 10)
    scale = torch.tensor([100.0, 30.0] * 5, dtype=torch.float64)
    obs, nobs = st / scale, st2 / scale
    reward = -(act[:, 0] ** 2) * 0.05
    cons = (st2[:, 4] - st2[:, 6] - 9.5).abs()
    mask = (r(n) > 0.02).double()
    t = torch.floor(r(n) * 300) * 0.02
    f32 = lambda x: x.float().double().numpy()  # noqa: E731
    return (f32(obs), f32(act), f32(reward), f32(cons), f32(st[:, 4:8]), f32(st2[:, 4:8]), f32(nobs), f32(mask), f32(t),
            f32(t + 0.02))


def _batch_P(gen, n, dm):
    """Pvtol minibatch (P replay layout: obs 11, action 2, reward, constraint, lya_pre_term 11, lya_next_term 11,
    next_obs 11, mask, None, None); observations through the reference's own DynamicsModel.get_obs."""
    r = lambda *s: torch.rand(*s, generator=gen)  # noqa: E731
    hz = torch.tensor([[-2.5, -2.5], [-2.5, 2.5], [0.0, -3.5], [0.0, 3.5], [-4.5, 0.0]])
    x, y = r(n) * 14 - 7, r(n) * 12 - 6
    k = torch.randint(0, 5, (n,), generator=gen)
    near = r(n) < 0.34
    x = torch.where(near, hz[k, 0] + (r(n) * 0.7 - 0.35), x)
    y = torch.where(near, hz[k, 1] + (r(n) * 0.7 - 0.35), y)
    th, vx, vy, thr = (r(n) * 2 - 1) * np.pi, r(n) * 4 - 2, r(n) * 4 - 2, r(n) * 2
    st = torch.stack([x, y, th, vx, vy, thr, x + (r(n) * 3 - 1.5)], 1)
    act = torch.stack([r(n) * 7 - 3.5, r(n) * 30 - 15], 1)
    st2 = torch.stack([x + 0.02 * vx, y + 0.02 * vy, th + 0.02 * act[:, 1] * 0.1, vx + 0.02 * torch.randn(n, generator=gen),
                       vy + 0.02 * torch.randn(n, generator=gen), (thr + 0.02 * act[:, 0]).clamp(0, 2),
                       st[:, 6] + 0.7 * (x + 0.02 * vx - st[:, 6])], 1)
    obs = dm.get_obs(st.unsqueeze(-1), torch.device("cpu")).squeeze(-1).detach()
    nobs = dm.get_obs(st2.unsqueeze(-1), torch.device("cpu")).squeeze(-1).detach()
    reward = -0.1 * (st2[:, :2] - 4.5).norm(dim=1)
    cons = (st2[:, :2] - 4.5).norm(dim=1)
    mask = (r(n) > 0.02).float()
    f = lambda t: t.float().double().numpy()  # noqa: E731
    return (f(obs), f(act), f(reward), f(cons), f(obs), f(nobs), f(nobs), f(mask), None, None)


def _batch_BU(gen, n):
    """BU minibatch: the Unicycle fields with barrier_signal (0 / -20 per hazard hit, BU unicycle_env.py:51-52) after
    constraint (BU replay_memory.py:12)."""
    b = list(_batch(gen, n))
    sig = -20.0 * (torch.rand(n, generator=gen) < 0.15).double().numpy()
    return tuple(b[:4] + [sig] + b[4:])


def _batch_BP(gen, n, dm):
    b = list(_batch_P(gen, n, dm))
    sig = -0.1 * (torch.rand(n, generator=gen) < 0.15).double().numpy()       # BP pvtol_env.py:28-29
    return tuple(b[:4] + [sig] + b[4:])


VARIANT = {
    "BP": dict(B=256, obs=11, act=2, lya_in=11, node=(6, 6, 12), low=[-3.5, -15.0], high=[3.5, 15.0], gamma_b=0.6, n_eps=3,
               num_cbfs=1, mode="Pvtol"),
    "BU": dict(B=128, obs=7, act=2, lya_in=2, node=(3, 3, 6), low=[-3.5, -12.0], high=[3.5, 12.0], gamma_b=5.0, n_eps=3,
               num_cbfs=1, mode="Unicycle"),
    "P": dict(B=256, obs=11, act=2, lya_in=11, node=(6, 6, 12), low=[-3.5, -15.0], high=[3.5, 15.0], gamma_b=0.6, n_eps=7,
              num_cbfs=9, mode="Pvtol"),
    "U": dict(B=128, obs=7, act=2, lya_in=2, node=(3, 3, 6), low=[-3.5, -12.0], high=[3.5, 12.0], gamma_b=50.0, n_eps=3,
              num_cbfs=7, mode="Unicycle"),
    "C": dict(B=256, obs=10, act=1, lya_in=4, node=(12, 10), low=[-3.0], high=[3.0], gamma_b=0.5, n_eps=5, num_cbfs=2,
              mode="SimulatedCars"),
}


def main(tag="U"):
    V = VARIANT[tag]
    B = V["B"]
    model, agent_mod, dyn_mod = import_variant(tag)
    gen = torch.Generator().manual_seed(777)
    torch.manual_seed(11)
    env = _Env()
    env.dynamics_mode, env.dt, env.hazards_radius = V["mode"], 0.02, 0.5
    env.hazards_locations = np.array([[0., 0.], [0., 1.], [0., -1.], [-1., 1.], [-1., -1.], [1., -1.], [1., 1.]]) * 1.5
    space = _Box(V["low"], V["high"])
    a = object.__new__(agent_mod.SAC_CBF_CLF)
    dev = torch.device("cpu")
    a.env, a.device, a.solver = env, dev, "euler"
    a.gamma, a.gamma_b, a.tau, a.alpha, a.backup_alpha = 0.99, V["gamma_b"], 0.005, 0.2, 0.2
    a.center_pos_num, a.policy_type, a.batch_size = V["lya_in"], "Gaussian", B
    a.target_update_interval, a.Lagrangian_multiplier_update_interval = 1, 8
    a.automatic_entropy_tuning, a.critic_lyapunov_lr = True, 0.0004
    no, na = V["obs"], V["act"]
    a.critic = model.QNetwork(no, na, 256)
    a.critic_target = model.QNetwork(no, na, 256)
    a.lyapunovNet = model.LyaNetwork(V["lya_in"], 256)
    a.lyapunovNet_target = model.LyaNetwork(V["lya_in"], 256)
    a.policy = model.GaussianPolicy(no, na, 256, space)
    a.backup_policy = model.GaussianPolicy(no, na, 256, space)
    for net in (a.critic, a.lyapunovNet, a.policy, a.backup_policy):
        _perturb_biases(net, gen)
    nets_names = ["critic", "lyapunovNet", "policy", "backup_policy", "neural_ode_model"]
    targets = ["critic_target", "lyapunovNet_target"]
    if tag in ("BU", "BP"):      # learned barrier: its own net / target / optimiser (BU sac_cbf_clf.py:52-61); no backup controller
        a.BarrierNet, a.BarrierNet_target = model.BarrierNetwork(no, na, 256), model.BarrierNetwork(no, na, 256)
        _perturb_biases(a.BarrierNet, gen)
        agent_mod.hard_update(a.BarrierNet_target, a.BarrierNet)
        a.BarrierNet_optim = torch.optim.Adam(a.BarrierNet.parameters(), lr=a.critic_lyapunov_lr)
        nets_names = ["critic", "lyapunovNet", "policy", "BarrierNet", "neural_ode_model"]
        targets = ["critic_target", "lyapunovNet_target", "BarrierNet_target"]
    agent_mod.hard_update(a.critic_target, a.critic)
    agent_mod.hard_update(a.lyapunovNet_target, a.lyapunovNet)
    a.critic_optim = torch.optim.Adam(a.critic.parameters(), lr=a.critic_lyapunov_lr)
    a.lyaNet_optim = torch.optim.Adam(a.lyapunovNet.parameters(), lr=a.critic_lyapunov_lr)
    a.policy_optim = torch.optim.Adam(a.policy.parameters(), lr=3e-4)
    a.backup_policy_optim = torch.optim.Adam(a.backup_policy.parameters(), lr=3e-4)
    a.target_entropy = -float(na)
    a.log_alpha = torch.zeros(1, requires_grad=True)
    a.alpha_optim = torch.optim.Adam([a.log_alpha], lr=3e-4)
    a.backup_log_alpha = torch.zeros(1, requires_grad=True)
    a.backup_alpha_optim = torch.optim.Adam([a.backup_log_alpha], lr=3e-4)
    a.cost_limit, a.augmented_term, a.augmented_ratio = 0.0, 1.0, 1.0005
    a.num_cbfs, a.l_p, a.action_dim, a.num_constraints = V["num_cbfs"], 0.03, na, V["num_cbfs"] + 1
    a.lambda_values = [0.0] * (V["num_cbfs"] + 1)
    a.backup_lambda_values = [0.0] * V["num_cbfs"]
    a.neural_ode_model = model.NeuralODEModel(*V["node"])
    a.neural_ode_model_optimizer = torch.optim.Adam(a.neural_ode_model.parameters(), lr=1e-3)
    a.model_loss_func = agent_mod.PoseLoss()
    if tag == "BP":
        env.safety_operator_follow = 0.7
    if tag == "P":   # the env attributes P's constraint block reads (as in make_golden_constraints.case_P) + P's own knobs
        env.hazards_radius = 0.25
        env.hazard_locations = np.array([[-2.5, -2.5], [-2.5, 2.5], [0.0, -3.5], [0.0, 3.5], [-4.5, 0.0]])
        env.safety_operator_follow, env.operator_dist = 0.7, 0.001
        env.y_max, env.y_min = 13.0, -13.0
        a.backup_update_interval, a.backup_augmented_term = 4, 1.0      # default 20: 4 puts three backup updates in the trace
    dm = dyn_mod.DynamicsModel(env, _Args())

    init = {k: _sd(getattr(a, k)) for k in nets_names}
    mkb = {"U": _batch, "BU": _batch_BU, "C": _batch_C, "P": lambda g_, n_: _batch_P(g_, n_, dm),
           "BP": lambda g_, n_: _batch_BP(g_, n_, dm)}[tag]
    batches = [mkb(gen, B) for _ in range(N_UPDATES)]
    # U / C: 8, 16: multipliers due; 10: NODE fit due.  P: 28, 32, 36 update the backup controller (interval 4), 32: both
    # multiplier sets due (backup: every 8 * 4), 30: NODE fit
    first = 28 if tag == "P" else 8
    updates_list = list(range(first, first + N_UPDATES))
    node_batches = [mkb(gen, NODE_N) for u in updates_list if u % 10 == 0]
    n_eps = lambda u: (7 if u % 4 == 0 else 4) if tag == "P" else V["n_eps"]  # noqa: E731
    eps = [[torch.randn(B, na, generator=gen) for _ in range(n_eps(u))] for u in updates_list]
    mem, node_mem = _Memory(batches, B * 4), _Memory(node_batches, NODE_N)

    import torch.distributions.normal as tdn
    orig = tdn._standard_normal
    queue = []

    def pop_eps(shape, dtype, device):
        e = queue.pop(0)
        assert tuple(shape) == tuple(e.shape), (shape, e.shape)
        return e.to(dtype)

    tdn._standard_normal = pop_eps
    returns, lam_trace, blam_trace, rho_trace = [], [], [], []
    try:
        for i, u in enumerate(updates_list):
            queue.extend(eps[i])             # next_state sample, pi, backup_pi (+ C: the two re-samples inside the rollouts)
            returns.append(a.update_parameters(mem, B, u, dm, node_mem, 10, 1) if tag in ("P", "BP") else
                           a.update_parameters(mem, B, u, dm, node_mem, 10))
            lam_trace.append([float(v) for v in a.lambda_values])
            blam_trace.append([float(v) for v in a.backup_lambda_values] if tag not in ("BU", "BP") else [0.0])
            rho_trace.append(float(a.augmented_term))
            if i == 0:
                after_first = {k: _sd(getattr(a, k)) for k in nets_names[:-1] + ["critic_target"]}
            assert not queue
    finally:
        tdn._standard_normal = orig

    to_t = lambda b: [None if x is None else torch.tensor(x, dtype=torch.float32) for x in b]  # noqa: E731
    out = {
        "B": B, "node_N": NODE_N, "updates": updates_list, "init": init, "batches": [to_t(b) for b in batches],
        "node_batches": [to_t(b) for b in node_batches], "eps": eps,
        "returns": torch.tensor(returns, dtype=torch.float64),        # qf1, qf2, lf, policy_loss_1, alpha_loss, alpha
        "final": {k: _sd(getattr(a, k)) for k in nets_names + targets},
        "log_alpha": a.log_alpha.detach().clone(), "backup_log_alpha": a.backup_log_alpha.detach().clone(),
        "lambda_values": [float(v) for v in a.lambda_values],
        "backup_lambda_values": [float(v) for v in getattr(a, "backup_lambda_values", [])],
        "augmented_term": float(a.augmented_term),
        "backup_augmented_term": float(getattr(a, "backup_augmented_term", 0.0)),
        "lambda_trace": torch.tensor(lam_trace, dtype=torch.float64),
        "backup_lambda_trace": torch.tensor(blam_trace, dtype=torch.float64),
        "rho_trace": torch.tensor(rho_trace, dtype=torch.float64), "after_first": after_first,
        "consts": {"dt": 0.02, "gamma_b": V["gamma_b"],
                   "hazards": env.hazard_locations if tag == "P" else getattr(env, "hazards_locations", None),
                   "hazards_radius": env.hazards_radius, "safety_operator_follow": 0.7, "operator_dist": 0.001,
                   "y_max": 13.0, "y_min": -13.0, "backup_update_interval": 4,
                   "l_p": 0.03, "gamma": 0.99, "tau": 0.005, "alpha": 0.2, "target_entropy": -float(na), "lr": 3e-4,
                   "critic_lr": 4e-4, "node_lr": 1e-3, "action_low": V["low"], "action_high": V["high"]},
    }
    torch.save(out, os.path.join(HERE, "update_%s.pt" % tag))
    print("update_%s.pt: returns\n" % tag, out["returns"])
    print("lambda", out["lambda_values"], "rho", out["augmented_term"])


if __name__ == "__main__":
    for tag_ in (sys.argv[1:] or ["U", "C", "P", "BU", "BP"]):
        main(tag_)
