"""GPU parity of the constraint block, obs/state maps, Adam and Polyak kernels (nlbac_constraints_*,
nlbac_lookahead_*, nlbac_get_state, nlbac_get_obs_*, nlbac_adam_step, nlbac_soft_update) against the golden
vectors produced by the reference's own get_policy_loss_2 / DynamicsModel and against the CPU oracle.
Tolerance rel 1e-5 unless stated (max-norm per tensor)."""
import os

import pytest
import torch

from oracle import constraints as OC
from oracle import nlbac_ref
from constraint_cases import build, run_backup_sequence, run_sequence

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


class _Env:
    def __init__(self, mode):
        self.dynamics_mode = mode


@pytest.mark.parametrize("tag,env", [("U", "Unicycle"), ("C", "SimulatedCars"), ("BU", "Unicycle"), ("P", "Pvtol"),
                                     ("BP", "Pvtol")])
def test_policy_loss_2_matches_reference_and_oracle(tag, env, golden_dir):
    import nlbac_b200
    g = torch.load(os.path.join(golden_dir, "constraints_%s.pt" % tag), weights_only=False)
    nets = build(g, nlbac_b200, "cuda")
    dyn = nlbac_b200.DynamicsModel(_Env(env))
    losses, grads, st = run_sequence(g, nets, nlbac_b200.constraints, "cuda", dyn)
    o_losses, o_grads, _ = run_sequence(g, build(g, nlbac_ref, "cpu"), OC, "cpu")
    for got, want, ow in zip(losses, g["losses"], o_losses):
        assert _rel(got, want.reshape(-1)) <= 2 * TOL and _rel(got, ow) <= 2 * TOL
    for got, want, ow in zip(grads, g["grad_action"], o_grads):
        assert _rel(got, want) <= 5 * TOL and _rel(got, ow) <= 5 * TOL
    lam = torch.tensor([float(v) for v in st.lambda_values])
    assert _rel(lam, torch.tensor(g["lambda_after"])) <= 2 * TOL


@pytest.mark.parametrize("tag,env", [("U", "Unicycle"), ("C", "SimulatedCars"), ("P", "Pvtol")])
def test_backup_policy_loss_2_matches_reference_and_oracle(tag, env, golden_dir):
    import nlbac_b200
    g = torch.load(os.path.join(golden_dir, "constraints_backup_%s.pt" % tag), weights_only=False)
    dyn = nlbac_b200.DynamicsModel(_Env(env))
    losses, grads, st = run_backup_sequence(g, build(g, nlbac_b200, "cuda"), nlbac_b200.constraints, "cuda", dyn)
    o_losses, o_grads, _ = run_backup_sequence(g, build(g, nlbac_ref, "cpu"), OC, "cpu")
    for got, want, ow in zip(losses, g["losses"], o_losses):
        assert _rel(got, want.reshape(-1)) <= 2 * TOL and _rel(got, ow) <= 2 * TOL
    # action gradient = two / three chained NODE backward passes of a 1e-10-sized quantity: rel 1e-4 (measured 6e-5)
    for got, want, ow in zip(grads, g["grad_action"], o_grads):
        assert _rel(got, want) <= 10 * TOL and _rel(got, ow) <= 10 * TOL
    lam = torch.tensor([float(v) for v in st.backup_lambda_values])
    assert _rel(lam, torch.tensor(g["lambda_after"])) <= 2 * TOL


def test_get_state_and_get_obs(golden_dir):
    import nlbac_b200
    for tag, env in (("U", "Unicycle"), ("C", "SimulatedCars"), ("BU", "Unicycle")):
        g = torch.load(os.path.join(golden_dir, "constraints_%s.pt" % tag), weights_only=False)
        dm = nlbac_b200.DynamicsModel(_Env(env))
        st = dm.get_state(g["obs"].cuda())
        # float64 atan2 rounded to fp32, like the reference's numpy round trip: exact or 1 ulp (libm vs CUDA fp64 atan2)
        assert _rel(st, g["state_from_obs"]) <= 2e-7
        one = dm.get_state(g["obs"][0].cuda())
        assert one.shape == g["state_from_obs"][0].shape
    g = torch.load(os.path.join(golden_dir, "constraints_BU.pt"), weights_only=False)
    dm = nlbac_b200.DynamicsModel(_Env("Unicycle"))
    s = g["state"].cuda().requires_grad_(True)
    obs = dm.get_obs(s.unsqueeze(-1))
    assert obs.shape == (s.shape[0], 7, 1)
    assert _rel(obs.squeeze(-1), g["obs"]) <= TOL
    (obs.squeeze(-1) * g["obs_cot"].cuda()).sum().backward()
    assert _rel(s.grad, g["grad_state_get_obs"]) <= 2 * TOL
    p = torch.load(os.path.join(golden_dir, "constraints_Pmaps.pt"), weights_only=False)
    dm = nlbac_b200.DynamicsModel(_Env("Pvtol"))
    s = p["state7"].cuda().requires_grad_(True)
    obs = dm.get_obs(s, "cuda")
    assert _rel(obs, p["obs"]) <= TOL
    (obs * p["obs_cot"].cuda()).sum().backward()
    assert _rel(s.grad, p["grad_state_get_obs"]) <= 2 * TOL
    st, dyn = dm.get_state(p["obs"].cuda())
    assert _rel(st, p["state_from_obs"]) <= 2e-7 and _rel(dyn, p["state_dyn_from_obs"]) <= 2e-7
    c = torch.load(os.path.join(golden_dir, "constraints_C.pt"), weights_only=False)
    dm = nlbac_b200.DynamicsModel(_Env("SimulatedCars"))
    assert _rel(dm.get_obs(c["state"].cuda()), c["obs"]) <= 1e-7
    import numpy as np
    assert np.allclose(dm.get_state(c["obs"].numpy()), c["state_from_obs"].numpy(), rtol=1e-6)


@pytest.mark.parametrize("kind,B", [("unicycle", 1), ("unicycle", 100000), ("cars", 3), ("cars", 70001), ("barrier", 2),
                                    ("barrier", 65537), ("pvtol", 1), ("pvtol", 90001)])
def test_constraint_sums_vs_oracle_rows(kind, B):
    """Kernel rows + filter + reduction (and gradients) on synthetic tensors, including B=1 and multi-CTA sizes."""
    import nlbac_b200
    g = torch.Generator().manual_seed(B)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    V, Vn = r(B, 1), r(B, 1).requires_grad_(True)
    bs = 256.0
    if kind == "unicycle":
        hz = torch.tensor([[0., 0.], [0., 1.], [0., -1.], [-1., 1.], [-1., -1.], [1., -1.], [1., 1.]]) * 1.5
        a, b = r(B, 2) * 1.5, (r(B, 2) * 1.5).requires_grad_(True)
        desc = nlbac_b200.make_constraint_desc("unicycle", dt=0.02, gamma_b=50.0, gamma_l=1.0, batch_size=bs,
                                               collision_radius=0.525, hazards=hz.numpy())
        hs = 0.5 * (((a[:, None] - hz[None]) ** 2).sum(2) - 0.525 ** 2)
        hn = 0.5 * (((b[:, None] - hz[None]) ** 2).sum(2) - 0.525 ** 2)
        rows = torch.cat((-((hn - hs) / 0.02) - 50.0 * hs, (Vn - V) / 0.02 + V), 1)
        c = None
    elif kind == "cars":
        base = torch.tensor([42.0, 3, 36, 3, 31, 3, 26, 3, 21, 3])
        a, b, c = base + r(B, 10), (base + r(B, 10)).requires_grad_(True), (base + r(B, 10)).requires_grad_(True)
        desc = nlbac_b200.make_constraint_desc("cars", dt=0.02, gamma_b=0.5, gamma_l=0.15, batch_size=bs,
                                               lya_inv_dt=1.0, collision_radius=4.5)
        rr = []
        for ia, ib in ((4, 6), (6, 8)):
            h0, h1, h2 = (x[:, ia:ia + 1] - x[:, ib:ib + 1] - 4.5 for x in (a, b, c))
            l1, l2 = h1 - h0 + 0.5 * h0, h2 - h1 + 0.5 * h1
            rr.append(-(l2 - l1) - 0.5 * l1)
        rows = torch.cat(rr + [(Vn - V) + 0.15 * V], 1)
    elif kind == "pvtol":
        hz = torch.tensor([[-2.5, -2.5], [-2.5, 2.5], [0.0, -3.5], [0.0, 3.5], [-4.5, 0.0]])
        a = r(B, 7) * 2.5
        b, c, e = ((a + 0.3 * r(B, 7)).requires_grad_(True) for _ in range(3))
        desc = nlbac_b200.make_constraint_desc("pvtol", dt=0.02, gamma_b=0.6, gamma_l=0.1, batch_size=bs, lya_inv_dt=1.0,
                                               collision_radius=0.3, hazards=hz.numpy(), operator_dist=0.9, y_max=13.0,
                                               y_min=-13.0)
        seq = (a, b, c, e)
        rr = [OC.third_order_cbf(*[0.5 * (((s[:, None, :2] - hz[None]) ** 2).sum(2) - 0.3 ** 2) for s in seq], 0.6)]
        for f in (lambda s: s[:, 0:1] - s[:, 6:7] + 0.9, lambda s: -s[:, 0:1] + s[:, 6:7] + 0.9,
                  lambda s: -s[:, 1:2] + 13.0 - 10.0, lambda s: s[:, 1:2] + 13.0 - 10.0):
            rr.append(OC.third_order_cbf(*[f(s) for s in seq], 0.6))
        rows = torch.cat(rr + [(Vn - V) + 0.1 * V], 1)
    else:
        a, b, c = r(B, 1), r(B, 1).requires_grad_(True), None
        desc = nlbac_b200.make_constraint_desc("barrier", dt=0.02, gamma_b=5.0, gamma_l=1.0, batch_size=bs)
        rows = torch.cat((-(b - a) - 5.0 * a, (Vn - V) / 0.02 + V), 1)
    want = OC.required_matrix_from_rows(rows.double(), bs)
    w = torch.arange(1, want.shape[0] + 1, dtype=torch.float64).unsqueeze(-1)
    (want * w).sum().backward()
    e4 = e if kind == "pvtol" else None
    leaves = [t for t in (b, c, e4, Vn) if t is not None]
    want_g = [t.grad.clone() for t in leaves]
    dev = [t.detach().cuda().requires_grad_(True) for t in leaves]
    bb = dev[0]
    cc = dev[1] if c is not None else None
    ee = dev[2] if e4 is not None else None
    got = nlbac_b200.constraint_sums(desc, a.cuda(), bb, cc, V.cuda(), dev[-1], e=ee)
    assert got.shape == want.shape
    assert _rel(got, want) <= 2 * TOL     # fp32 atomics over up to 400 CTAs vs an fp64 reference
    (got * w.float().cuda()).sum().backward()
    for d_, wg in zip(dev, want_g):
        assert _rel(d_.grad, wg) <= TOL


def test_fused_adam_and_soft_update_match_torch():
    import nlbac_b200
    torch.manual_seed(0)
    a = torch.nn.Linear(37, 53).cuda()
    b = torch.nn.Linear(37, 53).cuda()
    b.load_state_dict(a.state_dict())
    oa = torch.optim.Adam(a.parameters(), lr=1e-3)
    ob = nlbac_b200.FusedAdam(b.parameters(), lr=1e-3)
    for i in range(5):
        x = torch.randn(16, 37, device="cuda")
        for net, opt in ((a, oa), (b, ob)):
            opt.zero_grad()
            net(x).pow(2).sum().backward()
            opt.step()
    for p, q in zip(a.parameters(), b.parameters()):
        assert _rel(q, p) <= 1e-6
    tgt_a, tgt_b = torch.nn.Linear(37, 53).cuda(), torch.nn.Linear(37, 53).cuda()
    tgt_b.load_state_dict(tgt_a.state_dict())
    for tp, sp in zip(tgt_a.parameters(), a.parameters()):
        tp.data.copy_(tp.data * (1.0 - 0.005) + sp.data * 0.005)
    nlbac_b200.soft_update(tgt_b, a, 0.005)
    for p, q in zip(tgt_a.parameters(), tgt_b.parameters()):
        assert _rel(q, p) <= 1e-7
    nlbac_b200.hard_update(tgt_b, a)
    assert all(torch.equal(p, q) for p, q in zip(tgt_b.parameters(), a.parameters()))


def test_multi_tensor_adam_chunks_and_ragged_sizes():
    """40 tensors of different sizes (two launches of <= 32), one of them empty, one without a gradient."""
    import nlbac_b200
    torch.manual_seed(1)
    shapes = [(i % 7 + 1, 3 * i + 1) for i in range(38)] + [(0,), (5,)]
    pa = [torch.nn.Parameter(torch.randn(*s, device="cuda")) for s in shapes]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    oa, ob = torch.optim.Adam(pa, lr=3e-4), nlbac_b200.FusedAdam(pb, lr=3e-4)
    for it in range(3):
        for ps, opt in ((pa, oa), (pb, ob)):
            opt.zero_grad()
            torch.manual_seed(10 + it)
            sum((p * torch.randn_like(p)).sum() for p in ps[:-1]).backward()     # last tensor: grad None
            opt.step()
    for p, q in zip(pa, pb):
        assert p.numel() == 0 or _rel(q, p) <= 1e-6
    assert torch.equal(pa[-1], pb[-1])


def test_train_step_with_fused_adam_matches_golden(golden_dir):
    """train_step (zero_grad -> fused solve -> MSE -> fused backward -> Adam) with the fused optimiser reproduces
    the reference's own 3-step trace."""
    import nlbac_b200
    g = torch.load(os.path.join(golden_dir, "node_U.pt"), weights_only=False)
    m = nlbac_b200.NeuralODEModel(3, 3, 6)
    m.load_state_dict(g["state_dict"])
    m.cuda()
    opt = nlbac_b200.FusedAdam(m.parameters(), lr=1e-3)
    y0, tr = g["y0"].cuda(), g["train"]
    for want in tr["losses"]:
        got = nlbac_b200.train_step(m, y0[:, :3], y0[:, 3:5], tr["next_state"].cuda(), opt, torch.nn.MSELoss(),
                                    tr["horizon"], "euler", 0.02)
        assert abs(got - want) <= 1e-5 * abs(want) + 1e-9
    for k, v in m.state_dict().items():
        assert _rel(v, tr["state_dict_after"][k]) <= 1e-4, k


@pytest.mark.parametrize("shape", ["U", "C"])
def test_graphed_train_step_matches_eager(shape):
    """GraphedTrainStep (one CUDA-graph replay per model update) vs the eager train_step with FusedAdam: same
    kernels, bias corrections formed on the device instead of the host -> rel 1e-6 after 6 updates."""
    import nlbac_b200
    torch.manual_seed(0)
    if shape == "U":
        mk, n_s, n_u, has_t, B = (lambda: nlbac_b200.NeuralODEModel(3, 3, 6)), 3, 2, False, 128
    else:
        mk, n_s, n_u, has_t, B = (lambda: nlbac_b200.NeuralODEModel(12, 10)), 10, 1, True, 256
    a, b = mk().cuda(), mk().cuda()
    b.load_state_dict(a.state_dict())
    opt = nlbac_b200.FusedAdam(a.parameters(), lr=1e-3)
    step = nlbac_b200.GraphedTrainStep(b, B, n_s, n_u, solver="euler", time_interval=0.02, lr=1e-3, has_time=has_t)
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.equal(p, q)                      # warm-up left no trace
    loss_fn = torch.nn.MSELoss()
    for it in range(6):
        g = torch.Generator(device="cuda").manual_seed(it)
        st, ac = torch.randn(B, n_s, generator=g, device="cuda"), torch.randn(B, n_u, generator=g, device="cuda")
        nx = st + 0.02 * torch.randn(B, n_s, generator=g, device="cuda")
        tb = torch.rand(B, 1, generator=g, device="cuda") if has_t else None
        if has_t:
            want = nlbac_b200.train_step(a, st, ac, nx, tb, opt, loss_fn, B, "euler", 0.02)
        else:
            want = nlbac_b200.train_step(a, st, ac, nx, opt, loss_fn, B, "euler", 0.02)
        got = step(st, ac, nx, tb)
        assert abs(got.item() / B - want) <= 1e-6 * abs(want) + 1e-12
    for p, q in zip(a.parameters(), b.parameters()):
        assert _rel(q, p) <= 1e-6
    # an eager solve after the replays sees the updated weights (packed-image cache invalidated)
    y = torch.randn(B, n_s + n_u + (1 if has_t else 0), device="cuda")
    sa = nlbac_b200.odeint(a, y, [0.0, 0.02], method="euler")
    sb = nlbac_b200.odeint(b, y, [0.0, 0.02], method="euler")
    assert _rel(sb, sa) <= 1e-6


def test_critic_update_captured_in_a_cuda_graph():
    """Twin-Q forward + MSE + backward + FusedAdam(capturable=True) + Polyak update as ONE CUDA graph; three replays
    against the same three updates done eagerly with torch.optim.Adam on a reference copy (rel 1e-6)."""
    import nlbac_b200
    torch.manual_seed(0)
    q, qt = nlbac_b200.QNetwork(7, 2, 256).cuda(), nlbac_b200.QNetwork(7, 2, 256).cuda()
    r, rt = nlbac_ref.QNetwork(7, 2, 256).cuda(), nlbac_ref.QNetwork(7, 2, 256).cuda()
    r.load_state_dict(q.state_dict()); rt.load_state_dict(qt.state_dict())
    opt = nlbac_b200.FusedAdam(q.parameters(), lr=4e-4, capturable=True)
    ropt = torch.optim.Adam(r.parameters(), lr=4e-4)
    obs, act, tgt = (torch.zeros(128, 7, device="cuda"), torch.zeros(128, 2, device="cuda"),
                     torch.zeros(128, 1, device="cuda"))

    def update():
        q1, q2 = q(obs, act)
        loss = torch.nn.functional.mse_loss(q1, tgt) + torch.nn.functional.mse_loss(q2, tgt)
        grads = torch.autograd.grad(loss, list(q.parameters()))
        for p, g_ in zip(q.parameters(), grads):
            p.grad = g_
        opt.step()
        nlbac_b200.soft_update(qt, q, 0.005)
        return loss

    snap = [p.detach().clone() for p in list(q.parameters()) + list(qt.parameters())]
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        update()
    torch.cuda.current_stream().wait_stream(s)
    with torch.no_grad():
        for p, v in zip(list(q.parameters()) + list(qt.parameters()), snap):
            p.copy_(v)
        for st in opt.state.values():
            st["exp_avg"].zero_(); st["exp_avg_sq"].zero_()
        for v in opt._step_dev.values():
            v.zero_()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        loss = update()
    for it in range(3):
        g = torch.Generator(device="cuda").manual_seed(it)
        o, a_, t_ = (torch.randn(128, 7, generator=g, device="cuda"), torch.randn(128, 2, generator=g, device="cuda"),
                     torch.randn(128, 1, generator=g, device="cuda"))
        obs.copy_(o); act.copy_(a_); tgt.copy_(t_)
        graph.replay()
        ropt.zero_grad()
        r1, r2 = r(o, a_)
        rl = torch.nn.functional.mse_loss(r1, t_) + torch.nn.functional.mse_loss(r2, t_)
        rl.backward()
        ropt.step()
        with torch.no_grad():
            for tp, sp in zip(rt.parameters(), r.parameters()):
                tp.copy_(tp * (1.0 - 0.005) + sp * 0.005)
        assert _rel(loss, rl) <= 1e-5
    # Adam divides by sqrt(v): entries whose gradient is ~0 turn the 1e-6 difference between the fused MLP backward
    # and cuBLAS into an O(lr) difference of the step; 3 steps x lr 4e-4 against |p| ~ 0.1 -> bar 1e-4 (measured 2e-5)
    for p, w in zip(q.parameters(), r.parameters()):
        assert _rel(p, w) <= 1e-4
    for p, w in zip(qt.parameters(), rt.parameters()):
        assert _rel(p, w) <= 1e-5


def test_device_side_lagrangian_and_graphed_actor_update(golden_dir):
    """augmented_lagrangian_device (multipliers / rho / ratio on the device) against the host-float version over ten
    updates, then the whole Unicycle CBF/CLF actor-loss update (NODE step, Lyapunov net, constraint kernel,
    Lagrangian, backward to the action, Adam on a stand-in actor) captured in ONE CUDA graph and replayed."""
    import nlbac_b200
    C = nlbac_b200.constraints
    g = torch.load(os.path.join(golden_dir, "constraints_U.pt"), weights_only=False)
    nets = build(g, nlbac_b200, "cuda")
    k, B = g["consts"], g["B"]
    state, V = g["state"].cuda(), g["V"].cuda()
    kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards_locations=k["hazards"], hazards_radius=k["hazards_radius"],
              l_p=k["l_p"], batch_size=B)
    host, dev = C.ConstraintState(8, B), C.DeviceConstraintState(8, B, "cuda")
    for updates in range(6, 16):
        act = (g["action"].cuda() * (1.0 + 0.05 * (updates - 6))).requires_grad_(True)
        req = C.unicycle_required_matrix(nets["node"], nets["lya"], state, act, V, **kw)
        want = C.augmented_lagrangian(req, host, updates)
        dev.set_updates(updates)
        got = C.augmented_lagrangian_device(req, dev)
        assert _rel(got, want) <= 1e-6
        gw, = torch.autograd.grad(want, act, retain_graph=True)
        gg, = torch.autograd.grad(got, act)
        assert _rel(gg, gw) <= 1e-6
    lam = torch.tensor([float(v) for v in host.lambda_values], dtype=torch.float64)
    assert _rel(dev.lambda_values, lam) <= 1e-6 and abs(dev.augmented_term.item() - host.augmented_term) < 1e-12

    # --- the same update as one CUDA graph: the actor is a stand-in linear map obs -> action with capturable Adam
    torch.manual_seed(0)
    actor = torch.nn.Linear(3, 2).cuda()
    opt = nlbac_b200.FusedAdam(actor.parameters(), lr=3e-4, capturable=True)
    ref_actor = torch.nn.Linear(3, 2).cuda()
    ref_actor.load_state_dict(actor.state_dict())
    ref_opt = torch.optim.Adam(ref_actor.parameters(), lr=3e-4)
    st_g, st_h = C.DeviceConstraintState(8, B, "cuda"), C.ConstraintState(8, B)
    s_in = torch.zeros_like(state)

    def update():
        pi = actor(s_in)
        req = C.unicycle_required_matrix(nets["node"], nets["lya"], s_in, pi, V, **kw)
        loss = C.augmented_lagrangian_device(req, st_g)
        grads = torch.autograd.grad(loss, list(actor.parameters()))
        for p, g_ in zip(actor.parameters(), grads):
            p.grad = g_
        opt.step()
        return loss

    snap = [p.detach().clone() for p in actor.parameters()]
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        s_in.copy_(state)
        update()
    torch.cuda.current_stream().wait_stream(side)
    with torch.no_grad():
        for p, v in zip(actor.parameters(), snap):
            p.copy_(v)
        for s_ in opt.state.values():
            s_["exp_avg"].zero_(); s_["exp_avg_sq"].zero_()
        for v in opt._step_dev.values():
            v.zero_()
        st_g.lambda_values.zero_(); st_g.augmented_term.fill_(1.0)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        loss_g = update()
    for updates in range(8, 12):
        s_now = state + 0.01 * (updates - 8)
        s_in.copy_(s_now)
        st_g.set_updates(updates)
        graph.replay()
        ref_opt.zero_grad()
        req = C.unicycle_required_matrix(nets["node"], nets["lya"], s_now, ref_actor(s_now), V, **kw)
        want = C.augmented_lagrangian(req, st_h, updates)
        want.backward()
        ref_opt.step()
        assert _rel(loss_g, want) <= 1e-5
    for p, w in zip(actor.parameters(), ref_actor.parameters()):
        assert _rel(p, w) <= 1e-4


def test_config4_quadrotor_shaped_horizon10_rollout():
    """BASELINE config 4: Quadrotor-shaped black-box NODE (8 -> 64 -> 64 -> 64 -> 6, hidden 64 ASSUMED: the source is a
    missing submodule), 10 chained Euler predictions with the policy re-sampled on the detached prediction between steps
    (eps[9, 256, 2] injected), B = 256: states and gradients to the initial action / the NODE parameters vs the oracle.
    Tolerance 5e-5: ten chained VJPs."""
    import nlbac_b200
    torch.manual_seed(0)
    ref_node, ref_pol = nlbac_ref.BlackBoxNODE(8, 6), nlbac_ref.GaussianPolicy(6, 2, 256)
    node, pol = nlbac_b200.NeuralODEModel(8, 6), nlbac_b200.GaussianPolicy(6, 2, 256)
    node.load_state_dict(ref_node.state_dict()); pol.load_state_dict(ref_pol.state_dict())
    node, pol = node.cuda(), pol.cuda()
    g = torch.Generator().manual_seed(4)
    B, H = 256, 10
    x0, eps = torch.randn(B, 6, generator=g), torch.randn(H - 1, B, 2, generator=g)
    cot = torch.randn(H, B, 6, generator=g)
    ua = torch.randn(B, 2, generator=g).requires_grad_(True)
    want = OC.chained_rollout(ref_node, ref_pol, x0, ua, H, dt=0.02, eps=eps)
    (want * cot).sum().backward()
    ub = ua.detach().clone().cuda().requires_grad_(True)
    got = nlbac_b200.constraints.chained_rollout(node, pol, x0.cuda(), ub, H, dt=0.02, eps=eps.cuda())
    (got * cot.cuda()).sum().backward()
    assert got.shape == (H, B, 6)
    assert _rel(got, want) <= 2e-5, _rel(got, want)
    assert _rel(ub.grad, ua.grad) <= 5e-5, _rel(ub.grad, ua.grad)
    for p, q in zip(node.parameters(), ref_node.parameters()):
        assert _rel(p.grad, q.grad) <= 5e-5, _rel(p.grad, q.grad)


def test_config3_learned_barrier_block_with_dopri5_inside_a_cuda_graph(golden_dir):
    """BASELINE config 3 as named (Pvtol + learned barrier, dopri5 adaptive solve): the required matrix and its action
    gradient with ``solver="dopri5"``, recorded ONCE into a CUDA graph (the device-side step controller needs no host
    read-back) and replayed, against the CPU oracle running the restated torchdiffeq dopri5."""
    import nlbac_b200
    C = nlbac_b200.constraints
    g = torch.load(os.path.join(golden_dir, "constraints_BP.pt"), weights_only=False)
    k, B = g["consts"], g["B"]
    nets, onets = build(g, nlbac_b200, "cuda"), build(g, nlbac_ref, "cpu")
    dyn = nlbac_b200.DynamicsModel(_Env("Pvtol"))
    w = torch.linspace(0.5, 1.5, 2).reshape(2, 1)
    kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], batch_size=B, solver="dopri5")

    act_o = g["action"].clone().requires_grad_(True)
    req_o = OC.barrier_pvtol_required_matrix(onets["node"], onets["lya"], onets["barrier"], onets["policy"], g["obs"],
                                             g["state"], act_o, g["V"], follow=k["safety_operator_follow"],
                                             eps_next=g["eps"], **kw)
    (req_o * w).sum().backward()

    obs, state, V, eps = (g[n].cuda() for n in ("obs", "state", "V", "eps"))
    act = torch.zeros_like(g["action"]).cuda().requires_grad_(True)
    wc = w.cuda()

    def run():
        req = C.barrier_pvtol_required_matrix(nets["node"], nets["lya"], nets["barrier"], nets["policy"], dyn, obs, state,
                                              act, V, safety_operator_follow=k["safety_operator_follow"], eps_next=eps,
                                              **kw)
        (ga,) = torch.autograd.grad((req * wc).sum(), act)
        return req, ga

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):
            run()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        req_g, ga_g = run()
    with torch.no_grad():
        act.copy_(g["action"].cuda())          # the graph was recorded on zeros: the replay must see the new input
    graph.replay()
    assert _rel(req_g, req_o) <= 2 * TOL
    assert _rel(ga_g, act_o.grad) <= 10 * TOL


@pytest.mark.parametrize("B", [1, 128, 256, 1000])
def test_scalar_glue_kernels_match_torch_expressions(B):
    """nlbac_td_target / nlbac_mse_mean_* / nlbac_sac_actor_term_* against the reference's own expressions
    (U sac_cbf_clf.py:234-238, :244-246, :270), forward and backward, incl. an exact tie in min(q1, q2)."""
    from nlbac_b200 import tail
    g = torch.Generator().manual_seed(B)
    r, q1, q2, lp = (torch.randn(B, 1, generator=g).cuda() for _ in range(4))
    mask = (torch.rand(B, 1, generator=g) > 0.1).float().cuda()
    alpha = torch.tensor([0.2], device="cuda")
    q2[0] = q1[0]                                                   # a tie
    want = r + mask * 0.99 * (torch.min(q1, q2) - alpha * lp)
    assert torch.equal(tail.td_target(r, mask, 0.99, q1, q2, lp, alpha), want)
    assert torch.equal(tail.td_target(r, mask, 0.99, q1), r + mask * 0.99 * q1)
    # mse mean
    p1 = q1.clone().requires_grad_(True)
    p2 = q1.clone().requires_grad_(True)
    l1, l2 = tail.mse_mean(p1, want), torch.nn.functional.mse_loss(p2, want)
    (3.0 * l1).backward()
    (3.0 * l2).backward()
    assert l1.shape == l2.shape and _rel(l1, l2) <= 1e-6 and _rel(p1.grad, p2.grad) <= 1e-6
    # SAC actor term
    a = [t.clone().requires_grad_(True) for t in (lp, q1, q2)]
    b = [t.clone().requires_grad_(True) for t in (lp, q1, q2)]
    s1 = tail.sac_actor_term(a[0], a[1], a[2], alpha)
    s2 = (alpha * b[0] - torch.min(b[1], b[2])).mean()
    (2.0 * s1).backward()
    (2.0 * s2).backward()
    assert _rel(s1, s2) <= 1e-6
    for x, y in zip(a, b):
        assert _rel(x.grad, y.grad) <= 1e-6
