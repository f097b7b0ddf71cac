"""CPU restatement of ``SAC_CBF_CLF.update_parameters`` for the Unicycle variant (U sac_cbf_clf/sac_cbf_clf.py:181-319).

TEST INFRASTRUCTURE (see oracle/__init__.py): the eager-PyTorch checker / baseline of ``nlbac_b200.GraphedUpdate``.
Pinned to a 12-update trace of the reference's OWN method (tests/golden/make_golden_update.py -> update_U.pt ->
tests/test_oracle_update.py).  Device-agnostic: bench.py also runs it eagerly on the B200 as the like-for-like bar.
"""
import torch
import torch.nn.functional as F

from . import constraints as OC
from . import nlbac_ref


class UnicycleAgent:
    """The attributes ``SAC_CBF_CLF.__init__`` sets (U sac_cbf_clf.py:28-134) on restated network classes."""

    def __init__(self, consts, batch_size, init=None, device="cpu"):
        k, dev = consts, torch.device(device)
        hi, lo = k["action_high"], k["action_low"]
        self.k, self.B, self.device = k, batch_size, dev
        self.critic = nlbac_ref.QNetwork(7, 2, 256)
        self.critic_target = nlbac_ref.QNetwork(7, 2, 256)
        self.lyapunovNet = nlbac_ref.LyaNetwork(2, 256)
        self.lyapunovNet_target = nlbac_ref.LyaNetwork(2, 256)
        self.policy = nlbac_ref.GaussianPolicy(7, 2, 256, hi, lo)
        self.backup_policy = nlbac_ref.GaussianPolicy(7, 2, 256, hi, lo)
        self.neural_ode_model = nlbac_ref.AffineNODE(3, 3, 6)
        if init is not None:
            for n in ("critic", "lyapunovNet", "policy", "backup_policy", "neural_ode_model"):
                getattr(self, n).load_state_dict(init[n])
        self.critic_target.load_state_dict(self.critic.state_dict())                  # hard_update, :56, :58
        self.lyapunovNet_target.load_state_dict(self.lyapunovNet.state_dict())
        for n in ("critic", "critic_target", "lyapunovNet", "lyapunovNet_target", "policy", "backup_policy",
                  "neural_ode_model"):
            getattr(self, n).to(dev)
        for pol in (self.policy, self.backup_policy):       # plain tensor attributes (the reference overrides .to(), U model.py:130-133)
            pol.action_scale, pol.action_bias = pol.action_scale.to(dev), pol.action_bias.to(dev)
        A = torch.optim.Adam
        self.critic_optim = A(self.critic.parameters(), lr=k["critic_lr"])            # :48
        self.lyaNet_optim = A(self.lyapunovNet.parameters(), lr=k["critic_lr"])       # :51
        self.policy_optim = A(self.policy.parameters(), lr=k["lr"])                   # :86
        self.backup_policy_optim = A(self.backup_policy.parameters(), lr=k["lr"])     # :90
        self.log_alpha = torch.zeros(1, requires_grad=True, device=dev)               # :78-82
        self.backup_log_alpha = torch.zeros(1, requires_grad=True, device=dev)
        self.alpha_optim, self.backup_alpha_optim = A([self.log_alpha], lr=k["lr"]), A([self.backup_log_alpha], lr=k["lr"])
        self.node_optim = A(self.neural_ode_model.parameters(), lr=k["node_lr"])      # :133
        self.alpha = self.backup_alpha = k["alpha"]
        self.cs = OC.ConstraintState(8)
        self.kw = dict(dt=k["dt"], gamma_b=k["gamma_b"], hazards=k["hazards"], hazards_radius=k["hazards_radius"],
                       l_p=k["l_p"], batch_size=batch_size)

    def update_parameters(self, batch, updates, node_batch=None, eps=None):
        """``batch`` / ``node_batch``: the sampled tuples as tensors; ``eps``: the three N(0,1) draws of
        ``policy.sample`` in call order (next state, pi, backup pi) or None."""
        k, dev = self.k, self.device
        eps = eps or [None, None, None]
        env = "Unicycle"
        if node_batch is not None:                                                                       # :205-219
            s0, s1 = OC.get_state(env, node_batch[0].to(dev)), OC.get_state(env, node_batch[6].to(dev))
            nlbac_ref.train_step(self.neural_ode_model, s0, node_batch[1].to(dev), s1, self.node_optim,
                                 torch.nn.MSELoss(), s0.shape[0], "euler", k["dt"])
        state, action, reward, cons, cp, ncp, nstate, mask = [b.to(dev) for b in batch[:8]]
        reward, cons, mask = reward.reshape(-1, 1), cons.reshape(-1, 1), mask.reshape(-1, 1)             # :225-229
        with torch.no_grad():                                                                            # :231-238
            na, nlogp, _ = self.policy.sample(nstate, eps[0])
            q1t, q2t = self.critic_target(nstate, na)
            next_q = reward + mask * k["gamma"] * (torch.min(q1t, q2t) - self.alpha * nlogp)
            next_l = cons + mask * k["gamma"] * self.lyapunovNet_target(ncp)
        qf1, qf2 = self.critic(state, action)                                                            # :240-246
        qf1_loss, qf2_loss = F.mse_loss(qf1, next_q), F.mse_loss(qf2, next_q)
        lf_loss = F.mse_loss(self.lyapunovNet(cp), next_l)
        self.critic_optim.zero_grad(); (qf1_loss + qf2_loss).backward(); self.critic_optim.step()        # :249-251
        self.lyaNet_optim.zero_grad(); lf_loss.backward(); self.lyaNet_optim.step()                      # :253-255
        pi, log_pi, _ = self.policy.sample(state, eps[1])                                                # :258-267
        min_qf_pi = torch.min(*self.critic(state, pi))
        bpi, blog_pi, _ = self.backup_policy.sample(state, eps[2])
        bmin_qf_pi = torch.min(*self.critic(state, bpi))
        policy_loss_1 = (self.alpha * log_pi - min_qf_pi).mean()                                         # :270-273
        backup_loss_1 = (self.backup_alpha * blog_pi - bmin_qf_pi).mean()
        x = OC.get_state(env, state)                                                                     # :380-384
        V = self.lyapunovNet(cp).detach()
        req = OC.unicycle_required_matrix(self.neural_ode_model, self.lyapunovNet, x, pi, V, **self.kw)
        policy_loss_2 = OC.augmented_lagrangian(req, self.cs, updates)
        breq = OC.unicycle_required_matrix(self.neural_ode_model, None, x, bpi, None, **self.kw)         # :279
        backup_loss_2 = OC.backup_augmented_lagrangian(breq, self.cs, updates)
        self.policy_optim.zero_grad(); (policy_loss_1 + policy_loss_2).backward(); self.policy_optim.step()
        self.backup_policy_optim.zero_grad(); (backup_loss_1 + backup_loss_2).backward(); self.backup_policy_optim.step()
        alpha_loss = -(self.log_alpha * (log_pi + k["target_entropy"]).detach()).mean()                  # :292-308
        self.alpha_optim.zero_grad(); alpha_loss.backward(); self.alpha_optim.step()
        self.alpha = self.log_alpha.exp()
        balpha_loss = -(self.backup_log_alpha * (blog_pi + k["target_entropy"]).detach()).mean()
        self.backup_alpha_optim.zero_grad(); balpha_loss.backward(); self.backup_alpha_optim.step()
        self.backup_alpha = self.backup_log_alpha.exp()
        with torch.no_grad():                                                                            # :315-317
            for tgt, src in ((self.critic_target, self.critic), (self.lyapunovNet_target, self.lyapunovNet)):
                for tp, sp in zip(tgt.parameters(), src.parameters()):
                    tp.copy_(tp * (1.0 - k["tau"]) + sp * k["tau"])
        return (qf1_loss.item(), qf2_loss.item(), lf_loss.item(), policy_loss_1.item(), alpha_loss.item(),
                self.alpha.item())
