"""``GraphedUpdate`` — one minibatch update of ``SAC_CBF_CLF.update_parameters`` as ONE CUDA-graph replay.

Mirrors U ``sac_cbf_clf/sac_cbf_clf.py:181-319`` (C :177-319 with the cars constraint block): TD targets, critic +
Lyapunov updates, actor and backup-actor losses = SAC term + CBF/CLF augmented Lagrangian (``get_cbf_clf_part`` /
``backup_get_cbf_clf_part``, :364-406 -> ``constraints.*_required_matrix``), Adam steps, the temperature updates
(:292-308), Polyak averaging (:315-317).  Everything the reference keeps as Python floats -- multipliers, rho, the
CBF/CLF ratio, alpha, the update counter -- lives on the device, so the six ``.item()`` of :319 and the ``float()``
read-backs of :510, :525-528 disappear: the update is captured once and replayed per call with zero host
synchronisation.  The NODE model fit (:205-219, every ``NODE_model_update_interval`` updates) is a second graph
(``GraphedTrainStep``) once the NODE batch has reached its steady size, the eager fused ``train_step`` before that.
"""
import torch

from . import constraints as C
from .odeint import odeint
from .tail import FusedAdam, mse_mean, sac_actor_term, soft_update, td_target


class GraphedUpdate:
    """``upd = GraphedUpdate(critic, critic_target, lyapunovNet, lyapunovNet_target, policy, backup_policy,
    neural_ode_model, dynamics_model, batch_size, primary=fn, backup=fn, ...)``

    ``primary(state, pi, V) -> required_matrix [m, 1]`` and ``backup(state, backup_pi) -> [m - 1, 1]`` are the bodies
    of ``get_policy_loss_2`` / ``backup_get_policy_loss_2`` up to the constraint sums (e.g. closures over
    ``constraints.unicycle_required_matrix``); the augmented Lagrangians are applied here on the device.
    ``update_parameters(memory, batch_size, updates, dynamics_model, NODE_memory, NODE_model_update_interval)`` keeps the
    reference's signature; it returns a CUDA tensor with the reference's six values (qf1, qf2, lf, policy_loss_1,
    alpha_loss, alpha) instead of six Python floats -- ``.tolist()`` it when the floats are wanted.
    """

    def __init__(self, critic, critic_target, lyapunovNet, lyapunovNet_target, policy, backup_policy, neural_ode_model,
                 dynamics_model, batch_size, *, primary, backup, num_constraints, obs_dim, act_dim, center_pos_num,
                 gamma=0.99, tau=0.005, alpha=0.2, lr=3e-4, critic_lyapunov_lr=4e-4, node_lr=1e-3,
                 target_update_interval=1, automatic_entropy_tuning=True, multiplier_interval=8, lambda_max=400.0,
                 ratio_floor=None, use_ratio=True, backup_update_interval=1, separate_backup_rho=False,
                 node_solver="euler", dt=0.02, node_n_s=3, node_n_u=2, node_graph_batch=32768, with_time=False,
                 n_extra_eps=0, lya_on_obs=False, node_fit_max_episode=None, barrier=None, barrier_target=None,
                 concurrent=True):
        dev = next(critic.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("GraphedUpdate: the networks must live on a CUDA device (no CPU fallback)")
        self.device, self.B = dev, int(batch_size)
        self.critic, self.critic_target = critic, critic_target
        self.lyapunovNet, self.lyapunovNet_target = lyapunovNet, lyapunovNet_target
        self.policy, self.backup_policy = policy, backup_policy      # backup_policy None: no backup controller (BU / BP)
        # learned-barrier variants (BU sac_cbf_clf.py:181-186, 214-216, 225-227, 237-240, 277): a barrier network with its
        # own target, TD loss on the replay's barrier_signal field (11-field layout), optimiser and Polyak update
        self.barrier, self.barrier_target = barrier, barrier_target
        self.neural_ode_model, self.dynamics_model = neural_ode_model, dynamics_model
        self.primary, self.backup = primary, backup
        self.gamma, self.tau, self.target_update_interval = gamma, tau, target_update_interval
        self.automatic_entropy_tuning = automatic_entropy_tuning
        self.target_entropy = -float(act_dim)                                   # U sac_cbf_clf.py:77
        f32 = dict(dtype=torch.float32, device=dev)
        self.alpha = torch.full((1,), float(alpha), **f32)                      # python float in the reference
        self.backup_alpha = torch.full((1,), float(alpha), **f32)
        self.log_alpha = torch.zeros(1, requires_grad=True, **f32)
        self.backup_log_alpha = torch.zeros(1, requires_grad=True, **f32)
        mk = lambda ps, lr_: FusedAdam(ps, lr=lr_, capturable=True)             # noqa: E731
        self.critic_optim = mk(critic.parameters(), critic_lyapunov_lr)
        self.lyaNet_optim = mk(lyapunovNet.parameters(), critic_lyapunov_lr)
        self.policy_optim = mk(policy.parameters(), lr)
        self.backup_policy_optim = mk(backup_policy.parameters(), lr) if backup_policy is not None else None
        self.barrier_optim = mk(barrier.parameters(), critic_lyapunov_lr) if barrier is not None else None
        self.alpha_optim = mk([self.log_alpha], lr)
        self.backup_alpha_optim = mk([self.backup_log_alpha], lr)
        self.cstate = C.DeviceConstraintState(num_constraints, batch_size, dev, lambda_max=lambda_max,
                                              update_interval=multiplier_interval, ratio_floor=ratio_floor,
                                              use_ratio=use_ratio, backup_update_interval=backup_update_interval,
                                              separate_backup_rho=separate_backup_rho)
        self.updates_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        B = self.B
        z = lambda *s: torch.zeros(*s, **f32)                                   # noqa: E731
        self.buf = {"state": z(B, obs_dim), "action": z(B, act_dim), "reward": z(B, 1), "constraint": z(B, 1),
                    "center_pos": z(B, center_pos_num), "next_center_pos": z(B, center_pos_num),
                    "next_state": z(B, obs_dim), "mask": z(B, 1)}
        # SimulatedCars (C sac_cbf_clf.py:49-51, :98-102): the time columns of the minibatch reach the constraint blocks, and
        # those re-sample the policies inside their rollouts -> `with_time` adds t / next_t buffers that are passed to the
        # closures as primary(state, pi, V, t, next_t) / backup(state, pi, t, next_t); `n_extra_eps` adds noise buffers
        # self.eps[3:] for the closures' policy.sample calls (in the reference's call order)
        # Pvtol (P sac_cbf_clf.py:239-247, :282-300, :205): the Lyapunov TD loss is on the observations (V(x_t) still on the
        # replay's lya_pre_term field = the center_pos slot), the backup policy / its temperature are updated only every
        # `backup_update_interval` updates -> TWO captured graphs (with / without the backup branch) share all state and
        # the host picks one per call from `updates`; the NODE fit stops after episode `node_fit_max_episode`
        self.lya_on_obs, self.backup_update_interval = bool(lya_on_obs), int(backup_update_interval)
        self.node_fit_max_episode = node_fit_max_episode
        self.graph_nobackup = None
        self.with_time = bool(with_time)
        if self.with_time:
            self.buf["t"], self.buf["next_t"] = z(B, 1), z(B, 1)
        if barrier is not None:
            self.buf["barrier_signal"] = z(B, 1)
        self.eps = [z(B, act_dim) for _ in range(3 + n_extra_eps)]              # next-state sample, pi, backup pi, extras
        self.out = z(6)
        # NODE fit (its own optimiser state; eager fused step until the batch reaches node_graph_batch, a graph after)
        self.node_optim = FusedAdam(neural_ode_model.parameters(), lr=node_lr, capturable=True)
        self.node_lr, self.node_solver, self.dt = node_lr, node_solver, dt
        self.node_n_s, self.node_n_u, self.node_graph_batch = node_n_s, node_n_u, node_graph_batch
        self._node_graph = None
        self.graph = None
        self.concurrent = bool(concurrent)
        self._streams = tuple(torch.cuda.Stream(device=dev) for _ in range(4))
        if self.concurrent:
            critic.twin_streams = critic_target.twin_streams = True

    # ------------------------------------------------------------------------------------------------
    @staticmethod
    def _step(optim, loss, params):
        grads = torch.autograd.grad(loss, params, retain_graph=True)
        for p, g in zip(params, grads):
            p.grad = g
        optim.step()

    def _body(self, with_backup=True):
        """U sac_cbf_clf.py:221-319, statement for statement, on the static buffers (P :221-331 with `lya_on_obs` and, on
        the updates that skip the backup controller, `with_backup=False`)."""
        b = self.buf
        st = self.cstate
        st.set_updates_device(self.updates_dev)
        # The statements of the reference's update, issued as a dependency graph instead of a sequence (captured, the
        # forked streams become parallel branches of the CUDA graph; at B = 128 / 256 a kernel fills < half of the SMs
        # and every launch gap of one branch hides behind another).  What depends on what in :221-319:
        #   main  TD targets -> critic update -> SAC term of the actor loss (needs the UPDATED critic) -> actor step
        #   sA    Lyapunov update -> V, pi = policy.sample, the actor's Lagrangian term (NODE rollout + constraint sums:
        #         needs the updated Lyapunov / barrier nets, NOT the critic)
        #   sB    backup actor: sample + NODE rollout need nothing that this update writes; its SAC term waits for the
        #         critic update, its Lagrangian for the primary one (U / C share rho)
        #   sC    barrier-certificate update (BU / BP)
        # `concurrent=False` issues the same statements on one stream.
        main = torch.cuda.current_stream()
        sA, sB, sC, _ = (self._streams if self.concurrent else (main,) * 4)
        forked = []

        def fork(s):
            if s is not main:
                s.wait_stream(main)
                forked.append(s)
            return torch.cuda.stream(s)

        def after(s, *deps):        # stream s continues only once the work issued so far on deps is done
            for d in deps:
                if d is not s:
                    s.wait_stream(d)
            return torch.cuda.stream(s)

        with_backup = with_backup and self.backup_policy is not None
        targs = (b["t"], b["next_t"]) if self.with_time else ()
        state = self.dynamics_model.get_state(b["state"])                                        # :380 (get_cbf_clf_part)
        if isinstance(state, tuple):
            state = state[0]
        with fork(sA):                                                                           # :236-238, :244, :253-255
            with torch.no_grad():
                next_l = td_target(b["constraint"], b["mask"], self.gamma, self.lyapunovNet_target(
                    b["next_state"] if self.lya_on_obs else b["next_center_pos"]))
            lf_loss = mse_mean(self.lyapunovNet(b["state"] if self.lya_on_obs else b["center_pos"]), next_l)
            self._step(self.lyaNet_optim, lf_loss, list(self.lyapunovNet.parameters()))
        if with_backup:
            with fork(sB):
                bpi, blog_pi, _ = self.backup_policy.sample(b["state"], self.eps[2])             # :264-267
                backup_req = self.backup(state, bpi, *targs)                                     # :279 -> :532-608
        with torch.no_grad():                                                                    # :231-235
            next_a, next_logp, _ = self.policy.sample(b["next_state"], self.eps[0])
            if self.barrier is not None:
                with fork(sC):                                                                   # BU :214-216, :225-227, :237-240
                    next_b = td_target(b["barrier_signal"], b["mask"], self.gamma,
                                       self.barrier_target(b["next_state"], next_a))
            q1t, q2t = self.critic_target(b["next_state"], next_a)
            next_q = td_target(b["reward"], b["mask"], self.gamma, q1t, q2t, next_logp, self.alpha)
        if self.barrier is not None:
            with torch.cuda.stream(sC):
                barrier_loss = mse_mean(self.barrier(b["state"], b["action"]), next_b)
                self._step(self.barrier_optim, barrier_loss, list(self.barrier.parameters()))
        with after(sA, *([sC] if self.barrier is not None else [])):
            V = self.lyapunovNet(b["center_pos"]).detach()                                       # :382-383
            pi, log_pi, _ = self.policy.sample(b["state"], self.eps[1])                          # :258-261
            policy_loss_2 = C.augmented_lagrangian_device(self.primary(state, pi, V, *targs), st)    # :384 -> :408-530
        qf1, qf2 = self.critic(b["state"], b["action"])                                          # :240-246
        qf1_loss, qf2_loss = mse_mean(qf1, next_q), mse_mean(qf2, next_q)
        self._step(self.critic_optim, qf1_loss + qf2_loss, list(self.critic.parameters()))       # :249-251
        if with_backup:
            with after(sB, main, sA):        # the updated critic (main) and the primary Lagrangian's rho (sA)
                backup_policy_loss_1 = sac_actor_term(blog_pi, *self.critic(b["state"], bpi), self.backup_alpha)  # :273
                backup_loss_2 = C.backup_augmented_lagrangian_device(backup_req, st)             # :611-640
                self._step(self.backup_policy_optim, backup_policy_loss_1 + backup_loss_2,
                           list(self.backup_policy.parameters()))                                # :288-290
                if self.automatic_entropy_tuning:                                                # :302-308
                    balpha_loss = -(self.backup_log_alpha * (blog_pi + self.target_entropy).detach()).mean()
                    self._step(self.backup_alpha_optim, balpha_loss, [self.backup_log_alpha])
                    with torch.no_grad():
                        self.backup_alpha.copy_(self.backup_log_alpha.exp())
        with after(main, sA):
            policy_loss_1 = sac_actor_term(log_pi, *self.critic(b["state"], pi), self.alpha)     # :270
            self._step(self.policy_optim, policy_loss_1 + policy_loss_2, list(self.policy.parameters()))     # :284-286
            if self.automatic_entropy_tuning:                                                    # :292-300
                alpha_loss = -(self.log_alpha * (log_pi + self.target_entropy).detach()).mean()
                self._step(self.alpha_optim, alpha_loss, [self.log_alpha])
                with torch.no_grad():
                    self.alpha.copy_(self.log_alpha.exp())
            else:
                alpha_loss = torch.zeros((), device=self.device)
        for s_ in forked:           # join every branch before the Polyak updates and the returned values
            main.wait_stream(s_)
        if self.target_update_interval == 1:                                                     # :315-317
            soft_update(self.critic_target, self.critic, self.tau)
            soft_update(self.lyapunovNet_target, self.lyapunovNet, self.tau)
            if self.barrier is not None:
                soft_update(self.barrier_target, self.barrier, self.tau)                         # BU :277
        with torch.no_grad():
            self.out.copy_(torch.stack([qf1_loss.reshape(()), qf2_loss.reshape(()), lf_loss.reshape(()),
                                        policy_loss_1.reshape(()), alpha_loss.reshape(()), self.alpha.reshape(())]))
            self.updates_dev.add_(1)
        return self.out

    # ------------------------------------------------------------------------------------------------
    def _nets(self):
        return [n for n in (self.critic, self.critic_target, self.lyapunovNet, self.lyapunovNet_target, self.policy,
                            self.backup_policy, self.barrier, self.barrier_target) if n is not None]

    def _state_tensors(self):
        ts = []
        for net in self._nets():
            ts += list(net.parameters())
        ts += [self.log_alpha, self.backup_log_alpha, self.alpha, self.backup_alpha, self.updates_dev,
               self.cstate.lambda_values, self.cstate.backup_lambda_values, self.cstate.augmented_term,
               self.cstate.backup_augmented_term, self.cstate.due, self.cstate.backup_due]
        return ts

    def _optims(self):
        return [o for o in (self.critic_optim, self.lyaNet_optim, self.policy_optim, self.backup_policy_optim,
                            self.alpha_optim, self.backup_alpha_optim, self.barrier_optim) if o is not None]

    def capture(self):
        """Warm-up on a side stream (lazy kernel attributes, allocator, Adam state), restore every tensor the update
        mutates, then capture ``_body`` as one CUDA graph."""
        snap = [t.detach().clone() for t in self._state_tensors()]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            self._body(True)
            if self.backup_update_interval > 1:
                self._body(False)
        torch.cuda.current_stream().wait_stream(side)
        with torch.no_grad():
            for t, v in zip(self._state_tensors(), snap):
                t.copy_(v)
            for opt in self._optims():
                for s_ in opt.state.values():
                    s_["exp_avg"].zero_()
                    s_["exp_avg_sq"].zero_()
                for v in opt._step_dev.values():
                    v.zero_()
        for net in self._nets():
            for p in net.parameters():
                torch.autograd.graph.increment_version(p)
        torch.cuda.synchronize()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body(True)
        if self.backup_update_interval > 1:      # the updates that skip the backup controller: a second graph, same state
            self.graph_nobackup = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_nobackup):
                self._body(False)
        return self

    def _batch_names(self):
        """Static-buffer names in the order of the reference's sampled tuple."""
        names = ("state", "action", "reward", "constraint", "center_pos", "next_center_pos", "next_state", "mask")
        if self.barrier is not None:     # BU / BP replay layout: barrier_signal after constraint (replay_memory.py:12)
            names = names[:4] + ("barrier_signal",) + names[4:]
        if self.with_time:
            names = names + ("t", "next_t")
        return names

    @torch.no_grad()
    def load_batch(self, batch, eps=None):
        """``batch`` = the reference's sampled tuple (state, action, reward, constraint, center_pos, next_center_pos,
        next_state, mask, ...) as CUDA tensors or host arrays; ``eps`` = three [B, n_u] noise tensors or None (drawn
        from torch's CUDA generator, like ``Normal.rsample`` in the reference)."""
        for n, v in zip(self._batch_names(), batch):
            t = v if torch.is_tensor(v) else torch.as_tensor(v, dtype=torch.float32)
            self.buf[n].copy_(t.reshape(self.buf[n].shape), non_blocking=True)
        self.load_eps(eps)

    @torch.no_grad()
    def load_eps(self, eps=None):
        for i in range(len(self.eps)):
            if eps is None:
                self.eps[i].normal_()
            else:
                self.eps[i].copy_(eps[i], non_blocking=True)

    def step(self, batch=None, eps=None, with_backup=True):
        """One update: load the minibatch into the static buffers and replay the graph.  No host synchronisation."""
        if batch is not None:
            self.load_batch(batch, eps)
        if self.graph is None:
            self.capture()
        (self.graph if (with_backup or self.graph_nobackup is None) else self.graph_nobackup).replay()
        for net in self._nets():
            for p in net.parameters():          # written through raw pointers inside the graph
                torch.autograd.graph.increment_version(p)
        return self.out

    # ------------------------------------------------------------------------------------------------
    def _node_body(self, state, action, next_state, time_batch=None):
        y0 = torch.cat((state, action) if time_batch is None else (state, action, time_batch), dim=-1)
        pred = odeint(self.neural_ode_model, y0, [0.0, self.dt], method=self.node_solver, atol=1e-7, rtol=1e-5)[-1]
        loss = torch.nn.functional.mse_loss(pred[:, :state.shape[1]], next_state)
        self._step(self.node_optim, loss, list(self.neural_ode_model.parameters()))
        return loss

    def node_fit(self, state, action, next_state, time_batch=None):
        """The NODE model fit of U sac_cbf_clf.py:205-219 (``train_step``): eager fused step, or -- once the batch has the
        steady size ``node_graph_batch`` -- one graph replay.  One Adam state (device step counter) serves both."""
        N = state.shape[0]
        srcs = (state, action, next_state) + (() if time_batch is None else (time_batch,))
        if N != self.node_graph_batch:
            return self._node_body(*srcs)
        if self._node_graph is None:
            self._nb = [torch.empty_like(x) for x in srcs]
            for d, s_ in zip(self._nb, srcs):
                d.copy_(s_)
            self._node_body(*self._nb)          # eager: this update itself (also warms the allocator)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._node_loss = self._node_body(*self._nb)
            self._node_graph = g
            return self._node_loss
        for d, s_ in zip(self._nb, srcs):
            d.copy_(s_, non_blocking=True)
        self._node_graph.replay()
        for p in self.neural_ode_model.parameters():
            torch.autograd.graph.increment_version(p)
        return self._node_loss

    def update_parameters(self, memory, batch_size, updates, dynamics_model=None, NODE_memory=None,
                          NODE_model_update_interval=10, i_episode=None, eps=None):
        """The reference's call (U sac_cbf_clf.py:181; P :181 with ``i_episode``) on a device-resident
        ``nlbac_b200.ReplayMemory``."""
        if hasattr(memory, "sample_into") and batch_size == self.B:
            # gather straight into the static inputs of the captured update (same indices as memory.sample)
            memory.sample_into(batch_size, [self.buf[n] for n in self._batch_names()])
            self.load_eps(eps)
            batch = None
        else:
            batch = memory.sample(batch_size=batch_size)
        fit = NODE_memory is not None and updates % NODE_model_update_interval == 0
        if fit and self.node_fit_max_episode is not None and i_episode is not None:
            fit = i_episode <= self.node_fit_max_episode                                  # P :205
        if fit:
            n = min(NODE_memory.position, 32768)
            nb = NODE_memory.sample(batch_size=n)
            dm = dynamics_model or self.dynamics_model
            to_dev = lambda v: v if torch.is_tensor(v) else torch.as_tensor(v, dtype=torch.float32, device=self.device)  # noqa: E731
            off = 1 if self.barrier is not None else 0
            s0, s1 = dm.get_state(to_dev(nb[0])), dm.get_state(to_dev(nb[6 + off]))
            if isinstance(s0, tuple):           # Pvtol: (state7, state6) -- the NODE is fitted on the six dynamic states
                s0, s1 = s0[1], s1[1]
            tb = to_dev(nb[8 + off]).reshape(-1, 1) if self.with_time else None    # C :32: node_t_batch.unsqueeze(1)
            self.node_fit(s0, to_dev(nb[1]), s1, tb)
        with torch.no_grad():
            self.updates_dev.fill_(int(updates))
        return self.step(batch, eps, with_backup=updates % self.backup_update_interval == 0)
