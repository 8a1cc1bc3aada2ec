"""TD3 trainer coupled to the batched engine on the same GPU.

Mirrors scripts/MachineLearning/TD3.py (the class `TD3`: clipped double Q-learning, target policy smoothing, delayed policy
updates, soft target updates, :120-334) and the loop of scripts/rover_training_1.py:101-125 -- trial, extend the replay buffer,
`train(ITER_PER_EP)`, hand the new actor to the simulator -- with the two changes batching makes possible:

* the experience never leaves the device: tfode_batch_drain_transitions writes the tuples into a device tensor and the replay
  buffer is a ring of device tensors (the reference pickles Python tuples through Boost.Python);
* the actor goes to the engine as a weight blob (tfode_batch_load_policy) instead of a SavedModel directory on disk
  (`td3.actor.save( session_dir + '/actor' )` -> `trial( session_dir + '/actor' )`, rover_training_1.py:90,107,121).

PyTorch is the trainer harness here, as TensorFlow is in the reference; the stepping path stays the CUDA engine.
"""
import copy

import numpy as np
import torch
from torch import nn

from . import abi

STATE_SCALE = [90, 45, 25, 25, 45] + ([100] * 3 + [30] * 3) * 2   # scripts/rover_training_1.py:58
ACTION_SCALE = [15, 25]                                            # :59


def actor_net(s_dim=17, a_dim=2):   # scripts/rover_training_1.py:26-35
    return nn.Sequential(nn.Linear(s_dim, 512), nn.ReLU(), nn.Linear(512, 512), nn.ReLU(), nn.Linear(512, a_dim), nn.Tanh())


class Critic(nn.Module):            # scripts/rover_training_1.py:39-50
    def __init__(self, s_dim=17, a_dim=2):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(s_dim + a_dim, 512), nn.ReLU(), nn.Linear(512, 512), nn.ReLU(), nn.Linear(512, 1))

    def forward(self, s, a):
        return self.net(torch.cat([s, a], 1))


def _keras_init(module):
    """Keras Dense defaults: glorot_uniform kernels, zero biases (what the reference's networks start from)."""
    for m in module.modules():
        if isinstance(m, nn.Linear):
            nn.init.xavier_uniform_(m.weight)
            nn.init.zeros_(m.bias)


def actor_blob(actor):
    """The engine's weight blob (W1 b1 W2 b2 W3 b3, Keras [in,out] layout, fp32) of a torch actor."""
    parts = []
    for m in actor:
        if isinstance(m, nn.Linear):
            parts.append(m.weight.detach().t().contiguous().reshape(-1).float().cpu().numpy())
            parts.append(m.bias.detach().float().cpu().numpy())
    w = np.concatenate(parts).astype(np.float32)
    assert w.size == abi.POLICY_NFLOATS
    return w


class ReplayRing:
    """deque(maxlen=buffer_size) of (s, a, r, terminal, s') as device tensors; rows arrive in the engine's transition layout."""

    def __init__(self, capacity, device, s_dim=17, a_dim=2):
        self.cap, self.n, self.head = int(capacity), 0, 0
        self.s = torch.zeros(self.cap, s_dim, device=device); self.a = torch.zeros(self.cap, a_dim, device=device)
        self.r = torch.zeros(self.cap, 1, device=device); self.t = torch.zeros(self.cap, 1, device=device)
        self.s2 = torch.zeros(self.cap, s_dim, device=device)

    def extend_rows(self, rows):
        """rows: [m, abi.TRANSITION_WIDTH] = env id, s[17], a[2], r, terminal, s'[17], pad (include/tfode_batch.h)."""
        m = rows.shape[0]
        if m == 0:
            return
        if m > self.cap:
            rows = rows[-self.cap:]; m = self.cap
        idx = (self.head + torch.arange(m, device=rows.device)) % self.cap
        self.s[idx] = rows[:, 1:18]; self.a[idx] = rows[:, 18:20]; self.r[idx] = rows[:, 20:21]; self.t[idx] = rows[:, 21:22]; self.s2[idx] = rows[:, 22:39]
        self.head = (self.head + m) % self.cap
        self.n = min(self.cap, self.n + m)

    def __len__(self):
        return self.n

    def sample(self, k, gen):
        idx = torch.randint(0, self.n, (min(k, self.n),), device=self.s.device, generator=gen)
        return self.s[idx], self.a[idx], self.r[idx], self.t[idx], self.s2[idx]


class TD3:
    """scripts/MachineLearning/TD3.py with the hyper-parameters of scripts/rover_training_1.py:52-68 as defaults."""

    def __init__(self, s_dim=17, a_dim=2, state_scale=STATE_SCALE, action_scale=ACTION_SCALE, gamma=0.99, tau=5e-3, policy_update_delay=2,
                 policy_reg_sigma=0.05, policy_reg_bound=0.2, buffer_size=1e6, minibatch_size=128, learning_rate=1e-4, seed=None, device="cuda"):
        self.dev = torch.device(device)
        self.gamma, self.tau, self.delay, self.sigma, self.bound, self.mb = gamma, tau, policy_update_delay, policy_reg_sigma, policy_reg_bound, minibatch_size
        self.gen = torch.Generator(device=self.dev)
        self.gen.manual_seed(0 if seed is None else int(seed))
        if seed is not None:
            torch.manual_seed(int(seed))
        self.ss = torch.tensor(state_scale, dtype=torch.float32, device=self.dev)
        self.as_ = torch.tensor(action_scale, dtype=torch.float32, device=self.dev)
        self.actor = actor_net(s_dim, a_dim).to(self.dev); _keras_init(self.actor)
        self.actor_target = copy.deepcopy(self.actor)
        self.critics = [Critic(s_dim, a_dim).to(self.dev) for _ in range(2)]
        for c in self.critics:
            _keras_init(c)
        self.critic_targets = [copy.deepcopy(c) for c in self.critics]
        self.actor_opt = torch.optim.Adam(self.actor.parameters(), lr=learning_rate, eps=1e-7)          # tf.optimizers.Adam defaults
        self.critic_opts = [torch.optim.Adam(c.parameters(), lr=learning_rate, eps=1e-7) for c in self.critics]
        self.replay_buffer = ReplayRing(buffer_size, self.dev, s_dim, a_dim)
        self.n_iter = 0

    # _infer_actions / _infer_Q_values (TD3.py:180-222): scaling lives outside the networks
    def _actions(self, net, s):
        return net(s / self.ss) * self.as_

    def _q(self, net, s, a):
        return net(s / self.ss, a / self.as_)

    def _train_q(self, batch):   # TD3.py:225-263
        s, a, r, term, s2 = batch
        with torch.no_grad():
            a2 = self._actions(self.actor_target, s2)
            noise = (torch.randn(a2.shape, device=self.dev, generator=self.gen) * self.sigma).clamp(-self.bound, self.bound) * self.as_
            a2 = a2 + noise
            q2 = torch.min(self._q(self.critic_targets[0], s2, a2), self._q(self.critic_targets[1], s2, a2))
            target = r + self.gamma * q2 * (1.0 - term)
        losses = []
        for c, opt in zip(self.critics, self.critic_opts):
            loss = ((self._q(c, s, a) - target) ** 2).mean()
            opt.zero_grad(set_to_none=True); loss.backward(); opt.step()
            losses.append(loss.detach())
        return torch.stack(losses).mean()

    def _train_actor(self, s):   # TD3.py:266-284
        a = self._actions(self.actor, s)
        loss = -torch.min(self._q(self.critics[0], s, a), self._q(self.critics[1], s, a)).mean()
        self.actor_opt.zero_grad(set_to_none=True); loss.backward(); self.actor_opt.step()
        return loss.detach()

    @torch.no_grad()
    def _update_targets(self):   # TD3.py:287-294
        pairs = [(self.actor_target, self.actor)] + list(zip(self.critic_targets, self.critics))
        for tgt, src in pairs:
            for pt, p in zip(tgt.parameters(), src.parameters()):
                pt.mul_(1 - self.tau).add_(p, alpha=self.tau)

    def train(self, iterations=1):   # TD3.py:311-334
        if len(self.replay_buffer) < self.mb:
            return 0.0
        q_loss = torch.zeros((), device=self.dev)
        for _ in range(iterations):
            self.n_iter += 1
            for _ in range(self.delay):
                batch = self.replay_buffer.sample(self.mb, self.gen)
                q_loss = q_loss + self._train_q(batch)
            self._train_actor(batch[0])
            self._update_targets()
        return float(q_loss) / (iterations * self.delay)

    @torch.no_grad()
    def get_action(self, s):
        s = torch.as_tensor(s, dtype=torch.float32, device=self.dev)
        return self._actions(self.actor, s.reshape(-1, s.shape[-1])).squeeze(0).cpu().numpy()


def training_loop(td3, n_envs=256, max_episodes=1024, iter_per_ep=200, device=0, log=print, batch_factory=None, **cfg_overrides):
    """scripts/rover_training_1.py:101-125 with n_envs rovers per trial: every round runs one episode per environment of the
    rover_training_1 scene with the current actor (exploration on, randomised step yaw / IC start), appends the experience to the
    replay buffer on the device, trains ITER_PER_EP iterations per finished episode and installs the new actor.
    Returns the list of per-round dicts (episodes, buffer size, LQ, mean return)."""
    from .config import make_config
    from .engine import RoverBatch
    cfg = make_config(abi.PRESET_TRAINING, n_envs, device=device, auto_reset=0, transition_capacity=n_envs * 130, **cfg_overrides)
    batch = (batch_factory or RoverBatch)(cfg)
    dev = td3.dev
    buf = torch.zeros(cfg.transition_capacity, abi.TRANSITION_WIDTH, device=dev)
    rows = torch.zeros(n_envs, abi.ROLLOUT_WIDTH, device=dev)
    history, n_ep = [], 0
    try:
        while n_ep < max_episodes:
            batch.load_policy(actor_blob(td3.actor))
            batch.reset()
            for _ in range(61):                      # timeout 60 s: every environment ends its episode inside 61 x 1000 steps
                batch.step(1000)
                batch.pack_rollout(rows)
                if bool((rows[:, 21] > 0).all()):    # episodes_done: frozen environments cost nothing, stop when all are
                    break
            n = batch.drain_transitions(buf, buf.shape[0])
            td3.replay_buffer.extend_rows(buf[:n])
            n_ep += n_envs
            lq = td3.train(iter_per_ep * n_envs)
            rec = {"it": td3.n_iter, "episodes": n_ep, "buffer": len(td3.replay_buffer), "LQ": lq, "mean_return": float(rows[:, 19].mean()), "rows": n}
            history.append(rec)
            log("It %i | Ep %i | Bs %i | LQ %+7.4f" % (rec["it"], rec["episodes"], rec["buffer"], rec["LQ"]))   # the reference's progress line
    finally:
        batch.close()
    return history
