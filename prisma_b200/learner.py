"""DQN learner for the dqn_buffer agents (learner half), above the device-side environment.

Mirrors prisma/source/learner.py:160-255 and prisma/source/trainer.py:45-100.  `signaling_type` "ideal" (and "NN",
whose neighbour copies are refreshed by `sync_targets` every `sync_step`): node i samples (s, a, r, s', done) from its replay ring (HBM, prb_replay_sample), the TD target of a transition
whose action pointed at neighbour j is  r + gamma * (1 - done) * min_{a' != back to i} Q_target_j (s'),  the
loss is Huber (Q_i (s)[a] - target), the optimiser Adam (one state per node, as in the reference).  The M
networks live in ONE flat float32 tensor [M, P] in exactly the layout prb_qnet_load takes
(prisma_b200/qnet.py), so that

With `signaling_type` "target" (trainer.py:52-55) the stored reward already IS the target — the next node computed
it with its own current network when the packet reached it (prb_replay_set_target_mode) — and no target network
is consulted.

* acting on the device needs no conversion (`env.qnet_load (learner.weights ())`),
* the data-parallel gradient exchange is a single all-reduce over one bucket (torch.distributed, NCCL over
  NVLink on a multi-GPU job; gloo in the CPU tests).

PyTorch autograd and its matmul kernels do the training math here: library code above the environment, not
part of the simulated path."""
from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

from . import qnet

H1, H2 = qnet.H1, qnet.H2


def _views(flat: torch.Tensor, m: int, k_max: int):
    """Named views of one node's flat parameter vector."""
    out, o = {}, 0
    for name, shape in qnet.shapes(m, k_max).items():
        n = int(np.prod(shape))
        out[name] = flat[o:o + n].view(*shape)
        o += n
    return out


def forward(flat: torch.Tensor, obs: torch.Tensor, m: int, k: int, k_max: int, hidden: int = 2) -> torch.Tensor:
    """Q-values [B, k] of the node whose parameters are `flat` for observations [B, 1 + k_max]
    (prisma/source/models.py:257-301)."""
    w = _views(flat, m, k_max)
    dst = obs[:, 0].long().clamp(0, m - 1)
    hd = F.elu(w["w_dst"][dst] + w["b_dst"])
    buf = obs[:, 1:1 + k].float()
    ln = F.layer_norm(buf, (k,), eps=1e-3)
    hb = F.elu(ln @ w["w_buf"][:k] + w["b_buf"])
    h = F.elu(torch.cat([hd, hb], dim=1) @ w["w1"] + w["b1"])
    if hidden > 1:
        h = F.elu(h @ w["w2"] + w["b2"])
    return F.elu(h @ w["w3"][:, :k] + w["b3"][:k])


def huber(x: torch.Tensor, delta: float = 1.0) -> torch.Tensor:
    return torch.where(x.abs() < delta, 0.5 * x * x, delta * (x.abs() - 0.5 * delta))


class DQNLearner:
    def __init__(self, neighbors: Sequence[Sequence[int]], lr: float = 1e-3, gamma: float = 1.0, seed: int = 0,
                 device: str = "cpu", betas=(0.9, 0.999), adam_eps: float = 1e-7, signaling_type: str = "ideal",
                 variant: str = "dqn_buffer"):
        """neighbors[i] = overlay neighbours of node i in action order (BatchedRoutingEnv.neighbors mapped to overlay
        indices).  Adam defaults are Keras' (learner.py:103: tf.keras.optimizers.Adam (lr))."""
        if signaling_type not in ("ideal", "NN", "target"):
            raise ValueError("signaling_type must be ideal, NN or target")
        self.signaling_type = signaling_type
        self.neighbors = [list(n) for n in neighbors]
        self.m = len(self.neighbors)
        self.degrees = [len(n) for n in self.neighbors]
        self.k_max = max(self.degrees)
        self.gamma, self.lr, self.betas, self.adam_eps = float(gamma), float(lr), betas, float(adam_eps)
        self.hidden = qnet.hidden_layers(variant)
        nets = qnet.init_weights(self.m, self.degrees, self.k_max, seed=seed, variant=variant)
        w = torch.from_numpy(qnet.pack_weights(nets)).view(self.m, -1)
        self.W = w.to(device).requires_grad_(True)
        self.W_target = self.W.detach().clone()
        self.exp_avg = torch.zeros_like(self.W_target)
        self.exp_avg_sq = torch.zeros_like(self.W_target)
        self.steps = torch.zeros(self.m, dtype=torch.float32, device=device)
        # filtered_indices (trainer.py:60): at neighbour j, every action except the one that leads back to i
        self.filtered = [[[a for a, nb in enumerate(self.neighbors[j]) if nb != i] for j in self.neighbors[i]]
                         for i in range(self.m)]

    # ---- parameters ----
    def weights(self) -> np.ndarray:
        return self.W.detach().cpu().numpy().ravel()

    def sync_targets(self):
        """update_target / sync_neighbor_target_q_network (learner.py:190-199,256-270)."""
        self.W_target.copy_(self.W.detach())

    def q_values(self, node: int, obs: torch.Tensor) -> torch.Tensor:
        return forward(self.W[node], obs, self.m, self.degrees[node], self.k_max, self.hidden)

    # ---- one gradient step of every node that has a batch ----
    def targets(self, node: int, act, rew, nxt, done) -> torch.Tensor:
        """get_neighbor_target_value (learner.py:232-255) for a batch of node `node`."""
        t = rew.clone().float()
        if self.signaling_type == "target":
            return t                      # trainer.py:52-53: targets_t = rewards_t
        with torch.no_grad():
            for a, j in enumerate(self.neighbors[node]):
                sel = (act == a).nonzero(as_tuple=True)[0]
                idx = self.filtered[node][a]
                if len(sel) == 0 or not idx:
                    continue
                q = forward(self.W_target[j], nxt[sel], self.m, self.degrees[j], self.k_max, self.hidden)[:, idx]
                t[sel] = rew[sel].float() + self.gamma * (1.0 - done[sel].float()) * q.min(dim=1).values
        return t

    def loss(self, node: int, obs, act, rew, nxt, done, weights: Optional[torch.Tensor] = None):
        q = self.q_values(node, obs)
        q_sel = q.gather(1, act.long().clamp(0, self.degrees[node] - 1).view(-1, 1)).squeeze(1)
        td = q_sel - self.targets(node, act, rew, nxt, done)
        err = huber(td)
        if weights is not None:
            err = err * weights
        return err.mean(), td.detach()

    def step(self, batches: List[Optional[tuple]], world_size: int = 1, all_reduce=None):
        """batches[i] = (obs, action, reward, next_obs, done) of node i or None.  Gradients of all nodes are summed
        over ranks with ONE all-reduce (`all_reduce (tensor)`, e.g. torch.distributed.all_reduce) and divided by
        world_size; Adam is applied per node (nodes without a batch on any rank keep their state)."""
        if self.W.grad is not None:
            self.W.grad.zero_()
        active = torch.zeros(self.m, dtype=torch.float32, device=self.W.device)
        td_mse = [None] * self.m
        total = None
        for i, b in enumerate(batches):
            if b is None or len(b[1]) == 0:
                continue
            l, td = self.loss(i, *b)
            total = l if total is None else total + l
            active[i] = 1.0
            td_mse[i] = float((td * td).mean())
        if total is not None:
            total.backward()
        grad = self.W.grad if self.W.grad is not None else torch.zeros_like(self.W_target)
        if all_reduce is not None and world_size > 1:
            bucket = torch.cat([grad.reshape(-1), active])
            all_reduce(bucket)
            grad = (bucket[:-self.m] / world_size).view_as(self.W_target)
            active = (bucket[-self.m:] > 0).float()
        with torch.no_grad():
            b1, b2 = self.betas
            on = active.view(-1, 1)
            self.steps += active
            self.exp_avg.mul_(1 - on * (1 - b1)).add_(on * (1 - b1) * grad)
            self.exp_avg_sq.mul_(1 - on * (1 - b2)).add_(on * (1 - b2) * grad * grad)
            t = self.steps.clamp(min=1).view(-1, 1)
            m_hat = self.exp_avg / (1 - b1 ** t)
            v_hat = self.exp_avg_sq / (1 - b2 ** t)
            self.W -= on * self.lr * m_hat / (v_hat.sqrt() + self.adam_eps)
        return td_mse

    # ---- glue to the device-side environment ----
    def train_on_env(self, env, batch_size: int, seed: int = 0, min_rows: Optional[int] = None, world_size: int = 1,
                     all_reduce=None):
        """One training step from the replay rings of `env` (prb_replay_sample), then the new weights go back to
        the device-side agents (prb_qnet_load)."""
        counts, _ = env.replay_counts()
        need = batch_size if min_rows is None else min_rows          # trainer.py:41: total_samples >= batch_size
        batches = []
        for i in range(self.m):
            if counts[i] < need:
                batches.append(None)
                continue
            if getattr(env, "_replay_prioritized", False):
                # PrioritizedReplayBuffer.sample: importance weights multiply the Huber errors (learner.py:176-179)
                batches.append(env.replay_sample(i, batch_size, seed=seed, weighted=True))
            else:
                batches.append(env.replay_sample(i, batch_size, seed=seed))
        out = self.step(batches, world_size=world_size, all_reduce=all_reduce)
        env.qnet_load(self.weights())
        if getattr(env, "_replay_prioritized", False):
            env.replay_set_gradient_steps(self.steps.to(torch.int32).cpu().numpy())     # the targets computed from now on
        return out


class DeviceLearner:
    """The same learner with the gradient step on the device (prb_learner_*, prisma_b200/csrc/learner_core.cuh): sampling
    from the HBM replay rings, TD targets through the neighbours' target networks, forward, hand-written backward and Adam
    are CUDA kernels of the library; the only thing that leaves it is the ONE gradient bucket [M * P + M] floats that a
    data-parallel job all-reduces between `grad` and `apply`.  The online weights are the ones the device-side agents act
    on (prb_qnet_load), updated in place.  No autograd, no per-node Python loop."""

    def __init__(self, env, lr: float = 1e-3, gamma: float = 1.0, betas=(0.9, 0.999), adam_eps: float = 1e-7,
                 signaling_type: str = "ideal", weights: Optional[np.ndarray] = None, variant: str = "dqn_buffer"):
        import ctypes as C
        from . import _abi
        if signaling_type not in ("ideal", "NN", "target"):
            raise ValueError("signaling_type must be ideal, NN or target")
        self._C, self._abi = C, _abi
        self.env = env
        self.m = len(env.neighbors)
        self.k_max = max(len(n) for n in env.neighbors)
        self.p = qnet.floats_per_node(self.m, self.k_max)
        if weights is not None:
            env.qnet_load(weights)
        env.qnet_set_arch(qnet.hidden_layers(variant))
        _abi.check(env.lib.prb_learner_init(env.h, lr, gamma, betas[0], betas[1], adam_eps, 1 if signaling_type == "target" else 0))
        self._buf = {}

    def _buffers(self, batch):
        if batch not in self._buf:
            dev, ow = self.env._out.device, 1 + self.k_max
            f = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
            self._buf[batch] = dict(obs=f(self.m, batch, ow), nxt=f(self.m, batch, ow), act=torch.zeros((self.m, batch), dtype=torch.int32, device=dev),
                                    rew=f(self.m, batch), done=f(self.m, batch), wts=f(self.m, batch), active=f(self.m),
                                    bucket=f(self.m * self.p + self.m))
        return self._buf[batch]

    def read(self, which: int = 0) -> np.ndarray:
        n = self.m if which == 4 else self.m * self.p
        out = np.zeros(n, dtype=np.float32)
        self._abi.check(self.env.lib.prb_learner_read(self.env.h, which, out.ctypes.data_as(self._C.c_void_p), n))
        return out

    def weights(self) -> np.ndarray:
        return self.read(0)

    def sync_targets(self):
        self._abi.check(self.env.lib.prb_learner_sync_targets(self.env.h))

    def grad(self, batch, obs, act, rew, nxt, done, active, wts=None, want_td=True):
        """Gradient bucket (a device tensor [M * P + M]) of the batches given as [M, B, ...] device tensors."""
        C = self._C
        b = self._buffers(batch)
        td = np.zeros(self.m, dtype=np.float32)
        ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        self._abi.check(self.env.lib.prb_learner_grad(self.env.h, batch, ptr(obs), ptr(act), ptr(rew), ptr(nxt), ptr(done), ptr(wts), ptr(active),
                                                      ptr(b["bucket"]), td.ctypes.data_as(C.c_void_p) if want_td else None))
        return b["bucket"], td

    def apply(self, bucket, world_size: int = 1):
        self._abi.check(self.env.lib.prb_learner_apply(self.env.h, self._C.c_void_p(bucket.data_ptr()), int(world_size)))

    def train_on_env(self, env=None, batch_size: int = 512, seed: int = 0, min_rows: Optional[int] = None, world_size: int = 1,
                     all_reduce=None, want_td=True):
        """One training step from the replay rings (Trainer.step): sample, gradients, all-reduce of the one bucket over the
        ranks, Adam.  Returns the mean squared TD error per node (None for nodes without a batch)."""
        C = self._C
        b = self._buffers(batch_size)
        prio = getattr(self.env, "_replay_prioritized", False)
        ptr = lambda t: C.c_void_p(t.data_ptr())
        self._abi.check(self.env.lib.prb_replay_sample_all(self.env.h, batch_size, seed, batch_size if min_rows is None else min_rows, ptr(b["obs"]),
                                                           ptr(b["act"]), ptr(b["rew"]), ptr(b["nxt"]), ptr(b["done"]),
                                                           ptr(b["wts"]) if prio else None, ptr(b["active"])))
        bucket, td = self.grad(batch_size, b["obs"], b["act"], b["rew"], b["nxt"], b["done"], b["active"], b["wts"] if prio else None, want_td)
        if all_reduce is not None and world_size > 1:
            all_reduce(bucket)
        self.apply(bucket, world_size)
        if prio:
            self.env.replay_set_gradient_steps(self.read(4).astype(np.int32))
        if not want_td:
            return None
        act = b["active"].cpu().numpy()
        return [float(td[i]) if act[i] > 0 else None for i in range(self.m)]
