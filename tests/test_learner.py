"""The DQN learner above the environment (prisma_b200/learner.py) on CPU: the torch forward against a numpy
statement of DQN_buffer_model (prisma/source/models.py:257-301) on the flat device layout, TD targets and Huber
loss against a by-hand computation (learner.py:160-255, trainer.py:45-100), Adam against torch.optim.Adam, and
the one-bucket gradient all-reduce with world_size 2 over gloo."""
import os
import sys

import numpy as np
import pytest
import torch

from prisma_b200 import learner, qnet

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NEIGH = [[1, 2], [0, 2, 3], [0, 1], [1]]          # a 4-node overlay, action order = ascending neighbour id


def elu(x):
    return np.where(x > 0, x, np.expm1(x))


def numpy_forward(net, obs, m, k):
    dst = int(obs[0])
    hd = elu(net["w_dst"][dst] + net["b_dst"])
    buf = obs[1:1 + k].astype(np.float64)
    ln = (buf - buf.mean()) / np.sqrt(buf.var() + 1e-3)
    hb = elu(ln @ net["w_buf"][:k] + net["b_buf"])
    h = elu(np.concatenate([hd, hb]) @ net["w1"] + net["b1"])
    h = elu(h @ net["w2"] + net["b2"])
    return elu(h @ net["w3"][:, :k] + net["b3"][:k])


def random_batch(rng, node, n, m, k_max, degrees, neigh):
    obs = np.zeros((n, 1 + k_max), dtype=np.float32)
    obs[:, 0] = rng.integers(0, m, n)
    obs[:, 1:1 + degrees[node]] = rng.integers(0, 16260, (n, degrees[node]))
    act = rng.integers(0, degrees[node], n).astype(np.int32)
    nxt = np.zeros((n, 1 + k_max), dtype=np.float32)
    nxt[:, 0] = obs[:, 0]
    for r in range(n):
        j = neigh[node][act[r]]
        nxt[r, 1:1 + degrees[j]] = rng.integers(0, 16260, degrees[j])
    rew = rng.uniform(0.001, 0.3, n).astype(np.float32)
    done = (rng.uniform(size=n) < 0.2).astype(np.float32)
    return tuple(torch.from_numpy(x) for x in (obs, act, rew, nxt, done))


def test_forward_matches_numpy_on_the_flat_layout():
    L = learner.DQNLearner(NEIGH, seed=3)
    nets = qnet.init_weights(L.m, L.degrees, L.k_max, seed=3)
    assert np.array_equal(qnet.pack_weights(nets), L.weights())
    assert L.W.shape == (L.m, qnet.floats_per_node(L.m, L.k_max))
    rng = np.random.default_rng(0)
    for node in range(L.m):
        obs = random_batch(rng, node, 16, L.m, L.k_max, L.degrees, NEIGH)[0]
        q = L.q_values(node, obs).detach().numpy()
        for r in range(16):
            ref = numpy_forward(nets[node], obs[r].numpy(), L.m, L.degrees[node])
            assert np.allclose(q[r], ref, rtol=1e-4, atol=1e-5)


def test_targets_and_huber_loss_by_hand():
    L = learner.DQNLearner(NEIGH, gamma=0.9, seed=1)
    rng = np.random.default_rng(1)
    node = 1
    obs, act, rew, nxt, done = random_batch(rng, node, 64, L.m, L.k_max, L.degrees, NEIGH)
    t = L.targets(node, act, rew, nxt, done).numpy()
    for r in range(64):
        j = NEIGH[node][int(act[r])]
        q = L.q_values(j, nxt[r:r + 1]).detach().numpy()[0]            # target net == online net before any step
        keep = [a for a, nb in enumerate(NEIGH[j]) if nb != node]
        exp = float(rew[r]) + 0.9 * (1.0 - float(done[r])) * (min(q[a] for a in keep) if keep else 0.0)
        if not keep:
            exp = float(rew[r])
        assert abs(t[r] - exp) < 1e-5
    loss, td = L.loss(node, obs, act, rew, nxt, done)
    x = td.numpy()
    hub = np.where(np.abs(x) < 1.0, 0.5 * x * x, np.abs(x) - 0.5)
    assert abs(float(loss.detach()) - hub.mean()) < 1e-6


def test_adam_step_equals_torch_adam_and_leaves_idle_nodes_alone():
    L = learner.DQNLearner(NEIGH, lr=1e-3, seed=2)
    ref = L.W.detach().clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=1e-3, eps=1e-7)
    rng = np.random.default_rng(2)
    for it in range(3):
        batches = [random_batch(rng, i, 32, L.m, L.k_max, L.degrees, NEIGH) if i != 3 else None for i in range(L.m)]
        # reference: same loss through the same functions on a copy, torch.optim.Adam
        opt.zero_grad()
        total = 0
        w_keep = L.W
        L.W = ref
        tgt_keep = L.W_target
        for i, b in enumerate(batches):
            if b is not None:
                total = total + L.loss(i, *b)[0]
        total.backward()
        L.W = w_keep
        before3 = ref[3].detach().clone()
        opt.step()
        with torch.no_grad():
            ref[3] = before3                                        # node 3 had no batch
        mse = L.step(batches)
        assert mse[3] is None and all(m is not None for m in mse[:3])
        assert torch.allclose(L.W.detach()[:3], ref.detach()[:3], rtol=1e-5, atol=1e-7)
        assert torch.equal(L.W.detach()[3], before3)


def test_training_reduces_td_error_on_a_fixed_batch():
    L = learner.DQNLearner(NEIGH, lr=3e-3, seed=4)
    rng = np.random.default_rng(4)
    batches = [random_batch(rng, i, 128, L.m, L.k_max, L.degrees, NEIGH) for i in range(L.m)]
    first = L.step(batches)
    for _ in range(150):
        last = L.step(batches)
    assert all(b < 0.5 * a for a, b in zip(first, last))


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    L = learner.DQNLearner(NEIGH, lr=1e-3, seed=5)
    rng = np.random.default_rng(100 + rank)
    for it in range(2):
        batches = [random_batch(rng, i, 16, L.m, L.k_max, L.degrees, NEIGH) if not (rank == 1 and i == 0) else None
                   for i in range(L.m)]
        L.step(batches, world_size=world, all_reduce=dist.all_reduce)
    np.save(os.path.join(out_dir, "w%d.npy" % rank), L.weights())
    dist.barrier()
    dist.destroy_process_group()


def test_gradient_all_reduce_world2_gloo(tmp_path):
    import torch.multiprocessing as mp
    port = 29600 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    w0 = np.load(os.path.join(str(tmp_path), "w0.npy"))
    w1 = np.load(os.path.join(str(tmp_path), "w1.npy"))
    assert np.array_equal(w0, w1)                     # replicas stay identical
    # and equal to one process that sees both ranks' batches (mean of the two gradients)
    L = learner.DQNLearner(NEIGH, lr=1e-3, seed=5)
    rngs = [np.random.default_rng(100), np.random.default_rng(101)]
    for it in range(2):
        per_rank = [[random_batch(rngs[r], i, 16, L.m, L.k_max, L.degrees, NEIGH) if not (r == 1 and i == 0) else None
                     for i in range(L.m)] for r in range(2)]
        if L.W.grad is not None:
            L.W.grad.zero_()
        total = 0
        for r in range(2):
            for i, b in enumerate(per_rank[r]):
                if b is not None:
                    total = total + L.loss(i, *b)[0] / 2
        total.backward()
        g = L.W.grad.clone()

        def fake_all_reduce(bucket, g=g):
            bucket[:-L.m] = g.reshape(-1) * 2                      # the sum over ranks
            bucket[-L.m:] = 1.0
        L.W.grad = None
        L.step([None] * L.m, world_size=2, all_reduce=fake_all_reduce)
    assert np.allclose(L.weights(), w0, rtol=1e-5, atol=1e-7)


def test_target_signalling_uses_the_stored_targets():
    """signaling_type "target" (trainer.py:52-53): targets_t = rewards_t, no target network is consulted."""
    import torch
    from prisma_b200 import learner
    L = learner.DQNLearner([[1, 2], [0, 2], [0, 1]], lr=1e-2, seed=1, signaling_type="target")
    obs = torch.tensor([[2.0, 100.0, 0.0], [1.0, 0.0, 542.0]])
    act = torch.tensor([0, 1])
    rew = torch.tensor([0.25, 0.5])
    nxt = torch.tensor([[2.0, 0.0, 0.0], [1.0, 0.0, 0.0]])
    done = torch.tensor([0.0, 0.0])
    assert torch.equal(L.targets(0, act, rew, nxt, done), rew)
    L.W_target.add_(100.0)                       # would change every ideal-mode target
    loss, td = L.loss(0, obs, act, rew, nxt, done)
    q = L.q_values(0, obs).detach()
    assert torch.allclose(td, q[torch.arange(2), act] - rew)
    with pytest.raises(ValueError):
        learner.DQNLearner([[1], [0]], signaling_type="digital_twin")


@pytest.mark.parametrize("variant", ["dqn_buffer_lite", "dqn_routing", "dqn_buffer_lighter", "dqn_buffer_lighter_3"])
def test_narrower_model_variants_are_the_padded_network(variant):
    """DQN_buffer_lite_model (16/16 -> 32 -> 32, models.py:184-227) and DQN_routing_model (dst only: 32 -> 64 -> 64,
    models.py:360-392) laid out with zeros in the unused rows and columns of the dqn_buffer layout: the padded forward equals
    the narrow network's own forward, and the padding receives exactly zero gradient, so it stays zero through training."""
    import torch.nn.functional as F
    from prisma_b200 import qnet
    m, degrees = 5, [2, 3, 1, 3, 2]
    k_max = max(degrees)
    nets = qnet.init_weights(m, degrees, k_max, seed=3, variant=variant)
    flat = torch.from_numpy(qnet.pack_weights(nets)).view(m, -1).clone().requires_grad_(True)
    mask = torch.from_numpy(qnet.padding_mask(m, degrees, k_max, variant)).view(m, -1)
    assert float(flat.detach()[~mask].abs().max()) == 0.0 and float(flat.detach()[mask].abs().min()) >= 0.0
    rng = np.random.default_rng(0)
    total = 0.0
    for i in range(m):
        k = degrees[i]
        obs = np.zeros((40, 1 + k_max), dtype=np.float32)
        obs[:, 0] = rng.integers(0, m, 40)
        obs[:, 1:1 + k] = rng.integers(0, 16000, (40, k))
        x = torch.from_numpy(obs)
        q = learner.forward(flat[i], x, m, k, k_max, hidden=qnet.hidden_layers(variant))
        w = {n: torch.from_numpy(np.ascontiguousarray(v)) for n, v in qnet.narrow_view(nets[i], k, variant).items()}
        parts = [F.elu(w["w_dst"][x[:, 0].long()] + w["b_dst"])]
        if "w_buf" in w:
            parts.append(F.elu(F.layer_norm(x[:, 1:1 + k], (k,), eps=1e-3) @ w["w_buf"] + w["b_buf"]))
        h = F.elu(torch.cat(parts, dim=1) @ w["w1"] + w["b1"])
        if "w2" in w:
            h = F.elu(h @ w["w2"] + w["b2"])
        ref = F.elu(h @ w["w3"] + w["b3"])
        assert float((q.detach() - ref).abs().max()) < 1e-5
        total = total + (q * q).sum()
    total.backward()
    assert float(flat.grad[~mask].abs().max()) == 0.0
    assert float(flat.grad[mask].abs().max()) > 0.0
