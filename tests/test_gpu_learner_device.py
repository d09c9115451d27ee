"""The learner's gradient step on the device (prb_learner_*, prisma_b200/csrc/learner_core.cuh) against the PyTorch autograd
statement of the same step (prisma_b200/learner.py:DQNLearner, itself held to numpy / torch.optim.Adam in tests/test_learner.py):
identical batches drawn from the HBM replay rings, gradients of every parameter of every node, TD errors, and the weights after
several Adam steps.  Tolerance: 2e-5 of the largest gradient entry (fp32, different summation order); weights 1e-5 absolute.
The reference's TensorFlow numerics are unpinned."""
import ctypes as C

import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu


def setup(signaling_type="ideal", num_envs=256, variant="dqn_buffer"):
    import torch
    from prisma_b200 import learner
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=num_envs, simTime=5.0, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    u2o = {int(u): i for i, u in enumerate(env.overlay_to_underlay)}
    neigh = [[u2o[int(u)] for u in nb] for nb in env.neighbors]
    L = learner.DQNLearner(neigh, lr=1e-3, gamma=0.9, seed=4, device="cuda", signaling_type=signaling_type, variant=variant)
    env.qnet_load(L.weights())
    env.replay_enable(rows_per_node=1 << 15)
    if signaling_type == "target":
        env.replay_set_target_mode(True, gamma=0.9)
    env.step_qnet(300, eps=0.3, seed=1)
    D = learner.DeviceLearner(env, lr=1e-3, gamma=0.9, signaling_type=signaling_type, variant=variant)
    # a target network that differs from the online one
    with torch.no_grad():
        L.W += 0.01 * torch.randn_like(L.W)
    env.qnet_load(L.weights())
    return env, L, D


@pytest.mark.parametrize("signaling_type,variant", [("ideal", "dqn_buffer"), ("target", "dqn_buffer"), ("ideal", "dqn_buffer_lighter")])
def test_device_gradients_and_adam_match_autograd(signaling_type, variant):
    import torch
    env, L, D = setup(signaling_type, variant=variant)
    B = 256
    for it in range(3):
        b = D._buffers(B)
        p = lambda t: C.c_void_p(t.data_ptr())
        from prisma_b200 import _abi
        _abi.check(env.lib.prb_replay_sample_all(env.h, B, 10 + it, B, p(b["obs"]), p(b["act"]), p(b["rew"]), p(b["nxt"]), p(b["done"]), None,
                                                 p(b["active"])))
        assert float(b["active"].sum()) == L.m
        assert (b["rew"] != 0).any() and (b["done"] > 0).any()
        bucket, td = D.grad(B, b["obs"], b["act"], b["rew"], b["nxt"], b["done"], b["active"])
        got = bucket[:L.m * D.p].view(L.m, D.p).clone()
        # the same batches through autograd
        batches = [(b["obs"][i], b["act"][i], b["rew"][i], b["nxt"][i], b["done"][i]) for i in range(L.m)]
        if L.W.grad is not None:
            L.W.grad.zero_()
        total = None
        ref_td = []
        for i, bt in enumerate(batches):
            l, tdv = L.loss(i, *bt)
            total = l if total is None else total + l
            ref_td.append(float((tdv * tdv).mean()))
        total.backward()
        ref = L.W.grad.detach()
        scale = float(ref.abs().max())
        assert scale > 1e-4
        assert float((got - ref).abs().max()) <= 2e-5 * max(scale, 1.0), (float((got - ref).abs().max()), scale)
        assert np.allclose(td, ref_td, rtol=1e-4, atol=1e-7)
        # one Adam step: k_learn_adam against numpy on the SAME gradients (Keras' defaults; an update is g / (|g| + 1e-7)-like, so
        # comparing two learners through their slightly different near-zero gradients would test the noise, not the kernel)
        w0, m0, v0 = D.read(0).astype(np.float64), D.read(2).astype(np.float64), D.read(3).astype(np.float64)
        gnp = got.cpu().numpy().astype(np.float64).ravel()
        D.apply(bucket, 1)
        t = it + 1
        m1 = 0.9 * m0 + 0.1 * gnp
        v1 = 0.999 * v0 + 0.001 * gnp * gnp
        w1 = w0 - 1e-3 * (m1 / (1 - 0.9 ** t)) / (np.sqrt(v1 / (1 - 0.999 ** t)) + 1e-7)
        assert np.abs(D.weights() - w1).max() < 2e-6
        # float32 sums of terms of opposite sign: absolute tolerance relative to the largest moment
        assert np.allclose(D.read(2), m1, rtol=1e-4, atol=1e-6 * np.abs(m1).max())
        assert np.allclose(D.read(3), v1, rtol=1e-4, atol=1e-6 * np.abs(v1).max())
        assert np.array_equal(D.read(4), np.full(L.m, it + 1, dtype=np.float32))
        with torch.no_grad():                       # the autograd learner follows the device's weights
            L.W.copy_(torch.from_numpy(D.weights()).view(L.m, -1).cuda())
        if it == 1:
            L.sync_targets()
            D.sync_targets()
            assert np.array_equal(D.read(1), D.read(0))
    env.close()


def test_device_learner_trains_and_the_agents_act_on_the_new_weights():
    """Collect -> train on the device -> collect: the weights the agents act on are the trained ones, inactive nodes keep their
    state, and the TD error of a fixed batch falls."""
    import torch
    env, L, D = setup("ideal", num_envs=128)
    w0 = D.weights().copy()
    first = last = None
    for it in range(30):
        mse = D.train_on_env(batch_size=128, seed=5)         # the same rows every time: the fit must improve
        vals = [m for m in mse if m is not None]
        assert len(vals) == L.m and np.isfinite(vals).all()
        first = first if first is not None else float(np.mean(vals))
        last = float(np.mean(vals))
    assert last < first
    w1 = D.weights()
    assert np.abs(w1 - w0).max() > 1e-4
    # the Q-values the device-side agents compute are those of the trained weights
    q = torch.zeros((env.num_envs, L.k_max), dtype=torch.float32, device="cuda")
    env.step_qnet(5, eps=0.0, seed=9)
    rec = env.last.numpy()
    env.qnet_act(q_out=q)
    torch.cuda.synchronize()
    with torch.no_grad():
        L.W.copy_(torch.from_numpy(w1).view(L.m, -1).cuda())
    need = np.nonzero((rec["valid"] != 0) & (rec["need_action"] != 0))[0][:32]
    assert len(need) > 5
    for e in need:
        node = int(rec["overlay"][e])
        obs = torch.from_numpy(rec["obs"][e][:1 + L.k_max].astype(np.float32)).cuda().view(1, -1)
        ref = L.q_values(node, obs).detach().cpu().numpy()[0]
        assert np.abs(q[e, :L.degrees[node]].cpu().numpy() - ref).max() < 1e-4
    # a node without enough rows is left alone
    mse = D.train_on_env(batch_size=128, seed=6, min_rows=1 << 30)
    assert all(m is None for m in mse) and np.array_equal(D.weights(), w1)
    env.check_errors()
    env.close()


@pytest.mark.parametrize("variant", ["dqn_buffer_lite", "dqn_routing", "dqn_buffer_lighter", "dqn_buffer_lighter_3"])
def test_narrower_variants_act_and_learn_on_the_device(variant):
    """DQN_buffer_lite_model / DQN_routing_model as zero-padded dqn_buffer layouts (prisma_b200/qnet.py): the device-side agents
    compute the narrow network's Q-values, and after training on the device the padding is still exactly zero."""
    import torch
    import torch.nn.functional as F
    from prisma_b200 import learner, qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=128, simTime=5.0, load_factor=1.3, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    degrees = [len(n) for n in env.neighbors]
    m, k_max = len(degrees), max(degrees)
    nets = qnet.init_weights(m, degrees, k_max, seed=5, variant=variant)
    mask = qnet.padding_mask(m, degrees, k_max, variant)
    env.replay_enable(rows_per_node=1 << 14)
    D = learner.DeviceLearner(env, lr=1e-3, weights=qnet.pack_weights(nets), variant=variant)
    env.step_qnet(200, eps=0.3, seed=1)
    for it in range(5):
        mse = D.train_on_env(batch_size=128, seed=it)
        assert all(x is not None and np.isfinite(x) for x in mse)
    w = D.weights()
    assert np.abs(w[~mask]).max() == 0.0                       # padding untouched by Adam
    assert np.abs(w - qnet.pack_weights(nets))[mask].max() > 1e-5
    # Q-values of the device-side agents == the narrow network evaluated by PyTorch on the trained weights
    q = torch.zeros((env.num_envs, k_max), dtype=torch.float32, device="cuda")
    env.qnet_act(q_out=q)
    torch.cuda.synchronize()
    rec = env.last.numpy()
    need = np.nonzero((rec["valid"] != 0) & (rec["need_action"] != 0))[0][:24]
    assert len(need) > 5
    P = qnet.floats_per_node(m, k_max)
    for e in need:
        i = int(rec["overlay"][e])
        k = degrees[i]
        flat, o, net = w[i * P:(i + 1) * P], 0, {}
        for name, shp in qnet.shapes(m, k_max).items():
            n = int(np.prod(shp))
            net[name] = flat[o:o + n].reshape(shp)
            o += n
        nv = {n_: torch.from_numpy(np.ascontiguousarray(v)) for n_, v in qnet.narrow_view(net, k, variant).items()}
        x = torch.from_numpy(rec["obs"][e][:1 + k_max].astype(np.float32)).view(1, -1)
        parts = [F.elu(nv["w_dst"][x[:, 0].long()] + nv["b_dst"])]
        if "w_buf" in nv:
            parts.append(F.elu(F.layer_norm(x[:, 1:1 + k], (k,), eps=1e-3) @ nv["w_buf"] + nv["b_buf"]))
        h = F.elu(torch.cat(parts, dim=1) @ nv["w1"] + nv["b1"])
        if "w2" in nv:
            h = F.elu(h @ nv["w2"] + nv["b2"])
        ref = F.elu(h @ nv["w3"] + nv["b3"]).numpy()[0]
        assert np.abs(q[e, :k].cpu().numpy() - ref).max() < 1e-4
    env.check_errors()
    env.close()
