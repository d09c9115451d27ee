"""Device-side dqn_buffer agents (prb_qnet_act / prb_step_qnet) against a plain PyTorch fp32 statement of the
reference's DQN_buffer_model (prisma/source/models.py:257-301) and its argmin / eps-greedy step
(prisma/source/learner.py:142-159).  Tolerance: 1e-5 absolute on Q-values (fp32, different summation order);
actions equal wherever the two smallest reference Q-values are more than 1e-4 apart."""
import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu


def torch_reference(nets, degrees, overlay, obs):
    import torch
    import torch.nn.functional as F
    q_all = torch.zeros((len(overlay), obs.shape[1] - 1), dtype=torch.float32)
    for e in range(len(overlay)):
        m = int(overlay[e])
        k = degrees[m]
        w = {n: torch.from_numpy(v) for n, v in nets[m].items()}
        dst = F.one_hot(torch.tensor(int(obs[e, 0])), num_classes=len(nets)).float()
        hd = F.elu(dst @ w["w_dst"] + w["b_dst"])
        buf = torch.from_numpy(obs[e, 1:1 + k].astype(np.float32))
        ln = F.layer_norm(buf, (k,), eps=1e-3)
        hb = F.elu(ln @ w["w_buf"][:k] + w["b_buf"])
        h = F.elu(torch.cat([hd, hb]) @ w["w1"] + w["b1"])
        h = F.elu(h @ w["w2"] + w["b2"])
        q_all[e, :k] = F.elu(h @ w["w3"][:, :k] + w["b3"][:k])
    return q_all.numpy()


def setup(num_envs=256, **over):
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=num_envs, simTime=1.0, load_factor=1.5, pingAsObs=0,
                            activateUnderlayTraffic=1, **over)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    degrees = [len(n) for n in env.neighbors]
    k_max = max(degrees)
    nets = qnet.init_weights(len(degrees), degrees, k_max, seed=7)
    rng = np.random.default_rng(1)
    for net in nets:                      # non-zero biases so that every term is exercised
        for b in ("b_dst", "b_buf", "b1", "b2", "b3"):
            net[b][:] = rng.normal(0, 0.1, size=net[b].shape).astype(np.float32)
    env.qnet_load(qnet.pack_weights(nets))
    return env, nets, degrees, k_max


@pytest.mark.parametrize("mode,tol", [("fused", 1e-5), ("mma", 2e-5)])
def test_q_values_and_argmin_match_torch_fp32(mode, tol):
    """mode 'mma': decisions bucketed by node, Dense64 -> Dense64 -> Dense(k) as mma.sync TF32 tiles with split operands
    (k_qnet_mma); tolerance 2e-5 absolute on Q-values of magnitude ~1, actions equal wherever the margin exceeds 1e-4."""
    import torch
    env, nets, degrees, k_max = setup()
    env.qnet_set_mode(mode)
    q = torch.zeros((env.num_envs, k_max), dtype=torch.float32, device="cuda")
    batch = env.step(None)
    checked = 0
    for it in range(60):
        actions = env.qnet_act(q_out=q)
        torch.cuda.synchronize()
        rec = batch.numpy()
        need = (rec["valid"] != 0) & (rec["need_action"] != 0)
        if it % 10 == 9:                 # after some steps the queues are non-empty
            idx = np.nonzero(need)[0]
            ref = torch_reference(nets, degrees, rec["overlay"][idx], rec["obs"][idx][:, :1 + k_max])
            got = q.cpu().numpy()[idx]
            assert np.abs(ref - got).max() < tol, np.abs(ref - got).max()
            act = actions.cpu().numpy()[idx]
            for j, e in enumerate(idx):
                k = degrees[int(rec["overlay"][e])]
                srt = np.sort(ref[j, :k])
                assert 0 <= act[j] < k
                if k == 1 or srt[1] - srt[0] > 1e-4:
                    assert act[j] == int(np.argmin(ref[j, :k]))
            assert (actions.cpu().numpy()[~need] == 0).all()
            checked += len(idx)
            assert rec["obs"][idx][:, 1:].max() > 0
        batch = env.step(actions)
    assert checked > 500
    env.check_errors()
    env.close()


def test_step_qnet_equals_step_plus_act():
    """The fused device loop is the same sequence of launches as the explicit one: identical counters."""
    res = []
    for fused in (False, True):
        env, nets, degrees, k_max = setup(num_envs=64)
        if fused:
            env.step_qnet(400)
        else:
            batch = env.step(None)
            for _ in range(399):
                batch = env.step(env.qnet_act())
        c = env.counters()
        env.check_errors(c)
        res.append(c)
        env.close()
    for k in common.COUNTER_INTS:
        assert np.array_equal(res[0][k], res[1][k]), k
    assert int(res[0]["packet_hops"].sum()) > 0


def test_eps_greedy_explores_uniformly():
    import torch
    env, nets, degrees, k_max = setup(num_envs=2048)
    batch = env.step(None)
    greedy = env.qnet_act().clone()
    rnd = env.qnet_act(eps=1.0, seed=11).clone()
    half = env.qnet_act(eps=0.5, seed=11).clone()
    torch.cuda.synchronize()
    rec = batch.numpy()
    need = (rec["valid"] != 0) & (rec["need_action"] != 0)
    k = np.array([degrees[int(m)] for m in rec["overlay"]])
    g, r, h = (x.cpu().numpy() for x in (greedy, rnd, half))
    assert ((r >= 0) & (r < k))[need].all()
    multi = need & (k > 1)
    assert 0.3 < np.mean(r[multi] != g[multi]) < 0.9          # a uniform draw differs from argmin (k-1)/k of the time
    changed = np.mean(h[multi] != g[multi])
    assert 0.1 < changed < 0.5
    env.close()


def test_learner_trains_from_the_device_replay_rings():
    """Collect with the device-side agents, train with prisma_b200.learner from the HBM replay rings, load the
    new weights back: the device forward and the learner's torch forward stay the same function."""
    import torch
    from prisma_b200 import learner
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=256, simTime=3.0, load_factor=1.2, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    u2o = {int(u): i for i, u in enumerate(env.overlay_to_underlay)}
    neigh = [[u2o[int(u)] for u in nb] for nb in env.neighbors]
    L = learner.DQNLearner(neigh, lr=1e-3, seed=2, device="cuda")
    env.qnet_load(L.weights())
    env.replay_enable(rows_per_node=1 << 15)
    mse_first = mse_last = None
    for it in range(12):
        env.step_qnet(200, eps=0.2, seed=it)
        mse = L.train_on_env(env, batch_size=256, seed=it)
        if it % 4 == 3:
            L.sync_targets()
        vals = [m for m in mse if m is not None]
        assert len(vals) == L.m and all(np.isfinite(vals))
        mse_first = mse_first or float(np.mean(vals))
        mse_last = float(np.mean(vals))
    counts, missed = env.replay_counts()
    assert missed == 0 and counts.min() > 256
    # device forward == learner forward on the current weights
    q = torch.zeros((env.num_envs, L.k_max), dtype=torch.float32, device="cuda")
    batch = env.last
    env.qnet_act(q_out=q)
    torch.cuda.synchronize()
    rec = batch.numpy()
    need = np.nonzero((rec["valid"] != 0) & (rec["need_action"] != 0))[0]
    assert len(need) > 50
    for e in need[:64]:
        node = int(rec["overlay"][e])
        obs = torch.from_numpy(rec["obs"][e][:1 + L.k_max].astype(np.float32)).cuda().view(1, -1)
        ref = L.q_values(node, obs).detach().cpu().numpy()[0]
        assert np.abs(q[e, :L.degrees[node]].cpu().numpy() - ref).max() < 1e-4
    env.check_errors()
    env.close()


def test_batched_tensor_core_loop_keeps_the_books_like_the_warp_path():
    """prb_step_qnet in 'mma' mode (bucket + k_qnet_mma + k_advance per step, replay bookkeeping inside k_qnet_mma) against
    the one-warp-per-environment path on the same seeds.  The Q-values differ in the last bits, so a near-tie may pick
    another neighbour and the trajectories may part: the comparison is statistical (packet hops and stored transitions
    within 2 %), the invariants are exact (no error flag, no missed transition, packet conservation)."""
    out = {}
    for mode in ("fused", "mma"):
        env, nets, degrees, k_max = setup(num_envs=512)
        env.qnet_set_mode(mode)
        env.replay_enable(rows_per_node=1 << 16)
        env.step_qnet(300, eps=0.05, seed=3)
        c = env.counters()
        env.check_errors(c)
        counts, missed = env.replay_counts()
        assert missed == 0
        inflight = c["overlay_injected"] - c["overlay_arrived"] - c["overlay_lost"]
        assert (inflight >= 0).all()
        rows = env.replay_rows(int(np.argmax(counts)))
        assert np.isfinite(rows).all() and (rows[:, 1 + k_max] < k_max).all()
        out[mode] = (int(c["packet_hops"].sum()), int(counts.sum()), int(c["env_steps"].sum()))
        env.close()
    for a, b in zip(out["fused"], out["mma"]):
        assert abs(a - b) <= 0.02 * a, out
    assert out["mma"][2] == 512 * 300 or out["mma"][2] > 0


def test_one_hidden_layer_models_on_every_forward_path():
    """DQN_buffer_lighter_model (one hidden layer, models.py:39-84) as a zero-padded layout with prb_qnet_set_arch (1): the
    fused / warp kernel and the tensor-core kernel give the Q-values of the narrow network (PyTorch fp32)."""
    import torch
    import torch.nn.functional as F
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=256, simTime=1.0, load_factor=1.5, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    degrees = [len(n) for n in env.neighbors]
    m, k_max = len(degrees), max(degrees)
    nets = qnet.init_weights(m, degrees, k_max, seed=9, variant="dqn_buffer_lighter")
    env.qnet_load(qnet.pack_weights(nets))
    env.qnet_set_arch(1)
    env.step_qnet(40, eps=0.1, seed=2)                  # fused path with the switch
    rec = env.last.numpy()
    need = np.nonzero((rec["valid"] != 0) & (rec["need_action"] != 0))[0]
    assert len(need) > 50
    q = torch.zeros((env.num_envs, k_max), dtype=torch.float32, device="cuda")
    for mode, tol in (("warp", 1e-5), ("mma", 2e-5)):
        env.qnet_set_mode(mode)
        env.qnet_act(q_out=q)
        torch.cuda.synchronize()
        got = q.cpu().numpy()
        for e in need[:48]:
            i = int(rec["overlay"][e])
            k = degrees[i]
            nv = {n_: torch.from_numpy(np.ascontiguousarray(v)) for n_, v in qnet.narrow_view(nets[i], k, "dqn_buffer_lighter").items()}
            x = torch.from_numpy(rec["obs"][e][:1 + k_max].astype(np.float32)).view(1, -1)
            hd = F.elu(nv["w_dst"][x[:, 0].long()] + nv["b_dst"])
            hb = F.elu(F.layer_norm(x[:, 1:1 + k], (k,), eps=1e-3) @ nv["w_buf"] + nv["b_buf"])
            h = F.elu(torch.cat([hd, hb], dim=1) @ nv["w1"] + nv["b1"])
            ref = F.elu(h @ nv["w3"] + nv["b3"]).numpy()[0]
            assert np.abs(got[e, :k] - ref).max() < tol, (mode, np.abs(got[e, :k] - ref).max())
    env.check_errors()
    env.close()
