"""Replay buffers in HBM (prb_replay_*): the device-side producer against a Python restatement of the
Forwarder's bookkeeping (prisma/source/forwarder.py:135-161 temp_obs, :214-244 lost packets, :352-379 transit
packets, 'ideal' signalling) run on the same step records.  Rows of different environments interleave in a
node's ring in any order, so stored rows are compared as multisets, bit for bit (float32)."""
import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu


def expected_rows(history, k_max, degrees, loss_penalty, num_envs, m):
    """history: list of (records numpy [E], actions numpy [E]); returns per-node lists of rows."""
    ow = 1 + k_max
    rows = [[] for _ in range(m)]
    temp = [dict() for _ in range(num_envs)]
    for rec, act in history:
        for e in range(num_envs):
            r = rec[e]
            if not r["valid"] or r["ptype"] != 0:
                continue
            me = int(r["overlay"])
            for j in range(int(r["n_lost"])):
                info = temp[e].pop(int(r["lost_uid"][j]))
                nxt = np.zeros(ow, dtype=np.float32)
                nxt[0] = info["obs"][0]
                rows[me].append(np.concatenate([info["obs"], [info["action"], np.float32(loss_penalty)], nxt, [1.0]]).astype(np.float32))
            uid = int(r["uid"])
            if not r["is_new"]:
                info = temp[e].pop(uid)
                nxt = np.zeros(ow, dtype=np.float32)
                nxt[:int(r["nobs"])] = r["obs"][:int(r["nobs"])]
                rows[info["node"]].append(np.concatenate([info["obs"], [info["action"], np.float32(r["hop_time"])], nxt,
                                                          [float(r["done"])]]).astype(np.float32))
            if r["need_action"]:
                temp[e][uid] = {"node": me, "obs": r["obs"][:ow].astype(np.float32), "action": int(act[e])}
    return rows


def canon(rows, width):
    a = np.asarray(rows, dtype=np.float32).reshape(-1, width)
    v = a.view(np.uint32)
    return v[np.lexsort(v.T[::-1])]


def test_device_replay_rows_match_forwarder_bookkeeping():
    import torch
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    # small buffers: packets are lost, so the loss transitions are exercised too
    cfg = common.env_config("abilene", num_envs=48, simTime=1.5, load_factor=2.0, pingAsObs=0, activateUnderlayTraffic=1,
                            max_out_buffer_size=2168)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    degrees = [len(n) for n in env.neighbors]
    k_max, m = max(degrees), len(degrees)
    env.qnet_load(qnet.pack_weights(qnet.init_weights(m, degrees, k_max, seed=3)))
    env.replay_enable(rows_per_node=1 << 16, pending_slots=4096)
    history = []
    batch = env.step(None)
    for it in range(1500):
        actions = env.qnet_act(eps=0.3, seed=5)
        env.replay_record(actions)
        torch.cuda.synchronize()
        history.append((batch.numpy().copy(), actions.cpu().numpy().copy()))
        batch = env.step(actions)
    env.check_errors()
    penalty = cfg.derived_loss_penalty(arrays.n_overlay)
    exp = expected_rows(history, k_max, degrees, penalty, cfg.num_envs, m)
    counts, missed = env.replay_counts()
    assert missed == 0
    width = 2 * (1 + k_max) + 3
    n_loss = 0
    for node in range(m):
        got = env.replay_rows(node)
        assert counts[node] == len(exp[node]) == len(got)
        assert np.array_equal(canon(exp[node], width), canon(got, width)), node
        n_loss += int((got[:, 1 + k_max + 1] == np.float32(penalty)).sum())
    assert counts.sum() > 5000 and n_loss > 0

    # sampling: every sampled row is a stored row of that node; indices cover the ring
    node = int(np.argmax(counts))
    stored = env.replay_rows(node)
    obs, act, rew, nxt, done = env.replay_sample(node, 4096, seed=9)
    torch.cuda.synchronize()
    samp = np.concatenate([obs.cpu().numpy(), act.cpu().numpy()[:, None].astype(np.float32), rew.cpu().numpy()[:, None],
                           nxt.cpu().numpy(), done.cpu().numpy()[:, None]], axis=1)
    keys = {r.tobytes() for r in stored}
    assert all(r.tobytes() in keys for r in samp)
    assert len({r.tobytes() for r in samp}) > 0.5 * min(len(stored), 4096) * 0.6
    obs2 = env.replay_sample(node, 4096, seed=9)[0]
    assert not torch.equal(obs, obs2)                     # a new call draws new indices
    env.close()


def test_ring_overwrites_oldest_and_fused_loop_feeds_it():
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=64, simTime=1.0, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    degrees = [len(n) for n in env.neighbors]
    env.qnet_load(qnet.pack_weights(qnet.init_weights(len(degrees), degrees, max(degrees), seed=3)))
    env.replay_enable(rows_per_node=256, pending_slots=2048)
    env.step_qnet(600, eps=0.1, seed=1)
    counts, missed = env.replay_counts()
    assert missed == 0 and counts.max() > 256            # the ring wrapped
    for node in range(len(degrees)):
        rows = env.replay_rows(node)
        assert len(rows) == min(counts[node], 256)
        assert (rows[:, 1 + max(degrees) + 1] > 0).all()  # every stored reward is a positive hop time
    env.check_errors()
    env.close()


def test_target_signalling_rows_hold_the_next_nodes_target():
    """prb_replay_set_target_mode: the reward column of a transit transition is hop_time + gamma * min Q of the NEXT
    node over its actions that do not lead back (trainer.py:52-60, learner.py:232-255), from the same Q-values that
    choose the action; terminal and lost-packet rows keep the plain reward.  Checked against the Q-values the
    device reports (q_out), 1e-6 relative (one fp32 multiply-add); the fused loop stores the same rows."""
    import torch
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    gamma = 0.9
    cfg = common.env_config("abilene", num_envs=32, simTime=1.0, load_factor=1.5, pingAsObs=0, activateUnderlayTraffic=1,
                            max_out_buffer_size=4000)
    arrays = cfg.load()
    results = []
    for fused in (False, True):
        env = BatchedRoutingEnv(cfg, arrays=arrays)
        degrees = [len(n) for n in env.neighbors]
        k_max, m = max(degrees), len(degrees)
        env.qnet_load(qnet.pack_weights(qnet.init_weights(m, degrees, k_max, seed=3)))
        env.replay_enable(rows_per_node=1 << 15, pending_slots=4096)
        env.replay_set_target_mode(True, gamma)
        if fused:
            env.step(None)
            env.step_qnet(800, eps=0.0, seed=5)
            results.append([env.replay_rows(n) for n in range(m)])
            env.close()
            continue
        q = torch.zeros((env.num_envs, k_max), dtype=torch.float32, device="cuda")
        history, tmins = [], []
        batch = env.step(None)
        for it in range(800):
            actions = env.qnet_act(eps=0.0, seed=5, q_out=q)
            env.replay_record(actions)
            torch.cuda.synchronize()
            rec = batch.numpy().copy()
            qq = q.cpu().numpy()
            tm = np.full(env.num_envs, np.inf, dtype=np.float32)
            for e in range(env.num_envs):
                if rec["valid"][e] and rec["need_action"][e]:
                    me = int(rec["overlay"][e])
                    ok = [a for a, nb in enumerate(env.neighbors[me]) if int(nb) != int(rec["prev_node"][e])]
                    if ok:
                        tm[e] = qq[e, ok].min()
            history.append((rec, actions.cpu().numpy().copy()))
            tmins.append(tm)
            batch = env.step(actions)
        env.check_errors()
        # expected rows: the `ideal` bookkeeping with the transit reward replaced by the target
        for (rec, _), tm in zip(history, tmins):
            for e in range(env.num_envs):
                if rec["valid"][e] and rec["ptype"][e] == 0 and not rec["is_new"][e] and rec["need_action"][e] and np.isfinite(tm[e]):
                    rec["hop_time"][e] = np.float32(np.float32(rec["hop_time"][e]) + np.float32(gamma) * tm[e])
        exp = expected_rows(history, k_max, degrees, cfg.derived_loss_penalty(arrays.n_overlay), cfg.num_envs, m)
        counts, missed = env.replay_counts()
        assert missed == 0 and counts.sum() > 3000
        n_target = 0
        for node in range(m):
            got = env.replay_rows(node)
            a, b = canon(exp[node], 2 * (1 + k_max) + 3).view(np.float32), canon(got, 2 * (1 + k_max) + 3).view(np.float32)
            assert a.shape == b.shape
            rc = 1 + k_max + 1
            keep = [c for c in range(a.shape[1]) if c != rc]
            # sort by everything except the reward, then compare rewards within tolerance
            oa = np.lexsort(np.round(a, 4).T[::-1])
            ob = np.lexsort(np.round(b, 4).T[::-1])
            assert np.array_equal(a[oa][:, keep], b[ob][:, keep]), node
            assert np.allclose(a[oa][:, rc], b[ob][:, rc], rtol=1e-6, atol=1e-7), node
            n_target += int((got[:, -1] == 0).sum())
        assert n_target > 1000                       # non-terminal transitions carry targets
        results.append([env.replay_rows(n) for n in range(m)])
        env.close()
    for a, b in zip(*results):
        assert np.array_equal(canon(a, a.shape[1]), canon(b, b.shape[1]))     # fused loop == explicit loop, bit for bit


def test_prioritized_variant_weights_rows_by_gradient_step_staleness():
    """PrioritizedReplayBuffer (replay_buffer.py:393-534) as the reference uses it with `target` signalling
    (forwarder.py:494-505): rows remember the gradient step of the node that computed their target, the node the latest
    step seen per action; sample = uniform draws with weight p / sum (p) * len, p = (row step / latest step) ^ alpha."""
    import torch
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = common.env_config("abilene", num_envs=32, simTime=1.0, load_factor=1.5, pingAsObs=0, activateUnderlayTraffic=1,
                            max_out_buffer_size=4000)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, arrays=arrays)
    degrees = [len(n) for n in env.neighbors]
    k_max, m = max(degrees), len(degrees)
    env.qnet_load(qnet.pack_weights(qnet.init_weights(m, degrees, k_max, seed=3)))
    env.replay_enable(rows_per_node=1 << 15, pending_slots=4096)
    env.replay_set_target_mode(True, 1.0)
    alpha = 0.5
    env.replay_enable_prioritized(alpha)
    u2o = {int(u): i for i, u in enumerate(env.overlay_to_underlay)}
    steps = np.ones(m, dtype=np.int32)
    for phase in range(4):                       # the learners advance between collection phases, node i by i + 1 steps
        env.replay_set_gradient_steps(steps)
        env.step_qnet(250, eps=0.2, seed=phase)
        steps = steps + np.arange(1, m + 1, dtype=np.int32)
    env.check_errors()
    counts, missed = env.replay_counts()
    assert missed == 0
    node = int(np.argmax(counts))
    rows = env.replay_rows(node)
    g, latest = env.replay_priorities(node)
    assert len(g) == len(rows) > 500
    act = rows[:, 1 + k_max].astype(int)
    done_or_lost = rows[:, -1] == 1
    # a transit row carries the step count of the neighbour its action pointed at, at the time the target was computed
    for a, nb in enumerate(env.neighbors[node]):
        j = u2o[int(nb)]
        allowed = {1 + p * (j + 1) for p in range(4)} | {1}
        sel = (act == a) & ~done_or_lost
        assert sel.any() and set(np.unique(g[sel]).astype(int)) <= allowed, (a, np.unique(g[sel]))
        assert latest[a] == g[act == a].max()
    assert len(np.unique(g)) > 2
    p = (g / latest[act]) ** alpha
    obs, a_s, rew, nxt, done, w = env.replay_sample(node, 2048, seed=4, weighted=True)
    torch.cuda.synchronize()
    samp = np.concatenate([obs.cpu().numpy(), a_s.cpu().numpy()[:, None].astype(np.float32), rew.cpu().numpy()[:, None],
                           nxt.cpu().numpy(), done.cpu().numpy()[:, None]], axis=1)
    expect = {}
    for r, pi in zip(rows, p):
        expect.setdefault(r.tobytes(), set()).add(np.float32(pi / p.sum() * len(rows)))
    w = w.cpu().numpy()
    for r, wi in zip(samp, w):
        cands = expect[r.tobytes()]
        assert min(abs(float(c) - float(wi)) for c in cands) < 1e-4 * max(1.0, float(wi)), (cands, wi)
    assert 0.5 < w.mean() < 1.5 and w.std() > 0.01          # the weights average to ~1 and do differ
    # the learner multiplies the Huber errors by them (learner.py:176-179)
    from prisma_b200 import learner
    neigh = [[u2o[int(u)] for u in nb] for nb in env.neighbors]
    L = learner.DQNLearner(neigh, lr=1e-3, seed=2, device="cuda", signaling_type="target")
    mse = L.train_on_env(env, batch_size=256, seed=1)
    assert all(v is not None and np.isfinite(v) for v in mse)
    g2, _ = env.replay_priorities(node)
    assert np.array_equal(g2, g)                 # sampling does not touch the stored steps
    # without the prioritized variant the weights are 1
    env2 = BatchedRoutingEnv(cfg, arrays=arrays)
    env2.qnet_load(qnet.pack_weights(qnet.init_weights(m, degrees, k_max, seed=3)))
    env2.replay_enable(rows_per_node=4096)
    env2.step_qnet(200, eps=0.1, seed=0)
    assert bool((env2.replay_sample(node, 64, weighted=True)[5] == 1).all())
    env.close()
    env2.close()
