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
