"""The flag-compatible session driver (prisma_b200/main.py) on the GPU: a test session with the sp agent against
the restatement, and a short dqn_buffer training session with TensorBoard scalars and model snapshots."""
import os

import numpy as np
import pytest

from oracle import oracle
from tests import common

pytestmark = pytest.mark.gpu

FIX = common.topo_dir("abilene")
TOPO_FLAGS = ["--physical_adjacency_matrix_path=" + os.path.join(FIX, "topology_files", "physical_adjacency_matrix.txt"),
              "--overlay_adjacency_matrix_path=" + os.path.join(FIX, "topology_files", "overlay_adjacency_matrix.txt"),
              "--map_overlay_path=" + os.path.join(FIX, "topology_files", "map_overlay.txt"),
              "--traffic_matrix_root_path=" + os.path.join(FIX, "traffic_matrices")]


def test_sp_test_session_matches_the_restatement(tmp_path, capsys):
    from prisma_b200 import main as session
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    res = session.main(TOPO_FLAGS + ["--train=0", "--agent_type=sp", "--simTime=3", "--num_envs=1", "--load_factor=0.9",
                                     "--activateUnderlayTraffic=1", "--pingAsObs=0", "--session_name=sp_test",
                                     "--logs_parent_folder=" + str(tmp_path)])
    out = capsys.readouterr().out
    assert "Summary of the Episode 0" in out and "Overlay lost ratio" in out
    cfg = common.env_config("abilene", simTime=3, load_factor=0.9, pingAsObs=0, activateUnderlayTraffic=1, train=0, agent_type="sp")
    arrays = cfg.load()
    c = common.oracle_run(cfg, arrays, oracle, policy="table", table=common.sp_table(arrays), env_id=0).counters()
    assert res["total_data_pkts"] == c["overlay_injected"] and res["total_arrived_pkts"] == c["overlay_arrived"]
    assert res["avg_e2e_delay"] == pytest.approx(c["avg_e2e"], rel=1e-6)
    assert res["avg_hops"] == pytest.approx(c["data_hops"] / c["overlay_arrived"])
    acc = EventAccumulator(os.path.join(str(tmp_path), "sp_test", "test_results", "final"))
    acc.Reload()
    ev = acc.Scalars("test_overlay_arrived_pkts")[0]
    assert ev.step == 90 and ev.value == c["overlay_arrived"]


def test_dqn_buffer_training_session(tmp_path, capsys):
    from prisma_b200 import main as session, metrics
    res = session.main(TOPO_FLAGS + ["--train=1", "--agent_type=dqn_buffer", "--simTime=3", "--num_envs=128",
                                     "--load_factor=1.0", "--activateUnderlayTraffic=1", "--pingAsObs=0",
                                     "--batch_size=128", "--training_step=0.25", "--sync_step=1.0", "--lr=0.001",
                                     "--iterationNum=600", "--logging_timestep=0", "--snapshot_interval=2",
                                     "--save_models=1", "--replay_buffer_max_size=32768", "--session_name=train",
                                     "--logs_parent_folder=" + str(tmp_path)])
    assert res["total_data_pkts"] > 128 * 1000 and 0.0 <= res["loss_ratio"] < 0.9 and res["avg_e2e_delay"] > 0   # exploring agents
    out = capsys.readouterr().out
    assert "Total number of Transitions" in out
    root = os.path.join(str(tmp_path), "saved_models", "train")
    assert os.path.isdir(os.path.join(root, "final")) and os.path.isdir(os.path.join(root, "episode_1_step_1"))
    nets = metrics.load_model(os.path.join(root, "final"))
    assert len(nets) == 11 and all(np.isfinite(n["w1"]).all() for n in nets)
    first = metrics.load_model(os.path.join(root, "episode_1_step_1"))
    assert any(not np.array_equal(a["w1"], b["w1"]) for a, b in zip(first, nets))     # the learner kept training
    sess = os.path.join(str(tmp_path), "train")
    assert os.path.isdir(os.path.join(sess, "session")) and os.path.isdir(os.path.join(sess, "td_error", "node_0"))
    # the stored models load into a test session (--load_path)
    res2 = session.main(TOPO_FLAGS + ["--train=0", "--agent_type=dqn_buffer", "--simTime=1", "--num_envs=16",
                                      "--activateUnderlayTraffic=1", "--pingAsObs=0", "--session_name=eval",
                                      "--load_path=" + os.path.join(root, "final"), "--logs_parent_folder=" + str(tmp_path)])
    assert res2["total_arrived_pkts"] > 0


def test_target_signalling_training_session(tmp_path):
    from prisma_b200 import main as session
    res = session.main(TOPO_FLAGS + ["--train=1", "--agent_type=dqn_buffer", "--signaling_type=target", "--simTime=2",
                                     "--num_envs=64", "--activateUnderlayTraffic=1", "--pingAsObs=0", "--batch_size=128",
                                     "--training_step=0.25", "--lr=0.001", "--iterationNum=400", "--gamma=0.95",
                                     "--session_name=target", "--logs_parent_folder=" + str(tmp_path)])
    assert res["total_data_pkts"] > 64 * 500 and res["signaling_overhead"] > 0      # 24-byte target packets per hop


def test_prioritized_target_training_session(tmp_path):
    from prisma_b200 import main as session
    res = session.main(TOPO_FLAGS + ["--train=1", "--agent_type=dqn_buffer", "--signaling_type=target",
                                     "--prioritizedReplayBuffer=1", "--simTime=2", "--num_envs=64", "--pingAsObs=0",
                                     "--activateUnderlayTraffic=1", "--batch_size=128", "--training_step=0.25", "--lr=0.001",
                                     "--iterationNum=400", "--session_name=prio", "--logs_parent_folder=" + str(tmp_path)])
    assert res["total_data_pkts"] > 64 * 500 and res["avg_e2e_delay"] > 0
