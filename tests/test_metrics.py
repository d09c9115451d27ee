"""Session statistics, TensorBoard scalars (names of prisma/source/tb_logger.py:70-164), model snapshots (folder
layout of prisma/source/utils.py:14-99) and the flag parser of the session driver (prisma/source/argument_parser.py)."""
import os

import numpy as np
import pytest

from prisma_b200 import _abi, main as session, metrics, qnet


def counters(n=3):
    c = np.zeros(n, dtype=_abi.COUNTERS_DTYPE)
    c["overlay_injected"], c["overlay_arrived"], c["overlay_lost"] = [100, 200, 300], [80, 150, 250], [5, 10, 20]
    c["underlay_injected"], c["underlay_arrived"], c["underlay_lost"] = [10, 20, 30], [8, 15, 25], [1, 2, 3]
    c["n_e2e"], c["sum_e2e"] = c["overlay_arrived"], [8.0, 30.0, 75.0]
    c["n_cost"], c["sum_cost"] = c["overlay_arrived"] + c["overlay_lost"], [17.0, 64.0, 135.0]
    c["n_underlay_e2e"], c["sum_underlay_e2e"] = c["underlay_arrived"], [0.8, 3.0, 7.5]
    c["n_underlay_cost"], c["sum_underlay_cost"] = c["underlay_arrived"] + c["underlay_lost"], [0.9, 3.4, 8.4]
    c["bytes_data"], c["bytes_signaling"] = [54000, 108000, 162000], [5400, 10800, 16200]
    c["data_hops"], c["env_steps"], c["end_ns"] = [240, 450, 750], [300, 500, 800], 2_000_000_000
    return c


def test_session_stats_follow_the_forwarder_bookkeeping():
    s = metrics.SessionStats.from_counters(counters())
    assert (s.sim_injected_packets, s.sim_delivered_packets, s.sim_dropped_packets) == (600, 480, 35)
    assert s.sim_buffered_packets == 600 - 480 - 35
    # forwarder.py:280-283: global = underlay + overlay
    assert s.sim_global_injected_packets == 660 and s.sim_global_delivered_packets == 528 and s.sim_global_dropped_packets == 41
    assert s.sim_avg_e2e_delay == pytest.approx(113.0 / 480)
    assert s.sim_cost == pytest.approx(216.0 / 515)
    # forwarder.py:285-286: delivery-weighted mean of the underlay and overlay averages
    assert s.sim_global_avg_e2e_delay == pytest.approx((11.3 + 113.0) / 528)
    assert s.sim_signaling_overhead == pytest.approx(0.1)
    assert s.curr_time == 2.0
    one = metrics.SessionStats.from_counters(counters()[:1])
    assert one.sim_avg_e2e_delay == pytest.approx(0.1) and one.results()["avg_hops"] == pytest.approx(3.0)


def test_tensorboard_scalars_carry_the_reference_tags(tmp_path):
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    s = metrics.SessionStats.from_counters(counters())
    w = metrics.StatsWriter(str(tmp_path), "sess")
    w.write_train(s, iteration=1234, base_time=60.0, td_errors=[0.5, None, 0.25], epsilon=0.4, replay_lengths=[10, 20, 30])
    w.write_test(s, load_factor=0.9, model_version="final")
    w.close()

    def tags(sub):
        acc = EventAccumulator(os.path.join(str(tmp_path), "sess", sub))
        acc.Reload()
        return acc, set(acc.Tags()["scalars"])
    acc, t = tags("session")
    for base in ("total_e2e_delay", "total_rewards_with_loss", "loss_ratio", "total_hops", "avg_hops", "ma_avg_hops",
                 "nb_buffered_pkts", "avg_cost", "avg_delay", "ma_delays"):
        assert base + "_over_iterations" in t and base + "_over_time" in t, base
    assert {"signalling ratio", "sim_second_per_real_seconds", "exploaration_value_over_time"} <= t
    ev = acc.Scalars("loss_ratio_over_iterations")[0]
    assert ev.step == 1234 and ev.value == pytest.approx(35 / 600)
    assert acc.Scalars("avg_delay_over_time")[0].step == int(62.0 * 1e6)       # (base_curr_time + curr_time) * 1e6
    for sub, v in (("nb_arrived_pkts", 480), ("nb_lost_pkts", 35), ("nb_new_pkts", 600)):
        a, t = tags(sub)
        assert t == {"pkts_over_iterations", "pkts_over_time"} and a.Scalars("pkts_over_time")[0].value == v
    a, t = tags("td_error/node_2")
    assert a.Scalars("MSE_loss_over_time")[0].value == pytest.approx(0.25)
    assert not os.path.exists(os.path.join(str(tmp_path), "sess", "td_error", "node_1"))
    a, t = tags("test_results/final")
    assert {"test_global_injected_pkts", "test_overlay_injected_pkts", "test_global_lost_pkts", "test_overlay_lost_pkts",
            "test_global_arrived_pkts", "test_overlay_arrived_pkts", "test_global_e2e_delay", "test_overlay_e2e_delay",
            "test_global_loss_rate", "test_overlay_loss_rate", "test_global_cost", "test_overlay_cost"} == t
    assert a.Scalars("test_overlay_loss_rate")[0].step == 90


def test_model_snapshots_round_trip_in_the_reference_layout(tmp_path):
    m, degrees = 4, [2, 3, 3, 2]
    k_max = max(degrees)
    nets = qnet.init_weights(m, degrees, k_max, seed=3)
    flat = qnet.pack_weights(nets)
    root = str(tmp_path / "saved_models")
    final = metrics.save_all_models(flat, m, k_max, "sess", root=root)
    snap = metrics.save_all_models(flat, m, k_max, "sess/", t=7, num_episodes=1, root=root, snapshot=True)
    assert final == os.path.join(root, "sess", "final") and snap == os.path.join(root, "sess", "episode_1_step_7")
    assert sorted(os.listdir(final)) == ["node%d.npz" % i for i in range(m)]
    back = metrics.load_model(final)
    assert np.array_equal(qnet.pack_weights(back), flat)
    only = metrics.load_model(snap, node_index=2)
    assert only[0] is None and np.array_equal(only[2]["w3"], nets[2]["w3"])


def test_linear_schedule():
    s = metrics.LinearSchedule(3000, 0.1, 1.0)
    assert s.value(0) == 1.0 and s.value(1500) == pytest.approx(0.55) and s.value(10 ** 6) == pytest.approx(0.1)


def test_driver_takes_the_reference_flags():
    """Every option of prisma/source/argument_parser.py:33-98 with its default; the derived paths of :100-118."""
    p = session.parse_arguments([])
    ref_defaults = dict(numEpisodes=1, simTime=60, basePort=6555, seed=100, train=1, max_nb_arrived_pkts=-1, signalingSim=0,
                        activateOverlay=1, movingAverageObsSize=5, activateUnderlayTraffic=0, pingAsObs=1,
                        pingPacketIntervalTime=0.2, load_factor=1.0, traffic_matrix_index=0, max_out_buffer_size=16260,
                        link_delay=1, packet_size=512, link_cap=500000, logging_timestep=5, profile_session=0,
                        agent_type="dqn_buffer", signaling_type="ideal", lr=1e-4, bigSignalingSize=512,
                        prioritizedReplayBuffer=0, smart_exploration=0, batch_size=512, gamma=1, iterationNum=3000,
                        exploration_initial_eps=1.0, exploration_final_eps=0.1, load_path=None, save_models=0,
                        snapshot_interval=0, training_step=0.05, sync_step=1.0, sync_ratio=0.1,
                        replay_buffer_max_size=50000, loss_penalty_type="fixed", start_tensorboard=0, tensorboard_port=16666)
    for k, v in ref_defaults.items():
        assert p[k] == v, k
    assert os.path.isabs(p["physical_adjacency_matrix_path"]) and p["traffic_matrix_path"].endswith(
        "examples/abilene/traffic_matrices/node_intensity_normalized_0.txt")
    assert p["logs_folder"] == os.path.join(p["logs_parent_folder"], p["session_name"])
    q = session.parse_arguments(["--train=0", "--agent_type=sp", "--simTime=5", "--num_envs=64", "--load_factor=0.7",
                                 "--session_name=s1"])
    assert (q["train"], q["agent_type"], q["simTime"], q["num_envs"], q["session_name"]) == (0, "sp", 5, 64, "s1")
    with pytest.raises(SystemExit):
        session.parse_arguments(["--agent_type=unknown"])
