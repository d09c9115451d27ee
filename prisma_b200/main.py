"""Flag-compatible session driver: `python -m prisma_b200.main [options]` takes the options of
prisma/main.py (prisma/source/argument_parser.py:33-98, same names and defaults) and runs the session on the
device-side environment instead of ns-3 + ZMQ + one Python thread per node:

* `--train=0`, or an agent without learning (`sp`, `dqn_routing` loaded as a next-hop table; `opt` is refused: it is a
  caller-side Gym-mode hook, `prisma_b200.agents.OptAgent`): every
  environment of the batch runs `simTime` seconds inside `prb_run`; the summary of prisma/main.py:152-172 is
  printed and the `stats_writer_test` scalars are written (prisma/source/tb_logger.py:145-164).
* `--train=1 --agent_type=dqn_buffer`: the Gym loop runs on the device (`prb_step_qnet`: Q-network forward,
  eps-greedy argmin, replay rows into the HBM rings); every `training_step` simulated seconds the learner takes
  one gradient step per node from the rings (prisma/source/trainer.py:28-99), the target networks are
  synchronised every `sync_step` seconds (ideal signalling: prisma/source/learner.py:256-270), the
  `stats_writer_train` scalars are written every `logging_timestep` seconds of wall clock, snapshots every
  `snapshot_interval` simulated seconds, the final models when `--save_models=1`
  (prisma/main.py:143-184, prisma/source/utils.py:14-66).

Extra options of this driver: `--num_envs` (environments in the batch, default 1 = the reference's single
simulation) and `--device`.  Options that only concern the ns-3 process (`--basePort`, `--ns3_sim_path`,
`--start_tensorboard`, `--tensorboard_port`, `--profile_session`) are accepted and ignored."""
import argparse
import datetime
import json
import os
import sys
import time

import numpy as np

from .config import EnvConfig
from . import metrics

AGENTS = ["dqn_buffer", "dqn_routing", "dqn_buffer_fp", "dqn_buffer_lite", "dqn_buffer_lighter", "dqn_buffer_lighter_2",
          "dqn_buffer_lighter_3", "dqn_buffer_ff", "dqn_buffer_with_throughputs", "sp", "opt"]


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="python -m prisma_b200.main", allow_abbrev=False,
                                description="PRISMA session on the B200 environment (flags of prisma/main.py)")
    g = p.add_argument_group("Global simulation arguments")
    g.add_argument("--numEpisodes", type=int, default=1)
    g.add_argument("--simTime", type=int, default=60)
    g.add_argument("--basePort", type=int, default=6555)
    g.add_argument("--seed", type=int, default=100)
    g.add_argument("--train", type=int, default=1)
    g.add_argument("--max_nb_arrived_pkts", type=int, default=-1)
    g.add_argument("--ns3_sim_path", type=str, default="../ns3-gym/")
    g.add_argument("--signalingSim", type=int, default=0)
    g.add_argument("--activateOverlay", type=int, default=1)
    g.add_argument("--movingAverageObsSize", type=int, default=5)
    g.add_argument("--activateUnderlayTraffic", type=int, default=0)
    g.add_argument("--map_overlay_path", type=str, default="examples/abilene/topology_files/map_overlay.txt")
    g.add_argument("--pingAsObs", type=int, default=1)
    g.add_argument("--pingPacketIntervalTime", type=float, default=0.2)
    g = p.add_argument_group("Network parameters")
    g.add_argument("--load_factor", type=float, default=1.0)
    g.add_argument("--topology_name", type=str, default="abilene")
    g.add_argument("--physical_adjacency_matrix_path", type=str,
                   default="examples/abilene/topology_files/physical_adjacency_matrix.txt")
    g.add_argument("--overlay_adjacency_matrix_path", type=str,
                   default="examples/abilene/topology_files/overlay_adjacency_matrix.txt")
    g.add_argument("--traffic_matrix_root_path", type=str, default="examples/abilene/traffic_matrices/")
    g.add_argument("--traffic_matrix_index", type=int, default=0)
    g.add_argument("--max_out_buffer_size", type=int, default=16260)
    g.add_argument("--link_delay", type=int, default=1)
    g.add_argument("--packet_size", type=int, default=512)
    g.add_argument("--link_cap", type=int, default=500000)
    g = p.add_argument_group("Storing session logs arguments")
    g.add_argument("--session_name", type=str, default=None)
    g.add_argument("--logs_parent_folder", type=str, default="examples/abilene/results/")
    g.add_argument("--logging_timestep", type=int, default=5)
    g.add_argument("--profile_session", type=int, default=0)
    g = p.add_argument_group("DRL Agent arguments")
    g.add_argument("--agent_type", choices=AGENTS, type=str, default="dqn_buffer")
    g.add_argument("--signaling_type", type=str, choices=["NN", "target", "ideal"], default="ideal")
    g.add_argument("--lr", type=float, default=1e-4)
    g.add_argument("--bigSignalingSize", type=int, default=512)
    g.add_argument("--prioritizedReplayBuffer", type=int, default=0)
    g.add_argument("--smart_exploration", type=int, default=0)
    g.add_argument("--batch_size", type=int, default=512)
    g.add_argument("--gamma", type=float, default=1)
    g.add_argument("--iterationNum", type=int, default=3000)
    g.add_argument("--exploration_initial_eps", type=float, default=1.0)
    g.add_argument("--exploration_final_eps", type=float, default=0.1)
    g.add_argument("--load_path", type=str, default=None)
    g.add_argument("--save_models", type=int, default=0)
    g.add_argument("--snapshot_interval", type=int, default=0)
    g.add_argument("--training_step", type=float, default=0.05)
    g.add_argument("--sync_step", type=float, default=1.0)
    g.add_argument("--sync_ratio", type=float, default=0.1)
    g.add_argument("--replay_buffer_max_size", type=int, default=50000)
    g.add_argument("--loss_penalty_type", type=str, choices=["None", "fixed"], default="fixed")
    g = p.add_argument_group("Other parameters")
    g.add_argument("--start_tensorboard", type=int, default=0)
    g.add_argument("--tensorboard_port", type=int, default=16666)
    g = p.add_argument_group("prisma_b200")
    g.add_argument("--num_envs", type=int, default=1, help="independent environments simulated side by side")
    g.add_argument("--device", type=int, default=0)
    return p


def parse_arguments(argv=None) -> dict:
    """The `params` dict of prisma/source/argument_parser.py:100-164 (absolute paths, session name and log folders)."""
    params = vars(build_parser().parse_args(argv))
    # combinations this driver has no device path for are refused here, with the way out, instead of failing in set_agent
    if params["agent_type"] == "opt":
        raise SystemExit("--agent_type opt: the optimal-solution agent is a per-(source, destination) probabilistic choice "
                         "(utils.py:161-199) evaluated by the caller in Gym mode (prisma_b200.agents.OptAgent); this driver "
                         "runs table and dqn_buffer agents only")
    if params["agent_type"] == "dqn_routing" and not params["load_path"] and not params["train"]:
        raise SystemExit("--agent_type dqn_routing --train 0 needs --load_path (a routing_table.txt: a frozen destination-only "
                         "network is a next-hop table); with --train 1 it is trained on the device")
    for k in ("logs_parent_folder", "physical_adjacency_matrix_path", "overlay_adjacency_matrix_path", "map_overlay_path"):
        params[k] = os.path.abspath(params[k])
    if params["load_path"]:
        params["load_path"] = os.path.abspath(params["load_path"])
    params["traffic_matrix_path"] = os.path.abspath("%s/node_intensity_normalized_%d.txt" % (
        params["traffic_matrix_root_path"].rstrip("/"), params["traffic_matrix_index"]))
    if params["session_name"] is None:
        params["session_name"] = datetime.datetime.now().strftime("%Y%m%d-%H%M%S")
    params["logs_folder"] = os.path.join(params["logs_parent_folder"], params["session_name"])
    return params


def summary_text(episode: int, s: metrics.SessionStats, iterations: int, transitions: int) -> str:
    """prisma/main.py:152-172."""
    ratio = lambda a, b: a / b if b else float("nan")
    return "\n".join([
        " Summary of the Episode %d:" % episode,
        "    Simulation time = %s," % s.curr_time,
        "    Total Iterations = %d," % iterations,
        "    Total number of Transitions = %d," % transitions,
        "    Overlay Total injected packets = %d," % s.sim_injected_packets,
        "    Global Total injected packets = %d," % s.sim_global_injected_packets,
        "    Overlay arrived packets = %d," % s.sim_delivered_packets,
        "    Global arrived packets = %d," % s.sim_global_delivered_packets,
        "    Overlay lost packets = %d," % s.sim_dropped_packets,
        "    Global lost packets = %d," % s.sim_global_dropped_packets,
        "    Overlay buffered packets = %d," % s.sim_buffered_packets,
        "    Global buffered packets = %d," % s.sim_global_buffered_packets,
        "    Overlay lost ratio = %s," % ratio(s.sim_dropped_packets, s.sim_injected_packets),
        "    Global lost ratio = %s," % ratio(s.sim_global_dropped_packets, s.sim_global_injected_packets),
        "    Overlay e2e delay = %s," % s.sim_avg_e2e_delay,
        "    Global e2e delay = %s," % s.sim_global_avg_e2e_delay,
        "    Overlay Cost = %s," % s.sim_cost,
        "    Global Cost = %s," % s.sim_global_cost,
        "    Hops = %s," % ratio(s.total_hops, s.sim_delivered_packets),
        "    OverheadRatio = %s" % s.sim_signaling_overhead])


def qnet_variants():
    """Agent types whose Q-network the device kernels serve (prisma_b200/qnet.py: VARIANTS)."""
    from . import qnet
    return tuple(qnet.VARIANTS)


def overlay_neighbor_indices(env):
    u2o = {int(u): i for i, u in enumerate(env.overlay_to_underlay)}
    return [[u2o[int(u)] for u in nb] for nb in env.neighbors]


def run_fixed_agent(env, params, episode):
    """sp / a loaded dq_routing table: the agent hook is evaluated on the device inside prb_run."""
    env.reset()
    table = None
    if params["agent_type"] == "dqn_routing" and params["load_path"]:
        table = np.loadtxt(os.path.join(params["load_path"], "routing_table.txt"), dtype=np.int32)
    env.set_agent(params["agent_type"] if table is None else None, table=table)
    env.run()
    c = env.counters()
    env.check_errors(c)
    s = metrics.SessionStats.from_counters(c)
    return s, int(s.env_steps), 0


def run_dqn_session(env, params, episode, learner, writer, state):
    """One training (or evaluation) episode of the dqn_buffer agents with the Gym loop on the device."""
    import torch
    train = bool(params["train"])
    env.reset()
    env.qnet_load(learner.weights())
    eps_schedule = metrics.LinearSchedule(params["iterationNum"], params["exploration_final_eps"],
                                          params["exploration_initial_eps"])
    sim_time = float(params["simTime"])
    next_train = float(params["training_step"])
    next_sync = float(params["sync_step"]) if params["sync_step"] > 0 else float("inf")
    next_snapshot = float(params["snapshot_interval"]) if params["snapshot_interval"] > 0 else float("inf")
    last_log = time.time()
    # Gym steps per prb_step_qnet call: about one training interval of simulated time for the whole batch
    stride = 64
    td = None
    while True:
        eps = eps_schedule.value(state["iterations"]) if train else 0.0
        batch = env.step_qnet(stride, eps=eps, seed=params["seed"] + state["calls"])
        state["calls"] += 1
        state["iterations"] += stride
        torch.cuda.synchronize()
        rec_t = batch.raw.view(torch.int64)[:, 0]
        valid = batch.valid != 0
        if not bool(valid.any()):
            break
        now = float(rec_t[valid].min()) * 1e-9             # the slowest environment of the batch
        if train:
            while now >= next_train:
                td = learner.train_on_env(env, batch_size=params["batch_size"], seed=params["seed"] + state["grad_steps"])
                state["grad_steps"] += 1
                next_train += float(params["training_step"])
            if now >= next_sync:
                learner.sync_targets()
                next_sync += float(params["sync_step"])
            if state["base_time"] + now >= next_snapshot:
                metrics.save_all_models(learner.weights(), learner.m, learner.k_max, params["session_name"],
                                        t=state["snapshots"] + 1, num_episodes=1,
                                        root=os.path.join(params["logs_parent_folder"], "saved_models"), snapshot=True)
                state["snapshots"] += 1
                next_snapshot += float(params["snapshot_interval"])
        if writer is not None and train and time.time() - last_log >= params["logging_timestep"]:
            s = metrics.SessionStats.from_counters(env.counters())
            counts, _ = env.replay_counts()
            writer.write_train(s, state["iterations"], base_time=state["base_time"], td_errors=td, epsilon=eps,
                               replay_lengths=[int(min(x, params["replay_buffer_max_size"])) for x in counts])
            last_log = time.time()
        if params["max_nb_arrived_pkts"] > 0 and int(env.counters()["overlay_arrived"].sum()) >= params["max_nb_arrived_pkts"]:
            break
        if now >= sim_time:
            break
    c = env.counters()
    env.check_errors(c)
    s = metrics.SessionStats.from_counters(c)
    counts, _ = env.replay_counts() if train else (np.zeros(learner.m), 0)
    if writer is not None and train:
        writer.write_train(s, state["iterations"], base_time=state["base_time"], td_errors=td,
                           replay_lengths=[int(min(x, params["replay_buffer_max_size"])) for x in counts])
    return s, state["iterations"], int(np.sum(counts))


def main(argv=None) -> dict:
    params = parse_arguments(argv)
    from .env import BatchedRoutingEnv
    cfg = EnvConfig.from_params(params)
    env = BatchedRoutingEnv(cfg)
    os.makedirs(params["logs_folder"], exist_ok=True)
    with open(os.path.join(params["logs_folder"], "params.json"), "w") as f:
        json.dump({k: v for k, v in params.items() if isinstance(v, (int, float, str, type(None)))}, f, indent=1)
    writer = metrics.StatsWriter(params["logs_parent_folder"], params["session_name"])
    # Q-network agents run (and learn) on the device; a dqn_routing agent that only replays a loaded table does not need one
    learning = params["agent_type"].startswith("dqn_buffer") or (params["agent_type"] == "dqn_routing" and bool(params["train"]))
    learner = None
    state = {"iterations": 0, "calls": 0, "grad_steps": 0, "snapshots": 0, "base_time": 0.0}
    if learning:
        if params["agent_type"] not in qnet_variants():
            raise SystemExit("device-side agents exist for %s; %s is served through prb_step with the caller's own network"
                             % (", ".join(qnet_variants()), params["agent_type"]))
        if params["prioritizedReplayBuffer"] and params["train"] and params["signaling_type"] != "target":
            print("note: the reference feeds its prioritized buffer from `target` signalling only (forwarder.py:494-505); "
                  "with `%s` the weights stay 1" % params["signaling_type"], file=sys.stderr)
        from . import learner as learner_mod, qnet
        nbrs = overlay_neighbor_indices(env)
        if os.environ.get("PRB_LEARNER", "device") == "torch":
            # round 1's learner: PyTorch autograd above the environment (kept for A/B runs)
            learner = learner_mod.DQNLearner(nbrs, lr=params["lr"], gamma=params["gamma"], seed=params["seed"],
                                             device="cuda:%d" % params["device"], signaling_type=params["signaling_type"])
            if params["load_path"]:
                nets = metrics.load_model(params["load_path"])
                import torch
                w = torch.from_numpy(qnet.pack_weights(nets)).view(learner.m, -1).to(learner.W.device)
                with torch.no_grad():
                    learner.W.copy_(w)
                learner.sync_targets()
        else:
            # the gradient step on the device (prb_learner_*): no autograd, no per-node Python loop
            degrees = [len(n) for n in nbrs]
            nets = metrics.load_model(params["load_path"]) if params["load_path"] else \
                qnet.init_weights(len(nbrs), degrees, max(degrees), seed=params["seed"], variant=params["agent_type"])
            learner = learner_mod.DeviceLearner(env, lr=params["lr"], gamma=params["gamma"], signaling_type=params["signaling_type"],
                                                weights=qnet.pack_weights(nets), variant=params["agent_type"])
        if params["train"]:
            env.replay_enable(rows_per_node=int(params["replay_buffer_max_size"]))
            if params["signaling_type"] == "target":
                # the next node computes the target with its own network when the packet reaches it (trainer.py:52-55)
                env.replay_set_target_mode(True, params["gamma"])
            if params["prioritizedReplayBuffer"]:
                env.replay_enable_prioritized(alpha=1.0)       # agent.py:67: PrioritizedReplayBuffer (size, 1, ...)
    stats = None
    for episode in range(params["numEpisodes"]):
        t0 = time.time()
        if learning:
            stats, iters, transitions = run_dqn_session(env, params, episode, learner, writer, state)
        else:
            stats, iters, transitions = run_fixed_agent(env, params, episode)
        print(summary_text(episode, stats, iters, transitions))
        print("    (%d environments, %.2f s wall clock)" % (env.num_envs, time.time() - t0))
        state["base_time"] += stats.curr_time
    if params["train"] == 0 or not learning:
        writer.write_test(stats, params["load_factor"], model_version=os.path.basename(params["load_path"] or "final"))
    if learning and params["save_models"]:
        metrics.save_all_models(learner.weights(), learner.m, learner.k_max, params["session_name"], t=1, num_episodes=1,
                                root=os.path.join(params["logs_parent_folder"], "saved_models"), snapshot=False)
    writer.close()
    env.close()
    return stats.results()


if __name__ == "__main__":
    print(json.dumps(main()))
