#!/usr/bin/env python3
"""Headline benchmark: packet-hops/s of the batched environment step loop (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W]              # this repo's CUDA path
  python bench.py --impl reference [--gpus N] [--steps K] ...      # the reference's own CPU code, host cores

Workload (BASELINE.json configs[1]): Abilene 11-node underlay from the reference's examples, frozen
dq_routing-style agent (a next-hop table indexed by (node, destination); the sp-pretrained table), 4096
parallel environments per GPU, load_factor 0.9, seeded synthetic Poisson traffic (Philox), pings on.
One "step" advances every environment by --chunk simulated seconds (one k_advance launch).  Multi-GPU:
environments are independent, each rank owns 4096 of them, no collective in the step ("weak" scaling);
torch.distributed is used for the barrier and the max-over-ranks / sum reductions of the report only.

Timing: one episode of the reference's length (60 simulated seconds); its last K steps are timed, each bracketed
by CUDA events on the launch stream with the hand-over check inside the pair, everything before is untimed
warm-up (>= W steps); L2 flushed (256 MB memset) between steps outside the event pairs; max over ranks of the
summed step times.  The e2e leg runs a second, identical episode through the host-buffer calls and times the
same K steps by wall clock.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FIX = os.path.join(ROOT, "tests", "fixtures")
METRIC = "packet_hops_per_s"
UNIT = "packet-hops/s"
B_HOP_BASE = 144      # algorithmic bytes per packet hop = 144 + 4*(1+k)


def workload_cfg(envs, sim_time, topo="abilene", load_factor=0.9):
    from prisma_b200.config import EnvConfig
    d = os.path.join(FIX, topo)
    return EnvConfig(physical_adjacency_matrix_path=os.path.join(d, "topology_files", "physical_adjacency_matrix.txt"),
                     overlay_adjacency_matrix_path=os.path.join(d, "topology_files", "overlay_adjacency_matrix.txt"),
                     map_overlay_path=os.path.join(d, "topology_files", "map_overlay.txt"),
                     traffic_matrix_root_path=os.path.join(d, "traffic_matrices"),
                     num_envs=envs, simTime=sim_time, load_factor=load_factor, pingAsObs=0, activateUnderlayTraffic=1,
                     agent_type="dqn_routing", seed=100)


def config_dict(args, world):
    return {"workload": "abilene_11node_dq_routing_table_%denvs_per_gpu_lf0.9" % args.envs,
            "topology": "prisma/examples/abilene (TM 0)", "agent": "dq_routing frozen as next-hop table (sp-pretrained)",
            "envs_per_gpu": args.envs, "total_envs": args.envs * world, "load_factor": 0.9,
            "sim_seconds_per_step": args.chunk, "link_cap": 500000, "link_delay_ms": 1, "packet_size": 512,
            "max_out_buffer_size": 16260, "traffic": "Philox-4x32-10 Poisson per flow",
            "episode_sim_s": max(args.episode, (args.warmup + args.steps) * args.chunk),
            "timed_window_sim_s": [max(args.episode, (args.warmup + args.steps) * args.chunk) - args.steps * args.chunk,
                                   max(args.episode, (args.warmup + args.steps) * args.chunk)],
            "l2": "flushed between steps (256 MB memset outside the event pairs)", "parallelism": "env-sharded x%d" % world}


class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "20"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        lines = []
        if self.p is not None:
            time.sleep(0.15)
            self.p.terminate()
            try:
                out, _ = self.p.communicate(timeout=5)
                lines = out.strip().splitlines()
            except Exception:
                self.p.kill()
        if not lines:
            try:
                lines = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                        "--format=csv,noheader,nounits"], stdout=subprocess.PIPE, text=True,
                                       timeout=10).stdout.strip().splitlines()
            except Exception:
                lines = []
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in lines:
            t = [x.strip() for x in ln.split(",")]
            if len(t) < 6:
                continue
            try:
                sm.append(float(t[0]))
                mx.append(float(t[1]))
            except ValueError:
                continue
            for n, v in zip(names, t[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(args, kernel_name):
    """dram bytes per launch of the dominant kernel from the committed ncu capture of this exact workload."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    try:
        j = json.load(open(p))
        j = j.get(kernel_name, j)
        if j.get("envs") == args.envs and abs(j.get("chunk", -1) - args.chunk) < 1e-12:
            return float(j["dram_bytes_per_launch"])
    except Exception:
        pass
    return None


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own ns-3 program on the host cores
# ------------------------------------------------------------------------------------------------

def ref_sample(sim_time, procs, load_factor=0.9):
    """One prisma_ref process per core, each one environment for `sim_time` simulated seconds of the same
    workload (sp table agent evaluated in-process: no ZMQ, no Python in the loop).  Returns
    (hops, env_steps, wall_seconds, kind)."""
    from oracle import oracle, run_ref
    from prisma_b200.env import sp_table
    cfg = workload_cfg(1, sim_time, load_factor=load_factor)
    arrays = cfg.load()
    table = sp_table(arrays)
    if run_ref.available():
        tmp = tempfile.mkdtemp(prefix="prb_ref_")
        rej = os.path.join(tmp, "rej.txt")
        np.savetxt(rej, np.zeros((arrays.n_nodes, arrays.n_nodes)), fmt="%.6f")
        files = {"adj": cfg.physical_adjacency_matrix_path, "overlay": cfg.overlay_adjacency_matrix_path,
                 "map": cfg.map_overlay_path, "rej": rej,
                 "tm": os.path.join(cfg.traffic_matrix_root_path, "node_intensity_normalized_0.txt")}
        res = [None] * procs

        def work(i):
            out = os.path.join(tmp, "o%d" % i)
            run_ref.run(files, out, sim_time=sim_time, seed=100 + i, load_factor=load_factor, policy="table",
                        policy_data=table, events=False, loss_penalty=cfg.derived_loss_penalty(arrays.n_overlay),
                        activate_underlay_traffic=True, ping_as_obs=False)
            res[i] = run_ref.parse_summary(out)
        th = [threading.Thread(target=work, args=(i,)) for i in range(procs)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        dt = time.perf_counter() - t0
        hops = sum(r["packet_hops"] for r in res)
        steps = sum(r["env_steps"] for r in res)
        return hops, steps, dt, "reference"
    # the reference binary did not travel: time the C restatement instead (ctypes releases the GIL)
    from tests import common
    res = [None] * procs

    def work(i):
        p = common.oracle_params(cfg, arrays, oracle, env_id=i, inject_mode=1)
        sim = oracle.OracleSim(p, log_events=False)
        sim.set_policy("table", table)
        sim.run()
        res[i] = sim.counters()
    th = [threading.Thread(target=work, args=(i,)) for i in range(procs)]
    t0 = time.perf_counter()
    for t in th:
        t.start()
    for t in th:
        t.join()
    dt = time.perf_counter() - t0
    return sum(r["packet_hops"] for r in res), sum(r["env_steps"] for r in res), dt, "port"


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sim_time = args.ref_sim_time
    for _ in range(args.warmup if args.warmup < 2 else 1):
        ref_sample(min(sim_time, 2.0), cores)
    hops = steps = 0
    total = 0.0
    kind = "reference"
    for _ in range(args.steps):
        h, s, dt, kind = ref_sample(sim_time, cores)
        hops += h
        steps += s
        total += dt
    value = hops / total
    sample = "%d steps x %d processes (one per core) x 1 env x %g simulated s, Abilene lf 0.9, sp table agent in-process" % (
        args.steps, cores, sim_time)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000.0 * total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": config_dict(args, world),
            "env_steps_per_s": steps / total,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# secondary legs: BASELINE.json configs[2], [3], [4] at this rank's share, and the learner's gradient exchange
# ------------------------------------------------------------------------------------------------

def _timed(torch, fn):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / 1000.0


def _reduce(dist, torch, world, sums, maxes):
    """Whole-job figures: counts are summed over the ranks, times are the slowest rank's."""
    if world > 1:
        s = torch.tensor(sums, dtype=torch.float64, device="cuda")
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        m = torch.tensor(maxes, dtype=torch.float64, device="cuda")
        dist.all_reduce(m, op=dist.ReduceOp.MAX)
        return [float(x) for x in s], [float(x) for x in m]
    return list(sums), list(maxes)


def leg_table_run(torch, dist, rank, local_rank, world, topo, envs, load_factor, lead_s, timed_s, peak, extra=None):
    """prb_run with the shortest-path table on `envs` environments per GPU: `lead_s` untimed simulated seconds (queues
    fill), then `timed_s` one-second steps timed on the device with the hand-over check inside the timed region."""
    from prisma_b200.env import BatchedRoutingEnv, sp_table
    cfg = workload_cfg(envs, lead_s + timed_s, topo=topo, load_factor=load_factor)
    for k, v in (extra or {}).items():
        setattr(cfg, k, v)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, device=local_rank, arrays=arrays, env_id_base=rank * envs)
    env.set_policy("table", sp_table(arrays))
    env.run(until_s=float(lead_s))
    c0 = env.counters()

    def go():
        for t in range(int(timed_s)):
            env.run(until_s=float(lead_s + t + 1))
            env.handovers()
    dt = _timed(torch, go)
    c1 = env.counters()
    env.check_errors(c1)
    engine = env.engine
    handed, replay_ms = env.handover_stats() if engine == "parallel" else (0, 0.0)
    kdeg = max(len(x) for x in arrays.overlay_neighbors())
    d = lambda k: float(c1[k].sum() - c0[k].sum())
    lost = float(c1["overlay_lost"].sum()) / max(float(c1["overlay_injected"].sum()), 1.0)
    (hops, steps, handed), (dt,) = _reduce(dist, torch, world, [d("packet_hops"), d("env_steps"), handed], [dt])
    env.close()
    b_hop = B_HOP_BASE + 4 * (1 + kdeg)
    return {"topology": topo, "nodes": int(arrays.n_nodes), "envs_per_gpu": envs, "total_envs": envs * world,
            "load_factor": load_factor, "engine": engine, "timed_window_sim_s": [lead_s, lead_s + timed_s],
            "packet_hops_per_s": hops / dt, "env_steps_per_s": steps / dt, "overlay_loss": lost,
            "envs_replayed_by_the_serial_engine": int(handed),
            "roofline_frac": hops / dt * b_hop / 1e9 / (peak * world), "bytes_per_packet_hop": b_hop}


def leg_device_agents(torch, dist, rank, local_rank, world, topo, envs, load_factor, calls, peak, extra=None):
    """The Gym loop with dqn_buffer agents on the device (prb_step_qnet) and the replay rows written to the HBM rings."""
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = workload_cfg(envs, 60.0, topo=topo, load_factor=load_factor)
    cfg.agent_type = "dqn_buffer"
    for k, v in (extra or {}).items():
        setattr(cfg, k, v)
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, device=local_rank, arrays=arrays, env_id_base=rank * envs)
    deg = [len(n) for n in env.neighbors]
    env.qnet_load(qnet.pack_weights(qnet.init_weights(len(deg), deg, max(deg), seed=1)))
    env.replay_enable(rows_per_node=1 << 18)
    env.step_qnet(calls // 2, eps=0.05, seed=rank)
    c0 = env.counters()
    dt = _timed(torch, lambda: env.step_qnet(calls, eps=0.05, seed=1000 + rank))
    c1 = env.counters()
    env.check_errors(c1)
    d = lambda k: float(c1[k].sum() - c0[k].sum())
    stored = float(env.replay_counts()[0].sum())
    (steps, hops, stored), (dt,) = _reduce(dist, torch, world, [d("env_steps"), d("packet_hops"), stored], [dt])
    env.close()
    return {"topology": topo, "envs_per_gpu": envs, "total_envs": envs * world, "load_factor": load_factor,
            "agent": "dqn_buffer on the device (Q-network forward + eps-greedy argmin per step), pingAsObs=%d" % cfg.pingAsObs,
            "engine": "event-serial (k_advance, Gym mode)", "gym_calls": calls, "agent_steps_per_s": steps / dt,
            "packet_hops_per_s": hops / dt, "replay_rows_stored": int(stored)}

def leg_learner(torch, dist, rank, local_rank, world, envs, iters):
    """Data-parallel DQN training as trainer.py:45-99 runs it, on this rank's shard: collect with the device-side agents
    -> sample from the HBM replay rings -> TD targets, forward, backward and Adam as kernels of the library
    (prb_learner_*), the ranks' gradients exchanged as ONE bucket per step (NCCL all-reduce; none on one GPU); the agents
    act on the updated weights at once.  The PyTorch-autograd learner of round 1 is timed next to it on the same rings."""
    from prisma_b200 import learner
    from prisma_b200.env import BatchedRoutingEnv
    cfg = workload_cfg(envs, 60.0, load_factor=1.1)
    cfg.agent_type = "dqn_buffer"
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, device=local_rank, arrays=arrays, env_id_base=rank * envs)
    u2o = {int(u): i for i, u in enumerate(env.overlay_to_underlay)}
    neigh = [[u2o[int(u)] for u in nb] for nb in env.neighbors]
    L = learner.DQNLearner(neigh, lr=1e-3, seed=2, device="cuda:%d" % local_rank)      # same seed: same initial weights
    env.qnet_load(L.weights())
    env.replay_enable(rows_per_node=1 << 16)
    D = learner.DeviceLearner(env, lr=1e-3)
    ar_ms = []

    def ar(t):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        dist.all_reduce(t)
        b.record()
        ar_ms.append((a, b))
    use_ar = ar if world > 1 else None
    env.step_qnet(128, eps=0.2, seed=rank)
    D.train_on_env(batch_size=512, seed=rank, world_size=world, all_reduce=use_ar, want_td=False)   # warm-up
    torch.cuda.synchronize()
    ar_ms.clear()
    collect_s = train_s = 0.0
    t_all = time.perf_counter()
    for it in range(iters):
        collect_s += _timed(torch, lambda: env.step_qnet(64, eps=0.2, seed=100 * rank + it))
        t0 = time.perf_counter()
        D.train_on_env(batch_size=512, seed=1000 * rank + it, world_size=world, all_reduce=use_ar, want_td=False)
        torch.cuda.synchronize()
        train_s += time.perf_counter() - t0
    wall = time.perf_counter() - t_all
    env.check_errors()
    w = torch.from_numpy(D.weights()).cuda()
    same = True
    if world > 1:
        ref = w.clone()
        dist.broadcast(ref, 0)
        flag = torch.tensor([1.0 if torch.equal(w, ref) else 0.0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        same = bool(flag.item() == 1.0)
    ar_all = [a.elapsed_time(b) for a, b in ar_ms] or [0.0]
    ar_mean, ar_min = sum(ar_all) / len(ar_all), min(ar_all)     # the mean holds the wait for the slowest rank to arrive
    # the autograd learner on the same rings (one rank's step, no exchange): what the device step replaces
    L.train_on_env(env, batch_size=512, seed=1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for it in range(3):
        L.train_on_env(env, batch_size=512, seed=2 + it)
    torch.cuda.synchronize()
    autograd_ms = 1000.0 * (time.perf_counter() - t0) / 3
    (_,), (wall, train_s, collect_s, ar_mean, ar_min) = _reduce(dist, torch, world, [0.0], [wall, train_s, collect_s, ar_mean, ar_min])
    env.close()
    return {"what": "collect 64 Gym steps per environment (device-side dqn_buffer agents) -> sample 512 rows per node from the "
                    "HBM rings -> TD targets, forward, backward, Adam on the device (k_learn_grad, k_learn_adam) -> the agents act "
                    "on the new weights",
            "envs_per_gpu": envs, "iterations": iters, "training_steps_per_s": iters / wall,
            "train_step_ms": 1000.0 * train_s / iters, "collect_ms": 1000.0 * collect_s / iters,
            "pytorch_autograd_train_step_ms": autograd_ms,
            "allreduce_bytes": int((D.m * D.p + D.m) * 4) if world > 1 else 0, "allreduce_ms": ar_mean if world > 1 else 0.0,
            "allreduce_ms_min": ar_min if world > 1 else 0.0,
            "allreduce_share_of_train_step": (ar_mean / (1000.0 * train_s / iters)) if world > 1 else 0.0,
            "allreduce_share_of_iteration": (ar_mean / (1000.0 * wall / iters)) if world > 1 else 0.0,
            "backend": "nccl" if world > 1 else "none (one GPU)", "weights_identical_across_ranks": same}


def leg_qnet_forward(torch, local_rank, envs):
    """The Q-network forward of all pending decisions as its own call (prb_qnet_act): one warp per environment in fp32
    (k_qnet_act) against the node-bucketed tensor-core kernel (k_qnet_bucket + k_qnet_mma, mma.sync TF32 split operands)."""
    from prisma_b200 import qnet
    from prisma_b200.env import BatchedRoutingEnv
    cfg = workload_cfg(envs, 60.0)
    cfg.agent_type = "dqn_buffer"
    arrays = cfg.load()
    env = BatchedRoutingEnv(cfg, device=local_rank, arrays=arrays)
    deg = [len(n) for n in env.neighbors]
    env.qnet_load(qnet.pack_weights(qnet.init_weights(len(deg), deg, max(deg), seed=1)))
    env.step_qnet(64, eps=0.05, seed=1)
    rec = env.last.numpy()
    pending = int(((rec["valid"] != 0) & (rec["need_action"] != 0)).sum())
    out = {"envs": envs, "pending_decisions": pending, "flop_per_decision": 17700}
    for mode in ("warp", "mma"):
        env.qnet_set_mode(mode)
        env.qnet_act()

        def go():
            for _ in range(50):
                env.qnet_act()
        out["%s_us_per_call" % mode] = 1e6 * _timed(torch, go) / 50
    env.close()
    return out


def secondary_legs(args, torch, dist, rank, local_rank, world, peak):
    """BASELINE.json configs[2], [3], [4] (this rank's share of each, 1-GPU sized at --gpus 1) and the learner."""
    out = {}
    # configs[3]: GEANT-scale 23-node underlay, load sweep, 64k envs over 8 GPUs = 8192 per GPU
    out["c3_geant_load_sweep"] = [leg_table_run(torch, dist, rank, local_rank, world, "geant", args.c3_envs, lf, 30, 4, peak)
                                  for lf in (0.5, 0.9, 1.5)]
    # configs[2]: overlay + map_overlay with dynamic underlay traffic, dqn agents on the device, pingAsObs=1, 16k envs
    out["c2_overlay_dqn_pingobs"] = leg_device_agents(torch, dist, rank, local_rank, world, "overlay_full_mesh_3n_abilene",
                                                      args.c2_envs, 0.9, 256, peak, extra={"pingAsObs": 1})
    # configs[4]: 100-node random underlay, 4-packet buffers (drop-heavy), control packets (train=1, NN signalling)
    c5 = {"max_out_buffer_size": 2168, "train": 1, "signaling_type": "NN", "agent_type": "dqn_buffer"}
    out["c5_synthetic100_drop_heavy"] = leg_table_run(torch, dist, rank, local_rank, world, "synthetic100", args.c5_envs, 1.3,
                                                      2, 2, peak, extra=c5)
    out["c5_synthetic100_drop_heavy"]["note"] = ("1M envs over 8 GPUs = 131072 per GPU; measured at %d per GPU (state: see "
                                                 "DESIGN.md); table agent, control packets on" % args.c5_envs)
    if rank == 0:
        out["qnet_forward"] = leg_qnet_forward(torch, local_rank, 65536)
    out["learner"] = leg_learner(torch, dist, rank, local_rank, world, 1024, args.learner_iters)
    return out


# ------------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------------

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="prisma_b200", choices=["prisma_b200", "reference"])
    ap.add_argument("--envs", type=int, default=4096, help="environments per GPU")
    ap.add_argument("--chunk", type=float, default=1.0, help="simulated seconds per step")
    ap.add_argument("--episode", type=float, default=60.0,
                    help="simulated seconds per episode (the reference's simTime default); the timed steps are its last K")
    ap.add_argument("--lanes", type=int, default=0, help="serial engine: lanes per environment (0 = library default)")
    ap.add_argument("--engine", default="auto", choices=["auto", "serial", "parallel"])
    ap.add_argument("--team", type=int, default=0, help="parallel engine: lanes per environment (0 = from the node count)")
    ap.add_argument("--ref-sim-time", type=float, default=60.0, dest="ref_sim_time")
    ap.add_argument("--cpu-sim-time", type=float, default=60.0, dest="cpu_sim_time",
                    help="simulated seconds per reference process of the cpu_baseline leg (60 = the reference's default episode; "
                         "its statistics are re-summed at every step, so longer episodes run slower per simulated second)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gym-steps", type=int, default=512, dest="gym_steps",
                    help="iterations of the supplementary Gym-mode leg with device-side dqn_buffer agents (0 = skip)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the legs on BASELINE.json configs[2..4] and the learner")
    ap.add_argument("--c3-envs", type=int, default=8192, dest="c3_envs")
    ap.add_argument("--c2-envs", type=int, default=16384, dest="c2_envs")
    ap.add_argument("--c5-envs", type=int, default=4096, dest="c5_envs")
    ap.add_argument("--learner-iters", type=int, default=32, dest="learner_iters")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.warmup < 3 and args.impl != "reference":
        args.warmup = 3

    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (prisma_b200 has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from prisma_b200 import build as pb
    if rank == 0:
        pb.build()
    if world > 1:
        dist.barrier()
    from prisma_b200.env import BatchedRoutingEnv, sp_table

    K, W = args.steps, args.warmup
    # One episode of the reference's own length (simTime = 60 s, argument_parser.py:35-37) in steps of --chunk simulated
    # seconds; the K timed steps are the LAST K of the episode (queues saturated, drops, lock-step links), everything
    # before them is untimed warm-up (>= W steps).
    ep_steps = max(int(round(args.episode / args.chunk)), W + K)
    lead = ep_steps - K
    cfg = workload_cfg(args.envs, ep_steps * args.chunk)
    arrays = cfg.load()
    table = sp_table(arrays)
    kdeg = max(len(x) for x in arrays.overlay_neighbors())
    b_hop = B_HOP_BASE + 4 * (1 + kdeg)
    env = BatchedRoutingEnv(cfg, device=local_rank, arrays=arrays, env_id_base=rank * args.envs, lanes_per_env=args.lanes)
    env.set_engine(args.engine, args.team)
    env.set_policy("table", table)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    lib, h = env.lib, env.h

    def hops_now():
        c = env.counters()
        env.check_errors(c)
        return int(c["packet_hops"].sum()), int(c["env_steps"].sum()), int(c["events"].sum())

    t_sim = 0.0
    for _ in range(lead):
        t_sim += args.chunk
        env.run(until_s=t_sim)
    torch.cuda.synchronize()
    hops0, steps0, events0 = hops_now()
    window = [t_sim, t_sim + K * args.chunk]

    # ---- timed region: K steps, device-resident.  The hand-over check (which makes the serial engine replay what the
    # parallel engine could not finish) sits INSIDE the event pair: no work whose hops are counted is left out. ----
    sampler = ClockSampler(local_rank)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    launches0 = env.launch_count
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    for i in range(K):
        flush.zero_()
        t_sim += args.chunk
        ev[i][0].record()
        env.run(until_s=t_sim)
        env.handovers()
        ev[i][1].record()
    torch.cuda.synchronize()
    launches = env.launch_count - launches0
    clocks = sampler.stop()
    if world > 1:
        dist.barrier()
    step_ms = [a.elapsed_time(b) for a, b in ev]
    elapsed = sum(step_ms) / 1000.0
    hops1, steps1, events1 = hops_now()
    d_hops, d_steps, d_events = hops1 - hops0, steps1 - steps0, events1 - events0
    run_engine = env.engine
    round_stats = env.round_stats() if run_engine == "parallel" else (1, 1)
    handovers, replay_ms = env.handover_stats() if run_engine == "parallel" else (0, 0.0)

    # ---- e2e: a second episode through the public host-buffer calls; the same K steps of it are timed by wall clock,
    # host->device and device->host copies inside the timed region ----
    tab_host = np.ascontiguousarray(table, dtype=np.int32)
    cnt = np.empty(args.envs, dtype=env.counters().dtype)
    import ctypes as C
    env.reset()
    t_sim = 0.0
    for _ in range(lead):
        t_sim += args.chunk
        env.set_policy("table", tab_host)
        env.run(until_s=t_sim)
        lib.prb_counters_host(h, cnt.ctypes.data_as(C.c_void_p))
    hops_lead = int(cnt["packet_hops"].sum())
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e2e_wall = 0.0
    for i in range(K):
        flush.zero_()
        torch.cuda.synchronize()
        t_sim += args.chunk
        t0 = time.perf_counter()
        env.set_policy("table", tab_host)                      # H2D: the agent's table
        env.run(until_s=t_sim)
        lib.prb_counters_host(h, cnt.ctypes.data_as(C.c_void_p))   # D2H: per-env statistics (synchronises)
        e2e_wall += time.perf_counter() - t0
    hops2 = int(cnt["packet_hops"].sum())
    e2e_hops = hops2 - hops_lead
    assert hops2 == hops1, "the two episodes are the same seeded traffic: %d vs %d packet hops" % (hops2, hops1)
    h2d = int(tab_host.nbytes)
    d2h = int(cnt.nbytes)
    e2e_handovers, e2e_replay_ms = env.handover_stats() if run_engine == "parallel" else (0, 0.0)

    # ---- Gym mode with the dqn_buffer agents on the device (supplementary; single GPU runs only) ----
    gym = None
    if world == 1 and args.gym_steps > 0:
        from prisma_b200 import qnet
        degrees = [len(x) for x in arrays.overlay_neighbors()]
        env.qnet_load(qnet.pack_weights(qnet.init_weights(len(degrees), degrees, kdeg, seed=1)))
        env.reset()
        env.step_qnet(64, eps=0.05, seed=1)
        torch.cuda.synchronize()
        s0 = int(env.counters()["env_steps"].sum())
        l0 = env.launch_count
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        env.step_qnet(args.gym_steps, eps=0.05, seed=1)
        g1.record()
        torch.cuda.synchronize()
        c = env.counters()
        env.check_errors(c)
        gym = {"value": (int(c["env_steps"].sum()) - s0) / (g0.elapsed_time(g1) / 1000.0), "unit": "agent-steps/s",
               "launches": env.launch_count - l0, "calls": args.gym_steps,
               "what": "prb_step_qnet: every environment takes `calls` Gym steps; its warp evaluates the dqn_buffer Q-network "
                       "(forward + eps-greedy argmin) for each step inside k_advance; records and actions never leave the GPU"}

    # ---- the other BASELINE.json configurations and the learner's gradient exchange (all ranks take part) ----
    secondary = None
    if not args.no_secondary:
        env.close()
        env = None
        secondary = secondary_legs(args, torch, dist, rank, local_rank, world, measured_peak_gbs()[0])

    # ---- reductions over ranks ----
    if world > 1:
        t = torch.tensor([elapsed, e2e_wall], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        elapsed, e2e_wall = float(t[0]), float(t[1])
        s = torch.tensor([d_hops, d_steps, d_events, e2e_hops, launches, handovers, e2e_handovers], dtype=torch.int64, device="cuda")
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        d_hops, d_steps, d_events, e2e_hops, launches, handovers, e2e_handovers = (int(x) for x in s)

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        # dominant kernel: k_advance, one launch per step; algorithmic bytes of THIS rank's launch
        kernel_name = "k_run_par" if run_engine == "parallel" else "k_advance"
        my_hops_per_launch = (hops1 - hops0) / K
        kern_s = (sum(step_ms) / K) / 1000.0
        achieved = my_hops_per_launch * b_hop / kern_s / 1e9
        line = {"metric": METRIC, "value": d_hops / elapsed, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": 1000.0 * elapsed / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "int64", "data": "synthetic", "config": config_dict(args, world),
                "env_steps_per_s": d_steps / elapsed, "sim_events_per_s": d_events / elapsed,
                "packet_hops_per_step": d_hops / K,
                "clocks": clocks,
                "e2e": {"value": e2e_hops / e2e_wall, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "what": "prb_set_policy(host table) + prb_run + prb_counters_host per step, wall clock"},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": recorded_traffic(args, kernel_name), "peak_source": peak_src, "kernel": kernel_name,
                             "bytes_per_packet_hop": b_hop, "packet_hops_per_launch": my_hops_per_launch,
                             "kernel_ms": kern_s * 1000.0,
                             "note": "discrete-event simulator: issue/latency bound, not HBM bound (DESIGN.md)"}}
        if run_engine == "parallel":
            rounds, executed = round_stats
            line["engine"] = {"name": "node-parallel (k_run_par)", "events_per_round": executed / max(rounds, 1),
                              "envs_replayed_by_the_serial_engine": handovers,
                              "serial_replay_ms_in_episode": replay_ms,
                              "e2e_envs_replayed_by_the_serial_engine": e2e_handovers,
                              "e2e_serial_replay_ms_in_episode": e2e_replay_ms}
        else:
            line["engine"] = {"name": "event-serial (k_advance)"}
        if gym:
            line["gym_dqn_buffer"] = gym
        if secondary:
            line["configs"] = secondary
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            hc = sc_ = dt = 0.0
            episodes = 0
            while episodes < 3 and dt < 12.0:          # a bounded sample: 10-30 s of host work
                h1, s1, d1, kind = ref_sample(args.cpu_sim_time, cores)
                hc, sc_, dt, episodes = hc + h1, sc_ + s1, dt + d1, episodes + 1
            line["cpu_baseline"] = {"value": hc / dt, "unit": UNIT, "cores": cores, "kind": kind,
                                    "env_steps_per_s": sc_ / dt,
                                    "sample": "%d processes (one per core) x %d episode(s) x 1 env x %g simulated s of the same "
                                              "workload, sp table agent in-process, %.1f s wall" % (cores, episodes,
                                                                                                   args.cpu_sim_time, dt)}
        print(json.dumps(line), flush=True)
    if env is not None:
        env.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
