"""Generates tests/golden/*.npz from the reference's own ns-3 program (oracle/_ref/prisma_ref).

Run in the build container (needs /root/reference-built oracle/_ref):  python tests/golden/make_golden.py
Each file holds one short run: the injection trace, the complete event log, every Gym step, final
counters (float32 statistics as raw bits), the routing table and the policy that drove it.
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle, run_ref   # noqa: E402
from tests import common            # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

GOLDEN = {
    "abilene_sp_4s": ("abilene", dict(simTime=4, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1), "table"),
    "abilene_random_drops_3s": ("abilene", dict(simTime=3, load_factor=1.5, max_out_buffer_size=2168, pingAsObs=0,
                                                activateUnderlayTraffic=1), "random"),
    "overlay_ping_underlay_6s": ("overlay_full_mesh_3n_abilene", dict(simTime=6, load_factor=2.0, pingAsObs=1,
                                                                     activateUnderlayTraffic=1), "random"),
    "abilene_train_nn_2s": ("abilene", dict(simTime=2, load_factor=1.2, pingAsObs=0, train=1, signaling_type="NN",
                                            agent_type="dqn_buffer", activateUnderlayTraffic=1), "random"),
    # NN-weight segments between overlay neighbours on top of the small signalling (sim.cc:634-647)
    "abilene_train_nn_bigsig_2s": ("abilene", dict(simTime=2, load_factor=1.0, pingAsObs=0, train=1, signaling_type="NN",
                                                   signalingSim=1, sync_step=0.5, bigSignalingSize=8192,
                                                   agent_type="dqn_buffer", activateUnderlayTraffic=1), "random"),
    "geant_sp_2s": ("geant", dict(simTime=2, load_factor=1.0, pingAsObs=0, activateUnderlayTraffic=1), "table"),
}
COUNTERS = ("overlay_injected", "overlay_arrived", "overlay_lost", "underlay_injected", "underlay_arrived",
            "underlay_lost", "env_steps", "decisions", "n_e2e", "n_cost")
FLOATS = ("avg_e2e_delay", "avg_cost", "avg_underlay_e2e_delay", "avg_underlay_cost", "signaling_overhead")


def big_info(info):
    """(NN Index, segment Index, NodeId Signaled) of a big-signalling step (big-signaling-packet-manager.cc:104-121)."""
    f = dict(t.strip().split("=") for t in info.split(",") if "=" in t)
    return int(f["NN Index"]), int(f["segment Index"]), int(f["NodeId Signaled"])


def main(only=()):
    for name, (topo, over, policy) in GOLDEN.items():
        if only and name not in only:
            continue
        cfg = common.env_config(topo, **over)
        arrays = cfg.load()
        table = common.sp_table(arrays)
        with tempfile.TemporaryDirectory() as tmp:
            files = common.ref_files(cfg, tmp)
            out = os.path.join(tmp, "out")
            run_ref.run(files, out, sim_time=cfg.simTime, load_factor=cfg.load_factor, link_cap=cfg.link_cap,
                        link_delay_ms=cfg.link_delay, packet_size=cfg.packet_size,
                        max_out_buffer_size=cfg.max_out_buffer_size, agent_type=cfg.agent_type,
                        signaling_type=cfg.signaling_type, train=bool(cfg.train),
                        loss_penalty=cfg.derived_loss_penalty(arrays.n_overlay),
                        moving_average=cfg.movingAverageObsSize,
                        activate_underlay_traffic=bool(cfg.activateUnderlayTraffic), ping_as_obs=bool(cfg.pingAsObs),
                        ping_interval=cfg.pingPacketIntervalTime, policy=policy, policy_data=table, policy_seed=7,
                        seed=cfg.seed, signaling=cfg.effective_signaling(), sync_step=cfg.sync_step,
                        big_signaling_size=cfg.bigSignalingSize, log_info=True)
            summ = run_ref.parse_summary(out)
            fr, to = oracle.build_links(arrays.phys_adj)
            _, route = run_ref.link_maps(summ, fr, to)
            ev = run_ref.parse_events(out, summ, fr, to)
            dec = [d for d in run_ref.parse_decisions(out, with_info=True) if d["ptype"] >= 0]
        big = np.array([big_info(d["info"]) if d["ptype"] == 1 else (-1, -1, -1) for d in dec], dtype=np.int32)
        kmax = max(len(d["obs"]) for d in dec)
        obs = np.full((len(dec), kmax), -1, dtype=np.int32)
        for i, d in enumerate(dec):
            obs[i, :len(d["obs"])] = d["obs"]
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            topology=topo, overrides=repr(over), policy=policy, policy_seed=7, table=table, route=route, events=ev,
            dec_t=np.array([d["t_ns"] for d in dec], dtype=np.int64),
            dec_node=np.array([d["node"] for d in dec], dtype=np.int32),
            dec_uid=np.array([d["uid"] for d in dec], dtype=np.int64),
            dec_ptype=np.array([d["ptype"] for d in dec], dtype=np.int32),
            dec_done=np.array([d["done"] for d in dec], dtype=np.int32),
            dec_action=np.array([d["action"] for d in dec], dtype=np.int32),
            dec_nobs=np.array([len(d["obs"]) for d in dec], dtype=np.int32), dec_obs=obs,
            **({"dec_big": big} if (big[:, 0] >= 0).any() else {}),
            counters=np.array([summ[k] for k in COUNTERS], dtype=np.int64),
            float_bits=np.array([summ[k + "_bits"] for k in FLOATS], dtype=np.uint32))
        print(name, len(ev), "events", len(dec), "steps")


if __name__ == "__main__":
    main(sys.argv[1:])
