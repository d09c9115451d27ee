"""TEST INFRASTRUCTURE — runs oracle/_ref/prisma_ref (the reference's own ns-3 program built by
oracle/build_ref.py) on one configuration and parses what the harness wrote.

Command-line mapping follows prisma/source/run_ns3.py:41-73.  Used by the tests that pin the CPU
restatement, by tests/golden/make_golden.py and by `bench.py --impl reference`.
"""
import os
import shutil
import subprocess
import tempfile
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "_ref", "prisma_ref")


def available():
    return os.path.exists(BIN) and os.access(BIN, os.X_OK)


def _write_matrix(path, a, fmt="%d"):
    np.savetxt(path, np.asarray(a), fmt=fmt)


def write_inputs(workdir, arrays, traffic_strings=None):
    """Materialise the four input files the reference reads (sim.cc:274-302)."""
    os.makedirs(workdir, exist_ok=True)
    paths = {k: os.path.join(workdir, k + ".txt") for k in ("adj", "overlay", "map", "tm", "rej")}
    _write_matrix(paths["adj"], arrays["phys_adj"])
    _write_matrix(paths["overlay"], arrays["overlay_adj"])
    with open(paths["map"], "w") as f:
        for v in arrays["map_overlay"]:
            f.write("%d\n" % int(v))
    n = int(arrays["n_nodes"])
    with open(paths["tm"], "w") as f:
        for i in range(n):
            if traffic_strings is not None:
                f.write(" ".join(traffic_strings[i]) + "\n")
            else:
                f.write(" ".join("%dbps" % int(arrays["tm_bps"][i][j]) for j in range(n)) + "\n")
    rej = arrays.get("opt_rejected")
    _write_matrix(paths["rej"], rej if rej is not None else np.zeros((n, n)), fmt="%.6f")
    return paths


def run(files, out_dir, *, sim_time=60, seed=100, load_factor=1.0, link_cap=500000, link_delay_ms=1,
        packet_size=512, max_out_buffer_size=16260, agent_type="sp", signaling_type="ideal",
        signaling=False, train=False, sync_step=1.0, loss_penalty=0.0, moving_average=5,
        activate_underlay_traffic=True, ping_as_obs=False, big_signaling_size=35328,
        ping_interval=0.2, policy="zero", policy_data=None, policy_seed=1, events=True, log_info=False,
        timeout=3600):
    """Run the reference binary; returns wall-clock seconds spent inside it."""
    os.makedirs(out_dir, exist_ok=True)
    args = [BIN, "--h_out=" + out_dir, "--h_policy=" + policy, "--h_events=%d" % (1 if events else 0),
            "--h_log_info=%d" % (1 if log_info else 0), "--h_policy_seed=%d" % policy_seed]
    if policy in ("table", "trace"):
        pf = os.path.join(out_dir, "policy.txt")
        with open(pf, "w") as f:
            if policy == "table":
                for row in policy_data:
                    f.write(" ".join(str(int(x)) for x in row) + "\n")
            else:
                f.write("\n".join(str(int(x)) for x in policy_data) + "\n")
        args.append("--h_policy_file=" + pf)
    b = lambda v: "True" if v else "False"
    args += ["--simSeed=%d" % seed, "--openGymPort=6555", "--simTime=%s" % sim_time,
             "--AvgPacketSize=%d" % packet_size, "--LinkDelay=%sms" % link_delay_ms,
             "--LinkRate=%dbps" % link_cap, "--MaxBufferLength=%dB" % max_out_buffer_size,
             "--load_factor=%s" % load_factor, "--adj_mat_file_name=" + files["adj"],
             "--overlay_mat_file_name=" + files["overlay"], "--node_intensity_file_name=" + files["tm"],
             "--signaling=" + b(signaling), "--AgentType=" + agent_type, "--signalingType=" + signaling_type,
             "--syncStep=%s" % sync_step, "--lossPenalty=%s" % loss_penalty, "--activateOverlaySignaling=True",
             "--train=" + b(train), "--movingAverageObsSize=%d" % moving_average,
             "--activateUnderlayTraffic=" + b(activate_underlay_traffic),
             "--opt_rejected_file_name=" + files["rej"], "--map_overlay_file_name=" + files["map"],
             "--pingAsObs=" + b(ping_as_obs), "--bigSignalingSize=%d" % big_signaling_size,
             "--pingPacketIntervalTime=%s" % ping_interval]
    t0 = time.perf_counter()
    # the reference writes e2e_delay_per_flow_*.txt into its cwd (point-to-point-net-device.cc PrintDelayInfo)
    r = subprocess.run(args, cwd=out_dir, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE, timeout=timeout)
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        raise RuntimeError("prisma_ref failed (%d): %s" % (r.returncode, r.stderr.decode()[-2000:]))
    return dt


def kat(lines):
    """Known-answer queries answered by ns-3's own Time/DataRate code (ref_harness.cc RunKat)."""
    with tempfile.NamedTemporaryFile("w", suffix=".kat", delete=False) as f:
        f.write("\n".join(lines) + "\n")
        name = f.name
    try:
        out = subprocess.run([BIN, "--h_kat=" + name], stdout=subprocess.PIPE, check=True).stdout.decode()
    finally:
        os.unlink(name)
    return [l.split() for l in out.strip().splitlines()]


EV_KIND = {"INJ": 0, "ARR": 1, "ENQ": 2, "DRP": 3, "TXB": 4, "RCV": 5}


def parse_summary(out_dir):
    s = {"dev_peer": {}, "route_dev": {}}
    with open(os.path.join(out_dir, "summary.txt")) as f:
        for line in f:
            t = line.split()
            if not t:
                continue
            if t[0] == "dev":
                s["dev_peer"][(int(t[1]), int(t[2]))] = int(t[4])
            elif t[0] == "route":
                s["route_dev"][(int(t[1]), int(t[2]))] = int(t[4])
            elif t[0].startswith("avg_") or t[0] == "signaling_overhead":
                s[t[0]] = float(t[1])
                s[t[0] + "_bits"] = int(t[2], 16)
            else:
                s[t[0]] = int(t[1])
    return s


def link_maps(summary, link_from, link_to):
    """(node, dev) of the reference <-> our directed link ids (access link of node i = L+i)."""
    n = summary["n_switch"]
    L = len(link_from)
    by_pair = {(int(a), int(b)): l for l, (a, b) in enumerate(zip(link_from, link_to))}
    nd2link = {}
    for (node, dev), peer in summary["dev_peer"].items():
        if dev == 1:
            continue
        nd2link[(node, dev)] = by_pair[(node, peer)]
    route = np.full((n, n), -1, dtype=np.int32)
    for (x, d), dev in summary["route_dev"].items():
        if x != d and dev >= 2:
            route[x, d] = nd2link[(x, dev)]
    return nd2link, route


def parse_events(out_dir, summary, link_from, link_to):
    """events.csv -> structured array with our link ids.  Switch records: link = the device's own
    outgoing link for ENQ/DRP/TXB, the ARRIVAL link for RCV (peer->node); INJ/ARR: access link."""
    nd2link, _ = link_maps(summary, link_from, link_to)
    n = summary["n_switch"]
    L = len(link_from)
    dt = np.dtype([("t_ns", "i8"), ("kind", "i4"), ("node", "i4"), ("link", "i4"), ("uid", "i8"), ("size", "i4"),
                   ("type", "i4"), ("src", "i4"), ("dst", "i4"), ("last_hop", "i4"), ("next_hop", "i4"),
                   ("valable", "i4"), ("rejected", "i4"), ("qbytes", "i4")])
    rows = []
    with open(os.path.join(out_dir, "events.csv")) as f:
        next(f)
        for line in f:
            k, t, node, dev, uid, size, typ, src, dst, lh, nh, val, rej, qb = line.rstrip("\n").split(",")
            node, dev, typ = int(node), int(dev), int(typ)
            if typ < 0 or typ > 4:
                continue   # untagged packets (ICMP, quirk Q6) never touch a switch<->switch queue
            kind = EV_KIND[k]
            if kind == 0:
                link = L + (node - n)
            elif kind == 1:
                link = L + node
            elif kind == 5:
                link = nd2link[(node, dev)] ^ 1
            else:
                link = nd2link[(node, dev)]
            data = typ == 0
            rows.append((int(t), kind, node, link, int(uid) if data else -1, int(size), typ, int(src), int(dst),
                         int(lh), int(nh), int(val) if data else -1, int(rej) if data else 0, int(qb)))
    return np.array(rows, dtype=dt)


def parse_decisions(out_dir, with_info=False):
    rows = []
    with open(os.path.join(out_dir, "decisions.csv")) as f:
        next(f)
        for line in f:
            info = None
            if with_info and ',"' in line:
                line, info = line.split(',"', 1)
                info = info.rstrip().rstrip('"')
            t = line.rstrip("\n").split(",")
            nobs = int(t[8])
            rows.append({"t_ns": int(t[0]), "overlay": int(t[1]), "node": int(t[2]), "uid": int(t[3]),
                         "ptype": int(t[4]), "done": int(t[5]), "reward_bits": int(t[6], 16), "action": int(t[7]),
                         "obs": [int(x) for x in t[9:9 + nobs]], "info": info})
    return rows


def injection_trace(events):
    """The INJ records, in execution order = the replay input of the restatement and of the CUDA path."""
    e = events[events["kind"] == 0]
    return e
