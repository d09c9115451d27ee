"""Generates tests/fixtures/synthetic100: BASELINE.json configs[4] / SURVEY.md §8 C5 — a seeded 100-node random
underlay in the reference's file formats (prisma/examples/*/topology_files, traffic_matrices) — and
tests/fixtures/synthetic56, the same construction at the largest size the reference itself can run: sim.cc:446-447
numbers every switch device out of 10.1.1.0/24, so N + 2|E| <= 254 (at 100 nodes ns-3 aborts with
"Ipv4AddressGeneratorImpl::Add(): Address Collision"); 56 nodes with mean degree 3 stay below that.

  python tests/fixtures/make_synthetic100.py

Erdős–Rényi G(100, 3.5/99) from numpy's PCG64 seeded with 100, made connected by joining every further
component to the first one through its lowest-numbered nodes; every node is an overlay node and the overlay
adjacency equals the physical one (as in the Abilene and GEANT examples).  Traffic matrix: U(0, 2) per ordered
pair, scaled so that the busiest link under ns-3 global routing (oracle.global_routes, the restatement of
Ipv4GlobalRouting pinned by tests/test_oracle_vs_ref.py) carries exactly link_cap = 500 kbit/s at load_factor 1,
written with two decimals in Kbps like the reference's node_intensity_normalized_*.txt files."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle   # noqa: E402

SEED, LINK_CAP = 100, 500000.0
HERE = os.path.dirname(os.path.abspath(__file__))


def components(adj):
    n = len(adj)
    comp = -np.ones(n, dtype=int)
    c = 0
    for s in range(n):
        if comp[s] >= 0:
            continue
        stack = [s]
        comp[s] = c
        while stack:
            x = stack.pop()
            for y in np.nonzero(adj[x])[0]:
                if comp[y] < 0:
                    comp[y] = c
                    stack.append(y)
        c += 1
    return comp, c


def main(N, MEAN_DEGREE, OUT):
    rng = np.random.Generator(np.random.PCG64(SEED))
    upper = np.triu(rng.random((N, N)) < MEAN_DEGREE / (N - 1), 1)
    adj = (upper | upper.T).astype(np.uint8)
    comp, nc = components(adj)
    for c in range(1, nc):
        a, b = int(np.nonzero(comp == 0)[0][c % int((comp == 0).sum())]), int(np.nonzero(comp == c)[0][0])
        adj[a, b] = adj[b, a] = 1
    assert components(adj)[1] == 1
    tm = rng.random((N, N)) * 2.0
    np.fill_diagonal(tm, 0.0)
    # link loads under ns-3 global routing
    fr, to = oracle.build_links(adj)
    route = oracle.global_routes(adj)
    load = np.zeros(len(fr))
    for s in range(N):
        for d in range(N):
            if s == d:
                continue
            x = s
            while x != d:
                l = route[x, d]
                load[l] += tm[s, d]
                x = to[l]
    tm *= LINK_CAP / load.max()
    os.makedirs(os.path.join(OUT, "topology_files"), exist_ok=True)
    os.makedirs(os.path.join(OUT, "traffic_matrices"), exist_ok=True)
    np.savetxt(os.path.join(OUT, "topology_files", "physical_adjacency_matrix.txt"), adj, fmt="%2d")
    np.savetxt(os.path.join(OUT, "topology_files", "overlay_adjacency_matrix.txt"), adj, fmt="%2d")
    np.savetxt(os.path.join(OUT, "topology_files", "map_overlay.txt"), np.arange(N), fmt="%d")
    with open(os.path.join(OUT, "traffic_matrices", "node_intensity_normalized_0.txt"), "w") as f:
        for s in range(N):
            f.write(" ".join("0bps" if s == d else "%.2fKbps" % (tm[s, d] / 1000.0) for d in range(N)) + "\n")
    deg = adj.sum(1)
    assert N + int(adj.sum()) <= 254 or N > 85, "the reference cannot address this many devices"
    print(os.path.basename(OUT), "links", int(adj.sum()), "degree min/mean/max", deg.min(), deg.mean(), deg.max(), "flows", N * (N - 1),
          "max link load at lf 1: %.0f bit/s" % (load.max() * LINK_CAP / load.max()))


if __name__ == "__main__":
    main(100, 3.5, os.path.join(HERE, "synthetic100"))
    main(56, 3.0, os.path.join(HERE, "synthetic56"))
