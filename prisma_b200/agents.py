"""Agent hooks evaluated by the caller in Gym mode (`prb_step`): the ones whose decision needs more than the device-side
table / Q-network hooks offer.

`OptAgent` restates `optimal_routing_decision` (prisma/source/utils.py:161-199) and the way `Forwarder.take_action` calls
it (prisma/source/forwarder.py:192-194): the optimal solution gives, per (source, destination), the share of the flow on
every directed overlay edge; a node picks a neighbour with probability proportional to the shares of its out-edges, and a
packet that already carries the share it was routed with (`tag`) follows the out-edge with that same share when there
is one (split flows stay on their path).  The step record provides what it needs: `src`, `obs[0]` (destination) and
`uid`.  The optimal-solution files themselves are not part of the reference checkout (`agent.py:135-137` reads them
from `examples/<topology>/optimal_solution/`), so this hook is exercised with hand-made solutions only."""
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np


class OptAgent:
    def __init__(self, neighbors: Sequence[Sequence[int]], routing_mat: np.ndarray, rejected_mat: Optional[np.ndarray] = None,
                 seed: int = 0):
        """neighbors[i]: overlay neighbours of overlay node i in action order (= list(G.neighbors(i)), agent.py:165);
        routing_mat[src][dst][e]: share of flow (src, dst) on directed overlay edge e, edges in the order of
        list(G.edges) of the overlay DiGraph: for i ascending, for j in neighbors[i]."""
        self.neighbors = [list(n) for n in neighbors]
        self.edges: List[Tuple[int, int]] = [(i, j) for i, nb in enumerate(self.neighbors) for j in nb]
        self.out_edges = [np.array([k for k, (a, _) in enumerate(self.edges) if a == i], dtype=np.int64)
                          for i in range(len(self.neighbors))]
        self.routing = np.asarray(routing_mat, dtype=np.float64)
        self.rejected = None if rejected_mat is None else np.asarray(rejected_mat, dtype=np.float64)
        self.rng = np.random.default_rng(seed)
        self.tags: Dict[Tuple[int, int], float] = {}      # (environment, packet uid) -> tag (pkt_tracking_dict[...]["tag"])

    def decide(self, actual: int, src: int, dst: int, tag: Optional[float]) -> Tuple[int, float]:
        """optimal_routing_decision: (action, new tag)."""
        prob = self.routing[src][dst][self.out_edges[actual]]
        if tag:
            hit = np.nonzero(prob == tag)[0]
            if len(hit) and tag < 1:
                return int(hit[0]), tag
        p = prob / prob.sum()
        a = int(self.rng.choice(len(prob), p=p))
        return a, float(prob[a])

    def act(self, records: np.ndarray, overlay_of: np.ndarray) -> np.ndarray:
        """Actions for a batch of prb_step records (one per environment); overlay_of = map_overlay."""
        out = np.zeros(len(records), dtype=np.int32)
        for e, r in enumerate(records):
            if not r["valid"] or r["ptype"] != 0:
                continue
            key = (e, int(r["uid"]))
            if not r["need_action"]:
                self.tags.pop(key, None)                  # delivered
                continue
            a, tag = self.decide(int(r["overlay"]), int(overlay_of[int(r["src"])]), int(r["obs"][0]), self.tags.get(key))
            self.tags[key] = tag
            out[e] = a
        return out
