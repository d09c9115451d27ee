"""OptAgent: the restatement of optimal_routing_decision (prisma/source/utils.py:161-199)."""
import numpy as np

from prisma_b200 import _abi
from prisma_b200.agents import OptAgent


def ring4():
    # overlay ring 0-1-2-3-0; edges in list(G.edges) order: (0,1) (0,3) (1,0) (1,2) (2,1) (2,3) (3,0) (3,2)
    nb = [[1, 3], [0, 2], [1, 3], [0, 2]]
    r = np.zeros((4, 4, 8))
    r[0][2][[0, 1]] = [0.25, 0.75]       # flow 0 -> 2 splits at node 0: 25 % via 1, 75 % via 3
    r[0][2][3] = 0.25                    # ... and continues 1 -> 2 with the same share
    r[0][2][7] = 0.75                    # 3 -> 2
    return nb, r


def test_choice_follows_the_shares_and_the_tag_keeps_split_flows_on_their_path():
    nb, r = ring4()
    ag = OptAgent(nb, r, seed=1)
    picks = [ag.decide(0, 0, 2, None) for _ in range(4000)]
    share = np.mean([a == 1 for a, _ in picks])
    assert 0.72 < share < 0.78 and all(t == (0.25, 0.75)[a] for a, t in picks)
    assert ag.decide(1, 0, 2, 0.25) == (1, 0.25)       # node 1: out-edges (1,0) (1,2); the tag 0.25 is on (1,2)
    assert ag.decide(3, 0, 2, 0.75) == (1, 0.75)
    # a tag of 1 (unsplit flow) or an unknown tag falls back to the draw
    r[1][3][[3, 5]] = [1.0, 1.0]
    ag = OptAgent(nb, r, seed=2)
    assert ag.decide(1, 1, 3, 1.0) == (1, 1.0) and ag.decide(2, 1, 3, 0.3) == (1, 1.0)


def test_batch_interface_tracks_tags_per_environment_and_packet():
    nb, r = ring4()
    ag = OptAgent(nb, r, seed=3)
    rec = np.zeros(3, dtype=_abi.STEP_DTYPE)
    rec["valid"], rec["ptype"], rec["need_action"] = 1, 0, 1
    rec["overlay"], rec["src"], rec["uid"] = 0, 0, [10, 11, 12]
    rec["obs"][:, 0] = 2
    rec["valid"][2] = 0
    a0 = ag.act(rec, np.arange(4))
    assert a0[2] == 0 and set(ag.tags) == {(0, 10), (1, 11)}
    nxt = rec.copy()
    nxt["overlay"] = [1 if a0[0] == 0 else 3, 1 if a0[1] == 0 else 3, 0]
    a1 = ag.act(nxt, np.arange(4))
    assert a1[0] == 1 and a1[1] == 1                   # both branches reach node 2 through their second out-edge
    done = nxt.copy()
    done["need_action"] = 0
    ag.act(done, np.arange(4))
    assert ag.tags == {}
