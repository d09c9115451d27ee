"""Weights of the dqn_buffer agents for the device-side Q-network forward (prb_qnet_*).

The network is the reference's `DQN_buffer_model` (prisma/source/models.py:257-301): one-hot destination ->
Dense(32, elu); LayerNormalization (no affine) of the k output-queue occupancies -> Dense(32, elu); concat ->
Dense(64, elu) -> Dense(64, elu) -> Dense(k, elu), he_uniform kernels and biases; one network per overlay
node (prisma/source/agent.py).  The forward pass and the argmin / eps-greedy choice (learner.py:142-159) run in
prisma_b200/csrc/prisma_b200.cu:k_qnet_act; this module only lays the weights out for prb_qnet_load."""
from typing import Dict, List

import numpy as np

H1, H2 = 32, 64
LAYERS = ["w_dst", "b_dst", "w_buf", "b_buf", "w1", "b1", "w2", "b2", "w3", "b3"]


def shapes(m: int, k_max: int) -> Dict[str, tuple]:
    return {"w_dst": (m, H1), "b_dst": (H1,), "w_buf": (k_max, H1), "b_buf": (H1,), "w1": (2 * H1, H2), "b1": (H2,),
            "w2": (H2, H2), "b2": (H2,), "w3": (H2, k_max), "b3": (k_max,)}


def floats_per_node(m: int, k_max: int) -> int:
    return int(sum(int(np.prod(s)) for s in shapes(m, k_max).values()))


def he_uniform(rng, fan_in: int, shape) -> np.ndarray:
    lim = np.sqrt(6.0 / fan_in)
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


# Widths (destination branch, buffer branch, first hidden, second hidden) of the reference's models that share the
# two-branch, two-hidden-layer shape (prisma/source/models.py): DQN_buffer_model :258-306, DQN_buffer_lite_model :184-227,
# DQN_routing_model :360-392 (no buffer branch).  A narrower network is the SAME function as the full-width one with
# zeros in the unused rows and columns: elu (0) = 0, so a padded unit outputs nothing, feeds nothing and receives a zero
# gradient — the device kernels (acting and learning) serve these variants exactly, without a code path of their own.
# The one-hidden-layer variants (lighter, lighter_2, lighter_3) and the `ff` model cannot be embedded (elu is no identity).
# u2 == 0: ONE hidden layer after the concatenation (DQN_buffer_lighter_model :39-84, _2 :87-132, _3 :135-181): the output
# layer reads the first hidden layer and W_2 / b_2 of the layout stay zero (prb_qnet_set_arch (1)).
VARIANTS = {"dqn_buffer": (32, 32, 64, 64), "dqn_buffer_lite": (16, 16, 32, 32), "dqn_routing": (32, 0, 64, 64),
            "dqn_buffer_lighter": (16, 16, 32, 0), "dqn_buffer_lighter_2": (8, 8, 16, 0), "dqn_buffer_lighter_3": (4, 4, 8, 0)}


def hidden_layers(variant: str) -> int:
    return 2 if VARIANTS[variant][3] else 1


def init_weights(m: int, degrees: List[int], k_max: int, seed: int = 0, variant: str = "dqn_buffer") -> List[Dict[str, np.ndarray]]:
    """Fresh networks as the reference builds them (he_uniform kernels and biases; Keras takes fan_in =
    units for a 1-D bias).  Node i has degrees[i] actions; its arrays are padded to k_max columns / rows and — for the
    narrower variants — to the full layer widths with zeros."""
    ud, ub, u1, u2 = VARIANTS[variant]
    rng = np.random.default_rng(seed)
    nets = []
    for i in range(m):
        k = int(degrees[i])
        sh = shapes(m, k_max)
        net = {name: np.zeros(s, dtype=np.float32) for name, s in sh.items()}
        net["w_dst"][:, :ud] = he_uniform(rng, m, (m, ud))
        net["b_dst"][:ud] = he_uniform(rng, ud, (ud,))
        if ub:
            net["w_buf"][:k, :ub] = he_uniform(rng, k, (k, ub))
            net["b_buf"][:ub] = he_uniform(rng, ub, (ub,))
        w1 = he_uniform(rng, ud + ub, (ud + ub, u1))          # rows: the concatenated branch outputs
        net["w1"][:ud, :u1] = w1[:ud]
        net["w1"][H1:H1 + ub, :u1] = w1[ud:]
        net["b1"][:u1] = he_uniform(rng, u1, (u1,))
        if u2:
            net["w2"][:u1, :u2] = he_uniform(rng, u1, (u1, u2))
            net["b2"][:u2] = he_uniform(rng, u2, (u2,))
        last = u2 or u1
        net["w3"][:last, :k] = he_uniform(rng, last, (last, k))
        net["b3"][:k] = he_uniform(rng, k, (k,))
        nets.append(net)
    return nets


def narrow_view(net: Dict[str, np.ndarray], k: int, variant: str) -> Dict[str, np.ndarray]:
    """The variant's own (unpadded) arrays inside a padded network — what the reference's Keras model would hold."""
    ud, ub, u1, u2 = VARIANTS[variant]
    out = {"w_dst": net["w_dst"][:, :ud], "b_dst": net["b_dst"][:ud],
           "w1": np.concatenate([net["w1"][:ud, :u1], net["w1"][H1:H1 + ub, :u1]]), "b1": net["b1"][:u1],
           "w3": net["w3"][:(u2 or u1), :k], "b3": net["b3"][:k]}
    if u2:
        out["w2"] = net["w2"][:u1, :u2]
        out["b2"] = net["b2"][:u2]
    if ub:
        out["w_buf"] = net["w_buf"][:k, :ub]
        out["b_buf"] = net["b_buf"][:ub]
    return out


def padding_mask(m: int, degrees: List[int], k_max: int, variant: str) -> np.ndarray:
    """Flat boolean array [M * P]: True where a parameter belongs to the variant (False: padding that must stay zero)."""
    ud, ub, u1, u2 = VARIANTS[variant]
    parts = []
    for i in range(m):
        k = int(degrees[i])
        sh = shapes(m, k_max)
        net = {name: np.zeros(s, dtype=bool) for name, s in sh.items()}
        net["w_dst"][:, :ud] = True; net["b_dst"][:ud] = True
        net["w_buf"][:k, :ub] = True; net["b_buf"][:ub] = True
        net["w1"][:ud, :u1] = True; net["w1"][H1:H1 + ub, :u1] = True; net["b1"][:u1] = True
        net["w2"][:u1, :u2] = True; net["b2"][:u2] = True
        net["w3"][:(u2 or u1), :k] = True; net["b3"][:k] = True
        parts += [net[name].ravel() for name in LAYERS]
    return np.concatenate(parts)


def pack_weights(nets: List[Dict[str, np.ndarray]]) -> np.ndarray:
    """Flat float32 array in the order prb_qnet_load documents."""
    return np.concatenate([np.ascontiguousarray(net[name], dtype=np.float32).ravel() for net in nets for name in LAYERS])
