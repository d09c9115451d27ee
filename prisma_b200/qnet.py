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


def init_weights(m: int, degrees: List[int], k_max: int, seed: int = 0) -> List[Dict[str, np.ndarray]]:
    """Fresh networks as the reference builds them (he_uniform kernels and biases; Keras takes fan_in =
    units for a 1-D bias).  Node i has degrees[i]
    actions; its arrays are padded to k_max columns / rows with zeros."""
    rng = np.random.default_rng(seed)
    nets = []
    for i in range(m):
        k = int(degrees[i])
        sh = shapes(m, k_max)
        net = {name: np.zeros(s, dtype=np.float32) for name, s in sh.items()}
        net["w_dst"][:] = he_uniform(rng, m, sh["w_dst"])
        net["w_buf"][:k] = he_uniform(rng, k, (k, H1))
        net["w1"][:] = he_uniform(rng, 2 * H1, sh["w1"])
        net["w2"][:] = he_uniform(rng, H2, sh["w2"])
        net["w3"][:, :k] = he_uniform(rng, H2, (H2, k))
        for b, fan in (("b_dst", H1), ("b_buf", H1), ("b1", H2), ("b2", H2)):
            net[b][:] = he_uniform(rng, fan, sh[b])
        net["b3"][:k] = he_uniform(rng, k, (k,))
        nets.append(net)
    return nets


def pack_weights(nets: List[Dict[str, np.ndarray]]) -> np.ndarray:
    """Flat float32 array in the order prb_qnet_load documents."""
    return np.concatenate([np.ascontiguousarray(net[name], dtype=np.float32).ravel() for net in nets for name in LAYERS])
