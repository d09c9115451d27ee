"""Worst deviations of the device learner's gradients / TD errors from autograd over repeated runs (tolerance evidence)."""
import sys, ctypes as C
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from tests.test_gpu_learner_device import setup
from prisma_b200 import _abi
for sig in ("ideal", "target"):
    worst_g = worst_td = 0.0
    for rep in range(6):
        env, L, D = setup(sig)
        B = 256
        b = D._buffers(B)
        p = lambda t: C.c_void_p(t.data_ptr())
        for it in range(3):
            _abi.check(env.lib.prb_replay_sample_all(env.h, B, 10 + it, B, p(b["obs"]), p(b["act"]), p(b["rew"]), p(b["nxt"]), p(b["done"]), None, p(b["active"])))
            bucket, td = D.grad(B, b["obs"], b["act"], b["rew"], b["nxt"], b["done"], b["active"])
            got = bucket[:L.m * D.p].view(L.m, D.p).clone()
            if L.W.grad is not None: L.W.grad.zero_()
            total = None; ref_td = []
            for i in range(L.m):
                l, tdv = L.loss(i, b["obs"][i], b["act"][i], b["rew"][i], b["nxt"][i], b["done"][i])
                total = l if total is None else total + l
                ref_td.append(float((tdv * tdv).mean()))
            total.backward()
            ref = L.W.grad.detach()
            scale = float(ref.abs().max())
            worst_g = max(worst_g, float((got - ref).abs().max()) / max(scale, 1.0))
            worst_td = max(worst_td, float(np.max(np.abs(np.array(td) - np.array(ref_td)) / np.maximum(np.abs(ref_td), 1e-12))))
            D.apply(bucket, 1)
            with torch.no_grad(): L.W.copy_(torch.from_numpy(D.weights()).view(L.m, -1).cuda())
        env.close()
    print(sig, "worst grad dev / max(scale,1):", worst_g, " worst td rel:", worst_td, flush=True)
