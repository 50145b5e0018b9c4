import sys, numpy as np
sys.path.insert(0, "/root/repo")
from gravidy_b200.hermite4 import Hermite4B200
import torch
for nb in (262144, 2097152):
    rng = np.random.default_rng(7)
    hb = Hermite4B200(np.full(nb, 1.0 / nb), rng.standard_normal((nb, 3)), rng.standard_normal((nb, 3)), softening=1e-4, eta=0.01, device=0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3): hb.predict_timed(0.0)
    ms = []
    for k in range(20):
        if nb == 262144:
            flush.zero_(); torch.cuda.synchronize()
        ms.append(hb.predict_timed(0.125 * (k + 1)))
    ms = float(np.mean(ms)); print(nb, "us", round(ms * 1e3, 2), "GB/s", round(168.0 * nb / ms / 1e6, 1))
    hb.close()
