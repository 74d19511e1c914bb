"""Development helper: resident MSM timings under different scalar distributions (uniform,
Groth16-witness-like 0/1-heavy, short scalars, one repeated scalar).
usage: python tools/dist_bench.py [gid] [log2n]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid = int(sys.argv[1]) if len(sys.argv) > 1 else 0
lg = int(sys.argv[2]) if len(sys.argv) > 2 else 22
n = 1 << lg
m.init(0)
bases, _ = syn.random_bases(gid, n, seed=1)
full = syn.random_scalars(gid, n, seed=2)
sw = full.shape[1]
rng = np.random.RandomState(3)
one = syn.encode_scalars(gid, [1])[0]
kinds = rng.choice(4, size=n, p=[0.55, 0.3, 0.1, 0.05])
witness = full.copy()
witness[kinds == 0] = 0
witness[kinds == 1] = one
small = syn.encode_scalars(gid, [int(x) for x in rng.randint(2, 1 << 16, size=4096)])
witness[kinds == 2] = small[rng.randint(0, 4096, size=int((kinds == 2).sum()))]
short64 = np.zeros_like(full)
short64[:] = small[rng.randint(0, 4096, size=n)]
same = np.repeat(full[:1], n, axis=0)
rb = m.ResidentBases(gid, bases)
for name, sc in [("uniform", full), ("witness 55%/30%/10%/5%", witness), ("16-bit", short64),
                 ("one repeated scalar", same)]:
    rs = m.ResidentScalars(gid, np.ascontiguousarray(sc))
    best = None
    for it in range(3):
        m.multi_exp_resident(rb, rs)
        t = m.last_timings()
        if best is None or t["total"] < best["total"]:
            best = t
    print(f"{name:28s} n=2^{lg} Mpts/s={n / best['total'] / 1e3:8.1f} " +
          " ".join(f"{k}={v:.2f}" for k, v in best.items()), flush=True)
    rs.release()
