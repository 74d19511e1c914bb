"""Development helper: time the resident MSM at a few sizes and print phase timings.
usage: python tools/quick_bench.py [gid] [log2n ...]   (env WINDOW=c to force the window)"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from oracle import pymodel as pm  # noqa: E402
from oracle import ref  # noqa: E402

gid = int(sys.argv[1]) if len(sys.argv) > 1 else 0
sizes = [int(x) for x in sys.argv[2:]] or [20]
m.init(0)
g = pm.GROUPS[gid]
for lg in sizes:
    n = 1 << lg
    t = time.time()
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n)
    tg = time.time() - t
    rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
    for c in [int(x) for x in os.environ.get("WINDOW", "0").split(",")]:
        m.set_option("window_bits", c)
        best = None
        for it in range(3):
            out = m.multi_exp_resident(rb, rs)
            tm = m.last_timings()
            if best is None or tm["total"] < best["total"]:
                best = tm
        print(f"gid={gid} n=2^{lg} c={m.window_bits(gid, n)} gen={tg:.1f}s "
              f"Mpts/s={n / best['total'] / 1e3:.2f} " +
              " ".join(f"{k}={v:.2f}" for k, v in best.items()), flush=True)
    if lg <= 22:
        t = time.time()
        sp = g.decode_scalars(scalars) if lg <= 18 else None
        if sp is not None:
            k = pm.walk_expected_scalar(g, sp)
            want = g.encode_affine(g.mul(k, g.gen))
        else:
            ref.set_threads = None
            want = ref.multi_exp(gid, bases, scalars, chunks=os.cpu_count())
        got = m.batch_to_special(gid, out.reshape(1, -1))[0]
        print("  parity:", bool(np.array_equal(got, want)), f"(check {time.time() - t:.1f}s)", flush=True)
    rb.release()
    rs.release()
