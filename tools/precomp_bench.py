"""Development helper: resident MSM over a table of precomputed multiples against the plain
resident MSM.  usage: python tools/precomp_bench.py [gid] [log2n] [c ...]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid = int(sys.argv[1]) if len(sys.argv) > 1 else 0
lg = int(sys.argv[2]) if len(sys.argv) > 2 else 24
cs = [int(x) for x in sys.argv[3:]] or [0]
m.init(0)
n = 1 << lg
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
for it in range(3):
    m.timer_start()
    want = m.multi_exp_resident(rb, rs)
    ms = m.timer_stop()
print(f"plain: {ms:.2f} ms = {n / ms / 1e3:.1f} M points/s", m.last_timings(), flush=True)
want = m.batch_to_special(gid, want.reshape(1, -1))[0]
for c in cs:
    t0 = time.time()
    t = m.PrecomputedBases(gid, bases=rb, c=c)
    tb = time.time() - t0
    best, tm = 1e9, None
    for it in range(4):
        m.timer_start()
        got = m.multi_exp_precomputed(t, rs)
        ms = m.timer_stop()
        if ms < best:
            best, tm = ms, m.last_timings()
    ok = bool(np.array_equal(m.batch_to_special(gid, got.reshape(1, -1))[0], want))
    print(f"precomputed c={t.c} digits={m.precompute_digits(gid, t.c)} build={tb:.2f}s: {best:.2f} ms = "
          f"{n / best / 1e3:.1f} M points/s same={ok} "
          + " ".join(f"{k}={v:.2f}" for k, v in tm.items()), flush=True)
    t.release()
