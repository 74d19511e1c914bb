"""Development helper: resident MSM time for several values of one engine option.
usage: python tools/opt_sweep.py gid log2n option v1 [v2 ...]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid, lg, opt = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
vals = [int(x) for x in sys.argv[4:]]
m.init(0)
n = 1 << lg
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
first = None
for v in vals:
    m.set_option(opt, v)
    best, tm = 1e9, None
    for it in range(4):
        m.timer_start()
        got = m.multi_exp_resident(rb, rs)
        ms = m.timer_stop()
        if ms < best:
            best, tm = ms, m.last_timings()
    aff = m.batch_to_special(gid, got.reshape(1, -1))[0]
    if first is None:
        first = aff
    print(f"gid={gid} n=2^{lg} {opt}={v}: {best:.2f} ms = {n / best / 1e3:.1f} M points/s "
          f"same={bool(np.array_equal(aff, first))} "
          + " ".join(f"{k}={x:.2f}" for k, x in tm.items()), flush=True)
