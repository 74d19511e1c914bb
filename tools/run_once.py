"""Development helper: one resident MSM (for ncu captures).  usage: run_once.py gid log2n [option value]..."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid, lg = int(sys.argv[1]), int(sys.argv[2])
m.init(0)
for k in range(3, len(sys.argv), 2):
    m.set_option(sys.argv[k], int(sys.argv[k + 1]))
n = 1 << lg
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
for _ in range(2):
    m.multi_exp_resident(rb, rs)
print(m.last_timings())
