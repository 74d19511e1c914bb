"""Development helper: e2e (pinned host buffers -> result) time of b200msm_multi_exp for several
h2d_parts settings, next to the resident time.  usage: python tools/e2e_sweep.py [gid] [log2n] [parts ...]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid = int(sys.argv[1]) if len(sys.argv) > 1 else 0
lg = int(sys.argv[2]) if len(sys.argv) > 2 else 24
parts = [int(x) for x in sys.argv[3:]] or [0, 4, 5, 6]
m.init(0)
n = 1 << lg
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
want = None
for it in range(3):
    m.timer_start()
    want = m.multi_exp_resident(rb, rs)
    ms = m.timer_stop()
print(f"resident: {ms:.2f} ms", m.last_timings(), flush=True)
rb.release()
rs.release()
want = m.batch_to_special(gid, want.reshape(1, -1))[0]
hb = torch.from_numpy(bases).pin_memory()
hs = torch.from_numpy(scalars).pin_memory()
growths = [int(x) for x in os.environ.get("GROWTH", "0").split(",")]
shareds = [int(x) for x in os.environ.get("SHARED", "1").split(",")]
tapers = [int(x) for x in os.environ.get("TAPER", "0").split(",")]
windows = [int(x) for x in os.environ.get("WINDOW", "0").split(",")]
for sh, p, gr, tp, wb in [(sh, p, gr, tp, wb) for wb in windows for sh in shareds for p in parts for gr in growths
                          for tp in tapers]:
    m.set_option("window_bits", wb)
    m.set_option("h2d_shared_buckets", sh)
    m.set_option("h2d_taper", tp)
    m.set_option("h2d_parts", p)
    m.set_option("h2d_growth", gr)
    best = 1e9
    for it in range(4):
        m.timer_start()
        got = m.multi_exp_host_ptr(gid, hb.data_ptr(), hs.data_ptr(), n)
        best = min(best, m.timer_stop())
    ok = bool(np.array_equal(m.batch_to_special(gid, got.reshape(1, -1))[0], want))
    print(f"window_bits={wb} shared={sh} h2d_parts={p} growth={gr} taper={tp}: e2e {best:.2f} ms = {n / best / 1e3:.1f} M points/s  same={ok}", flush=True)
