"""Development helper: shared-bucket upload ranges against the resident MSM on tiny inputs, then
per-range phase times at 2^24.  usage: python tools/range_debug.py"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

m.init(0)
gid = 0
for n in (2, 3, 5, 16, 64, 1000):
    bases, _ = syn.random_bases(gid, n, seed=1000)
    scalars = syn.random_scalars(gid, n, seed=2000)
    for wb in (0, 2, 5, 11):
        m.set_option("window_bits", wb)
        m.set_option("h2d_parts", 1)
        want = m.batch_to_special(gid, m.multi_exp(gid, bases, scalars).reshape(1, -1))[0]
        for parts in (2, 3, 4):
            for sh in (0, 1):
                m.set_option("h2d_parts", parts)
                m.set_option("h2d_shared_buckets", sh)
                res = []
                for rep in range(3):
                    got = m.batch_to_special(gid, m.multi_exp(gid, bases, scalars).reshape(1, -1))[0]
                    res.append(bool(np.array_equal(got, want)))
                if not all(res):
                    print(f"n={n} wb={wb} parts={parts} shared={sh}: {res}", flush=True)
print("tiny cases done", flush=True)
m.set_option("window_bits", 0)
n = 1 << 24
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
hb = torch.from_numpy(bases).pin_memory()
hs = torch.from_numpy(scalars).pin_memory()
for sh, parts in ((1, 8), (0, 5)):
    m.set_option("h2d_shared_buckets", sh)
    m.set_option("h2d_parts", parts)
    m.set_option("h2d_growth", 100 if sh else 130)
    for it in range(2):
        m.multi_exp_host_ptr(gid, hb.data_ptr(), hs.data_ptr(), n)
    m.set_option("trace", 1)
    print(f"--- shared={sh} parts={parts}", file=sys.stderr, flush=True)
    m.multi_exp_host_ptr(gid, hb.data_ptr(), hs.data_ptr(), n)
    m.set_option("trace", 0)
