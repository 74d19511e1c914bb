"""Development helper: e2e time of b200msm_multi_exp from pinned and pageable host buffers for the
h2d_pack / stage_lanes settings.  usage: python tools/pack_sweep.py [gid] [log2n]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid = int(sys.argv[1]) if len(sys.argv) > 1 else 0
lg = int(sys.argv[2]) if len(sys.argv) > 2 else 24
m.init(0)
n = 1 << lg
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
hb = torch.from_numpy(bases).pin_memory()
hs = torch.from_numpy(scalars).pin_memory()
first = None
for src in ("pinned", "pageable"):
    for pack in (0, 2) if src == "pinned" else (0, 1):
        for lanes in (4, 8, 12, 16):
            m.set_option("h2d_pack", pack)
            m.set_option("stage_lanes", lanes)
            best = 1e9
            for it in range(3):
                m.timer_start()
                if src == "pinned":
                    got = m.multi_exp_host_ptr(gid, hb.data_ptr(), hs.data_ptr(), n)
                else:
                    got = m.multi_exp(gid, bases, scalars)
                best = min(best, m.timer_stop())
            aff = m.batch_to_special(gid, got.reshape(1, -1))[0]
            first = aff if first is None else first
            print(f"{src} h2d_pack={pack} stage_lanes={lanes}: {best:.2f} ms = {n / best / 1e3:.1f} M points/s "
                  f"same={bool(np.array_equal(aff, first))}", flush=True)
            if src == "pinned" and pack == 0:
                break
