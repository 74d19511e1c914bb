"""Development helper: does the host -> device copy of a 2^24-point MSM's inputs (2 GiB from pinned
memory) run at the same rate while the accumulation kernels are running?  Times the copy alone and
under a resident MSM of the same size (engine stream), with CUDA events on the copy stream."""
import os
import sys
import threading

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from libff_b200 import synthetic as syn  # noqa: E402

gid, lg = 0, int(sys.argv[1]) if len(sys.argv) > 1 else 24
n = 1 << lg
m.init(0)
bases, _ = syn.random_bases(gid, n, seed=1000)
scalars = syn.random_scalars(gid, n, seed=2000)
rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
hb = torch.from_numpy(bases).pin_memory()
hs = torch.from_numpy(scalars).pin_memory()
db = torch.empty_like(hb, device="cuda")
ds = torch.empty_like(hs, device="cuda")
cs = torch.cuda.Stream()
parts = 12


def copy_all():
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(cs):
        e0.record()
        step = n // parts
        for k in range(parts):
            lo, hi = k * step, n if k == parts - 1 else (k + 1) * step
            ds[lo:hi].copy_(hs[lo:hi], non_blocking=True)
            db[lo:hi].copy_(hb[lo:hi], non_blocking=True)
        e1.record()
    return e0, e1


gib = (hb.numel() + hs.numel()) * 8 / 2 ** 30
for it in range(3):
    e0, e1 = copy_all()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"copy alone: {ms:.2f} ms = {gib / ms * 1e3 * 1.0737:.1f} GB/s", flush=True)
for it in range(3):
    t = threading.Thread(target=lambda: [m.multi_exp_resident(rb, rs) for _ in range(2)])
    t.start()
    e0, e1 = copy_all()
    t.join()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"copy under the resident MSM: {ms:.2f} ms = {gib / ms * 1e3 * 1.0737:.1f} GB/s", flush=True)
