"""Every entry point family once at small sizes, each result checked against the reference: a
one-minute end-to-end check of a build, small enough to run under compute-sanitizer
(`compute-sanitizer --tool memcheck python tools/sanitize_small.py`) where that tool is available
(it is closed on this pool's boxes)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import libff_b200 as m  # noqa: E402
from oracle import pymodel as pm  # noqa: E402
from oracle import ref  # noqa: E402

gids = [int(x) for x in sys.argv[1:]] or [0, 1, 3]
m.init(0)
for gid in gids:
    n = 1500 if gid != 3 else 300
    g = pm.GROUPS[gid]
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=3)
    want = ref.multi_exp(gid, bases, scalars)
    aff = lambda p: m.batch_to_special(gid, np.asarray(p).reshape(1, -1))[0]  # noqa: E731
    ok = [np.array_equal(aff(m.multi_exp(gid, bases, scalars)), want)]
    for opt, val in (("reduce_cluster", 1), ("reduce_cluster", 8), ("h2d_parts", 3), ("window_bits", 7)):
        m.set_option(opt, val)
        ok.append(np.array_equal(aff(m.multi_exp(gid, bases, scalars)), want))
        m.set_option(opt, 0)
    rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
    ok.append(np.array_equal(aff(m.multi_exp_resident(rb, rs)), want))
    ok.append(np.array_equal(aff(m.multi_exp_resident_bases(rb, scalars)), want))
    zero, one = ref.group_constants(gid)
    bits, window = g.fr.bits, 4
    table = m.WindowTable(gid, bits, window, g=one)
    pts = m.batch_exp(gid, bits, window, table, scalars[:200])
    _, wantb = ref.batch_exp(gid, bits, window, ref.window_table(gid, bits, window, one), scalars[:200],
                             normalize=True)
    ok.append(np.array_equal(m.batch_to_special(gid, pts), wantb))
    table.release()
    import torch
    raw = np.ascontiguousarray(pts).view(np.uint8).reshape(-1)
    d = torch.from_numpy(raw.copy()).cuda()
    m.batch_to_special_noalloc_ptr(gid, d.data_ptr(), d.data_ptr(), raw.nbytes, pts.shape[0], 0)
    torch.cuda.synchronize()
    ok.append(np.array_equal(d.cpu().numpy().view(np.uint64).reshape(pts.shape), wantb))
    rb.release()
    rs.release()
    print(f"gid {gid}: {ok}", flush=True)
    assert all(ok)
m.shutdown()
print("sanitize_small ok")
