"""ONE process driving 1, 2, 4, ... D GPUs through the C ABI (b200msm_init_devices): the reference's
own shape of the call -- multi_exp(bases, scalars, chunks) with one chunk per device
(multiexp.tcc:655-687) -- on a FIXED total (strong scaling, BASELINE.json "2^24 at 1/2/4/8 B200").

usage: inprocess_multi_gpu.py [group] [log2n_total] [steps]
Per device count one JSON line: resident inputs (value), host buffers (e2e), resident bases + host
scalars; every result is checked against the closed form (sum s_i (a + i d)) G.  Times are wall
clock around the call (it returns after the partials are summed on device 0 and copied back) next
to the engine's own per-device CUDA-event totals.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import libff_b200 as m  # noqa: E402


def main():
    group = sys.argv[1] if len(sys.argv) > 1 else "alt_bn128_g1"
    log2n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    gid = bench.GROUP_IDS[group]
    n = 1 << log2n
    cores = len(os.sched_getaffinity(0))
    import torch
    have = torch.cuda.device_count()
    t0 = time.time()
    bases, scalars, src = bench.workload(gid, n, 0, cores)
    want = bench.expected_point(gid, [bench.closed_form_scalar(gid, scalars, 0)])
    gen_s = time.time() - t0

    def affine(pt):
        return m.batch_to_special(gid, np.asarray(pt, dtype=np.uint64).reshape(1, -1))[0]

    counts = [d for d in (1, 2, 4, 8) if d <= have]
    for D in counts:
        m.shutdown()
        assert m.init_devices(list(range(D))) == D
        rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
        res = {}
        for name, call in (("resident", lambda: m.multi_exp_resident(rb, rs)),
                           ("host_buffers", lambda: m.multi_exp(gid, bases, scalars)),
                           ("resident_bases", lambda: m.multi_exp_resident_bases(rb, scalars))):
            for _ in range(2):
                out = call()
            ok = bool(np.array_equal(affine(out), want))
            ts = []
            for _ in range(steps):
                t = time.perf_counter()
                call()
                ts.append((time.perf_counter() - t) * 1e3)
            dev = [round(m.device_timings(k).get("total", 0.0), 3) for k in range(D)]
            ms = sorted(ts)[len(ts) // 2]
            res[name] = {"ms_median": round(ms, 3), "ms_min": round(min(ts), 3),
                         "points_per_s": n / ms * 1e3, "device_total_ms": dev, "verified": ok}
        rb.release()
        rs.release()
        print(json.dumps({"tool": "inprocess_multi_gpu", "group": group, "points_total": n,
                          "devices": D, "process": "one process, one host thread per device",
                          "steps": steps, "inputs": src, "input_gen_s": round(gen_s, 2),
                          "host_threads": cores, **res}), flush=True)
    m.shutdown()


if __name__ == "__main__":
    main()
