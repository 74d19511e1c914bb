#!/usr/bin/env python3
"""bench.py -- MSM points/sec on 1..8 B200 (BASELINE.json metric), one JSON line.

A "step" is one multi-scalar multiplication (libff::multi_exp, multiexp.tcc:643-688) over the
rank's shard of the synthetic workload SURVEY.md §8d / BASELINE.md §3 define -- the SAME inputs for
the GPU arm and the CPU reference arm: scalars SHA512_rng<Fr>(i) (rng.tcc:26-71), bases the
arithmetic walk P_i = (a + i d) G, produced by the reference's own generator (oracle/_ref) and
cached under baseline/cache/; rank k of N holds indices [k n, (k + 1) n).

  value        inputs resident in HBM (b200msm_multi_exp_resident), whole job = all ranks' points
               per step; for N > 1 each step ends with the gather of the N partial points (NCCL
               all_gather of sizeof(G) bytes) and their sum on rank 0 (multiexp.tcc:681-685)
  e2e          the same through b200msm_multi_exp with HOST (pinned) buffers: H2D of bases and
               scalars and D2H of the result inside the timed region; e2e.resident_bases = the
               prover's shape (bases stay in HBM, only the scalars travel)
  verified     the FULL result of the timed calls -- every rank's partial and the combined point,
               resident, e2e and pageable -- against the closed form (sum s_i (a + i d)) G computed
               with numpy / Python integers and one scalar multiplication of the reference
  strong       the same N GPUs on a FIXED total (2^log2n points cut into N ranges): the metric as
               BASELINE.json quotes it ("2^24 at 1/2/4/8 B200")
  roofline     dominant kernel = bucket accumulation (k_accum_first), timed with CUDA events on
               the engine's own streams; bound = the int32 multiply pipe (not HBM, not tensor
               cores: SURVEY.md §8d), peak measured live by libff_b200/microbench
  cpu_baseline libff's own CPU multi_exp (oracle/_ref, the unmodified reference) on the host
               cores, on a bounded prefix of the same workload
`--impl reference` times only the CPU arm: the FULL single-GPU workload (n points) on all host
threads.  `--op batch_exp | batch_to_special` benches the other two entry points of the path.
Default: N = 1, alt_bn128 G1, n = 2^24.
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# the groups BASELINE.json's configs name come first; the default is alt_bn128_g1
GROUP_IDS = {"alt_bn128_g1": 0, "alt_bn128_g2": 1, "bls12_377_g1": 2, "bw6_761_g1": 3,
             "bls12_377_g2": 4, "bw6_761_g2": 5, "bls12_381_g1": 6, "bls12_381_g2": 7}
# limb-MACs per point = W_ref * A * L with the reference's own window choice
SCALAR_BITS = {0: 254, 1: 254, 2: 253, 3: 377, 4: 253, 5: 377, 6: 255, 7: 255}
# Fq multiplications per mixed addition (Fq2: 7 M x 3 + 4 S x 2)
ADD_COST = {0: 11, 1: 29, 2: 11, 3: 11, 4: 29, 5: 11, 6: 11, 7: 29}
# 2 N^2 + N limb-MACs per Fq multiplication
MUL_MACS = {0: 136, 1: 136, 2: 300, 3: 1176, 4: 300, 5: 1176, 6: 300, 7: 300}
# one pass over affine bases + scalars
MIN_BYTES = {0: 96, 1: 160, 2: 128, 3: 240, 4: 224, 5: 240, 6: 128, 7: 224}
NOMINAL_INT_PEAK = 148 * 64 * 1.965e9 / 1e12  # T IMAD/s: SMs x lanes x sm_max_mhz


def ncu_capture(group, log2n):
    """dram traffic and multiply-pipe utilisation of k_accum_first from the newest committed
    `ncu --set full` capture of this workload under profiles/ (read at run time: nothing pasted
    into the source).  Returns None when the workload has no capture."""
    import csv
    pats = [f"profiles/ncu_accum_{group}_2p{log2n}_*_raw.csv"]
    legacy = {("alt_bn128_g1", 24): "profiles/accum_first_r01e_raw.csv",
              ("bls12_377_g1", 22): "profiles/accum_first_bls12_377_g1_r01f_raw.csv",
              ("alt_bn128_g2", 20): "profiles/accum_first_alt_bn128_g2_r01f_raw.csv",
              ("bw6_761_g1", 19): "profiles/accum_first_bw6_761_g1_r01f_raw.csv"}
    files = sorted(f for p in pats for f in glob.glob(os.path.join(ROOT, p)))
    if not files and (group, log2n) in legacy:
        files = [os.path.join(ROOT, legacy[(group, log2n)])]
    files = [f for f in files if os.path.exists(f)]
    if not files:
        return None
    path = files[-1]
    try:
        rows = list(csv.reader(open(path)))
        hdr = rows[0]
        col = {k: hdr.index(k) for k in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum",
                                         "dram__bytes_write.sum",
                                         "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed")}
        units = rows[1]
        # the capture holds the kernel's launches of ONE step (one per window group): totals over
        # them, the pipe utilisation weighted by launch time
        picked = [r for r in rows[2:] if "k_accum" in r[col["Kernel Name"]]]
        if not picked:
            return None
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}

        def nbytes(r, k):
            return float(r[col[k]]) * scale.get(units[col[k]], 1.0)

        ms = [float(r[col["gpu__time_duration.sum"]]) for r in picked]
        fm = [float(r[col["sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"]]) for r in picked]
        return {"file": os.path.relpath(path, ROOT), "launches": len(picked),
                "kernel_ms_under_ncu": sum(ms),
                "traffic": sum(nbytes(r, "dram__bytes_read.sum") + nbytes(r, "dram__bytes_write.sum")
                               for r in picked),
                "fmaheavy": sum(a * b for a, b in zip(ms, fm)) / sum(ms) / 100}
    except Exception as e:  # a capture in another layout: say so instead of guessing
        return {"file": os.path.relpath(path, ROOT), "error": str(e)}


def executed_macs_per_item(gid):
    """Limb products this engine issues per accumulated item (one XYZZ mixed addition, 8M + 2S):
    six plain products (2N^2 + N), one two-product sum with a shared reduction (3N^2 + 2N) and two
    dedicated squarings (N(N+1)/2 + N^2 + N); over Fp2 a product is three and a squaring two
    base-field products (Karatsuba / complex squaring)."""
    N = {136: 8, 300: 12, 1176: 24}[MUL_MACS[gid]]
    mul, sqr, mul2 = 2 * N * N + N, N * (N + 1) // 2 + N * N + N, 3 * N * N + 2 * N
    if ADD_COST[gid] == 11:
        return 6 * mul + mul2 + 2 * sqr
    return (8 * 3 + 2 * 2) * mul


def ceil_log2(n):
    r = 1 if n & (n - 1) else 0
    while n > 1:
        n >>= 1
        r += 1
    return r


def limb_macs_per_point(gid, n):
    L = ceil_log2(n)
    c_ref = L - (L // 3 - 2)  # pippenger_optimal_c, multiexp.tcc:35-40
    W = -(-SCALAR_BITS[gid] // c_ref)
    return W * ADD_COST[gid] * MUL_MACS[gid], c_ref, W


class ClockSampler:
    """nvidia-smi clocks and throttle reasons sampled during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.thread = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.thread.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        mhz = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for k, nm in enumerate(names) if any(r[2 + k] == "Active" for r in self.rows)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None,
                "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def measure_int_peak():
    """int32 IMAD issue peak of this GPU (T IMAD/s), measured by libff_b200/microbench"""
    exe = os.path.join(ROOT, "libff_b200", "microbench")
    try:
        out = subprocess.run([exe, "400"], capture_output=True, text=True, timeout=120).stdout
        d = json.loads(out)
        return max(v for k, v in d.items() if k.startswith("imad32_")), d
    except Exception as e:  # fall back to the nominal 148 SM x 64 IMAD/clk x 1.965 GHz
        return NOMINAL_INT_PEAK, {"error": str(e)}


# ---- the workload: identical for both arms -------------------------------------------------------

def checker():
    """the compiled, unmodified reference (oracle/_ref) or, without it, the C++ restatement"""
    from oracle import ref
    if ref.available():
        ref.lib()
        return ref, "reference"
    from oracle import port
    port.lib()
    return port, "port"


def set_threads(chk, kind, t):
    (chk.lib().ref_set_num_threads if kind == "reference" else chk.lib().orc_set_num_threads)(int(t))


def workload(gid, n, start, threads):
    """(bases, scalars) of indices [start, start + n): generated by the reference's own code, cached
    under baseline/cache/ so that the two arms of one round read the same bytes"""
    chk, kind = checker()
    cdir = os.path.join(ROOT, "baseline", "cache")
    tag = f"g{gid}_n{n}_s{start}"
    fb, fs = os.path.join(cdir, f"bases_{tag}.npy"), os.path.join(cdir, f"scalars_{tag}.npy")
    try:
        if os.path.exists(fb) and os.path.exists(fs):
            bases, scalars = np.load(fb), np.load(fs)
            if bases.shape[0] == n and scalars.shape[0] == n:
                return bases, scalars, "cache"
    except Exception:
        pass
    set_threads(chk, kind, threads)
    bases = chk.gen_bases_walk(gid, n, start=start)
    scalars = chk.gen_scalars(gid, n, start=start)
    try:  # only the range both arms share (the CPU arm runs indices [0, n))
        if start == 0 and n * (bases.shape[1] + scalars.shape[1]) * 8 <= (3 << 30):
            os.makedirs(cdir, exist_ok=True)
            np.save(fb + ".tmp.npy", bases)
            os.replace(fb + ".tmp.npy", fb)
            np.save(fs + ".tmp.npy", scalars)
            os.replace(fs + ".tmp.npy", fs)
    except Exception:
        pass
    return bases, scalars, kind


def expected_point(gid, ks):
    """(sum of the closed-form scalars) G, affine, by ONE scalar multiplication of the checker"""
    from oracle import pymodel as pm
    chk, _ = checker()
    g = pm.GROUPS[gid]
    k = sum(ks) % g.fr.p
    return chk.scalar_mul(gid, chk.group_constants(gid)[1], g.encode_scalars([k])[0])


def closed_form_scalar(gid, scalars, start):
    from oracle import pymodel as pm
    return pm.walk_expected_scalar_fast(pm.GROUPS[gid], scalars, start=start)


def cpu_reference_arm(gid, bases, scalars, cores, steps, warmup, budget_s):
    """libff's own CPU multi_exp (unmodified reference, oracle/_ref): BDLO12, base_form_special,
    chunks = host threads on the given inputs.  The number of timed calls is
    bounded by `budget_s` (at least min(steps, 3)); the count actually run is reported."""
    chk, kind = checker()
    set_threads(chk, kind, cores)
    m = bases.shape[0]
    calls = []
    t_begin = time.time()
    min_timed = min(steps, 3)
    for it in range(warmup + steps):
        t, out = chk.multi_exp_timed(gid, bases, scalars, form=chk.FORM_SPECIAL, chunks=cores)
        calls.append(t)
        spent = time.time() - t_begin
        if it == 0 and spent > budget_s / 6:  # minutes per call: one timed call is all there is time for
            min_timed = 1
        if len(calls) - warmup >= min_timed and spent + spent / len(calls) > budget_s:
            break
    times = calls[warmup:]
    done_w = warmup
    sec = sum(times) / len(times)
    return {"value": m / sec, "unit": "points/s", "cores": cores, "kind": kind,
            "sample": f"{m} points of the workload (SHA512_rng scalars, walk bases), "
                      f"multi_exp<BDLO12, special> chunks={cores}, mean of {len(times)} calls "
                      f"after {done_w} warm-up",
            "ms": sec * 1e3, "steps_run": len(times), "warmup_run": done_w, "result": out}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--op", default="multi_exp", choices=["multi_exp", "batch_exp", "batch_to_special"])
    ap.add_argument("--group", default="alt_bn128_g1", choices=sorted(GROUP_IDS))
    ap.add_argument("--log2n", type=int, default=None, help="points per GPU (weak scaling; default 24)")
    ap.add_argument("--log2n-total", type=int, default=None,
                    help="fixed total cut over the GPUs (strong scaling as the main value)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=240.0,
                    help="seconds of CPU calls the reference arm may spend")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    gid = GROUP_IDS[args.group]
    cores = len(os.sched_getaffinity(0))
    if args.op != "multi_exp":
        return bench_batch_op(args, gid, rank, world, local, cores)
    strong_main = args.log2n_total is not None
    if strong_main:
        if (1 << args.log2n_total) % world:
            raise SystemExit("--log2n-total must be divisible by the number of GPUs")
        n = (1 << args.log2n_total) // world
        label = f"2^{args.log2n_total} points in total, cut into {world} ranges"
    else:
        log2n = args.log2n if args.log2n is not None else 24
        n = 1 << log2n
        label = f"2^{log2n} points per GPU"
    n_total = n * world
    macs, c_ref, w_ref = limb_macs_per_point(gid, n)
    config = {"workload": f"{args.group} multi_exp, {label} (SHA512_rng scalars, arithmetic-walk bases)",
              "group": args.group, "points_per_gpu": n, "points_total": n_total,
              "parallelism": f"point-range shards x{world}",
              "l2": "inputs (n x 64 B bases + n x 32 B scalars) exceed the 126 MB L2; no flush"
                    if n * MIN_BYTES[gid] > (126 << 20) else
                    "inputs fit the L2: a 256 MB buffer is overwritten between timed steps"}

    if args.impl == "reference":
        if rank != 0:
            return
        # the CPU cannot take N x n points in minutes: at N > 1 it runs ONE rank's range and the
        # line says so (points_total = what was run)
        bases, scalars, src = workload(gid, n, 0, cores)
        cb = cpu_reference_arm(gid, bases, scalars, cores, args.steps, min(args.warmup, 1),
                               args.cpu_budget)
        got = cb.pop("result")
        ok = bool(np.array_equal(got, expected_point(gid, [closed_form_scalar(gid, scalars, 0)])))
        config["points_total"] = n
        config["parallelism"] = f"{cores} host threads (OpenMP chunks)"
        if world > 1:
            config["workload"] += f" -- the CPU arm runs ONE GPU's range ({n} points), not all {world}"
        line = {"impl": "reference", "metric": "MSM points/sec", "value": cb["value"],
                "unit": "points/s", "n_gpus": args.gpus, "steps": cb["steps_run"],
                "warmup": cb["warmup_run"], "steps_requested": args.steps,
                "warmup_requested": args.warmup, "ms_per_step": cb["ms"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u64 limbs (GMP / x86-64 asm)",
                "data": "synthetic", "config": config, "cpu_baseline": cb,
                "verified": {"n": n, "ok": ok, "against": "closed form + scalar_mul"},
                "inputs": src,
                "e2e": {"value": cb["value"], "unit": "points/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return

    import torch
    import torch.distributed as dist

    import libff_b200 as m
    from libff_b200 import distributed as D
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the MSM engine has no CPU path")
    torch.cuda.set_device(local)
    # one process per GPU: run next to the GPU (first-touch placement of the pinned source and
    # staging buffers, copy threads) -- libff_b200/numa.py; B200MSM_NO_NUMA=1 for the A/B
    numa_info = {"bound": False, "why": "single process"}
    if world > 1 and not os.environ.get("B200MSM_NO_NUMA"):
        from libff_b200 import numa
        numa_info = numa.bind_to_gpu_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    m.init(local)
    gb, sb = m.group_bytes(gid), m.scalar_bytes(gid)

    # ---- inputs: this rank's range of the common workload
    t0 = time.time()
    start = rank * n
    bases, scalars, src = workload(gid, n, start, max(1, cores // world))
    gen_s = time.time() - t0
    k_mine = closed_form_scalar(gid, scalars, start)

    def affine(pt):
        return m.batch_to_special(gid, np.asarray(pt, dtype=np.uint64).reshape(1, -1))[0]

    rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)
    flush = None
    if n * MIN_BYTES[gid] <= (126 << 20):
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def all_ks(k):
        """every rank's closed-form scalar (Python integers through an object gather)"""
        if world == 1:
            return [k]
        out = [None] * world
        dist.all_gather_object(out, int(k))
        return out

    last = {}

    def combine(partial, tag):
        """gather the per-GPU partial points and add them (multiexp.tcc:681-685)"""
        last[tag + "_partial"] = partial
        if world == 1:
            last[tag] = partial
            return partial
        allp = D.gather_partials(partial)  # NCCL all_gather of sizeof(G) bytes per rank
        out = D.combine_partials(gid, allp) if rank == 0 else partial
        last[tag] = out
        return out

    def step_resident():
        return combine(m.multi_exp_resident(rb, rs), "resident")

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        phases = []
        barrier()
        l0 = m.launch_count()
        # CUDA events on the engine's own stream (every kernel and copy of the library is ordered
        # against it; torch.cuda.Event would only see torch's stream)
        dt_sum = 0.0
        if flush is None:
            m.timer_start()
        for _ in range(steps):
            if flush is not None:  # small inputs: evict them from the L2 between timed steps
                flush.fill_(1)
                torch.cuda.synchronize()
                m.timer_start()
            fn()
            if flush is not None:
                dt_sum += m.timer_stop() * 1e-3
            phases.append(m.last_timings())
        barrier()
        dt = dt_sum if flush is not None else m.timer_stop() * 1e-3
        launches = m.launch_count() - l0
        if world > 1:
            x = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(x, op=dist.ReduceOp.MAX)
            dt = float(x.item())
        return dt, phases, launches

    verified = {"n": n_total, "against": "closed form (sum s_i (a + i d)) G: numpy / Python integers + "
                                         "one scalar_mul of the compiled reference", "paths": {}}

    def verify(tag, ks):
        """the full result of the LAST timed call on path `tag`: this rank's partial against its
        own closed form, and on rank 0 the combined point against the sum over the ranks"""
        mine = bool(np.array_equal(affine(last[tag + "_partial"]), expected_point(gid, [k_mine])))
        ok_all = mine
        if world > 1:
            flags = [None] * world
            dist.all_gather_object(flags, mine)
            ok_all = all(flags)
            if rank == 0:
                ok_all = ok_all and bool(np.array_equal(affine(last[tag]), expected_point(gid, ks)))
        verified["paths"][tag] = ok_all
        if not ok_all:
            raise SystemExit(f"bench.py: the {tag} result of the timed call does not match the closed form")

    ks = all_ks(k_mine)

    # ---- value: inputs resident in HBM
    with ClockSampler(local) as clk:
        dt, phases, launches = timed(step_resident, args.steps, args.warmup)
    clocks = clk.summary()
    verify("resident", ks)
    ms_step = dt / args.steps * 1e3
    value = n_total / (dt / args.steps)
    accum_ms = float(np.mean([p["accumulate"] for p in phases]))
    phase_ms = {k: float(np.mean([p[k] for p in phases])) for k in phases[0]}

    # ---- strong scaling: the same GPUs on a fixed total of n points (BASELINE.json: "2^24 at
    # 1/2/4/8 B200"): every rank takes n / N of its resident points
    strong = None
    if not args.no_strong and not strong_main and n % world == 0:
        ns = n // world
        k_s = closed_form_scalar(gid, scalars[:ns], start)

        def step_strong():
            return combine(m.multi_exp_resident(rb, rs, 0, ns), "strong")

        dts, ph_s, _ = timed(step_strong, args.steps, args.warmup)
        ok = bool(np.array_equal(affine(last["strong_partial"]), expected_point(gid, [k_s])))
        ks_s = all_ks(k_s)
        if rank == 0 and world > 1:
            ok = ok and bool(np.array_equal(affine(last["strong"]), expected_point(gid, ks_s)))
        strong = {"n_gpus": world, "points_total": n, "points_per_gpu": ns,
                  "ms_per_step": dts / args.steps * 1e3, "value": n / (dts / args.steps),
                  "unit": "points/s", "verified": ok,
                  "phases_ms": {k: float(np.mean([p[k] for p in ph_s])) for k in ph_s[0]},
                  "note": "fixed total = one GPU's weak-scaling load; efficiency = value(N) / (N x value(1))"}
        if not ok:
            raise SystemExit("bench.py: the strong-scaling result does not match the closed form")

    # ---- e2e: host (pinned) buffers through b200msm_multi_exp, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        hb = torch.from_numpy(bases).pin_memory()
        hs = torch.from_numpy(scalars).pin_memory()

        def step_e2e():
            return combine(m.multi_exp_host_ptr(gid, hb.data_ptr(), hs.data_ptr(), n), "e2e")

        dte, ph_e, _ = timed(step_e2e, args.steps, args.warmup)
        verify("e2e", ks)
        e2e = {"value": n_total / (dte / args.steps), "unit": "points/s",
               "h2d_bytes_per_step": n * (gb + sb) * world, "d2h_bytes_per_step": gb * world,
               "ms_per_step": dte / args.steps * 1e3,
               "ingest_ms": float(np.mean([p["ingest"] for p in ph_e])),
               "call": "b200msm_multi_exp(host bases, host scalars) -> host result",
               "host_placement": numa_info}

        # the prover's shape: the bases (a proving key) stay resident, the scalars travel
        def step_rb():
            return combine(m.multi_exp_resident_bases_ptr(rb, hs.data_ptr(), n), "resident_bases")

        dtr, _, _ = timed(step_rb, args.steps, args.warmup)
        verify("resident_bases", ks)
        e2e["resident_bases"] = {"value": n_total / (dtr / args.steps), "unit": "points/s",
                                 "ms_per_step": dtr / args.steps * 1e3,
                                 "h2d_bytes_per_step": n * sb * world, "d2h_bytes_per_step": gb * world,
                                 "call": "b200msm_multi_exp_resident_bases(resident bases, host scalars)"}

        # the same call on pageable memory (what a libff std::vector is): staged through the
        # engine's pinned ring buffers, special-form bases leave their Z behind
        def step_pageable():
            return combine(m.multi_exp(gid, bases, scalars), "pageable")

        ps = max(2, args.steps // 2)
        dtp, _, _ = timed(step_pageable, ps, 1)
        verify("pageable", ks)
        e2e["pageable_value"] = n_total / (dtp / ps)
        e2e["pageable_h2d_bytes_per_step"] = n * (m.wire_point_bytes(gid) + sb) * world
        del hb, hs

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (bucket accumulation), integer-multiply pipe
    peak, micro = measure_int_peak()
    c_eng = m.window_bits(gid, n)
    engine_windows = -(-(SCALAR_BITS[gid] + 1) // c_eng)
    achieved = n * macs * 2 / (accum_ms * 1e-3) / 1e12  # T IMAD slots/s, SURVEY.md §8d convention
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm_peak = json.load(open(peaks_file))["hbm_gbs"] if os.path.exists(peaks_file) else 6650.0
    cap = ncu_capture(args.group, ceil_log2(n))
    roofline = {
        "bound": "int32-multiply pipe (IMAD)", "kernel": "k_accum_first",
        "achieved": achieved, "peak": peak, "unit": "T IMAD/s", "frac": achieved / peak,
        "peak_source": "measured live: libff_b200/microbench imad32 (MEASURED_PEAKS.json has no "
                       "integer peak)" if "error" not in micro else "nominal 148 x 64 x 1.965 GHz",
        "peak_nominal": NOMINAL_INT_PEAK, "frac_of_nominal": achieved / NOMINAL_INT_PEAK,
        "microbench_raw": micro,
        "algorithmic_unit": f"{macs} limb-MACs/point (W_ref={w_ref} x A={ADD_COST[gid]} x "
                            f"L={MUL_MACS[gid]}, c_ref={c_ref}) x 2 IMAD slots",
        "kernel_ms": accum_ms, "kernel_share_of_step": accum_ms / phase_ms["total"],
        "kernel_ms_note": "phase `accumulate`: all launches of k_accum_first of one step (one per "
                          "window group), with the next group's sort running beside them",
        # the limb products the kernel really issues (signed digits, this run's digit size, XYZZ
        # mixed additions with dedicated squarings) against the same peak: always <= 1
        "executed": {"frac": n * engine_windows * executed_macs_per_item(gid) * 2
                     / (accum_ms * 1e-3) / 1e12 / peak,
                     "unit": f"n x {engine_windows} digits x {executed_macs_per_item(gid)} "
                             "limb products x 2 IMAD slots / kernel time / peak"},
        # dram__bytes_read.sum + dram__bytes_write.sum and sm__pipe_fmaheavy_cycles_active of the
        # kernel's longest launch in the newest committed `ncu --set full` capture of this workload
        "traffic": cap.get("traffic") if cap else None,
        "pipe_utilisation_ncu": cap.get("fmaheavy") if cap else None,
        "ncu_capture": cap,
        "traffic_unit": "bytes per step = all launches of the kernel in one MSM (ncu dram read + "
                        "write); algorithmic gather = items x (affine point + 4-byte item)",
        "hbm_stream": {"achieved_GBps": n * MIN_BYTES[gid] / (phase_ms["total"] * 1e-3) / 1e9,
                       "peak_GBps": hbm_peak, "note": "secondary counter; the path is not HBM-bound"},
    }

    cpu_baseline = None
    if not args.no_cpu_baseline:
        try:
            # a bounded prefix of this rank's workload: ~10-30 s of CPU work
            rate = 70e3 * cores * (11 * 136) / (ADD_COST[gid] * MUL_MACS[gid])
            lg = max(10, min(ceil_log2(n), int(np.log2(max(2.0, rate * 5.0)))))
            ms_ = 1 << lg
            cpu_baseline = cpu_reference_arm(gid, bases[:ms_], scalars[:ms_], cores, 2, 1, 30.0)
            got = cpu_baseline.pop("result")
            cpu_baseline["verified"] = bool(np.array_equal(
                got, expected_point(gid, [closed_form_scalar(gid, scalars[:ms_], start)])))
        except Exception as e:
            cpu_baseline = {"value": None, "unit": "points/s", "cores": cores, "kind": "reference",
                            "sample": f"unavailable: {e}"}

    line = {"metric": "MSM points/sec", "value": value, "unit": "points/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong" if strong_main else "weak",
            "vs_baseline": None,
            "dtype": "u32 limbs (254-bit Montgomery)" if gid < 2 else "u32 limbs",
            "data": "synthetic", "config": config, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
            "phases_ms": phase_ms, "window_bits": c_eng, "verified": verified, "strong": strong,
            "inputs": src, "input_gen_s": gen_s}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bench_batch_op(args, gid, rank, world, local, cores):
    """The other two entry points of the path (north_star: "multi_exp / batch_exp / batch_to_special"):
    batch_exp = fixed-base exponentiation of n scalars from a window table (multiexp.tcc:875-915,
    the reference's own window choice get_exp_window_size :770-807), batch_to_special = batched
    normalisation (multiexp.tcc:949-974, field_utils.tcc:419-439).  value = elements/s through the
    C ABI with host buffers (both calls are host-in / host-out by their signature), kernel_ms =
    the device part; the reference's CPU time on the same inputs beside it."""
    if rank != 0:
        return
    import torch

    import libff_b200 as m
    chk, kind = checker()
    set_threads(chk, kind, cores)
    m.init(local)
    log2n = args.log2n if args.log2n is not None else 20
    n = 1 << log2n
    bits = SCALAR_BITS[gid]
    N = {136: 8, 300: 12, 1176: 24}[MUL_MACS[gid]]
    peak, micro = measure_int_peak()
    scalars = chk.gen_scalars(gid, n, start=0)
    _, one = chk.group_constants(gid)
    window = int(chk.lib().ref_exp_window_size(gid, n)) if kind == "reference" else 17
    rows = -(-bits // window)
    t0 = time.time()
    table = m.WindowTable(gid, bits, window, g=one)
    build_s = time.time() - t0

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        ph, wall = [], []
        for _ in range(args.steps):
            t = time.time()
            fn()
            wall.append(time.time() - t)
            ph.append(m.last_timings())
        return float(np.mean(wall)), {k: float(np.mean([p[k] for p in ph])) for k in ph[0]}

    if args.op == "batch_exp":
        out = {"pts": np.zeros((n, m.group_bytes(gid) // 8), dtype=np.uint64)}  # caller-owned, reused

        def fn():
            m.batch_exp(gid, bits, window, table, scalars, out=out["pts"])

        wall, ph = timed(fn)
        # CPU: the reference's batch_exp on a bounded prefix, same table shape
        ms_ = min(n, 1 << 16)
        tab_host = table.download()
        cpu_s, want = chk.batch_exp(gid, bits, window, tab_host, scalars[:ms_], normalize=True)
        ok = bool(np.array_equal(m.batch_to_special(gid, out["pts"][:ms_]), want))
        # one mixed addition per table row (8M + 2S) + the conversion back to Jacobian (2M)
        macs = (rows * 10 + 2) * (2 * N * N + N) * (3 if ADD_COST[gid] == 29 else 1)
        unit_desc = f"({rows} rows x 10 products + 2) x (2N^2+N) limb-MACs x 2 IMAD slots per scalar"
        items, cpu_items = n, ms_
        what = f"batch_exp, 2^{log2n} scalars, window {window} (get_exp_window_size), {rows} rows"
    else:
        pts = m.batch_exp(gid, bits, window, table, scalars)  # Jacobian outputs with Z != 1
        work = pts.copy()

        def fn():  # in place, as the reference's call is; re-normalising costs the same work
            m.batch_to_special(gid, work, inplace=True)

        wall, ph = timed(fn)
        ms_ = min(n, 1 << 20)
        t = time.time()
        want = chk.batch_to_special(gid, pts[:ms_])
        cpu_s = time.time() - t
        ok = bool(np.array_equal(m.batch_to_special(gid, pts[:ms_]), want))
        # Montgomery trick: 3 products per element for the inverses + 1S + 3M to apply them
        macs = 7 * (2 * N * N + N) * (3 if ADD_COST[gid] == 29 else 1)
        unit_desc = "7 products x (2N^2+N) limb-MACs x 2 IMAD slots per point (+ one inversion per batch)"
        items, cpu_items = n, ms_
        what = f"batch_to_special, 2^{log2n} Jacobian points"
    kernel_ms = ph["accumulate"]
    achieved = items * macs * 2 / (kernel_ms * 1e-3) / 1e12
    line = {"metric": f"{args.op} elements/sec", "op": args.op, "value": items / wall,
            "unit": "elements/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": wall * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32 limbs", "data": "synthetic",
            "config": {"workload": f"{args.group} {what}", "group": args.group},
            "e2e": {"value": items / wall, "unit": "elements/s",
                    "h2d_bytes_per_step": int(scalars.nbytes if args.op == "batch_exp" else pts.nbytes),
                    "d2h_bytes_per_step": int(items * m.group_bytes(gid)),
                    "note": "the call is host-in / host-out by its signature: value IS the e2e figure"},
            "kernel_ms": kernel_ms, "kernel_value": items / (kernel_ms * 1e-3),
            "phases_ms": ph, "table_build_s": build_s if args.op == "batch_exp" else None,
            "roofline": {"bound": "int32-multiply pipe (IMAD)", "achieved": achieved, "peak": peak,
                         "unit": "T IMAD/s", "frac": achieved / peak, "algorithmic_unit": unit_desc,
                         "traffic": None},
            "cpu_baseline": {"value": cpu_items / cpu_s, "unit": "elements/s", "cores": cores,
                             "kind": kind, "sample": f"{cpu_items} elements, reference {args.op}"},
            "verified": {"n": cpu_items, "ok": ok, "against": f"reference {args.op}, limb for limb after normalisation"},
            "gpu_launches": int(m.launch_count())}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
