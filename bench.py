#!/usr/bin/env python3
"""bench.py -- MSM points/sec on 1..8 B200 (BASELINE.json metric), one JSON line.

A "step" is one multi-scalar multiplication (libff::multi_exp, multiexp.tcc:643-688) over the
rank's shard of synthetic random bases/scalars:
  value        inputs resident in HBM (b200msm_multi_exp_resident), whole job = all ranks' points
               per step; for N > 1 each step ends with the gather of the N partial points (NCCL
               all_gather of sizeof(G) bytes) and their sum on rank 0 (multiexp.tcc:681-685)
  e2e          the same through b200msm_multi_exp with HOST (pinned) buffers: H2D of bases and
               scalars and D2H of the result inside the timed region
  roofline     dominant kernel = bucket accumulation (k_accum_first), timed with CUDA events on
               the engine's own stream; bound = the int32 multiply pipe (not HBM, not tensor
               cores: SURVEY.md §8d), peak measured live by libff_b200/microbench
  cpu_baseline libff's own CPU multi_exp (oracle/_ref, the unmodified reference) on the host
               cores, on a bounded sample of the same workload
`--impl reference` times only that CPU arm.  Default: N = 1, alt_bn128 G1, n = 2^24.
"""
import argparse
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


# ncu --set full, k_accum_first, per launch: 28.72 GB read + 1.28 GB written
# (profiles/accum_first_r01d_raw.csv)
NCU_TRAFFIC = {("alt_bn128_g1", 1 << 24): 29.85e9, ("bls12_377_g1", 1 << 22): 11.01e9,
               ("alt_bn128_g2", 1 << 20): 2.64e9, ("bw6_761_g1", 1 << 19): 2.02e9}
# sm__pipe_fmaheavy_cycles_active (the pipe IMAD / IMAD.WIDE issue on) of k_accum_first in the
# committed `ncu --set full` captures: the hardware-side utilisation figure next to `frac`
NCU_PIPE_UTIL = {("alt_bn128_g1", 1 << 24): (0.894, "profiles/accum_first_r01e_raw.csv"),
                 ("bls12_377_g1", 1 << 22): (0.831, "profiles/accum_first_bls12_377_g1_r01f_raw.csv"),
                 ("alt_bn128_g2", 1 << 20): (0.677, "profiles/accum_first_alt_bn128_g2_r01f_raw.csv"),
                 ("bw6_761_g1", 1 << 19): (0.626, "profiles/accum_first_bw6_761_g1_r01f_raw.csv")}


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
        return 18.6, {"error": str(e)}


def cpu_reference_arm(gid, n_total, cores, steps, warmup, budget_s=20.0):
    """libff's own CPU multi_exp (unmodified reference, oracle/_ref) on a bounded sample of the
    workload: BDLO12, base_form_special, chunks = host threads."""
    from oracle import ref  # the checker library; the only place bench.py executes oracle/
    ref.lib()
    ref.lib().ref_set_num_threads(cores)
    # ~40 k points/s/core for alt_bn128 G1, scaled by field-multiplication cost
    rate = 40e3 * cores * (11 * 136) / (ADD_COST[gid] * MUL_MACS[gid])
    lg = max(10, min(ceil_log2(n_total), int(np.log2(max(2.0, rate * budget_s / max(1, steps + warmup))))))
    m = 1 << lg
    bases = ref.gen_bases_walk(gid, m)
    scalars = ref.gen_scalars(gid, m)
    times = []
    for it in range(warmup + steps):
        t, _ = ref.multi_exp_timed(gid, bases, scalars, form=ref.FORM_SPECIAL, chunks=cores)
        if it >= warmup:
            times.append(t)
    sec = sum(times) / len(times)
    return {"value": m / sec, "unit": "points/s", "cores": cores, "kind": "reference",
            "sample": f"2^{lg} points of the workload (SHA512_rng scalars, walk bases), "
                      f"multi_exp<BDLO12, special> chunks={cores}, mean of {len(times)}",
            "ms": sec * 1e3}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--group", default="alt_bn128_g1", choices=sorted(GROUP_IDS))
    ap.add_argument("--log2n", type=int, default=24, help="points per GPU (weak scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    gid = GROUP_IDS[args.group]
    n = 1 << args.log2n
    n_total = n * world
    cores = len(os.sched_getaffinity(0))
    macs, c_ref, w_ref = limb_macs_per_point(gid, n)
    workload = f"{args.group} multi_exp, 2^{args.log2n} random points/scalars per GPU"
    config = {"workload": workload, "group": args.group, "points_per_gpu": n,
              "points_total": n_total, "parallelism": f"point-range shards x{world}",
              "l2": "inputs (n x 64 B bases + n x 32 B scalars) exceed the 126 MB L2; no flush"}

    if args.impl == "reference":
        if rank != 0:
            return
        cb = cpu_reference_arm(gid, n_total, cores, args.steps, max(1, args.warmup))
        line = {"impl": "reference", "metric": "MSM points/sec", "value": cb["value"],
                "unit": "points/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": cb["ms"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (u64 on the CPU)",
                "data": "synthetic", "config": config, "cpu_baseline": cb,
                "e2e": {"value": cb["value"], "unit": "points/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
        return

    import torch
    import torch.distributed as dist

    import libff_b200 as m
    from libff_b200 import distributed as D
    from libff_b200 import synthetic as syn
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the MSM engine has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    m.init(local)
    gb, sb = m.group_bytes(gid), m.scalar_bytes(gid)

    # ---- synthetic inputs, generated with the engine itself (libff_b200/synthetic.py)
    t0 = time.time()
    bases, kk = syn.random_bases(gid, n, seed=1000 + rank)
    scalars = syn.random_scalars(gid, n, seed=2000 + rank)
    gen_s = time.time() - t0

    # ---- property check before timing (no oracle): on a 2^12 prefix the MSM must equal
    # (sum k_i s_i) G, the right side computed with Python integers and one fixed-base exp
    chk = min(n, 1 << 12)
    r_mod = syn.CURVES[gid][0]
    kp, sp = syn.scalars_plain(gid, kk[:chk]), syn.scalars_plain(gid, scalars[:chk])
    expect_k = sum(a * b for a, b in zip(kp, sp)) % r_mod
    tab = syn.generator_window_table(gid, 8)
    expect = m.batch_to_special(gid, m.batch_exp(gid, r_mod.bit_length(), 8, list(tab),
                                                 syn.encode_scalars(gid, [expect_k])))[0]
    got = m.batch_to_special(gid, m.multi_exp(gid, bases[:chk], scalars[:chk]).reshape(1, -1))[0]
    verified = bool(np.array_equal(got, expect))
    if not verified:
        raise SystemExit("bench.py: MSM result does not match the closed form; not timing")

    rb, rs = m.ResidentBases(gid, bases), m.ResidentScalars(gid, scalars)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def combine(partial):
        """gather the per-GPU partial points and add them (multiexp.tcc:681-685)"""
        if world == 1:
            return partial
        allp = D.gather_partials(partial)  # NCCL all_gather of sizeof(G) bytes per rank
        return D.combine_partials(gid, allp) if rank == 0 else partial

    def step_resident():
        return combine(m.multi_exp_resident(rb, rs))

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        phases = []
        barrier()
        l0 = m.launch_count()
        # CUDA events on the engine's own stream (every kernel and copy of the library is issued
        # there; torch.cuda.Event would only see torch's stream)
        m.timer_start()
        for _ in range(steps):
            fn()
            phases.append(m.last_timings())
        barrier()
        dt = m.timer_stop() * 1e-3
        launches = m.launch_count() - l0
        if world > 1:
            x = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(x, op=dist.ReduceOp.MAX)
            dt = float(x.item())
        return dt, phases, launches

    # ---- value: inputs resident in HBM
    with ClockSampler(local) as clk:
        dt, phases, launches = timed(step_resident, args.steps, args.warmup)
    clocks = clk.summary()
    ms_step = dt / args.steps * 1e3
    value = n_total / (dt / args.steps)
    accum_ms = float(np.mean([p["accumulate"] for p in phases]))
    phase_ms = {k: float(np.mean([p[k] for p in phases])) for k in phases[0]}

    # ---- e2e: host (pinned) buffers through b200msm_multi_exp, H2D + D2H inside the timed region
    e2e = None
    if not args.no_e2e:
        hb = torch.from_numpy(bases).pin_memory()
        hs = torch.from_numpy(scalars).pin_memory()

        def step_e2e():
            return combine(m.multi_exp_host_ptr(gid, hb.data_ptr(), hs.data_ptr(), n))

        dte, ph_e, _ = timed(step_e2e, args.steps, args.warmup)
        e2e = {"value": n_total / (dte / args.steps), "unit": "points/s",
               "h2d_bytes_per_step": n * (gb + sb) * world, "d2h_bytes_per_step": gb * world,
               "ms_per_step": dte / args.steps * 1e3,
               "ingest_ms": float(np.mean([p["ingest"] for p in ph_e])),
               "call": "b200msm_multi_exp(host bases, host scalars) -> host result"}
        # the same call on pageable memory (what a libff std::vector is): staged through the
        # engine's pinned ring buffers
        def step_pageable():
            return combine(m.multi_exp(gid, bases, scalars))

        dtp, _, _ = timed(step_pageable, max(2, args.steps // 2), 1)
        e2e["pageable_value"] = n_total / (dtp / max(2, args.steps // 2))
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
    roofline = {
        "bound": "int32-multiply pipe (IMAD)", "kernel": "k_accum_first",
        "achieved": achieved, "peak": peak, "unit": "T IMAD/s", "frac": achieved / peak,
        "peak_source": "measured live: libff_b200/microbench imad32 (MEASURED_PEAKS.json has no "
                       "integer peak)" if "error" not in micro else "nominal 148 x 64 x 1.965 GHz",
        "algorithmic_unit": f"{macs} limb-MACs/point (W_ref={w_ref} x A={ADD_COST[gid]} x "
                            f"L={MUL_MACS[gid]}, c_ref={c_ref}) x 2 IMAD slots",
        "kernel_ms": accum_ms, "kernel_share_of_step": accum_ms / phase_ms["total"],
        # the limb products the kernel really issues (signed digits, this run's digit size, XYZZ
        # mixed additions with dedicated squarings) against the same peak: always <= 1
        "executed": {"frac": n * engine_windows * executed_macs_per_item(gid) * 2
                     / (accum_ms * 1e-3) / 1e12 / peak,
                     "unit": f"n x {engine_windows} digits x {executed_macs_per_item(gid)} "
                             "limb products x 2 IMAD slots / kernel time / peak"},
        "pipe_utilisation_ncu": dict(zip(("fmaheavy_cycles_active", "source"),
                                         NCU_PIPE_UTIL.get((args.group, n), (None, None)))),
        # dram__bytes_read.sum + dram__bytes_write.sum of one launch, from the committed
        # `ncu --set full` capture of this workload (profiles/accum_first*_r01e_raw.csv)
        "traffic": NCU_TRAFFIC.get((args.group, n)),
        "traffic_unit": "bytes per launch (ncu dram read + write); algorithmic gather = items x "
                        "(affine point + 4-byte item)",
        "hbm_stream": {"achieved_GBps": n * MIN_BYTES[gid] / (phase_ms["total"] * 1e-3) / 1e9,
                       "peak_GBps": hbm_peak, "note": "secondary counter; the path is not HBM-bound"},
        "mont_mul_Gmul_per_s": {k: v for k, v in micro.items() if k.startswith("mont")},
    }

    cpu_baseline = None
    if not args.no_cpu_baseline:
        try:
            cpu_baseline = cpu_reference_arm(gid, n_total, cores, 2, 1, budget_s=15.0)
        except Exception as e:
            cpu_baseline = {"value": None, "unit": "points/s", "cores": cores, "kind": "reference",
                            "sample": f"unavailable: {e}"}

    line = {"metric": "MSM points/sec", "value": value, "unit": "points/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u32 limbs (254-bit Montgomery)" if gid < 2 else "u32 limbs",
            "data": "synthetic", "config": config, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline,
            "phases_ms": phase_ms, "window_bits": m.window_bits(gid, n), "verified": verified,
            "input_gen_s": gen_s}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
