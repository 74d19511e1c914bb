#!/usr/bin/env python3
"""Generates tests/golden/reference_vectors.npz from the UNMODIFIED reference.

Run in the container that has /root/reference:   make -C oracle ref && python tests/golden/make_golden.py
Every array is an output of oracle/_ref/libffref.so (clearmatics/libff compiled as is, see
oracle/Makefile) on the inputs stored next to it, in the reference's in-memory layout (uint64
little-endian Montgomery limbs; points X|Y|Z).  The GPU box has no /root/reference: tests there use
this file, never the reference sources.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import pymodel as pm  # noqa: E402
from oracle import ref  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors.npz")
FIELD_NAMES = ["alt_bn128_fq", "alt_bn128_fr", "bls12_377_fq", "bls12_377_fr", "bw6_761_fq",
               "bw6_761_fr", "alt_bn128_fq2", "bls12_377_fq2", "bls12_381_fq", "bls12_381_fr",
               "bls12_381_fq2"]
FIELD_MODEL = [pm.ALT_BN128_Q, pm.ALT_BN128_R, pm.BLS12_377_Q, pm.BLS12_377_R, pm.BW6_761_Q,
               pm.BW6_761_R, pm.ALT_BN128_Q, pm.BLS12_377_Q, pm.BLS12_381_Q, pm.BLS12_381_R,
               pm.BLS12_381_Q]
MSM_SIZES = [0, 1, 2, 3, 5, 64, 257, 1000]


def field_elements(fid, count=10):
    """Montgomery-form elements: 0, 1, p-1, 2, then SHA512_rng-style pseudo-random values."""
    f = FIELD_MODEL[fid]
    deg = 2 if fid in ref.FQ2_FIELDS else 1
    vals = [0, 1, f.p - 1, 2] + [pm.sha512_rng(f, 7000 + 31 * fid + i) for i in range(count * deg)]
    rows = []
    for i in range(count):
        limbs = []
        for k in range(deg):
            m = f.to_mont(vals[(i * deg + k) % len(vals)] if i >= 4 or k == 0 else 0)
            limbs += [(m >> (64 * j)) & 0xFFFFFFFFFFFFFFFF for j in range(f.limbs64)]
        rows.append(limbs)
    return np.array(rows, dtype=np.uint64)


def main():
    ref.lib()
    out = {}
    # ---- field leaves (fp.tcc / fp2.tcc)
    for fid, name in enumerate(FIELD_NAMES):
        a = field_elements(fid)
        b = np.roll(a, 3, axis=0).copy()
        out[f"field/{name}/a"] = a
        out[f"field/{name}/b"] = b
        for op, opname in [(ref.OP_MUL, "mul"), (ref.OP_SQR, "sqr"), (ref.OP_ADD, "add"),
                           (ref.OP_SUB, "sub"), (ref.OP_NEG, "neg")]:
            out[f"field/{name}/{opname}"] = ref.field_op(fid, op, a, b)
        nz = a[1:]  # inverse of zero is undefined in the reference
        out[f"field/{name}/inv"] = ref.field_op(fid, ref.OP_INV, nz)
        if fid not in ref.FQ2_FIELDS:
            out[f"field/{name}/to_plain"] = ref.field_op(fid, ref.OP_TO_PLAIN, a)
    # ---- scalar generator, digits, window size
    for gid in range(8):
        idx = [0, 1, 2, 5]
        out[f"rng/{gid}/first"] = ref.gen_scalars(gid, 6)
        out[f"rng/{gid}/walk_params"] = ref.gen_scalars(gid, 2, start=1 << 63)
    for fid in (ref.F_ALT_BN128_FR, ref.F_BLS12_377_FR, ref.F_BW6_761_FR, ref.F_BLS12_381_FR):
        g = {ref.F_ALT_BN128_FR: 0, ref.F_BLS12_377_FR: 2, ref.F_BW6_761_FR: 3,
             ref.F_BLS12_381_FR: 6}[fid]
        f = FIELD_MODEL[fid]
        G = pm.GROUPS[g]
        sc = np.concatenate([G.encode_scalars([f.p - 1, f.p - 2, 0, 1, (1 << 200) - 1, 0x79D69]),
                             ref.gen_scalars(g, 4, start=40)])
        out[f"digits/{fid}/scalars"] = sc
        for c in (2, 3, 5, 8, 13, 16, 17, 20, 21):
            nd = ref.num_signed_digits(fid, c)
            out[f"digits/{fid}/num_signed/{c}"] = np.array([nd], dtype=np.int64)
            out[f"digits/{fid}/signed/{c}"] = np.array(
                [[ref.signed_digit(fid, s, c, i) for i in range(nd)] for s in sc], dtype=np.int64)
    ns = list(range(0, 70)) + [255, 256, 257, 1000, 65535, 65536, 65537, 1 << 20, (1 << 20) + 1,
                               1 << 22, 1 << 24, 1 << 26, 1 << 30]
    out["pippenger_c/n"] = np.array(ns, dtype=np.uint64)
    out["pippenger_c/c"] = np.array([ref.pippenger_optimal_c(n) for n in ns], dtype=np.uint64)
    # ---- group law (raw representatives) incl. the exceptional cases
    for gid in range(8):
        zero, one = ref.group_constants(gid)
        P = ref.gen_bases_walk(gid, 6)
        J = ref.group_op(gid, ref.G_ADD, P[:5], P[1:6])      # non-normalised points
        a = np.stack([J[0], J[1], J[2], zero, J[3], J[0], J[4], one, zero])
        b = np.stack([J[1], J[2], zero, J[3], J[3], ref.group_op(gid, ref.G_NEG, J[:1])[0], P[0],
                      one, zero])
        bs = ref.batch_to_special(gid, b)                     # special operands for mixed_add
        out[f"group/{gid}/zero"], out[f"group/{gid}/one"] = zero, one
        out[f"group/{gid}/a"], out[f"group/{gid}/b"], out[f"group/{gid}/b_special"] = a, b, bs
        out[f"group/{gid}/add"] = ref.group_op(gid, ref.G_ADD, a, b)
        out[f"group/{gid}/mixed_add"] = ref.group_op(gid, ref.G_MIXED_ADD, a, bs)
        out[f"group/{gid}/dbl"] = ref.group_op(gid, ref.G_DBL, a)
        out[f"group/{gid}/neg"] = ref.group_op(gid, ref.G_NEG, a)
        out[f"group/{gid}/to_special"] = ref.batch_to_special(gid, a)
        out[f"group/{gid}/add_affine"] = ref.batch_to_special(gid, out[f"group/{gid}/add"])
        out[f"group/{gid}/mixed_add_affine"] = ref.batch_to_special(gid, out[f"group/{gid}/mixed_add"])
        out[f"group/{gid}/dbl_affine"] = ref.batch_to_special(gid, out[f"group/{gid}/dbl"])
        out[f"group/{gid}/walk_bases"] = ref.gen_bases_walk(gid, 8)
        out[f"group/{gid}/walk_bases_at_5000"] = ref.gen_bases_walk(gid, 3, start=5000)
    # ---- multi_exp on the synthetic workload (gen_scalars + gen_bases_walk), affine results
    for gid in range(8):
        for n in MSM_SIZES:
            if gid == 3 and n > 300:
                continue
            bases, scalars = ref.gen_bases_walk(gid, n), ref.gen_scalars(gid, n)
            r = ref.multi_exp(gid, bases, scalars, method=ref.BDLO12, form=ref.FORM_SPECIAL)
            r2 = ref.multi_exp(gid, bases, scalars, method=ref.BDLO12_SIGNED, form=ref.FORM_NORMAL,
                               chunks=3)
            r3 = ref.multi_exp(gid, bases, scalars, method=ref.NAIVE_PLAIN)
            assert np.array_equal(r, r2) and np.array_equal(r, r3)
            out[f"msm/{gid}/{n}"] = r
        # raw (non-normalised) outputs pin the restatement's representative too
        bases, scalars = ref.gen_bases_walk(gid, 64), ref.gen_scalars(gid, 64)
        out[f"msm_raw/{gid}/64/special"] = ref.multi_exp(gid, bases, scalars, form=1, normalize=False)
        out[f"msm_raw/{gid}/64/normal"] = ref.multi_exp(gid, bases, scalars, form=0, normalize=False)
    # ---- exceptional inputs (SURVEY.md §8d edge-case suite)
    for gid in range(8):
        G = pm.GROUPS[gid]
        r = G.fr.p
        zero, _ = ref.group_constants(gid)
        P = ref.gen_bases_walk(gid, 8)
        negP = ref.group_op(gid, ref.G_NEG, P)
        enc = G.encode_scalars
        cases = {
            "all_zero_scalars": (P, enc([0] * 8)),
            "ones": (P, enc([1] * 8)),
            "r_minus_1": (P, enc([r - 1] * 8)),
            "powers": (P, enc([2, 1 << 31, 1 << 32, 1 << 63, 1 << 64, (1 << 100) - 1, 2, 3])),
            "cancel": (np.stack([P[0], negP[0]]), enc([12345, 12345])),
            "dups": (np.stack([P[0]] * 4), enc([99, 99, 99, 99])),
            "dups_mixed": (np.stack([P[0], P[0], negP[0], P[1], P[1]]), enc([5, 5, 5, r - 1, r - 1])),
            "infinity_bases": (np.stack([zero, P[0], zero, P[1]]), enc([3, 4, 5, 6])),
            "all_infinity": (np.stack([zero] * 3), enc([3, 4, 5])),
            "single": (P[:1], enc([r - 1])),
        }
        for name, (bases, scalars) in cases.items():
            out[f"edge/{gid}/{name}/bases"] = bases
            out[f"edge/{gid}/{name}/scalars"] = scalars
            out[f"edge/{gid}/{name}/result"] = ref.multi_exp(gid, bases, scalars, method=ref.BDLO12)
    # ---- TestEdgeCase1 (test_multiexp.cpp:344-390): the reference's literal vectors
    g = pm.ALT_BN128_G1
    P0 = (0x157cbb5c3e5ff87af54877aefd2b3f95393ee3004a91c59c5c5023ca85aee7d0,
          0x162bb95fdd346c7276bb20ad526863d63c0eca0776fb57c71b76bd02dab6c554)
    P1 = (0x277a1d218ecc7a3ea26cec72b46c4f13b8b6594fcc080fa6d8ae61d43c34e084,
          0x13c48130e8519a34162bdc018409e98ca81683c23556314401a3acad24b1b9bc)
    eb, es = g.encode_points([P0, P1]), g.encode_scalars([1, 0x79D69])
    out["edge_case1/bases"], out["edge_case1/scalars"] = eb, es
    out["edge_case1/result"] = ref.multi_exp(0, eb, es, method=ref.BDLO12_SIGNED, form=ref.FORM_NORMAL)
    # ---- batch_to_special and batch_exp
    for gid in range(8):
        P = ref.gen_bases_walk(gid, 13)
        J = ref.group_op(gid, ref.G_ADD, P[:12], P[1:13])
        zero, one = ref.group_constants(gid)
        J[3] = zero
        J[7] = P[7]
        out[f"to_special/{gid}/in"] = J
        out[f"to_special/{gid}/out"] = ref.batch_to_special(gid, J)
        bits, window = G_BITS[gid], 4
        table = ref.window_table(gid, bits, window, one)
        sc = np.concatenate([pm.GROUPS[gid].encode_scalars([0, 1, pm.GROUPS[gid].fr.p - 1]),
                             ref.gen_scalars(gid, 5, start=9)])
        out[f"batch_exp/{gid}/table_row1"] = table[1]       # 16 entries of the second row (raw)
        out[f"batch_exp/{gid}/table_last_row"] = table[-1]
        out[f"batch_exp/{gid}/scalars"] = sc
        out[f"batch_exp/{gid}/out_raw"] = ref.batch_exp(gid, bits, window, table, sc, normalize=False)[1]
        out[f"batch_exp/{gid}/out"] = ref.batch_exp(gid, bits, window, table, sc, normalize=True)[1]
    # ---- wire format of the reference's ffi layer (ffi/ffi_serialization.tcc:98-187)
    for gid in range(8):
        zero, one = ref.group_constants(gid)
        P = ref.gen_bases_walk(gid, 3)
        J = ref.group_op(gid, ref.G_ADD, P[:2], P[1:3])  # non-normalised: the writer normalises
        pts = np.stack([P[0], J[0], zero, one])
        sc = np.concatenate([pm.GROUPS[gid].encode_scalars([0, 1, pm.GROUPS[gid].fr.p - 1]),
                             ref.gen_scalars(gid, 2, start=77)])
        out[f"wire/{gid}/points"] = pts
        out[f"wire/{gid}/point_bytes"] = np.frombuffer(ref.wire_group_write(gid, pts), dtype=np.uint8)
        out[f"wire/{gid}/scalars"] = sc
        out[f"wire/{gid}/scalar_bytes"] = np.frombuffer(ref.wire_scalar_write(gid, sc), dtype=np.uint8)
    # ---- streamed / precomputed MSM and the binary stream encoding (multiexp_stream.tcc:164-223,
    # curve_serialization.tcc:79-101).  "defined" = 0 where the reference sums an empty bucket set.
    for gid in range(8):
        G = pm.GROUPS[gid]
        r = G.fr.p
        zero, _ = ref.group_constants(gid)
        P = ref.gen_bases_walk(gid, 24)
        P[4] = zero
        sc = ref.gen_scalars(gid, 24, start=4000)
        sc[:6] = G.encode_scalars([0, 1, r - 1, r - 2, 1 << (r.bit_length() - 1), 2])
        out[f"stream/{gid}/bases"] = P
        out[f"stream/{gid}/scalars"] = sc
        for form in (0, 1):
            data = ref.stream_write(gid, P, form=form)
            out[f"stream/{gid}/bytes_form{form}"] = np.frombuffer(data, dtype=np.uint8)
            res = ref.multi_exp_stream(gid, data, sc, form=form, normalize=False)
            out[f"stream/{gid}/defined_form{form}"] = np.array([res is not None], dtype=np.uint8)
            if res is not None:
                out[f"stream/{gid}/result_raw_form{form}"] = res
        out[f"stream/{gid}/plain_msm"] = ref.multi_exp(gid, P, sc)
        for c in (9, DIVIDING_C[gid]):
            m = 3
            data = ref.stream_write(gid, P[:m], form=1, precompute_c=c)
            out[f"stream/{gid}/pre{c}_bytes"] = np.frombuffer(data, dtype=np.uint8)
            res = ref.multi_exp_stream(gid, data, sc[1:1 + m], form=1, precompute_c=c, normalize=False)
            out[f"stream/{gid}/pre{c}_result_raw"] = res
            out[f"stream/{gid}/pre{c}_plain_msm"] = ref.multi_exp(gid, P[:m], sc[1:1 + m])
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes,", len(out), "arrays")


G_BITS = {0: 254, 1: 254, 2: 253, 3: 377, 4: 253, 5: 377, 6: 255, 7: 255}
# digit sizes that divide the scalar width: the reference drops the carry out of the top digit
DIVIDING_C = {0: 2, 1: 2, 2: 11, 3: 13, 4: 11, 5: 13, 6: 15, 7: 17}

if __name__ == "__main__":
    main()
