"""CPU tests that pin the oracle (no GPU needed).

  * oracle/port.py (C++ restatement, oracle/msm_oracle.cpp) against tests/golden/reference_vectors.npz
    -- outputs of the UNMODIFIED reference produced by tests/golden/make_golden.py -- and against
    the literal known-answer vectors of the reference's own tests;
  * oracle/pymodel.py (Python big-int model) against the same vectors;
  * when oracle/_ref/libffref.so is present, the restatement against the live reference on fresh
    seeded inputs, including the raw (non-normalised) representative.
Bit-exact everywhere: integer arithmetic only on this path.
"""
import os

import numpy as np
import pytest

from oracle import port
from oracle import pymodel as pm

HERE = os.path.dirname(os.path.abspath(__file__))
FIELD_NAMES = ["alt_bn128_fq", "alt_bn128_fr", "bls12_377_fq", "bls12_377_fr", "bw6_761_fq",
               "bw6_761_fr", "alt_bn128_fq2", "bls12_377_fq2", "bls12_381_fq", "bls12_381_fr",
               "bls12_381_fq2"]
PRIME_FIELDS = {0: pm.ALT_BN128_Q, 1: pm.ALT_BN128_R, 2: pm.BLS12_377_Q, 3: pm.BLS12_377_R,
                4: pm.BW6_761_Q, 5: pm.BW6_761_R, 8: pm.BLS12_381_Q, 9: pm.BLS12_381_R}
GROUPS = [0, 1, 2, 3, 4, 5, 6, 7]
SCALAR_FIELD = port.FR_OF_GROUP


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "reference_vectors.npz"))


@pytest.fixture(scope="module")
def live_ref():
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref/libffref.so not built here")
    ref.lib()
    return ref


# ------------------------------------------------------------------------------------------------
# fields (fp.tcc / fp2.tcc)

@pytest.mark.parametrize("fid", range(11))
def test_port_field_ops_match_reference(golden, fid):
    name = FIELD_NAMES[fid]
    a, b = golden[f"field/{name}/a"], golden[f"field/{name}/b"]
    for op, opname in [(port.OP_MUL, "mul"), (port.OP_SQR, "sqr"), (port.OP_ADD, "add"),
                       (port.OP_SUB, "sub"), (port.OP_NEG, "neg")]:
        np.testing.assert_array_equal(port.field_op(fid, op, a, b), golden[f"field/{name}/{opname}"],
                                      err_msg=f"{name} {opname}")
    np.testing.assert_array_equal(port.field_op(fid, port.OP_INV, a[1:]), golden[f"field/{name}/inv"])
    if fid not in port.FQ2_FIELDS:
        plain = port.field_op(fid, port.OP_TO_PLAIN, a)
        np.testing.assert_array_equal(plain, golden[f"field/{name}/to_plain"])
        np.testing.assert_array_equal(port.field_op(fid, port.OP_FROM_PLAIN, plain), a)


@pytest.mark.parametrize("fid", sorted(PRIME_FIELDS))
def test_pymodel_field_matches_reference(golden, fid):
    name = FIELD_NAMES[fid]
    f = PRIME_FIELDS[fid]

    def dec(rows):
        return [f.from_mont(sum(int(r[j]) << (64 * j) for j in range(f.limbs64))) for r in rows]

    a, b = dec(golden[f"field/{name}/a"]), dec(golden[f"field/{name}/b"])
    assert dec(golden[f"field/{name}/mul"]) == [(x * y) % f.p for x, y in zip(a, b)]
    assert dec(golden[f"field/{name}/add"]) == [(x + y) % f.p for x, y in zip(a, b)]
    assert dec(golden[f"field/{name}/sub"]) == [(x - y) % f.p for x, y in zip(a, b)]
    assert dec(golden[f"field/{name}/neg"]) == [(-x) % f.p for x in a]
    assert dec(golden[f"field/{name}/inv"]) == [pow(x, -1, f.p) for x in a[1:]]
    plain = golden[f"field/{name}/to_plain"]
    assert [sum(int(r[j]) << (64 * j) for j in range(f.limbs64)) for r in plain] == a


def test_montgomery_constants_match_reference_literals():
    """inv = -p^-1 mod 2^64 literals of the init files (alt_bn128_init.cpp:62,103;
    bls12_377_init.cpp:81,127; bw6_761_init.cpp:115) -- SURVEY.md appendix A."""
    assert pm.ALT_BN128_R.inv64() == 0xc2e1f593efffffff
    assert pm.ALT_BN128_Q.inv64() == 0x87d20782e4866389
    assert pm.BLS12_377_R.inv64() == 0x0a117fffffffffff
    assert pm.BLS12_377_Q.inv64() == 0x8508bfffffffffff
    assert pm.BW6_761_Q.inv64() == 0x0a5593568fa798dd
    assert (pm.ALT_BN128_Q.bits, pm.BLS12_377_R.bits, pm.BLS12_377_Q.bits, pm.BW6_761_Q.bits) == \
        (254, 253, 377, 761)


# ------------------------------------------------------------------------------------------------
# scalar generator, digits, window size

@pytest.mark.parametrize("gid", GROUPS)
def test_sha512_rng(golden, gid):
    g = pm.GROUPS[gid]
    np.testing.assert_array_equal(port.gen_scalars(gid, 6), golden[f"rng/{gid}/first"])
    np.testing.assert_array_equal(port.gen_scalars(gid, 2, start=1 << 63),
                                  golden[f"rng/{gid}/walk_params"])
    np.testing.assert_array_equal(g.encode_scalars([pm.sha512_rng(g.fr, i) for i in range(6)]),
                                  golden[f"rng/{gid}/first"])
    assert g.decode_scalars(golden[f"rng/{gid}/walk_params"]) == list(pm.walk_params(g))


def test_field_get_digit_reference_kats():
    """test_field_get_digit_alt_bn128 (test_fields.cpp:283-346): digits of r - 1."""
    g = pm.ALT_BN128_G1
    v = g.fr.p - 1
    s = g.encode_scalars([v])[0]
    fid = port.F_ALT_BN128_FR
    assert [port.get_digit(fid, s, 2, i) for i in (127, 126, 125)] == [0, 3, 0]
    hex16 = [0x3064, 0x4e72, 0xe131, 0xa029, 0xb850, 0x45b6, 0x8181, 0x585d, 0x2833, 0xe848, 0x79b9,
             0x7091, 0x43e1, 0xf593, 0xf000, 0x0000]
    assert [port.get_digit(fid, s, 16, 15 - i) for i in range(16)] == hex16
    hex12 = [0x003, 0x064, 0x4e7, 0x2e1, 0x31a, 0x029, 0xb85, 0x045, 0xb68, 0x181, 0x585, 0xd28,
             0x33e, 0x848, 0x79b, 0x970, 0x914, 0x3e1, 0xf59, 0x3f0, 0x000, 0x000]
    assert [port.get_digit(fid, s, 12, 21 - i) for i in range(22)] == hex12
    assert [pm.get_digit(v, 12, 21 - i) for i in range(22)] == hex12


@pytest.mark.parametrize("fid", [port.F_ALT_BN128_FR, port.F_BLS12_377_FR, port.F_BW6_761_FR,
                                 port.F_BLS12_381_FR])
def test_signed_digits(golden, fid):
    """field_get_signed_digit / field_get_num_signed_digits (field_utils.tcc:102-239) and the
    recomposition property of do_test_signed_digits (test_fields.cpp:348-398)."""
    f = PRIME_FIELDS[fid]
    sc = golden[f"digits/{fid}/scalars"]
    plain = [f.from_mont(sum(int(r[j]) << (64 * j) for j in range(f.limbs64))) for r in sc]
    for c in (2, 3, 5, 8, 13, 16, 17, 20, 21):
        nd = int(golden[f"digits/{fid}/num_signed/{c}"][0])
        want = golden[f"digits/{fid}/signed/{c}"]
        assert port.num_signed_digits(fid, c) == nd
        got = np.array([[port.signed_digit(fid, s, c, i) for i in range(nd)] for s in sc])
        np.testing.assert_array_equal(got, want)
        for v, row in zip(plain, want):
            assert pm.signed_digits(v, c, nd) == list(row)
            assert all(-(1 << (c - 1)) <= d < (1 << (c - 1)) for d in row)
            assert sum(int(d) << (c * i) for i, d in enumerate(row)) == v


def test_pippenger_optimal_c(golden):
    """multiexp.tcc:35-40 incl. the size_t wrap-around for tiny n; SURVEY.md §8 row a2 values."""
    for n, c in zip(golden["pippenger_c/n"], golden["pippenger_c/c"]):
        assert port.pippenger_optimal_c(int(n)) == int(c)
        assert pm.pippenger_optimal_c(int(n)) == int(c)
    table = {1 << 16: 13, 1 << 20: 16, 1 << 21: 16, 1 << 22: 17, 1 << 23: 18, 1 << 24: 18,
             1 << 25: 19, 1 << 26: 20, 0: 2}
    for n, c in table.items():
        assert port.pippenger_optimal_c(n) == c


# ------------------------------------------------------------------------------------------------
# group law

@pytest.mark.parametrize("gid", GROUPS)
def test_port_group_law_matches_reference(golden, gid):
    """Raw X|Y|Z outputs (the representative, not only the element) of add / mixed_add / dbl / neg,
    including infinity operands, P + P and P + (-P) (test_groups.cpp:25-116 cases)."""
    k = f"group/{gid}/"
    zero, one = port.group_constants(gid)
    np.testing.assert_array_equal(zero, golden[k + "zero"])
    np.testing.assert_array_equal(one, golden[k + "one"])
    a, b, bs = golden[k + "a"], golden[k + "b"], golden[k + "b_special"]
    np.testing.assert_array_equal(port.group_op(gid, port.G_ADD, a, b), golden[k + "add"])
    np.testing.assert_array_equal(port.group_op(gid, port.G_MIXED_ADD, a, bs), golden[k + "mixed_add"])
    np.testing.assert_array_equal(port.group_op(gid, port.G_DBL, a), golden[k + "dbl"])
    np.testing.assert_array_equal(port.group_op(gid, port.G_NEG, a), golden[k + "neg"])
    np.testing.assert_array_equal(port.batch_to_special(gid, a), golden[k + "to_special"])
    np.testing.assert_array_equal(port.batch_to_special(gid, b), bs)
    assert list(port.group_eq(gid, a, b)) == [False] * 4 + [True, False, False, True, True]
    np.testing.assert_array_equal(port.gen_bases_walk(gid, 8), golden[k + "walk_bases"])
    np.testing.assert_array_equal(port.gen_bases_walk(gid, 3, start=5000),
                                  golden[k + "walk_bases_at_5000"])


@pytest.mark.parametrize("gid", GROUPS)
def test_pymodel_group_law_matches_reference(golden, gid):
    g = pm.GROUPS[gid]
    k = f"group/{gid}/"
    a = [g.decode(r) for r in golden[k + "a"]]
    b = [g.decode(r) for r in golden[k + "b"]]
    assert all(g.is_on_curve(P) for P in a + b)
    assert [g.decode(r) for r in golden[k + "add"]] == [g.add(P, Q) for P, Q in zip(a, b)]
    assert [g.decode(r) for r in golden[k + "mixed_add"]] == [g.add(P, Q) for P, Q in zip(a, b)]
    assert [g.decode(r) for r in golden[k + "dbl"]] == [g.add(P, P) for P in a]
    for row, P in zip(golden[k + "add_affine"], [g.add(P, Q) for P, Q in zip(a, b)]):
        np.testing.assert_array_equal(row, g.encode_affine(P))
    # arithmetic walk: P_i = (a + i d) G
    for i, row in enumerate(golden[k + "walk_bases"][:3]):
        np.testing.assert_array_equal(row, g.encode_affine(pm.walk_point(g, i)))
    np.testing.assert_array_equal(golden[k + "walk_bases_at_5000"][1],
                                  g.encode_affine(pm.walk_point(g, 5001)))
    np.testing.assert_array_equal(golden[k + "one"], g.encode_affine(g.gen))


# ------------------------------------------------------------------------------------------------
# multi_exp, batch_to_special, batch_exp

@pytest.mark.parametrize("gid", GROUPS)
def test_port_multi_exp_matches_reference(golden, gid):
    """The synthetic workload (SHA512_rng scalars, walk bases) at the reference test's sizes
    (test_multiexp.cpp:258-326) plus n = 0, for both methods, both base forms, several chunkings."""
    for n in [0, 1, 2, 3, 5, 64, 257, 1000]:
        key = f"msm/{gid}/{n}"
        if key not in golden:
            continue
        bases, scalars = port.gen_bases_walk(gid, n), port.gen_scalars(gid, n)
        want = golden[key]
        for method in (port.BDLO12, port.BDLO12_SIGNED):
            for form, chunks in ((port.FORM_SPECIAL, 1), (port.FORM_NORMAL, 1), (port.FORM_SPECIAL, 4)):
                if n > 300 and (form == port.FORM_NORMAL or chunks > 1):
                    continue
                got = port.multi_exp(gid, bases, scalars, method=method, form=form, chunks=chunks)
                np.testing.assert_array_equal(got, want, err_msg=f"n={n} m={method} f={form}")
    bases, scalars = port.gen_bases_walk(gid, 64), port.gen_scalars(gid, 64)
    for form, name in ((1, "special"), (0, "normal")):
        np.testing.assert_array_equal(
            port.multi_exp(gid, bases, scalars, form=form, normalize=False),
            golden[f"msm_raw/{gid}/64/{name}"])


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_closed_form(golden, gid):
    """sum s_i (a + i d) G computed with Python integers only (no MSM code at all)."""
    g = pm.GROUPS[gid]
    for n in (5, 64, 257):
        scalars = g.decode_scalars(port.gen_scalars(gid, n))
        k = pm.walk_expected_scalar(g, scalars)
        np.testing.assert_array_equal(golden[f"msm/{gid}/{n}"], g.encode_affine(g.mul(k, g.gen)))
    np.testing.assert_array_equal(golden[f"msm/{gid}/0"], g.encode_affine(None))


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_edge_cases(golden, gid):
    g = pm.GROUPS[gid]
    names = ["all_zero_scalars", "ones", "r_minus_1", "powers", "cancel", "dups", "dups_mixed",
             "infinity_bases", "all_infinity", "single"]
    for name in names:
        k = f"edge/{gid}/{name}/"
        bases, scalars, want = golden[k + "bases"], golden[k + "scalars"], golden[k + "result"]
        for method in (port.BDLO12, port.BDLO12_SIGNED):
            np.testing.assert_array_equal(port.multi_exp(gid, bases, scalars, method=method), want,
                                          err_msg=name)
        if gid != 3 or name in ("cancel", "all_infinity", "all_zero_scalars", "single"):
            pts = [g.decode(b) for b in bases]
            np.testing.assert_array_equal(g.encode_affine(g.msm(g.decode_scalars(scalars), pts)),
                                          want, err_msg=name)


def test_edge_case1_vectors(golden):
    """TestEdgeCase1 (test_multiexp.cpp:344-390): top signed digit negative and absorbing a carry."""
    g = pm.ALT_BN128_G1
    P0 = (0x157cbb5c3e5ff87af54877aefd2b3f95393ee3004a91c59c5c5023ca85aee7d0,
          0x162bb95fdd346c7276bb20ad526863d63c0eca0776fb57c71b76bd02dab6c554)
    P1 = (0x277a1d218ecc7a3ea26cec72b46c4f13b8b6594fcc080fa6d8ae61d43c34e084,
          0x13c48130e8519a34162bdc018409e98ca81683c23556314401a3acad24b1b9bc)
    bases, scalars = g.encode_points([P0, P1]), g.encode_scalars([1, 0x79D69])
    np.testing.assert_array_equal(bases, golden["edge_case1/bases"])
    want = g.encode_affine(g.add(P0, g.mul(0x79D69, P1)))
    np.testing.assert_array_equal(golden["edge_case1/result"], want)
    for method in (port.BDLO12, port.BDLO12_SIGNED):
        for form in (0, 1):
            np.testing.assert_array_equal(
                port.multi_exp(0, bases, scalars, method=method, form=form), want)


@pytest.mark.parametrize("gid", GROUPS)
def test_port_batch_to_special_and_batch_exp(golden, gid):
    np.testing.assert_array_equal(port.batch_to_special(gid, golden[f"to_special/{gid}/in"]),
                                  golden[f"to_special/{gid}/out"])
    bits, window = pm.GROUPS[gid].fr.bits, 4
    _, one = port.group_constants(gid)
    table = port.window_table(gid, bits, window, one)
    assert table.shape[0] == (bits + window - 1) // window
    np.testing.assert_array_equal(table[1], golden[f"batch_exp/{gid}/table_row1"])
    np.testing.assert_array_equal(table[-1], golden[f"batch_exp/{gid}/table_last_row"])
    sc = golden[f"batch_exp/{gid}/scalars"]
    np.testing.assert_array_equal(port.batch_exp(gid, bits, window, table, sc, normalize=False)[1],
                                  golden[f"batch_exp/{gid}/out_raw"])
    out = port.batch_exp(gid, bits, window, table, sc, normalize=True)[1]
    np.testing.assert_array_equal(out, golden[f"batch_exp/{gid}/out"])
    g = pm.GROUPS[gid]
    for s, row in zip(g.decode_scalars(sc)[:4], out[:4]):
        np.testing.assert_array_equal(row, g.encode_affine(g.mul(s, g.gen)))


# ------------------------------------------------------------------------------------------------
# live cross-check against the compiled reference (this container; skipped where it is absent)

@pytest.mark.parametrize("gid", GROUPS)
def test_port_equals_live_reference(live_ref, gid):
    ref = live_ref
    n = {0: 3000, 1: 700, 2: 1200, 3: 300, 4: 300, 5: 300, 6: 1200, 7: 300}[gid]
    bases = ref.gen_bases_walk(gid, n, start=123)
    scalars = ref.gen_scalars(gid, n, start=999)
    np.testing.assert_array_equal(port.gen_bases_walk(gid, n, start=123), bases)
    np.testing.assert_array_equal(port.gen_scalars(gid, n, start=999), scalars)
    for form in (0, 1):
        np.testing.assert_array_equal(
            port.multi_exp(gid, bases, scalars, form=form, normalize=False),
            ref.multi_exp(gid, bases, scalars, form=form, normalize=False))
    np.testing.assert_array_equal(
        port.multi_exp(gid, bases, scalars, method=port.BDLO12_SIGNED, chunks=5, normalize=False),
        ref.multi_exp(gid, bases, scalars, method=ref.BDLO12_SIGNED, chunks=5, normalize=False))
    J = ref.group_op(gid, ref.G_ADD, bases[:-1], bases[1:])
    np.testing.assert_array_equal(port.group_op(gid, port.G_ADD, bases[:-1], bases[1:]), J)
    np.testing.assert_array_equal(port.group_op(gid, port.G_DBL, J), ref.group_op(gid, ref.G_DBL, J))
    np.testing.assert_array_equal(port.batch_to_special(gid, J), ref.batch_to_special(gid, J))
    fid = ref.FQ_OF_GROUP[gid]
    w = ref.lib().ref_field_bytes(fid) // 8
    a = np.ascontiguousarray(J[:, :w])
    b = np.ascontiguousarray(J[:, w:2 * w])
    for op in (ref.OP_MUL, ref.OP_SQR, ref.OP_ADD, ref.OP_SUB, ref.OP_NEG, ref.OP_INV):
        np.testing.assert_array_equal(port.field_op(fid, op, a, b), ref.field_op(fid, op, a, b))


# ------------------------------------------------------------------------------------------------
# wire format of the reference's ffi layer

@pytest.mark.parametrize("gid", GROUPS)
def test_wire_format_model_matches_reference_serializer(golden, gid):
    """ffi::group_element_write / field_element_write (ffi/ffi_serialization.tcc:98-187): big-endian
    plain values, Fq2 as c1 | c0, affine X | Y, (0, 1) for infinity."""
    g = pm.GROUPS[gid]
    pts = [g.decode(r) for r in golden[f"wire/{gid}/points"]]
    want = golden[f"wire/{gid}/point_bytes"].tobytes()
    assert b"".join(pm.wire_point(g, P) for P in pts) == want
    assert pts[2] is None and pm.wire_point(g, None) == want[2 * len(want) // 4:3 * len(want) // 4]
    sc = g.decode_scalars(golden[f"wire/{gid}/scalars"])
    assert b"".join(pm.wire_scalar(g, s) for s in sc) == golden[f"wire/{gid}/scalar_bytes"].tobytes()
    assert len(want) // 4 == {0: 64, 1: 128, 2: 96, 3: 192, 4: 192, 5: 192, 6: 96, 7: 192}[gid]  # ffi/ffi.h:13-17,55-59 sizes


# ------------------------------------------------------------------------------------------------
# streamed / precomputed MSM (multiexp_stream.tcc:164-223) and the binary stream encoding

DIVIDING_C = {0: 2, 1: 2, 2: 11, 3: 13, 4: 11, 5: 13, 6: 15, 7: 17}


@pytest.mark.parametrize("gid", GROUPS)
def test_stream_encoding_model_matches_reference_writer(golden, gid):
    """group_write<encoding_binary, form, compression_off> (curve_serialization.tcc:79-101) against
    the Python big-int model and the C++ restatement"""
    g = pm.GROUPS[gid]
    bases = golden[f"stream/{gid}/bases"]
    pts = [g.decode(r) for r in bases]
    assert pts[4] is None
    for form in (0, 1):
        want = golden[f"stream/{gid}/bytes_form{form}"].tobytes()
        assert b"".join(pm.stream_point(g, P, bool(form)) for P in pts) == want
        assert port.stream_write(gid, bases, form=form) == want
    assert len(want) == len(pts) * 2 * g.coord_bytes


@pytest.mark.parametrize("gid", GROUPS)
def test_port_stream_msm_matches_reference(golden, gid):
    """multi_exp_stream: the raw (non-normalised) representative of the restatement equals the
    compiled reference's limb for limb, and both are the plain multi_exp element; where the
    reference sums an empty bucket set (undefined there) the restatement says so too"""
    sc = golden[f"stream/{gid}/scalars"]
    for form in (0, 1):
        data = golden[f"stream/{gid}/bytes_form{form}"].tobytes()
        got = port.multi_exp_stream(gid, data, sc, form=form, normalize=False)
        if golden[f"stream/{gid}/defined_form{form}"][0]:
            np.testing.assert_array_equal(got, golden[f"stream/{gid}/result_raw_form{form}"])
            np.testing.assert_array_equal(port.multi_exp_stream(gid, data, sc, form=form),
                                          golden[f"stream/{gid}/plain_msm"])
        else:
            assert got is None
    # 24 points, c = 8: 377-bit scalars leave the top digit position empty, 253..255-bit ones do not
    assert bool(golden[f"stream/{gid}/defined_form1"][0]) == (gid not in (3, 5))


@pytest.mark.parametrize("gid", GROUPS)
def test_port_precomputed_stream_msm_matches_reference(golden, gid):
    """multi_exp_stream_with_precompute on the reference's own precompute stream (each base followed
    by its multiples).  For a digit size dividing the scalar width the reference drops the carry
    out of the top digit of r - 1 and r - 2 (field_utils.tcc:205-239): pinned here as the
    reference's behaviour; the engine folds such scalars and returns the true sum."""
    sc = golden[f"stream/{gid}/scalars"][1:4]  # 1, r - 1, r - 2
    for c in (9, DIVIDING_C[gid]):
        data = golden[f"stream/{gid}/pre{c}_bytes"].tobytes()
        assert port.stream_write(gid, golden[f"stream/{gid}/bases"][:3], form=1, precompute_c=c) == data
        got = port.multi_exp_stream(gid, data, sc, form=1, precompute_c=c, normalize=False)
        np.testing.assert_array_equal(got, golden[f"stream/{gid}/pre{c}_result_raw"])
        norm = port.multi_exp_stream(gid, data, sc, form=1, precompute_c=c)
        same = np.array_equal(norm, golden[f"stream/{gid}/pre{c}_plain_msm"])
        assert same == (c != DIVIDING_C[gid])
