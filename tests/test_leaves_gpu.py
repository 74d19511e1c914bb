"""GPU parity of the leaves of the path against the unmodified reference (SURVEY.md §8a rows
a7-a13): Montgomery field arithmetic (fp.tcc:50-727, fp2.tcc:79-168), as_bigint, and the group law
incl. its exceptional cases (test_groups.cpp:25-116 is the model: 0+0, 0+P, P+0, P+Q, P+P, P-P)."""
import os

import numpy as np
import pytest

from oracle import pymodel as pm

pytestmark = pytest.mark.gpu
GROUPS = [0, 1, 2, 3, 4, 5, 6, 7]


def rand_field(ref, fid, n, seed):
    """n canonical Montgomery elements of field fid (any limbs < p are valid elements)."""
    words = ref.lib().ref_field_bytes(fid) // 8
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(n, words), dtype=np.uint64) * np.uint64(2) + \
        rng.integers(0, 2, size=(n, words), dtype=np.uint64)
    # reduce through the reference itself: x -> x * 1 keeps it canonical (mul_reduce output < p)
    a[:, -1] &= np.uint64(0x00FFFFFFFFFFFFFF)
    if fid in ref.FQ2_FIELDS:
        a[:, words // 2 - 1] &= np.uint64(0x00FFFFFFFFFFFFFF)
    one = ref.field_op(fid, ref.OP_FROM_PLAIN, np.tile(np.eye(1, words, 0, dtype=np.uint64), (n, 1))) \
        if fid not in ref.FQ2_FIELDS else None
    return ref.field_op(fid, ref.OP_MUL, a, one) if one is not None else ref.field_op(fid, ref.OP_ADD, ref.field_op(fid, ref.OP_SQR, a), a)


@pytest.mark.parametrize("gid", GROUPS)
def test_coordinate_field_ops(engine, ref, gid):
    fid = ref.FQ_OF_GROUP[gid]
    n = 257
    a = rand_field(ref, fid, n, 1)
    b = rand_field(ref, fid, n, 2)
    a[0] = 0
    b[1] = 0
    b[2] = a[2]
    for eop, rop in [(engine.FQ_ADD, ref.OP_ADD), (engine.FQ_SUB, ref.OP_SUB),
                     (engine.FQ_MUL, ref.OP_MUL)]:
        np.testing.assert_array_equal(engine.debug_op(gid, eop, a, b), ref.field_op(fid, rop, a, b),
                                      err_msg=f"op {eop}")
    for eop, rop in [(engine.FQ_SQR, ref.OP_SQR), (engine.FQ_NEG, ref.OP_NEG)]:
        np.testing.assert_array_equal(engine.debug_op(gid, eop, a), ref.field_op(fid, rop, a),
                                      err_msg=f"op {eop}")
    nz = a[3:40]
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_INV, nz),
                                  ref.field_op(fid, ref.OP_INV, nz))


def test_fp64_pipe_product_equals_integer_product():
    """The measured-and-rejected DFMA Montgomery product (tools/fp64_product/fp_dfma.cuh; not in
    the library) still computes the same function as Fp::mul -- itself pinned to mul_reduce
    (fp.tcc:50-228) by the test above -- for all seven moduli, bit for bit, incl. 0, 1, p-1."""
    import subprocess
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools",
                       "fp64_product", "dfma_check")
    if not os.path.exists(exe):
        pytest.skip("tools/fp64_product/dfma_check not built (make -C libff_b200/csrc dfma_check)")
    out = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "dfma_check ok" in out.stdout, out.stdout + out.stderr


@pytest.mark.parametrize("gid", GROUPS)
def test_shared_reduction_product_and_squaring(engine, ref, gid):
    """Fp::mul2 (a b + c d with one Montgomery reduction) and the dedicated Fp::sqr against the
    reference's field operations (fp.tcc:50-228, 632-677), incl. 0, 1 and p-1 operands."""
    fid = ref.FQ_OF_GROUP[gid]
    n = 4099
    a = rand_field(ref, fid, n, 21)
    b = rand_field(ref, fid, n, 22)
    minus_one = ref.field_op(fid, ref.OP_NEG, _one(ref, fid))
    a[0] = 0
    b[1] = 0
    b[2] = a[2]
    a[3], b[3] = minus_one[0], minus_one[0]
    a[4], b[4] = minus_one[0], _one(ref, fid)[0]
    a[5] = _one(ref, fid)[0]
    want = ref.field_op(fid, ref.OP_ADD, ref.field_op(fid, ref.OP_MUL, a, b),
                        ref.field_op(fid, ref.OP_MUL, ref.field_op(fid, ref.OP_ADD, a, b),
                                     ref.field_op(fid, ref.OP_SUB, a, b)))
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_MUL2, a, b), want)
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_SQR, a),
                                  ref.field_op(fid, ref.OP_SQR, a))


@pytest.mark.parametrize("gid", GROUPS)
def test_product_code_shapes_agree(engine, ref, gid):
    """The Montgomery product exists in three code shapes chosen per kernel (fp.cuh CALL_MIN:
    fully unrolled, rolled row pairs, one shared called copy); each must give the reference's
    product (fp.tcc:50-228) bit for bit, incl. 0, 1 and p-1 operands, also through the
    shared-reduction form a b + c d and the dedicated squaring."""
    fid = ref.FQ_OF_GROUP[gid]
    n = 2051
    a = rand_field(ref, fid, n, 31)
    b = rand_field(ref, fid, n, 32)
    one = _one(ref, fid)[0]
    minus_one = ref.field_op(fid, ref.OP_NEG, _one(ref, fid))[0]
    a[0] = 0
    b[1] = 0
    b[2] = a[2]
    a[3], b[3] = minus_one, minus_one
    a[4], b[4] = minus_one, one
    a[5] = one
    prod = ref.field_op(fid, ref.OP_MUL, a, b)
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_MUL_ROLLED, a, b), prod)
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_MUL_CALLED, a, b), prod)
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_SQR_CALLED, a, b),
                                  ref.field_op(fid, ref.OP_SQR, a))
    want2 = ref.field_op(fid, ref.OP_ADD, prod,
                         ref.field_op(fid, ref.OP_MUL, ref.field_op(fid, ref.OP_ADD, a, b),
                                      ref.field_op(fid, ref.OP_SUB, a, b)))
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_MUL2_ROLLED, a, b), want2)


BASE_GROUP = {0: 0, 1: 0, 2: 2, 3: 3, 4: 2, 5: 3, 6: 6, 7: 6}  # a group whose Fq is the base prime field


@pytest.mark.parametrize("gid", GROUPS)
def test_separated_product_and_reduction(engine, ref, gid):
    """Fp::mul_wide (plain 2N-limb integer product) followed by Fp::redc_wide (Montgomery reduction
    of a double-width value) -- the halves the lazy-reduction Fq2 product is built from -- give the
    reference's product (fp.tcc:50-228) on the base prime field, incl. 0, 1, p-1 operands."""
    fid = ref.FQ_OF_GROUP[gid]
    bfid = ref.FQ_OF_GROUP[BASE_GROUP[gid]]
    w = ref.lib().ref_field_bytes(bfid) // 8
    n = 1025
    a = rand_field(ref, fid, n, 41)
    b = rand_field(ref, fid, n, 42)
    minus_one = ref.field_op(bfid, ref.OP_NEG, _one(ref, bfid))[0]
    a2, b2 = a.reshape(-1, w), b.reshape(-1, w)
    a2[0] = 0
    b2[1] = 0
    a2[2], b2[2] = minus_one, minus_one
    a2[3], b2[3] = minus_one, _one(ref, bfid)[0]
    got = engine.debug_op(gid, engine.FQ_MUL_SEPARATED, a2.reshape(a.shape), b2.reshape(b.shape))
    np.testing.assert_array_equal(got.reshape(-1, w), ref.field_op(bfid, ref.OP_MUL, a2, b2))


@pytest.mark.parametrize("gid", [1, 4, 7])
def test_lazy_fq2_product_corner_values(engine, ref, gid):
    """The lazy-reduction Fq2 product (fp2.cuh mul_lazy: differences of unreduced products, p R added
    back on a borrow, five-fold for bls12_377's non-residue) on coordinates 0, 1 and p - 1 in every
    combination, against Fp2_model::operator* (fp2.tcc:101-114)."""
    fid = ref.FQ_OF_GROUP[gid]
    bfid = ref.FQ_OF_GROUP[BASE_GROUP[gid]]
    w = ref.lib().ref_field_bytes(bfid) // 8
    one = _one(ref, bfid)[0]
    minus_one = ref.field_op(bfid, ref.OP_NEG, _one(ref, bfid))[0]
    vals = [np.zeros(w, dtype=np.uint64), one, minus_one, rand_field(ref, bfid, 1, 5)[0]]
    rows_a, rows_b = [], []
    for x0 in vals:
        for x1 in vals:
            for y0 in vals:
                for y1 in vals:
                    rows_a.append(np.concatenate([x0, x1]))
                    rows_b.append(np.concatenate([y0, y1]))
    a, b = np.array(rows_a, dtype=np.uint64), np.array(rows_b, dtype=np.uint64)
    prod = ref.field_op(fid, ref.OP_MUL, a, b)
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_MUL, a, b), prod)
    # a b + (a + b)(a - b) through the two-product form that shares its reductions (mul2_lazy)
    want2 = ref.field_op(fid, ref.OP_ADD, prod,
                         ref.field_op(fid, ref.OP_MUL, ref.field_op(fid, ref.OP_ADD, a, b),
                                      ref.field_op(fid, ref.OP_SUB, a, b)))
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FQ_MUL2, a, b), want2)


def _one(ref, fid):
    """Montgomery one of field fid as a (1, words) array (Fq2: (1, 0))."""
    words = ref.lib().ref_field_bytes(fid) // 8
    if fid in ref.FQ2_FIELDS:
        half = words // 2
        plain = np.zeros((1, half), dtype=np.uint64)
        plain[0, 0] = 1
        one = ref.field_op(ref.FQ2_BASE[fid], ref.OP_FROM_PLAIN, plain)
        return np.concatenate([one, np.zeros_like(one)], axis=1)
    plain = np.zeros((1, words), dtype=np.uint64)
    plain[0, 0] = 1
    return ref.field_op(fid, ref.OP_FROM_PLAIN, plain)


@pytest.mark.parametrize("gid", GROUPS)
def test_scalar_field_as_bigint(engine, ref, gid):
    g = pm.GROUPS[gid]
    fid = ref.FR_OF_GROUP[gid]
    vals = [0, 1, 2, g.fr.p - 1, g.fr.p - 2, 1 << 64, (1 << 200) + 12345] + \
        [pm.sha512_rng(g.fr, i) for i in range(50)]
    s = g.encode_scalars(vals)
    got = engine.debug_op(gid, engine.FR_FROM_MONT, s)
    np.testing.assert_array_equal(got, ref.field_op(fid, ref.OP_TO_PLAIN, s))
    n64 = g.fr.limbs64
    for row, v in zip(got, vals):
        assert sum(int(row[j]) << (64 * j) for j in range(n64)) == v % g.fr.p
    np.testing.assert_array_equal(engine.debug_op(gid, engine.FR_MUL, s, s[::-1].copy()),
                                  ref.field_op(fid, ref.OP_MUL, s, s[::-1].copy()))


@pytest.mark.parametrize("gid", GROUPS)
def test_group_law(engine, ref, gid):
    n = 64
    aff = ref.gen_bases_walk(gid, n + 1)
    P = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])   # non-normalised
    Q = ref.group_op(gid, ref.G_DBL, aff[1:])
    zero, _ = ref.group_constants(gid)
    # exceptional rows: 0+0, 0+Q, P+0, P+P (different representatives), P+(-P)
    P[0], Q[0] = zero, zero
    P[1] = zero
    negP = ref.group_op(gid, ref.G_NEG, P)
    Q[2] = zero
    Q[3] = ref.batch_to_special(gid, P[3:4])[0]
    Q[4] = negP[4]
    Q[5] = ref.batch_to_special(gid, negP[5:6])[0]

    def same(got, want):
        assert ref.group_eq(gid, got, want).all()
        np.testing.assert_array_equal(ref.batch_to_special(gid, got), ref.batch_to_special(gid, want))

    same(engine.debug_op(gid, engine.G_ADD, P, Q), ref.group_op(gid, ref.G_ADD, P, Q))
    same(engine.debug_op(gid, engine.G_DBL, P), ref.group_op(gid, ref.G_DBL, P))
    same(engine.debug_op(gid, engine.G_NEG, P), negP)
    Qs = ref.batch_to_special(gid, Q)
    same(engine.debug_op(gid, engine.G_MADD, P, Qs), ref.group_op(gid, ref.G_MIXED_ADD, P, Qs))
