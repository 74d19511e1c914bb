"""Streamed and precomputed MSM on the GPU (b200msm_stream_*, b200msm_precompute*,
b200msm_multi_exp_precomputed) against the reference's multi_exp_stream and
multi_exp_stream_with_precompute (multiexp_stream.tcc:164-223) run on the reference's own binary
stream encoding (curve_serialization.tcc:79-101), both through oracle/_ref.  Bit-exact on the
affine result."""
import numpy as np
import pytest

from oracle import pymodel as pm

pytestmark = pytest.mark.gpu
GROUPS = [0, 1, 2, 3, 4, 5, 6, 7]
SMALL_N = {0: 600, 1: 200, 2: 300, 3: 120, 4: 120, 5: 120, 6: 300, 7: 120}
# digit sizes c with W * c == scalar bits: the top signed digit can carry out (the reference drops
# that carry, field_utils.tcc:205-239; the engine folds the scalar to r - s)
DIVIDING_C = {0: 2, 1: 2, 2: 11, 3: 13, 4: 11, 5: 13, 6: 15, 7: 17}


def affine(engine, gid, pt):
    return engine.batch_to_special(gid, np.asarray(pt).reshape(1, -1))[0]


def edge_scalars(ref, gid, n):
    """SHA512_rng scalars with the extremes mixed in (0, 1, r - 1, r - 2, 2^k)"""
    g = pm.GROUPS[gid]
    s = ref.gen_scalars(gid, n, start=900)
    r = g.fr.p
    special = [0, 1, r - 1, r - 2, 1 << (r.bit_length() - 1), (1 << (r.bit_length() - 1)) - 1, 2]
    k = min(len(special), n)
    if k:
        s[:k] = g.encode_scalars(special[:k])
    return s


@pytest.mark.parametrize("gid", GROUPS)
def test_precomputed_on_device_matches_plain_msm(engine, ref, gid):
    n = SMALL_N[gid]
    bases = ref.gen_bases_walk(gid, n)
    zero, _ = ref.group_constants(gid)
    bases[5] = zero  # an infinity among the bases: skipped, as are its multiples
    scalars = edge_scalars(ref, gid, n)
    want = ref.multi_exp(gid, bases, scalars)
    rb, rs = engine.ResidentBases(gid, bases), engine.ResidentScalars(gid, scalars)
    for c in (0, 5, 8, DIVIDING_C[gid]):
        if gid in (0, 1) and c == 2:
            rb2 = engine.ResidentBases(gid, bases[:64])  # 127 multiples per base
            t = engine.PrecomputedBases(gid, bases=rb2, c=c)
            got = engine.multi_exp_precomputed(t, scalars[:64])
            np.testing.assert_array_equal(affine(engine, gid, got),
                                          ref.multi_exp(gid, bases[:64], scalars[:64]))
            continue
        t = engine.PrecomputedBases(gid, bases=rb, c=c)
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_precomputed(t, rs)), want,
                                      err_msg=f"c={c} resident scalars")
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_precomputed(t, scalars)),
                                      want, err_msg=f"c={c} host scalars")
        t.release()


@pytest.mark.parametrize("gid", GROUPS)
@pytest.mark.parametrize("form", [0, 1])
def test_stream_binary_encoding_matches_reference_stream_msm(engine, ref, gid, form):
    """the bytes multi_exp_stream reads, pushed in ragged chunks and decoded on the device"""
    n = SMALL_N[gid] // 2
    bases = ref.gen_bases_walk(gid, n)
    zero, _ = ref.group_constants(gid)
    bases[3] = zero
    scalars = edge_scalars(ref, gid, n)
    data = ref.stream_write(gid, bases, form=form)
    pb = engine.wire_point_bytes(gid)
    assert len(data) == n * pb
    want = ref.multi_exp(gid, bases, scalars)
    streamed = ref.multi_exp_stream(gid, data, scalars, form=form)
    if streamed is not None:  # None: the reference itself is undefined here (see oracle/ref.py)
        np.testing.assert_array_equal(streamed, want)
    enc = engine.STREAM_BINARY_MONTGOMERY if form else engine.STREAM_BINARY_PLAIN
    st = engine.MsmStream(gid, scalars, encoding=enc)
    cuts = [0, 0, 1, 2, 50, n]
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        st.push(data[lo * pb:hi * pb], hi - lo)
    np.testing.assert_array_equal(affine(engine, gid, st.finish()), want)


@pytest.mark.parametrize("gid", GROUPS)
def test_stream_with_precompute_matches_reference(engine, ref, gid):
    """the precompute stream (each base followed by its multiples, multiexp_stream.hpp:30-36)"""
    n = 40 if gid in (3, 5) else 64
    bases = ref.gen_bases_walk(gid, n)
    scalars = edge_scalars(ref, gid, n)
    plain = ref.multi_exp(gid, bases, scalars)
    pb = engine.wire_point_bytes(gid)
    for c in (ref.bdlo12_signed_optimal_c(n), 7, DIVIDING_C[gid]):
        if c == 2:
            continue
        per = engine.precompute_digits(gid, c)
        data = ref.stream_write(gid, bases, form=1, precompute_c=c)
        assert len(data) == n * per * pb
        if c != DIVIDING_C[gid]:  # otherwise the reference loses the top carry of large scalars
            streamed = ref.multi_exp_stream(gid, data, scalars, form=1, precompute_c=c)
            assert streamed is not None
            np.testing.assert_array_equal(streamed, plain)
        st = engine.MsmStream(gid, scalars, encoding=engine.STREAM_BINARY_MONTGOMERY, precompute_c=c)
        cuts = [0, 1, 17, n]
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            st.push(data[lo * per * pb:hi * per * pb], hi - lo)
        np.testing.assert_array_equal(affine(engine, gid, st.finish()), plain, err_msg=f"c={c}")
        # reference_carry: the reference's own recoding with the top carry dropped -- identical to
        # what multi_exp_stream_with_precompute returns for EVERY c, including the dividing one
        # where that differs from the true sum (field_utils.tcc:205-239)
        streamed = ref.multi_exp_stream(gid, data, scalars, form=1, precompute_c=c)
        if streamed is not None:
            st = engine.MsmStream(gid, scalars, encoding=engine.STREAM_BINARY_MONTGOMERY,
                                  precompute_c=c, reference_carry=True)
            for lo, hi in zip(cuts[:-1], cuts[1:]):
                st.push(data[lo * per * pb:hi * per * pb], hi - lo)
            np.testing.assert_array_equal(affine(engine, gid, st.finish()), streamed,
                                          err_msg=f"c={c} reference carry")
            if c == DIVIDING_C[gid]:
                assert not np.array_equal(streamed, plain), "edge scalars should lose their top carry"


@pytest.mark.parametrize("gid", GROUPS)
def test_stream_in_memory_elements_and_uploaded_table(engine, ref, gid):
    n = 48
    aff = ref.gen_bases_walk(gid, n + 1)
    jac = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])  # Z != 1
    scalars = ref.gen_scalars(gid, n, start=77)
    for enc, pts in ((engine.STREAM_MEMORY_SPECIAL, aff[:n]), (engine.STREAM_MEMORY_NORMAL, jac)):
        st = engine.MsmStream(gid, scalars, encoding=enc)
        st.push(pts[:7])
        st.push(pts[7:])
        np.testing.assert_array_equal(affine(engine, gid, st.finish()),
                                      ref.multi_exp(gid, pts, scalars, form=ref.FORM_NORMAL))
    # table in the reference's stream order, as in-memory elements: P, 2^c P, 2^2c P, ...
    c = 9
    per = engine.precompute_digits(gid, c)
    rows = [jac]
    for _ in range(per - 1):
        cur = rows[-1]
        for _ in range(c):
            cur = ref.group_op(gid, ref.G_DBL, cur)
        rows.append(cur)
    table = np.stack(rows, axis=1).reshape(n * per, -1)  # base-major
    want = ref.multi_exp(gid, jac, scalars, form=ref.FORM_NORMAL)
    t = engine.PrecomputedBases(gid, c=c, elements=table, base_form=engine.FORM_NORMAL)
    np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_precomputed(t, scalars)), want)
    st = engine.MsmStream(gid, scalars, encoding=engine.STREAM_MEMORY_NORMAL, precompute_c=c)
    st.push(table[:5 * per])
    st.push(table[5 * per:])
    np.testing.assert_array_equal(affine(engine, gid, st.finish()), want)


def test_stream_many_chunks_fold_partials(engine, ref):
    """more pushes than partial-result slots: the partials are folded on the device"""
    gid, n = 0, 300
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=5)
    st = engine.MsmStream(gid, scalars)
    for i in range(n):
        st.push(bases[i:i + 1])
    np.testing.assert_array_equal(affine(engine, gid, st.finish()), ref.multi_exp(gid, bases, scalars))


def test_stream_empty_and_errors(engine, ref):
    gid = 0
    zero, _ = ref.group_constants(gid)
    st = engine.MsmStream(gid, np.zeros((0, 4), dtype=np.uint64))
    np.testing.assert_array_equal(st.finish(), zero)
    bases = ref.gen_bases_walk(gid, 4)
    scalars = ref.gen_scalars(gid, 3)
    st = engine.MsmStream(gid, scalars)
    with pytest.raises(engine.B200MsmError, match="holds 3 scalars"):
        st.push(bases)
    st.push(bases[:2])
    with pytest.raises(engine.B200MsmError, match="2 of 3"):
        st.finish()
    assert not engine.load().b200msm_stream_push(12345, None, 0, 0)
    assert b"unknown stream" in engine.load().b200msm_last_error()
    with pytest.raises(engine.B200MsmError, match="digit size"):
        engine.MsmStream(gid, scalars, precompute_c=30)
    rb = engine.ResidentBases(gid, bases)
    t = engine.PrecomputedBases(gid, bases=rb, c=6)
    with pytest.raises(engine.B200MsmError, match="need 4 scalars"):
        engine.multi_exp_precomputed(t, scalars)


def test_precomputed_large_window_and_timing(engine, ref):
    """2^16 points, engine-chosen digit size (sliced single bucket set), against the CPU reference"""
    gid, n = 0, 1 << 16
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n)
    want = ref.multi_exp(gid, bases, scalars, chunks=8)
    rb = engine.ResidentBases(gid, bases)
    for c in (0, 13, 16):
        t = engine.PrecomputedBases(gid, bases=rb, c=c)
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_precomputed(t, scalars)),
                                      want, err_msg=f"c={c}")
        t.release()
