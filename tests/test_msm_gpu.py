"""GPU parity tests: the CUDA path (through the C ABI) against the unmodified reference
(oracle/_ref) and the pure-Python model, on the same inputs.  Bit-exact: every comparison is on
affine-normalised limbs."""
import numpy as np
import pytest

from oracle import pymodel as pm

pytestmark = pytest.mark.gpu

GROUPS = [0, 1, 2, 3, 4, 5, 6, 7]
SMALL_N = {0: 1000, 1: 300, 2: 500, 3: 200, 4: 200, 5: 200, 6: 500, 7: 200}


def affine(ref, gid, pts):
    pts = np.ascontiguousarray(pts, dtype=np.uint64).reshape(-1, ref.group_words(gid))
    return ref.batch_to_special(gid, pts)


def assert_same_point(ref, gid, got, want):
    got = np.asarray(got, dtype=np.uint64).reshape(1, -1)
    want = np.asarray(want, dtype=np.uint64).reshape(1, -1)
    assert ref.group_eq(gid, got, want)[0], "operator== says the points differ"
    np.testing.assert_array_equal(affine(ref, gid, got), affine(ref, gid, want))


@pytest.mark.parametrize("gid", GROUPS)
@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 255, 256, 257])
def test_multi_exp_small_sizes(engine, ref, gid, n):
    """Sizes of the reference's own test (test_multiexp.cpp:258-326) plus n = 0."""
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n)
    want = ref.multi_exp(gid, bases, scalars, method=ref.BDLO12, form=ref.FORM_SPECIAL)
    got = engine.multi_exp(gid, bases, scalars, base_form=engine.FORM_SPECIAL)
    assert_same_point(ref, gid, got, want)


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_reference_pattern(engine, ref, gid):
    """test_multi_exp_config (test_multiexp.cpp:205-256): bases (i+1) G, scalars n - i."""
    g = pm.GROUPS[gid]
    n = 64
    _, one = ref.group_constants(gid)
    bases = np.stack([ref.scalar_mul(gid, one, g.encode_scalars([i + 1])[0]) for i in range(n)])
    scalars = g.encode_scalars([n - i for i in range(n)])
    k = sum((i + 1) * (n - i) for i in range(n))
    want = ref.scalar_mul(gid, one, g.encode_scalars([k])[0])
    for form in (engine.FORM_SPECIAL, engine.FORM_NORMAL):
        got = engine.multi_exp(gid, bases, scalars, base_form=form)
        assert_same_point(ref, gid, got, want)


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_random(engine, ref, gid):
    n = SMALL_N[gid]
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=77)
    want = ref.multi_exp(gid, bases, scalars)
    got = engine.multi_exp(gid, bases, scalars)
    assert_same_point(ref, gid, got, want)
    # closed form, independent of any MSM code
    g = pm.GROUPS[gid]
    k = pm.walk_expected_scalar(g, g.decode_scalars(scalars))
    np.testing.assert_array_equal(affine(ref, gid, got)[0], g.encode_affine(g.mul(k, g.gen)))


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_jacobian_bases(engine, ref, gid):
    """base_form_normal: bases with arbitrary Z (results of group additions)."""
    n = 100
    aff = ref.gen_bases_walk(gid, n + 1)
    bases = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])  # non-normalised sums
    scalars = ref.gen_scalars(gid, n, start=5)
    want = ref.multi_exp(gid, bases, scalars, form=ref.FORM_NORMAL)
    got = engine.multi_exp(gid, bases, scalars, base_form=engine.FORM_NORMAL)
    assert_same_point(ref, gid, got, want)


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_edge_cases(engine, ref, gid):
    """Exceptional inputs (SURVEY.md §8d edge-case suite): zero/one/r-1 scalars, infinity among the
    bases, duplicated points (in-bucket doubling), P and -P with the same scalar (cancellation),
    the reference profiler's repeated-points distribution (profile_multiexp.cpp:15,31-46)."""
    g = pm.GROUPS[gid]
    r = g.fr.p
    zero, _ = ref.group_constants(gid)
    P = ref.gen_bases_walk(gid, 8)
    negP = ref.group_op(gid, ref.G_NEG, P)
    enc = g.encode_scalars
    cases = {
        "all_zero_scalars": (P, enc([0] * 8)),
        "ones": (P, enc([1] * 8)),
        "r_minus_1": (P, enc([r - 1] * 8)),
        "powers": (P, enc([1 << 1, 1 << 31, 1 << 32, 1 << 63, 1 << 64, (1 << 100) - 1, 2, 3])),
        "cancel": (np.stack([P[0], negP[0]]), enc([12345, 12345])),
        "cancel_big": (np.stack([P[0], negP[0], P[1]]), enc([r - 2, r - 2, 7])),
        "dups": (np.stack([P[0]] * 4), enc([99, 99, 99, 99])),
        "dups_mixed": (np.stack([P[0], P[0], negP[0], P[1], P[1]]), enc([5, 5, 5, r - 1, r - 1])),
        "infinity_bases": (np.stack([zero, P[0], zero, P[1]]), enc([3, 4, 5, 6])),
        "all_infinity": (np.stack([zero] * 3), enc([3, 4, 5])),
        "single": (P[:1], enc([r - 1])),
    }
    rep = np.concatenate([P] * 40)[:300]  # few distinct points, many repeats
    cases["repeated_points"] = (rep, ref.gen_scalars(gid, len(rep), start=1000))
    cases["repeated_same_scalar"] = (rep, np.repeat(ref.gen_scalars(gid, 1, start=3), len(rep), 0))
    for name, (bases, scalars) in cases.items():
        want = ref.multi_exp(gid, bases, scalars, method=ref.NAIVE_PLAIN)
        want2 = ref.multi_exp(gid, bases, scalars, method=ref.BDLO12)
        got = engine.multi_exp(gid, bases, scalars)
        try:
            assert_same_point(ref, gid, want2, want)
            assert_same_point(ref, gid, got, want)
        except AssertionError as e:
            raise AssertionError(f"case {name}: {e}")


def test_edge_case1_vectors(engine, ref):
    """TestEdgeCase1 (test_multiexp.cpp:344-390): the reference's literal vectors -- top signed
    digit negative and absorbing a carry.  Expected value from Python integers."""
    g = pm.ALT_BN128_G1
    P0 = (0x157cbb5c3e5ff87af54877aefd2b3f95393ee3004a91c59c5c5023ca85aee7d0,
          0x162bb95fdd346c7276bb20ad526863d63c0eca0776fb57c71b76bd02dab6c554)
    P1 = (0x277a1d218ecc7a3ea26cec72b46c4f13b8b6594fcc080fa6d8ae61d43c34e084,
          0x13c48130e8519a34162bdc018409e98ca81683c23556314401a3acad24b1b9bc)
    assert g.is_on_curve(P0) and g.is_on_curve(P1)
    bases = g.encode_points([P0, P1])
    scalars = g.encode_scalars([1, 0x79D69])
    want = g.encode_affine(g.add(P0, g.mul(0x79D69, P1)))
    for form in (engine.FORM_NORMAL, engine.FORM_SPECIAL):
        got = engine.multi_exp(0, bases, scalars, base_form=form)
        np.testing.assert_array_equal(affine(ref, 0, got)[0], want)
    assert_same_point(ref, 0, got, ref.multi_exp(0, bases, scalars, method=ref.BDLO12_SIGNED,
                                                 form=ref.FORM_NORMAL))
    assert_same_point(ref, 0, got, ref.multi_exp(0, bases, scalars, method=ref.NAIVE,
                                                 form=ref.FORM_NORMAL, chunks=2))


@pytest.mark.parametrize("gid", GROUPS)
@pytest.mark.parametrize("cbits", [2, 3, 7, 11, 16])
def test_multi_exp_window_sizes(engine, ref, gid, cbits):
    """Any window size gives the same group element."""
    n = 200 if gid != 3 else 80
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=31)
    want = ref.multi_exp(gid, bases, scalars)
    engine.set_option("window_bits", cbits)
    try:
        got = engine.multi_exp(gid, bases, scalars)
    finally:
        engine.set_option("window_bits", 0)
    assert_same_point(ref, gid, got, want)


@pytest.mark.parametrize("gid", GROUPS)
def test_batch_to_special(engine, ref, gid):
    n = 700 if gid != 3 else 200
    aff = ref.gen_bases_walk(gid, n + 1)
    pts = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])
    zero, _ = ref.group_constants(gid)
    pts[3] = zero
    pts[n - 1] = zero
    pts[10] = aff[10]  # already special
    want = ref.batch_to_special(gid, pts)
    got = engine.batch_to_special(gid, pts)
    np.testing.assert_array_equal(got, want)
    for k in (1, 2, 16, 17):
        np.testing.assert_array_equal(engine.batch_to_special(gid, pts[:k]), want[:k])


@pytest.mark.parametrize("gid", GROUPS)
def test_group_sum(engine, ref, gid):
    aff = ref.gen_bases_walk(gid, 9)
    pts = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])
    zero, _ = ref.group_constants(gid)
    pts[2] = zero
    acc = zero.reshape(1, -1)
    for p in pts:
        acc = ref.group_op(gid, ref.G_ADD, acc, p.reshape(1, -1))
    assert_same_point(ref, gid, engine.group_sum(gid, pts), acc[0])
    # P + P and P + (-P)
    two = np.stack([pts[0], pts[0]])
    assert_same_point(ref, gid, engine.group_sum(gid, two),
                      ref.group_op(gid, ref.G_DBL, pts[:1])[0])
    pm_ = np.stack([pts[0], ref.group_op(gid, ref.G_NEG, pts[:1])[0]])
    assert_same_point(ref, gid, engine.group_sum(gid, pm_), zero)


@pytest.mark.parametrize("gid", GROUPS)
def test_batch_exp(engine, ref, gid):
    g = pm.GROUPS[gid]
    bits = g.fr.bits
    window = 5
    _, one = ref.group_constants(gid)
    table = ref.window_table(gid, bits, window, one)
    n = 120 if gid != 3 else 40
    scalars = ref.gen_scalars(gid, n, start=9)
    scalars[0] = g.encode_scalars([0])[0]
    scalars[1] = g.encode_scalars([g.fr.p - 1])[0]
    _, want = ref.batch_exp(gid, bits, window, table, scalars, normalize=True)
    got = engine.batch_exp(gid, bits, window, list(table), scalars)
    np.testing.assert_array_equal(affine(ref, gid, got), want)


def test_multi_exp_resident_and_shards(engine, ref):
    """Resident inputs and the point-range sharding used across GPUs (multiexp.tcc:663-687)."""
    gid, n = 0, 3000
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n)
    want = ref.multi_exp(gid, bases, scalars)
    rb = engine.ResidentBases(gid, bases)
    rs = engine.ResidentScalars(gid, scalars)
    assert_same_point(ref, gid, engine.multi_exp_resident(rb, rs), want)
    for world in (2, 3, 8):
        parts = []
        for rank in range(world):
            lo, hi = engine.shard_range(n, rank, world)
            parts.append(engine.multi_exp_resident(rb, rs, lo, hi - lo))
        assert_same_point(ref, gid, engine.group_sum(gid, np.stack(parts)), want)
    t = engine.last_timings()
    assert t["total"] > 0 and engine.launch_count() > 0
    rb.release()
    rs.release()


def test_errors(engine):
    b = np.zeros((4, 12), dtype=np.uint64)
    s = np.zeros((3, 4), dtype=np.uint64)
    with pytest.raises(engine.B200MsmError):
        engine.multi_exp(0, b, s)
    with pytest.raises(engine.B200MsmError):
        engine.multi_exp(9, b, s)


@pytest.mark.parametrize("seg", [3, 4, 12, 64, 256])
def test_multi_exp_segment_lengths(engine, ref, seg):
    """Accumulation segment length (items per thread) is a tuning knob: any value gives the same
    element, including segments that span many small buckets and buckets that span many segments
    (few distinct scalars -> long runs and multi-round carries)."""
    gid = 0
    n = 3000
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=11)
    heavy = np.repeat(ref.gen_scalars(gid, 3, start=5), n // 3, axis=0)  # three huge buckets/window
    engine.set_option("segment", seg)
    try:
        for cbits in (0, 4, 12):
            engine.set_option("window_bits", cbits)
            for sc in (scalars, heavy):
                want = ref.multi_exp(gid, bases, sc)
                assert_same_point(ref, gid, engine.multi_exp(gid, bases, sc), want)
    finally:
        engine.set_option("segment", 0)
        engine.set_option("window_bits", 0)


@pytest.mark.parametrize("gid", [0, 1, 2, 3])
def test_window_totals_over_a_cluster(engine, ref, gid):
    """The per-window totals run on one CTA or spread over a thread-block cluster of 2 / 4 / 8 CTAs
    that combine through distributed shared memory (option reduce_cluster; 0 = automatic): same
    element for every setting, for digit sizes that give fewer chunks than threads and more."""
    n = 5000 if gid != 3 else 900
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=21)
    want = ref.multi_exp(gid, bases, scalars)
    try:
        for cbits in (0, 7, 14, 16):
            engine.set_option("window_bits", cbits)
            for cl in (1, 2, 4, 8, 0):
                engine.set_option("reduce_cluster", cl)
                assert_same_point(ref, gid, engine.multi_exp(gid, bases, scalars), want)
    finally:
        engine.set_option("reduce_cluster", 0)
        engine.set_option("window_bits", 0)


@pytest.mark.parametrize("gid", GROUPS)
@pytest.mark.parametrize("shared", [1, 0])
@pytest.mark.parametrize("parts", [1, 2, 3, 7])
def test_multi_exp_pipelined_upload(engine, ref, gid, parts, shared):
    """Host-buffer calls cut the point range into `h2d_parts` independently uploaded ranges (upload
    of range k + 1 under the kernels of range k).  shared = 1 (default): all ranges are sorted with
    the digit size of the whole call and accumulate into ONE bucket set that is reduced once;
    shared = 0: every range is an MSM of its own and the partial results are added
    (multiexp.tcc:663-687 does the same split across threads).  Same element either way."""
    n = {0: 1003, 1: 301, 2: 502, 3: 99, 4: 99, 5: 99, 6: 502, 7: 99}[gid]
    aff = ref.gen_bases_walk(gid, n + 1)
    jac = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])
    scalars = ref.gen_scalars(gid, n, start=21)
    engine.set_option("h2d_parts", parts)
    engine.set_option("h2d_shared_buckets", shared)
    try:
        assert_same_point(ref, gid, engine.multi_exp(gid, aff[:n], scalars), ref.multi_exp(gid, aff[:n], scalars))
        assert_same_point(ref, gid, engine.multi_exp(gid, jac, scalars, base_form=engine.FORM_NORMAL),
                          ref.multi_exp(gid, jac, scalars, form=ref.FORM_NORMAL))
        assert_same_point(ref, gid, engine.multi_exp(gid, aff[:2], scalars[:2]), ref.multi_exp(gid, aff[:2], scalars[:2]))
    finally:
        engine.set_option("h2d_parts", 0)
        engine.set_option("h2d_shared_buckets", 1)


@pytest.mark.parametrize("gid", GROUPS)
@pytest.mark.parametrize("window_bits", [0, 3, 9])
def test_multi_exp_shared_bucket_ranges_edge_cases(engine, ref, gid, window_bits):
    """The shared bucket set across upload ranges: buckets that an earlier range filled and a later
    one extends, ranges whose scalars are all zero (no items at all, first / middle / last range),
    a later range that only uses low windows after an earlier one used all of them, repeated
    bases and P + (-P) meeting in one bucket across two ranges, infinity among the bases."""
    g = pm.GROUPS[gid]
    n = {0: 600, 1: 240, 2: 360, 3: 96, 4: 96, 5: 96, 6: 360, 7: 96}[gid]
    parts = 4
    q = n // parts
    bases = ref.gen_bases_walk(gid, n)
    full = g.decode_scalars(ref.gen_scalars(gid, n, start=77))
    r = g.fr.p
    zero, _ = ref.group_constants(gid)
    cases = []
    for zero_part in (0, 2, 3):            # one range contributes nothing
        cases.append((bases, [0 if i // q == zero_part else v for i, v in enumerate(full)]))
    # high windows only in the first range, tiny scalars afterwards
    cases.append((bases, [v if i < q else (v % 7) for i, v in enumerate(full)]))
    # the same base with the same digit in every range; then cancelling pairs across ranges
    same = np.repeat(bases[:1], n, axis=0)
    cases.append((same, [5] * n))
    cases.append((same, [(5 if (i // q) % 2 == 0 else r - 5) for i in range(n)]))
    # infinity bases (encoded as the reference does) in every range
    with_inf = bases.copy()
    with_inf[::q] = zero
    cases.append((with_inf, full))
    engine.set_option("h2d_parts", parts)
    engine.set_option("window_bits", window_bits)
    try:
        for b, vals in cases:
            sc = g.encode_scalars(vals)
            assert_same_point(ref, gid, engine.multi_exp(gid, b, sc), ref.multi_exp(gid, b, sc))
    finally:
        engine.set_option("h2d_parts", 0)
        engine.set_option("window_bits", 0)


@pytest.mark.parametrize("gid", GROUPS)
def test_multi_exp_groth16_shaped_scalars(engine, ref, gid):
    """Witness-like scalars (what multi_exp_filter_one_zero, multiexp.tcc:690-757, is written for):
    mostly 0 and 1, some short, few full-width.  Zero scalars give no bucket items, unit scalars
    pile into one bucket of the lowest window, and the windows above the longest scalar are
    dropped from the reduction (the reference sizes its loop by num_bits, multiexp.tcc:298-306)."""
    g = pm.GROUPS[gid]
    n = {0: 2000, 1: 400, 2: 800, 3: 150, 4: 150, 5: 150, 6: 800, 7: 150}[gid]
    rng = np.random.RandomState(5 + gid)
    kinds = rng.choice(4, size=n, p=[0.55, 0.3, 0.1, 0.05])
    full = g.decode_scalars(ref.gen_scalars(gid, n, start=300))
    vals = [0 if k == 0 else 1 if k == 1 else int(rng.randint(2, 1 << 16)) if k == 2 else full[i]
            for i, k in enumerate(kinds)]
    bases = ref.gen_bases_walk(gid, n)
    for sc in (vals, [v if k != 3 else 0 for v, k in zip(vals, kinds)], [1] * n, [0] * (n - 1) + [3]):
        scalars = g.encode_scalars(sc)
        want = ref.multi_exp(gid, bases, scalars)
        assert_same_point(ref, gid, engine.multi_exp(gid, bases, scalars), want)
