"""Parity at the sizes of BASELINE.json's configurations ("GPU == CPU BDLO12, limb
for limb, for every timed configuration up to 2^24, and == closed form for all n including 2^26").

The small-n tests never reach the code paths a prover-scale MSM runs: the automatic digit size
(c = 15..20), 128-entry accumulation segments, the multi-pass tile scan (more than 1024 tiles),
window groups whose sort overlaps the previous group's accumulation, the 12-range shared-bucket
upload pipeline of host-buffer calls and position arithmetic near 2^32.  Here the CUDA path runs
with DEFAULT options on

  A  alt_bn128 G1 2^16           C  bls12_377 G1 2^20 (2^22 is the timed size; same code path)
  B  alt_bn128 G1 2^20           D  alt_bn128 G2 2^20
                                 E  bw6_761 G1 2^19 (the per-GPU share of 2^22 over 8 GPUs)

against the unmodified reference's multi_exp<BDLO12, special>(chunks = host threads) on the same
SHA512_rng scalars and arithmetic-walk bases, limb for limb after normalisation -- resident inputs,
pageable host buffers (plain and packed upload) and the ranged upload pipeline -- and at 2^22 / 2^24 / 2^26
against the closed form (sum_i s_i (a + i d)) G, computed with numpy / Python integers and ONE
reference scalar multiplication, independent of any MSM code.
"""
import os

import numpy as np
import pytest

from oracle import pymodel as pm

pytestmark = pytest.mark.gpu

CONFIGS = [(0, 16), (0, 20), (2, 20), (1, 20), (3, 19)]
IDS = ["A-alt_bn128_g1-2p16", "B-alt_bn128_g1-2p20", "C-bls12_377_g1-2p20", "D-alt_bn128_g2-2p20",
       "E-bw6_761_g1-2p19"]


def affine(engine, gid, pt):
    return engine.batch_to_special(gid, np.asarray(pt, dtype=np.uint64).reshape(1, -1))[0]


def threads():
    return max(1, len(os.sched_getaffinity(0)))


@pytest.fixture(scope="module")
def workloads(ref):
    cache = {}

    def get(gid, log2n):
        key = (gid, log2n)
        if key not in cache:
            n = 1 << log2n
            ref.lib().ref_set_num_threads(threads())
            bases = ref.gen_bases_walk(gid, n)
            scalars = ref.gen_scalars(gid, n)
            want = ref.multi_exp(gid, bases, scalars, method=ref.BDLO12, form=ref.FORM_SPECIAL,
                                 chunks=threads())
            cache.clear()  # one workload at a time: they are up to hundreds of MB
            cache[key] = (bases, scalars, want)
        return cache[key]

    return get


@pytest.mark.parametrize("gid,log2n", CONFIGS, ids=IDS)
def test_config_size_matches_reference_multi_exp(engine, ref, workloads, gid, log2n):
    """every host-side path of the drop-in at the configuration's size, default options"""
    bases, scalars, want = workloads(gid, log2n)
    n = 1 << log2n
    g = pm.GROUPS[gid]
    # the reference's own result equals the closed form (pins the workload generator too)
    k = pm.walk_expected_scalar_fast(g, scalars)
    np.testing.assert_array_equal(want, ref.scalar_mul(gid, ref.group_constants(gid)[1],
                                                       g.encode_scalars([k])[0]))
    # 1. resident inputs (the timed `value` path of bench.py)
    rb, rs = engine.ResidentBases(gid, bases), engine.ResidentScalars(gid, scalars)
    np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident(rb, rs)), want,
                                  err_msg="resident")
    # a sub-range that starts and ends off every alignment
    lo, cnt = 12345, n - 2 * 12345 - 1
    part = ref.multi_exp(gid, bases[lo:lo + cnt], scalars[lo:lo + cnt], chunks=threads())
    np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident(rb, rs, lo, cnt)),
                                  part, err_msg="resident sub-range")
    # 2. host buffers, default options (pageable numpy memory: staged, full elements travel)
    np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp(gid, bases, scalars)), want,
                                  err_msg="host buffers")
    # 3. the ranged upload pipeline (what n >= 2^23 takes by default): 12 ranges, one bucket set
    engine.set_option("h2d_parts", 12)
    try:
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp(gid, bases, scalars)),
                                      want, err_msg="12 upload ranges, shared buckets")
        engine.set_option("h2d_pack", 1)  # special-form bases packed to X|Y on the host
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp(gid, bases, scalars)),
                                      want, err_msg="12 upload ranges, packed")
        engine.set_option("h2d_pack", 0)
        # 4. the prover's shape: resident bases, host scalars in ranges
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident_bases(rb, scalars)),
                                      want, err_msg="resident bases, host scalars")
        engine.set_option("h2d_shared_buckets", 0)
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp(gid, bases, scalars)),
                                      want, err_msg="12 independent ranges")
    finally:
        engine.set_option("h2d_parts", 0)
        engine.set_option("h2d_pack", 0)
        engine.set_option("h2d_shared_buckets", 1)
    # 5. one sort for all windows against the overlapped window groups
    engine.set_option("sort_overlap", 1)
    try:
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident(rb, rs)), want,
                                      err_msg="single sort")
        engine.set_option("sort_overlap", 5)
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident(rb, rs)), want,
                                      err_msg="five window groups")
    finally:
        engine.set_option("sort_overlap", 0)
    rb.release()
    rs.release()


def engine_walk_bases(engine, ref, gid, n, batch=1 << 22):
    """Arithmetic-walk bases P_i = (a + i d) G for sizes the CPU generator would take minutes for:
    fixed-base exponentiation of the generator on the device, then checked against the reference's
    own generator at a prefix and at scattered positions (so the inputs are the reference's).
    a + i d = d (a / d + i): the rows handed to batch_exp_with_coeff are the plain integers
    t0 + j (read as Montgomery forms they stand for (t0 + j) / R), the coefficient is d R."""
    g = pm.GROUPS[gid]
    a, d = pm.walk_params(g)
    r = g.fr.p
    _, one = ref.group_constants(gid)
    bits = g.fr.bits
    limbs = g.fr.limbs64
    R = 1 << (64 * limbs)
    table = engine.WindowTable(gid, bits, 8, g=one)
    coeff = g.encode_scalars([d * R % r])[0]
    out = np.zeros((n, ref.group_words(gid)), dtype=np.uint64)
    for lo in range(0, n, batch):
        cnt = min(batch, n - lo)
        t0 = (a * pow(d, -1, r) + lo) % r
        assert t0 + cnt < r, "walk wraps inside a batch"
        rows = np.zeros((cnt, limbs), dtype=np.uint64)
        j = np.arange(cnt, dtype=np.uint64)
        low = np.uint64(t0 & 0xFFFFFFFFFFFFFFFF)
        rows[:, 0] = low + j
        carry = (rows[:, 0] < low).astype(np.uint64)
        for q in range(1, limbs):
            limb = np.uint64((t0 >> (64 * q)) & 0xFFFFFFFFFFFFFFFF)
            rows[:, q] = limb + carry
            carry = ((rows[:, q] < limb) & (carry > 0)).astype(np.uint64)
        pts = engine.batch_exp_with_coeff(gid, bits, 8, table, coeff, rows)
        out[lo:lo + cnt] = engine.batch_to_special(gid, pts)
    table.release()
    np.testing.assert_array_equal(out[:4096], ref.gen_bases_walk(gid, 4096))
    rng = np.random.RandomState(7)
    for pos in list(rng.randint(0, n, size=24)) + [n - 1]:
        np.testing.assert_array_equal(out[pos], ref.gen_bases_walk(gid, 1, start=int(pos))[0])
    return out


@pytest.mark.parametrize("log2n", [22, 24, 26])
def test_closed_form_at_scale_alt_bn128_g1(engine, ref, log2n):
    """2^22 / 2^24 / 2^26 points: resident, host-buffer (12 upload ranges sharing one bucket set)
    and resident-bases calls against (sum s_i (a + i d)) G."""
    if log2n == 26 and os.environ.get("B200MSM_SKIP_2P26"):
        pytest.skip("B200MSM_SKIP_2P26 set")
    gid = 0
    n = 1 << log2n
    g = pm.GROUPS[gid]
    ref.lib().ref_set_num_threads(threads())
    scalars = ref.gen_scalars(gid, n)
    if log2n <= 22:
        bases = ref.gen_bases_walk(gid, n)
    else:
        bases = engine_walk_bases(engine, ref, gid, n)
    k = pm.walk_expected_scalar_fast(g, scalars)
    want = ref.scalar_mul(gid, ref.group_constants(gid)[1], g.encode_scalars([k])[0])
    rb = engine.ResidentBases(gid, bases)
    rs = engine.ResidentScalars(gid, scalars)
    np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident(rb, rs)), want,
                                  err_msg="resident")
    rs.release()
    np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp_resident_bases(rb, scalars)),
                                  want, err_msg="resident bases, host scalars")
    rb.release()
    if log2n <= 24:
        np.testing.assert_array_equal(affine(engine, gid, engine.multi_exp(gid, bases, scalars)), want,
                                      err_msg="host buffers")
    if log2n == 22:  # the CPU reference agrees at this size too (a few seconds)
        np.testing.assert_array_equal(ref.multi_exp(gid, bases, scalars, chunks=threads()), want)


def test_batch_to_special_and_batch_exp_at_2p16(engine, ref):
    """the fixed-base path at n = 2^16 with the reference's own window choice
    (get_exp_window_size, multiexp.tcc:770-807), table built by the reference and on the device,
    with and without a coefficient (batch_exp_with_coeff, multiexp.tcc:917-947)"""
    n = 1 << 16
    for gid in (0, 1, 2):
        g = pm.GROUPS[gid]
        bits = g.fr.bits
        window = int(ref.lib().ref_exp_window_size(gid, n))
        if gid != 0:
            window = min(window, 10)  # keeps the CPU-built table of the wide groups small
        _, one = ref.group_constants(gid)
        table = ref.window_table(gid, bits, window, one)
        m = n if gid == 0 else 1 << 12
        scalars = ref.gen_scalars(gid, m, start=9)
        secs, want = ref.batch_exp(gid, bits, window, table, scalars, normalize=False)
        want_aff = ref.batch_to_special(gid, want)
        got = engine.batch_exp(gid, bits, window, list(table), scalars)
        # batch_to_special on the device against the reference's, limb for limb, at this size
        np.testing.assert_array_equal(engine.batch_to_special(gid, got), want_aff)
        np.testing.assert_array_equal(engine.batch_to_special(gid, want), want_aff)
        # resident table uploaded once, and the table built on the device from g
        up = engine.WindowTable(gid, bits, window, table=list(table))
        dev = engine.WindowTable(gid, bits, window, g=one)
        np.testing.assert_array_equal(dev.download(), ref.batch_to_special(
            gid, table.reshape(-1, table.shape[-1])).reshape(table.shape))
        for t in (up, dev):
            np.testing.assert_array_equal(
                engine.batch_to_special(gid, engine.batch_exp(gid, bits, window, t, scalars)), want_aff)
        # coefficient: (coeff * v_i) g == batch_exp of the products the reference computes
        coeff = ref.gen_scalars(gid, 1, start=4242)
        prod = ref.field_op(ref.FR_OF_GROUP[gid], ref.OP_MUL, np.repeat(coeff, m, axis=0), scalars)
        _, want_c = ref.batch_exp(gid, bits, window, table, prod, normalize=True)
        for t in (list(table), dev):
            got_c = engine.batch_exp_with_coeff(gid, bits, window, t, coeff[0], scalars)
            np.testing.assert_array_equal(engine.batch_to_special(gid, got_c), want_c)
        up.release()
        dev.release()
