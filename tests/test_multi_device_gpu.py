"""Several GPUs inside ONE process, through the C ABI (b200msm_init(num_gpus)).

The reference splits the index range into `chunks` contiguous ranges, runs them on OpenMP threads
and adds the partial results (multiexp.tcc:655-687); the engine does the same with one engine
context and one host thread per bound device.  With fewer than the wanted number of GPUs on the box
a device is bound more than once -- independent contexts (streams, scratch, host threads) on one
GPU -- so the split, the concurrent contexts and the combination of the partials are exercised on a
single-GPU box too; with several GPUs visible the contexts sit on distinct devices.  Everything is
compared with the unmodified reference (oracle/_ref), limb for limb after normalisation.
"""
import os
import subprocess

import numpy as np
import pytest

from oracle import pymodel as pm

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def visible_gpus():
    import torch
    return torch.cuda.device_count()


@pytest.fixture(scope="module", params=[2, 3, 8])
def multi(request, engine):
    """the engine bound to `param` contexts: distinct GPUs where the box has them"""
    want = request.param
    have = visible_gpus()
    devices = [k % max(have, 1) for k in range(want)]
    engine.shutdown()
    assert engine.init_devices(devices) == want
    assert engine.device_count() == want
    yield engine, want
    engine.shutdown()
    engine.init(0)


def affine(ref, gid, pt):
    return ref.batch_to_special(gid, np.asarray(pt, dtype=np.uint64).reshape(1, -1))[0]


@pytest.mark.parametrize("gid", [0, 1, 2, 3])
def test_multi_exp_split_over_devices(multi, ref, gid):
    engine, D = multi
    for n in (0, 1, D - 1, D, D + 1, 257, 1000 if gid != 3 else 120):
        bases = ref.gen_bases_walk(gid, n)
        scalars = ref.gen_scalars(gid, n, start=40)
        want = ref.multi_exp(gid, bases, scalars)
        np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp(gid, bases, scalars)), want,
                                      err_msg=f"n={n}")
    # arbitrary Z
    n = 300 if gid != 3 else 60
    aff = ref.gen_bases_walk(gid, n + 1)
    jac = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])
    scalars = ref.gen_scalars(gid, n, start=3)
    np.testing.assert_array_equal(
        affine(ref, gid, engine.multi_exp(gid, jac, scalars, base_form=engine.FORM_NORMAL)),
        ref.multi_exp(gid, jac, scalars, form=ref.FORM_NORMAL))


def test_every_device_takes_its_range(multi, ref):
    """the reference's split: one = n / D per range, the last takes the remainder"""
    engine, D = multi
    gid, n = 0, (1 << 18) + 5
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n)
    want = ref.multi_exp(gid, bases, scalars, chunks=max(1, len(os.sched_getaffinity(0))))
    l0 = engine.launch_count()
    np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp(gid, bases, scalars)), want)
    assert engine.launch_count() > l0
    totals = [engine.device_timings(k).get("total", 0.0) for k in range(D)]
    assert all(t > 0 for t in totals), totals
    # resident inputs are cut the same way; sub-ranges cross the cuts
    rb, rs = engine.ResidentBases(gid, bases), engine.ResidentScalars(gid, scalars)
    np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp_resident(rb, rs)), want)
    one = n // D
    for lo, cnt in ((0, one), (one - 3, 7), (5, n - 9), (n - 1, 1), (one, 0)):
        np.testing.assert_array_equal(
            affine(ref, gid, engine.multi_exp_resident(rb, rs, lo, cnt)),
            ref.multi_exp(gid, bases[lo:lo + cnt], scalars[lo:lo + cnt], chunks=8), err_msg=f"{lo}+{cnt}")
    # the prover's shape
    np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp_resident_bases(rb, scalars)), want)
    # option num_gpus limits a call to the first k devices
    engine.set_option("num_gpus", 1)
    try:
        np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp(gid, bases, scalars)), want)
    finally:
        engine.set_option("num_gpus", 0)
    with pytest.raises(engine.B200MsmError, match="num_gpus"):
        engine.set_option("num_gpus", D + 1)
    rb.release()
    rs.release()


def test_sharded_result_equals_closed_form_at_2p22(multi, ref):
    """2^22 alt_bn128 G1 points cut over the bound devices (2, 3 -- an uneven split -- and 8):
    the combined point equals (sum s_i (a + i d)) G computed with numpy / Python integers and ONE
    reference scalar multiplication, for resident inputs, host buffers and resident bases."""
    engine, D = multi
    gid, n = 0, 1 << 22
    g = pm.GROUPS[gid]
    ref.lib().ref_set_num_threads(max(1, len(os.sched_getaffinity(0))))
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n)
    k = pm.walk_expected_scalar_fast(g, scalars)
    want = ref.scalar_mul(gid, ref.group_constants(gid)[1], g.encode_scalars([k])[0])
    rb, rs = engine.ResidentBases(gid, bases), engine.ResidentScalars(gid, scalars)
    np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp_resident(rb, rs)), want)
    assert all(engine.device_timings(d).get("total", 0.0) > 0 for d in range(D))
    np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp_resident_bases(rb, scalars)), want)
    np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp(gid, bases, scalars)), want)
    rb.release()
    rs.release()


@pytest.mark.parametrize("gid", [0, 3])
def test_precomputed_tables_are_sharded_too(multi, ref, gid):
    engine, D = multi
    n = 500 if gid == 0 else 90
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=8)
    want = ref.multi_exp(gid, bases, scalars)
    rb, rs = engine.ResidentBases(gid, bases), engine.ResidentScalars(gid, scalars)
    for c in (0, 6):
        t = engine.PrecomputedBases(gid, bases=rb, c=c)
        np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp_precomputed(t, rs)), want)
        np.testing.assert_array_equal(affine(ref, gid, engine.multi_exp_precomputed(t, scalars)), want)
        t.release()


@pytest.mark.parametrize("gid", [0, 1, 2, 3])
def test_batch_ops_shard_by_output_range(multi, ref, gid):
    """batch_exp / batch_to_special: independent outputs, no exchange (multiexp.tcc:898-908)"""
    engine, D = multi
    g = pm.GROUPS[gid]
    n = 333 if gid != 3 else 77
    aff = ref.gen_bases_walk(gid, n + 1)
    pts = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])
    zero, one = ref.group_constants(gid)
    pts[0] = zero
    pts[n - 1] = zero
    np.testing.assert_array_equal(engine.batch_to_special(gid, pts), ref.batch_to_special(gid, pts))
    bits, window = g.fr.bits, 5
    table = ref.window_table(gid, bits, window, one)
    scalars = ref.gen_scalars(gid, n, start=9)
    _, want = ref.batch_exp(gid, bits, window, table, scalars, normalize=True)
    np.testing.assert_array_equal(
        ref.batch_to_special(gid, engine.batch_exp(gid, bits, window, list(table), scalars)), want)
    dev = engine.WindowTable(gid, bits, window, g=one)
    np.testing.assert_array_equal(
        ref.batch_to_special(gid, engine.batch_exp(gid, bits, window, dev, scalars)), want)
    dev.release()


def test_cpp_dropin_on_several_devices(ref):
    """The reference's own test code (tests/cpp/dropin_test.cpp: libff headers + libff_b200.hpp)
    in one process over 2 and 8 engine contexts ($B200MSM_NUM_GPUS; more contexts than GPUs bind
    devices repeatedly through $B200MSM_DEVICES)."""
    exe = os.path.join(ROOT, "oracle", "_ref", "dropin_test")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/dropin_test not built")
    have = max(visible_gpus(), 1)
    for want in (2, 8):
        env = dict(os.environ)
        env["B200MSM_DEVICES"] = ",".join(str(k % have) for k in range(want))
        out = subprocess.run([exe], capture_output=True, text=True, timeout=900, env=env)
        assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
        assert f"engine contexts: {want}" in out.stdout, out.stdout[-2000:]
