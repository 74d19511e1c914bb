"""Wire-format MSM (b200msm_multi_exp_wire) on the GPU against the reference's own serializer
(ffi::group_element_read/write, ffi::field_element_read/write, ffi/ffi_serialization.tcc:98-187,
called through oracle/_ref) and its CPU multi_exp.  Byte-exact."""
import numpy as np
import pytest

from oracle import pymodel as pm

pytestmark = pytest.mark.gpu
GROUPS = [0, 1, 2, 3, 4, 5, 6, 7]


@pytest.mark.parametrize("gid", GROUPS)
@pytest.mark.parametrize("n", [0, 1, 2, 257])
def test_wire_msm_matches_reference(engine, ref, gid, n):
    if gid == 3 and n > 100:
        n = 90
    bases = ref.gen_bases_walk(gid, n)
    scalars = ref.gen_scalars(gid, n, start=17)
    pw, sw = ref.wire_group_write(gid, bases), ref.wire_scalar_write(gid, scalars)
    assert len(pw) == n * engine.wire_point_bytes(gid)
    want = ref.wire_group_write(gid, ref.multi_exp(gid, bases, scalars).reshape(1, -1))
    skip_both = engine.WIRE_SKIP_CURVE_CHECK | engine.WIRE_SKIP_SUBGROUP_CHECK
    # default = both checks of group_element_read; then the opt-outs
    for flags in (0, engine.WIRE_SKIP_SUBGROUP_CHECK, skip_both):
        assert engine.multi_exp_wire(gid, pw, sw, flags) == want


@pytest.mark.parametrize("gid", GROUPS)
def test_wire_infinity_and_cancellation(engine, ref, gid):
    g = pm.GROUPS[gid]
    zero, _ = ref.group_constants(gid)
    P = ref.gen_bases_walk(gid, 2)
    negP = ref.group_op(gid, ref.G_NEG, P)
    bases = np.stack([P[0], zero, negP[0], P[1]])
    inf_bytes = pm.wire_point(g, None)
    assert ref.wire_group_write(gid, zero.reshape(1, -1)) == inf_bytes  # (0, 1) encodes infinity
    pw = ref.wire_group_write(gid, bases)
    sc = g.encode_scalars([5, 7, 5, 0])
    assert engine.multi_exp_wire(gid, pw, ref.wire_scalar_write(gid, sc)) == inf_bytes
    sc = g.encode_scalars([5, 7, 4, 0])  # 5 P - 4 P = P
    assert engine.multi_exp_wire(gid, pw, ref.wire_scalar_write(gid, sc)) == pw[:len(inf_bytes)]


@pytest.mark.parametrize("gid", GROUPS)
def test_wire_rejects_what_the_reference_rejects(engine, ref, gid):
    g = pm.GROUPS[gid]
    P = ref.gen_bases_walk(gid, 2)
    pw = bytearray(ref.wire_group_write(gid, P))
    sw = bytearray(ref.wire_scalar_write(gid, ref.gen_scalars(gid, 2)))
    cb = g.fq.nbytes
    skip_sub = engine.WIRE_SKIP_SUBGROUP_CHECK
    # coordinate >= modulus (ffi_serialization.tcc:64-75)
    bad = bytearray(pw)
    bad[-cb:] = g.fq.p.to_bytes(cb, "big")
    assert not ref.wire_group_read(gid, bytes(bad))[1].all()
    with pytest.raises(engine.B200MsmError, match="modulus"):
        engine.multi_exp_wire(gid, bytes(bad), bytes(sw), 0)
    # scalar >= r
    bads = bytearray(sw)
    bads[:g.fr.nbytes] = g.fr.p.to_bytes(g.fr.nbytes, "big")
    assert not ref.wire_scalar_read(gid, bytes(bads))[1].all()
    with pytest.raises(engine.B200MsmError, match="modulus"):
        engine.multi_exp_wire(gid, bytes(pw), bytes(bads), 0)
    # not on the curve (is_well_formed, ffi_serialization.tcc:165): y -> y + 1
    off = bytearray(pw)
    off[-1] ^= 1
    assert not ref.wire_group_read(gid, bytes(off))[1].all()
    with pytest.raises(engine.B200MsmError, match="curve"):
        engine.multi_exp_wire(gid, bytes(off), bytes(sw))
    with pytest.raises(engine.B200MsmError, match="curve"):
        engine.multi_exp_wire(gid, bytes(off), bytes(sw), skip_sub)
    # on the curve but outside the prime-order subgroup (is_in_safe_subgroup); alt_bn128 G1 has
    # cofactor 1, so every curve point passes there
    if gid in (2, 3, 5, 6):  # prime coordinate field and cofactor > 1
        b = g.b
        x = 5
        while True:
            y = pm.sqrt_mod(x * x * x + b, g.fq.p)
            if y is not None:
                break
            x += 1
        stray = pm.wire_point(g, (x, y)) + bytes(pw[len(pw) // 2:])
        accepted = ref.wire_group_read(gid, stray)[1]
        assert not accepted[0] and accepted[1]
        # rejected BY DEFAULT, as the reference's reader rejects it; accepted only on opt-out
        with pytest.raises(engine.B200MsmError, match="subgroup"):
            engine.multi_exp_wire(gid, stray, bytes(sw))
        engine.multi_exp_wire(gid, stray, bytes(sw), skip_sub)  # on the curve: accepted
    # sizes must match exactly
    with pytest.raises(engine.B200MsmError):
        engine.multi_exp_wire(gid, bytes(pw[:-1]), bytes(sw), 0)
