"""TEST INFRASTRUCTURE ONLY — ctypes binding of oracle/libmsmoracle.so, the CPU restatement
("port") of the reference's MSM path written in plain C++ (oracle/msm_oracle.cpp).

Same function names and array conventions as oracle/ref.py (the compiled, unmodified reference),
so tests and bench.py can use either as the checker; `cpu_baseline.kind == "port"` when this one
is timed.  tests/test_oracle.py pins it against tests/golden/ (vectors produced by the reference)
and, where oracle/_ref is present, against the reference itself.  The product package
(libff_b200/) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmsmoracle.so")

# method ids of ref_multi_exp
BDLO12, BDLO12_SIGNED = 0, 1
FORM_NORMAL, FORM_SPECIAL = 0, 1
# field ids of ref_field_op
F_ALT_BN128_FQ, F_ALT_BN128_FR, F_BLS12_377_FQ, F_BLS12_377_FR, F_BW6_761_FQ, F_BW6_761_FR, \
    F_ALT_BN128_FQ2, F_BLS12_377_FQ2, F_BLS12_381_FQ, F_BLS12_381_FR, F_BLS12_381_FQ2 = range(11)
FR_OF_GROUP = {0: F_ALT_BN128_FR, 1: F_ALT_BN128_FR, 2: F_BLS12_377_FR, 3: F_BW6_761_FR,
               4: F_BLS12_377_FR, 5: F_BW6_761_FR, 6: F_BLS12_381_FR, 7: F_BLS12_381_FR}
FQ_OF_GROUP = {0: F_ALT_BN128_FQ, 1: F_ALT_BN128_FQ2, 2: F_BLS12_377_FQ, 3: F_BW6_761_FQ,
               4: F_BLS12_377_FQ2, 5: F_BW6_761_FQ, 6: F_BLS12_381_FQ, 7: F_BLS12_381_FQ2}
FQ2_FIELDS = (F_ALT_BN128_FQ2, F_BLS12_377_FQ2, F_BLS12_381_FQ2)
FQ2_BASE = {F_ALT_BN128_FQ2: F_ALT_BN128_FQ, F_BLS12_377_FQ2: F_BLS12_377_FQ,
            F_BLS12_381_FQ2: F_BLS12_381_FQ}
OP_MUL, OP_SQR, OP_ADD, OP_SUB, OP_NEG, OP_INV, OP_TO_PLAIN, OP_FROM_PLAIN = range(8)
G_ADD, G_ADD_FN, G_MIXED_ADD, G_DBL, G_NEG, G_TO_AFFINE = range(6)

_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError(
                f"{LIB_PATH} missing: run `make -C oracle oracle`")
        L = C.CDLL(LIB_PATH)
        vp, sz, i32, u64, dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_double
        sigs = {
            "orc_init": (i32, []),
            "orc_num_threads": (i32, []),
            "orc_set_num_threads": (None, [i32]),
            "orc_group_bytes": (sz, [i32]),
            "orc_scalar_bytes": (sz, [i32]),
            "orc_scalar_bits": (sz, [i32]),
            "orc_gen_scalars": (None, [i32, u64, sz, vp]),
            "orc_gen_bases_walk": (None, [i32, sz, sz, vp]),
            "orc_scalar_mul": (None, [i32, vp, vp, vp, i32]),
            "orc_multi_exp": (dbl, [i32, i32, i32, vp, vp, sz, sz, vp, i32]),
            "orc_batch_to_special": (None, [i32, vp, sz]),
            "orc_window_table_rows": (sz, [i32, sz, sz]),
            "orc_window_table": (None, [i32, sz, sz, vp, vp]),
            "orc_batch_exp": (dbl, [i32, sz, sz, vp, sz, sz, vp, sz, vp, i32]),
            "orc_group_op": (None, [i32, i32, vp, vp, vp, sz]),
            "orc_group_eq": (None, [i32, vp, vp, vp, sz]),
            "orc_group_constants": (None, [i32, vp, vp]),
            "orc_field_bytes": (sz, [i32]),
            "orc_field_op": (None, [i32, i32, vp, vp, vp, sz]),
            "orc_signed_digit": (C.c_long, [i32, vp, sz, sz]),
            "orc_num_signed_digits": (sz, [i32, sz]),
            "orc_get_digit": (sz, [i32, vp, sz, sz]),
            "orc_pippenger_optimal_c": (sz, [sz]),
            "orc_stream_write": (sz, [i32, i32, vp, sz, sz, vp]),
            "orc_multi_exp_stream": (i32, [i32, i32, vp, sz, vp, sz, sz, vp, i32]),
            "orc_bdlo12_signed_optimal_c": (sz, [sz]),
        }
        for name, (res, args) in sigs.items():
            f = getattr(L, name)
            f.restype = res
            f.argtypes = args
        L.orc_init()
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a):
    return np.ascontiguousarray(a, dtype=np.uint64)


def group_words(gid: int) -> int:
    return lib().orc_group_bytes(gid) // 8


def scalar_words(gid: int) -> int:
    return lib().orc_scalar_bytes(gid) // 8


def gen_scalars(gid: int, n: int, start: int = 0) -> np.ndarray:
    out = np.zeros((n, scalar_words(gid)), dtype=np.uint64)
    lib().orc_gen_scalars(gid, start, n, _p(out))
    return out


def gen_bases_walk(gid: int, n: int, start: int = 0) -> np.ndarray:
    out = np.zeros((n, group_words(gid)), dtype=np.uint64)
    lib().orc_gen_bases_walk(gid, start, n, _p(out))
    return out


def scalar_mul(gid: int, base, scalar, normalize: bool = True) -> np.ndarray:
    base, scalar = _c(base), _c(scalar)
    out = np.zeros(group_words(gid), dtype=np.uint64)
    lib().orc_scalar_mul(gid, _p(base), _p(scalar), _p(out), int(normalize))
    return out


def multi_exp(gid, bases, scalars, method=BDLO12, form=FORM_SPECIAL, chunks=1,
              normalize=True) -> np.ndarray:
    bases, scalars = _c(bases), _c(scalars)
    n = bases.shape[0] if bases.ndim == 2 else bases.size // group_words(gid)
    out = np.zeros(group_words(gid), dtype=np.uint64)
    t = lib().orc_multi_exp(gid, method, form, _p(bases), _p(scalars), n, chunks, _p(out),
                            int(normalize))
    assert t >= 0
    return out


def multi_exp_timed(gid, bases, scalars, form=FORM_SPECIAL, chunks=1):
    """Returns (seconds inside the restated multi_exp<BDLO12>, affine result)."""
    bases, scalars = _c(bases), _c(scalars)
    n = bases.shape[0]
    out = np.zeros(group_words(gid), dtype=np.uint64)
    t = lib().orc_multi_exp(gid, BDLO12, form, _p(bases), _p(scalars), n, chunks, _p(out), 1)
    return t, out


def batch_to_special(gid, pts) -> np.ndarray:
    out = _c(pts).copy()
    lib().orc_batch_to_special(gid, _p(out), out.shape[0])
    return out


def window_table(gid, scalar_bits, window, g) -> np.ndarray:
    g = _c(g)
    rows = lib().orc_window_table_rows(gid, scalar_bits, window)
    out = np.zeros((rows, 1 << window, group_words(gid)), dtype=np.uint64)
    lib().orc_window_table(gid, scalar_bits, window, _p(g), _p(out))
    return out


def batch_exp(gid, scalar_bits, window, table, scalars, normalize=True):
    """Returns (seconds inside libff::batch_exp, results)."""
    table, scalars = _c(table), _c(scalars)
    n = scalars.shape[0]
    out = np.zeros((n, group_words(gid)), dtype=np.uint64)
    t = lib().orc_batch_exp(gid, scalar_bits, window, _p(table), table.shape[0], table.shape[1],
                            _p(scalars), n, _p(out), int(normalize))
    return t, out


def group_op(gid, op, a, b=None) -> np.ndarray:
    a = _c(a)
    b = None if b is None else _c(b)
    out = np.zeros_like(a)
    lib().orc_group_op(gid, op, _p(a), _p(b), _p(out), a.shape[0])
    return out


def group_eq(gid, a, b) -> np.ndarray:
    a, b = _c(a), _c(b)
    out = np.zeros(a.shape[0], dtype=np.uint8)
    lib().orc_group_eq(gid, _p(a), _p(b), _p(out), a.shape[0])
    return out.astype(bool)


def group_constants(gid):
    z = np.zeros(group_words(gid), dtype=np.uint64)
    o = np.zeros(group_words(gid), dtype=np.uint64)
    lib().orc_group_constants(gid, _p(z), _p(o))
    return z, o


def field_op(fid, op, a, b=None) -> np.ndarray:
    a = _c(a)
    b = None if b is None else _c(b)
    out = np.zeros_like(a)
    lib().orc_field_op(fid, op, _p(a), _p(b), _p(out), a.shape[0])
    return out


def signed_digit(fid, scalar_mont, c, idx) -> int:
    s = _c(scalar_mont)
    return int(lib().orc_signed_digit(fid, _p(s), c, idx))


def get_digit(fid, scalar_mont, c, idx) -> int:
    s = _c(scalar_mont)
    return int(lib().orc_get_digit(fid, _p(s), c, idx))


def num_signed_digits(fid, c) -> int:
    return int(lib().orc_num_signed_digits(fid, c))


def pippenger_optimal_c(n) -> int:
    return int(lib().orc_pippenger_optimal_c(n))


# ---- streamed / precomputed MSM (multiexp_stream.tcc:164-223), same shapes as oracle/ref.py ------
FORM_PLAIN, FORM_MONTGOMERY = 0, 1


def stream_write(gid, pts, form=FORM_MONTGOMERY, precompute_c=0) -> bytes:
    a = _c(pts)
    n = a.shape[0]
    size = lib().orc_stream_write(gid, form, _p(a), n, precompute_c, None)
    out = np.zeros(size, dtype=np.uint8)
    lib().orc_stream_write(gid, form, _p(a), n, precompute_c, out.ctypes.data_as(C.c_void_p))
    return out.tobytes()


def multi_exp_stream(gid, data: bytes, scalars, form=FORM_MONTGOMERY, precompute_c=0,
                     normalize=True):
    """None where the reference sums an empty bucket set (undefined there, multiexp.tcc:88-106)."""
    s = _c(scalars)
    buf = np.frombuffer(data, dtype=np.uint8)
    out = np.zeros(group_words(gid), dtype=np.uint64)
    rc = lib().orc_multi_exp_stream(gid, form, buf.ctypes.data_as(C.c_void_p), len(data), _p(s),
                                    s.shape[0], precompute_c, _p(out), int(normalize))
    if rc < 0:
        raise ValueError("orc_multi_exp_stream: unknown group")
    return out if rc == 1 else None


def bdlo12_signed_optimal_c(n) -> int:
    return int(lib().orc_bdlo12_signed_optimal_c(n))
