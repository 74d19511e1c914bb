"""libff_b200 -- Python-side mirror of the libff MSM surface over the b200msm C ABI.

The product is the C ABI shared library `libff_b200/libb200msm.so` (include/b200msm.h), built from
the hand-written sm_100a kernels in `libff_b200/csrc`.  This module is a thin ctypes binding whose
function names, argument meaning and error behaviour follow the reference templates

    libff::multi_exp<G, Fr, Method, BaseForm>(bases, scalars, chunks)   multiexp.tcc:643-688
    libff::batch_exp<G, Fr>(scalar_size, window, table, v)              multiexp.tcc:875-915
    libff::batch_to_special<G>(vec)                                     multiexp.tcc:949-974

(paths under libff/algebra/scalar_multiplication/).  Arrays are numpy uint64 in the reference's
in-memory layout: field elements as little-endian 64-bit Montgomery limbs, group elements as
X | Y | Z.  There is no CPU fallback: importing works anywhere, but every compute call raises
`B200MsmError` unless the CUDA library is present and a device is bound.

The C++ drop-in (template specialisations for existing libff call sites) is
`libff_b200/include/libff_b200.hpp`; see INTEGRATION.md.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

__all__ = [
    "ALT_BN128_G1", "ALT_BN128_G2", "BLS12_377_G1", "BW6_761_G1", "BLS12_377_G2", "BW6_761_G2",
    "BLS12_381_G1", "BLS12_381_G2", "GROUP_NAMES",
    "FORM_NORMAL", "FORM_SPECIAL", "B200MsmError", "lib_path", "load", "init", "shutdown",
    "group_bytes", "scalar_bytes", "multi_exp", "batch_exp", "batch_to_special", "group_sum",
    "ResidentBases", "ResidentScalars", "multi_exp_resident", "last_timings", "launch_count",
    "timer_start", "timer_stop", "PrecomputedBases", "multi_exp_precomputed", "precompute_digits",
    "precompute_window_bits", "MsmStream", "STREAM_MEMORY_SPECIAL", "STREAM_MEMORY_NORMAL",
    "STREAM_BINARY_PLAIN", "STREAM_BINARY_MONTGOMERY",
    "set_option", "window_bits", "shard_range", "multi_exp_wire", "wire_point_bytes",
    "WIRE_SKIP_CURVE_CHECK", "WIRE_SKIP_SUBGROUP_CHECK", "init_devices", "init_all",
    "device_count", "multi_exp_resident_bases", "batch_exp_with_coeff", "WindowTable",
    "STREAM_REFERENCE_CARRY", "device_timings",
]

ALT_BN128_G1, ALT_BN128_G2, BLS12_377_G1, BW6_761_G1 = 0, 1, 2, 3
BLS12_377_G2, BW6_761_G2, BLS12_381_G1, BLS12_381_G2 = 4, 5, 6, 7
GROUP_NAMES = {0: "alt_bn128_G1", 1: "alt_bn128_G2", 2: "bls12_377_G1", 3: "bw6_761_G1",
               4: "bls12_377_G2", 5: "bw6_761_G2", 6: "bls12_381_G1", 7: "bls12_381_G2"}
FORM_NORMAL, FORM_SPECIAL = 0, 1  # multi_exp_base_form (multiexp.hpp:49-58)

# every symbol include/b200msm.h declares (tests check the library exports all of them)
ABI_SYMBOLS = [
    "b200msm_init", "b200msm_shutdown", "b200msm_last_error", "b200msm_group_bytes",
    "b200msm_scalar_bytes", "b200msm_multi_exp", "b200msm_bases_upload",
    "b200msm_scalars_upload", "b200msm_multi_exp_resident", "b200msm_release", "b200msm_sum",
    "b200msm_batch_to_special", "b200msm_batch_exp", "b200msm_debug_op", "b200msm_set_option",
    "b200msm_last_timings", "b200msm_timing_name", "b200msm_launch_count", "b200msm_window_bits",
    "b200msm_wire_point_bytes", "b200msm_multi_exp_wire", "b200msm_timer_start",
    "b200msm_timer_stop", "b200msm_precompute_window_bits", "b200msm_precompute_digits",
    "b200msm_precompute", "b200msm_precomputed_upload", "b200msm_multi_exp_precomputed",
    "b200msm_stream_begin", "b200msm_stream_push", "b200msm_stream_finish",
    "b200msm_init_devices", "b200msm_init_all", "b200msm_device_count",
    "b200msm_multi_exp_resident_bases", "b200msm_batch_exp_with_coeff", "b200msm_table_upload",
    "b200msm_table_build", "b200msm_table_shape", "b200msm_table_download",
    "b200msm_batch_exp_resident", "b200msm_device_timings", "b200msm_batch_to_special_noalloc",
]


class B200MsmError(RuntimeError):
    pass


_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None
_bound_device: Optional[int] = None


def lib_path() -> str:
    # B200MSM_LIB: development aid for A/B runs of differently compiled engines
    return os.environ.get("B200MSM_LIB") or os.path.join(_HERE, "libb200msm.so")


def load():
    """Loads the C ABI library (no CUDA call is made).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise B200MsmError(
            f"{path} is missing: build it with `make -C libff_b200/csrc -j` (or "
            "`python -c 'import __graft_entry__ as g; g.build()'`); there is no CPU fallback")
    L = C.CDLL(path)
    vp, sz, u32, u64, i64 = C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_int64
    sigs = {
        "b200msm_init": (C.c_bool, [C.c_int]),
        "b200msm_shutdown": (None, []),
        "b200msm_last_error": (C.c_char_p, []),
        "b200msm_group_bytes": (sz, [u32]),
        "b200msm_scalar_bytes": (sz, [u32]),
        "b200msm_multi_exp": (C.c_bool, [u32, vp, sz, vp, sz, sz, u32, vp, sz]),
        "b200msm_bases_upload": (C.c_bool, [u32, vp, sz, sz, u32, C.POINTER(u64)]),
        "b200msm_scalars_upload": (C.c_bool, [u32, vp, sz, sz, C.POINTER(u64)]),
        "b200msm_multi_exp_resident": (C.c_bool, [u32, u64, u64, sz, sz, vp, sz]),
        "b200msm_release": (C.c_bool, [u64]),
        "b200msm_sum": (C.c_bool, [u32, vp, sz, sz, vp, sz]),
        "b200msm_batch_to_special": (C.c_bool, [u32, vp, sz, sz]),
        "b200msm_batch_to_special_noalloc": (C.c_bool, [u32, vp, vp, sz, sz, vp]),
        "b200msm_batch_exp": (C.c_bool, [u32, sz, sz, C.POINTER(vp), sz, sz, vp, sz, sz, vp, sz]),
        "b200msm_debug_op": (C.c_bool, [u32, u32, vp, vp, vp, sz, sz]),
        "b200msm_set_option": (C.c_bool, [C.c_char_p, i64]),
        "b200msm_last_timings": (sz, [C.POINTER(C.c_float), sz]),
        "b200msm_timing_name": (C.c_char_p, [sz]),
        "b200msm_launch_count": (u64, []),
        "b200msm_window_bits": (u32, [u32, sz]),
        "b200msm_wire_point_bytes": (sz, [u32]),
        "b200msm_multi_exp_wire": (C.c_bool, [u32, vp, sz, vp, sz, sz, u32, vp, sz]),
        "b200msm_precompute_window_bits": (u32, [u32, sz]),
        "b200msm_precompute_digits": (sz, [u32, u32]),
        "b200msm_precompute": (C.c_bool, [u32, u64, u32, C.POINTER(u64)]),
        "b200msm_precomputed_upload": (C.c_bool, [u32, vp, sz, sz, u32, u32, C.POINTER(u64)]),
        "b200msm_multi_exp_precomputed": (C.c_bool, [u32, u64, u64, vp, sz, vp, sz]),
        "b200msm_stream_begin": (C.c_bool, [u32, vp, sz, sz, u32, u32, C.POINTER(u64)]),
        "b200msm_stream_push": (C.c_bool, [u64, vp, sz, sz]),
        "b200msm_stream_finish": (C.c_bool, [u64, vp, sz]),
        "b200msm_timer_start": (C.c_bool, []),
        "b200msm_timer_stop": (C.c_float, []),
        "b200msm_init_devices": (C.c_bool, [C.POINTER(C.c_int), sz]),
        "b200msm_init_all": (C.c_int, []),
        "b200msm_device_count": (sz, []),
        "b200msm_multi_exp_resident_bases": (C.c_bool, [u32, u64, vp, sz, sz, vp, sz]),
        "b200msm_batch_exp_with_coeff": (C.c_bool, [u32, sz, sz, C.POINTER(vp), sz, sz, vp, sz, vp,
                                                    sz, sz, vp, sz]),
        "b200msm_table_upload": (C.c_bool, [u32, sz, sz, C.POINTER(vp), sz, sz, C.POINTER(u64)]),
        "b200msm_table_build": (C.c_bool, [u32, sz, sz, vp, sz, C.POINTER(u64)]),
        "b200msm_table_shape": (C.c_bool, [u64, C.POINTER(sz), C.POINTER(sz), C.POINTER(sz),
                                           C.POINTER(sz)]),
        "b200msm_table_download": (C.c_bool, [u64, C.POINTER(vp), sz, sz]),
        "b200msm_batch_exp_resident": (C.c_bool, [u32, u64, vp, sz, sz, vp, sz, vp, sz]),
        "b200msm_device_timings": (sz, [sz, C.POINTER(C.c_float), sz]),
    }
    for name, (res, args) in sigs.items():
        f = getattr(L, name)
        f.restype = res
        f.argtypes = args
    _lib = L
    return L


def _fail(what: str):
    msg = load().b200msm_last_error()
    raise B200MsmError(f"{what}: {msg.decode() if msg else 'unknown error'}")


def init(device: Optional[int] = None) -> int:
    """Binds this process to one CUDA device (default: LOCAL_RANK, else 0)."""
    global _bound_device
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    if not load().b200msm_init(int(device)):
        _fail("b200msm_init")
    _bound_device = int(device)
    return _bound_device


def init_devices(devices: Sequence[int]) -> int:
    """Binds several GPUs of the box (b200msm_init_devices): every later multi_exp / batch call
    cuts its range into one contiguous range per device, as the reference cuts it into `chunks`
    (multiexp.tcc:655-687).  A device may be listed twice (two contexts on one GPU; a test aid)."""
    global _bound_device
    ids = (C.c_int * len(devices))(*[int(d) for d in devices])
    if not load().b200msm_init_devices(ids, len(devices)):
        _fail("b200msm_init_devices")
    _bound_device = int(devices[0])
    return len(devices)


def init_all() -> int:
    """Binds every visible CUDA device; returns their number."""
    global _bound_device
    k = int(load().b200msm_init_all())
    if k <= 0:
        _fail("b200msm_init_all")
    _bound_device = 0
    return k


def device_count() -> int:
    return int(load().b200msm_device_count())


def shutdown() -> None:
    global _bound_device
    if _lib is not None:
        _lib.b200msm_shutdown()
    _bound_device = None


def _ensure_init():
    if _bound_device is None:
        init()


def group_bytes(group: int) -> int:
    return int(load().b200msm_group_bytes(group))


def scalar_bytes(group: int) -> int:
    return int(load().b200msm_scalar_bytes(group))


def _as_u64(a, what: str) -> np.ndarray:
    a = np.ascontiguousarray(a)
    if a.dtype != np.uint64:
        raise B200MsmError(f"{what}: expected uint64 limbs, got {a.dtype}")
    return a


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _count(a: np.ndarray, elem_bytes: int, what: str) -> int:
    if a.nbytes % elem_bytes:
        raise B200MsmError(f"{what}: {a.nbytes} bytes is not a multiple of {elem_bytes}")
    return a.nbytes // elem_bytes


def shard_range(n: int, rank: int, world: int):
    """Point range of one GPU shard: contiguous ranges, the last takes the remainder -- the
    reference's own `chunks` split (multiexp.tcc:663-678)."""
    one = n // world
    lo = rank * one
    hi = n if rank == world - 1 else lo + one
    return lo, hi


def multi_exp(group: int, bases, scalars, base_form: int = FORM_SPECIAL,
              chunks: int = 1) -> np.ndarray:
    """libff::multi_exp<G, Fr, multi_exp_method_BDLO12, base_form>(bases, scalars, chunks).

    `chunks` is accepted for call-site compatibility and ignored: the reference uses it to split
    the range over OpenMP threads (multiexp.tcc:655-687); one device consumes the whole range.
    Returns one group element (3 coordinates) in the reference's layout."""
    _ensure_init()
    L = load()
    b = _as_u64(bases, "bases")
    s = _as_u64(scalars, "scalars")
    gb, sb = group_bytes(group), scalar_bytes(group)
    if gb == 0:
        raise B200MsmError(f"unknown group id {group}")
    n = _count(b, gb, "bases")
    if _count(s, sb, "scalars") != n:
        raise B200MsmError(f"multi_exp: {n} bases but {_count(s, sb, 'scalars')} scalars")
    out = np.zeros(gb // 8, dtype=np.uint64)
    if not L.b200msm_multi_exp(group, _ptr(b), b.nbytes, _ptr(s), s.nbytes, n, base_form,
                               _ptr(out), out.nbytes):
        _fail("b200msm_multi_exp")
    return out


def multi_exp_host_ptr(group: int, bases_ptr: int, scalars_ptr: int, n: int,
                       base_form: int = FORM_SPECIAL) -> np.ndarray:
    """multi_exp on raw host addresses (e.g. pinned buffers owned by the caller)."""
    _ensure_init()
    gb, sb = group_bytes(group), scalar_bytes(group)
    out = np.zeros(gb // 8, dtype=np.uint64)
    if not load().b200msm_multi_exp(group, C.c_void_p(bases_ptr), n * gb, C.c_void_p(scalars_ptr),
                                    n * sb, n, base_form, _ptr(out), out.nbytes):
        _fail("b200msm_multi_exp")
    return out


# opt-outs of the point checks group_element_read always makes (ffi_serialization.tcc:165)
WIRE_SKIP_CURVE_CHECK, WIRE_SKIP_SUBGROUP_CHECK = 1, 2


def wire_point_bytes(group: int) -> int:
    return int(load().b200msm_wire_point_bytes(group))


def multi_exp_wire(group: int, points: bytes, scalars: bytes, flags: int = 0) -> bytes:
    """MSM on the byte format of the reference's ffi layer (ffi/ffi_serialization.hpp:12-16): affine
    big-endian plain X | Y points ((0, 1) = infinity) and big-endian plain scalars; returns the
    affine result in the same format.  Raises on malformed input -- a value >= its modulus, a point
    off the curve or outside the prime-order subgroup -- as the reference's readers return false
    (ffi_serialization.tcc:64-75,165); `flags` opts out of a point check."""
    _ensure_init()
    pb, sb = wire_point_bytes(group), scalar_bytes(group)
    if pb == 0:
        raise B200MsmError(f"unknown group id {group}")
    if len(points) % pb or len(scalars) % sb or len(points) // pb != len(scalars) // sb:
        raise B200MsmError("multi_exp_wire: buffer sizes do not describe the same number of elements")
    n = len(points) // pb
    pa = np.frombuffer(points, dtype=np.uint8)
    sa = np.frombuffer(scalars, dtype=np.uint8)
    out = np.zeros(pb, dtype=np.uint8)
    if not load().b200msm_multi_exp_wire(group, pa.ctypes.data if n else None, pa.nbytes,
                                         sa.ctypes.data if n else None, sa.nbytes, n, flags,
                                         out.ctypes.data, out.nbytes):
        _fail("b200msm_multi_exp_wire")
    return out.tobytes()


def group_sum(group: int, points) -> np.ndarray:
    """Sum of group elements: the partial-result reduction that ends libff::multi_exp
    (multiexp.tcc:681-685); combines the per-GPU partial MSMs."""
    _ensure_init()
    p = _as_u64(points, "points")
    gb = group_bytes(group)
    cnt = _count(p, gb, "points")
    out = np.zeros(gb // 8, dtype=np.uint64)
    if not load().b200msm_sum(group, _ptr(p), p.nbytes, cnt, _ptr(out), out.nbytes):
        _fail("b200msm_sum")
    return out


def batch_to_special(group: int, points, inplace: bool = False) -> np.ndarray:
    """libff::batch_to_special<G>(vec): returns the normalised copy (the C ABI works in place, as
    the reference does; `inplace=True` normalises the caller's own contiguous array)."""
    _ensure_init()
    p = _as_u64(points, "points")
    if not inplace or p is not points:
        p = p.copy()
    gb = group_bytes(group)
    n = _count(p, gb, "points")
    if not load().b200msm_batch_to_special(group, _ptr(p), p.nbytes, n):
        _fail("b200msm_batch_to_special")
    return p


NOALLOC_MAX_POINTS = 32768  # B200MSM_NOALLOC_MAX_POINTS


def batch_to_special_noalloc_ptr(group: int, d_in: int, d_out: int, nbytes: int, n: int,
                                 stream: int = 0) -> None:
    """b200msm_batch_to_special_noalloc: batch_to_special on DEVICE pointers of the current
    device with the scratch in the library's static device array (no cudaMalloc, no engine
    context needed); enqueued on `stream` (a cudaStream_t handle), not synchronised.  At most
    NOALLOC_MAX_POINTS per call; overlapping calls on one device race on the shared scratch."""
    if not load().b200msm_batch_to_special_noalloc(group, d_in, d_out, nbytes, n, stream or None):
        _fail("b200msm_batch_to_special_noalloc")


# element-wise leaves (b200msm_debug_op)
FQ_ADD, FQ_SUB, FQ_MUL, FQ_SQR, FQ_NEG, FQ_INV, FR_FROM_MONT, FR_MUL = range(8)
G_ADD, G_DBL, G_MADD, G_NEG, FQ_MUL2 = 10, 11, 12, 13, 14
FQ_MUL_ROLLED, FQ_MUL2_ROLLED, FQ_MUL_CALLED, FQ_SQR_CALLED, FQ_MUL_SEPARATED = 15, 16, 17, 18, 19


def debug_op(group: int, op: int, a, b=None) -> np.ndarray:
    """One field / group operation per element on the device (parity tests of the leaves)."""
    _ensure_init()
    a = _as_u64(a, "a")
    n = a.shape[0]
    out = np.zeros_like(a)
    bp = None
    if b is not None:
        b = _as_u64(b, "b")
        if b.shape != a.shape:
            raise B200MsmError("debug_op: operand shapes differ")
        bp = _ptr(b)
    if not load().b200msm_debug_op(group, op, _ptr(a), bp, _ptr(out), a.nbytes, n):
        _fail("b200msm_debug_op")
    return out


def batch_exp(group: int, scalar_size: int, window: int, table, scalars, out=None) -> np.ndarray:
    """libff::batch_exp<G, Fr>(scalar_size, window, table, v).  `table` is a sequence of rows
    (window_table<G> = vector<vector<G>>), a (rows, 1 << window, words) array, or a resident
    WindowTable."""
    return batch_exp_with_coeff(group, scalar_size, window, table, None, scalars, out)


def batch_exp_with_coeff(group: int, scalar_size: int, window: int, table, coeff,
                         scalars, out=None) -> np.ndarray:
    """libff::batch_exp_with_coeff<G, Fr>(scalar_size, window, table, coeff, v) (multiexp.tcc:
    917-947): out[i] = (coeff * v[i]) g.  coeff: one Fr element (Montgomery limbs) or None.
    `out`: a caller-owned (n, words) uint64 array to write into (else a new one)."""
    _ensure_init()
    gb, sb = group_bytes(group), scalar_bytes(group)
    s = _as_u64(scalars, "scalars")
    n = _count(s, sb, "scalars")
    if out is None:
        out = np.zeros((n, gb // 8), dtype=np.uint64)
    elif out.dtype != np.uint64 or not out.flags.c_contiguous or out.nbytes != n * gb:
        raise B200MsmError("batch_exp: `out` must be a contiguous uint64 array of n elements")
    cf = None if coeff is None else _as_u64(coeff, "coeff").reshape(-1)
    cp, cb = (None, 0) if cf is None else (_ptr(cf), cf.nbytes)
    if isinstance(table, WindowTable):
        if (table.scalar_size, table.window) != (scalar_size, window):
            raise B200MsmError("batch_exp: the resident table was built for another shape")
        if not load().b200msm_batch_exp_resident(group, table.handle, _ptr(s), s.nbytes, n, cp, cb,
                                                 _ptr(out), out.nbytes):
            _fail("b200msm_batch_exp_resident")
        return out
    rows = [_as_u64(r, "table row") for r in table]
    row_len = _count(rows[0], gb, "table row") if rows else 0
    ptrs = (C.c_void_p * len(rows))(*[r.ctypes.data for r in rows])
    if not load().b200msm_batch_exp_with_coeff(group, scalar_size, window, ptrs, len(rows), row_len,
                                               cp, cb, _ptr(s), s.nbytes, n, _ptr(out), out.nbytes):
        _fail("b200msm_batch_exp_with_coeff")
    return out


class _Resident:
    def __init__(self, handle: int, group: int, n: int):
        self.handle, self.group, self.n = handle, group, n

    def release(self):
        if self.handle and _lib is not None:
            _lib.b200msm_release(self.handle)
        self.handle = 0

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


class WindowTable(_Resident):
    """A fixed-base window table kept in HBM, normalised once (get_window_table, multiexp.tcc:
    809-846).  `g=`: built on the device from the base; `table=`: uploaded from host rows."""

    def __init__(self, group: int, scalar_size: int, window: int, g=None, table=None):
        _ensure_init()
        h = C.c_uint64(0)
        gb = group_bytes(group)
        self.scalar_size, self.window = scalar_size, window
        self.rows = (scalar_size + window - 1) // window
        if g is not None:
            ga = _as_u64(g, "g").reshape(-1)
            if not load().b200msm_table_build(group, scalar_size, window, _ptr(ga), ga.nbytes,
                                              C.byref(h)):
                _fail("b200msm_table_build")
        else:
            rows = [_as_u64(r, "table row") for r in table]
            row_len = _count(rows[0], gb, "table row") if rows else 0
            ptrs = (C.c_void_p * len(rows))(*[r.ctypes.data for r in rows])
            if not load().b200msm_table_upload(group, scalar_size, window, ptrs, len(rows), row_len,
                                               C.byref(h)):
                _fail("b200msm_table_upload")
        super().__init__(h.value, group, self.rows << window)

    def download(self) -> np.ndarray:
        """(rows, 1 << window, words) array in the reference's layout, special form."""
        words = group_bytes(self.group) // 8
        out = np.zeros((self.rows, 1 << self.window, words), dtype=np.uint64)
        ptrs = (C.c_void_p * self.rows)(*[out[r].ctypes.data for r in range(self.rows)])
        if not load().b200msm_table_download(self.handle, ptrs, self.rows, 1 << self.window):
            _fail("b200msm_table_download")
        return out


class ResidentBases(_Resident):
    """Bases kept in HBM (a prover's fixed bases): uploaded and normalised once."""

    def __init__(self, group: int, bases, base_form: int = FORM_SPECIAL):
        _ensure_init()
        b = _as_u64(bases, "bases")
        n = _count(b, group_bytes(group), "bases")
        h = C.c_uint64(0)
        if not load().b200msm_bases_upload(group, _ptr(b), b.nbytes, n, base_form, C.byref(h)):
            _fail("b200msm_bases_upload")
        super().__init__(h.value, group, n)


class ResidentScalars(_Resident):
    def __init__(self, group: int, scalars):
        _ensure_init()
        s = _as_u64(scalars, "scalars")
        n = _count(s, scalar_bytes(group), "scalars")
        h = C.c_uint64(0)
        if not load().b200msm_scalars_upload(group, _ptr(s), s.nbytes, n, C.byref(h)):
            _fail("b200msm_scalars_upload")
        super().__init__(h.value, group, n)


def multi_exp_resident(bases: ResidentBases, scalars: ResidentScalars, first: int = 0,
                       n: Optional[int] = None) -> np.ndarray:
    group = bases.group
    if n is None:
        n = bases.n - first
    out = np.zeros(group_bytes(group) // 8, dtype=np.uint64)
    if not load().b200msm_multi_exp_resident(group, bases.handle, scalars.handle, first, n,
                                             _ptr(out), out.nbytes):
        _fail("b200msm_multi_exp_resident")
    return out


def multi_exp_resident_bases(bases: ResidentBases, scalars) -> np.ndarray:
    """The prover-shaped call: resident bases, host scalars uploaded under the computation."""
    group = bases.group
    s = _as_u64(scalars, "scalars")
    n = _count(s, scalar_bytes(group), "scalars")
    out = np.zeros(group_bytes(group) // 8, dtype=np.uint64)
    if not load().b200msm_multi_exp_resident_bases(group, bases.handle, _ptr(s), s.nbytes, n,
                                                   _ptr(out), out.nbytes):
        _fail("b200msm_multi_exp_resident_bases")
    return out


def multi_exp_resident_bases_ptr(bases: ResidentBases, scalars_ptr: int, n: int) -> np.ndarray:
    """multi_exp_resident_bases on a raw host address (e.g. a pinned buffer owned by the caller)."""
    group = bases.group
    out = np.zeros(group_bytes(group) // 8, dtype=np.uint64)
    if not load().b200msm_multi_exp_resident_bases(group, bases.handle, C.c_void_p(scalars_ptr),
                                                   n * scalar_bytes(group), n, _ptr(out), out.nbytes):
        _fail("b200msm_multi_exp_resident_bases")
    return out


# ---- precomputed multiples (multi_exp_stream_with_precompute, multiexp_stream.hpp:27-41) ---------

def precompute_window_bits(group: int, n: int) -> int:
    return int(load().b200msm_precompute_window_bits(group, n))


def precompute_digits(group: int, c: int) -> int:
    """multiples per base for digit size c: (Fr::num_bits + c - 1) / c (multiexp_stream.tcc:205)"""
    return int(load().b200msm_precompute_digits(group, c))


class PrecomputedBases(_Resident):
    """Table of the multiples 2^(k c) P_i kept in HBM.  Either built on the device from resident
    bases (`bases=`), or uploaded in the order of the reference's precompute stream (`elements=`:
    n * digits in-memory elements, for each base its multiples)."""

    def __init__(self, group: int, bases: Optional[ResidentBases] = None, c: int = 0,
                 elements=None, base_form: int = FORM_SPECIAL):
        _ensure_init()
        h = C.c_uint64(0)
        if bases is not None:
            if not load().b200msm_precompute(group, bases.handle, c, C.byref(h)):
                _fail("b200msm_precompute")
            n = bases.n
            self.c = c or precompute_window_bits(group, n)
        else:
            e = _as_u64(elements, "elements")
            per = precompute_digits(group, c)
            total = _count(e, group_bytes(group), "elements")
            if per == 0 or total % per:
                raise B200MsmError(f"elements: {total} is not a multiple of {per} digits")
            n = total // per
            if not load().b200msm_precomputed_upload(group, _ptr(e), e.nbytes, n, c, base_form,
                                                     C.byref(h)):
                _fail("b200msm_precomputed_upload")
            self.c = c
        super().__init__(h.value, group, n)


def multi_exp_precomputed(table: PrecomputedBases, scalars) -> np.ndarray:
    """sum s_i P_i from the table; `scalars` is a ResidentScalars or a host array of n scalars."""
    group = table.group
    out = np.zeros(group_bytes(group) // 8, dtype=np.uint64)
    if isinstance(scalars, ResidentScalars):
        ok = load().b200msm_multi_exp_precomputed(group, table.handle, scalars.handle, None, 0,
                                                  _ptr(out), out.nbytes)
    else:
        s = _as_u64(scalars, "scalars")
        ok = load().b200msm_multi_exp_precomputed(group, table.handle, 0, _ptr(s), s.nbytes,
                                                  _ptr(out), out.nbytes)
    if not ok:
        _fail("b200msm_multi_exp_precomputed")
    return out


# ---- streamed bases (multi_exp_stream, multiexp_stream.tcc:164-193) ------------------------------
STREAM_MEMORY_SPECIAL, STREAM_MEMORY_NORMAL, STREAM_BINARY_PLAIN, STREAM_BINARY_MONTGOMERY = 0, 1, 2, 3
# or'ed into precompute_c: reproduce the reference's dropped top carry (field_utils.tcc:205-239)
STREAM_REFERENCE_CARRY = 0x100


class MsmStream:
    """The scalars are known up front, the bases arrive in chunks (`push`), `finish` returns the sum.
    encoding: in-memory elements, or the bytes of the reference's uncompressed binary stream
    encoding (decoded on the device).  precompute_c != 0: every base comes as its multiples."""

    def __init__(self, group: int, scalars, encoding: int = STREAM_MEMORY_SPECIAL,
                 precompute_c: int = 0, reference_carry: bool = False):
        _ensure_init()
        s = _as_u64(scalars, "scalars")
        self.group, self.n = group, _count(s, scalar_bytes(group), "scalars")
        self.encoding, self.c = encoding, precompute_c
        h = C.c_uint64(0)
        flag = STREAM_REFERENCE_CARRY if (reference_carry and precompute_c) else 0
        if not load().b200msm_stream_begin(group, _ptr(s), s.nbytes, self.n, encoding,
                                           precompute_c | flag, C.byref(h)):
            _fail("b200msm_stream_begin")
        self.handle = h.value

    def push(self, elements, count: Optional[int] = None) -> None:
        if isinstance(elements, (bytes, bytearray, memoryview)):
            buf = np.frombuffer(elements, dtype=np.uint8)
        else:
            buf = np.ascontiguousarray(elements)
        if count is None:
            per = precompute_digits(self.group, self.c) if self.c else 1
            eb = group_bytes(self.group) if self.encoding <= STREAM_MEMORY_NORMAL \
                else wire_point_bytes(self.group)
            count = buf.nbytes // (eb * per)
        if not load().b200msm_stream_push(self.handle, buf.ctypes.data_as(C.c_void_p), buf.nbytes, count):
            _fail("b200msm_stream_push")

    def finish(self) -> np.ndarray:
        out = np.zeros(group_bytes(self.group) // 8, dtype=np.uint64)
        h, self.handle = self.handle, 0
        if not load().b200msm_stream_finish(h, _ptr(out), out.nbytes):
            _fail("b200msm_stream_finish")
        return out

    def __del__(self):
        try:
            if self.handle and _lib is not None:
                tmp = np.zeros(group_bytes(self.group) // 8, dtype=np.uint64)
                _lib.b200msm_stream_finish(self.handle, _ptr(tmp), tmp.nbytes)
        except Exception:
            pass


def last_timings() -> dict:
    """CUDA-event phase timings (ms) of the most recent multi_exp, on the engine's stream."""
    L = load()
    buf = (C.c_float * 16)()
    k = L.b200msm_last_timings(buf, 16)
    return {L.b200msm_timing_name(i).decode(): float(buf[i]) for i in range(k)}


def device_timings(index: int) -> dict:
    """phase timings of one bound device (last_timings is the maximum over the devices)"""
    L = load()
    buf = (C.c_float * 16)()
    k = L.b200msm_device_timings(index, buf, 16)
    return {L.b200msm_timing_name(i).decode(): float(buf[i]) for i in range(k)}


def launch_count() -> int:
    return int(load().b200msm_launch_count())


def timer_start() -> None:
    """Records a CUDA event on the engine's stream (device-side stopwatch for benchmarks)."""
    _ensure_init()
    if not load().b200msm_timer_start():
        _fail("timer_start")


def timer_stop() -> float:
    """Milliseconds on the device between timer_start() and now (waits for the stream)."""
    ms = float(load().b200msm_timer_stop())
    if ms < 0:
        _fail("timer_stop")
    return ms


def set_option(name: str, value: int) -> None:
    _ensure_init()
    if not load().b200msm_set_option(name.encode(), int(value)):
        _fail("b200msm_set_option")


def window_bits(group: int, n: int) -> int:
    _ensure_init()
    return int(load().b200msm_window_bits(group, n))
