"""CPU-side checks of the drop-in boundary: the C ABI library loads and exports every symbol
include/b200msm.h declares, the size queries need no device, the host mirror validates arguments
like the reference's ffi layer (ffi/ffi_serialization.tcc:101-104: exact sizes or failure), the
product never imports the oracle, and without a CUDA device every compute entry point fails loudly
(no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

import libff_b200 as m

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "b200msm.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200msm_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    syms = declared_symbols()
    assert len(syms) >= 19
    assert set(syms) == set(m.ABI_SYMBOLS), "libff_b200.ABI_SYMBOLS out of sync with b200msm.h"
    lib = ctypes.CDLL(m.lib_path())
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/b200msm.h but not exported"


def test_size_queries_match_reference_sizeof():
    """sizeof(GroupT) / sizeof(Fr) of the reference (SURVEY.md §8 sizes table)."""
    assert [m.group_bytes(g) for g in range(8)] == [96, 192, 144, 288, 288, 288, 144, 288]
    assert [m.scalar_bytes(g) for g in range(8)] == [32, 32, 32, 48, 32, 48, 32, 32]
    assert m.group_bytes(8) == 0 and m.scalar_bytes(8) == 0


def test_shard_range_is_the_reference_chunk_split():
    """multiexp.tcc:663-678: `chunks` contiguous ranges of n / chunks, the last takes the rest."""
    for n in (0, 1, 7, 8, 1000, (1 << 24) + 5):
        for world in (1, 2, 3, 8):
            r = [m.shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            assert all(hi - lo == n // world for lo, hi in r[:-1])


def test_product_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "libff_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", "Makefile")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), f
                assert "libmsmoracle" not in src and "libffref" not in src, f


def test_argument_validation_happens_before_any_device_work():
    b = np.zeros((4, 12), dtype=np.uint64)
    s = np.zeros((3, 4), dtype=np.uint64)
    with pytest.raises(m.B200MsmError):
        m._count(np.zeros(13, dtype=np.uint64), 96, "bases")
    with pytest.raises(m.B200MsmError):
        m._as_u64(np.zeros(4, dtype=np.uint32), "x")
    assert m._count(b, 96, "bases") == 4 and m._count(s, 32, "scalars") == 3


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    lib = m.load()
    out = np.zeros(12, dtype=np.uint64)
    b = np.zeros((1, 12), dtype=np.uint64)
    s = np.zeros((1, 4), dtype=np.uint64)
    # not initialised: the C ABI refuses
    ok = lib.b200msm_multi_exp(0, b.ctypes.data, b.nbytes, s.ctypes.data, s.nbytes, 1, 1,
                               out.ctypes.data, out.nbytes)
    assert not ok
    assert b"no CPU path" in lib.b200msm_last_error()
    with pytest.raises(m.B200MsmError):
        m.init(0)
    with pytest.raises(m.B200MsmError):
        m.multi_exp(0, b, s)
    with pytest.raises(m.B200MsmError):
        m.batch_to_special(0, b)
