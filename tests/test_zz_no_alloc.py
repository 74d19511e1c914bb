"""b200msm_batch_to_special_noalloc (include/b200msm.h): batch_to_special on device memory with the
inversion tree's scratch in the library's file-scope `static __device__` array -- no cudaMalloc, for
callers under CUDA graph capture or in loops that must not allocate.  It has no counterpart in the
reference; it is checked here, through the C ABI, against the existing entry point
b200msm_batch_to_special (and that one against the unmodified reference), bit for bit, on calls in
a row that share the static scratch.

Collected last on purpose (the file name sorts after every other test file).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def jacobian_points(ref, gid, n, start):
    """n elements with arbitrary Z (sums of neighbouring walk points) and a few zeros"""
    aff = ref.gen_bases_walk(gid, n + 1 + start)[start:]
    pts = ref.group_op(gid, ref.G_ADD, aff[:-1], aff[1:])
    zero, _ = ref.group_constants(gid)
    for k in (0, n // 2, n - 1):
        pts[k] = zero
    return np.ascontiguousarray(pts)


def noalloc(engine, gid, pts, inplace):
    """device buffers from torch (allocated BEFORE the call; the entry point itself allocates
    nothing), the call on torch's current stream, the result copied back after a synchronise"""
    import torch
    torch.cuda.set_device(0)
    raw = np.ascontiguousarray(pts).view(np.uint8).reshape(-1)
    d_in = torch.from_numpy(raw.copy()).cuda()
    d_out = d_in if inplace else torch.full_like(d_in, 0xA5)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    before = torch.cuda.memory_allocated()
    engine.batch_to_special_noalloc_ptr(gid, d_in.data_ptr(), d_out.data_ptr(), raw.nbytes,
                                        pts.shape[0], stream)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    if not inplace:  # the input is left alone
        np.testing.assert_array_equal(d_in.cpu().numpy(), raw)
    return d_out.cpu().numpy().view(np.uint64).reshape(pts.shape)


@pytest.mark.parametrize("gid", list(range(8)))
def test_noalloc_matches_batch_to_special_on_calls_in_a_row(engine, ref, gid):
    wide = gid in (3, 4, 5, 7)
    sizes = [(700, 0), (37, 5), (1, 2), (257, 9), (700, 0)] if wide else \
            [(5000, 0), (333, 7), (16, 1), (17, 3), (5000, 0)]
    for call, (n, start) in enumerate(sizes):
        pts = jacobian_points(ref, gid, n, start)
        want = engine.batch_to_special(gid, pts)  # the existing entry point
        got = noalloc(engine, gid, pts, inplace=bool(call & 1))
        np.testing.assert_array_equal(got, want, err_msg=f"call {call}: n={n}")
        if call == 0:
            np.testing.assert_array_equal(want, ref.batch_to_special(gid, pts))


@pytest.mark.parametrize("gid", [0, 3])
def test_noalloc_at_the_maximum_and_beyond(engine, ref, gid):
    nmax = engine.NOALLOC_MAX_POINTS
    pts = jacobian_points(ref, gid, nmax, 0)
    want = engine.batch_to_special(gid, pts)
    for call in range(2):  # two full-size calls in a row over the same scratch
        np.testing.assert_array_equal(noalloc(engine, gid, pts, inplace=bool(call)), want)
    # a small call right after a full one reads nothing of what the full one left behind
    np.testing.assert_array_equal(noalloc(engine, gid, pts[100:119], inplace=False),
                                  engine.batch_to_special(gid, pts[100:119]))
    gb = engine.group_bytes(gid)
    with pytest.raises(engine.B200MsmError, match="exceeds the static scratch"):
        engine.batch_to_special_noalloc_ptr(gid, 256, 256, (nmax + 1) * gb, nmax + 1, 0)
    with pytest.raises(engine.B200MsmError, match="does not match"):
        engine.batch_to_special_noalloc_ptr(gid, 256, 256, 5 * gb + 8, 5, 0)
    with pytest.raises(engine.B200MsmError, match="unknown group"):
        engine.batch_to_special_noalloc_ptr(99, 256, 256, 0, 0, 0)
    # n == 0 is a no-op
    engine.batch_to_special_noalloc_ptr(gid, 0, 0, 0, 0, 0)
