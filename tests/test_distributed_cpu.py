"""world_size-2 (and 3) runs of the multi-GPU host logic over the gloo backend on CPU: shard
ranges, the all_gather of the partial points and their rank-ordered sum (multiexp.tcc:655-687).
The device engine is replaced by the checker (oracle/port.py) through the injection points of
libff_b200.distributed -- only the plumbing is under test here; the kernels are covered by the
`-m gpu` tests."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port_no, gid, n, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no), RANK=str(rank),
                      WORLD_SIZE=str(world))
    import torch.distributed as dist
    from libff_b200 import distributed as D
    from oracle import port
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bases, scalars = port.gen_bases_walk(gid, n), port.gen_scalars(gid, n)

        def partial_fn(group, b, s, form):  # non-normalised partial, like the device engine's
            return port.multi_exp(group, b, s, form=form, normalize=False)

        def sum_fn(group, pts):
            zero, _ = port.group_constants(group)
            acc = zero.reshape(1, -1)
            for p in pts:
                acc = port.group_op(group, port.G_ADD, acc, p.reshape(1, -1))
            return acc[0]

        assert D.world_info() == (rank, world)
        got = D.multi_exp_sharded(gid, bases, scalars, partial_fn=partial_fn, sum_fn=sum_fn)
        got = port.batch_to_special(gid, got.reshape(1, -1))[0]
        want = port.multi_exp(gid, bases, scalars)
        parts = D.gather_partials(np.full(3, rank, dtype=np.uint64))
        ok = np.array_equal(got, want) and parts.tolist() == [[r] * 3 for r in range(world)]
        ret[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,gid,n", [(2, 0, 1001), (3, 2, 200), (2, 3, 7), (2, 1, 1)])
def test_sharded_multi_exp_over_gloo(world, gid, n):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port_no = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port_no, gid, n, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=240)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    assert [ret.get(r) for r in range(world)] == [True] * world
