"""world_size-2 gloo test (CPU) of the sharded path's host logic: slab partition,
size all-gather + exclusive prefix, gather / scatter of the packed stream. Every
rank encodes its slab of segments with the CPU oracle (tests may use it as the
stand-in encoder); the gathered stream must equal the single-process stream and
each rank's scattered slab must decode to the oracle's floats (shard invariance,
SURVEY.md section 8d/8e)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

N_SEG = 5      # segments of the snapshot (odd: ragged split over 2 ranks)
SEG = 3000     # particles per segment (ragged last block)
NGRID = 32


def _segments():
    from _data import zeldovich_host
    out = []
    for s in range(N_SEG):
        pos, vel = zeldovich_host(NGRID, start=s * SEG, count=SEG)
        out.append((pos.reshape(-1), vel.reshape(-1)))
    return out


def _encode_slab(orc, segs, first, count):
    parts, offs = [], [0]
    for s in range(first, first + count):
        for code, data, delta, width in ((0x506F736E, segs[s][0], 1e-3, 100.0), (0x56656C63, segs[s][1], 1.0, 0.0)):
            b = orc.encode_field(code, data, delta, width, seed=s)
            parts.append(b)
            offs.append(offs[-1] + b.size)
    return (np.concatenate(parts) if parts else np.zeros(0, np.uint8)), np.array(offs, np.int64)


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import _oracle
        from minnow_c_b200 import sharded
        orc = _oracle.port()
        segs = _segments()
        slabs = sharded.partition(N_SEG, world)
        first, count = slabs[rank]
        local, offs = _encode_slab(orc, segs, first, count)
        t_local = torch.from_numpy(local.copy())
        sizes, base = sharded.exchange_sizes(torch.tensor([local.size], dtype=torch.int64))
        goffs = sharded.global_offsets(torch.from_numpy(offs), base)
        stream = sharded.gather_stream(t_local, sizes, dst=0)
        full, full_offs = _encode_slab(orc, segs, 0, N_SEG)
        ok = True
        if rank == 0:
            ok = ok and np.array_equal(stream.numpy(), full)
        # this rank's global offsets are the single-process offsets of its fields
        ok = ok and np.array_equal(goffs.numpy(), full_offs[2 * first:2 * (first + count) + 1])
        ok = ok and int(sizes.sum()) == full.size
        # scatter back and decode the slab
        back = sharded.scatter_stream(stream, sizes, src=0)
        ok = ok and np.array_equal(back.numpy(), local)
        for i in range(2 * count):
            payload = back.numpy()[offs[i]:offs[i + 1]]
            want = orc.decode_field(full[full_offs[2 * first + i]:full_offs[2 * first + i + 1]])
            got = orc.decode_field(payload)
            ok = ok and np.array_equal(got.view(np.uint32), want.view(np.uint32))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_partition_is_contiguous_and_complete():
    from minnow_c_b200 import sharded
    for n in (0, 1, 5, 256, 2048):
        for w in (1, 2, 3, 4, 8):
            slabs = sharded.partition(n, w)
            assert slabs[0][0] == 0 and sum(c for _, c in slabs) == n
            for (a, ca), (b, _) in zip(slabs, slabs[1:]):
                assert a + ca == b


@pytest.mark.timeout(180)
def test_sharded_stream_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=150) for _ in procs]
    for p in procs:
        p.join(timeout=30)
    assert sorted(results) == [(0, True), (1, True)]
