"""World-size-2 `gloo` test of the N>1 host path: stream sharding by rank, packed result
tables gathered to rank 0, device time max-reduced -- the same module bench.py drives over NCCL."""
import os
import socket
import zlib

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rows(cam, frame, n):
    rng = np.random.default_rng(zlib.crc32(("%s/%d" % (cam, frame)).encode()))
    return rng.random((n, 7)).astype(np.float32)


CAMS = ["cam3", "cam1", "cam4", "cam2", "cam5"]
MAX_OUT = 16


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from telescope_cam_detection_b200 import stream_shard as ss

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = ss.my_streams(CAMS, rank, world)
    B = 3  # fixed table height per step: ranks with fewer cameras send empty frames
    like = torch.zeros((B, MAX_OUT + 1, 7), dtype=torch.float32)
    g = ss.ResultGather(world, rank, like)
    seen = {}
    for step in range(3):
        rows = [_rows(c, step, (rank + 2 * i + step) % (MAX_OUT + 1)) for i, c in enumerate(mine)]
        rows += [None] * (B - len(rows))
        t = torch.from_numpy(ss.pack_table(rows, MAX_OUT))
        got = g.gather(t)
        if rank == 0:
            for r, tab in enumerate(got):
                cams_r = ss.my_streams(CAMS, r, world)
                out = ss.unpack_table(tab.numpy())
                assert len(out) == B
                for i, c in enumerate(cams_r):
                    seen[(c, step)] = out[i]
                for i in range(len(cams_r), B):
                    assert out[i].shape == (0, 7)
    # ring of 4 steps in one collective (what bench.py does over NCCL every 8 steps)
    rg = ss.RingGather(world, 4, 5, "cpu")
    for step in range(6):
        half, i = rg.slot(step)
        assert (half, i) == ((step // 4) & 1, step % 4)
        rg.buf[half][i] = torch.full((5,), float(10 * rank + step))
        if i == 3 or step == 5:
            out = rg.collect(half)
            for r in range(world):
                n = 4 if step == 3 else 2
                exp = torch.tensor([10 * r + s for s in range(step - n + 1, step + 1)], dtype=torch.float32)
                assert torch.equal(out[r, :n, 0], exp), (rank, r, step, out[r, :, 0])
    ms = ss.max_over_ranks(1.0 + rank, world)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        q.put((sorted(k for k in seen), {k: v.tolist() for k, v in seen.items()}, ms))


def test_stream_assignment_is_a_partition():
    from telescope_cam_detection_b200 import stream_shard as ss

    for world in (1, 2, 3, 8):
        a = ss.assign_streams(CAMS, world)
        flat = [c for r in range(world) for c in a[r]]
        assert sorted(flat) == sorted(CAMS)
        assert max(len(v) for v in a.values()) - min(len(v) for v in a.values()) <= 1
        for c in CAMS:
            assert c in a[ss.owner_of(c, CAMS, world)]
    assert ss.assign_streams(CAMS, 2) == ss.assign_streams(list(reversed(CAMS)), 2)   # order-independent


def test_pack_unpack_round_trip_and_cap():
    from telescope_cam_detection_b200 import stream_shard as ss

    rows = [_rows("a", 0, 0), _rows("a", 1, 5), None, _rows("a", 2, MAX_OUT)]
    out = ss.unpack_table(ss.pack_table(rows, MAX_OUT))
    for r, o in zip(rows, out):
        assert o.shape == ((0, 7) if r is None else r.shape)
        if r is not None:
            assert np.array_equal(r, o)
    buf = np.full((4, MAX_OUT + 1, 7), 9.0, np.float32)     # reuse of a dirty pinned buffer
    assert np.array_equal(ss.pack_table(rows, MAX_OUT, out=buf), ss.pack_table(rows, MAX_OUT))
    with pytest.raises(ValueError):
        ss.pack_table([_rows("a", 0, MAX_OUT + 1)], MAX_OUT)
    assert ss.aggregate_fps(8, 10, 2, 20.0) == 8000.0


@pytest.mark.timeout(120)
def test_two_rank_gather_over_gloo():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    keys, seen, ms = q.get(timeout=100)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert keys == sorted((c, s) for c in CAMS for s in range(3))      # every stream of every rank, every step
    from telescope_cam_detection_b200 import stream_shard as ss
    for c in CAMS:
        r = ss.owner_of(c, CAMS, 2)
        i = ss.my_streams(CAMS, r, 2).index(c)
        for s in range(3):
            exp = _rows(c, s, (r + 2 * i + s) % (MAX_OUT + 1))
            assert np.array_equal(np.asarray(seen[(c, s)], np.float32).reshape(-1, 7), exp)
    assert ms == 2.0                                                    # max over ranks of (1 + rank)
