"""CPU tests of the multi-GPU host logic with world_size-2 gloo processes: camera sharding covers
every camera exactly once and the flat gradient buffer is summed across ranks in place."""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

from d3ga_b200 import dist as ddist


def test_shard_cameras_partition():
    for world in (1, 2, 4, 8):
        seen = sorted(sum((ddist.shard_cameras(160, r, world) for r in range(world)), []))
        assert seen == list(range(160))
        sizes = [len(ddist.shard_cameras(160, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1


def test_flat_grads_are_views():
    a = torch.zeros(5, 3, requires_grad=True)
    b = torch.zeros(7, requires_grad=True)
    fg = ddist.FlatGrads([a, b])
    ((a * 2).sum() + (b * 3).sum()).backward()
    ((a * 2).sum() + (b * 3).sum()).backward()      # accumulates in place, still inside the flat buffer
    assert a.grad.data_ptr() == fg.flat.data_ptr() and torch.all(fg.flat[:15] == 4) and torch.all(fg.flat[15:] == 6)
    fg.zero_()
    assert float(a.grad.abs().sum()) == 0.0


def _worker(rank, world, port, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    r, _, w = ddist.init_from_env(backend="gloo")
    p = torch.zeros(10, requires_grad=True)
    q = torch.zeros(4, 2, requires_grad=True)
    fg = ddist.FlatGrads([p, q])
    cams = ddist.shard_cameras(6, r, w)
    for c in cams:   # "views" whose gradient is the camera id + 1
        ((p * float(c + 1)).sum() + (q * float(c + 1)).sum()).backward()
    fg.all_reduce()
    ok = bool(torch.all(fg.flat == float(sum(range(1, 7)))))
    # the chunked lane-sum + all-reduce (training.PipelinedViews.finish): the same bits as add-everything-then-one-call
    g = torch.Generator().manual_seed(r)
    big = torch.zeros(300_000, requires_grad=True)
    fb = ddist.FlatGrads([big])
    lanes = [torch.randn(300_000, generator=g) for _ in range(3)]
    ref = lanes[0] + lanes[1]
    ref += lanes[2]
    torch.distributed.all_reduce(ref)
    fb.sum_lanes_and_ranks(lanes, chunks=4)
    ok = ok and bool(torch.equal(fb.flat, ref)) and big.grad.data_ptr() == fb.flat.data_ptr()
    fb.sum_lanes_and_ranks(lanes[:1], chunks=3)
    one = lanes[0].clone()
    torch.distributed.all_reduce(one)
    ok = ok and bool(torch.equal(fb.flat, one))
    mx = ddist.max_over_ranks(float(r), torch.device("cpu"))
    ddist.barrier()
    out.put((r, ok, mx, len(cams)))
    torch.distributed.destroy_process_group()


@pytest.mark.timeout(120)
def test_two_rank_gloo_allreduce():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=90) for _ in procs)
    for p in procs:
        p.join(30)
    assert [r[1] for r in res] == [True, True]
    assert [r[2] for r in res] == [1.0, 1.0]
    assert [r[3] for r in res] == [3, 3]
