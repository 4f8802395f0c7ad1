"""CPU, world_size 2 over gloo: the N>1 host logic (frame sharding, max-over-ranks timing, count
gather) that bench.py uses under torchrun."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from voxelnet_pytorch_b200 import sharding

    mine = sharding.shard_frames(64, rank, world)
    ms = sharding.reduce_max([10.0 + rank, 5.0 - rank])
    counts = sharding.gather_ints(len(mine))
    allf = [None] * world
    dist.all_gather_object(allf, mine)
    # training config: per-rank gradients of ragged shapes, one of them absent -> one flat all-reduce, averaged
    params = [torch.nn.Parameter(torch.zeros(s)) for s in ((3, 4), (5,), (2, 2, 2))]
    params[0].grad = torch.full((3, 4), float(rank + 1))
    params[2].grad = torch.arange(8.0).view(2, 2, 2) * (rank + 1)
    n = sharding.allreduce_gradients(params)
    grads = [None if p.grad is None else p.grad.clone() for p in params]
    torch.save(dict(mine=mine, ms=ms, counts=counts, allf=allf, n=n, grads=grads), os.path.join(out_dir, "r%d.pt" % rank))
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_sharding_world2(tmp_path):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    r = [torch.load(os.path.join(tmp_path, "r%d.pt" % i)) for i in range(world)]
    assert r[0]["ms"] == r[1]["ms"] == [11.0, 5.0]          # MAX over ranks, per element
    assert r[0]["counts"] == r[1]["counts"] == [32, 32]       # stress config: 64 frames split evenly
    flat = sorted(r[0]["allf"][0] + r[0]["allf"][1])
    assert flat == list(range(64))                           # disjoint cover
    assert r[1]["mine"][:3] == [1, 3, 5]
    for rr in r:  # mean over the two ranks: (1 + 2) / 2
        assert rr["n"] == 20 and rr["grads"][1] is None
        assert torch.equal(rr["grads"][0], torch.full((3, 4), 1.5))
        assert torch.equal(rr["grads"][2], torch.arange(8.0).view(2, 2, 2) * 1.5)


def test_sharding_single_process_and_ragged():
    from voxelnet_pytorch_b200 import sharding

    assert sharding.shard_frames(5, 0, 2) == [0, 2, 4] and sharding.shard_frames(5, 1, 2) == [1, 3]
    assert sharding.shard_frames(3, 3, 4) == [3][:0] or sharding.shard_frames(3, 3, 4) == []
    assert sharding.reduce_max([1.5, 2.5]) == [1.5, 2.5]
    assert sharding.gather_ints(7) == [7]
    p = torch.nn.Parameter(torch.zeros(2))
    p.grad = torch.ones(2)
    assert sharding.allreduce_gradients([p]) == 0 and torch.equal(p.grad, torch.ones(2))  # no group: untouched
    with pytest.raises(ValueError):
        sharding.shard_frames(4, 2, 2)
