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
    torch.save(dict(mine=mine, ms=ms, counts=counts, allf=allf), os.path.join(out_dir, "r%d.pt" % rank))
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


def test_sharding_single_process_and_ragged():
    from voxelnet_pytorch_b200 import sharding

    assert sharding.shard_frames(5, 0, 2) == [0, 2, 4] and sharding.shard_frames(5, 1, 2) == [1, 3]
    assert sharding.shard_frames(3, 3, 4) == [3][:0] or sharding.shard_frames(3, 3, 4) == []
    assert sharding.reduce_max([1.5, 2.5]) == [1.5, 2.5]
    assert sharding.gather_ints(7) == [7]
    with pytest.raises(ValueError):
        sharding.shard_frames(4, 2, 2)
