"""Multi-GPU plumbing for the hot path: frames are independent units (no stage of the forward
couples frames), so N GPUs run N replicas on disjoint frames with NO data-path collective
(SURVEY.md section 8e).  The only cross-rank traffic is timing/bookkeeping: a MAX-reduction of
device-measured milliseconds and a gather of per-rank frame counts, over whatever backend the
process group uses (NCCL on the GPU box, gloo in the CPU tests).
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_frames(n_frames: int, rank: int, world: int) -> list:
    """Frame i of a global batch goes to rank i mod world (round-robin keeps ragged frame sizes
    balanced and is stable when world changes by a factor of two)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world %d" % (rank, world))
    return list(range(rank, n_frames, world))


def _device_for_backend() -> torch.device:
    if dist.get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_max(values: list) -> list:
    """Element-wise MAX over ranks of a short list of floats (device-timed milliseconds)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(v) for v in values]
    t = torch.tensor(values, dtype=torch.float64, device=_device_for_backend())
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(v) for v in t.cpu()]


def gather_ints(value: int) -> list:
    """Every rank's integer (e.g. frames or voxels processed) on every rank."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [int(value)]
    t = torch.tensor([int(value)], dtype=torch.int64, device=_device_for_backend())
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [int(o.item()) for o in out]


def gather_floats(value: float) -> list:
    """Every rank's float (e.g. its own device-timed ms per step) on every rank, in rank order."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [float(value)]
    t = torch.tensor([float(value)], dtype=torch.float64, device=_device_for_backend())
    out = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [float(o.item()) for o in out]


def allreduce_gradients(params, average: bool = True) -> int:
    """Data-parallel gradient exchange of the training config: the .grad tensors of
    `params` are packed into ONE flat fp32 buffer, summed over the ranks with a single all-reduce (NCCL over
    NVLink/NVSwitch on the GPU box, gloo in the CPU tests), divided by the world size and unpacked in place.
    Returns the number of gradient elements exchanged (0 without a process group or at world size 1)."""
    grads = [p.grad for p in params if p.grad is not None]
    if not grads or not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 0
    flat = torch.cat([g.reshape(-1).float() for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM)
    if average:
        flat.div_(dist.get_world_size())
    off = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[off:off + n].view_as(g))
        off += n
    return int(flat.numel())
