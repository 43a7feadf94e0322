"""Multi-GPU host logic: one process per GPU, `torch.distributed` for the plumbing.

* volpath render: pixel samples are independent.  Every rank renders its own block of sample indices
  of every pixel into a private film; ONE reduce of the film (sum, fp64) to rank 0 ends the step.
  Under the IndependentSampler the union of the blocks is exactly the 1-GPU sample set, so the result
  equals the 1-GPU film up to fp64 summation order.  No collective on the data path.
* graph propagation (LightingCalculator::ComputeFinalLight, graph/lighting_calculator.cpp:23-59):
  1-D vertex-range partition of the CSR rows; per bounce every rank multiplies its rows with the full
  `cur` vector and the slices are exchanged with ONE all-gather (N * 4 bytes); the NaN/Inf early-out of
  the reference is agreed on with a 1-int all-reduce(max).

The functions take the compute step as a callable so the same code drives the CUDA library (NCCL,
device tensors) and the CPU checker in the world_size-2 `gloo` tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def init_comm(ctx, group=None):
    """The library's own communicator (avr_comm, NCCL behind the C ABI) for this rank of the torch.distributed job:
    rank 0 creates the unique id, torch.distributed only carries its 128 bytes to the other ranks."""
    from .api import Comm, comm_unique_id
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    box = [comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    return Comm(ctx, box[0], rank, world)


def shard_samples(render_desc, rank: int, world: int, spp_per_rank: int):
    """Weak scaling by sample index: rank r owns samples [r*spp, (r+1)*spp) of a (world*spp)-sample render."""
    if not 0 <= rank < world:
        raise ValueError("rank outside [0, world)")
    return render_desc.with_(spp=spp_per_rank * world, sample_begin=rank * spp_per_rank,
                             sample_end=(rank + 1) * spp_per_rank)


def split_samples(render_desc, rank: int, world: int):
    """Strong scaling: the spp of `render_desc` are split into `world` contiguous blocks (sizes differ by <= 1)."""
    spp = render_desc.spp
    base, extra = divmod(spp, world)
    begin = rank * base + min(rank, extra)
    end = begin + base + (1 if rank < extra else 0)
    if end == begin:
        return None  # more ranks than samples: this rank has nothing to do
    return render_desc.with_(sample_begin=begin, sample_end=end)


def reduce_film(film: torch.Tensor, dst: int = 0, group=None) -> torch.Tensor:
    """Sum the per-rank films (H, W, 4) fp64 into rank `dst` (ncclReduce on device tensors, gloo on CPU)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return film


def row_partition(n: int, world: int):
    """Equal row chunks (the last ranks' chunks are padded): chunk size and [r0, r1) per rank."""
    chunk = (n + world - 1) // world
    return chunk, [(min(r * chunk, n), min((r + 1) * chunk, n)) for r in range(world)]


def propagate_sharded(spmv_rows, L0: torch.Tensor, bounces: int, rank: int, world: int, group=None):
    """total = L0 + T L0 + ... + T^bounces L0 with the rows of T split over the ranks.

    spmv_rows(cur, r0, r1) -> tensor of r1-r0 floats = (T @ cur)[r0:r1], summed per row in column order
    (the order Eigen's row-major product uses, so slices are bit-identical to the 1-GPU result).
    Returns (total, iterations) on every rank; `iterations` is the reference's curIteration.
    """
    n = L0.numel()
    chunk, parts = row_partition(n, world)
    r0, r1 = parts[rank]
    cur = L0.clone()
    total = L0.clone()
    padded = torch.zeros(chunk * world, dtype=L0.dtype, device=L0.device)
    mine = torch.zeros(chunk, dtype=L0.dtype, device=L0.device)
    flag = torch.zeros(1, dtype=torch.int32, device=L0.device)
    it = 0
    for it in range(bounces + 1):
        if it == bounces:
            break
        out = spmv_rows(cur, r0, r1)
        mine.zero_()
        mine[: r1 - r0] = out
        flag[0] = 0 if bool(torch.isfinite(out).all()) else 1
        if world > 1:
            dist.all_gather_into_tensor(padded, mine, group=group)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=group)
            # padded holds rank-major chunks; rows of rank r start at r*chunk
            for r, (a, b) in enumerate(parts):
                cur[a:b] = padded[r * chunk: r * chunk + (b - a)]
        else:
            cur[r0:r1] = out
        if int(flag[0]) != 0:
            break  # reference: invalid light after this bounce -> keep the total of the bounces before
        total += cur
    return total, it
