"""Multi-GPU plumbing: one process per GPU, samples (or rows) sharded over ranks, ONE reduce of
the float4 sum buffer to rank 0 per frame.

The collective itself lives in the C ABI (b200pt_comm_init_rank / b200pt_render_frame: ncclReduce inside libb200pt.so,
the same entry points `path_tracer --gpus N` uses).  torch.distributed only carries the 128-byte NCCL id from rank 0 to
the other ranks (`init_comm`).  `reduce_sum_to_rank0` on a torch tensor remains for the gloo CPU tests and for callers
that want the reduce on their own process group."""
from __future__ import annotations

import numpy as np


class _DevPtr:
    """Expose a raw device pointer through __cuda_array_interface__ so torch can wrap it."""

    def __init__(self, ptr, n_floats):
        self.__cuda_array_interface__ = dict(shape=(int(n_floats),), typestr="<f4", data=(int(ptr), False), version=3,
                                             strides=None)


def accum_tensor(ctx):
    """The context's float4 sum buffer as a torch CUDA tensor (no copy)."""
    import torch
    ptr, n = ctx.accum_buffer()
    return torch.as_tensor(_DevPtr(ptr, n), device=f"cuda:{torch.cuda.current_device()}")


def reduce_sum_to_rank0(tensor, group=None):
    """The frame's single collective.  In place; only rank 0's result is meaningful."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(tensor, dst=0, op=dist.ReduceOp.SUM, group=group)
    return tensor


def init_comm(ctx, group=None):
    """Create the library's NCCL communicator for this rank: rank 0 makes the id, torch.distributed ships it."""
    import torch.distributed as dist
    if ctx.world_size == 1:
        return
    box = [ctx.comm_unique_id() if ctx.rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    ctx.comm_init_rank(box[0])
    ctx._comm_ready = True


def render_frame(ctx, group=None):
    """render this rank's share -> reduce -> rank 0 resolves (scale by 1/total spp, RGBA8 pack).  After init_comm the
    whole sequence is one C call (b200pt_render_frame); otherwise the reduce runs on the torch process group."""
    if getattr(ctx, "_comm_ready", False) or ctx.world_size == 1:
        ctx.render_frame()
        return
    ctx.render()
    if ctx.world_size > 1:
        import torch
        t = accum_tensor(ctx)
        torch.cuda.current_stream().synchronize()
        reduce_sum_to_rank0(t, group)
        torch.cuda.current_stream().synchronize()
        if ctx.rank == 0:
            ctx.resolve()


def render_pass(ctx, n_samples, group=None):
    """One progressive pass on every rank (n_samples of each rank's stream) -> reduce of the copies -> rank 0 resolves
    the frame over all samples so far and returns the pass info with the combined frame's variance figure."""
    info = ctx.render_pass(n_samples)
    if ctx.world_size > 1:
        import torch
        t = accum_tensor(ctx)
        torch.cuda.current_stream().synchronize()
        reduce_sum_to_rank0(t, group)
        torch.cuda.current_stream().synchronize()
        if ctx.rank == 0:
            ctx.resolve()
            info["rel_variance"] = ctx.pass_stats()["rel_variance"]
    return info


def shard_samples(total_spp: int, rank: int, world: int) -> int:
    """Samples per pixel rank `rank` renders in SPP mode (same rule as api.cu rank_samples)."""
    return total_spp // world + (1 if rank < total_spp % world else 0)
