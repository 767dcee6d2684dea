"""Image-space multi-GPU mode: one process per GPU, every rank holds the volumes, the image is cut
into tiles dealt round-robin to the ranks, partial images are combined on rank 0.

The reference has no multi-GPU path; its only "combine" step is image addition in
MATLAB (examples/example1.m:85).  Tile ownership lives in the kernel (vr_types.h: tile j belongs
to rank j % nranks); this module only configures it and runs the one exchange step.  Partial
images are zero outside a rank's tiles, so a SUM reduction to rank 0 reproduces the single-GPU
image bit for bit (x + 0 == x).
"""
import numpy as np

from . import _abi


def tile_grid(W, H, tile_w, tile_h):
    return (W + tile_w - 1) // tile_w, (H + tile_h - 1) // tile_h


def owned_mask(H, W, tile_w, tile_h, rank, nranks):
    """Boolean H x W mask of the pixels `rank` renders - the host twin of the kernel's
    work-item -> pixel mapping (raymarch_kernel.cu): tile index = ty * ntx + tx."""
    ntx, _ = tile_grid(W, H, tile_w, tile_h)
    ys, xs = np.mgrid[0:H, 0:W]
    tile = (ys // tile_h) * ntx + (xs // tile_w)
    return (tile % nranks) == rank


def configure_partition(ctx, rank, nranks, tile_w=64, tile_h=64):
    lib = _abi.lib()
    for opt, val in ((_abi.OPT_TILE_W, tile_w), (_abi.OPT_TILE_H, tile_h), (_abi.OPT_PART_NRANKS, nranks), (_abi.OPT_PART_RANK, rank)):
        _abi.check(lib.vr_set_option(ctx, opt, val), ctx)


def combine_partials(partial, dst=0, group=None):
    """The exchange step: SUM-reduce the partial images (torch tensors, CUDA over NCCL or CPU over
    gloo) onto rank `dst`.  Returns the tensor (complete on `dst` only)."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(partial, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return partial
