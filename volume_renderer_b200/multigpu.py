"""Image-space multi-GPU mode: one process per GPU, every rank holds the volumes, the image is cut
into tiles dealt round-robin to the ranks, partial images are combined on rank 0.

The reference has no multi-GPU path; its only "combine" step is image addition in
MATLAB (examples/example1.m:85).  Tile ownership lives in the kernel (vr_types.h: tile j belongs
to rank j % nranks); this module only configures it and runs the one exchange step.  Partial
images are zero outside a rank's tiles, so a SUM reduction to rank 0 reproduces the single-GPU
image bit for bit (x + 0 == x).

Two exchange steps are provided:
  * direct send (`SharedImage`, the default on NVLink): the root's image is mapped into every
    rank with CUDA IPC and each rank's ray-march kernel stores its tiles straight into the
    root's HBM over NVLink - compute and "gather" are one kernel, the only collective left is
    a barrier that tells the root the frame is complete;
  * `combine_partials`: every rank renders into its own zeroed image and the partials are
    SUM-reduced to the root with NCCL (also the gloo path of the CPU tests).
"""
import ctypes as C

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


def exchange_handle(handle, src=0, group=None):
    """Broadcast the 64-byte IPC handle of the root's image from rank `src` (any backend)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return bytes(handle)
    box = [bytes(handle) if dist.get_rank(group) == src else None]
    dist.broadcast_object_list(box, src=src, group=group)
    assert len(box[0]) == _abi.IPC_HANDLE_BYTES
    return box[0]


class SharedImage:
    """One H x W x 3 fp32 image in the root's HBM that every rank renders its tiles into.

    rank == root: allocates (vr_device_alloc) and exports the buffer; other ranks map it.
    `ptr` is the device pointer to pass as d_out_image of vr_render_device on this rank.
    """

    def __init__(self, ctx, H, W, rank, world, root=0, group=None, count=1):
        lib = _abi.lib()
        self.ctx, self.rank, self.root, self.world = ctx, rank, root, world
        self.frame_bytes = int(H) * int(W) * 3 * 4
        self.count = int(count)  # frames of a batch live back to back in the one allocation
        self.nbytes = self.frame_bytes * self.count
        self._owned = rank == root
        p = C.c_void_p()
        handle = C.create_string_buffer(_abi.IPC_HANDLE_BYTES)
        if self._owned:
            _abi.check(lib.vr_device_alloc(ctx, self.nbytes, C.byref(p)), ctx)
            if world > 1:
                _abi.check(lib.vr_ipc_export(ctx, p, handle), ctx)
        raw = exchange_handle(handle.raw, root, group) if world > 1 else handle.raw
        if not self._owned:
            _abi.check(lib.vr_ipc_open(ctx, raw, C.byref(p)), ctx)
        self.ptr = p
        self.ptrs = [C.c_void_p(p.value + i * self.frame_bytes) for i in range(self.count)]
        # several ranks write disjoint tiles of this one image: nobody may clear the whole of it
        _abi.check(lib.vr_set_option(ctx, _abi.OPT_CLEAR_IMAGE, 0 if world > 1 else 1), ctx)

    def download(self, host_array, frames=None):
        """Root only: copy the first `frames` finished frames (default all) to a host numpy array."""
        assert self._owned
        n = self.frame_bytes * (self.count if frames is None else int(frames))
        _abi.check(_abi.lib().vr_device_download(self.ctx, host_array.ctypes.data, self.ptr, n), self.ctx)
        return host_array

    def close(self):
        lib = _abi.lib()
        if self.ptr:
            if self._owned:
                _abi.check(lib.vr_device_free(self.ctx, self.ptr), self.ctx)
            else:
                _abi.check(lib.vr_ipc_close(self.ctx, self.ptr), self.ctx)
            self.ptr = None


def frame_barrier(stream_tensor=None, group=None):
    """End of a direct-send frame: every rank's kernel has finished storing into the root's image.
    With NCCL the all-reduce is enqueued on the current CUDA stream behind the ray-march kernel, so
    no host synchronisation is needed on the senders."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        if stream_tensor is not None:
            dist.all_reduce(stream_tensor, group=group)
        else:
            dist.barrier(group=group)
