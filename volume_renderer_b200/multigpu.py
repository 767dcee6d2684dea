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


def owned_mask(H, W, tile_w, tile_h, rank, nranks, mode=0):
    """Boolean H x W mask of the pixels `rank` renders - the host twin of the kernel's
    work-item -> pixel mapping (raymarch_kernel.cu): tile index = ty * ntx + tx.
    mode 0: tiles dealt round-robin; mode 1 (VR_OPT_PART_MODE): tile COLUMNS dealt round-robin."""
    ntx, _ = tile_grid(W, H, tile_w, tile_h)
    ys, xs = np.mgrid[0:H, 0:W]
    if mode == 1:
        return ((xs // tile_w) % nranks) == rank
    tile = (ys // tile_h) * ntx + (xs // tile_w)
    return (tile % nranks) == rank


def configure_partition(ctx, rank, nranks, tile_w=64, tile_h=64, mode=0):
    lib = _abi.lib()
    for opt, val in ((_abi.OPT_TILE_W, tile_w), (_abi.OPT_TILE_H, tile_h), (_abi.OPT_PART_NRANKS, nranks), (_abi.OPT_PART_RANK, rank), (_abi.OPT_PART_MODE, mode)):
        _abi.check(lib.vr_set_option(ctx, opt, val), ctx)


class SharedHostFrames:
    """`count` H x W x 3 fp32 frames in ONE host buffer that every rank of the node maps (POSIX shared memory) and registers
    with CUDA (page-locked), so that each rank's device copies the pixels it rendered straight into the assembled frames
    over its own PCIe link (column partition, VR_OPT_PART_MODE = 1) - the read-back of a frame is spread over all GPUs.

        frames = SharedHostFrames(H, W, count, rank, world)      # collective: rank 0 creates, the others attach
        frames.ptrs[i]                                           # host pointer of frame i in THIS process
        frames.array(i)                                          # numpy view (H, W, 3), Fortran order
    """

    def __init__(self, H, W, count, rank, world, group=None, name=None):
        import torch
        import torch.distributed as dist
        from multiprocessing import shared_memory

        self.H, self.W, self.count, self.rank = int(H), int(W), int(count), rank
        self.frame_bytes = self.H * self.W * 3 * 4
        nbytes = self.frame_bytes * self.count
        box = [None]
        if rank == 0:
            self.shm = shared_memory.SharedMemory(create=True, size=nbytes, name=name)
            box[0] = self.shm.name
        if world > 1:
            dist.broadcast_object_list(box, src=0, group=group)
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=box[0])
            try:  # the creator unlinks; keep this process's resource tracker out of it
                from multiprocessing import resource_tracker

                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.buf = np.frombuffer(self.shm.buf, dtype=np.float32, count=nbytes // 4)
        self.base = self.buf.ctypes.data
        self._registered = False
        if torch.cuda.is_available():
            rc = torch.cuda.cudart().cudaHostRegister(self.base, nbytes, 0)
            if int(rc) != 0:
                raise RuntimeError("cudaHostRegister of the shared frames failed: %s" % rc)
            self._registered = True
        self.ptrs = [self.base + i * self.frame_bytes for i in range(self.count)]
        if world > 1:
            dist.barrier(group=group)

    def array(self, i):
        n = self.H * self.W * 3
        return self.buf[i * n:(i + 1) * n].reshape((self.H, self.W, 3), order="F")

    def close(self, world=1, group=None):
        import torch

        if self._registered:
            torch.cuda.cudart().cudaHostUnregister(self.base)
            self._registered = False
        self.buf = None
        if world > 1:
            import torch.distributed as dist

            dist.barrier(group=group)
        try:
            self.shm.close()
        except BufferError:
            pass
        if self.rank == 0:
            self.shm.unlink()


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


# ---- volume-space mode (sort-last) -------------------------------------------------------------
def brick_grid(dims, splits):
    """Cut a (w, h, d) volume into splits[0] x splits[1] x splits[2] voxel boxes.
    Returns [(lo, hi), ...] with lo/hi 3-tuples, brick index = (bz * sy + by) * sx + bx."""
    edges = [[(n * i) // s for i in range(s + 1)] for n, s in zip(dims, splits)]
    out = []
    for bz in range(splits[2]):
        for by in range(splits[1]):
            for bx in range(splits[0]):
                out.append(((edges[0][bx], edges[1][by], edges[2][bz]), (edges[0][bx + 1], edges[1][by + 1], edges[2][bz + 1])))
    return out


def splits_for(n):
    """Brick grid for n ranks: powers of two are split octant-wise (8 -> 2x2x2), anything else along z."""
    s = [1, 1, 1]
    axis = 2
    while n > 1 and n % 2 == 0:
        s[axis] *= 2
        n //= 2
        axis = (axis - 1) % 3
    s[2] *= n
    return tuple(s)


def boxes_array(bricks):
    import numpy as np

    return np.array([list(lo) + list(hi) for lo, hi in bricks], dtype=np.uint64).reshape(-1)


def compose(ctx, partial_ptrs, n_floats, d_out, stream=None):
    """vr_compose_partials_device: out = sum of the bricks' shares, in brick order."""
    arr = (C.c_void_p * len(partial_ptrs))(*[int(p) for p in partial_ptrs])
    _abi.check(_abi.lib().vr_compose_partials_device(ctx, len(partial_ptrs), arr, int(n_floats), C.c_void_p(int(d_out)), C.c_void_p(stream or 0)), ctx)


def render_bricks_one_gpu(renderers, bricks):
    """A whole sort-last frame with every brick on THIS GPU (virtual shards; tests and single-GPU
    runs of volumes cut into bricks).  `renderers[b]` is an api.VolumeRender that called
    setRegion(*bricks[b]) before its volumes were assigned.  Returns the H x W x 3 image."""
    import numpy as np
    import torch

    W, H = int(renderers[0].ImageResolution[0]), int(renderers[0].ImageResolution[1])
    n = len(bricks)
    opac = torch.zeros(n * H * W, dtype=torch.float32, device="cuda")
    part = torch.zeros(n * H * W * 3, dtype=torch.float32, device="cuda")
    out = torch.zeros(H * W * 3, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    for b, r in enumerate(renderers):
        r.renderOpacity(opac.data_ptr() + 4 * b * H * W)
    for r in renderers:
        r.stats()  # synchronise: every brick's opacity plane is complete ("all-gather")
    for b, r in enumerate(renderers):
        r.renderPartial(bricks, b, opac.data_ptr(), part.data_ptr() + 4 * b * H * W * 3)
    for r in renderers:
        r.stats()
    compose(renderers[0].objectHandle, [part.data_ptr() + 4 * b * H * W * 3 for b in range(n)], H * W * 3, out.data_ptr())
    renderers[0].stats()
    torch.cuda.synchronize()
    return out.cpu().numpy().reshape((H, W, 3), order="F")


class VolumeSpaceRank:
    """One process = one GPU = one brick of a sort-last render (torch.distributed, NCCL or gloo-free:
    the exchanges are an all-gather of opacity planes and a gather of the shares).

        vs = VolumeSpaceRank(renderer, bricks, rank)     # renderer.setRegion(*bricks[rank]) already done
        img = vs.render()                                # H x W x 3 numpy on rank 0, None elsewhere
    """

    def __init__(self, renderer, bricks, rank, group=None, exchange="p2p"):
        import torch

        self.r, self.bricks, self.rank, self.group = renderer, bricks, rank, group
        self.n = len(bricks)
        W, H = int(renderer.ImageResolution[0]), int(renderer.ImageResolution[1])
        self.H, self.W = H, W
        dev = torch.device("cuda", torch.cuda.current_device())
        # the context works on torch's current stream (handle 0 = the legacy default stream), where the
        # collectives are enqueued too: passing 0 per call would mean "the context's own stream"
        _abi.check(_abi.lib().vr_set_stream(renderer.objectHandle, C.c_void_p(torch.cuda.current_stream().cuda_stream)), renderer.objectHandle)
        self.opac_mine = torch.zeros(H * W, dtype=torch.float32, device=dev)
        self.opac_all = torch.zeros(self.n * H * W, dtype=torch.float32, device=dev)
        self.out = torch.zeros(H * W * 3, dtype=torch.float32, device=dev) if rank == 0 else None
        self.exchange = exchange if self.n > 1 else "local"
        if self.exchange == "p2p":
            # the n shares live back to back in rank 0's HBM; rank b's shading kernel stores slot b over NVLink
            self.shared = SharedImage(renderer.objectHandle, H, W, rank, self.n, root=0, group=group, count=self.n)
            renderer.setOption(_abi.OPT_CLEAR_IMAGE, 0)
            self.token = torch.zeros(1, dtype=torch.float32, device=dev)
        else:
            self.shared = None
            self.part = torch.zeros(H * W * 3, dtype=torch.float32, device=dev)

    def _mark(self, name):
        """Phase timing (CUDA events on the stream the passes and collectives run on); read with phase_ms()."""
        import torch

        if self.timing is None:
            return
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self.timing.append((name, ev))

    def phase_ms(self):
        """{phase: total ms} over the frames rendered since timing was switched on (self.timing = [])."""
        import torch

        torch.cuda.synchronize()
        out = {}
        for (n0, e0), (n1, e1) in zip(self.timing[:-1], self.timing[1:]):
            if n1 != "start":
                out[n1] = out.get(n1, 0.0) + e0.elapsed_time(e1)
        return out

    timing = None

    def render(self, download=True):
        import torch
        import torch.distributed as dist

        stream = None  # = the context's stream = torch's current stream (set in __init__)
        packed = self.r._packed_mono_args()  # one 'sync_volumes' + one marshalling per frame, shared by both passes
        self._mark("start")
        self.r.renderOpacity(self.opac_mine.data_ptr(), stream, packed)
        self._mark("opacity_pass")
        if self.n > 1:
            dist.all_gather_into_tensor(self.opac_all, self.opac_mine, group=self.group)
        else:
            self.opac_all.copy_(self.opac_mine)
        self._mark("all_gather")
        if self.shared is not None:
            self.r.renderPartial(self.bricks, self.rank, self.opac_all.data_ptr(), self.shared.ptrs[self.rank].value, stream, packed)
            self._mark("shading_pass")
            frame_barrier(self.token, self.group)
            if self.rank == 0:
                compose(self.r.objectHandle, [p.value for p in self.shared.ptrs], self.H * self.W * 3, self.out.data_ptr(), stream)
            self._mark("barrier_and_compose")
        else:
            self.r.renderPartial(self.bricks, self.rank, self.opac_all.data_ptr(), self.part.data_ptr(), stream, packed)
            self._mark("shading_pass")
            if self.n > 1:
                dist.reduce(self.part, dst=0, op=dist.ReduceOp.SUM, group=self.group)  # the sum IS the composite
            if self.rank == 0:
                self.out.copy_(self.part)
            self._mark("barrier_and_compose")
        if self.rank == 0 and download:
            torch.cuda.synchronize()
            return self.out.cpu().numpy().reshape((self.H, self.W, 3), order="F")
        return None

    def close(self):
        if self.shared is not None:
            self.shared.close()
            self.shared = None
