"""Image-space direct send on the GPU: several ranks render disjoint tiles into ONE image.

(a) in one process: two contexts, one shared device buffer, no clear -> bit-identical to one GPU;
(b) across two processes on cuda:0 through the CUDA IPC entry points of the C ABI (the same
    plumbing bench.py uses over NVLink with one process per GPU), handle exchanged over gloo.
No kernel waits on another kernel: ranks render one after the other or independently."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

import vrt
from volume_renderer_b200 import _abi, api, multigpu

pytestmark = pytest.mark.gpu
RES = (192, 160)  # ImageResolution = [W, H]


def _scene(vol, lut):
    return vrt.example1_scene(api.VolumeRender, vol, RES, lights=1, lut=lut)


def _render_into(r, ptr, rank, world, tile=32):
    """p_render's argument packing (VolumeRender.m:497-583) but through vr_render_device into `ptr`."""
    r.syncVolumes()
    multigpu.configure_partition(r.objectHandle, rank, world, tile, tile)
    r.setOption(_abi.OPT_CLEAR_IMAGE, 0)
    args, keep = api._render_args(r.LightSources, r.VolumeIllumination, [r.FactorEmission, r.FactorReflection, r.FactorAbsorption],
                                  r.ElementSizeUm, [RES[1], RES[0]], np.flipud(np.asarray(r.RotationMatrix)),
                                  [0.0, r.FocalLength, r.DistanceToObject], r.OpacityThreshold, r.Color)
    _abi.check(_abi.lib().vr_render_device(r.objectHandle, *args, ptr, None), r.objectHandle)
    r.stats()  # synchronises the context's stream


def test_two_contexts_one_image_no_clear():
    import torch

    vol = vrt.synth_brain((48, 48, 48))
    lut = api.HenyeyGreenstein(64)
    full = _scene(vol, lut).render()
    # poisoned buffer: every pixel must be written by exactly one rank, misses as 0
    buf = torch.full((full.size,), 7.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    ranks = [_scene(vol, lut) for _ in range(3)]
    for rank, r in enumerate(ranks):
        _render_into(r, C.c_void_p(buf.data_ptr()), rank, 3)
    out = buf.cpu().numpy().reshape(full.shape, order="F")
    assert np.array_equal(out, full)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_path):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    vol = vrt.synth_brain((48, 48, 48))
    lut = vrt.oracle_lut(64)
    r = _scene(vol, lut)
    shared = multigpu.SharedImage(r.objectHandle, RES[1], RES[0], rank, world, root=0)
    for turn in range(world):  # one rank at a time: nothing runs concurrently on the one GPU
        if turn == rank:
            _render_into(r, shared.ptr, rank, world)
        dist.barrier()
    if rank == 0:
        out = np.zeros((RES[1], RES[0], 3), dtype=np.float32, order="F")
        shared.download(out)
        np.save(out_path, out)
    dist.barrier()
    if rank != 0:
        shared.close()
    dist.barrier()
    if rank == 0:
        shared.close()
    dist.destroy_process_group()


def test_two_processes_share_the_root_image_through_cuda_ipc(tmp_path):
    import torch.multiprocessing as mp

    vol = vrt.synth_brain((48, 48, 48))
    full = _scene(vol, vrt.oracle_lut(64)).render()
    out = str(tmp_path / "shared.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    assert np.array_equal(np.load(out), full)


def _vs_worker(rank, world, port, out_path):
    """One process per GPU: sort-last render of a volume cut into `world` bricks, both exchanges."""
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    vol = vrt.synth_brain((64, 56, 72))
    lut = vrt.oracle_lut(64)
    bricks = multigpu.brick_grid(vol.shape, multigpu.splits_for(world))
    imgs = {}
    # both exchanges with the load path forced (bricks then fetch the very bits of a single pass), and the default - the
    # texture unit on the brick's own arrays - once
    for exchange, tu in (("p2p", 0), ("nccl", 0), ("p2p", 2)):
        r = api.VolumeRender()
        r.setRegion(*bricks[rank])
        r = vrt.example1_scene(lambda: r, vol, RES, lights=2, lut=lut)
        r.setOption(_abi.OPT_TEXTURE_UNIT, tu)
        r.FactorAbsorption = 30.0
        vs = multigpu.VolumeSpaceRank(r, bricks, rank, exchange=exchange)
        for _ in range(2):  # a second frame reuses every buffer
            r.rotate(0, 10, 0)
            img = vs.render()
        assert r.stats()["fetch_mode"] == (1 if tu else 0)
        if rank == 0:
            imgs["%s_tu%d" % (exchange, tu)] = img
        torch.cuda.synchronize()
        dist.barrier()
        if rank != 0:
            vs.close()
        dist.barrier()
        if rank == 0:
            vs.close()
    if rank == 0:
        full = {}
        for tu in (0, 2):
            whole = vrt.example1_scene(api.VolumeRender, vol, RES, lights=2, lut=lut)
            whole.setOption(_abi.OPT_TEXTURE_UNIT, tu)  # compare like with like
            whole.FactorAbsorption = 30.0
            whole.rotate(0, 20, 0)
            full["full_tu%d" % tu] = whole.render()
        np.savez(out_path, **full, **imgs)
    dist.barrier()
    dist.destroy_process_group()


def test_volume_space_one_brick_per_gpu(tmp_path):
    import torch
    import torch.multiprocessing as mp

    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least 2 GPUs (gpurun --gpus 2)")
    out = str(tmp_path / "vs.npz")
    mp.spawn(_vs_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    z = np.load(out)
    for k in ("p2p_tu0", "nccl_tu0"):
        m = vrt.compare(z[k], z["full_tu0"])
        print(k, m)
        assert m["max_abs_norm"] < 2e-5, (k, m)
    # through the texture unit a brick's samples meet the filter's 8-bit weights a hair differently than the single pass's
    # (remapped coordinate); at a rare pixel that moves the threshold break by a sample - isolated pixels within the bar
    m = vrt.compare(z["p2p_tu2"], z["full_tu2"])
    print("p2p_tu2", m)
    assert m["max_abs_norm"] <= vrt.TOL and m["psnr_db"] >= 95.0, m


def _e2e_worker(rank, world, port, out_path):
    """bench.py's end-to-end leg in small: every rank renders its tile COLUMNS of a movie in pipelined batches and its device
    copies exactly those pixels into frames that live in ONE host buffer shared by the ranks (POSIX shared memory, page-locked
    in every process)."""
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    vol = vrt.synth_brain((48, 48, 48))
    lut = vrt.oracle_lut(64)
    r = api.VolumeRender()
    r = vrt.example1_scene(lambda: r, vol, RES, lights=2, lut=lut)
    frames = multigpu.SharedHostFrames(RES[1], RES[0], 5, rank, world)
    for i in range(5):
        if rank == 0:
            frames.array(i)[...] = 7.0  # poisoned: every pixel must be written by exactly one rank
    dist.barrier()
    multigpu.configure_partition(r.objectHandle, rank, world, 32, 32, mode=1)
    rots = []
    for _ in range(5):
        r.rotate(0, 13, 0)
        rots.append(r.RotationMatrix.copy())
    r.renderMoviePipelined(rots, batch=2, out=[frames.array(i) for i in range(5)])
    dist.barrier()  # every rank's columns have landed
    if rank == 0:
        np.save(out_path, np.stack([np.array(frames.array(i)) for i in range(5)]))
    frames.close(world)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_ranks_copy_their_own_columns_into_shared_host_frames(tmp_path, world):
    import torch.multiprocessing as mp

    vol = vrt.synth_brain((48, 48, 48))
    single = _scene(vol, vrt.oracle_lut(64))
    single.LightSources = [api.LightSource([500, 1000, 550], [0, 1, 1]), api.LightSource([550, 1000, 500], [1, 0, 1])]
    want = []
    for _ in range(5):
        single.rotate(0, 13, 0)
        want.append(single.render())
    out = str(tmp_path / "frames.npy")
    mp.spawn(_e2e_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    got = np.load(out)
    for k in range(5):
        assert np.array_equal(got[k], want[k]), (k, np.abs(got[k] - want[k]).max())
