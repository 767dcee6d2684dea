"""World-size-2 run of the image-space exchange step on CPU (gloo): each rank renders only the
tiles it owns (here: the oracle image masked by the kernel's ownership rule), partial images are
SUM-reduced to rank 0, and the result must be bit-identical to the single-rank image."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import vrt
from volume_renderer_b200 import multigpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, full, tile, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, W = full.shape[:2]
    mask = multigpu.owned_mask(H, W, tile, tile, rank, world)
    partial = torch.from_numpy(np.ascontiguousarray(full * mask[:, :, None]))
    multigpu.combine_partials(partial, 0)
    if rank == 0:
        np.save(out, partial.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_tile_split_reassembles_bit_identically(tmp_path):
    vol = vrt.synth_brain((24, 24, 24))
    full = np.ascontiguousarray(vrt.example1_scene(vrt.OracleVolumeRender, vol, (96, 80), lights=1).render())
    out = str(tmp_path / "combined.npy")
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), full, 16, out), nprocs=world, join=True)
    assert np.array_equal(np.load(out), full)


def _frames_worker(rank, world, port, full, tile, out):
    """The end-to-end exchange of bench.py on CPU: frames live in ONE host buffer shared by the ranks (POSIX shared memory,
    multigpu.SharedHostFrames); every rank writes exactly the tile COLUMNS it owns (VR_OPT_PART_MODE = 1) - on a GPU box its
    device does that with one strided copy per image plane."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, W = full.shape[:2]
    frames = multigpu.SharedHostFrames(H, W, 3, rank, world)  # no CUDA here: mapped, not page-locked
    if rank == 0:
        for i in range(3):
            frames.array(i)[...] = 7.0  # poisoned: every pixel must be written by exactly one rank
    dist.barrier()
    mask = multigpu.owned_mask(H, W, tile, tile, rank, world, mode=1)
    for i in range(3):
        frames.array(i)[mask] = (full * np.float32(i + 1))[mask]
    dist.barrier()
    if rank == 0:
        np.save(out, np.stack([np.array(frames.array(i)) for i in range(3)]))
    frames.close(world)
    dist.destroy_process_group()


def test_column_partition_covers_every_pixel_once_and_shared_frames_assemble(tmp_path):
    # the ownership rule itself: columns of tile_w pixels dealt round-robin, every pixel owned by exactly one rank, and the owned
    # pixels of a plane are runs of tile_w * H contiguous floats in the MATLAB (column-major) layout
    for W, H, tile, world in ((96, 80, 16, 2), (203, 97, 32, 3), (64, 64, 64, 8)):
        masks = [multigpu.owned_mask(H, W, tile, tile, r, world, mode=1) for r in range(world)]
        assert np.array_equal(np.sum(masks, axis=0), np.ones((H, W), dtype=int))
        for r, m in enumerate(masks):
            cols = np.where(m.any(axis=0))[0]
            assert m[:, cols].all() and all((c // tile) % world == r for c in cols)
    vol = vrt.synth_brain((24, 24, 24))
    full = np.asfortranarray(vrt.example1_scene(vrt.OracleVolumeRender, vol, (96, 80), lights=1).render())
    out = str(tmp_path / "frames.npy")
    mp.spawn(_frames_worker, args=(2, _free_port(), full, 16, out), nprocs=2, join=True)
    got = np.load(out)
    for i in range(3):
        assert np.array_equal(got[i], full * np.float32(i + 1))
