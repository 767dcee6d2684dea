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
