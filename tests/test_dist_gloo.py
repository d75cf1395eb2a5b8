"""CPU suite: the N>1 host logic with world_size-2 gloo — stripe assignment, all-gather of the RGB8 stripes into
the PPM writer's buffer, max all-reduce of the per-rank luminance (the auto-max-white variant).
Each rank's stripes come from the CPU checker (there is no GPU here); on the GPU box the same code paths run
in bench.py with NCCL and librte_b200.so."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from raytracingengine_b200 import scenes, sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, stripe, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pyoracle
        orc = pyoracle.load_port()
        arrays, cam = scenes.config3(scale=0.02)
        W, H = cam.width, cam.height
        rows = sharding.stripe_row_list(rank, world, H, stripe)
        counts = [sharding.stripe_row_count(r, world, H, stripe) for r in range(world)]
        rows_max = max(counts)
        # render my stripes (row by row through the checker), tonemap with ACES
        hdr = np.concatenate([orc.render(arrays, cam, row0=y, row1=y + 1)["hdr"] for y in rows]) if rows else np.zeros((0, 3))
        rgb = orc.tonemap_one(6, hdr)
        local = torch.zeros(rows_max * W * 3, dtype=torch.uint8)
        local[:rgb.size] = torch.from_numpy(rgb.reshape(-1))
        gathered = torch.empty(world * rows_max * W * 3, dtype=torch.uint8)
        dist.all_gather_into_tensor(gathered, local)
        lum = hdr @ np.array([0.2126, 0.7152, 0.0722]) if len(hdr) else np.zeros(1)
        mx = torch.tensor([float(np.nanmax(lum))], dtype=torch.float32)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        if rank == 0:
            frame = sharding.deinterleave(gathered.numpy().reshape(-1, 3), world, H, W, stripe, rows_max)
            np.save(os.path.join(out_dir, "frame.npy"), frame)
            np.save(os.path.join(out_dir, "maxlum.npy"), mx.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,stripe", [(2, 4), (2, 8)])
def test_striped_frame_reassembles_over_gloo(tmp_path, world, stripe):
    mp.spawn(_worker, args=(world, _free_port(), stripe, str(tmp_path)), nprocs=world, join=True)
    from oracle import pyoracle
    orc = pyoracle.load_port()
    arrays, cam = scenes.config3(scale=0.02)
    full = orc.render(arrays, cam)["hdr"]
    want = orc.tonemap_one(6, full)
    assert np.array_equal(np.load(tmp_path / "frame.npy"), want)
    lum = full @ np.array([0.2126, 0.7152, 0.0722])
    assert np.load(tmp_path / "maxlum.npy")[0] == np.float32(lum.max())
