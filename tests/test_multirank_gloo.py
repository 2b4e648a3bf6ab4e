"""N > 1 host-side logic on CPU (gloo, world_size 2 and 3): band ownership, the gather to rank 0 and the band
un-interleave of chaos-ray-tracing-course-2025_b200/distributed.py.  Pixels come from the oracle port here (there is no
GPU in this container); on the GPU box the same BandGather moves the CUDA framebuffers over NCCL (bench.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_tools as otools

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
needs_scenes = pytest.mark.skipif(not (otools.have_reference_build() or os.path.isdir("/root/reference/src")),
                                  reason="scene fixtures live in oracle/_ref/scenes")


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, scene_text, band_rows, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as entry
    import oracle_tools as ot
    crt = entry.load_package()
    from crt_b200 import distributed as D

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene = crt.Scene.load_json(scene_text)
        flat = scene.flatten()
        W, H = flat.camera.width, flat.camera.height
        params = crt.default_params(rank=rank, world=world, band_rows=band_rows)
        part = ot.port_render(flat, params, threads=2, primary=False)
        assert part["rgb8"].shape[0] == len(D.owned_rows(H, rank, world, band_rows))
        for bpp, buf in ((3, part["rgb8"]), (12, part["rgb"])):
            g = D.BandGather(W, H, band_rows, bpp, "cpu")
            frame = g.gather(torch.from_numpy(np.ascontiguousarray(buf).view(np.uint8).reshape(-1)))
            if rank == 0:
                np.save(out_path + f".{bpp}.npy", frame.numpy())
            else:
                assert frame is None
        # animation: frames are dealt round-robin and every frame is owned exactly once
        mine = D.owned_frames(10, rank, world)
        counts = torch.zeros(10, dtype=torch.int32)
        counts[mine] = 1
        dist.all_reduce(counts)
        assert (counts == 1).all()
    finally:
        dist.destroy_process_group()


@needs_scenes
@pytest.mark.parametrize("world,band_rows,size", [(2, 8, (96, 54)), (3, 5, (64, 37)), (2, 64, (48, 20))])
def test_band_gather_reassembles_the_frame(crt, ot, tmp_path, world, band_rows, size):
    text = ot.scene_variant("15-01-conclusion/scene1.crtscene", *size)
    out = str(tmp_path / "frame")
    mp.spawn(_worker, args=(world, _free_port(), text, band_rows, out), nprocs=world, join=True)
    scene = crt.Scene.load_json(text)
    full = ot.port_render(scene.flatten(), crt.default_params(), primary=False)
    got8 = np.load(out + ".3.npy").reshape(size[1], size[0], 3)
    gotf = np.load(out + ".12.npy").view(np.float32).reshape(size[1], size[0], 3)
    assert np.array_equal(got8, full["rgb8"])
    assert np.array_equal(ot.bits(gotf), ot.bits(full["rgb"]))


def test_owned_rows_partition(crt):
    from crt_b200 import distributed as D
    for height in (0, 1, 7, 8, 9, 1080, 2160):
        for world in (1, 2, 3, 4, 8):
            for band in (1, 5, 8, 24):
                rows = [D.owned_rows(height, r, world, band) for r in range(world)]
                allrows = np.sort(np.concatenate(rows)) if rows else np.zeros(0)
                assert np.array_equal(allrows, np.arange(height))
                for r in range(world):
                    assert np.array_equal(rows[r], crt.Renderer.owned_rows(height, r, world, band))
                    assert (np.diff(rows[r]) > 0).all()


def _shared_frame_worker(rank, world, port, scene_text, band_rows, out_path):
    """what bench.py's e2e leg does at N > 1, with the oracle port standing in for the GPU: every rank writes its row
    bands to their place in ONE frame in shared host memory (bench.SharedHostFrame), nothing is gathered"""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import __graft_entry__ as entry
    import bench
    import oracle_tools as ot
    crt = entry.load_package()
    from crt_b200 import distributed as D

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        flat = crt.Scene.load_json(scene_text).flatten()
        W, H = flat.camera.width, flat.camera.height
        name = [f"crt_test_{os.getpid()}"] if rank == 0 else [None]
        dist.broadcast_object_list(name, src=0)
        shared = bench.SharedHostFrame(name[0], W * H * 12, create=True) if rank == 0 else None
        dist.barrier()
        if rank != 0:
            shared = bench.SharedHostFrame(name[0], W * H * 12, create=False)
        rows = D.owned_rows(H, rank, world, band_rows)
        shared.touch([(int(r) * W * 12, int(min(H, r + band_rows)) * W * 12) for r in rows[::band_rows]])
        part = ot.port_render(flat, crt.default_params(rank=rank, world=world, band_rows=band_rows), threads=2, primary=False)
        frame = shared.array.view(np.float32).reshape(H, W, 3)
        frame[rows] = part["rgb"]  # (the GPU path: crt_cuda_render_frame, one strided copy per rank)
        dist.barrier()
        if rank == 0:
            np.save(out_path, frame.copy())
        del frame
        dist.barrier()
        shared.close()
    finally:
        dist.destroy_process_group()


@needs_scenes
@pytest.mark.parametrize("world,band_rows,size", [(2, 8, (96, 54)), (3, 8, (64, 37))])
def test_ranks_assemble_one_frame_in_shared_host_memory(crt, ot, tmp_path, world, band_rows, size):
    text = ot.scene_variant("15-01-conclusion/scene1.crtscene", *size)
    out = str(tmp_path / "frame.npy")
    mp.spawn(_shared_frame_worker, args=(world, _free_port(), text, band_rows, out), nprocs=world, join=True)
    full = ot.port_render(crt.Scene.load_json(text).flatten(), crt.default_params(), primary=False)
    assert np.array_equal(ot.bits(np.load(out)), ot.bits(full["rgb"]))
