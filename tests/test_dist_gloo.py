"""N>1 host logic on CPU: world_size-2 gloo run of the sample split + the one reduce per frame.  The per-rank
compute is stood in for by the C oracle (a checker, allowed in tests) so that the reduced frame can be compared
with a single-rank frame of all samples."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sexy_raytracer_b200.dist import sample_split, reduce_frame

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sample_split_partitions():
    for spp in (0, 1, 7, 16, 5000):
        for world in (1, 2, 3, 4, 8):
            parts = [sample_split(spp, world, r) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == spp
            for (b0, c0), (b1, _) in zip(parts, parts[1:]):
                assert b1 == b0 + c0
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
    assert sample_split(5000, 8, 3) == (1875, 625)
    with pytest.raises(ValueError):
        sample_split(8, 2, 2)


def _worker(rank, world, port_no, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.refbind import PortOracle
    from sexy_raytracer_b200.scene import Scene
    s = Scene.load(os.path.join(ROOT, "scenes", "c2_cube.npz"))
    P = PortOracle()
    h = P.from_scene(s)
    begin, count = sample_split(6, world, rank)
    acc, st = P.render(h, s, s.render_params(width=64, height=36, spp=6, sample_begin=begin, sample_count=count, seed=8))
    t = torch.from_numpy(acc)
    reduce_frame(t, dst=0)
    seg = torch.tensor([int(st.segments)], dtype=torch.int64)
    dist.reduce(seg, dst=0)
    if rank == 0:
        full, st_full = P.render(h, s, s.render_params(width=64, height=36, spp=6, seed=8))
        np.save(out, np.array([np.abs(t.numpy() - full).max() / np.abs(full).max(), float(seg.item() - st_full.segments)]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_frame_equals_single_rank(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port_no = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "r.npy")
    mp.spawn(_worker, args=(2, port_no, out), nprocs=2, join=True)
    err, dseg = np.load(out)
    assert err <= 1e-5 and dseg == 0
