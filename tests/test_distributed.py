"""Multi-GPU host logic on the CPU: sample sharding, and the film / counter reduce over a world-size-2 gloo group."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pathtrace_b200.distributed import reduce_counts, reduce_film, shard_samples


def test_shard_samples_partitions():
    for total in (1, 2, 7, 64, 1024, 1025):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_samples(total, world, r) for r in range(world)]
            assert sum(c for _, c in spans) == total
            pos = 0
            for first, count in spans:
                assert first == pos
                pos += count
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1
    with pytest.raises(ValueError):
        shard_samples(8, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    total = 10
    first, count = shard_samples(total, world, rank)
    # stand-in film: every global sample s adds (s + 1) to channel 0 and 1 to channel 3 of every pixel
    film = torch.zeros(4, 6, 4)
    for s in range(first, first + count):
        film[..., 0] += s + 1
        film[..., 3] += 1
    reduce_film(film, 0)
    counts = reduce_counts([count, 100 * (rank + 1)], "cpu", 0)
    if rank == 0:
        np.save(os.path.join(out_dir, "film.npy"), film.numpy())
        np.save(os.path.join(out_dir, "counts.npy"), np.array(counts))
    dist.barrier()
    dist.destroy_process_group()


def test_reduce_film_gloo_world2(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    film = np.load(tmp_path / "film.npy")
    counts = np.load(tmp_path / "counts.npy")
    assert np.all(film[..., 0] == sum(range(1, 11))) and np.all(film[..., 3] == 10)
    assert counts.tolist() == [10, 300]


def test_render_sharded_raises_on_every_rank_when_samples_are_fewer_than_ranks():
    """with config.samples < world_size only the empty ranks used to raise and the others hung in the reduce.
    The check now comes before the sharding, so every rank raises -- and before any GPU work, so it can be tested here."""
    import pathtrace_b200 as pt
    from pathtrace_b200.distributed import render_sharded
    from conftest import scene_path
    scene = pt.Scene.from_file(scene_path("cornell_box"), 8, 8, device=-1)
    cfg = pt.Config.from_json(pt.default_config_dict(width=8, height=8, samples=1, scene=scene_path("cornell_box")))
    for rank in (0, 1):
        with pytest.raises(ValueError, match="fewer samples than ranks"):
            render_sharded(scene, cfg, rank, 2)
