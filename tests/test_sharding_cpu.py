"""The N>1 path on CPU: two gloo ranks shard the samples of a frame by index, render their shares (the CPU oracle stands in
for the GPU kernels here), combine the sum buffers with the frame's single reduce and resolve — equal to the 1-rank frame."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_samples_partition():
    from gpuspectral_b200.sharding import shard_samples
    for world in (1, 2, 3, 4, 8):
        for first, count in [(0, 16), (5, 7), (100, 1), (3, 0), (0, 1024)]:
            seen = []
            for r in range(world):
                f, n, st = shard_samples(first, count, r, world)
                seen += [f + k * st for k in range(n)]
            assert sorted(seen) == list(range(first, first + count))
    with pytest.raises(ValueError):
        shard_samples(0, 4, 2, 2)


def test_resolve_keeps_nan_skip_semantics():
    from gpuspectral_b200.sharding import resolve
    acc = np.zeros((2, 2, 4), np.float32)
    acc[0, 0] = [4, 8, 12, 4]
    acc[0, 1] = [3, 3, 3, 3]      # one of 4 samples was NaN and skipped: mean over the 3 that were added
    img = resolve(acc)
    assert np.allclose(img[0, 0], [1, 2, 3, 1]) and np.allclose(img[0, 1], [1, 1, 1, 1]) and np.allclose(img[1, 1], [0, 0, 0, 1])


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import gpuspectral_b200 as g
    from gpuspectral_b200.sharding import reduce_frame, resolve, shard_samples
    from oracle import oracle as orc
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = g.Scene.builtin("cornell")
    o = orc.OracleScene.from_host_scene(s)
    o.set_params(max_depth=6)
    w, h, spp = 40, 24, 7
    total = torch.zeros(h, w, 4)
    for first in (0, spp):                      # two progressive frames
        f, n, st = shard_samples(first, spp, rank, world)
        part, _, _ = o.render(w, h, f, n, st, sum_mode=True, threads=2)
        acc = torch.from_numpy(part)
        if rank == 0:
            acc += total
        reduce_frame(acc, 0)
        if rank == 0:
            total = acc.clone()
    if rank == 0:
        np.save(out_path, resolve(total.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_render_equals_single_rank(tmp_path, orc):
    import torch.multiprocessing as mp
    out = str(tmp_path / "img.npy")
    port = 29600 + os.getpid() % 300
    mp.start_processes(_worker, args=(2, port, out), nprocs=2, join=True, start_method="spawn")
    got = np.load(out)
    import gpuspectral_b200 as g
    s = g.Scene.builtin("cornell")
    o = orc.OracleScene.from_host_scene(s)
    o.set_params(max_depth=6)
    ref, _, _ = o.render(40, 24, 0, 14, sum_mode=True)
    from gpuspectral_b200.sharding import resolve
    assert np.allclose(got, resolve(ref), rtol=2e-5, atol=1e-6)
