"""ShardedRenderer on the GPU: the render, the frame's reduce and the buffer reset share ONE explicit CUDA stream
(torch's default-stream handle 0 would make gsb_set_stream create a private stream and un-order the reduce).
1 rank: equal to a plain Context render.  2 ranks over NCCL (needs 2 visible GPUs): the reduced frame equals the 1-rank frame."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, SPP = 320, 180, 12


def test_sharded_renderer_single_rank_equals_context(gsb):
    from gpuspectral_b200.sharding import ShardedRenderer
    s = gsb.Scene.builtin("cornell")
    r = ShardedRenderer(s, W, H, 0, params={"max_depth": 8})
    assert r.stream.cuda_stream != 0
    r.render(SPP // 2)
    r.render(SPP - SPP // 2)
    got = r.read_hdr().copy()
    r.close()
    with gsb.Context(0) as c:
        c.upload(s)
        c.set_params(max_depth=8)
        c.render(W, H, 0, 0); c.clear(); c.render(W, H, 0, SPP)
        want = c.read_hdr()
    assert np.allclose(got, want, rtol=2e-5, atol=1e-6)


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist

    import gpuspectral_b200 as g
    from gpuspectral_b200.sharding import ShardedRenderer
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    r = ShardedRenderer(g.Scene.builtin("cornell"), W, H, rank, params={"max_depth": 8})
    # several frames back to back without a host synchronisation in between: a reduce that is not ordered after its
    # render (or a reset racing with the next render's atomics) shows up as a wrong sample count or image
    for n in (5, 4, 3):
        r.render(n)
    if rank == 0:
        acc = r.ctx.read_sum()
        np.save(out_path, np.concatenate([r.read_hdr()[None], acc[None]]))
    r.synchronize()
    dist.barrier()
    r.close()
    dist.destroy_process_group()


def test_two_rank_nccl_render_equals_single_rank(gsb, tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    out = str(tmp_path / "img.npy")
    port = 29700 + os.getpid() % 200
    mp.start_processes(_worker, args=(2, port, out), nprocs=2, join=True, start_method="spawn")
    got, acc = np.load(out)
    assert np.all(acc[..., 3] == SPP)        # every pixel holds exactly the 12 samples of the three frames
    with gsb.Context(0) as c:
        c.upload(gsb.Scene.builtin("cornell"))
        c.set_params(max_depth=8)
        c.render(W, H, 0, 0); c.clear(); c.render(W, H, 0, SPP)
        want = c.read_hdr()
    assert np.allclose(got, want, rtol=2e-5, atol=1e-6)
