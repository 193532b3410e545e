"""Sample-index sharding of a render across the GPUs of one box (one process per GPU, torch.distributed).

The reference is single-GPU; a pixel-sample depends only on (x, y, timestamp) through its seed (raygen.rgen:37) and samples
combine by a mean (raygen.rgen:84-108), so GPU g of G renders timestamps first + g, first + g + G, ... for every pixel — the
RNG streams are those of a 1-GPU run — and the float4 sum buffers (rgb = sum, a = samples added) are combined by ONE reduce
per frame.  Dividing by the alpha channel after the reduce keeps the NaN-skip semantics.
"""
import numpy as np


def shard_samples(first, count, rank, world):
    """(first_sample, num_samples, stride) of `rank` for the timestamps first .. first+count-1 split over `world` ranks."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    mine = (count - rank + world - 1) // world if count > rank else 0
    return first + rank, mine, world


def reduce_frame(accum, dst=0, group=None):
    """In-place sum of every rank's sum buffer (a torch tensor viewing it) onto rank `dst`: the frame's single collective."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def resolve(accum_rgba):
    """Mean image from a (reduced) sum buffer: rgb / a, alpha 1 (what gsb_read_hdr does on the device)."""
    a = np.asarray(accum_rgba, np.float32).reshape(-1, 4)
    out = np.zeros_like(a)
    n = a[:, 3:4]
    np.divide(a[:, :3], n, out=out[:, :3], where=n > 0)
    out[:, 3] = 1.0
    return out.reshape(np.shape(accum_rgba))


class DeviceBuffer:
    """__cuda_array_interface__ view of a device allocation (used to hand the context's sum buffer to torch / NCCL, no copy)."""

    def __init__(self, ptr, nfloats):
        self.__cuda_array_interface__ = {"shape": (nfloats,), "typestr": "<f4", "data": (ptr, False), "version": 3, "strides": None}


class ShardedRenderer:
    """One rank's share of a multi-GPU progressive render: Context + torch.distributed (NCCL) on the same CUDA stream."""

    def __init__(self, scene, width, height, device_index=0, params=None):
        import torch
        import torch.distributed as dist

        from .api import Context
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.width, self.height = width, height
        self.device = torch.device("cuda", device_index)
        self.ctx = Context(device_index)
        # An explicit stream shared by the render, the NCCL reduce and the buffer resets.  torch's default stream has the
        # handle 0, which gsb_set_stream reads as "create a private stream": the reduce would then not be ordered after
        # the (asynchronous) render.  Every torch call below therefore runs under `with torch.cuda.stream(self.stream)`.
        self.stream = torch.cuda.Stream(device=self.device)
        assert self.stream.cuda_stream != 0
        self.ctx.set_stream(self.stream.cuda_stream)
        self.ctx.upload(scene)
        if params:
            self.ctx.set_params(**params)
        self.ctx.render(width, height, 0, 0)
        self.ctx.clear()
        ptr, nbytes = self.ctx.device_sum_ptr()
        self.accum = torch.as_tensor(DeviceBuffer(ptr, nbytes // 4), device=self.device)
        self.next_sample = 0

    def render(self, spp_total):
        """Adds `spp_total` samples per pixel to the frame, split over the ranks; rank 0 ends up holding the global sum."""
        first, n, stride = shard_samples(self.next_sample, spp_total, self.rank, self.world)
        import torch
        self.ctx.render(self.width, self.height, first, n, stride)    # asynchronous, on self.stream
        with torch.cuda.stream(self.stream):
            reduce_frame(self.accum, 0)
            if self.rank != 0:
                self.accum.zero_()
        self.next_sample += spp_total

    def synchronize(self):
        self.stream.synchronize()

    def read_hdr(self, out=None):
        return self.ctx.read_hdr(out)

    def close(self):
        self.ctx.close()
