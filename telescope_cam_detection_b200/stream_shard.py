"""Host side of the multi-GPU layout: camera streams are sharded across ranks, one
process per GPU, no exchange on the compute path.  The reference is single-GPU
(src/shared_inference_coordinator.py:18-338 serves every camera from one model); here each rank runs
that coordinator over its own cameras and only the small result tables travel to rank 0 for serving.

Backend-agnostic (`nccl` on the GPU box, `gloo` in the CPU tests): tensors live wherever the caller
puts them.  Nothing here touches the detection kernels.
"""
import numpy as np

ROW = 7  # x1, y1, x2, y2, obj_conf, class_conf, class_id  (yolox.utils.postprocess row)


def assign_streams(camera_ids, world):
    """Stable round-robin over the SORTED camera ids: {rank: [camera_id, ...]}.  Per-stream state (MOG2
    model, empty-frame reference image) lives on the owning rank and never migrates."""
    out = {r: [] for r in range(world)}
    for i, cam in enumerate(sorted(camera_ids)):
        out[i % world].append(cam)
    return out


def my_streams(camera_ids, rank, world):
    return assign_streams(camera_ids, world)[rank]


def owner_of(camera_id, camera_ids, world):
    return sorted(camera_ids).index(camera_id) % world


def pack_table(rows, max_out, out=None):
    """List of B arrays [n_i, 7] -> fixed table [B, max_out + 1, 7] f32; slot [b, 0, 0] carries n_i (a count
    row, so one tensor travels), rows beyond n_i are zero.  Raises if a frame has more than max_out rows
    (the kernels cap at max_out, so that is a caller bug, not data)."""
    B = len(rows)
    if out is None:
        out = np.zeros((B, max_out + 1, ROW), np.float32)
    for b, r in enumerate(rows):
        n = 0 if r is None else int(r.shape[0])
        if n > max_out:
            raise ValueError("frame %d has %d rows > max_out %d" % (b, n, max_out))
        out[b, 0] = 0
        out[b, 0, 0] = n
        if n:
            out[b, 1:1 + n] = r
        out[b, 1 + n:] = 0
    return out


def unpack_table(table):
    """Inverse of pack_table: -> list of B arrays [n_i, 7]."""
    table = np.asarray(table)
    return [table[b, 1:1 + int(table[b, 0, 0])].copy() for b in range(table.shape[0])]


class DeviceView:
    """A device buffer owned by libtcd, exposed through __cuda_array_interface__ so torch can wrap it without a copy
    (torch.as_tensor(DeviceView(...), device="cuda"))."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}


class ResultGather:
    """Per-step gather of the packed result tables to rank 0.  `gather(t)` returns, on rank 0, the list of
    every rank's table (rank order); None elsewhere.  With world == 1 it is a pass-through."""

    def __init__(self, world, rank, like):
        import torch

        self.world, self.rank = world, rank
        self.bufs = [torch.zeros_like(like) for _ in range(world)] if (world > 1 and rank == 0) else None

    def gather(self, t):
        if self.world == 1:
            return [t]
        import torch.distributed as dist

        dist.gather(t, self.bufs, dst=0)
        return self.bufs


class RingGather:
    """Result tables of `ring` consecutive steps travel in ONE collective (the per-call host cost of a collective,
    not its 67 KB, is what a 0.7 ms step cannot afford: measured 19 746 -> 9 722 frames/s at N = 2 with one
    gather per step).  all_gather_into_tensor = a single kernel with a contiguous output; every rank receives the
    ring (world x ring x 67 KB), rank 0 is the one that serves it.  Latency to rank 0 <= ring steps."""

    def __init__(self, world, ring, row_len, device, dtype=None):
        import torch

        dtype = dtype or torch.float32
        self.world, self.ring = world, ring
        self.buf = [torch.zeros((ring, row_len), dtype=dtype, device=device) for _ in range(2)]
        self.out = [torch.zeros((world, ring, row_len), dtype=dtype, device=device) for _ in range(2)]

    def slot(self, step):
        """-> (half, index): where the table of global step `step` goes."""
        return (step // self.ring) & 1, step % self.ring

    def collect(self, half):
        """One collective over ring buffer `half`; returns the [world, ring, row_len] tensor (valid on every rank)."""
        if self.world == 1:
            self.out[half][0].copy_(self.buf[half])
            return self.out[half]
        import torch.distributed as dist

        if dist.get_backend() == "nccl":
            dist.all_gather_into_tensor(self.out[half], self.buf[half])
        else:  # gloo (CPU tests)
            parts = [self.out[half][r] for r in range(self.world)]
            dist.all_gather(parts, self.buf[half])
        return self.out[half]


def max_over_ranks(value, world, device="cpu"):
    """Device-timed milliseconds -> max over ranks (the job's step time is its slowest rank's)."""
    if world == 1:
        return float(value)
    import torch
    import torch.distributed as dist

    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_fps(frames_per_rank_step, steps, world, ms):
    """Whole-job frames/s: the frames ALL ranks processed / the slowest rank's device time."""
    return world * frames_per_rank_step * steps / (ms * 1e-3)
