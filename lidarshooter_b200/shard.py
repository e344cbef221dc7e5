"""Multi-GPU partitioning of the trace path: one process per GPU, the scene replicated, (device, frame) scans
sharded with NO data-path collective.  The only collective is the OPTIONAL gather of the per-GPU point
clouds (NCCL over NVLink on GPUs; gloo in the CPU tests) and the max-over-ranks reduction of the timings.

The reference has no multi-GPU code: it runs one tracer per device in one process (SRC/mainwindow.cpp:256-274).
"""
import torch
import torch.distributed as dist


def shard_round_robin(n_units, rank, world):
    """Units (frames, or devices) owned by `rank`: rank, rank + world, ...  Every unit has exactly one owner."""
    return list(range(rank, n_units, world))


def shard_scans(n_devices, n_frames, rank, world, major="frame"):
    """(device, frame) scans of this rank.  major='frame': whole frames per rank (each rank builds the BVH of its own
    frames only -- config 5); major='device': every rank sees every frame but traces its own devices (config 4 when
    the frame must be finished as early as possible)."""
    if major == "frame":
        return [(d, f) for f in shard_round_robin(n_frames, rank, world) for d in range(n_devices)]
    return [(d, f) for f in range(n_frames) for d in shard_round_robin(n_devices, rank, world)]


def reduce_max(values, device=None):
    """Max over ranks of a list of floats (timings are reported as the slowest rank's)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=device or ("cuda" if dist.get_backend() == "nccl" else "cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def gather_clouds(points, dst=0):
    """Optional gather of variable-length point clouds: `points` is this rank's (n, record_bytes) uint8 tensor (on the
    GPU under NCCL -- the transfers then run over NVLink / NVSwitch peer to peer; on the CPU under gloo).
    One all_gather of the counts (8 bytes per rank), then every rank sends exactly its own count * record_bytes to
    `dst` (batched isend / irecv: one NCCL group, no padding to the largest cloud).
    Returns the concatenation in rank order on rank `dst`, None elsewhere.
    Counterpart in the reference: each device's cloud is published on its own ROS topic (SRC/MeshProjector.cpp:380-390)
    and merged by whoever subscribes; there is no collective."""
    world = dist.get_world_size()
    rank = dist.get_rank()
    n = torch.tensor([points.shape[0]], dtype=torch.int64, device=points.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n)
    counts = [int(c.item()) for c in counts]
    width = points.shape[1]
    points = points.contiguous()
    if rank == dst:
        out = torch.empty((sum(counts), width), dtype=torch.uint8, device=points.device)
        offs = [0]
        for c in counts:
            offs.append(offs[-1] + c)
        out[offs[rank]:offs[rank + 1]] = points
        ops = [dist.P2POp(dist.irecv, out[offs[r]:offs[r + 1]], r) for r in range(world) if r != dst and counts[r] > 0]
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        return out
    if points.shape[0] > 0:
        for req in dist.batch_isend_irecv([dist.P2POp(dist.isend, points, dst)]):
            req.wait()
    return None


class DeviceArray:
    """A raw device pointer (lsb200_device_results) dressed as a __cuda_array_interface__ object, so that
    torch.as_tensor(DeviceArray(...), device='cuda') wraps the library's buffer without a copy."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}
