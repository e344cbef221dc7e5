"""World-size-2 run of the multi-rank host logic on the CPU (gloo): scan sharding, the optional cloud gather and the
max-over-ranks timing reduction that bench.py uses under torchrun."""
import os
import socket
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys
    sys.path.insert(0, %r)
    import numpy as np, torch, torch.distributed as dist
    from lidarshooter_b200 import shard
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    # sharding: every (device, frame) scan has exactly one owner, for both partitionings
    for major in ("frame", "device"):
        mine = shard.shard_scans(5, 7, rank, world, major)
        flat = torch.zeros(35, dtype=torch.int64)
        for d, f in mine:
            flat[f * 5 + d] += 1
        dist.all_reduce(flat)
        assert bool((flat == 1).all()), (major, flat)
    # cloud gather: rank r contributes 3 + 2 r records whose bytes identify (rank, row)
    n = 3 + 2 * rank
    pts = (torch.arange(n * 32, dtype=torch.int64).reshape(n, 32) %% 251 + rank).to(torch.uint8)
    out = shard.gather_clouds(pts, dst=0)
    if rank == 0:
        exp = torch.cat([(torch.arange((3 + 2 * r) * 32, dtype=torch.int64).reshape(-1, 32) %% 251 + r).to(torch.uint8) for r in range(world)])
        assert out.shape == exp.shape and bool((out == exp).all())
    else:
        assert out is None
    empty = shard.gather_clouds(torch.zeros((0, 32), dtype=torch.uint8), dst=0)      # ragged / empty inputs
    assert (empty is None) if rank else empty.shape == (0, 32)
    t = shard.reduce_max([1.0 + rank, 5.0 - rank])
    assert t == [float(world), 5.0], t
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "rank" + str(rank) + ".ok"), "w").write("ok")
""") % ROOT


def test_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script)]
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert (tmp_path / "rank0.ok").exists() and (tmp_path / "rank1.ok").exists(), r.stdout


def test_shard_is_a_partition():
    from lidarshooter_b200 import shard
    for world in (1, 2, 4, 8):
        seen = sorted(u for r in range(world) for u in shard.shard_round_robin(4096, r, world))
        assert seen == list(range(4096))
        sizes = [len(shard.shard_round_robin(4096, r, world)) for r in range(world)]
        assert max(sizes) - min(sizes) <= 1
