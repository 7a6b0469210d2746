"""CPU, world_size 2 over gloo: the data-parallel host logic of the trainer (ray sharding + gradient
mean-allreduce with in-place big tensors and one flat bucket for the small ones)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nerf_kbs_b200 import trainer as T
    # full-batch "gradients" are the mean over ranks of per-shard gradients
    g = torch.Generator().manual_seed(0)
    full = [torch.randn(T.BIG_TENSOR + 5, 2, generator=g), torch.randn(64, 32, generator=g), torch.randn(7, generator=g)]
    shard = [t * (rank + 1) for t in full]  # rank-dependent contribution
    T.allreduce_mean_(shard)
    want = [t * (sum(range(1, world + 1)) / world) for t in full]
    ok = all(torch.allclose(a, b, rtol=1e-6, atol=1e-6) for a, b in zip(shard, want))
    rng = T.shard_rays(4096, rank, world)
    q.put((rank, ok, rng.start, rng.stop))
    dist.barrier()
    dist.destroy_process_group()


def test_allreduce_mean_and_ray_sharding_world2():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(r[1] for r in res)
    assert [(r[2], r[3]) for r in res] == [(0, 2048), (2048, 4096)]


def test_shard_rays_covers_batch():
    from nerf_kbs_b200.trainer import shard_rays
    for ws in (1, 2, 4, 8):
        idx = [i for r in range(ws) for i in shard_rays(4096, r, ws)]
        assert idx == list(range(4096))


def test_sharded_adam_flat_layout():
    """Slices of the fused NVLink optimizer step: aligned, disjoint, and covering every group on every world size."""
    from nerf_kbs_b200.dp_optim import flat_layout
    groups = {"proposal_networks": [1310720 * 2, 160, 16, 16, 1] * 2, "fields": [16777216, 2048, 64, 1024, 16, 3680, 4032, 64, 4096, 64, 192, 3]}
    for world in (1, 2, 4, 8):
        slices = {}
        for rank in range(world):
            lay = flat_layout(groups, world, rank)
            total = lay.pop("__total__")
            for name, g in lay.items():
                assert g["lo"] % 4 == 0 and g["n"] % 4 == 0 and g["len"] == g["n"] * world
                assert all(o % 4 == 0 for o in g["offs"]) and g["offs"][0] == g["start"]
                assert g["offs"][-1] + groups[name][-1] <= g["start"] + g["len"] <= total
                slices.setdefault(name, []).append((g["lo"], g["lo"] + g["n"]))
        for name, sl in slices.items():
            sl.sort()
            assert sl[0][0] == lay[name]["start"] and sl[-1][1] == lay[name]["start"] + lay[name]["len"]
            assert all(a[1] == b[0] for a, b in zip(sl, sl[1:]))


def test_rank_ray_ranges_partition_a_frame():
    """Full-frame render over several ranks: contiguous, ordered, balanced blocks that cover every ray exactly once."""
    from nerf_kbs_b200.render import rank_ray_range
    N = 375 * 1242
    for ws in (1, 2, 4, 8):
        blocks = [rank_ray_range(N, r, ws) for r in range(ws)]
        assert [i for b in blocks for i in (b.start, b.stop)][0] == 0 and blocks[-1].stop == N
        assert all(a.stop == b.start for a, b in zip(blocks, blocks[1:]))
        assert max(len(b) for b in blocks) - min(len(b) for b in blocks) < ws
    # ceil-sized blocks: a frame much smaller than the rank count leaves the last ranks empty (the driver refuses that)
    assert [len(rank_ray_range(5, r, 4)) for r in range(4)] == [2, 2, 1, 0]
