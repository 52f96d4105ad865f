"""Host-side multi-process logic on CPU: gloo backend, world size 2.
The data path has no collective; what is tested is the sharding arithmetic, the gradient
all-reduce (mean) that replaces nn.DataParallel, and the end-of-stage gathers."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from stablekeypoints_b200 import distributed as D
    r, lr, ws = D.init_from_env("gloo")
    assert (r, ws) == (rank, world) and D.is_dist()
    # 1. image sharding: disjoint, complete, balanced
    n = 7
    mine = list(D.shard_range(n))
    # 2. gradient all-reduce == mean over ranks (the reference averages the per-GPU losses, optimize.py:428-429)
    g = torch.full((1, 5, 3), float(rank + 1))
    D.allreduce_mean_(g)
    # 3. broadcast of the initial context from rank 0
    ctx = torch.full((2, 2), float(rank))
    D.broadcast_(ctx)
    # 4. gather of per-rank keypoints back into dataset order
    per = (n + world - 1) // world
    rows = torch.zeros(per, 2)
    for j, i in enumerate(mine):
        rows[j] = torch.tensor([float(i), float(i) * 10])
    full = D.interleave_shards(D.all_gather_cat(rows), n)
    # 5. sharded dataset
    ds = D.shard_dataset(list(range(n)))
    # 6. per-rank random streams (the reference draws an independent affine per image, invertable_transform.py:42-57;
    #    every process starts from the same default seed, so ranks must derive their own)
    from stablekeypoints_b200.invertable_transform import RandomAffineWithInverse
    g_cpu, g_dev = D.rank_generators("cpu", seed=11)
    theta = RandomAffineWithInverse(degrees=15, scale=(0.8, 1.0), translate=(0.25, 0.25)).sample_theta(2, g_cpu)
    out[rank] = dict(mine=mine, g=g, ctx=ctx, full=full, ds=[ds[i] for i in range(len(ds))], theta=theta,
                     noise=torch.randn(3, generator=g_dev))
    D.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_gloo_world2_sharding_allreduce_and_gathers():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    a, b = out[0], out[1]
    assert sorted(a["mine"] + b["mine"]) == list(range(7)) and not set(a["mine"]) & set(b["mine"])
    assert abs(len(a["mine"]) - len(b["mine"])) <= 1
    assert a["ds"] == a["mine"] and b["ds"] == b["mine"]
    assert not torch.equal(a["theta"], b["theta"]) and not torch.equal(a["noise"], b["noise"])
    assert not torch.equal(a["theta"][0], a["theta"][1])           # and independent per image within a rank
    for r in (a, b):
        assert torch.allclose(r["g"], torch.full((1, 5, 3), 1.5))
        assert torch.equal(r["ctx"], torch.zeros(2, 2))
        assert torch.equal(r["full"][:, 0], torch.arange(7.0)) and torch.equal(r["full"][:, 1], torch.arange(7.0) * 10)


def test_single_process_helpers_are_noops():
    from stablekeypoints_b200 import distributed as D
    assert D.world_size() == 1 and D.rank() == 0 and not D.is_dist()
    t = torch.ones(3)
    D.allreduce_mean_(t)
    D.broadcast_(t)
    assert torch.equal(D.all_gather_cat(t), t) and list(D.shard_range(5)) == [0, 1, 2, 3, 4]
    assert D.init_from_env() == (0, 0, 1)
    assert D.rank_generators("cpu") == (None, None)       # single process, no seed: the global generators stay in charge
    g1, _ = D.rank_generators("cpu", seed=3)
    g2, _ = D.rank_generators("cpu", seed=3)
    assert torch.equal(torch.rand(4, generator=g1), torch.rand(4, generator=g2))


def test_vote_top_k_counts_and_tie_rule():
    from stablekeypoints_b200.keypoint_regressor import vote_top_k
    votes = torch.tensor([[5, 2, 9], [5, 9, 1], [5, 2, 7]])
    got = vote_top_k(votes, 3)
    assert got.tolist() == [5, 2, 9]          # counts 3, 2, 2; the tie 2 vs 9 goes to the smaller id
    assert got.dtype == torch.int64
