"""N > 1 host logic on CPU: world_size-2 gloo processes: image sharding balanced by cells, ONE
all-reduce for the loss normalisers (reference head.py:167-172 semantics: sum over ranks / world size), bucketed
gradient averaging, max-over-ranks timing."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from skghoi_b200 import shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = {}
        # 1. image shards: ragged batch, cost = cells per image
        n_h = [16, 4, 8, 40, 2, 16, 64, 3]; n = [32, 8, 16, 80, 5, 32, 64, 9]
        cost = shard.cells_of(n_h, n)
        mine = shard.shard_of(list(range(len(cost))), cost)
        res["mine"] = mine
        # 2. loss normalisers: one all-reduce, rank mean
        n_p = torch.tensor([3.0 + rank, 10.0 * (rank + 1)])
        res["n_p"] = shard.average_normalisers(n_p).tolist()
        assert n_p.tolist() == [3.0 + rank, 10.0 * (rank + 1)]          # input left untouched
        n_q = n_p.clone()
        n_avg = shard.average_normalisers(n_q, in_place=True)           # the training pre-pass gives its buffer up
        assert n_avg.data_ptr() == n_q.data_ptr() and n_avg.tolist() == res["n_p"]
        # 3. gradient buckets: small bucket size -> several buckets, one param without grad
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(7, 5), torch.nn.ReLU(), torch.nn.Linear(5, 3), torch.nn.Linear(3, 2))
        x = torch.full((4, 7), float(rank + 1))
        net[2](net[1](net[0](x))).sum().backward()                      # net[3] gets no gradient
        local = [None if p.grad is None else p.grad.clone() for p in net.parameters()]
        gb = shard.GradBuckets(net.parameters(), bucket_bytes=64)
        res["buckets"] = len(gb.buckets)
        gb.allreduce()
        res["grads"] = [p.grad.clone() for p in net.parameters()]
        res["local"] = local
        # 3b. attached mode: gradients live in the flat buckets; zero_grad + backward + all-reduce without packing
        gb2 = shard.GradBuckets(net.parameters(), bucket_bytes=64).attach()
        gb2.zero_grad()
        net[2](net[1](net[0](x))).sum().backward()
        assert gb2._attached()
        gb2.allreduce()
        assert gb2._direct
        res["grads_attached"] = [p.grad.clone() for p in net.parameters()]
        # 4. timing: max over ranks
        res["tmax"] = shard.max_over_ranks([1.0 + rank, 5.0 - rank])
        out[rank] = res
    finally:
        dist.destroy_process_group()


def test_world2_gloo_host_logic():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    r0, r1 = out[0], out[1]
    # shards: a partition of all images, balanced by cells
    n_h = [16, 4, 8, 40, 2, 16, 64, 3]; n = [32, 8, 16, 80, 5, 32, 64, 9]
    cost = shard.cells_of(n_h, n)
    assert sorted(r0["mine"] + r1["mine"]) == list(range(8))
    assert r0["mine"] == sorted(r0["mine"]) and r1["mine"] == sorted(r1["mine"])
    l0, l1 = sum(cost[i] for i in r0["mine"]), sum(cost[i] for i in r1["mine"])
    assert abs(l0 - l1) <= max(cost)
    assert [r0["mine"], r1["mine"]] == shard.partition_images(cost, 2)
    # normalisers: (sum over ranks) / world, identical on both ranks
    assert r0["n_p"] == r1["n_p"] == [3.5, 15.0]
    # gradients: rank mean; parameters without a local gradient on any rank come back as zeros
    assert r0["buckets"] > 1
    for g0, g1, a, b in zip(r0["grads"], r1["grads"], r0["local"], r1["local"]):
        torch.testing.assert_close(g0, g1, rtol=0, atol=0)
        want = torch.zeros_like(g0) if a is None else (a + b) / 2
        torch.testing.assert_close(g0, want, rtol=1e-6, atol=1e-7)
    for g0, ga in zip(r0["grads"], r0["grads_attached"]):          # attached mode gives the same rank mean
        torch.testing.assert_close(ga, g0, rtol=1e-6, atol=1e-7)
    assert r0["tmax"] == r1["tmax"] == [2.0, 5.0]


def test_partition_properties():
    # single rank: everything; more ranks than images: some shards empty, none lost
    assert shard.partition_images([5, 1, 3], 1) == [[0, 1, 2]]
    p = shard.partition_images([5, 1, 3], 5)
    assert sorted(i for s in p for i in s) == [0, 1, 2] and len(p) == 5
    assert shard.partition_images([], 3) == [[], [], []]
    with pytest.raises(ValueError):
        shard.partition_images([1], 0)
    # equal costs -> equal counts
    p = shard.partition_images([512] * 16, 8)
    assert all(len(s) == 2 for s in p)
    # without a process group the collectives are identities
    t = torch.tensor([2.0, 4.0])
    assert shard.average_normalisers(t) is t
    assert shard.max_over_ranks([3.0]) == [3.0]


def test_flat_params_follow_a_fused_style_optimizer():
    """GradBuckets.attach(flatten_params=True): parameters and gradients live in flat buffers, an optimiser over
    ``flat_params`` gives the same update as the per-tensor one, and state_dict / module access are unchanged."""
    import copy
    torch.manual_seed(0)
    net = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.ReLU(), torch.nn.Linear(7, 3))
    ref = copy.deepcopy(net)
    x = torch.randn(4, 5)
    from skghoi_b200.shard import GradBuckets
    holder = torch.nn.Module()
    buckets = GradBuckets(net.parameters(), bucket_bytes=64).attach(flatten_params=True, module=holder)
    assert len(buckets.flat_params) >= 2 and holder.__dict__["_skg_flat_params"] is buckets.flat_params
    opt = torch.optim.SGD(buckets.flat_params, lr=0.1, momentum=0.9)
    opt_ref = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
    for _ in range(3):
        buckets.zero_grad(); opt_ref.zero_grad()
        net(x).pow(2).sum().backward(); ref(x).pow(2).sum().backward()
        v0 = sum(f._version for f in buckets.flat_params)
        opt.step(); opt_ref.step()
        assert sum(f._version for f in buckets.flat_params) > v0          # the change detector of the packed operands
    for (k, a), (_, b) in zip(net.state_dict().items(), ref.state_dict().items()):
        torch.testing.assert_close(a, b, rtol=1e-6, atol=1e-7)
