"""BASELINE configs[3] on N > 1 (north_star: "by triple batch for TransH"), host logic on CPU with world_size-2 gloo:
``kg_train.Trainer(distributed=True)`` slices every global batch [B positives | B negatives x rounds] per rank, averages the
table gradients with one all-reduce and must reproduce the single-process trajectory on the same global batches.
The scoring module here is a plain-torch stand-in (the product's TransH runs on CUDA only); the Trainer, the slicing and
the bucketed all-reduce are the product code under test."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

from skghoi_b200.kg_train import Trainer

ENT, REL, DIM, B, NEG = 11, 5, 6, 8, 3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class _Loader:
    """Deterministic global batches in the sampler's layout (Base.cpp:161-197)."""

    def __init__(self, nbatches=2, seed=5):
        self.nb, self.seed = nbatches, seed
        self.epoch = 0

    def get_batch_size(self):
        return B

    def __len__(self):
        return self.nb

    def __iter__(self):
        g = torch.Generator().manual_seed(self.seed + self.epoch)
        self.epoch += 1
        for _ in range(self.nb):
            n = B * (1 + NEG)
            yield dict(batch_h=torch.randint(0, ENT, (n,), generator=g), batch_t=torch.randint(0, ENT, (n,), generator=g),
                       batch_r=torch.randint(0, REL, (n,), generator=g), batch_y=torch.ones(n), mode="normal")


class _Strategy(nn.Module):
    """TransE-style stand-in with the strategy contract of kg_train.NegativeSamplingStrategy (``batch_size`` attribute,
    ``forward(data)`` -> margin loss over [B, 1] positives against [B, NEG] negatives)."""

    def __init__(self):
        super().__init__()
        torch.manual_seed(3)
        self.ent = nn.Embedding(ENT, DIM); self.rel = nn.Embedding(REL, DIM)
        self.batch_size = B

    def forward(self, data):
        s = (self.ent(data["batch_h"]) + self.rel(data["batch_r"]) - self.ent(data["batch_t"])).abs().sum(1)
        bs = self.batch_size
        p = s[:bs].view(-1, bs).permute(1, 0); n = s[bs:].view(-1, bs).permute(1, 0)
        return torch.clamp(p - n, min=-2.0).mean() + 2.0


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m = _Strategy()
        tr = Trainer(m, _Loader(), train_times=3, alpha=0.1)
        assert tr.distributed
        hist = tr.run()
        out[rank] = dict(hist=hist, ent=m.ent.weight.detach().clone(), rel=m.rel.weight.detach().clone(),
                         span=tr._span, idx=tr._shard_idx.clone())
    finally:
        dist.destroy_process_group()


def test_trainer_shards_triple_batch_over_two_ranks():
    world = 2
    mgr = mp.Manager(); out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    ref = _Strategy()
    tr = Trainer(ref, _Loader(), train_times=3, alpha=0.1, distributed=False)
    hist = tr.run()
    assert out[0]["span"] == (0, 4) and out[1]["span"] == (4, 8)
    # rank 1's slice: positives 4..7 and, from each negative round k, entries B*k + 4..7
    assert out[1]["idx"].tolist() == [k * B + i for k in range(1 + NEG) for i in range(4, 8)]
    for r in range(world):
        torch.testing.assert_close(torch.tensor(out[r]["hist"]), torch.tensor(hist), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(out[r]["ent"], ref.ent.weight.detach(), rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(out[r]["rel"], ref.rel.weight.detach(), rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(out[0]["ent"], out[1]["ent"], rtol=0, atol=0)     # replicas stay bit-identical
