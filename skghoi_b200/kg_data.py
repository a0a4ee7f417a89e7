"""Knowledge-graph training data on the device: the index structures OpenKE's importTrainFiles builds
(OpenKE/openke/base/Reader.h:52-160) and the negative sampler behind ``skg_kg_sample`` (csrc/kgsample.cu), with the
``sampling`` contract of OpenKE/openke/base/Base.cpp:161-197 (the Python ``openke.data`` loader is not part of the
reference tree, so that C entry point is the interface mirrored here).  Bit-exact with the reference's sampler for any
``work_threads`` given the same per-thread seeds."""
from typing import Dict, Optional, Sequence

import numpy as np
import torch
from torch import Tensor

from . import _lib


class KGIndex:
    """De-duplicated training triples in head order and tail order, entity ranges, and the bern statistics."""

    def __init__(self, h: Sequence[int], t: Sequence[int], r: Sequence[int], ent_total: int, rel_total: int, device="cuda"):
        trip = np.unique(np.stack([np.asarray(h, np.int64), np.asarray(r, np.int64), np.asarray(t, np.int64)], 1), axis=0)
        # np.unique sorts rows lexicographically by (h, r, t) = Triple::cmp_head
        head = trip
        tail = trip[np.lexsort((trip[:, 0], trip[:, 1], trip[:, 2]))]                    # (t, r, h) = Triple::cmp_tail
        T, E, R = len(trip), int(ent_total), int(rel_total)
        if T == 0:
            raise ValueError("the knowledge graph has no training triples")
        self.train_total, self.ent_total, self.rel_total = T, E, R

        def ranges(col):
            lef = np.zeros(E, np.int64); rig = np.full(E, -1, np.int64)
            ents, first = np.unique(col, return_index=True)
            last = np.concatenate([first[1:], [len(col)]]) - 1
            lef[ents] = first; rig[ents] = last
            return lef, rig
        lef_head, rig_head = ranges(head[:, 0])
        lef_tail, rig_tail = ranges(tail[:, 2])
        freq = np.bincount(trip[:, 1], minlength=R).astype(np.float32)
        # left_mean[r] = triples of r / distinct heads of r, right_mean[r] = triples of r / distinct tails of r
        lm = np.bincount(np.unique(trip[:, [0, 1]], axis=0)[:, 1], minlength=R).astype(np.float32)
        rm = np.bincount(np.unique(trip[:, [2, 1]], axis=0)[:, 1], minlength=R).astype(np.float32)
        with np.errstate(divide="ignore", invalid="ignore"):
            self.left_mean_host = (freq / lm).astype(np.float32); self.right_mean_host = (freq / rm).astype(np.float32)
        dev = torch.device(device)
        to = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)                 # noqa: E731
        self.head = to(head.T); self.tail = to(tail.T)                                   # [3, T] rows h | r | t
        self.lef_head, self.rig_head, self.lef_tail, self.rig_tail = to(lef_head), to(rig_head), to(lef_tail), to(rig_tail)
        self.left_mean, self.right_mean = to(self.left_mean_host), to(self.right_mean_host)
        self.device = dev


class TrainDataSampler:
    """``sampling(batch_h, batch_t, batch_r, batch_y, batchSize, negRate, negRelRate=0, mode, ...)`` (Base.cpp:161-197)
    as an iterator of device batches: positives drawn uniformly from the training triples, ``neg_rate`` filtered
    corruptions each, head/tail chosen by a fair or "bern" coin (mode 0) or fixed (mode -1 / +1)."""

    def __init__(self, index: KGIndex, batch_size: int, neg_rate: int = 1, bern: bool = False, work_threads: int = 8,
                 seeds: Optional[Sequence[int]] = None, mode: int = 0):
        _lib.require_cuda(index.head)
        self.index, self.batch_size, self.neg_rate, self.bern, self.W, self.mode = index, int(batch_size), int(neg_rate), bool(bern), int(work_threads), int(mode)
        if seeds is None:                                      # randReset(): next_random[i] = rand()
            seeds = np.random.randint(0, 2 ** 31 - 1, size=self.W)
        s = np.asarray([int(x) & ((1 << 64) - 1) for x in seeds], dtype=np.uint64)
        if len(s) != self.W:
            raise ValueError("one seed per worker thread")
        dev = index.device
        self._seeds = [torch.from_numpy(s.view(np.int64).copy()).to(dev), torch.empty(self.W, dtype=torch.int64, device=dev)]
        n = self.batch_size * (1 + self.neg_rate)
        self.batch_h = torch.empty(n, dtype=torch.int64, device=dev); self.batch_t = torch.empty_like(self.batch_h)
        self.batch_r = torch.empty_like(self.batch_h); self.batch_y = torch.empty(n, dtype=torch.float32, device=dev)

    def seeds(self):
        return [int(v) & ((1 << 64) - 1) for v in self._seeds[0].cpu().tolist()]

    def sample(self) -> Dict[str, Tensor]:
        ix, p = self.index, _lib.ptr
        lm, rm = (ix.left_mean, ix.right_mean) if self.bern else (None, None)
        with torch.cuda.device(ix.device):
            _lib.call("skg_kg_sample", p(ix.head), p(ix.tail), p(ix.lef_head), p(ix.rig_head), p(ix.lef_tail), p(ix.rig_tail),
                      p(lm), p(rm), ix.train_total, ix.ent_total, p(self._seeds[0]), p(self._seeds[1]), self.W, self.batch_size,
                      self.neg_rate, self.mode, p(self.batch_h), p(self.batch_t), p(self.batch_r), p(self.batch_y),
                      _lib.stream_ptr(ix.device))
        self._seeds.reverse()
        return dict(batch_h=self.batch_h, batch_t=self.batch_t, batch_r=self.batch_r, batch_y=self.batch_y, mode="normal")

    def __iter__(self):
        while True:
            yield self.sample()
