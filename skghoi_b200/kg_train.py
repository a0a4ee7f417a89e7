"""KG-embedding training loop on the device (BASELINE configs[3]): the upstream OpenKE pieces the head's
TransH tables would be trained with -- the dict-of-batches ``forward(data)`` strategy
(OpenKE/openke/module/strategy/NegativeSampling.py:13-32), the data loader contract around ``sampling``
(OpenKE/openke/base/Base.cpp:161-197) and ``Trainer`` (OpenKE/openke/config/Trainer.py:43-99).  Sampling
(``skg_kg_sample``), scoring and its gradient (``skg_transh_forward`` / ``_backward``) and the margin loss
(``skg_margin_loss``) are CUDA kernels of this library; the optimiser is torch.optim, as in the reference."""
from typing import Iterator, List, Optional, Sequence

import torch
import torch.distributed as dist
from torch import nn

from .kg_data import KGIndex, TrainDataSampler
from .kge import MarginLoss, TransH


class TrainDataLoader:
    """``nbatches`` batches per epoch of ``batch_size = train_total // nbatches`` positives (or an explicit batch size),
    ``neg_ent`` filtered corruptions each; iterating yields dicts ``{batch_h, batch_t, batch_r, batch_y, mode}`` on the
    device (the reference's loader hands the same dict over host numpy buffers)."""

    def __init__(self, index: KGIndex, nbatches: Optional[int] = None, batch_size: Optional[int] = None, neg_ent: int = 1,
                 bern_flag: bool = False, threads: int = 8, seeds: Optional[Sequence[int]] = None):
        if batch_size is None:
            if not nbatches:
                raise ValueError("give nbatches or batch_size")
            batch_size = index.train_total // nbatches
        if nbatches is None:
            nbatches = max(index.train_total // batch_size, 1)
        self.nbatches, self.batch_size, self.neg_ent = int(nbatches), int(batch_size), int(neg_ent)
        self.sampler = TrainDataSampler(index, self.batch_size, self.neg_ent, bern_flag, threads, seeds)

    def get_batch_size(self) -> int:
        return self.batch_size

    def __len__(self) -> int:
        return self.nbatches

    def __iter__(self) -> Iterator[dict]:
        for _ in range(self.nbatches):
            yield self.sampler.sample()


class NegativeSamplingStrategy(nn.Module):
    """OpenKE/openke/module/strategy/NegativeSampling.py:13-32: ``forward(data)`` scores the batch, splits it into the
    positives [B, 1] and their negatives [B, neg], applies the loss and the optional L2 regulariser."""

    def __init__(self, model: TransH, loss: nn.Module, batch_size: int = 256, regul_rate: float = 0.0):
        super().__init__()
        self.model, self.loss, self.batch_size, self.regul_rate = model, loss, int(batch_size), float(regul_rate)

    def forward(self, data: dict) -> torch.Tensor:
        h, t, r = data["batch_h"], data["batch_t"], data["batch_r"]
        score = self.model._score(h, r, t, False)[4]
        if self.model.margin_flag:
            score = self.model.margin - score
        B = self.batch_size
        p_score = score[:B].view(-1, B).permute(1, 0)
        n_score = score[B:].view(-1, B).permute(1, 0)
        loss = self.loss(p_score, n_score)
        if self.regul_rate != 0:
            loss = loss + self.regul_rate * self.model.regularization(h, r, t)
        return loss


class Trainer:
    """OpenKE/openke/config/Trainer.py:43-99: ``run()`` builds the optimiser named by ``opt_method`` and makes
    ``train_times`` passes over the loader; returns the per-epoch summed losses.

    ``distributed=True`` (or an initialised ``torch.distributed`` with world size > 1 and ``distributed=None``) shards the
    TRIPLE BATCH across ranks (north_star: "by triple batch for TransH"): every rank draws the same global batch
    (same sampler seeds -- the device sampler is ~0.1 ms), scores / back-propagates only its contiguous slice of the
    positives together with their negatives, and the table gradients (ent | rel | norm_vector: 62.8 KB for the HOI label
    graph) are averaged with ONE all-reduce per step (they live in one flat bucket).  With equal shards the update equals
    the single-process update on the global batch up to fp32 summation order."""

    def __init__(self, model: nn.Module, data_loader, train_times: int = 1000, alpha: float = 0.5, opt_method: str = "sgd",
                 weight_decay: float = 0.0, lr_decay: float = 0.0, optimizer=None, graph: bool = False,
                 distributed: Optional[bool] = None, group=None, fused: bool = False):
        self.model, self.data_loader, self.train_times = model, data_loader, int(train_times)
        self.graph = bool(graph)      # replay whole epochs as one CUDA graph (sampler + forward + loss + backward + update)
        self.fused = bool(fused)      # the whole run as ONE persistent cooperative kernel (skg_kg_train_steps, csrc/kgtrain.cu)
        self.alpha, self.opt_method, self.weight_decay, self.lr_decay, self.optimizer = alpha, opt_method, weight_decay, lr_decay, optimizer
        self.group = group
        if distributed is None:
            distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        self.distributed = bool(distributed)
        self._shard_idx = None
        self._buckets = None

    # ---- data-parallel split of the triple batch -------------------------------------------------------------------
    def _setup_shards(self) -> None:
        world, rank = dist.get_world_size(self.group), dist.get_rank(self.group)
        B = int(self.data_loader.get_batch_size())
        if B % world:
            raise ValueError("batch size %d does not divide over %d ranks" % (B, world))
        self._global_bs, self._local_bs = B, B // world
        self._span = (rank * self._local_bs, (rank + 1) * self._local_bs)
        from .shard import GradBuckets
        self._buckets = GradBuckets([p for p in self.model.parameters() if p.requires_grad], group=self.group).attach()

    def _local(self, data: dict) -> dict:
        """This rank's slice of a global batch laid out [B positives | B negatives (round 1) | B negatives (round 2) | ...]
        (Base.cpp:161-197): positives lo..hi and, from every negative round, the corruptions of those positives."""
        lo, hi = self._span
        if self._shard_idx is None or self._shard_idx.device != data["batch_h"].device:
            rounds = data["batch_h"].numel() // self._global_bs
            ar = torch.arange(lo, hi, device=data["batch_h"].device)
            self._shard_idx = torch.cat([ar + k * self._global_bs for k in range(rounds)])
        return {k: (v[self._shard_idx] if torch.is_tensor(v) else v) for k, v in data.items()}

    def train_one_step(self, data: dict) -> torch.Tensor:
        if not self.distributed:
            self.optimizer.zero_grad(set_to_none=True)
            loss = self.model(data)
            loss.backward()
            self.optimizer.step()
            return loss.detach()
        self._buckets.zero_grad()
        old = getattr(self.model, "batch_size", None)
        if old is not None:
            self.model.batch_size = self._local_bs
        try:
            loss = self.model(self._local(data))
        finally:
            if old is not None:
                self.model.batch_size = old
        loss.backward()
        self._buckets.allreduce()                      # one flat bucket: ONE all-reduce of the table gradients
        self.optimizer.step()
        out = loss.detach().clone()
        dist.all_reduce(out, group=self.group)         # the reported loss is the global-batch mean
        return out / dist.get_world_size(self.group)

    def _graph_safe(self) -> bool:
        """Optimisers whose step captures correctly into a CUDA graph: SGD; Adagrad without lr decay (its decayed rate is
        computed on the host from the step count and would be frozen at capture); Adam only when built ``capturable``."""
        o = self.optimizer
        if isinstance(o, torch.optim.SGD):
            return True
        if isinstance(o, torch.optim.Adagrad):
            return all(g.get("lr_decay", 0) == 0 for g in o.param_groups)
        if isinstance(o, torch.optim.Adam):
            return all(g.get("capturable", False) for g in o.param_groups)
        return False

    def run(self) -> List[float]:
        if self.optimizer is None:
            m = self.opt_method.lower()
            p = self.model.parameters()
            if m == "adagrad":
                self.optimizer = torch.optim.Adagrad(p, lr=self.alpha, lr_decay=self.lr_decay, weight_decay=self.weight_decay)
            elif m == "adadelta":
                self.optimizer = torch.optim.Adadelta(p, lr=self.alpha, weight_decay=self.weight_decay)
            elif m == "adam":
                on_gpu = next(self.model.parameters()).is_cuda
                self.optimizer = torch.optim.Adam(p, lr=self.alpha, weight_decay=self.weight_decay,
                                                  capturable=bool(self.graph and on_gpu))
            else:
                self.optimizer = torch.optim.SGD(p, lr=self.alpha, weight_decay=self.weight_decay)
        if self.fused:
            why = self._fused_blocker()
            if why is None:
                return self._run_fused()
            import warnings
            warnings.warn("Trainer(fused=True): %s -- running the eager loop instead" % why)
        if self.distributed and self._buckets is None:
            self._setup_shards()
        if self.graph and next(self.model.parameters()).is_cuda:
            if self.distributed or not self._graph_safe():
                import warnings
                warnings.warn("Trainer(graph=True): %s -- running the eager loop instead"
                              % ("collectives are not captured" if self.distributed else
                                 "%s does not capture correctly with these settings" % type(self.optimizer).__name__))
            else:
                return self._run_graphed()
        return [float(self._epoch()) for _ in range(self.train_times)]

    # ---- the fused trainer ------------------------------------------------------------------------------------------
    def _fused_blocker(self) -> Optional[str]:
        """None when the run is the configuration skg_kg_train_steps implements, else the reason it is not."""
        st, o = self.model, self.optimizer
        if self.distributed:
            return "the fused kernel is single-GPU"
        if not isinstance(st, NegativeSamplingStrategy) or not isinstance(st.model, TransH) or not isinstance(self.data_loader, TrainDataLoader):
            return "it needs NegativeSamplingStrategy(TransH) over a TrainDataLoader"
        if not isinstance(st.loss, MarginLoss) or st.loss.adv_flag or st.regul_rate != 0 or st.model.margin_flag:
            return "it implements the plain MarginLoss without regulariser"
        if not isinstance(o, torch.optim.SGD) or any(g["momentum"] or g["weight_decay"] or g["nesterov"] or g["dampening"] or g.get("maximize")
                                                      for g in o.param_groups):
            return "it implements plain SGD"
        m = st.model
        tabs = (m.ent_embeddings.weight, m.rel_embeddings.weight, m.norm_vector.weight)
        if not all(t.is_cuda and t.dtype == torch.float32 and t.is_contiguous() for t in tabs):
            return "the tables must be contiguous fp32 CUDA tensors"
        if sum(t.numel() for t in tabs) * 8 > 200 * 1024:
            return "the tables and their gradients must fit in 200 KB of shared memory"
        if int(st.batch_size) != int(self.data_loader.get_batch_size()):
            return "the strategy's batch_size must be the loader's"
        return None

    def _run_fused(self) -> List[float]:
        """``train_times`` epochs of ``nbatches`` steps as one launch of the persistent cooperative kernel: the same batches
        (the sampler's LCG streams continue where they were and are handed back where they end), the same updates up to the
        fp32 summation order of the gradient atomics."""
        import ctypes
        from . import _lib
        st, smp = self.model, self.data_loader.sampler
        m, ix = st.model, smp.index
        dev = ix.device
        steps = self.train_times * len(self.data_loader)
        if steps == 0:
            return []
        a = _lib.KgTrainArgs()
        p = lambda t: t.data_ptr() if t is not None else None                        # noqa: E731
        a.head_hrt, a.tail_hrt = p(ix.head), p(ix.tail)
        a.lef_head, a.rig_head, a.lef_tail, a.rig_tail = p(ix.lef_head), p(ix.rig_head), p(ix.lef_tail), p(ix.rig_tail)
        a.left_mean, a.right_mean = (p(ix.left_mean), p(ix.right_mean)) if smp.bern else (None, None)
        a.train_total, a.ent_total, a.rel_total = ix.train_total, ix.ent_total, ix.rel_total
        seeds = torch.stack([smp._seeds[0], smp._seeds[1]]).contiguous()
        a.seeds = seeds.data_ptr()
        a.work_threads, a.neg_rate, a.mode = smp.W, smp.neg_rate, smp.mode
        a.dim, a.p_norm, a.norm_flag, a.num_steps = m.dim, int(m.p_norm), int(bool(m.norm_flag)), steps
        a.batch_size = smp.batch_size
        ent, rel, nrm = m.ent_embeddings.weight.data, m.rel_embeddings.weight.data, m.norm_vector.weight.data
        a.ent, a.rel, a.nrm = ent.data_ptr(), rel.data_ptr(), nrm.data_ptr()
        a.margin, a.lr = float(st.loss.margin), float(self.optimizer.param_groups[0]["lr"])
        loss = torch.empty(steps, dtype=torch.float32, device=dev)
        nbytes = int(_lib.lib().skg_kg_train_scratch_bytes(smp.batch_size, smp.neg_rate, ix.ent_total, ix.rel_total, m.dim))
        scratch = torch.empty(nbytes + 256, dtype=torch.uint8, device=dev)
        off = (-scratch.data_ptr()) % 256
        a.loss_out, a.scratch, a.scratch_bytes = loss.data_ptr(), scratch.data_ptr() + off, nbytes
        with torch.cuda.device(dev):
            _lib.call("skg_kg_train_steps", ctypes.byref(a), _lib.stream_ptr(dev))
        smp._seeds[0].copy_(seeds[steps & 1])
        for t in (m.ent_embeddings.weight, m.rel_embeddings.weight, m.norm_vector.weight):
            t.data.add_(0)                                       # bump the version counters: the tables changed in place
        per_epoch = loss.view(self.train_times, len(self.data_loader)).sum(1)
        return [float(v) for v in per_epoch.cpu().tolist()]

    def _epoch(self) -> torch.Tensor:
        res = torch.zeros((), device=next(self.model.parameters()).device)
        for data in self.data_loader:
            res = res + self.train_one_step(data)              # one host read per epoch, not per step (Trainer.py:57 .item())
        return res

    def _run_graphed(self) -> List[float]:
        """The step is ~25 small launches (0.6-0.8 ms of launch latency for ~0.1 ms of GPU work): capture whole epochs
        -- the device sampler included, its seeds and batch buffers are static device tensors -- into ONE CUDA graph and
        replay it.  The sampler ping-pongs between two seed buffers, so the graph spans an even number of steps and
        leaves the seeds where a replay expects them; the first epochs run eagerly on a side stream (allocator warm-up),
        a remainder that does not fill a graph runs eagerly at the end.  Same batches, same updates as the eager loop."""
        dev = next(self.model.parameters()).device
        period = 1 if len(self.data_loader) % 2 == 0 else 2
        warm = min(2 * period, self.train_times)
        losses: List[torch.Tensor] = []
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(warm):
                losses.append(self._epoch())
        cur.wait_stream(side)
        replays = (self.train_times - warm) // period
        if replays > 0:
            g = torch.cuda.CUDAGraph()
            self.optimizer.zero_grad(set_to_none=True)
            with torch.cuda.graph(g):
                static = [self._epoch() for _ in range(period)]
            for _ in range(replays):
                g.replay()
                losses.extend(l.clone() for l in static)
        for _ in range(self.train_times - warm - replays * period):
            losses.append(self._epoch())
        return [float(v) for v in torch.stack([l.reshape(()) for l in losses]).cpu().tolist()] if losses else []
