"""Multi-GPU host logic of the interaction head: one process per GPU, images sharded across ranks.

The path shards by IMAGE with no data-path collective -- every per-image quantity (pair grid, softmax over
that image's adjacency, LayerNorm per node, priors) is local to the image (reference head.py:822-982; the
reference shards with a ``DistributedSampler``, main.py:46-64).  Training adds one exchange step:

* the loss normalisers ``n_p`` (positives per loss) are averaged over ranks -- the reference does
  ``barrier(); all_reduce(n_p); n_p / world_size`` three times (head.py:167-172, 194-199, 223-228); here ONE
  all-reduce of a 2-float tensor carries them all (``average_normalisers``);
* the gradient all-reduce (NCCL over NVLink on the GPU box, gloo in the CPU tests): ``GradBuckets`` packs the
  parameters' gradients into flat buckets in reverse registration order (the order backward produces them),
  launches one asynchronous all-reduce per bucket and scatters the averaged values back.

Nothing here touches a kernel; the module works on any ``torch.distributed`` backend.
"""
from typing import Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def cells_of(n_h: Sequence[int], n: Sequence[int]) -> List[int]:
    """Grid cells per image, N2 = n_h * n: the unit the head's cost is linear in (SURVEY 8 notation)."""
    return [int(a) * int(b) for a, b in zip(n_h, n)]


def partition_images(cost: Sequence[int], world_size: int) -> List[List[int]]:
    """Splits image indices over ranks, balancing the summed ``cost`` (cells per image), not the image count:
    pairs grow quadratically with detections (SURVEY 8(e) note).  Longest-processing-time greedy, deterministic
    (ties -> lower image index, lower rank); every image lands on exactly one rank and each rank's list is
    sorted so the original relative order is preserved inside a shard."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    order = sorted(range(len(cost)), key=lambda i: (-int(cost[i]), i))
    load = [0] * world_size
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], len(shards[k]), k))
        shards[r].append(i)
        load[r] += int(cost[i])
    return [sorted(s) for s in shards]


def shard_of(items: Sequence, cost: Sequence[int], rank: Optional[int] = None, world_size: Optional[int] = None) -> list:
    """This rank's share of ``items`` (detections, targets, image shapes ...) under ``partition_images``."""
    if world_size is None:
        world_size = dist.get_world_size() if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    return [items[i] for i in partition_images(cost, world_size)[rank]]


def average_normalisers(n_p: torch.Tensor, group=None, in_place: bool = False) -> torch.Tensor:
    """Rank-mean of the loss normalisers (reference head.py:167-172: all_reduce then divide by the world size).
    One collective for all of them, no barrier; returns ``n_p`` unchanged when not distributed.  On NCCL the mean is the
    collective's own reduction (``ReduceOp.AVG``: no clone, no divide launch); ``in_place`` lets the caller give up ``n_p``."""
    if not (dist.is_available() and dist.is_initialized()):
        return n_p
    world = dist.get_world_size(group)
    if world == 1:
        return n_p
    out = n_p if in_place else n_p.clone()
    if out.is_cuda and dist.get_backend(group) == "nccl":
        dist.all_reduce(out, op=dist.ReduceOp.AVG, group=group)
        return out
    dist.all_reduce(out, group=group)
    return out.div_(world) if in_place else out / world


def max_over_ranks(values: Sequence[float], device=None, group=None) -> List[float]:
    """Element-wise maximum over ranks of a few host floats (bench.py: a step is as slow as its slowest rank)."""
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return [float(v) for v in t]


class GradBuckets:
    """Bucketed, asynchronous gradient averaging for the head's parameters (29.57 M fp32 -> 118 MB).

    Parameters are walked in REVERSE registration order (predictor / suppressor first, box_head last): that is
    the order backward finishes them, so the first buckets can be on the wire while the box_head GEMM gradients
    are still being computed.  ``bucket_bytes`` is sized for launch latency, not link count (NVSwitch gives every
    GPU full bandwidth to every peer): the default 32 MiB yields four buckets for the head.
    """

    def __init__(self, params: Iterable[torch.nn.Parameter], bucket_bytes: int = 32 << 20, group=None):
        self.group = group
        self.params = [p for p in params if p.requires_grad][::-1]
        self.buckets: List[List[torch.nn.Parameter]] = []
        cur, size = [], 0
        for p in self.params:
            nb = p.numel() * p.element_size()
            if cur and (size + nb > bucket_bytes or p.dtype != cur[0].dtype or p.device != cur[0].device):
                self.buckets.append(cur); cur, size = [], 0
            cur.append(p); size += nb
        if cur:
            self.buckets.append(cur)
        self._flat = [None] * len(self.buckets)
        self._work = []

    @property
    def world(self) -> int:
        return dist.get_world_size(self.group) if dist.is_initialized() else 1

    def _pack(self, b: int) -> torch.Tensor:
        ps = self.buckets[b]
        n = sum(p.numel() for p in ps)
        flat = self._flat[b]
        if flat is None or flat.numel() != n:
            flat = self._flat[b] = torch.empty(n, dtype=ps[0].dtype, device=ps[0].device)
        o = 0
        for p in ps:
            k = p.numel()
            if p.grad is None:
                flat[o:o + k].zero_()
            else:
                flat[o:o + k].copy_(p.grad.reshape(-1))
            o += k
        return flat

    # -- attached mode: the gradients LIVE in the flat buckets -----------------------------------------------------
    def attach(self, flatten_params: bool = False, module=None) -> "GradBuckets":
        """Makes every parameter's ``.grad`` a view into its bucket's flat buffer (autograd accumulates into an existing
        ``.grad`` in place), so an averaging step is one all-reduce + one scale per BUCKET -- no per-parameter pack / unpack
        (~400 small launches for the head).  Use ``zero_grad()`` of this object instead of the optimiser's
        ``zero_grad(set_to_none=True)``, which would drop the views (``start`` then falls back to packing).

        ``flatten_params``: the parameters' storage moves into one flat buffer per bucket as well (each ``p`` keeps its
        shape and its place in the module / ``state_dict``), and ``flat_params`` holds one ``nn.Parameter`` per bucket whose
        ``.grad`` is the bucket's gradient buffer: an optimiser built over ``flat_params`` updates the whole head with one
        (fused) kernel per bucket instead of ~400 per-tensor updates -- element-wise optimisers (SGD, Adam) give the same
        result.  ``module`` (the interaction head) is told about the flat buffers so that its packed operands follow the
        flat tensors' version counters."""
        self._views = []
        self.flat_params = []
        al = lambda k: (k + 63) // 64 * 64            # noqa: E731  every tensor starts 256-byte aligned (float4 / TMA readers)
        for b, ps in enumerate(self.buckets):
            n = sum(al(p.numel()) for p in ps)
            flat = self._flat[b] = torch.zeros(n, dtype=ps[0].dtype, device=ps[0].device)
            fp = None
            if flatten_params:
                fp = torch.nn.Parameter(torch.zeros(n, dtype=ps[0].dtype, device=ps[0].device))
                fp.grad = flat
                self.flat_params.append(fp)
            o = 0
            for p in ps:
                k = p.numel()
                v = flat[o:o + k].view_as(p)
                if p.grad is not None:
                    v.copy_(p.grad)
                p.grad = v
                self._views.append((p, v.data_ptr()))
                if fp is not None:
                    pv = fp.data[o:o + k].view_as(p)
                    pv.copy_(p.data)
                    p.data = pv
                o += al(k)
        if flatten_params and module is not None:
            module.__dict__["_skg_flat_params"] = self.flat_params
        return self

    def zero_grad(self) -> None:
        if not self._attached():
            for p in self.params:
                p.grad = None
            return
        for flat in self._flat:
            flat.zero_()

    def _attached(self, full: bool = False) -> bool:
        """Are the gradients still the bucket views?  Checked on a sample of the parameters per call (every step asks
        twice; the full walk over ~400 tensors is kept for ``full=True``)."""
        views = getattr(self, "_views", None)
        if not views:
            return False
        if not full:
            step = max(len(views) // 8, 1)
            views = views[::step] + views[-1:]
        return all(p.grad is not None and p.grad.data_ptr() == ptr for p, ptr in views)

    def start(self) -> None:
        """Launches every bucket's all-reduce (async); packs the buckets first unless the gradients already live there."""
        self._work = []
        self._direct = False
        if self.world == 1:
            return
        self._direct = self._attached()
        for b in range(len(self.buckets)):
            flat = self._flat[b] if self._direct else self._pack(b)
            self._work.append(dist.all_reduce(flat, group=self.group, async_op=True))

    def finish(self) -> None:
        """Waits for the buckets in launch order and writes the rank-mean back into ``p.grad``."""
        if self.world == 1:
            return
        inv = 1.0 / self.world
        if self._direct:
            for b, w in enumerate(self._work):
                w.wait()
                self._flat[b].mul_(inv)
            self._work = []
            return
        for b, w in enumerate(self._work):
            w.wait()
            flat = self._flat[b]
            o = 0
            for p in self.buckets[b]:
                k = p.numel()
                g = flat[o:o + k].view_as(p) * inv
                if p.grad is None:
                    p.grad = g.clone()
                else:
                    p.grad.copy_(g)
                o += k
        self._work = []

    def allreduce(self) -> None:
        self.start(); self.finish()
