"""Training pass of the interaction head on the native kernels: BASELINE configs[2], SURVEY 8(b) "fwd/bwd", 8(e).

The reference trains the head with autograd (utils.py:213-229: forward, ``loss.backward()``, optimiser step; losses
head.py:153-235).  Here one training step is three C-ABI calls around the loss bookkeeping:

    skg_head_forward(SKG_HEAD_TRAIN)   the inference chain, keeping what the backward needs (csrc/forward.cu)
    skg_head_post + loss kernels        result dicts, GT association, focal losses, TransH margin loss (unchanged)
    skg_head_backward                   every dX / dW / db of the chain (csrc/backward.cu, csrc/trainops.cu)

and two descriptor-driven block-copy launches (``skg_block_copy``): packing the reference's per-branch fp32 parameters
into the concatenated bf16 operands (+ the transposed copies the dX GEMMs read) once per optimiser step, and unpacking
the packed fp32 gradients into the per-branch ``.grad`` tensors.  PyTorch is the plumbing: it owns the parameters, the
``.grad`` buffers and the optimiser, and a one-node autograd Function makes ``loss.backward()`` run the native backward
(the gradients are ACCUMULATED into ``p.grad`` like autograd does; ``transH_loss`` keeps its own tape through the TransH
tables).  Gradients of the head's parameters w.r.t. ``hoi_loss`` and ``interactiveness_loss``; the TransH entity rows
that enter fc_head / fc_tail are constants of the step (the reference re-draws the tables on every call, head.py:574-578).
"""
import ctypes
import os
import weakref
from typing import List, Optional

import numpy as np
import torch
from torch import Tensor

from . import _lib

_SRC_BF16, _DST_BF16, _ACC, _SCALE = 1, 2, 4, 8


class BlockDesc(ctypes.Structure):
    """Mirror of ``skg_block_desc``."""
    _fields_ = [("src", ctypes.c_void_p), ("dst", ctypes.c_void_p), ("dst_t", ctypes.c_void_p),
                ("rows", ctypes.c_int32), ("cols", ctypes.c_int32),
                ("ld_src", ctypes.c_int64), ("ld_dst", ctypes.c_int64), ("ld_dst_t", ctypes.c_int64),
                ("flags", ctypes.c_int32), ("scale", ctypes.c_float)]


class HeadWeightsT(ctypes.Structure):
    """Mirror of ``skg_head_weights_t``."""
    _fields_ = [(k, ctypes.c_void_p) for k in ("Wb2T", "Wsp1T", "Wsp2T", "WfcT", "W_hprojT", "W_oprojT", "W2catT", "W3attT",
                                               "W3osT", "W3soT", "W3gT", "WpredT")]


_GRAD_FIELDS = ("Wb1", "bb1", "Wb2", "bb2", "Wsp0", "bsp0", "Wsp1", "bsp1", "Wsp2", "bsp2", "Wfc", "bfc", "W_hproj", "b_hproj",
                "W_oproj", "b_oproj", "Wg1", "bg1", "W2cat", "b2cat", "W3att", "b3att", "W3os", "b3os", "W3so", "b3so", "W3g",
                "b3g", "w_adj", "b_adj", "ln_h_w", "ln_h_b", "ln_o_w", "ln_o_b", "Wpred", "bpred")


# gradient groups in the order skg_head_backward_staged finishes them: the packed gradient buffer is laid out group by group,
# so every group is ONE contiguous range to all-reduce as soon as its stage event has fired
_GRAD_STAGES = (("Wpred", "bpred", "W3g", "b3g", "Wg1", "bg1", "ln_h_w", "ln_h_b", "ln_o_w", "ln_o_b", "W3os", "b3os", "W3so", "b3so",
                 "w_adj", "b_adj", "W3att", "b3att"),
                ("W_hproj", "b_hproj", "W_oproj", "b_oproj", "Wfc", "bfc", "Wb2", "bb2", "Wb1", "bb1"),
                ("W2cat", "b2cat"),
                ("Wsp2", "bsp2", "Wsp1", "bsp1", "Wsp0", "bsp0"))
assert sorted(k for g in _GRAD_STAGES for k in g) == sorted(_GRAD_FIELDS)


class HeadGrads(ctypes.Structure):
    """Mirror of ``skg_head_grads``."""
    _fields_ = [(k, ctypes.c_void_p) for k in _GRAD_FIELDS]


class HeadLossArgs(ctypes.Structure):
    """Mirror of ``skg_head_loss_args``."""
    _fields_ = [("tables", ctypes.POINTER(_lib.PostTables)), ("gt_labels", ctypes.c_void_p), ("n_p", ctypes.c_void_p),
                ("g_hoi", ctypes.c_void_p), ("g_int", ctypes.c_void_p), ("prior_power", ctypes.c_float)]


class _HadamardBwd(ctypes.Structure):
    """Mirror of ``skg_hadamard_bwd`` (only the tests call the kernel on its own)."""
    _fields_ = [("S", ctypes.c_void_p), ("ld_s", ctypes.c_int64),
                ("dP_att", ctypes.c_void_p), ("dP_fin", ctypes.c_void_p), ("dP_os", ctypes.c_void_p), ("dP_so", ctypes.c_void_p),
                ("dP_g", ctypes.c_void_p), ("ld_dp", ctypes.c_int64),
                ("hproj", ctypes.c_void_p), ("oproj", ctypes.c_void_p), ("U2h", ctypes.c_void_p), ("U2o", ctypes.c_void_p),
                ("GT", ctypes.c_void_p), ("dS", ctypes.c_void_p), ("ld_ds", ctypes.c_int64),
                ("dhproj", ctypes.c_void_p), ("doproj", ctypes.c_void_p), ("dU2h", ctypes.c_void_p), ("dU2o", ctypes.c_void_p),
                ("dGT", ctypes.c_void_p), ("box_off", ctypes.c_void_p), ("hum_off", ctypes.c_void_p), ("cell_off", ctypes.c_void_p),
                ("num_images", ctypes.c_int32), ("R", ctypes.c_int32)]


class _BlockTable:
    """A list of 2-D block copies executed as ONE ``skg_block_copy`` launch (descriptor + tile tables live on the device)."""

    def __init__(self, device):
        self.device = device
        self.descs: List[BlockDesc] = []
        self.keep = []
        self.d_descs = self.d_tiles = None
        self.num_tiles = 0

    def add(self, src: Tensor, dst: Optional[Tensor], dst_t: Optional[Tensor] = None, flags: int = 0, scale: float = 1.0):
        """src [rows, cols] -> dst [rows, cols] (+ dst_t [cols, rows]); 2-D views with unit inner stride."""
        assert src.dim() == 2 and src.stride(1) == 1 and (dst is None or (dst.shape == src.shape and dst.stride(1) == 1))
        d = BlockDesc()
        d.src, d.dst = src.data_ptr(), (dst.data_ptr() if dst is not None else None)
        d.rows, d.cols = int(src.shape[0]), int(src.shape[1])
        d.ld_src, d.ld_dst = int(src.stride(0)), (int(dst.stride(0)) if dst is not None else 0)
        if dst_t is not None:
            assert tuple(dst_t.shape) == (src.shape[1], src.shape[0]) and dst_t.stride(1) == 1 and dst_t.dtype == torch.bfloat16
            d.dst_t, d.ld_dst_t = dst_t.data_ptr(), int(dst_t.stride(0))
        f = flags
        if src.dtype == torch.bfloat16:
            f |= _SRC_BF16
        if dst is not None and dst.dtype == torch.bfloat16:
            f |= _DST_BF16
        if scale != 1.0:
            f |= _SCALE
        d.flags, d.scale = f, float(scale)
        self.descs.append(d)
        self.keep.append((src, dst, dst_t))

    def finalize(self):
        n = len(self.descs)
        arr = (BlockDesc * n)(*self.descs)
        raw = np.frombuffer(arr, dtype=np.uint8).copy()
        self.d_descs = torch.from_numpy(raw).to(self.device)
        tiles = []
        for i, d in enumerate(self.descs):
            rr = np.arange(0, d.rows, 32, dtype=np.int32); cc = np.arange(0, d.cols, 32, dtype=np.int32)
            g = np.stack(np.meshgrid(rr, cc, indexing="ij"), -1).reshape(-1, 2)
            tiles.append(np.concatenate([np.full((len(g), 1), i, np.int32), g], 1))
        t = np.concatenate(tiles).astype(np.int32)
        self.num_tiles = int(t.shape[0])
        self.d_tiles = torch.from_numpy(np.ascontiguousarray(t)).to(self.device)
        return self

    def run(self):
        with _lib.on_device(self.device):
            _lib.call("skg_block_copy", _lib.ptr(self.d_descs), _lib.ptr(self.d_tiles), self.num_tiles, _lib.stream_ptr(self.device))


def _v(t: Tensor) -> Tensor:
    """1-D tensor as a [1, n] row block."""
    return t.reshape(1, -1)


class TrainState:
    """Everything the native training step keeps per head: the packed bf16 operands (shared layout with the inference
    engine), their transposed copies, the packed fp32 gradient buffer and the two block-copy tables.

    ``refresh()`` re-packs after an optimiser step (one launch, detected through the parameters' version counters;
    ``force=True`` after writes the counters cannot see).  ``grads_into_params()`` adds the packed gradients of the last
    backward into the per-branch ``p.grad`` tensors (allocated as views of one flat buffer when missing)."""

    def __init__(self, ih):
        from . import engine
        self.ih = weakref.ref(ih)
        self.pk = engine.PackedHead(ih)                 # initial pack through the reference-layout tensors
        pk, G = self.pk, ih.box_pair_head
        dev = pk.device
        self.device = dev
        R, E = pk.R, pk.E
        bf = lambda *shape: torch.zeros(*shape, dtype=torch.bfloat16, device=dev)   # noqa: E731
        self.T = dict(Wb2T=bf(E, E), Wsp1T=bf(128, 256), Wsp2T=bf(256, R), WfcT=bf(pk.fc_k, 2 * R), W_hprojT=bf(R, 2 * R),
                      W_oprojT=bf(R, 2 * R), W2catT=bf(R, 4 * R), W3attT=bf(R, R), W3osT=bf(R, R), W3soT=bf(R, R), W3gT=bf(R, R),
                      WpredT=bf(2 * R, pk.Wpred.shape[0]))
        self.wt = HeadWeightsT()
        for k, t in self.T.items():
            setattr(self.wt, k, t.data_ptr())
        # packed fp32 gradients: one flat buffer, regions in the layout of the packed operands
        shapes = dict(Wb1=pk.Wb1.shape, bb1=(E,), Wb2=pk.Wb2.shape, bb2=(E,), Wsp0=pk.Wsp0.shape, bsp0=(128,), Wsp1=pk.Wsp1.shape,
                      bsp1=(256,), Wsp2=pk.Wsp2.shape, bsp2=(R,), Wfc=pk.Wfc.shape, bfc=(2 * R,), W_hproj=pk.W_hproj.shape,
                      b_hproj=(2 * R,), W_oproj=pk.W_oproj.shape, b_oproj=(2 * R,), Wg1=pk.Wg1.shape, bg1=(R,), W2cat=pk.W2cat.shape,
                      b2cat=(4 * R,), W3att=(R, R), b3att=(R,), W3os=(R, R), b3os=(R,), W3so=(R, R), b3so=(R,), W3g=(R, R), b3g=(R,),
                      w_adj=(R,), b_adj=(4,), ln_h_w=(R,), ln_h_b=(R,), ln_o_w=(R,), ln_o_b=(R,), Wpred=pk.Wpred.shape,
                      bpred=(pk.Wpred.shape[0],))
        total = sum((int(np.prod(s)) + 63) // 64 * 64 for s in shapes.values()) + 256
        self.gflat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.g = {}
        self.cg = HeadGrads()
        o = 0
        self.stage_ranges = []                          # (begin, end) element ranges of the gradient groups inside gflat
        for group in _GRAD_STAGES:
            o0 = o
            for k in group:
                n = int(np.prod(shapes[k]))
                self.g[k] = self.gflat[o:o + n].view(*shapes[k])
                setattr(self.cg, k, self.g[k].data_ptr())
                o += (n + 63) // 64 * 64
            self.stage_ranges.append((o0, o))
        self.dist = None                                # set by enable_gradient_allreduce
        self.graph = True                               # forward / backward launch sequences replayed as CUDA graphs when their
        #                                                 keys repeat (fixed batch geometry, stable buffers); else plain launches
        self.loss_scalars = torch.zeros(4, dtype=torch.float32, device=dev)
        self._gt_buf = None
        # side stream of the pre-pass (labels, normalisers, TransH loss: geometry + targets only) and its two events
        self.side = torch.cuda.Stream(dev, priority=-1)          # high priority: its small kernels slot in between the GEMMs
        self.ev_maps, self.ev_pre = torch.cuda.Event(), torch.cuda.Event()
        self.ev_maps.record(torch.cuda.current_stream(dev))      # materialises the cudaEvent_t handle
        self.prepass = True
        self._build_pack_table(ih)
        self._unpack = None
        self._grad_key = None
        self._grad_skey = None
        self.params = list(ih.parameters())
        self.vsum = -1
        self.refresh(force=True)                        # fills the transposed copies (and re-derives the operands)
        self.anchor = torch.zeros(1, device=dev, requires_grad=True)

    def gt_buffer(self, pairs: int, num_cls: int) -> Tensor:
        """Persistent [pairs, num_cls] label buffer (a stable address for the backward's graph key)."""
        n = max(pairs, 1) * num_cls
        if self._gt_buf is None or self._gt_buf.numel() < n:
            self._gt_buf = torch.empty(int(n * 1.25) + 64, dtype=torch.float32, device=self.device)
        return self._gt_buf[:pairs * num_cls].view(pairs, num_cls)

    # ---- parameter -> operand map (the inverse, operand gradient -> parameter gradient, is built from the same list) ----
    def _blocks(self, ih):
        """Yields (parameter view [rows, cols], operand name, row offset, col offset, transposed-operand name or None,
        kind) with kind 'w' (bf16 operand), 'f' (fp32 operand / bias), 'bsum' (fp32, summed over the branches)."""
        G = ih.box_pair_head
        pk = self.pk
        R, E = pk.R, pk.E
        out = []
        w = lambda p, name, r0=0, c0=0, tname=None: out.append((p, name, r0, c0, tname, "w"))      # noqa: E731
        f = lambda p, name, c0=0, kind="f": out.append((_v(p), name, 0, c0, None, kind))            # noqa: E731
        w(G.box_head[1].weight, "Wb1"); f(G.box_head[1].bias, "bb1")
        w(G.box_head[3].weight, "Wb2", tname="Wb2T"); f(G.box_head[3].bias, "bb2")
        out.append((G.spatial_head[0].weight, "Wsp0", 0, 0, None, "f")); f(G.spatial_head[0].bias, "bsp0")
        w(G.spatial_head[2].weight, "Wsp1", tname="Wsp1T"); f(G.spatial_head[2].bias, "bsp1")
        w(G.spatial_head[4].weight, "Wsp2", tname="Wsp2T"); f(G.spatial_head[4].bias, "bsp2")
        fh, ft = G.fc_head[0], G.fc_tail[0]
        w(fh.weight[:, :E], "Wfc", 0, 0, "WfcT")
        out.append((fh.weight[:, E:], "Wfh_ent", 0, 0, None, "f"))
        f(fh.bias, "bfh")
        w(ft.weight, "Wfc", R, 0, "WfcT"); f(ft.bias, "bfc", R)
        sub = R // G.attention_head.cardinality
        for c in range(G.attention_head.cardinality):
            a1 = G.attention_head.fc_1[c]
            w(a1.weight[:, :E], "W_hproj", c * sub, 0, "W_hprojT"); w(a1.weight[:, E:], "W_oproj", c * sub, 0, "W_oprojT")
            f(a1.bias, "b_oproj", c * sub)
            w(G.sub_to_obj.fc_1[c].weight, "W_hproj", R + c * sub, 0, "W_hprojT"); f(G.sub_to_obj.fc_1[c].bias, "b_hproj", R + c * sub)
            w(G.obj_to_sub.fc_1[c].weight, "W_oproj", R + c * sub, 0, "W_oprojT"); f(G.obj_to_sub.fc_1[c].bias, "b_oproj", R + c * sub)
            w(G.attention_head_g.fc_1[c].weight, "Wg1", c * sub, 0); f(G.attention_head_g.fc_1[c].bias, "bg1", c * sub)
            for gi, m in enumerate((G.attention_head, G.obj_to_sub, G.sub_to_obj, G.attention_head_g)):
                w(m.fc_2[c].weight, "W2cat", gi * R + c * sub, 0, "W2catT"); f(m.fc_2[c].bias, "b2cat", gi * R + c * sub)
            for m, nm in ((G.attention_head, "att"), (G.obj_to_sub, "os"), (G.sub_to_obj, "so"), (G.attention_head_g, "g")):
                w(m.fc_3[c].weight, "W3" + nm, 0, c * sub, "W3" + nm + "T"); f(m.fc_3[c].bias, "b3" + nm, 0, "bsum")
        f(G.adjacency.weight.reshape(-1), "w_adj"); f(G.adjacency.bias, "b_adj")
        f(G.norm_h.weight, "ln_h_w"); f(G.norm_h.bias, "ln_h_b"); f(G.norm_o.weight, "ln_o_w"); f(G.norm_o.bias, "ln_o_b")
        nc = pk.num_cls
        w(ih.box_pair_predictor.weight, "Wpred", 0, 0, "WpredT"); f(ih.box_pair_predictor.bias, "bpred", 0)
        w(ih.box_pair_suppressor.weight, "Wpred", nc, 0, "WpredT"); f(ih.box_pair_suppressor.bias, "bpred", nc)
        return out

    def _operand(self, name: str) -> Optional[Tensor]:
        pk = self.pk
        if name in ("ln_h_w", "ln_h_b", "ln_o_w", "ln_o_b"):
            return _v(pk.ln[("ln_h_w", "ln_h_b", "ln_o_w", "ln_o_b").index(name)])
        if name == "b_adj":
            return None                                  # a host float of the descriptor: softmax is shift invariant
        t = getattr(pk, name)
        return t if t.dim() > 1 else _v(t)

    def _build_pack_table(self, ih):
        tab = _BlockTable(self.device)
        self._bsum = []
        for p, name, r0, c0, tname, kind in self._blocks(ih):
            dst_full = self._operand(name)
            if dst_full is None:
                continue
            src = p.detach()
            rows, cols = src.shape
            dst = dst_full[r0:r0 + rows, c0:c0 + cols]
            if dst.data_ptr() == src.data_ptr():
                continue                                 # the operand aliases the parameter (fp32 biases, LayerNorm affine)
            dst_t = self.T[tname][c0:c0 + cols, r0:r0 + rows] if tname else None
            if kind == "bsum":
                if not any(d.data_ptr() == dst_full.data_ptr() for d in self._bsum):     # (_operand hands out a fresh view per call)
                    self._bsum.append(dst_full)
                tab.add(src, dst, None, flags=_ACC)
            else:
                tab.add(src, dst, dst_t)
        self.pack = tab.finalize()

    def refresh(self, force: bool = False) -> None:
        from . import engine
        flats = self.ih().__dict__.get("_skg_flat_params") if self.ih() is not None else None
        # parameters living in flat buffers (shard.GradBuckets.attach(flatten_params=True)): an optimiser step on the flat
        # tensors bumps THEIR version counters -- three reads instead of ~400
        vs = (engine._version_sum(flats) if flats else engine._version_sum(self.params), engine._OPT_STEPS[0])
        if not force and vs == self.vsum:
            return
        for t in self._bsum:
            t.zero_()
        self.pack.run()
        self.vsum = vs

    # ---- gradients --------------------------------------------------------------------------------------------------
    def _ensure_unpack(self):
        params = self.params
        # fast path: the unpack table stands while every .grad is the very tensor object it was built for (identity walk,
        # no data_ptr calls) and a sample of them still sits at the same address
        sample = params[::max(len(params) // 8, 1)] + params[-1:]
        if self._unpack is not None and self._grad_skey is not None:
            objs = self._grad_objs
            if all(p.grad is g for p, g in zip(params, objs)) and \
                    tuple(p.grad.data_ptr() if p.grad is not None else 0 for p in sample) == self._grad_skey:
                return
        missing = [p for p in params if p.requires_grad and p.grad is None]
        if missing:
            flat = torch.zeros(sum((p.numel() + 63) // 64 * 64 for p in missing), dtype=torch.float32, device=self.device)
            o = 0
            for p in missing:
                p.grad = flat[o:o + p.numel()].view_as(p)
                o += (p.numel() + 63) // 64 * 64
        key = tuple(p.grad.data_ptr() if p.grad is not None else 0 for p in params)
        self._grad_skey = tuple(p.grad.data_ptr() if p.grad is not None else 0 for p in sample)
        self._grad_objs = [p.grad for p in params]
        if self._unpack is not None and key == self._grad_key:
            return
        ih = self.ih()
        tab = _BlockTable(self.device)
        scale = 1.0 / self.dist.world if self.dist is not None else 1.0      # rank mean of the all-reduced packed gradients
        self._unpack_scale = scale
        for p, name, r0, c0, tname, kind in self._blocks(ih):
            base = p._base if p._base is not None else p
            if not getattr(base, "requires_grad", False) or base.grad is None:
                continue
            if name in ("Wfh_ent", "bfh"):               # fc_head bias / entity columns: both come from dbfc[:R]
                if name == "bfh":
                    tab.add(self.g["bfc"].reshape(1, -1)[:, :p.shape[1]], base.grad.reshape(1, -1), None, flags=_ACC, scale=scale)
                continue
            gsrc = self.g[name] if self.g[name].dim() > 1 else _v(self.g[name])
            rows, cols = p.shape
            src = gsrc[r0:r0 + rows, c0:c0 + cols]
            # the parameter view p (a slice / reshape of the parameter) maps onto the same slice of its .grad
            gview = base.grad.as_strided(p.shape, p.stride(), base.grad.storage_offset() + p.storage_offset() - base.storage_offset())
            tab.add(src, gview, None, flags=_ACC, scale=scale)
        self._unpack = tab.finalize()
        self._grad_key = key

    def grads_into_params(self) -> None:
        """p.grad += packed gradients of the last skg_head_backward (one block-copy launch + the rank-1 entity block)."""
        self._ensure_unpack()
        self._unpack.run()
        ih = self.ih()
        G = ih.box_pair_head
        fh = G.fc_head[0].weight
        if fh.requires_grad and fh.grad is not None:
            ent = G.transh_head.transh.ent_embeddings.weight.detach()
            # d fc_head.weight[:, E:] = sum_rows g_row (x) E[human_idx] = dbias (x) E[human_idx]   (head_stack rows are constant)
            fh.grad[:, self.pk.E:].addmm_(self.g["bfc"][:self.pk.R].unsqueeze(1), ent[G.human_idx].float().unsqueeze(0),
                                          alpha=getattr(self, "_unpack_scale", 1.0))


class _GradAllReduce:
    """Overlapped gradient averaging of the native training step (the reference trains under DDP, utils.py:200-229): one
    all-reduce per gradient group of skg_head_backward_staged on the PACKED gradient buffer (identical layout on every rank),
    started behind the group's stage event on a side stream while the backward continues; the unpack that follows scales by
    1 / world.  NCCL on the box, gloo in the CPU tests of the host logic."""

    def __init__(self, ts: "TrainState", group=None):
        import torch.distributed as dist
        self.group = group
        self.world = dist.get_world_size(group)
        dev = ts.device
        self.side = torch.cuda.Stream(dev)
        self.events = [torch.cuda.Event() for _ in range(len(ts.stage_ranges))]
        for e in self.events:
            e.record(torch.cuda.current_stream(dev))         # materialises the cudaEvent_t handle
        self.event_array = (ctypes.c_void_p * len(self.events))(*[e.cuda_event for e in self.events])
        self.work = []
        # the collectives go straight through the process group with cached views of the packed buffer: the python of
        # dist.all_reduce (argument checks, group lookup, a fresh slice) was ~40 us per group of a 2.5 ms step
        self._views = None
        self._gflat_ptr = 0
        try:
            if os.environ.get("SKG_DIST_LEGACY_ISSUE"):      # diagnostics: the plain dist.all_reduce issue path (A/B runs)
                raise RuntimeError("legacy issue path requested")
            self._pg = group if group is not None else dist.distributed_c10d._get_default_group()
        except Exception:                                    # an older / newer torch without the private accessor
            self._pg = None

    def launch(self, ts: "TrainState") -> None:
        import torch.distributed as dist
        self.work = []
        if os.environ.get("SKG_DIST_SYNC"):                  # diagnostics: no overlap, collectives after the whole backward
            torch.cuda.synchronize(ts.device)
        if self._views is None or self._gflat_ptr != ts.gflat.data_ptr():
            self._views = [ts.gflat[a:b] for a, b in ts.stage_ranges]
            self._gflat_ptr = ts.gflat.data_ptr()
        pg = self._pg
        with torch.cuda.stream(self.side):
            for ev, view in zip(self.events, self._views):
                self.side.wait_event(ev)
                if pg is not None:
                    self.work.append(pg.allreduce([view]))
                else:
                    self.work.append(dist.all_reduce(view, group=self.group, async_op=True))
        for w in self.work:
            w.wait()                                         # the current stream waits for the collective, not the host
        self.work = []
        ts.gflat.record_stream(self.side)


def enable_gradient_allreduce(ih, group=None) -> None:
    """Data-parallel training of the head: from the next backward on, the packed parameter gradients are averaged over the
    ranks of ``group`` INSIDE ``loss.backward()`` (overlapped with it) -- do not all-reduce the head's ``p.grad`` again."""
    ts = train_state(ih)
    ts.dist = _GradAllReduce(ts, group)
    ts._unpack = None                                        # rebuild the unpack table with the 1 / world scale
    ts._grad_key = ts._grad_skey = None


def train_state(ih) -> TrainState:
    ts = ih.__dict__.get("_skg_train")
    if ts is None or ts.device != next(ih.parameters()).device or ts.params[0].data_ptr() != next(ih.parameters()).data_ptr():
        ts = TrainState(ih)
        ih.__dict__["_skg_train"] = ts
    else:
        ts.refresh()
    return ts


class _StepRecord:
    """What ``loss.backward()`` needs from the forward of the same step."""

    def __init__(self, ts, run, b, flags, ws, tables, gt, n_p, keep):
        self.ts, self.run, self.b, self.flags, self.ws, self.tables, self.gt, self.n_p, self.keep = ts, run, b, flags, ws, tables, gt, n_p, keep
        self.done = False

    def backward(self, g_hoi: Optional[Tensor], g_int: Optional[Tensor]) -> None:
        ts = self.ts
        dev = ts.device
        zero = None
        if g_hoi is None or g_int is None:
            zero = torch.zeros(1, dtype=torch.float32, device=dev)
        gh = (g_hoi if g_hoi is not None else zero).reshape(-1).float()
        gi = (g_int if g_int is not None else zero).reshape(-1).float()
        # normalisers and upstream gradients into ONE persistent 4-float buffer: stable addresses for the backward's graph key
        sc = ts.loss_scalars
        torch.cat([self.n_p.reshape(-1), gh, gi], out=sc)
        la = HeadLossArgs()
        la.tables = ctypes.pointer(self.tables)
        la.gt_labels, la.n_p, la.g_hoi, la.g_int, la.prior_power = self.gt.data_ptr(), sc.data_ptr(), sc[2:].data_ptr(), sc[3:].data_ptr(), 1.0
        bflags = self.flags | (_lib.HEAD_GRAPH if ts.graph else 0)
        dd = ts.dist
        with _lib.on_device(dev):
            ts.gflat.zero_()
            if dd is None:
                _lib.call("skg_head_backward_staged", ctypes.byref(ts.pk.cw), ctypes.byref(ts.wt), ctypes.byref(self.b), ctypes.byref(la),
                          ctypes.byref(ts.cg), bflags, _lib.ptr(self.ws), self.ws.numel(), _lib.stream_ptr(dev), None, 0)
            else:
                # data parallel: the packed gradient groups are all-reduced as the backward finishes them (stage events recorded
                # inside the C call; the collectives are issued behind them on a side stream) and averaged by the unpack's scale
                _lib.call("skg_head_backward_staged", ctypes.byref(ts.pk.cw), ctypes.byref(ts.wt), ctypes.byref(self.b), ctypes.byref(la),
                          ctypes.byref(ts.cg), bflags, _lib.ptr(self.ws), self.ws.numel(), _lib.stream_ptr(dev), dd.event_array,
                          len(dd.events))
                dd.launch(ts)
        ts.grads_into_params()
        self.done = True


class _HeadLossFn(torch.autograd.Function):
    """One autograd node for the whole head: forward hands out the two loss values the kernels computed, backward runs
    skg_head_backward and ADDS the parameter gradients into ``p.grad`` (so the node itself has no differentiable inputs
    besides a dummy anchor)."""

    @staticmethod
    def forward(ctx, anchor, hoi, inter, rec):
        ctx.rec = rec
        return hoi.clone(), inter.clone()

    @staticmethod
    def backward(ctx, g_hoi, g_int):
        ctx.rec.backward(g_hoi, g_int)
        return None, None, None, None


def _prepass(ih, ts: "TrainState", run, boxes_cat: Tensor, labels_dev: Tensor, scores_dev: Tensor, targets) -> dict:
    """Everything of the training step that depends on the batch geometry and the targets only -- GT association
    (head.py:703-719), the loss normalisers (head.py:167-205, with their rank average), TransH positive / negative selection
    and the margin loss (head.py:207-235, 926-963) -- on a side stream behind the pair-enumeration event of the forward, so
    it runs UNDER the GEMM chain.  The one host synchronisation of the step (the selection's sizes) then waits for this side
    stream only: the main stream keeps the whole forward queued, and the rank average of the normalisers has the forward
    as cover instead of sitting between the loss and the backward."""
    from . import engine
    G = ih.box_pair_head
    dev = ts.device
    cur = torch.cuda.current_stream(dev)
    tables = engine._run_post_tables(G, run, labels_dev, scores_dev)
    Pn = int(run["seg"].total_pairs)
    with torch.cuda.stream(ts.side):
        ts.side.wait_event(ts.ev_maps)
        gt = engine._gt_labels(G, run, tables, boxes_cat, targets)
        n_p = torch.empty(2, dtype=torch.float32, device=dev)
        with _lib.on_device(dev):
            _lib.call("skg_loss_normalisers", _lib.ptr(gt), ctypes.byref(tables), Pn, G.num_cls, _lib.ptr(n_p), _lib.stream_ptr(dev))
        if ih.distributed:
            from .shard import average_normalisers
            n_p = average_normalisers(n_p, in_place=True)
        with torch.enable_grad():
            transh = engine.transh_loss(ih, run, gt, targets) / n_p[1]
        ts.ev_pre.record(ts.side)
    cur.wait_event(ts.ev_pre)
    n_p.record_stream(cur); transh.record_stream(cur)
    return dict(tables=tables, gt=gt, n_p=n_p, transh=transh)


def training_forward(ih, features, image_shapes, box_features: Tensor, box_coords: List[Tensor], box_labels, box_scores,
                     targets, meta=None) -> List[dict]:
    """InteractionHead.forward in training mode from the pooled box features on (head.py:389-427): the per-image
    result dicts plus the loss dict, with ``hoi_loss`` / ``interactiveness_loss`` differentiable w.r.t. every head
    parameter (native backward) and ``transH_loss`` w.r.t. the TransH tables."""
    from . import engine
    assert targets is not None, "Targets should be passed during training"                 # head.py:380-381
    G = ih.box_pair_head
    meta, boxes_cat, labels_dev, scores_dev = engine._prep_inputs(G, features, box_features, box_coords, box_labels,
                                                                  box_scores, meta)
    ts = train_state(ih)
    with torch.no_grad():
        run = engine.head_forward(ih, features["3"], image_shapes, box_features.detach(), boxes_cat, meta, precise_logits=False,
                                  graph=ts.graph, train=True, pk=ts.pk, maps_event=ts.ev_maps.cuda_event if ts.prepass else 0)
        dict.__setitem__(run, "_gt_out", ts.gt_buffer(int(run["seg"].total_pairs), G.num_cls))
    # (called from forward_from_pooled only in training mode: no train() / eval() round trip over the ~470 sub-modules)
    assert ih.training, "training_forward needs the head in training mode"
    pre = _prepass(ih, ts, run, boxes_cat, labels_dev, scores_dev, targets) if ts.prepass and int(run["seg"].total_pairs) > 0 else None
    if pre is not None:
        pre["prepped"] = (meta, boxes_cat, labels_dev, scores_dev)
    with torch.enable_grad():               # the TransH loss is differentiable w.r.t. the tables (kge autograd Functions)
        results = engine.interaction_head_from_pooled(ih, features, image_shapes, box_features.detach(), box_coords,
                                                      box_labels, box_scores, targets=targets, meta=meta, run=run, pre=pre)
    losses = results.pop()
    post = ih.__dict__["_skg_last_post"]
    rec = _StepRecord(ts, run, run["_batch"], run["_flags"], run._ws, post["tables"], post["gt"], post["n_p"],
                      keep=(boxes_cat, labels_dev, scores_dev, run.get("_keep")))
    hoi, inter = _HeadLossFn.apply(ts.anchor, losses["hoi_loss"].detach(), losses["interactiveness_loss"].detach(), rec)
    ih.__dict__["_skg_last_train"] = dict(logits=run["logits"], run=run, rec=rec)
    results.append(dict(hoi_loss=hoi, interactiveness_loss=inter, transH_loss=losses["transH_loss"]))
    return results
