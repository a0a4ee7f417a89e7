"""Host-side mirror of the reference interaction head's module interface.

Same class names, constructor signatures, sub-module / ``state_dict`` names and construction order
as heads/adamixer_transH_spatial_r50_head.py (InteractionHead head.py:29-429, MultiBranchFusion
head.py:431-474, MessageMBF head.py:476-530, transH_head head.py:533-582, GraphHead head.py:586-993),
so a reference checkpoint loads unchanged and ``torch.manual_seed(s)`` followed by construction
yields bit-identical initial weights.  The modules hold fp32 master parameters in the reference's
per-branch layout; all arithmetic runs in libskghoi_b200.so through ``engine.HeadEngine`` (packed
bf16 weights, tcgen05 GEMMs with fused epilogues).  CUDA only -- no CPU fallback.
"""
from collections import OrderedDict
from typing import List, Optional, Tuple

import torch
from torch import Tensor, nn
from torch.nn import Module

from . import _lib
from .kge import TransH, MarginLoss, NegativeSampling


class MultiBranchFusion(Module):
    """Reference head.py:431-474: ``ReLU(sum_c fc_3[c](ReLU(fc_1[c](app) * fc_2[c](sp))))``."""

    def __init__(self, appearance_size: int, spatial_size: int, representation_size: int, cardinality: int) -> None:
        super().__init__()
        self.cardinality = cardinality
        sub_repr_size = int(representation_size / cardinality)
        assert sub_repr_size * cardinality == representation_size, \
            "The given representation size should be divisible by cardinality"
        self.fc_1 = nn.ModuleList([nn.Linear(appearance_size, sub_repr_size) for _ in range(cardinality)])
        self.fc_2 = nn.ModuleList([nn.Linear(spatial_size, sub_repr_size) for _ in range(cardinality)])
        self.fc_3 = nn.ModuleList([nn.Linear(sub_repr_size, representation_size) for _ in range(cardinality)])
        self._outer_relu = True

    def packed(self):
        """Concatenated-branch weights: W1 [R, A], b1 [R], W2 [R, S], b2 [R], W3 [R, R], b3 [R]
        (sum over branches of fc_3 = one [R, R] matrix applied to the concatenated hidden vector;
        equivalent to the 16-branch form up to fp32 summation order)."""
        W1 = torch.cat([l.weight for l in self.fc_1], 0); b1 = torch.cat([l.bias for l in self.fc_1], 0)
        W2 = torch.cat([l.weight for l in self.fc_2], 0); b2 = torch.cat([l.bias for l in self.fc_2], 0)
        W3 = torch.cat([l.weight for l in self.fc_3], 1); b3 = torch.stack([l.bias for l in self.fc_3], 0).sum(0)
        return W1, b1, W2, b2, W3, b3

    def forward(self, appearance: Tensor, spatial: Tensor) -> Tensor:
        from .engine import mbf_forward
        return mbf_forward(self, appearance, spatial)


class MessageMBF(MultiBranchFusion):
    """Reference head.py:476-530: anisotropic messages, no outer ReLU."""

    def __init__(self, appearance_size: int, spatial_size: int, representation_size: int, node_type: str,
                 cardinality: int) -> None:
        super().__init__(appearance_size, spatial_size, representation_size, cardinality)
        if node_type not in ("human", "object"):
            raise ValueError("Unknown node type \"{}\"".format(node_type))          # head.py:507
        self.node_type = node_type
        self._outer_relu = False

    def forward(self, appearance: Tensor, spatial: Tensor) -> Tensor:
        from .engine import message_mbf_forward
        return message_mbf_forward(self, appearance, spatial)


class transH_head(Module):
    """Reference head.py:533-582.  The reference builds a NEW randomly initialised TransH on every
    call (head.py:574-578), so its tables are never trained or checkpointed; here the tables live
    in ``self.transh`` (not part of the reference state_dict -- saved under the extra key prefix
    ``transh``) and are re-drawn once per batch with ``resample=True`` for reference-like behaviour (honoured by this
    module's ``forward``, by ``engine.head_forward`` and by the training pass)."""

    def __init__(self, transh_dim: int = 200, transh_p_norm: int = 2, transh_norm_flag: bool = True,
                 human_idx: int = 49, num_object: int = 80, num_cls: int = 117) -> None:
        super().__init__()
        self.transh_dim = transh_dim
        self.transh_p_norm = transh_p_norm
        self.transh_norm_flag = transh_norm_flag
        self.device = "cuda"
        self.human_idx = human_idx
        self.num_object = num_object
        self.num_cls = num_cls
        self.resample = False
        # False: negatives drawn like the reference, torch.randperm on the CPU generator per image (head.py:939, seed
        # compatible); True: same distribution from the device generator for the whole batch at once (no per-image syncs)
        self.negatives_on_device = False
        # built lazily (and outside the seeded construction order of the reference modules)
        self.__dict__["_transh"] = None

    @property
    def transh(self) -> TransH:
        if self.__dict__["_transh"] is None:
            rng = torch.random.get_rng_state()
            m = TransH(ent_tot=self.num_object, rel_tot=self.num_cls, dim=self.transh_dim,
                       p_norm=self.transh_p_norm, norm_flag=self.transh_norm_flag)
            torch.random.set_rng_state(rng)      # do not disturb the caller's RNG stream
            self.__dict__["_transh"] = m
        return self.__dict__["_transh"]

    def set_tables(self, ent: Tensor, rel: Tensor, nrm: Tensor, device=None) -> None:
        m = self.transh
        with torch.no_grad():
            m.ent_embeddings.weight.copy_(ent); m.rel_embeddings.weight.copy_(rel); m.norm_vector.weight.copy_(nrm)
        if device is not None:
            m.to(device)

    def maybe_resample(self) -> None:
        """``resample=True``: re-draws the three tables like the reference, which constructs a fresh Xavier-initialised
        TransH on every call (head.py:574-578).  In place on the parameters (no ``.data`` detour: the version counters
        move, and every forward re-reads the table on the device anyway).  Called by ``forward`` and by the engine /
        training paths once per batch."""
        if self.resample:
            m = self.transh
            with torch.no_grad():
                for emb in (m.ent_embeddings, m.rel_embeddings, m.norm_vector):
                    nn.init.xavier_uniform_(emb.weight)

    def forward(self, ind_x, ind_y, n_h: int, n: int, x_keep, y_keep, labels: Optional[Tensor] = None):
        """head.py:558-582: 117 triples per grid cell, head = human_idx, tail = box index ``ind_y``."""
        m = self.transh
        dev = ind_y.device
        _lib.require_cuda(ind_y)
        if m.ent_embeddings.weight.device != dev:
            m.to(dev)
        self.maybe_resample()
        cells = len(ind_x)
        T = cells * self.num_cls
        e, r, w = m.ent_embeddings.weight, m.rel_embeddings.weight, m.norm_vector.weight
        outs = [torch.empty(T, self.transh_dim, dtype=torch.float32, device=dev) for _ in range(4)]
        score = torch.empty(T, dtype=torch.float32, device=dev)
        tails = ind_y.to(torch.int64).contiguous()
        if cells and int(tails.max()) >= self.num_object:
            raise IndexError("index out of range in self")     # head.py:572 indexes the 80-row table with box indices
        with torch.cuda.device(dev):
            _lib.call("skg_transh_forward_grid", _lib.ptr(e), _lib.ptr(r), _lib.ptr(w), self.num_object, self.num_cls,
                      self.transh_dim, self.human_idx, _lib.ptr(tails), cells, self.transh_p_norm,
                      int(self.transh_norm_flag), _lib.ptr(outs[0]), _lib.ptr(outs[1]), _lib.ptr(outs[2]),
                      _lib.ptr(outs[3]), _lib.ptr(score), _lib.stream_ptr(dev))
        return outs[0], outs[1], outs[2], outs[3], score


class GraphHead(Module):
    """Reference head.py:586-993 (constructor head.py:609-701)."""

    def __init__(self, out_channels: int, roi_pool_size: int, node_encoding_size: int, representation_size: int,
                 num_cls: int, human_idx: int, object_class_to_target_class: List[list],
                 fg_iou_thresh: float = 0.5, num_iter: int = 2) -> None:
        super().__init__()
        self.out_channels = out_channels
        self.roi_pool_size = roi_pool_size
        self.node_encoding_size = node_encoding_size
        self.representation_size = representation_size
        self.num_cls = num_cls
        self.human_idx = human_idx
        self.object_class_to_target_class = object_class_to_target_class
        self.fg_iou_thresh = fg_iou_thresh
        self.num_iter = num_iter
        # construction order below is the reference's (it fixes the RNG stream of the initialisers)
        self.box_head = nn.Sequential(
            nn.Flatten(start_dim=1),
            nn.Linear(out_channels * roi_pool_size ** 2, node_encoding_size), nn.ReLU(),
            nn.Linear(node_encoding_size, node_encoding_size), nn.ReLU())
        self.adjacency = nn.Linear(representation_size, 1)
        self.sub_to_obj = MessageMBF(node_encoding_size, 1024, representation_size, node_type="human", cardinality=16)
        self.obj_to_sub = MessageMBF(node_encoding_size, 1024, representation_size, node_type="object", cardinality=16)
        self.norm_h = nn.LayerNorm(node_encoding_size)
        self.norm_o = nn.LayerNorm(node_encoding_size)
        self.spatial_head = nn.Sequential(
            nn.Linear(46, 128), nn.ReLU(), nn.Linear(128, 256), nn.ReLU(), nn.Linear(256, 1024), nn.ReLU())
        self.attention_head = MultiBranchFusion(node_encoding_size * 2, 1024, representation_size, cardinality=16)
        self.avg_pool = nn.AdaptiveAvgPool2d(output_size=1)
        self.attention_head_g = MultiBranchFusion(256, 1024, representation_size, cardinality=16)
        self.transh_head = transH_head(transh_dim=50, transh_p_norm=2, transh_norm_flag=True,
                                       human_idx=self.human_idx, num_object=80, num_cls=self.num_cls)
        self.fc_head = nn.Sequential(nn.Linear(1074, 1024), nn.ReLU())
        self.fc_tail = nn.Sequential(nn.Linear(1074, 1024), nn.ReLU())

    def forward(self, features: OrderedDict, image_shapes: List[Tuple[int, int]], box_features: Tensor,
                box_coords: List[Tensor], box_labels: List[Tensor], box_scores: List[Tensor],
                targets: Optional[List[dict]] = None):
        """Reference head.py:769-993: returns the 8-tuple (eval) / 12-tuple (train) of per-image lists."""
        from .engine import graph_head_forward
        return graph_head_forward(self, features, image_shapes, box_features, box_coords, box_labels,
                                  box_scores, targets)


class InteractionHead(Module):
    """Reference head.py:29-429 (constructor head.py:58-90)."""

    def __init__(self, box_roi_pool: Module, box_pair_head: Module, box_pair_suppressor: Module,
                 box_pair_predictor: Module, human_idx: int, num_classes: int, box_nms_thresh: float = 0.5,
                 box_score_thresh: float = 0.2, max_human: int = 15, max_object: int = 15,
                 distributed: bool = False) -> None:
        super().__init__()
        self.box_roi_pool = box_roi_pool
        self.box_pair_head = box_pair_head
        self.box_pair_suppressor = box_pair_suppressor
        self.box_pair_predictor = box_pair_predictor
        self.num_classes = num_classes
        self.human_idx = human_idx
        self.box_nms_thresh = box_nms_thresh
        self.box_score_thresh = box_score_thresh
        self.max_human = max_human
        self.max_object = max_object
        self.distributed = distributed

    def _apply(self, fn, *args, **kwargs):
        # .to() / .cuda() / .float() re-home the parameters: drop the packed bf16 operands
        self.__dict__.pop("_skg_packed", None)
        return super()._apply(fn, *args, **kwargs)

    def preprocess(self, detections: List[dict], targets: List[dict], append_gt: Optional[bool] = None):
        from .engine import preprocess_detections, preprocess_detections_device
        if len(detections) and detections[0]["boxes"].is_cuda and getattr(self, "device_preprocess", True):
            return preprocess_detections_device(self, detections, targets, append_gt)[0]      # csrc/preprocess.cu
        return preprocess_detections(self, detections, targets, append_gt)                    # stock torchvision ops

    def forward(self, features: OrderedDict, detections: List[dict], image_shapes: List[Tuple[int, int]],
                targets: Optional[List[dict]] = None) -> List[dict]:
        """Reference head.py:341-429."""
        from .engine import interaction_head_forward
        return interaction_head_forward(self, features, detections, image_shapes, targets)

    def forward_from_pooled(self, features: OrderedDict, image_shapes, box_features: Tensor, box_coords,
                            box_labels, box_scores, targets=None, meta=None) -> List[dict]:
        """The hot path proper: head.py:389-427 given already-preprocessed detections and pooled
        box features (what remains of ``forward`` after the stock torchvision NMS / RoIAlign steps)."""
        if self.training and torch.is_grad_enabled():
            from .train import training_forward      # losses differentiable w.r.t. the head parameters
            return training_forward(self, features, image_shapes, box_features, box_coords, box_labels, box_scores,
                                    targets, meta=meta)
        from .engine import interaction_head_from_pooled
        return interaction_head_from_pooled(self, features, image_shapes, box_features, box_coords, box_labels,
                                            box_scores, targets, meta=meta)

    def host_meta(self, box_labels, box_scores, n=None):
        """Host-side summary of a batch's detections (boxes / humans per image, labels, scores) that sizes the launch
        and the outputs.  Build it from HOST copies of the labels / scores (the reference reads its detections from a
        host-side cache) and pass it as ``meta=`` to skip the device->host read ``forward`` otherwise needs.  ``box_labels`` / ``box_scores`` are
        per-image lists, or two PACKED tensors with the per-image counts in ``n``."""
        from .engine import _host_meta
        return _host_meta(box_labels, box_scores, self.human_idx, n=n)

    def results_to_host(self, results: List[dict], wait: bool = True):
        """Host copies of the dicts ``forward`` just returned (two pinned copies for the whole batch).  ``wait=False``
        returns a handle whose ``get()`` waits only for this step's copies."""
        from .engine import results_to_host
        return results_to_host(self, results, wait)


def build_interaction_head(object_to_verb: List[list], human_idx: int = 49, num_classes: int = 117,
                           num_iterations: int = 2, box_roi_pool: Optional[Module] = None,
                           box_nms_thresh: float = 0.5, box_score_thresh: float = 0.2, max_human: int = 15,
                           max_object: int = 15, distributed: bool = False, seed: Optional[int] = None,
                           output_size: int = 7, sampling_ratio: int = 2, out_channels: int = 256,
                           node_encoding_size: int = 1024, representation_size: int = 1024,
                           fg_iou_thresh: float = 0.5) -> InteractionHead:
    """Builds the head exactly as the reference network shell does (models.py:158-191), without
    the mmdet detector (which stays stock PyTorch and outside this package)."""
    if seed is not None:
        torch.manual_seed(seed)
    if box_roi_pool is None:
        from torchvision.ops import MultiScaleRoIAlign
        box_roi_pool = MultiScaleRoIAlign(featmap_names=["0", "1", "2", "3"], output_size=output_size,
                                          sampling_ratio=sampling_ratio)
    graph = GraphHead(out_channels=out_channels, roi_pool_size=output_size, node_encoding_size=node_encoding_size,
                      representation_size=representation_size, num_cls=num_classes, human_idx=human_idx,
                      object_class_to_target_class=object_to_verb, fg_iou_thresh=fg_iou_thresh,
                      num_iter=num_iterations)
    predictor = nn.Linear(representation_size * 2, num_classes)
    suppressor = nn.Linear(representation_size * 2, 1)
    return InteractionHead(box_roi_pool=box_roi_pool, box_pair_head=graph, box_pair_suppressor=suppressor,
                           box_pair_predictor=predictor, num_classes=num_classes, human_idx=human_idx,
                           box_nms_thresh=box_nms_thresh, box_score_thresh=box_score_thresh, max_human=max_human,
                           max_object=max_object, distributed=distributed)
