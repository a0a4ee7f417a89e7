"""Import harness for the UNMODIFIED reference head (test infrastructure only).

This module exists only to (a) validate the CPU restatement in ``oracle/head_oracle.py`` and
(b) generate the golden vectors under ``tests/golden/`` in the build container, where the
read-only reference checkout is mounted at ``/root/reference``.  It never travels to the GPU box
(``/root/reference`` is absent there) and nothing in the product package imports it.

Recipe; no reference file is edited:
  * stub ``mmdet.utils`` (``get_root_logger``/``get_device``; head.py:20, ops.py:17, TransH.py:5)
    and ``pocket.ops.Flatten`` (head.py:12);
  * replace the hard-coded cluster ``sys.path`` entries (head.py:24, MarginLoss.py:7, Loss.py:2);
  * inject ONE fixed, seeded TransH so the per-image random construction (head.py:574-578) is
    deterministic;
  * map the literal ``'cuda'`` device of ``transH_head`` (head.py:548) to CPU.
"""
import logging
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("SKGHOI_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "heads", "adamixer_transH_spatial_r50_head.py"))


_loaded = {}


def load():
    """Returns (head_module, ops_module, TransH_cls, MarginLoss_cls, NegativeSampling_cls)."""
    if _loaded:
        return _loaded["mods"]
    if not available():
        raise RuntimeError("reference checkout not present at %s" % REF_ROOT)
    # --- stubs for packages that are not installable here
    mmdet = types.ModuleType("mmdet")
    mmdet_utils = types.ModuleType("mmdet.utils")
    _log = logging.getLogger("skghoi_ref")
    _log.addHandler(logging.NullHandler())
    _log.propagate = False
    mmdet_utils.get_root_logger = lambda *a, **k: _log
    mmdet_utils.get_device = lambda *a, **k: "cpu"
    mmdet.utils = mmdet_utils
    sys.modules.setdefault("mmdet", mmdet)
    sys.modules.setdefault("mmdet.utils", mmdet_utils)
    pocket = types.ModuleType("pocket")
    pocket_ops = types.ModuleType("pocket.ops")
    pocket_ops.Flatten = torch.nn.Flatten
    pocket.ops = pocket_ops
    sys.modules.setdefault("pocket", pocket)
    sys.modules.setdefault("pocket.ops", pocket_ops)
    # --- import paths replacing the hard-coded /users/PCS0256/... entries
    for p in ("", "heads", "heads/TransH", "OpenKE/openke/module", "OpenKE/openke/module/loss"):
        full = os.path.join(REF_ROOT, p) if p else REF_ROOT
        if full not in sys.path:
            sys.path.append(full)
    import importlib
    ops = importlib.import_module("ops")
    assert ops.__file__.startswith(REF_ROOT), ops.__file__
    head = importlib.import_module("adamixer_transH_spatial_r50_head")
    TransH = importlib.import_module("TransH").TransH
    MarginLoss = importlib.import_module("MarginLoss").MarginLoss
    NegativeSampling = importlib.import_module("NegativeSampling").NegativeSampling
    _loaded["mods"] = (head, ops, TransH, MarginLoss, NegativeSampling)
    return _loaded["mods"]


def build_heads(object_to_verb, num_cls=117, human_idx=49, num_iter=2, seed=42,
                max_human=15, max_object=15, box_nms_thresh=0.5):
    """Builds the reference GraphHead + InteractionHead exactly as models.py:158-191 does."""
    from torchvision.ops import MultiScaleRoIAlign
    head, _, _, _, _ = load()
    torch.manual_seed(seed)
    graph = head.GraphHead(256, 7, 1024, 1024, num_cls, human_idx, object_to_verb,
                           fg_iou_thresh=0.5, num_iter=num_iter)
    predictor = torch.nn.Linear(2048, num_cls)      # models.py:176
    suppressor = torch.nn.Linear(2048, 1)           # models.py:177
    ih = head.InteractionHead(MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2), graph, suppressor,
                              predictor, human_idx, num_cls, box_nms_thresh=box_nms_thresh,
                              max_human=max_human, max_object=max_object)
    graph.transh_head.device = "cpu"   # head.py:548 hard-codes 'cuda'
    return graph, ih


class fixed_transh:
    """Context manager: the head's ``TransH(...)`` constructor returns one pre-built instance."""

    def __init__(self, tables):
        head, _, TransH, _, _ = load()
        self.head = head
        self.orig = head.TransH
        ent, rel, nrm = tables
        m = TransH(ent.shape[0], rel.shape[0], dim=ent.shape[1], p_norm=2, norm_flag=True)
        with torch.no_grad():
            m.ent_embeddings.weight.copy_(ent)
            m.rel_embeddings.weight.copy_(rel)
            m.norm_vector.weight.copy_(nrm)
        self.model = m

    def __enter__(self):
        self.head.TransH = lambda **kw: self.model
        return self.model

    def __exit__(self, *exc):
        self.head.TransH = self.orig
        return False
