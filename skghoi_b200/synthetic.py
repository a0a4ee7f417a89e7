"""Seeded synthetic HICO-DET-shaped inputs for the interaction head.

Everything is generated on the CPU generator so that the reference (build container), the CPU
oracle and the CUDA path all see identical tensors; callers move them to the device.
There is no dataset or checkpoint on the benchmark box: detections are synthetic boxes with
strictly positive area, humans (label ``human_idx``) sorted first, scores in [0.3, 1].
"""
import json
import os
from collections import OrderedDict
from typing import List, Optional, Sequence, Tuple

import torch

_DATA = os.path.join(os.path.dirname(__file__), "data")


def hico_object_to_verb() -> List[List[int]]:
    """The 80 x {verbs} HICO-DET table (600 entries, person = 49); see
    tests/golden/make_object_to_verb.py for its derivation (reference hicodet/hicodet.py:168-179)."""
    with open(os.path.join(_DATA, "hico_object_to_verb.json")) as f:
        return json.load(f)


def make_detections(shapes_hn: Sequence[Tuple[int, int]], image_shapes: Sequence[Tuple[int, int]],
                    gen: torch.Generator, human_idx: int = 49, num_obj_cls: int = 80):
    """Per image (n_h, n): boxes[n,4] xyxy f32, labels[n] i64 (humans first), scores[n] f32."""
    boxes, labels, scores = [], [], []
    for (n_h, n), (h, w) in zip(shapes_hn, image_shapes):
        xy = torch.rand(n, 2, generator=gen) * torch.tensor([0.7 * w, 0.7 * h])
        wh = torch.rand(n, 2, generator=gen) * torch.tensor([0.25 * w, 0.25 * h]) + 8.0
        boxes.append(torch.cat([xy, xy + wh], 1).float())
        other = torch.randint(0, num_obj_cls - 1, (n - n_h,), generator=gen)
        other = other + (other >= human_idx).long()          # uniform over {0..79} \ {human_idx}
        labels.append(torch.cat([torch.full((n_h,), human_idx, dtype=torch.int64), other]))
        scores.append(torch.rand(n, generator=gen) * 0.7 + 0.3)
    return boxes, labels, scores


def make_targets(boxes, labels, shapes_hn, object_to_verb, gen: torch.Generator, num_gt: int = 3):
    """num_gt ground-truth pairs per image, sampled from the image's own boxes with a verb that is
    valid for the object class."""
    targets = []
    for b, l, (n_h, n) in zip(boxes, labels, shapes_hn):
        hi = torch.randint(0, n_h, (num_gt,), generator=gen)
        oi = torch.randint(0, n, (num_gt,), generator=gen)
        oi = torch.where(oi == hi, (oi + 1) % n, oi)
        verbs = []
        for o in oi.tolist():
            valid = object_to_verb[int(l[o])]
            verbs.append(valid[int(torch.randint(0, len(valid), (1,), generator=gen))])
        targets.append(dict(boxes_h=b[hi].clone(), boxes_o=b[oi].clone(), object=l[oi].clone(),
                            labels=torch.tensor(verbs, dtype=torch.int64)))
    return targets


def make_batch(shapes_hn: Sequence[Tuple[int, int]], seed: int = 42,
               image_shapes: Optional[Sequence[Tuple[int, int]]] = None, with_targets: bool = False,
               levels: Sequence[str] = ("3",), human_idx: int = 49):
    """Inputs at the ``GraphHead.forward`` boundary (reference head.py:769-777).

    Returns a dict: features (OrderedDict level -> [B,256,h/s,w/s]), image_shapes, box_features
    [sum n,256,7,7], box_coords / box_labels / box_scores (lists), targets (list or None).
    Only level '3' is read outside RoIAlign (head.py:811), so it is the only one built by default.
    """
    gen = torch.Generator().manual_seed(seed)
    B = len(shapes_hn)
    if image_shapes is None:
        image_shapes = [(800, 1216)] * B
    hmax = max(s[0] for s in image_shapes); wmax = max(s[1] for s in image_shapes)
    strides = {"0": 4, "1": 8, "2": 16, "3": 32}
    features = OrderedDict()
    for lv in levels:
        s = strides[lv]
        features[lv] = torch.randn(B, 256, (hmax + s - 1) // s, (wmax + s - 1) // s, generator=gen)
    boxes, labels, scores = make_detections(shapes_hn, image_shapes, gen, human_idx)
    total = sum(n for _, n in shapes_hn)
    box_features = torch.randn(total, 256, 7, 7, generator=gen)
    targets = None
    if with_targets:
        targets = make_targets(boxes, labels, shapes_hn, hico_object_to_verb(), gen)
    return dict(features=features, image_shapes=list(image_shapes), box_features=box_features,
                box_coords=boxes, box_labels=labels, box_scores=scores, targets=targets)


def make_transh_tables(seed: int = 42, ent_tot: int = 80, rel_tot: int = 117, dim: int = 50):
    """Xavier-uniform TransH tables as TransH.py:25-28 initialises them (ent, rel, norm_vector)."""
    gen = torch.Generator().manual_seed(seed)

    def xavier(rows):
        bound = (6.0 / (rows + dim)) ** 0.5
        return (torch.rand(rows, dim, generator=gen) * 2 - 1) * bound
    return xavier(ent_tot), xavier(rel_tot), xavier(rel_tot)


def edge_case_boxes():
    """Box pairs that exercise the NaN / log(eps) / nan_to_num semantics of ops.py:107-155 and
    head.py:866-868: zero-area, inverted, identical, touching and image-border boxes."""
    b1 = torch.tensor([
        [10., 20., 110., 220.],     # generic
        [10., 20., 110., 220.],     # identical pair (i == j rows of the grid)
        [50., 50., 50., 50.],       # zero-area human, zero-area object -> IoU 0/0 = NaN
        [50., 50., 50., 90.],       # zero-width human
        [0., 0., 100., 100.],       # touching edge
        [0., 0., 1216., 800.],      # whole image
        [200., 300., 100., 100.],   # inverted human box (negative w, h)
        [100., 100., 200., 200.],   # object left/up of human
        [100., 100., 200., 200.],   # zero-area object
        [0., 0., 0., 0.],           # degenerate at origin vs generic
    ])
    b2 = torch.tensor([
        [60., 40., 300., 260.],
        [10., 20., 110., 220.],
        [70., 70., 70., 70.],
        [40., 40., 80., 80.],
        [100., 0., 200., 100.],
        [1000., 700., 1216., 800.],
        [50., 60., 150., 160.],
        [10., 10., 60., 60.],
        [150., 150., 150., 150.],
        [60., 40., 300., 260.],
    ])
    return b1, b2, (800, 1216)
