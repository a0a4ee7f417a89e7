"""CPU restatement of the SKGHOI interaction head -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this file; it is the checker (and the timed CPU baseline), never the thing shipped.
The product package ``skghoi_b200`` never imports it and raises when its CUDA library is missing.

Parity status: the reference repository ships no tests or golden vectors for this path
("parity unpinned" by the reference itself).  This restatement is therefore
pinned against outputs of the UNMODIFIED reference executed in the build container
(``oracle/ref_harness.py`` -> ``tests/golden/make_golden.py`` -> ``tests/golden/*.pt``), and
``tests/test_oracle_vs_golden.py`` re-checks it against those fixtures on every run.

Everything is plain fp32 ``torch`` (on the CPU in the tests; tensor constructors follow the inputs' device so that
bench.py's ``gpu_eager_baseline`` leg can time the very same eager code on ``cuda:0``), written op-for-op in the order the reference executes
(16 separate branch Linears per MBF, 117*N2 materialised TransH triples, the per-row entity pick),
so that timing it is an honest stand-in for the reference's own CPU head.  Each function cites the
reference file:line it restates (paths relative to the reference root; ``head.py`` abbreviates
``heads/adamixer_transH_spatial_r50_head.py``).
"""
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor


# --------------------------------------------------------------------------------------------
# ops.py
# --------------------------------------------------------------------------------------------
def box_iou_diag(b1: Tensor, b2: Tensor) -> Tensor:
    """Diagonal of torchvision ``box_iou`` (torchvision 0.26 ``_box_inter_union``), as used at
    ops.py:55 and ops.py:119.  No eps in the division: two zero-area boxes give NaN."""
    a1 = (b1[:, 2] - b1[:, 0]) * (b1[:, 3] - b1[:, 1])
    a2 = (b2[:, 2] - b2[:, 0]) * (b2[:, 3] - b2[:, 1])
    lt = torch.max(b1[:, :2], b2[:, :2])
    rb = torch.min(b1[:, 2:], b2[:, 2:])
    wh = (rb - lt).clamp(min=0)
    inter = wh[:, 0] * wh[:, 1]
    union = a1 + a2 - inter
    return inter / union


def box_iou_matrix(b1: Tensor, b2: Tensor) -> Tensor:
    """Full torchvision ``box_iou`` matrix [len(b1), len(b2)] (head.py:712-713)."""
    a1 = (b1[:, 2] - b1[:, 0]) * (b1[:, 3] - b1[:, 1])
    a2 = (b2[:, 2] - b2[:, 0]) * (b2[:, 3] - b2[:, 1])
    lt = torch.max(b1[:, None, :2], b2[None, :, :2])
    rb = torch.min(b1[:, None, 2:], b2[None, :, 2:])
    wh = (rb - lt).clamp(min=0)
    inter = wh[..., 0] * wh[..., 1]
    return inter / (a1[:, None] + a2[None, :] - inter)


def _spatial_common(b1: Tensor, b2: Tensor, eps: float):
    c1_x = (b1[:, 0] + b1[:, 2]) / 2; c1_y = (b1[:, 1] + b1[:, 3]) / 2
    c2_x = (b2[:, 0] + b2[:, 2]) / 2; c2_y = (b2[:, 1] + b2[:, 3]) / 2
    b1_w = b1[:, 2] - b1[:, 0]; b1_h = b1[:, 3] - b1[:, 1]
    b2_w = b2[:, 2] - b2[:, 0]; b2_h = b2[:, 3] - b2[:, 1]
    d_x = torch.abs(c2_x - c1_x) / (b1_w + eps)
    d_y = torch.abs(c2_y - c1_y) / (b1_h + eps)
    iou = box_iou_diag(b1, b2)
    return c1_x, c1_y, c2_x, c2_y, b1_w, b1_h, b2_w, b2_h, d_x, d_y, iou


def compute_spatial_encodings(boxes_1: Sequence[Tensor], boxes_2: Sequence[Tensor],
                              shapes: Sequence[Tuple[int, int]], eps: float = 1e-10) -> Tensor:
    """ops.py:20-83 -- the 18-feature variant -> [N, 36].  The reference's in-loop
    ``features = torch.cat(features)`` (ops.py:81) only works for 1-image lists; the evident intent
    (concatenate over images) is what is restated, identical for 1-image lists."""
    out = []
    for b1, b2, (h, w) in zip(boxes_1, boxes_2, shapes):
        assert h > 0 and w > 0                                                # ops.py:43-44
        c1_x, c1_y, c2_x, c2_y, b1_w, b1_h, b2_w, b2_h, d_x, d_y, iou = _spatial_common(b1, b2, eps)
        f = torch.stack([
            c1_x / w, c1_y / h, c2_x / w, c2_y / h,
            b1_w / w, b1_h / h, b2_w / w, b2_h / h,
            b1_w * b1_h / (h * w), b2_w * b2_h / (h * w),
            b2_w * b2_h / (b1_w * b1_h + eps),
            b1_w / (b1_h + eps), b2_w / (b2_h + eps),
            iou,
            (c2_x > c1_x).float() * d_x, (c2_x < c1_x).float() * d_x,
            (c2_y > c1_y).float() * d_y, (c2_y < c1_y).float() * d_y,
        ], 1)                                                                 # ops.py:58-75
        out.append(torch.cat([f, torch.log(f + eps)], 1))                     # ops.py:77-79
    return torch.cat(out)


def compute_spatial_ratio_encodings(boxes_1: Sequence[Tensor], boxes_2: Sequence[Tensor],
                                    shapes: Sequence[Tuple[int, int]], eps: float = 1e-10) -> Tensor:
    """ops.py:85-157 -- the 23-feature variant the head calls (head.py:863) -> [N, 46]."""
    out = []
    for b1, b2, (h, w) in zip(boxes_1, boxes_2, shapes):
        c1_x, c1_y, c2_x, c2_y, b1_w, b1_h, b2_w, b2_h, d_x, d_y, iou = _spatial_common(b1, b2, eps)
        c1_xw = c1_x / w; c1_yh = c1_y / h; c2_xw = c2_x / w; c2_yh = c2_y / h
        b1_ww = b1_w / w; b1_hh = b1_h / h; b2_ww = b2_w / w; b2_hh = b2_h / h
        a1 = b1_w * b1_h / (h * w); a2 = b2_w * b2_h / (h * w)
        r1 = b1_w / (b1_h + eps); r2 = b2_w / (b2_h + eps)
        f = torch.stack([
            c1_xw, c1_yh, c2_xw, c2_yh, c1_xw / (c2_xw + eps), c1_yh / (c2_yh + eps),
            b1_ww, b1_hh, b2_ww, b2_hh, b1_ww / (b2_ww + eps), b1_hh / (b2_hh + eps),
            a1, a2, a1 / (a2 + eps),
            b2_w * b2_h / (b1_w * b1_h + eps),
            r1, r2,
            iou,
            (c2_x > c1_x).float() * d_x, (c2_x < c1_x).float() * d_x,
            (c2_y > c1_y).float() * d_y, (c2_y < c1_y).float() * d_y,
        ], 1)                                                                 # ops.py:134-151
        out.append(torch.cat([f, torch.log(f + eps)], 1))                     # ops.py:154-156
    return torch.cat(out)


def binary_focal_loss(x: Tensor, y: Tensor, alpha: float = 0.5, gamma: float = 2.0,
                      reduction: str = "mean", eps: float = 1e-6) -> Tensor:
    """ops.py:159-211."""
    loss = (1 - y - alpha).abs() * ((y - x).abs() + eps) ** gamma * \
        F.binary_cross_entropy(x, y, reduction="none")
    if reduction == "mean":
        return loss.mean()
    if reduction == "sum":
        return loss.sum()
    if reduction == "none":
        return loss
    raise ValueError("Unsupported reduction method {}".format(reduction))


# --------------------------------------------------------------------------------------------
# heads/TransH/TransH.py, heads/MarginLoss.py, heads/NegativeSampling.py
# --------------------------------------------------------------------------------------------
def transh_forward(ent: Tensor, rel: Tensor, nrm: Tensor, head: Tensor, relation: Tensor,
                   tail: Tensor, p_norm: int = 2, norm_flag: bool = True):
    """TransH.py:87-106 ('normal' mode; ``_transfer`` TransH.py:75-85, ``_calc`` TransH.py:56-73).
    Returns the raw gathered ``h_, r, r_norm, t_`` and the score, as the fork does."""
    h_ = ent[head]; t_ = ent[tail]; r = rel[relation]; w = nrm[relation]
    wn = F.normalize(w, p=2, dim=-1)
    h = h_ - torch.sum(h_ * wn, -1, True) * wn
    t = t_ - torch.sum(t_ * wn, -1, True) * wn
    rr = r
    if norm_flag:
        h = F.normalize(h, 2, -1); rr = F.normalize(r, 2, -1); t = F.normalize(t, 2, -1)
    score = torch.norm((h + rr) - t, p_norm, -1).flatten()
    return h_, r, w, t_, score


def margin_loss(p_score: Tensor, n_score: Tensor, margin: float = 6.0,
                adv_temperature: Optional[float] = None) -> Tensor:
    """MarginLoss.py:28-36."""
    m = torch.tensor([margin], dtype=p_score.dtype, device=p_score.device)
    if adv_temperature is not None:
        wgt = F.softmax(-n_score * adv_temperature, dim=-1).detach()
        return (wgt * torch.max(p_score - n_score, -m)).sum(dim=-1).mean() + m
    return torch.max(p_score - n_score, -m).mean() + m


def negative_sampling_loss(score: Tensor, margin: float = 6.0,
                           adv_temperature: Optional[float] = None) -> Tensor:
    """NegativeSampling.py:30-63 (regul_rate = 0): first half positives, second half negatives."""
    half = len(score) // 2
    p = score[:half].view(-1, half).permute(1, 0)
    n = score[half:].view(-1, half).permute(1, 0)
    return margin_loss(p, n, margin, adv_temperature)


# --------------------------------------------------------------------------------------------
# heads/adamixer_transH_spatial_r50_head.py
# --------------------------------------------------------------------------------------------
def _linear(P: Dict[str, Tensor], name: str, x: Tensor) -> Tensor:
    return F.linear(x, P[name + ".weight"], P[name + ".bias"])


def mbf_forward(P: Dict[str, Tensor], prefix: str, appearance: Tensor, spatial: Tensor,
                cardinality: int = 16) -> Tensor:
    """MultiBranchFusion.forward head.py:469-474 (16 separate branch Linears, stack, sum, ReLU)."""
    return F.relu(torch.stack([
        _linear(P, f"{prefix}.fc_3.{c}", F.relu(_linear(P, f"{prefix}.fc_1.{c}", appearance)
                                                * _linear(P, f"{prefix}.fc_2.{c}", spatial)))
        for c in range(cardinality)]).sum(dim=0))


def message_mbf_forward(P: Dict[str, Tensor], prefix: str, node_type: str, appearance: Tensor,
                        spatial: Tensor, cardinality: int = 16) -> Tensor:
    """MessageMBF head.py:509-527.  No outer ReLU."""
    n_h, n = spatial.shape[:2]
    outs = []
    for c in range(cardinality):
        a = _linear(P, f"{prefix}.fc_1.{c}", appearance)
        s = _linear(P, f"{prefix}.fc_2.{c}", spatial)
        if node_type == "human":
            assert len(appearance) == n_h
            prod = a.repeat(n, 1, 1) * s.permute([1, 0, 2])
        elif node_type == "object":
            assert len(appearance) == n
            prod = a.repeat(n_h, 1, 1) * s
        else:
            raise ValueError("Unknown node type \"{}\"".format(node_type))
        outs.append(_linear(P, f"{prefix}.fc_3.{c}", F.relu(prod)))
    return torch.stack(outs).sum(dim=0)


def associate_with_ground_truth(boxes_h: Tensor, boxes_o: Tensor, target: dict, num_cls: int,
                                fg_iou_thresh: float = 0.5) -> Tensor:
    """head.py:703-719."""
    labels = torch.zeros(boxes_h.shape[0], num_cls, device=boxes_h.device)
    x, y = torch.nonzero(torch.min(box_iou_matrix(boxes_h, target["boxes_h"]),
                                   box_iou_matrix(boxes_o, target["boxes_o"])) >= fg_iou_thresh).unbind(1)
    labels[x, target["labels"][y]] = 1
    return labels


def compute_prior_scores(x: Tensor, y: Tensor, scores: Tensor, object_class: Tensor,
                         object_to_verb: List[list], num_cls: int, training: bool) -> Tensor:
    """head.py:721-767."""
    prior_h = torch.zeros(len(x), num_cls, device=scores.device)
    prior_o = torch.zeros_like(prior_h)
    p = 1.0 if training else 2.8
    s_h = scores[x].pow(p); s_o = scores[y].pow(p)
    target_cls_idx = [object_to_verb[int(obj)] for obj in object_class[y]]
    pair_idx = [i for i, tar in enumerate(target_cls_idx) for _ in tar]
    flat_target_idx = [t for tar in target_cls_idx for t in tar]
    prior_h[pair_idx, flat_target_idx] = s_h[pair_idx]
    prior_o[pair_idx, flat_target_idx] = s_o[pair_idx]
    return torch.stack([prior_h, prior_o])


def graph_head_forward(P: Dict[str, Tensor], tables: Tuple[Tensor, Tensor, Tensor],
                       features3: Tensor, image_shapes: Sequence[Tuple[int, int]],
                       box_features: Tensor, box_coords: List[Tensor], box_labels: List[Tensor],
                       box_scores: List[Tensor], object_to_verb: List[list],
                       targets: Optional[List[dict]] = None, training: bool = False,
                       num_cls: int = 117, human_idx: int = 49, num_iter: int = 2,
                       fg_iou_thresh: float = 0.5, prefix: str = "box_pair_head",
                       row_pick: str = "loop", neg_perm: Optional[List[Tensor]] = None,
                       taps: Optional[dict] = None):
    """GraphHead.forward head.py:769-993, restated step by step.

    ``tables`` = (ent[80,50], rel[117,50], norm[117,50]) stands in for the TransH the reference
    constructs per image (head.py:574-578; injected on both sides for determinism).
    ``row_pick='loop'`` walks the 117*N2 rows in Python like head.py:877-883; ``'slice'`` takes
    ``[::num_cls]`` (same values).  ``neg_perm`` optionally supplies the per-image permutation that
    ``torch.randperm`` (head.py:939) would draw.  ``taps`` (dict) collects named intermediates.
    Skipped images advance ``counter`` (the reference does not, SURVEY defect 6; not exercised by
    any fixture) and the image loop runs over all images.
    """
    ent, rel, nrm = tables
    pf = prefix
    global_features = features3.mean(dim=(2, 3))                              # head.py:811
    bf = box_features.flatten(start_dim=1)
    bf = F.relu(_linear(P, f"{pf}.box_head.1", bf))
    bf = F.relu(_linear(P, f"{pf}.box_head.3", bf))                           # head.py:812
    out = dict(feat=[], boxes_h=[], boxes_o=[], obj=[], labels=[], prior=[], pos=[], neg=[],
               h_ent=[], t_ent=[], r=[], r_norm=[])
    counter = 0
    for b_idx, (coords, labels, scores) in enumerate(zip(box_coords, box_labels, box_scores)):
        n = len(coords)
        n_h = int(torch.sum(labels == human_idx))
        if n_h == 0 or n <= 1:                                                # head.py:829-839
            dv = bf.device
            out["feat"].append(torch.zeros(0, 2048, device=dv)); out["boxes_h"].append(torch.zeros(0, 4, device=dv))
            out["boxes_o"].append(torch.zeros(0, 4, device=dv)); out["obj"].append(torch.zeros(0, dtype=torch.int64, device=dv))
            out["prior"].append(torch.zeros(2, 0, num_cls, device=dv)); out["labels"].append(torch.zeros(0, num_cls, device=dv))
            counter += n
            continue
        if not torch.all(labels[:n_h] == human_idx):
            raise ValueError("Human detections are not permuted to the top")  # head.py:840-841
        node_enc = bf[counter: counter + n]
        h_node_enc = node_enc[:n_h]
        dv = coords.device
        x, y = torch.meshgrid(torch.arange(n_h, device=dv), torch.arange(n, device=dv), indexing="ij")   # head.py:847-850
        x_keep, y_keep = torch.nonzero(x != y).unbind(1)                      # head.py:852
        if len(x_keep) == 0:
            raise ValueError("There are no valid human-object pairs")
        x = x.flatten(); y = y.flatten()
        sp46 = compute_spatial_ratio_encodings([coords[x]], [coords[y]], [image_shapes[b_idx]])
        if torch.isnan(sp46).sum() > 0:                                       # head.py:866-868
            sp46 = torch.nan_to_num(sp46)
        # transH_head.forward head.py:570-580
        N2 = len(x)
        relations = torch.arange(num_cls, device=dv).repeat(N2)
        heads = torch.full((N2 * num_cls,), human_idx, dtype=torch.int64, device=dv)
        tails = y.repeat_interleave(num_cls)
        th, tr, tw, tt, tscore = transh_forward(ent, rel, nrm, heads, relations, tails, 2, True)
        if row_pick == "loop":                                                # head.py:875-883
            he, te = [], []
            for idx, (ex, ey) in enumerate(zip(th, tt)):
                if idx % num_cls == 0:
                    he.append(ex); te.append(ey)
            head_stack = torch.stack(he); tail_stack = torch.stack(te)
        else:
            head_stack = th[::num_cls]; tail_stack = tt[::num_cls]
        g_h = F.relu(_linear(P, f"{pf}.fc_head.0", torch.cat((h_node_enc[x], head_stack), 1)))
        g_o = F.relu(_linear(P, f"{pf}.fc_tail.0", torch.cat((node_enc[y], tail_stack), 1)))
        sp = sp46
        for l in (0, 2, 4):                                                   # head.py:662-669, 888
            sp = F.relu(_linear(P, f"{pf}.spatial_head.{l}", sp))
        sp_r = sp.reshape(n_h, n, -1)
        adj = torch.ones(n_h, n, device=dv)
        for _ in range(num_iter):                                             # head.py:892-925
            app = torch.cat([g_h, g_o], 1)
            weights = mbf_forward(P, f"{pf}.attention_head", app, sp)
            adj = _linear(P, f"{pf}.adjacency", weights).reshape(n_h, n)
            all_ent = g_o[0:int(g_o.shape[0] / n_h)]
            h_ent = g_h[[i for i in range(g_h.shape[0]) if i % n == 0]]
            o_t_s = message_mbf_forward(P, f"{pf}.obj_to_sub", "object", all_ent, sp_r)
            msg_h = F.relu(torch.sum(adj.softmax(dim=1)[..., None] * o_t_s, dim=1))
            h_node_enc = F.layer_norm(h_ent + msg_h, (h_ent.shape[-1],), P[f"{pf}.norm_h.weight"],
                                      P[f"{pf}.norm_h.bias"], 1e-5)
            s_t_o = message_mbf_forward(P, f"{pf}.sub_to_obj", "human", h_ent, sp_r)
            msg_o = F.relu(torch.sum(adj.t().softmax(dim=1)[..., None] * s_t_o, dim=1))
            node_enc = F.layer_norm(all_ent + msg_o, (all_ent.shape[-1],), P[f"{pf}.norm_o.weight"],
                                    P[f"{pf}.norm_o.bias"], 1e-5)
        score_keep = tscore.reshape(n_h, n, num_cls)[x_keep, y_keep]          # head.py:926
        if targets is not None:                                               # head.py:933-963
            tl = associate_with_ground_truth(coords[x_keep], coords[y_keep], targets[b_idx],
                                             num_cls, fg_iou_thresh)
            px, py = torch.nonzero(tl).unbind(1)
            neg_xy = (tl == 0).nonzero()
            perm = neg_perm[b_idx] if neg_perm is not None else torch.randperm(neg_xy.size(0))
            nx, ny = neg_xy[perm[:len(px)]].unbind(1)
            out["labels"].append(tl)
            out["pos"].append(score_keep[px, py]); out["neg"].append(score_keep[nx, ny])
            for key, t in (("h_ent", th), ("t_ent", tt), ("r", tr), ("r_norm", tw)):
                k = t.reshape(n_h, n, num_cls, -1)[x_keep, y_keep]
                out[key].append(torch.cat((k[px, py], k[nx, ny]), 0))
        sp_keep = sp_r[x_keep, y_keep]
        att1 = mbf_forward(P, f"{pf}.attention_head",
                           torch.cat([h_node_enc[x_keep], node_enc[y_keep]], 1), sp_keep)
        att2 = mbf_forward(P, f"{pf}.attention_head_g", global_features[b_idx, None], sp_keep)
        out["feat"].append(torch.cat([att1, att2], dim=1))                    # head.py:970-973
        out["boxes_h"].append(coords[x_keep]); out["boxes_o"].append(coords[y_keep])
        out["obj"].append(labels[y_keep])
        out["prior"].append(compute_prior_scores(x_keep, y_keep, scores, labels, object_to_verb,
                                                 num_cls, training))
        if taps is not None:
            taps.setdefault("x", []).append(x); taps.setdefault("y", []).append(y)
            taps.setdefault("x_keep", []).append(x_keep); taps.setdefault("y_keep", []).append(y_keep)
            taps.setdefault("spatial46", []).append(sp46); taps.setdefault("spatial1024", []).append(sp)
            taps.setdefault("graph_human_node", []).append(g_h[::n].clone())
            taps.setdefault("graph_object_node", []).append(g_o[:n].clone())
            taps.setdefault("adjacency", []).append(adj)
            taps.setdefault("h_node_enc", []).append(h_node_enc)
            taps.setdefault("node_enc", []).append(node_enc)
            taps.setdefault("transh_score", []).append(tscore.reshape(n_h, n, num_cls))
        counter += n
    return out


def postprocess(logits_p: Tensor, logits_s: Tensor, prior: List[Tensor], boxes_h: List[Tensor],
                boxes_o: List[Tensor], object_class: List[Tensor], labels: List[Tensor]) -> List[dict]:
    """InteractionHead.postprocess head.py:291-337."""
    num_boxes = [len(b) for b in boxes_h]
    weights = torch.sigmoid(logits_s).squeeze(1).split(num_boxes)
    scores = torch.sigmoid(logits_p).split(num_boxes)
    if len(labels) == 0:
        labels = [None for _ in num_boxes]
    results = []
    for w, s, p, b_h, b_o, o, l in zip(weights, scores, prior, boxes_h, boxes_o, object_class, labels):
        x, y = torch.nonzero(p[0]).unbind(1)
        action_score = s[x, y] * p[:, x, y].prod(dim=0) * w[x].detach()
        r = dict(boxes_h=b_h, boxes_o=b_o, index=x, prediction=y, scores=action_score, object=o,
                 prior=p[:, x, y], weights=w)
        if l is not None:
            r["labels"] = l[x, y]
            r["unary_labels"] = l.sum(dim=1).clamp(max=1)
        results.append(r)
    return results


def interaction_head_forward(P: Dict[str, Tensor], tables, features3: Tensor, image_shapes,
                             box_features: Tensor, box_coords, box_labels, box_scores,
                             object_to_verb, targets=None, training: bool = False,
                             num_cls: int = 117, human_idx: int = 49, num_iter: int = 2,
                             row_pick: str = "loop", neg_perm=None, taps: Optional[dict] = None):
    """InteractionHead.forward head.py:341-429 from the pooled box features onward (the detection
    preprocess and torchvision RoIAlign, head.py:382-387, stay stock and sit outside the path).
    Returns (results, logits_p, logits_s); in training a loss dict is appended to ``results``
    with the intended margin-loss call (SURVEY defect 2: ``NegativeSampling.forward(cat(pos,neg))``)."""
    g = graph_head_forward(P, tables, features3, image_shapes, box_features, box_coords, box_labels,
                           box_scores, object_to_verb, targets, training, num_cls, human_idx,
                           num_iter, row_pick=row_pick, neg_perm=neg_perm, taps=taps)
    feat = torch.cat(g["feat"])                                               # head.py:408
    logits_p = _linear(P, "box_pair_predictor", feat)                         # head.py:410
    logits_s = _linear(P, "box_pair_suppressor", feat)                        # head.py:411
    results = postprocess(logits_p, logits_s, g["prior"], g["boxes_h"], g["boxes_o"], g["obj"],
                          g["labels"])
    if training:
        scores = torch.cat([r["scores"] for r in results]); lab = torch.cat([r["labels"] for r in results])
        n_p = len(torch.nonzero(lab))
        hoi = binary_focal_loss(scores, lab, reduction="sum", gamma=0.2) / n_p          # head.py:153-177
        w = torch.cat([r["weights"] for r in results]); ul = torch.cat([r["unary_labels"] for r in results])
        n_pu = len(torch.nonzero(ul))
        inter = binary_focal_loss(w, ul, reduction="sum", gamma=2.0) / n_pu             # head.py:180-205
        flat = torch.cat([torch.cat(g["pos"]), torch.cat(g["neg"])])
        transh = negative_sampling_loss(flat, margin=1.0) / n_pu                        # head.py:207-235
        results.append(dict(hoi_loss=hoi, interactiveness_loss=inter, transH_loss=transh))
    return results, logits_p, logits_s
