"""Capacity-bucket mode of the whole-path forward (SKG_HEAD_CAPACITY, csrc/forward.cu): launches sized by a bucket, true
counts and input addresses read from the segment table on the device, so ONE captured CUDA graph serves every batch
geometry of the bucket (real detection streams change (n_h, n) per image every batch, head.py:826-841).  Parity of every
step of a stream of different geometries against the CPU oracle, and the graph-cache counters."""
import pytest
import torch

from helpers import build_head, head_params, synthetic, assert_logits_parity
from oracle import head_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda:0"

STREAM = [  # all in the 1024-cell / 128-box / 128-human bucket except the last two
    [(4, 9), (3, 7), (8, 20), (1, 2)], [(5, 11), (2, 30), (7, 8), (6, 13)], [(15, 30), (1, 5), (3, 3), (2, 6)],
    [(9, 21), (8, 17), (4, 4), (1, 30)], [(1, 3), (6, 9), (10, 25), (3, 12)], [(15, 30), (15, 30), (12, 28), (9, 9)],
    [(1, 2), (1, 3), (2, 2), (1, 2)]]


def _head():
    ih = build_head(42)
    ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
    ih = ih.to(DEV).eval()
    ih.box_pair_head.transh_head.transh.to(DEV)
    return ih


def test_capacity_stream_parity_and_graph_replays():
    from skghoi_b200 import _lib, engine
    assert engine.CAPACITY_BUCKETS
    ih = _head()
    otv = synthetic.hico_object_to_verb()
    tables = synthetic.make_transh_tables(42)
    refs, devs = [], []
    for i, shapes in enumerate(STREAM):
        batch = synthetic.make_batch(shapes, seed=300 + i)
        with torch.no_grad():
            res, lp, ls = O.interaction_head_forward(head_params(), tables, batch["features"]["3"], batch["image_shapes"],
                                                     batch["box_features"], batch["box_coords"], batch["box_labels"],
                                                     batch["box_scores"], otv, row_pick="slice")
        refs.append((batch, res, torch.cat([lp, ls], 1)))
        devs.append(dict(f={"3": batch["features"]["3"].to(DEV)}, bf=batch["box_features"].to(DEV),
                         boxes=[b.to(DEV) for b in batch["box_coords"]], labels=[l.to(DEV) for l in batch["box_labels"]],
                         scores=[s.to(DEV) for s in batch["box_scores"]]))
    with torch.no_grad():     # size the workspace for the largest bucket first (a new workspace address re-keys every graph)
        batch, _, _ = refs[5]; d = devs[5]
        ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"])
    g0 = _lib.graph_stats()
    g_rep2 = None
    for rep in range(3):
        if rep == 2:
            g_rep2 = _lib.graph_stats()
        for (batch, res, lg_ref), d in zip(refs, devs):
            if rep == 2:      # the same detections at NEW addresses: the graph must follow the address block, not the capture
                d = dict(f={"3": d["f"]["3"].clone()}, bf=d["bf"].clone(), boxes=[b.clone() for b in d["boxes"]],
                         labels=[l.clone() for l in d["labels"]], scores=[s.clone() for s in d["scores"]])
            with torch.no_grad():
                got = ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"])
            torch.cuda.synchronize()
            run = ih.__dict__["_skg_last_run"]
            assert run["logits"].shape == lg_ref.shape
            assert_logits_parity(run["logits"].cpu(), lg_ref, "capacity stream rep %d" % rep)
            assert len(got) == len(res)
            for g, r in zip(got, res):
                assert torch.equal(g["index"].cpu(), r["index"]) and torch.equal(g["prediction"].cpu(), r["prediction"])
                assert torch.equal(g["object"].cpu(), r["object"])
                torch.testing.assert_close(g["boxes_h"].cpu(), r["boxes_h"], rtol=0, atol=0)
                torch.testing.assert_close(g["boxes_o"].cpu(), r["boxes_o"], rtol=0, atol=0)
                if r["scores"].numel():
                    assert float((g["scores"].cpu() - r["scores"]).abs().max()) <= 1e-2 * float(r["scores"].abs().max()) + 1e-7
    g1 = _lib.graph_stats()
    replays, captures, plain = (a - b for a, b in zip(g1, g0))
    # every capacity bucket runs plain once, is captured once, every other step is a replay (7 geometries share <= 6 buckets)
    # the 7 geometries fall into 5 buckets (cells 128 / 256 / 384 / 512 / 1408)
    assert captures <= 5 and plain <= 5 and replays >= 3 * len(STREAM) - 10, (replays, captures, plain)
    # the third pass -- same detections in buffers at NEW addresses -- is replays only
    assert (g1[0] - g_rep2[0], g1[1] - g_rep2[1], g1[2] - g_rep2[2]) == (len(STREAM), 0, 0), (g_rep2, g1)


def test_capacity_matches_exact_geometry_path():
    """Same batch through the bucketed and through the exact-geometry launch sequence: identical integer outputs, logits
    equal up to the fp32 summation order of the two launch sequences (different tile shapes; split-K partials arrive in a
    different order every run).  A last-bit fp32 difference can flip the bf16 rounding of a hidden activation (2^-9 relative
    on that element), so single logits move by up to a few 1e-3 of the logits' scale while the bulk stays put: bound the
    relative L2 tightly and the worst element at half the parity tolerance."""
    from skghoi_b200 import engine
    ih = _head()
    batch = synthetic.make_batch([(7, 19), (3, 5), (12, 12)], seed=11)
    meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
    args = (batch["features"]["3"].to(DEV), batch["image_shapes"], batch["box_features"].to(DEV), torch.cat(batch["box_coords"]).to(DEV))
    outs = {}
    for mode in (True, False):
        engine.CAPACITY_BUCKETS = mode
        try:
            with torch.no_grad():
                for _ in range(3):
                    run = engine.head_forward(ih, *args, dict(meta), precise_logits=True, graph=True)
            torch.cuda.synchronize()
            outs[mode] = {k: run[k].clone() for k in ("logits", "x_keep", "y_keep", "cell2pair", "pair2cell", "enc")}
        finally:
            engine.CAPACITY_BUCKETS = True
    for k in ("x_keep", "y_keep", "cell2pair", "pair2cell", "enc"):
        assert torch.equal(outs[True][k], outs[False][k]), k
    diff = (outs[True]["logits"] - outs[False]["logits"]).double()
    ref = outs[False]["logits"].double()
    d, scale = float(diff.abs().max()), float(ref.abs().max())
    rel_l2 = float(diff.norm() / ref.norm())
    print("capacity vs exact: max|d| / max|ref| = %.3e, rel L2 = %.3e" % (d / scale, rel_l2))
    assert rel_l2 <= 1e-3, (rel_l2, d / scale)
    assert d <= 5e-3 * scale, (d / scale, rel_l2)


def test_session_error_behaviour_matches_reference():
    """The serving session keeps the reference's errors: IndexError for an image with more boxes than the entity table has
    rows (head.py:572), ValueError when humans are not first (head.py:840-841) -- also when the caller hands in its own
    host meta, i.e. when the check runs inside skg_session_prepare."""
    import numpy as np
    from skghoi_b200 import engine
    ih = _head()
    batch = synthetic.make_batch([(3, 81)], seed=5)
    args = ({"3": batch["features"]["3"].to(DEV)}, batch["image_shapes"], batch["box_features"].to(DEV),
            [b.to(DEV) for b in batch["box_coords"]], [l.to(DEV) for l in batch["box_labels"]], [s.to(DEV) for s in batch["box_scores"]])
    with torch.no_grad(), pytest.raises(IndexError):
        ih.forward_from_pooled(*args)
    batch = synthetic.make_batch([(3, 7)], seed=6)
    labels = batch["box_labels"][0].clone()
    labels[[0, 5]] = labels[[5, 0]]                           # a human behind an object
    args = ({"3": batch["features"]["3"].to(DEV)}, batch["image_shapes"], batch["box_features"].to(DEV),
            [b.to(DEV) for b in batch["box_coords"]], [labels.to(DEV)], [s.to(DEV) for s in batch["box_scores"]])
    with torch.no_grad(), pytest.raises(ValueError, match="not permuted"):
        ih.forward_from_pooled(*args)
    meta = dict(n=[7], n_h=[3], labels=labels.numpy().astype(np.int64), scores=batch["box_scores"][0].numpy().astype(np.float32))
    with torch.no_grad(), pytest.raises(ValueError, match="not permuted"):
        ih.forward_from_pooled(*args, meta=meta)


def test_session_skipped_and_empty_images_match_exact_path():
    """Images the reference skips (no human, or a single box: head.py:829-839) and an all-empty batch through the session
    path against the per-kernel orchestration (documented divergence from the reference: a skipped image yields an empty
    result instead of misaligned rows, DESIGN 2 defect 6)."""
    from skghoi_b200 import engine
    ih = _head()
    batch = synthetic.make_batch([(0, 3), (4, 9), (1, 1), (2, 6)], seed=31)
    args = ({"3": batch["features"]["3"].to(DEV)}, batch["image_shapes"], batch["box_features"].to(DEV),
            [b.to(DEV) for b in batch["box_coords"]], [l.to(DEV) for l in batch["box_labels"]], [s.to(DEV) for s in batch["box_scores"]])
    outs = {}
    for mode in (True, False):
        engine.SESSION_PATH = mode
        try:
            with torch.no_grad():
                res = ih.forward_from_pooled(*args)
            torch.cuda.synchronize()
            outs[mode] = [{k: v.clone() for k, v in r.items()} for r in res]
        finally:
            engine.SESSION_PATH = True
    assert len(outs[True]) == len(outs[False]) == 4
    for a, b in zip(outs[True], outs[False]):
        for k in ("index", "prediction", "object", "boxes_h", "boxes_o"):
            assert torch.equal(a[k], b[k]), k
        if b["scores"].numel():
            # (two launch sequences: a last-bit fp32 difference may flip a bf16 rounding upstream, see the test above)
            assert float((a["scores"] - b["scores"]).abs().max()) <= 5e-3 * float(b["scores"].abs().max()) + 1e-7
    assert outs[True][0]["index"].numel() == 0 and outs[True][2]["index"].numel() == 0
