"""On-disk formats: reference checkpoint round trip, cached-detection JSON, (object, verb) -> interaction map."""
import json
import os

import torch

from helpers import build_head, synthetic
from skghoi_b200 import io


def test_reference_checkpoint_round_trip(tmp_path):
    src = build_head(42)
    src.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(7))
    # a reference-shaped checkpoint: head under 'interaction_head.', detector weights and trainer state beside it
    base = {"model_state_dict": {"detector_backbone.conv1.weight": torch.ones(2, 2), "interaction_head.stale": torch.zeros(1)},
            "optim_state_dict": {"state": {}}, "scheduler_state_dict": {}, "epoch": 3, "iteration": 1234}
    path = os.path.join(tmp_path, "ckpt.pt")
    ckpt = io.save_reference_checkpoint(src, path, base=base, epoch=4)
    assert ckpt["epoch"] == 4 and ckpt["iteration"] == 1234 and "interaction_head.stale" not in ckpt["model_state_dict"]
    keys = [k for k in ckpt["model_state_dict"] if k.startswith("interaction_head.") and ".transh." not in k]
    assert len(keys) == len(src.state_dict()) == 408
    assert "interaction_head.box_pair_head.attention_head.fc_1.15.weight" in ckpt["model_state_dict"]
    assert torch.equal(ckpt["model_state_dict"]["detector_backbone.conv1.weight"], torch.ones(2, 2))
    dst = build_head(1)                                     # different initial weights
    missing = io.load_reference_checkpoint(dst, path)
    assert not missing.missing_keys and not missing.unexpected_keys
    for (k, a), (_, b) in zip(src.state_dict().items(), dst.state_dict().items()):
        assert torch.equal(a, b), k
    ta, tb = src.box_pair_head.transh_head.transh, dst.box_pair_head.transh_head.transh
    assert torch.equal(ta.ent_embeddings.weight, tb.ent_embeddings.weight)
    assert torch.equal(ta.norm_vector.weight, tb.norm_vector.weight)
    # a pure reference checkpoint (no TransH tables) and a bare state_dict load as well
    pure = {"model_state_dict": {k: v for k, v in ckpt["model_state_dict"].items() if ".transh." not in k}}
    io.load_reference_checkpoint(build_head(2), pure)
    io.load_reference_checkpoint(build_head(2), src.state_dict())


def test_detection_json_and_filter(tmp_path):
    det = {"boxes": [[0.5, 1.0, 10.0, 20.0], [3.0, 4.0, 30.0, 40.0], [5.0, 5.0, 6.0, 6.0]], "labels": [49, 17, 49],
           "scores": [0.9, 0.15, 0.1]}
    path = os.path.join(tmp_path, "HICO_train2015_00000001.json")
    with open(path, "w") as f:
        json.dump(det, f)
    d = io.read_detections(path)
    assert d["boxes"].dtype == torch.float32 and d["labels"].dtype == torch.int64 and d["scores"].dtype == torch.float32
    assert d["boxes"].shape == (3, 4) and d["labels"].tolist() == [49, 17, 49]
    f = io.filter_detections(d, 49, 0.2, 0.1)               # utils.py:91-113: humans first, thresholds per group
    assert f["labels"].tolist() == [49, 17] and torch.allclose(f["scores"], torch.tensor([0.9, 0.15]))
    empty = io.read_detections({"boxes": [], "labels": [], "scores": []})
    assert empty["boxes"].shape == (0, 4)
    io.write_detections(d, path)
    assert torch.equal(io.read_detections(path)["boxes"], d["boxes"])


def test_interaction_table_matches_reference_double_loop():
    otv = synthetic.hico_object_to_verb()
    tab = io.InteractionTable.from_object_to_verb(otv)
    assert tab.num_interactions == sum(len(set(v)) for v in otv)
    nested = tab.table.tolist()
    g = torch.Generator().manual_seed(0)
    obj = torch.randint(0, 80, (12,), generator=g)
    index = torch.randint(0, 12, (40,), generator=g)
    pred = torch.tensor([otv[int(o)][int(torch.randint(0, len(otv[int(o)]), (1,), generator=g))] for o in obj[index]])
    res = dict(object=obj, index=index, prediction=pred)
    want = torch.tensor([nested[int(o)][int(v)] for o, v in zip(obj[index], pred)])      # utils.py:177-180
    assert torch.equal(tab(res), want) and bool((want >= 0).all())
