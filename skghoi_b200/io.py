"""On-disk formats either side of the hot path: reference checkpoints, cached-detection JSON files, and the
(object, verb) -> interaction-class mapping the reference applies to the head's results at test time.

Nothing here touches a kernel; it is the glue a user of the reference needs to switch heads without converting files.
"""
import json
from typing import Dict, List, Optional, Sequence, Union

import torch
from torch import Tensor

PREFIX = "interaction_head."      # the reference network shell registers the head under this name (models.py:45)


def head_state_from_checkpoint(checkpoint: Union[str, dict]) -> Dict[str, Tensor]:
    """The head's ``state_dict`` out of a reference checkpoint (``{'model_state_dict': {'interaction_head.*': ...,
    'detector_backbone.*': ..., ...}, 'optim_state_dict': ..., 'epoch': ...}``, configures/hicodet/
    adamixer_transH_spatial_r50_main.py:85-93).  A bare state_dict (with or without the prefix) is accepted too."""
    if isinstance(checkpoint, str):
        checkpoint = torch.load(checkpoint, map_location="cpu", weights_only=False)
    sd = checkpoint.get("model_state_dict", checkpoint)
    if any(k.startswith(PREFIX) for k in sd):
        return {k[len(PREFIX):]: v for k, v in sd.items() if k.startswith(PREFIX)}
    return dict(sd)


def load_reference_checkpoint(ih, checkpoint: Union[str, dict], strict: bool = True):
    """Loads the interaction-head weights of a reference checkpoint into ``ih`` (same parameter names, SURVEY 8(a) state
    dict table).  The TransH tables are not part of a reference checkpoint (the reference re-draws them on every call,
    head.py:574-578); keys under ``transh.`` written by ``save_reference_checkpoint`` restore them when present.
    In-place copies bump the parameter versions, so the packed bf16 operands are rebuilt on the next forward."""
    sd = head_state_from_checkpoint(checkpoint)
    tables = {k[len("transh."):]: v for k, v in sd.items() if k.startswith("transh.")}
    sd = {k: v for k, v in sd.items() if not k.startswith("transh.")}
    out = ih.load_state_dict(sd, strict=strict)
    if tables:
        ih.box_pair_head.transh_head.set_tables(tables["ent_embeddings.weight"], tables["rel_embeddings.weight"],
                                                tables["norm_vector.weight"])
    ih.__dict__.pop("_skg_packed", None)
    return out


def save_reference_checkpoint(ih, path: Optional[str] = None, base: Optional[dict] = None, with_tables: bool = True,
                              **extra) -> dict:
    """Writes ``ih`` back in the reference layout: per-branch parameters under ``model_state_dict['interaction_head.*']``.
    ``base`` (a loaded reference checkpoint) keeps its detector weights, optimiser / scheduler state, epoch and iteration;
    ``extra`` overrides or adds top-level entries (``epoch=...``)."""
    ckpt = dict(base) if base is not None else {}
    msd = {k: v for k, v in ckpt.get("model_state_dict", {}).items() if not k.startswith(PREFIX)}
    for k, v in ih.state_dict().items():
        msd[PREFIX + k] = v.detach().cpu().clone()
    if with_tables:
        tr = ih.box_pair_head.transh_head.transh
        for k, v in tr.state_dict().items():
            msd[PREFIX + "transh." + k] = v.detach().cpu().clone()
    ckpt["model_state_dict"] = msd
    ckpt.update(extra)
    if path is not None:
        torch.save(ckpt, path)
    return ckpt


def read_detections(src: Union[str, dict], device=None) -> Dict[str, Tensor]:
    """One image's cached detections in the reference wire format -- a JSON object ``{"boxes": [[x1,y1,x2,y2],...],
    "labels": [...], "scores": [...]}`` (hicodet/detections/adamixer_preprocessing.py:104-135), read by the dataset as
    ``pocket.ops.to_tensor(json.load(f), input_format='dict')`` (utils.py:137) -- as float32 / int64 / float32 tensors."""
    if isinstance(src, str):
        with open(src, "r") as f:
            src = json.load(f)
    out = dict(boxes=torch.as_tensor(src["boxes"], dtype=torch.float32).reshape(-1, 4),
               labels=torch.as_tensor(src["labels"], dtype=torch.int64).reshape(-1),
               scores=torch.as_tensor(src["scores"], dtype=torch.float32).reshape(-1))
    if not (out["boxes"].shape[0] == out["labels"].shape[0] == out["scores"].shape[0]):
        raise ValueError("boxes / labels / scores of a detection file must have the same length")
    if device is not None:
        out = {k: v.to(device) for k, v in out.items()}
    return out


def write_detections(detection: Dict[str, Tensor], path: str) -> None:
    with open(path, "w") as f:
        json.dump({k: detection[k].detach().cpu().tolist() for k in ("boxes", "labels", "scores")}, f)


def filter_detections(detection: Dict[str, Tensor], human_idx: int, box_score_thresh_h: float,
                      box_score_thresh_o: float) -> Dict[str, Tensor]:
    """DataFactory.filter_detections (utils.py:91-113): humans above their threshold first, then the other classes above
    theirs (each group in file order)."""
    boxes, labels, scores = detection["boxes"], detection["labels"], detection["scores"]
    idx = torch.nonzero(labels == human_idx).squeeze(1)
    keep = idx[torch.nonzero(scores[idx] >= box_score_thresh_h).squeeze(1)]
    idx = torch.nonzero(labels != human_idx).squeeze(1)
    keep = torch.cat([keep, idx[torch.nonzero(scores[idx] >= box_score_thresh_o).squeeze(1)]])
    return dict(boxes=boxes[keep].view(-1, 4), labels=labels[keep].view(-1), scores=scores[keep].view(-1))


class InteractionTable:
    """``object_n_verb_to_interaction[o][v]`` (utils.py:177-180: a Python double loop over every detected pair) as ONE
    dense [num_objects, num_verbs] int64 lookup on the results' device; invalid (object, verb) combinations map to -1."""

    def __init__(self, object_n_verb_to_interaction: Sequence[Sequence[Optional[int]]]):
        rows = [[-1 if v is None else int(v) for v in r] for r in object_n_verb_to_interaction]
        self.table = torch.tensor(rows, dtype=torch.int64)

    @classmethod
    def from_object_to_verb(cls, object_to_verb: List[List[int]], num_verbs: int = 117) -> "InteractionTable":
        """HICO-DET numbering: interaction classes enumerate the valid (object, verb) combinations object by object,
        verbs ascending inside an object (600 classes for the 80 x 117 label graph)."""
        rows, k = [], 0
        for verbs in object_to_verb:
            row = [None] * num_verbs
            for v in sorted(set(int(x) for x in verbs)):
                row[v] = k; k += 1
            rows.append(row)
        return cls(rows)

    @property
    def num_interactions(self) -> int:
        return int(self.table.max()) + 1

    def __call__(self, result: Dict[str, Tensor]) -> Tensor:
        """Interaction class of every scored entry of one result dict (``object[index]`` x ``prediction``)."""
        dev = result["prediction"].device
        if self.table.device != dev:
            self.table = self.table.to(dev)
        return self.table[result["object"][result["index"]], result["prediction"]]
