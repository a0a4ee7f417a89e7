"""Derives the HICO-DET object->verb table (80 lists, 600 entries) from the KEYS of
``hicodet/hico_text_label.py:1-600`` of the reference checkout; equivalent to
``HICODet.object_to_verb`` (hicodet/hicodet.py:168-179) without the annotation JSON.
Run in the build container only:  python tests/golden/make_object_to_verb.py
"""
import json, os, runpy, sys

ref = os.environ.get("SKGHOI_REFERENCE", "/root/reference")
ns = runpy.run_path(os.path.join(ref, "hicodet", "hico_text_label.py"))
pairs = sorted(ns["hico_text_label"].keys(), key=lambda vo: (vo[1], vo[0]))
table = [[] for _ in range(80)]
for verb, obj in pairs:
    table[obj].append(verb)
assert sum(len(t) for t in table) == 600 and all(len(t) > 0 for t in table)
out = os.path.join(os.path.dirname(__file__), "..", "..", "skghoi_b200", "data", "hico_object_to_verb.json")
with open(out, "w") as f:
    json.dump(table, f)
print("wrote", os.path.abspath(out), [len(t) for t in table][:10])
