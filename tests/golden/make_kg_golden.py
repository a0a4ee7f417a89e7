"""Golden vectors of the KG negative sampler from the REAL reference code: oracle/_ref/Base.so, compiled by
oracle/Makefile from /root/reference/OpenKE/openke/base/Base.cpp.  Build container only:
    make -C oracle && python tests/golden/make_kg_golden.py
One process per configuration (Base.so keeps global state).  Writes tests/golden/kg_sampler.pt."""
import os
import subprocess
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.join(HERE, "..", "..")
sys.path.insert(0, ROOT)
CALLS = [(64, 1, 0), (100, 25, 0), (37, 3, -1), (37, 3, 1), (256, 5, 0)]     # (batch_size, neg_rate, mode), consecutive


def graph():
    """HOI label graph (human, verb, object) plus random extra triples and duplicates; (49, no_interaction, o) is dropped
    for every 8th object so that no (head, relation) is linked to ALL 80 entities (the reference divides by zero there)."""
    from skghoi_b200 import synthetic
    otv = synthetic.hico_object_to_verb()
    h, t, r = [], [], []
    for o in range(80):
        for v in otv[o]:
            if v == 57 and o % 8 == 0:
                continue
            h.append(49); r.append(v); t.append(o)
    rng = np.random.RandomState(0)
    for _ in range(400):
        h.append(int(rng.randint(0, 80))); t.append(int(rng.randint(0, 80))); r.append(int(rng.randint(0, 117)))
    return h + h[:50], t + t[:50], r + r[:50]


def seeds_for(W):
    return [12345 + 977 * i for i in range(W)]


def worker(bern, W, out):
    from oracle import kg_sampler_oracle as K
    h, t, r = graph()
    ref = K.ReferenceSampler(h, t, r, 80, 117, W, bool(bern))
    ref.set_seeds(seeds_for(W))
    calls = []
    for bs, neg, mode in CALLS:
        bh, bt, br, by = ref.sample(bs, neg, mode)
        calls.append(dict(batch_h=torch.from_numpy(bh), batch_t=torch.from_numpy(bt), batch_r=torch.from_numpy(br),
                          batch_y=torch.from_numpy(by), seeds_after=[int(s) for s in ref.get_seeds()]))
    torch.save(dict(train_total=ref.train_total, calls=calls), out)


if __name__ == "__main__":
    if len(sys.argv) == 4:
        worker(int(sys.argv[1]), int(sys.argv[2]), sys.argv[3])
        sys.exit(0)
    gold = {}
    for bern in (0, 1):
        for W in (1, 3, 8):
            tmp = "/tmp/kg_gold_%d_%d.pt" % (bern, W)
            subprocess.check_call([sys.executable, os.path.abspath(__file__), str(bern), str(W), tmp], stdout=subprocess.DEVNULL)
            gold[(bern, W)] = torch.load(tmp, weights_only=False)
    torch.save(dict(calls_spec=CALLS, cases=gold), os.path.join(HERE, "kg_sampler.pt"))
    print("wrote kg_sampler.pt", {k: v["train_total"] for k, v in gold.items()})
