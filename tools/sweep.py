"""BASELINE configs[4]: pair-count stress sweep on one GPU -- (n_h, n) x images per step -> pairs/s with device-resident
inputs (CUDA events, L2 flushed between steps) -- and configs[3]: TransH KG-embedding training step on the HOI label graph.
    python tools/sweep.py > gpurun_out/sweep.md      (on the GPU box)"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from skghoi_b200 import engine, synthetic                 # noqa: E402
from skghoi_b200.head import build_interaction_head       # noqa: E402
from skghoi_b200.kge import MarginLoss, NegativeSampling, TransH   # noqa: E402

dev = torch.device("cuda:0")
otv = synthetic.hico_object_to_verb()
ih = build_interaction_head(otv, seed=42)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).eval(); ih.box_pair_head.transh_head.transh.to(dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
print("| n_h | n | images/step | cells/step | pairs/step | ms/step | M pairs/s | GEMM TFLOP/s issued |")
print("|---:|---:|---:|---:|---:|---:|---:|---:|")
for nh, n in ((4, 8), (8, 16), (16, 32), (32, 64), (40, 80), (64, 64)):
    for B in (1, 8, 64, 512):
        cells = B * nh * n
        if cells > 600000:                                   # keep the sweep inside a minute of GPU time
            continue
        batch = synthetic.make_batch([(nh, n)] * B, seed=7)
        d = dict(f3={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev), boxes=[b.to(dev) for b in batch["box_coords"]],
                 labels=[l.to(dev) for l in batch["box_labels"]], scores=[s.to(dev) for s in batch["box_scores"]])
        meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)

        def step():
            with torch.no_grad():
                return engine.interaction_head_from_pooled(ih, d["f3"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"],
                                                           d["scores"], meta=meta)
        for _ in range(3):
            step(); flush.zero_()
        torch.cuda.synchronize()
        K = 10
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        for s, e in ev:
            s.record(); step(); e.record(); flush.zero_()
        torch.cuda.synchronize()
        ms = sum(s.elapsed_time(e) for s, e in ev) / K
        pairs = B * nh * (n - 1)
        flops = 2.0 * (cells * (128 * 256 + 256 * 1024 + 1024 * 4096 + 2 * 1024 * 1024) + pairs * (2 * 1024 * 1024 + 4096 * 128)
                       + B * n * (12544 * 1024 + 1024 * 1024 + 2 * 1074 * 1024 + 4 * 1024 * 1024))
        print("| %d | %d | %d | %d | %d | %.3f | %.2f | %.0f |" % (nh, n, B, cells, pairs, ms, pairs / ms / 1e3, flops / ms / 1e9))
        del d, batch

# ---- configs[3]: TransH training on the HOI label graph (80 objects, 117 verbs): positives (human, verb, object)
print()
print("| KG training: positives/step | negatives per positive | triples/step | ms/step | M triples/s |")
print("|---:|---:|---:|---:|---:|")
pos = torch.tensor([(49, v, o) for o in range(80) for v in otv[o]], dtype=torch.int64, device=dev)
for neg in (1, 25):
    model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to(dev)
    strat = NegativeSampling(model=None, loss=MarginLoss(margin=4.0), batch_size=pos.shape[0]).to(dev)
    opt = torch.optim.SGD(model.parameters(), lr=0.5)
    g = torch.Generator(device=dev).manual_seed(0)

    def kg_step():
        # uniform tail / head corruption (the reference's C++ sampler, Base.cpp:78-197, is SURVEY 8(f3) "next")
        npos = pos.shape[0]
        rep = pos.repeat(neg, 1)
        corrupt_tail = torch.rand(rep.shape[0], device=dev, generator=g) < 0.5
        rnd = torch.randint(0, 80, (rep.shape[0],), device=dev, generator=g)
        rep = torch.stack([torch.where(corrupt_tail, rep[:, 0], rnd), rep[:, 1], torch.where(corrupt_tail, rnd, rep[:, 2])], 1)
        allt = torch.cat([pos, rep])
        score = model(allt[:, 0], allt[:, 1], allt[:, 2])[4]
        p = score[:npos].view(-1, 1); nsc = score[npos:].view(neg, npos).t()
        loss = strat.loss(p, nsc)
        opt.zero_grad(set_to_none=True); loss.backward(); opt.step()
        return loss
    for _ in range(5):
        l0 = kg_step()
    torch.cuda.synchronize()
    K = 50
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(K):
        l1 = kg_step()
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / K
    T = pos.shape[0] * (1 + neg)
    print("| %d | %d | %d | %.3f | %.2f |   loss %.3f -> %.3f" % (pos.shape[0], neg, T, ms, T / ms / 1e3, float(l0), float(l1)))

# ---- KG negative sampler: skg_kg_sample vs the reference's own Base.so on the host cores (same graph, same seeds)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import make_kg_golden as MG                                  # noqa: E402
from skghoi_b200.kg_data import KGIndex, TrainDataSampler    # noqa: E402
from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer   # noqa: E402
h, t, r = MG.graph()
idx = KGIndex(h, t, r, 80, 117, device=dev)
print()
print("| KG sampler | batch | neg | triples/call | us/call | M triples/s |")
print("|---|---:|---:|---:|---:|---:|")
for bs, neg in ((990, 25), (1 << 16, 25), (1 << 20, 8)):
    smp = TrainDataSampler(idx, bs, neg, True, 8, MG.seeds_for(8))
    for _ in range(3):
        smp.sample()
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    K = 20
    s.record()
    for _ in range(K):
        smp.sample()
    e.record(); torch.cuda.synchronize()
    us = s.elapsed_time(e) / K * 1e3
    print("| skg_kg_sample (B200) | %d | %d | %d | %.1f | %.1f |" % (bs, neg, bs * (1 + neg), us, bs * (1 + neg) / us))
try:
    from oracle import kg_sampler_oracle as KO
    cores = len(os.sched_getaffinity(0))
    ref = KO.ReferenceSampler(h, t, r, 80, 117, cores, True)
    ref.set_seeds(MG.seeds_for(cores))
    for bs, neg in ((990, 25), (1 << 16, 25)):
        ref.sample(bs, neg, 0)
        t0 = time.perf_counter()
        K = 10
        for _ in range(K):
            ref.sample(bs, neg, 0)
        us = (time.perf_counter() - t0) / K * 1e6
        print("| reference Base.so, %d host threads | %d | %d | %d | %.1f | %.1f |" % (cores, bs, neg, bs * (1 + neg), us, bs * (1 + neg) / us))
except Exception as ex:                                      # noqa: BLE001
    print("| reference Base.so unavailable: %s |" % ex)
loader = TrainDataLoader(idx, nbatches=1, neg_ent=25, bern_flag=True, threads=8, seeds=MG.seeds_for(8))
model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to(dev)
strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to(dev)
tr = Trainer(strat, loader, train_times=20, alpha=0.5)
tr.run(); torch.cuda.synchronize()
t0 = time.perf_counter(); hist = tr.run(); torch.cuda.synchronize(); dt = time.perf_counter() - t0
T = loader.get_batch_size() * 26
print()
print("KG training with the device sampler (Trainer.run, %d positives x 25 negatives per step): %.3f ms/step, %.1f M triples/s, loss %.3f -> %.3f"
      % (loader.get_batch_size(), dt / 20 * 1e3, T * 20 / dt / 1e6, hist[0], hist[-1]))
