"""Trainer(graph=True) against the eager loop and the eager loop against itself (run-to-run noise of the fp32 atomics)."""
import sys, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests/golden")
import make_kg_golden as MG
from skghoi_b200.kg_data import KGIndex
from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
from skghoi_b200.kge import MarginLoss, TransH
h, t, r = MG.graph()
idx = KGIndex(h, t, r, 80, 117, device="cuda:0")
def run(graph, nb=1, epochs=11):
    loader = TrainDataLoader(idx, nbatches=nb, neg_ent=5, bern_flag=True, threads=4, seeds=MG.seeds_for(4))
    torch.manual_seed(0)
    model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to("cuda:0")
    strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to("cuda:0")
    hist = Trainer(strat, loader, train_times=epochs, alpha=0.5, opt_method="sgd", graph=graph).run()
    return torch.tensor(hist), model.ent_embeddings.weight.detach().cpu().clone(), loader.sampler.seeds()
for nb in (1, 4):
    a, b, g = run(False, nb), run(False, nb), run(True, nb)
    print("nb=%d eager/eager: loss %.2e emb %.2e | graph/eager: loss %.2e emb %.2e seeds %s" % (
        nb, float((a[0] - b[0]).abs().max()), float((a[1] - b[1]).abs().max()),
        float((g[0] - a[0]).abs().max()), float((g[1] - a[1]).abs().max()), g[2] == a[2]))
    print("  eager", [round(float(v), 4) for v in a[0]]); print("  graph", [round(float(v), 4) for v in g[0]])
