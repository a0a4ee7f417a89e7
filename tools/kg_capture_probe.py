import sys, torch
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests/golden")
import make_kg_golden as MG
from skghoi_b200.kg_data import KGIndex
from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
from skghoi_b200.kge import MarginLoss, TransH
h, t, r = MG.graph()
idx = KGIndex(h, t, r, 80, 117, device="cuda:0")
loader = TrainDataLoader(idx, nbatches=1, neg_ent=5, bern_flag=True, threads=4, seeds=MG.seeds_for(4))
model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to("cuda:0")
strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to("cuda:0")
opt = torch.optim.SGD(model.parameters(), lr=0.5)
def stage(name, fn):
    s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2): fn()
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g):
            fn()
        g.replay(); torch.cuda.synchronize(); print(name, "OK")
    except Exception as e:
        print(name, "FAILED:", str(e).splitlines()[0][:200])
        torch.cuda.synchronize()
data = loader.sampler.sample()
stage("sample", lambda: loader.sampler.sample())
stage("score", lambda: model._score(data["batch_h"], data["batch_r"], data["batch_t"], False)[4])
stage("loss fwd", lambda: strat(data))
def fb():
    opt.zero_grad(set_to_none=True); l = strat(data); l.backward(); return l
stage("fwd+bwd", fb)
def full():
    opt.zero_grad(set_to_none=True); l = strat(data); l.backward(); opt.step(); return l
stage("fwd+bwd+sgd", full)
