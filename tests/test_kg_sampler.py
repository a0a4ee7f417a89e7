"""KG negative sampler.  CPU: the oracle restatement against golden vectors produced by the REAL reference
(oracle/_ref/Base.so) and, when that binary is present, against the binary itself; the product's index builder against the
oracle's.  GPU: skg_kg_sample bit-exact against the goldens and the oracle."""
import importlib.util
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from helpers import load_golden, ROOT
from oracle import kg_sampler_oracle as K

_spec = importlib.util.spec_from_file_location("make_kg_golden", os.path.join(ROOT, "tests", "golden", "make_kg_golden.py"))
MG = importlib.util.module_from_spec(_spec); _spec.loader.exec_module(MG)


@pytest.mark.parametrize("bern", [0, 1])
@pytest.mark.parametrize("W", [1, 3, 8])
def test_oracle_matches_reference_golden(bern, W):
    gold = load_golden("kg_sampler")
    case = gold["cases"][(bern, W)]
    h, t, r = MG.graph()
    idx = K.KGIndexOracle(h, t, r, 80, 117)
    assert idx.train_total == case["train_total"]
    seeds = MG.seeds_for(W)
    for (bs, neg, mode), g in zip(gold["calls_spec"], case["calls"]):
        bh, bt, br, by, seeds = K.sample(idx, seeds, bs, neg, mode, bool(bern), W)
        assert np.array_equal(bh, g["batch_h"].numpy()) and np.array_equal(bt, g["batch_t"].numpy())
        assert np.array_equal(br, g["batch_r"].numpy()) and np.array_equal(by, g["batch_y"].numpy())
        assert seeds == g["seeds_after"]


@pytest.mark.skipif(not os.path.isfile(K.REF_SO), reason="oracle/_ref/Base.so not built (needs /root/reference)")
def test_golden_is_what_the_reference_binary_produces(tmp_path):
    out = os.path.join(tmp_path, "g.pt")
    subprocess.check_call([sys.executable, os.path.join(ROOT, "tests", "golden", "make_kg_golden.py"), "1", "3", out],
                          stdout=subprocess.DEVNULL)
    fresh = torch.load(out, weights_only=False)
    case = load_golden("kg_sampler")["cases"][(1, 3)]
    for a, b in zip(fresh["calls"], case["calls"]):
        assert all(torch.equal(a[k], b[k]) for k in ("batch_h", "batch_t", "batch_r", "batch_y")) and a["seeds_after"] == b["seeds_after"]


def test_product_index_builder_matches_oracle():
    from skghoi_b200.kg_data import KGIndex
    h, t, r = MG.graph()
    o = K.KGIndexOracle(h, t, r, 80, 117)
    p = KGIndex(h, t, r, 80, 117, device="cpu")
    assert p.train_total == o.train_total
    assert p.head.t().tolist() == [list(x) for x in o.head] and p.tail.t().tolist() == [list(x) for x in o.tail]
    for a, b in ((p.lef_head, o.lef_head), (p.rig_head, o.rig_head), (p.lef_tail, o.lef_tail), (p.rig_tail, o.rig_tail)):
        # entities that never occur keep (0, -1) in both
        assert a.tolist() == list(b)
    used = np.isfinite(o.left_mean)
    assert np.array_equal(p.left_mean_host[used], o.left_mean[used]) and np.array_equal(p.right_mean_host[used], o.right_mean[used])


def test_lcg_jump_ahead_identity():
    # the device kernel jumps to (item - lef) * draws; check the closed form against stepping, in Python integers
    def jump(s, k):
        am, ap, cm, cp = 1, 0, K.LCG_A, K.LCG_C
        while k:
            if k & 1:
                am = (am * cm) & K.MASK; ap = (ap * cm + cp) & K.MASK
            cp = ((cm + 1) * cp) & K.MASK; cm = (cm * cm) & K.MASK; k >>= 1
        return (am * s + ap) & K.MASK
    s = 0xDEADBEEFCAFEF00D
    x = s
    for k in range(0, 200):
        assert jump(s, k) == x
        x = K.lcg(x)


@pytest.mark.gpu
@pytest.mark.parametrize("bern", [0, 1])
@pytest.mark.parametrize("W", [1, 3, 8])
def test_device_sampler_bit_exact_vs_reference(bern, W):
    from skghoi_b200.kg_data import KGIndex, TrainDataSampler
    gold = load_golden("kg_sampler")
    case = gold["cases"][(bern, W)]
    h, t, r = MG.graph()
    idx = KGIndex(h, t, r, 80, 117, device="cuda:0")
    seeds = MG.seeds_for(W)
    for (bs, neg, mode), g in zip(gold["calls_spec"], case["calls"]):
        smp = TrainDataSampler(idx, bs, neg, bool(bern), W, seeds, mode)
        b = smp.sample()
        for k in ("batch_h", "batch_t", "batch_r", "batch_y"):
            assert torch.equal(b[k].cpu(), g[k]), (k, bs, neg, mode)
        seeds = smp.seeds()
        assert seeds == g["seeds_after"]
    # consecutive calls of ONE sampler continue the streams: equal to the oracle stepping the same seeds
    smp = TrainDataSampler(idx, 96, 4, bool(bern), W, MG.seeds_for(W))
    o = K.KGIndexOracle(h, t, r, 80, 117)
    s = MG.seeds_for(W)
    for _ in range(3):
        b = smp.sample()
        bh, bt, br, by, s = K.sample(o, s, 96, 4, 0, bool(bern), W)
        assert np.array_equal(b["batch_h"].cpu().numpy(), bh) and np.array_equal(b["batch_t"].cpu().numpy(), bt)
        assert np.array_equal(b["batch_r"].cpu().numpy(), br)


@pytest.mark.gpu
def test_device_sampler_properties_large_and_saturated_relation():
    """Size-independent properties at a large batch: positives are training triples, no negative is a training triple
    (filtered corruption), exactly one side differs from its positive; a (head, relation) linked to every entity keeps the
    positive where the reference would divide by zero."""
    from skghoi_b200.kg_data import KGIndex, TrainDataSampler
    h, t, r = MG.graph()
    idx = KGIndex(h, t, r, 80, 117, device="cuda:0")
    bs, neg = 1 << 16, 8
    b = TrainDataSampler(idx, bs, neg, True, 8, list(range(1, 9))).sample()
    key = lambda hh, rr, tt: (hh * 117 + rr) * 80 + tt                        # noqa: E731
    train = set(key(idx.head[0], idx.head[1], idx.head[2]).cpu().tolist())
    bh, bt, br = b["batch_h"].cpu(), b["batch_t"].cpu(), b["batch_r"].cpu()
    assert set(key(bh[:bs], br[:bs], bt[:bs]).tolist()) <= train
    negs = key(bh[bs:], br[bs:], bt[bs:]).tolist()
    assert not (set(negs) & train)
    ph, pt = bh[:bs].repeat(neg), bt[:bs].repeat(neg)
    assert bool(((bh[bs:] != ph) ^ (bt[bs:] != pt)).all()) and torch.equal(br[bs:], br[:bs].repeat(neg))
    assert torch.equal(b["batch_y"].cpu(), torch.cat([torch.ones(bs), -torch.ones(bs * neg)]))
    # saturated: relation 5 links head 3 to all 6 entities
    sat = KGIndex([3] * 6 + [0], list(range(6)) + [1], [5] * 6 + [2], 6, 7, device="cuda:0")
    b = TrainDataSampler(sat, 256, 2, False, 2, [7, 9], mode=1).sample()      # always corrupt the tail
    is_sat = b["batch_r"][:256] == 5
    assert bool(is_sat.any())
    kept = b["batch_t"][256:].view(2, 256)[:, is_sat] == b["batch_t"][:256][is_sat]
    assert bool(kept.all())


@pytest.mark.gpu
def test_kg_training_loop_first_loss_vs_oracle_and_descent():
    """configs[3]: TransH on the HOI label graph with negative sampling.  The first step's loss equals the oracle's
    (TransH score + margin loss on the very batch the reference sampler draws); the loop then drives the loss down."""
    from oracle import head_oracle as O
    from skghoi_b200 import synthetic
    from skghoi_b200.kg_data import KGIndex
    from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
    from skghoi_b200.kge import MarginLoss, TransH
    h, t, r = MG.graph()
    idx = KGIndex(h, t, r, 80, 117, device="cuda:0")
    W, neg, nb = 4, 5, 5
    loader = TrainDataLoader(idx, nbatches=nb, neg_ent=neg, bern_flag=True, threads=W, seeds=MG.seeds_for(W))
    B = loader.get_batch_size()
    assert B == idx.train_total // nb
    torch.manual_seed(0)
    model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to("cuda:0")
    ent, rel, nrm = (x.detach().cpu().clone() for x in (model.ent_embeddings.weight, model.rel_embeddings.weight, model.norm_vector.weight))
    strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=B).to("cuda:0")
    # oracle loss of the first batch
    o = K.KGIndexOracle(h, t, r, 80, 117)
    bh, bt, br, by, _ = K.sample(o, MG.seeds_for(W), B, neg, 0, True, W)
    sc = O.transh_forward(ent, rel, nrm, torch.from_numpy(bh), torch.from_numpy(br), torch.from_numpy(bt), 1, True)[4]
    want = O.margin_loss(sc[:B].view(-1, B).t(), sc[B:].view(-1, B).t(), margin=4.0)
    first = next(iter(loader))
    got = strat(first)
    assert abs(float(got) - float(want)) <= 1e-4 * max(1.0, abs(float(want))), (float(got), float(want))
    hist = Trainer(strat, loader, train_times=30, alpha=0.5, opt_method="sgd").run()
    assert hist[-1] < 0.7 * hist[0], hist[::6]


@pytest.mark.gpu
@pytest.mark.parametrize("nb", [1, 4])
def test_kg_trainer_graph_replay_equals_eager_loop(nb):
    """Trainer(graph=True) replays whole epochs (device sampler + forward + loss + backward + SGD) as one CUDA graph:
    same batches (the seed ping-pong returns to its start inside a graph), same losses, same final embeddings and
    sampler seeds as the eager loop -- odd and even batches per epoch, with an eager remainder."""
    from skghoi_b200.kg_data import KGIndex
    from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
    from skghoi_b200.kge import MarginLoss, TransH
    h, t, r = MG.graph()
    idx = KGIndex(h, t, r, 80, 117, device="cuda:0")

    def run(graph):
        loader = TrainDataLoader(idx, nbatches=nb, neg_ent=5, bern_flag=True, threads=4, seeds=MG.seeds_for(4))
        torch.manual_seed(0)
        model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to("cuda:0")
        strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to("cuda:0")
        hist = Trainer(strat, loader, train_times=11, alpha=0.5, opt_method="sgd", graph=graph).run()
        return hist, model.ent_embeddings.weight.detach().cpu().clone(), loader.sampler.seeds()

    eager, graphed = run(False), run(True)
    assert len(graphed[0]) == len(eager[0]) == 11
    assert graphed[2] == eager[2]                         # identical sampler state: the very same batches were drawn
    # one step per epoch: graph and eager agree to the last bits.  Four steps per epoch (44 updates): the TransH backward
    # accumulates with fp32 atomics whose order varies, the L1 score's sign gradients amplify that, and two EAGER runs
    # already differ by ~1e-2 in the losses and ~2e-2 in the embeddings (tools/kg_graph_vs_eager.py) -- same spread here
    if nb == 1:
        torch.testing.assert_close(torch.tensor(graphed[0]), torch.tensor(eager[0]), rtol=0, atol=1e-4)
        torch.testing.assert_close(graphed[1], eager[1], rtol=0, atol=1e-4)
    else:
        torch.testing.assert_close(torch.tensor(graphed[0]), torch.tensor(eager[0]), rtol=5e-2, atol=1e-2)
        torch.testing.assert_close(graphed[1], eager[1], rtol=0, atol=1e-1)


@pytest.mark.gpu
@pytest.mark.parametrize("nb,bern", [(1, True), (4, False)])
def test_kg_trainer_fused_kernel_equals_eager_loop(nb, bern):
    """Trainer(fused=True): the whole run as ONE persistent cooperative kernel (skg_kg_train_steps: sampler, TransH
    forward, margin loss, backward, SGD per step, grid-wide barriers between the phases).  Same batches as the eager loop
    (identical sampler state afterwards), same losses and embeddings up to the fp32 order of the gradient atomics (two
    EAGER runs already differ by ~1e-2 over 44 updates, see the graph-replay test above)."""
    from skghoi_b200.kg_data import KGIndex
    from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
    from skghoi_b200.kge import MarginLoss, TransH
    h, t, r = MG.graph()
    idx = KGIndex(h, t, r, 80, 117, device="cuda:0")

    def run(fused):
        loader = TrainDataLoader(idx, nbatches=nb, neg_ent=5, bern_flag=bern, threads=4, seeds=MG.seeds_for(4))
        torch.manual_seed(0)
        model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to("cuda:0")
        strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to("cuda:0")
        tr = Trainer(strat, loader, train_times=11, alpha=0.5, opt_method="sgd", fused=fused)
        hist = tr.run()
        hist2 = tr.run() if fused else None                  # a second call continues tables and sampler streams
        return hist, model.ent_embeddings.weight.detach().cpu().clone(), model.norm_vector.weight.detach().cpu().clone(), \
            loader.sampler.seeds(), hist2

    fused = run(True)
    eager = run(False)
    assert len(fused[0]) == len(eager[0]) == 11
    assert fused[4] is not None and len(fused[4]) == 11 and fused[4][-1] < fused[0][0]        # keeps training
    if nb == 1:
        torch.testing.assert_close(torch.tensor(fused[0]), torch.tensor(eager[0]), rtol=0, atol=2e-4)
    else:
        torch.testing.assert_close(torch.tensor(fused[0]), torch.tensor(eager[0]), rtol=5e-2, atol=1e-2)
    assert fused[0][-1] < 0.9 * fused[0][0]


@pytest.mark.gpu
def test_kg_trainer_fused_state_after_one_run():
    """After ONE fused run the tables and the sampler seeds equal the eager loop's (one step per epoch: 1e-4)."""
    from skghoi_b200.kg_data import KGIndex
    from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
    from skghoi_b200.kge import MarginLoss, TransH
    h, t, r = MG.graph()
    idx = KGIndex(h, t, r, 80, 117, device="cuda:0")
    out = {}
    for fused in (False, True):
        loader = TrainDataLoader(idx, nbatches=1, neg_ent=25, bern_flag=True, threads=8, seeds=MG.seeds_for(8))
        torch.manual_seed(0)
        model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to("cuda:0")
        strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to("cuda:0")
        hist = Trainer(strat, loader, train_times=7, alpha=0.5, opt_method="sgd", fused=fused).run()
        out[fused] = (hist, [w.detach().cpu().clone() for w in (model.ent_embeddings.weight, model.rel_embeddings.weight,
                                                                 model.norm_vector.weight)], loader.sampler.seeds())
    assert out[True][2] == out[False][2]
    torch.testing.assert_close(torch.tensor(out[True][0]), torch.tensor(out[False][0]), rtol=0, atol=2e-4)
    for a, b in zip(out[True][1], out[False][1]):
        torch.testing.assert_close(a, b, rtol=0, atol=2e-4)
