"""Timeline of ONE data-parallel training step (torch.profiler / kineto, the box has no nsys): which NCCL all-reduce kernels
run under which backward kernels.  Run under torchrun on >= 2 GPUs; rank 0 writes a text summary to stdout.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/dist_timeline.py"""
import json, os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from skghoi_b200 import engine, shard, synthetic, train
from skghoi_b200.head import build_interaction_head

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
otv = synthetic.hico_object_to_verb()
ih = build_interaction_head(otv, seed=42, distributed=True)
ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
ih = ih.to(dev).train(); ih.box_pair_head.transh_head.transh.to(dev)
ih.box_pair_head.transh_head.negatives_on_device = True
batch = synthetic.make_batch([(16, 32)] * 16, seed=1042 + rank, with_targets=True)
n_per = [int(b.shape[0]) for b in batch["box_coords"]]
d = dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev), boxes=list(torch.cat(batch["box_coords"]).to(dev).split(n_per)),
         labels=list(torch.cat(batch["box_labels"]).to(dev).split(n_per)), scores=list(torch.cat(batch["box_scores"]).to(dev).split(n_per)),
         targets=[{k: v.to(dev) for k, v in t.items()} for t in batch["targets"]])
meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
params = [p for p in ih.parameters() if p.requires_grad]
buckets = shard.GradBuckets(params).attach(flatten_params=True, module=ih)
opt = torch.optim.SGD(buckets.flat_params, lr=1e-4, momentum=0.9, fused=True)
train.enable_gradient_allreduce(ih)


def step():
    buckets.zero_grad()
    res = ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"], d["targets"], meta=meta)
    ls = res[-1]
    (ls["hoi_loss"] + ls["interactiveness_loss"] + ls["transH_loss"].sum()).backward()
    opt.step()


for _ in range(6):
    step()
torch.cuda.synchronize(); dist.barrier()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
if rank == 0:
    path = os.path.join(tempfile.mkdtemp(), "trace.json")
    prof.export_chrome_trace(path)
    ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
    ev.sort(key=lambda e: e["ts"])
    t0 = ev[0]["ts"]
    nccl = [e for e in ev if "nccl" in e["name"].lower()]
    comp = [e for e in ev if "nccl" not in e["name"].lower()]
    # the second profiled step: from the first kernel after the 4th-from-last NCCL group ... keep it simple: report every NCCL kernel
    print("world %d, two training steps, GPU activity %.3f ms, %d kernels (%d NCCL)" % (world, (ev[-1]["ts"] + ev[-1]["dur"] - t0) / 1e3, len(ev), len(nccl)))
    print("| NCCL kernel | start ms | dur us | overlapped by compute kernels us | share under compute | compute kernels running meanwhile |")
    print("|---|---:|---:|---:|---:|---|")
    for k in nccl:
        a, b = k["ts"], k["ts"] + k["dur"]
        ov, names = 0.0, {}
        for c in comp:
            lo, hi = max(a, c["ts"]), min(b, c["ts"] + c["dur"])
            if hi > lo:
                ov += hi - lo
                nm = c["name"].split("(")[0][-40:]
                names[nm] = names.get(nm, 0) + 1
        top = ", ".join("%s x%d" % kv for kv in sorted(names.items(), key=lambda kv: -kv[1])[:4])
        print("| %s | %.3f | %.1f | %.1f | %.0f%% | %s |" % (k["name"][:48], (a - t0) / 1e3, k["dur"], min(ov, k["dur"]), 100 * min(ov, k["dur"]) / max(k["dur"], 1e-9), top))
    # step boundaries: fused SGD kernels
    sgd = [e for e in ev if "multi_tensor_apply" in e["name"]]
    if len(sgd) >= 2:
        ends = sorted(e["ts"] + e["dur"] for e in sgd)
        print("optimizer kernels end at (ms):", ", ".join("%.3f" % ((t - t0) / 1e3) for t in ends))
dist.destroy_process_group()
