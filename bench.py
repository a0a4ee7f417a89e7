#!/usr/bin/env python
"""Benchmark of the interaction-head hot path (BASELINE.json metric: HOI-head human-object pairs/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One "step" = one pass of the interaction head (eval) over one batch of synthetic HICO-DET-shaped
detections: BASELINE.json configs[1] = 8 images x (16 humans, 32 boxes) -> 4096 grid cells, 3968
valid pairs, 117 verbs, from the pooled box features to the post-processed result dicts
(reference head.py:389-427).  For N > 1 (torchrun, one rank per GPU) every rank processes its own
batch -- images are independent, there is no data-path collective ("weak" scaling).

JSON keys beyond the base contract:
  value      pairs/s with the step's inputs already resident in HBM (CUDA events, per-step, L2
             flushed between steps, max over ranks);
  e2e        the same metric through the public module API with pinned HOST inputs: H2D of the
             step's inputs and D2H of its results are inside the timed region;
  roofline   the dominant kernel (the fused fc_2 GEMM of the four MBFs, M = cells, N = 4096, K = 1024)
             timed alone with CUDA events; spatial / transh hold the two bandwidth kernels;
  cpu_baseline  the CPU restatement of the reference head (oracle/head_oracle.py, kind "port")
             timed on this box's host cores on a bounded sample (rank 0, N = 1 only).
  training   supplementary: one interaction-head training step (BASELINE configs[2]), data-parallel when N > 1;
  kg_training  supplementary: TransH KG training on the HOI label graph with the device negative sampler (configs[3]).
`--impl reference` times that CPU head alone (the reference is eager PyTorch; its own checkout is
not on the GPU box, so the pinned op-for-op port stands in).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "hoi_head_pairs_per_sec"
UNIT = "pairs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--images", type=int, default=8, help="images per step per GPU (configs[1]: 8)")
    ap.add_argument("--humans", type=int, default=16)
    ap.add_argument("--boxes", type=int, default=32)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-rooflines", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the supplementary training-step line (configs[2])")
    ap.add_argument("--no-sweep", action="store_true", help="skip the C5 pair-count stress sweep (BASELINE configs[4])")
    ap.add_argument("--no-varying", action="store_true", help="skip the varying-geometry leg (fresh batch geometry every step)")
    ap.add_argument("--train-images", type=int, default=16, help="images per training step per GPU (configs[2]: 16)")
    ap.add_argument("--train-steps", type=int, default=40)
    ap.add_argument("--profile-e2e", action="store_true", help="cProfile of the (untimed) e2e loop to stderr, then exit")
    return ap.parse_args()


def workload_config(args, extra=None):
    cfg = {"workload": "HICO-DET-shaped inference, batch %d images x (%d humans, %d boxes), 117 verbs "
                       "(BASELINE configs[1])" % (args.images, args.humans, args.boxes),
           "images_per_step_per_gpu": args.images, "n_h": args.humans, "n": args.boxes,
           "cells_per_step_per_gpu": args.images * args.humans * args.boxes,
           "pairs_per_step_per_gpu": args.images * args.humans * (args.boxes - 1),
           "boundary": "InteractionHead.forward_from_pooled (head.py:389-427; RoIAlign/NMS stay stock torchvision)",
           "repeat": "value / e2e replay ONE batch (same geometry, same buffers: every host cache and the CUDA graph hit); "
                     "varying_geometry times a stream of never-seen batches",
           "parallelism": "dp%d (images sharded, no data-path collective)" % args.gpus}
    if extra:
        cfg.update(extra)
    return cfg


# ------------------------------------------------------------------------------------------------
# CPU reference arm / cpu_baseline  (the only place outside tests/ and smoke() that executes oracle/)
# ------------------------------------------------------------------------------------------------
def _oracle_inputs(args, images, device=None):
    import torch
    from skghoi_b200 import synthetic
    from skghoi_b200.head import build_interaction_head
    otv = synthetic.hico_object_to_verb()
    ih = build_interaction_head(otv, seed=42)
    P = {k: v.detach().clone() for k, v in ih.state_dict().items()}
    tables = synthetic.make_transh_tables(42)
    batch = synthetic.make_batch([(args.humans, args.boxes)] * images, seed=42)
    if device is not None:
        P = {k: v.to(device) for k, v in P.items()}
        tables = tuple(t.to(device) for t in tables)
        batch = dict(batch, features={k: v.to(device) for k, v in batch["features"].items()},
                     box_features=batch["box_features"].to(device), box_coords=[b.to(device) for b in batch["box_coords"]],
                     box_labels=[l.to(device) for l in batch["box_labels"]], box_scores=[x.to(device) for x in batch["box_scores"]])
    return P, tables, batch, otv


def cpu_head_throughput(args, steps, warmup, images=1, threads=None, budget_s=None):
    """Times the reference head's CPU restatement (reference-faithful per-row entity loop, head.py:877-883) on the host
    cores; one step = ``images`` images of the workload.  ``budget_s`` caps the timed steps so that the run ends in time
    (the first timed step sizes the rest).  Returns (pairs/s, threads, s/step, timed steps)."""
    import torch
    from oracle import head_oracle as O
    cores = len(os.sched_getaffinity(0))
    threads = cores if threads is None else threads
    torch.set_num_threads(threads)
    P, tables, batch, otv = _oracle_inputs(args, images)
    pairs = images * args.humans * (args.boxes - 1)

    def step():
        with torch.no_grad():
            O.interaction_head_forward(P, tables, batch["features"]["3"], batch["image_shapes"],
                                       batch["box_features"], batch["box_coords"], batch["box_labels"],
                                       batch["box_scores"], otv, row_pick="loop")
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    done = 0
    while done < steps:
        step(); done += 1
        if budget_s is not None and done < steps:
            per = (time.perf_counter() - t0) / done
            if per * (done + 1) > budget_s:
                break
    dt = time.perf_counter() - t0
    torch.set_num_threads(cores)
    return pairs * done / dt, threads, dt / done, done


def run_reference(args):
    """`--impl reference`: the reference head's own CPU implementation (pinned op-for-op port; the reference checkout is
    not on the GPU box and needs mmcv / mmdet / pocket) on the SAME batch as the b200 arm -- all `--images` images per step,
    all host threads, the requested warm-up and step counts (steps are cut only if the run would pass ~4 minutes)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, args.steps); warmup = max(0, args.warmup)
    v, cores, sps, done = cpu_head_throughput(args, steps, warmup, images=args.images, budget_s=200.0)
    sample = "%d images (%d humans x %d boxes, %d pairs) per step, %d timed steps after %d warm-up, fp32 torch CPU, %d threads" % (
        args.images, args.humans, args.boxes, args.images * args.humans * (args.boxes - 1), done, warmup, cores)
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": done,
            "warmup": warmup, "ms_per_step": sps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, {"sample": sample}),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def gpu_eager_baseline(args, dev):
    """SURVEY 8(d) / BASELINE.md 2(i): the reference head as stock eager PyTorch on the SAME B200 -- the op-for-op port
    (oracle/head_oracle.py, 16 branch Linears per MBF, 117*N2 materialised TransH triples, per-image Python loop with
    its device->host syncs) with its tensors on cuda:0.  fp32 (TF32 off, the PyTorch default for matmul) and TF32 on;
    the per-row entity pick as the reference's Python loop (head.py:877-883) and as the equivalent strided slice."""
    import torch
    from oracle import head_oracle as O
    P, tables, batch, otv = _oracle_inputs(args, args.images, device=dev)
    pairs = args.images * args.humans * (args.boxes - 1)
    out = {"workload": "same batch as `value` (%d images x (%d, %d)), reference head as eager PyTorch on cuda:0 (oracle port), "
                       "device-resident inputs, wall clock with synchronize" % (args.images, args.humans, args.boxes),
           "unit": UNIT}
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    try:
        for name, tf32, pick, steps in (("fp32_slice", False, "slice", 5), ("tf32_slice", True, "slice", 5),
                                        ("fp32_loop", False, "loop", 2)):
            torch.backends.cuda.matmul.allow_tf32 = tf32; torch.backends.cudnn.allow_tf32 = tf32

            def step():
                with torch.no_grad():
                    O.interaction_head_forward(P, tables, batch["features"]["3"], batch["image_shapes"], batch["box_features"],
                                               batch["box_coords"], batch["box_labels"], batch["box_scores"], otv, row_pick=pick)
            step(); torch.cuda.synchronize(dev)
            t0 = time.perf_counter()
            for _ in range(steps):
                step()
            torch.cuda.synchronize(dev)
            dt = (time.perf_counter() - t0) / steps
            out[name] = {"value": pairs / dt, "ms_per_step": dt * 1e3, "steps": steps,
                         "row_pick": "python loop over 117*N2 rows (head.py:877-883)" if pick == "loop" else "[::117] slice (same values)"}
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
    return out


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index; self.proc = None

    def start(self):
        if self.index is None:
            return
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.index is None:
            return None
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill(); out, _ = self.proc.communicate()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in out.strip().splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from skghoi_b200 import _lib, engine, synthetic
    from skghoi_b200.head import build_interaction_head

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.lib()
    peaks = {}
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(pk_path):
        peaks = json.load(open(pk_path))
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0)); tf_peak = float(peaks.get("bf16_tflops", 1590.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if peaks else "fallback (B200_PROFILING.md)"

    otv = synthetic.hico_object_to_verb()
    ih = build_interaction_head(otv, seed=42)
    ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
    ih = ih.to(dev).eval()
    ih.box_pair_head.transh_head.transh.to(dev)
    shapes = [(args.humans, args.boxes)] * args.images
    batch = synthetic.make_batch(shapes, seed=42 + rank)
    pairs_step = args.images * args.humans * (args.boxes - 1)
    # host (pinned) and device copies of the step's inputs
    host = dict(f3=batch["features"]["3"].pin_memory(), bf=batch["box_features"].pin_memory(),
                boxes=[b.pin_memory() for b in batch["box_coords"]], labels=[l.pin_memory() for l in batch["box_labels"]],
                scores=[s.pin_memory() for s in batch["box_scores"]])
    # the detection lists live in three packed device buffers; the module API gets per-image views of them
    n_per_d = [int(b.shape[0]) for b in batch["box_coords"]]
    d = dict(f3=host["f3"].to(dev), bf=host["bf"].to(dev),
             boxes=list(torch.cat(batch["box_coords"]).to(dev).split(n_per_d)),
             labels=list(torch.cat(batch["box_labels"]).to(dev).split(n_per_d)),
             scores=list(torch.cat(batch["box_scores"]).to(dev).split(n_per_d)))
    meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)   # host-side copy of the tiny label/score lists
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)          # > 126 MB L2

    def step_resident():
        with torch.no_grad():
            return engine.interaction_head_from_pooled(ih, {"3": d["f3"]}, batch["image_shapes"], d["bf"], d["boxes"],
                                                       d["labels"], d["scores"], meta=meta)

    h2d_bytes = (host["f3"].numel() + host["bf"].numel()) * 4 + sum(b.numel() * 4 for b in host["boxes"]) + \
        sum(l.numel() * 8 for l in host["labels"]) + sum(s.numel() * 4 for s in host["scores"])
    d2h_holder = {}
    copy_stream = torch.cuda.Stream(dev)

    # the per-image detection lists travel as ONE packed pinned buffer (boxes | labels | scores); the step's inputs land
    # in two alternating sets of persistent device buffers (a serving loop's double buffering: stable addresses, no
    # allocator traffic), and the module API gets per-image VIEWS of the device copies
    n_per = [int(b.shape[0]) for b in batch["box_coords"]]
    nb_tot = sum(n_per)
    small = torch.empty(nb_tot * 28, dtype=torch.uint8).pin_memory()
    small[:nb_tot * 16].view(torch.float32).copy_(torch.cat(batch["box_coords"]).reshape(-1))
    small[nb_tot * 16:nb_tot * 24].view(torch.int64).copy_(torch.cat(batch["box_labels"]))
    small[nb_tot * 24:].view(torch.float32).copy_(torch.cat(batch["box_scores"]))
    host["small"] = small
    host["labels_all"] = small[nb_tot * 16:nb_tot * 24].view(torch.int64)
    host["scores_all"] = small[nb_tot * 24:].view(torch.float32)

    def device_set():
        sm = torch.empty_like(small, device=dev)
        return dict(small=sm, f3=torch.empty_like(host["f3"], device=dev), bf=torch.empty_like(host["bf"], device=dev),
                    boxes=sm[:nb_tot * 16].view(torch.float32).view(nb_tot, 4).split(n_per),
                    labels=sm[nb_tot * 16:nb_tot * 24].view(torch.int64).split(n_per),
                    scores=sm[nb_tot * 24:].view(torch.float32).split(n_per), free=None)
    sets = [device_set(), device_set()]

    def e2e_h2d(i):
        """Step i's inputs, pinned host -> device set i % 2, on the copy stream (small buffer first)."""
        ds = sets[i % 2]
        with torch.cuda.stream(copy_stream):
            if ds["free"] is not None:
                copy_stream.wait_event(ds["free"])            # the step that last read this set has finished
            ds["small"].copy_(host["small"], non_blocking=True)
            ds["f3"].copy_(host["f3"], non_blocking=True)
            ds["bf"].copy_(host["bf"], non_blocking=True)
            ev = torch.cuda.Event(); ev.record(copy_stream)
        return ds, ev

    def e2e_compute(ds, ev):
        """Public module API on the device copies, then the results on their way to pinned host memory."""
        cur = torch.cuda.current_stream(dev)
        cur.wait_event(ev)
        with torch.no_grad():
            m = ih.host_meta(host["labels_all"], host["scores_all"], n=n_per)   # from the HOST copies: no device round trip
            res = ih.forward_from_pooled({"3": ds["f3"]}, batch["image_shapes"], ds["bf"], ds["boxes"], ds["labels"],
                                         ds["scores"], meta=m)
            pend = ih.results_to_host(res, wait=False)
        ds["free"] = torch.cuda.Event(); ds["free"].record(cur)
        return pend

    def run_e2e(steps):
        """Software pipeline over the K steps: the H2D copy of step i+1 is in flight while step i computes, and the
        host reads step i-1's results (waiting on THAT step's D2H event) after enqueueing step i.  Every step's H2D,
        forward, L2 flush and D2H + wait are inside the timed region."""
        nxt = e2e_h2d(0)
        prev = None
        for i in range(steps):
            cur = nxt
            if i + 1 < steps:
                nxt = e2e_h2d(i + 1)
            flush.zero_()
            pend = e2e_compute(*cur)
            if prev is not None:
                out = prev.get()
            prev = pend
        out = prev.get()
        d2h_holder["bytes"] = sum(t.numel() * t.element_size() for r in out for t in r.values())

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- warm-up
    for _ in range(max(args.warmup, 3)):
        step_resident(); flush.zero_()
    torch.cuda.synchronize(dev)
    # a step is ~0.3 ms: W of them end before clocks, allocator pools and the graph cache have settled, so keep stepping
    # (untimed) for a fixed 0.5 s on top of the W warm-up steps
    # the clock sampler starts here: nvidia-smi's start-up (NVML initialisation takes driver locks) then falls into the
    # spin-up instead of the timed region; its samples cover spin-up + timed steps, all under load
    sampler = ClockSampler(local if rank == 0 else None)
    sampler.start()
    t_spin = time.perf_counter()
    while time.perf_counter() - t_spin < 0.5:
        for _ in range(20):
            step_resident(); flush.zero_()
        torch.cuda.synchronize(dev)
    # ---- timed: device-resident inputs
    # clocks / throttle reasons are sampled by rank 0 only (its own GPU): eight concurrent nvidia-smi pollers contend for
    # the driver with the ranks' kernel launches and slow the host-bound step down
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    l0 = lib.skg_launch_count()
    barrier()
    for s, e in ev:
        s.record(); step_resident(); e.record()
        flush.zero_()
    barrier()
    launches = lib.skg_launch_count() - l0
    t_ms = sum(s.elapsed_time(e) for s, e in ev)
    # ---- timed: end to end from pinned host buffers (pipelined loop, host wall clock over all K steps)
    run_e2e(max(args.warmup, 3) + 20)
    barrier()
    if args.profile_e2e:
        import cProfile, pstats
        pr = cProfile.Profile(); pr.enable(); run_e2e(200); torch.cuda.synchronize(dev); pr.disable()
        pstats.Stats(pr, stream=sys.stderr).sort_stats("tottime").print_stats(40)
        print("graph stats (replays, captures, plain):", _lib.graph_stats(), file=sys.stderr)
        sampler.stop()
        return
    if os.environ.get("SKG_E2E_REPEAT"):
        for _ in range(int(os.environ["SKG_E2E_REPEAT"])):
            t0 = time.perf_counter(); run_e2e(args.steps); torch.cuda.synchronize(dev)
            print("e2e repeat: %.3f ms/step" % ((time.perf_counter() - t0) / args.steps * 1e3), file=sys.stderr)
    t0 = time.perf_counter()
    run_e2e(args.steps)
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    barrier()
    # the H2D of one step ALONE (the e2e floor): this step's own pinned buffers into one of the device sets, nothing else
    # running, best of five event-timed bursts of five steps' worth of copies
    h2d_gbs = 0.0
    for _ in range(5):
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(5):
            sets[0]["small"].copy_(host["small"], non_blocking=True)
            sets[0]["f3"].copy_(host["f3"], non_blocking=True)
            sets[0]["bf"].copy_(host["bf"], non_blocking=True)
        c1.record(); torch.cuda.synchronize(dev)
        h2d_gbs = max(h2d_gbs, 5 * h2d_bytes / (c0.elapsed_time(c1) * 1e-3) / 1e9)
    clocks = sampler.stop()
    tt = torch.tensor([t_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_ms, e2e_ms = float(tt[0]), float(tt[1])
    value = pairs_step * world * args.steps / (t_ms * 1e-3)
    e2e_v = pairs_step * world * args.steps / (e2e_ms * 1e-3)

    roof = None; extra_roofs = {}
    if rank == 0 and not args.no_rooflines:
        roof, extra_roofs = measure_rooflines(torch, _lib, engine, ih, dev, flush, hbm_peak, tf_peak, peak_src)
    varying = None
    if not args.no_varying:
        varying = measure_varying_geometry(torch, dist, args, ih, dev, rank, world)
    sweep = None
    if not args.no_sweep:
        sweep = measure_c5_sweep(torch, dist, args, ih, dev, rank, world, flush)
    train = None
    if not args.no_train:
        train = measure_training(torch, dist, args, dev, rank, world, flush)
    kg = None
    if not args.no_train:
        kg = measure_kg(torch, dist, dev, rank, world)
    cpu = None
    gpu_eager = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        gpu_eager = gpu_eager_baseline(args, dev)
        v, cores, sps, done = cpu_head_throughput(args, steps=10, warmup=1, images=args.images, budget_s=15.0)
        v1, _, sps1, done1 = cpu_head_throughput(args, steps=3, warmup=1, images=1, threads=1, budget_s=8.0)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "oracle/head_oracle.py (op-for-op fp32 restatement of the reference head), %d steps of the whole "
                         "%d-image batch (%d pairs) each, %.3f s/step" % (done, args.images, args.images * args.humans * (args.boxes - 1), sps),
               "single_thread": {"value": v1, "cores": 1, "sample": "%d steps of 1 image, %.3f s/step" % (done1, sps1)}}
    if rank == 0 and roof is not None:
        tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")
        if os.path.isfile(tp):
            tr_ = json.load(open(tp))
            roof["traffic"] = tr_.get("roofline"); roof["traffic_source"] = tr_.get("source")
            for k in extra_roofs:
                if k in tr_ and isinstance(extra_roofs[k], dict):
                    extra_roofs[k]["traffic"] = tr_[k]
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": t_ms / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16 (fp32 accumulate; fp32 spatial/TransH)",
                "data": "synthetic", "impl": "b200",
                "config": workload_config(args, {"l2": "256 MiB flush buffer written between timed steps",
                                                 "timing": "CUDA events per step on the launch stream, summed, max over ranks"}),
                "e2e": {"value": e2e_v, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes,
                        "d2h_bytes_per_step": d2h_holder.get("bytes", 0), "ms_per_step": e2e_ms / args.steps,
                        "h2d_copy_gbs_measured": h2d_gbs, "h2d_floor_ms_per_step": h2d_bytes / h2d_gbs / 1e6,
                        "timing": "host wall clock over all K steps; per step: H2D of its inputs from pinned memory (copy stream, one step ahead), forward through InteractionHead.forward_from_pooled, 256 MiB L2 flush write, D2H of its results to pinned memory, read on the host one step later (event wait)"},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu}
        line.update(extra_roofs)
        if varying is not None:
            line["varying_geometry"] = varying
        if sweep is not None:
            line["c5_sweep"] = sweep
        if train is not None:
            line["training"] = train
        if kg is not None:
            line["kg_training"] = kg
        if gpu_eager is not None:
            line["gpu_eager_baseline"] = gpu_eager
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def measure_varying_geometry(torch, dist, args, ih, dev, rank, world):
    """VERDICT r1 item 3 / ADVICE: the headline replays ONE batch, real detection streams change (n_h, n) per image every
    batch (head.py:136-139, 826-841).  Here every step is a batch the process has never seen: per image n_h ~ U[1, 15],
    n ~ U[n_h + 1, 30] (what InteractionHead.preprocess emits with max_human = max_object = 15), its own detections and
    its own input buffers (distinct device addresses).  `stream` = one pass over `fresh` never-seen batches after a
    warm-up on OTHER batches (capacity-bucket graphs, SKG_HEAD_CAPACITY); `cycle_capacity` / `cycle_exact` = a 16-batch
    cycle replayed with bucket graphs / with one exact-geometry graph per batch (the old best case: no padding)."""
    from skghoi_b200 import _lib, engine, synthetic
    n_img = args.images
    warm_n, fresh_n = 128, 256      # the warm-up meets (nearly) every capacity bucket twice; first sightings inside the timed stream stay in it
    g = torch.Generator().manual_seed(9000 + rank)

    def draw():
        shapes = []
        for _ in range(n_img):
            nh = int(torch.randint(1, 16, (1,), generator=g))
            shapes.append((nh, int(torch.randint(nh + 1, 31, (1,), generator=g))))
        return shapes
    img_shapes = [(800, 1216)] * n_img
    gd = torch.Generator(device=dev).manual_seed(77 + rank)
    batches = []
    for i in range(warm_n + fresh_n):
        shapes = draw()
        boxes, labels, scores = synthetic.make_detections(shapes, img_shapes, g)
        tot = sum(n for _, n in shapes)
        n_per = [n for _, n in shapes]
        # detections as per-image views of three packed device tensors (what preprocess_detections_device hands out)
        batches.append(dict(shapes=shapes, pairs=sum(h * (n - 1) for h, n in shapes), cells=sum(h * n for h, n in shapes),
                            f3={"3": torch.randn(n_img, 256, 25, 38, device=dev, generator=gd)},
                            bf=torch.randn(tot, 256, 7, 7, device=dev, generator=gd),
                            boxes=list(torch.cat(boxes).to(dev).split(n_per)), labels=list(torch.cat(labels).to(dev).split(n_per)),
                            scores=list(torch.cat(scores).to(dev).split(n_per)),
                            meta=engine._host_meta(labels, scores, 49)))

    def step(bt):
        bt["meta"].pop("_skg_fwd", None); bt["meta"].pop("_skg_post", None)     # no per-batch host cache carried over
        with torch.no_grad():
            return engine.interaction_head_from_pooled(ih, bt["f3"], img_shapes, bt["bf"], bt["boxes"], bt["labels"], bt["scores"],
                                                       meta=bt["meta"])

    def timed(seq, reps=1):
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        g0 = _lib.graph_stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(reps):
            for bt in seq:
                step(bt)
        e1.record()
        host_ms = (time.perf_counter() - t0) * 1e3
        torch.cuda.synchronize(dev)
        g1 = _lib.graph_stats()
        ms = e0.elapsed_time(e1)
        from skghoi_b200 import shard
        ms = shard.max_over_ranks([ms], device=dev)[0] if world > 1 else ms
        steps = reps * len(seq)
        pairs = reps * sum(bt["pairs"] for bt in seq)
        return {"value": pairs * world / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms / steps, "steps": steps,
                "pairs_per_step_mean": pairs / steps, "host_enqueue_ms_per_step": host_ms / steps,
                "graph": {"replays": g1[0] - g0[0], "captures": g1[1] - g0[1], "plain": g1[2] - g0[2]}}
    engine.CAPACITY_BUCKETS = True
    engine._geometry_cache.clear()
    for _ in range(2):
        for bt in batches[:warm_n]:
            step(bt)
    stream = timed(batches[warm_n:])
    cyc = batches[warm_n:warm_n + 16]
    timed(cyc, 2)
    cycle_cap = timed(cyc, 8)
    engine.CAPACITY_BUCKETS = False
    engine._geometry_cache.clear()
    timed(cyc, 3)
    cycle_exact = timed(cyc, 8)
    engine.CAPACITY_BUCKETS = True
    engine._geometry_cache.clear()
    cells = [bt["cells"] for bt in batches[warm_n:]]
    out = {"workload": "%d images/step, per image n_h ~ U[1,15], n ~ U[n_h+1,30]; every timed step is a batch geometry, detections "
                       "and input buffers (%d distinct sets, %.1f GB: larger than L2) the process has never seen"
                       % (n_img, fresh_n, sum(bt["bf"].numel() * 4 + bt["f3"]["3"].numel() * 4 for bt in batches[warm_n:]) / 1e9),
           "cells_per_step": {"min": min(cells), "mean": sum(cells) / len(cells), "max": max(cells)},
           "stream": stream, "cycle_capacity": cycle_cap, "cycle_exact": cycle_exact,
           "stream_vs_exact": stream["ms_per_step"] and (cycle_exact["value"] and stream["value"] / cycle_exact["value"]),
           "note": "stream / cycle_capacity: graphs keyed on a capacity bucket (cells to a 128-row tile, node rows to 128; true "
                   "counts and input addresses read from the segment table on the device); cycle_exact: one graph per exact "
                   "geometry on a 16-batch cycle (no padding, the headline's best case at this batch size)"}
    return out if rank == 0 else None


def measure_c5_sweep(torch, dist, args, ih, dev, rank, world, flush):
    """BASELINE configs[4]: pair-count stress sweep up to 64 x 64 detections per image (4096 cells), every rank its own
    batch (weak scaling, no data-path collective), device-resident inputs, CUDA events per step, L2 flushed between steps,
    max over ranks.  Also (16, 32) x 512 images: the same 262 144 cells as 64 x (64, 64) spread over 8 x more images
    (VERDICT r1 'batch-512 regression': 4 x more boxes through box_head plus per-image host work)."""
    from skghoi_b200 import engine, shard, synthetic
    gd = torch.Generator(device=dev).manual_seed(4242 + rank)
    g = torch.Generator().manual_seed(777 + rank)
    rows = []
    for nh, n, B in ((32, 64, 8), (32, 64, 64), (40, 80, 8), (40, 80, 64), (64, 64, 8), (64, 64, 64), (16, 32, 512)):
        shapes = [(nh, n)] * B
        img_shapes = [(800, 1216)] * B
        boxes, labels, scores = synthetic.make_detections(shapes, img_shapes, g)
        n_per = [n] * B
        d = dict(f3={"3": torch.randn(B, 256, 25, 38, device=dev, generator=gd)},
                 bf=torch.randn(B * n, 256, 7, 7, device=dev, generator=gd),
                 boxes=list(torch.cat(boxes).to(dev).split(n_per)), labels=list(torch.cat(labels).to(dev).split(n_per)),
                 scores=list(torch.cat(scores).to(dev).split(n_per)))
        meta = engine._host_meta(labels, scores, 49)

        def step():
            with torch.no_grad():
                return engine.interaction_head_from_pooled(ih, d["f3"], img_shapes, d["bf"], d["boxes"], d["labels"], d["scores"], meta=meta)
        for _ in range(3):
            step(); flush.zero_()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        K = 5
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        t0 = time.perf_counter()
        for s_, e in ev:
            s_.record(); step(); e.record(); flush.zero_()
        host_ms = (time.perf_counter() - t0) / K * 1e3
        torch.cuda.synchronize(dev)
        ms = sum(s_.elapsed_time(e) for s_, e in ev) / K
        ms = shard.max_over_ranks([ms], device=dev)[0] if world > 1 else ms
        cells, pairs = B * nh * n, B * nh * (n - 1)
        flops = 2.0 * (cells * (128 * 256 + 256 * 1024 + 1024 * 4096 + 2 * 1024 * 1024) + pairs * (2 * 1024 * 1024 + 6144 * 128)
                       + B * n * (12544 * 1024 + 1024 * 1024 + 2 * 1088 * 1024 + 4 * 1024 * 1024))
        rows.append({"n_h": nh, "n": n, "images_per_gpu": B, "cells_per_gpu": cells, "pairs_per_gpu": pairs, "ms_per_step": ms,
                     "value": pairs * world / (ms * 1e-3), "unit": UNIT, "gemm_tflops_issued_per_gpu": flops / (ms * 1e-3) / 1e12,
                     "host_loop_ms_per_step": host_ms})
        del d
        torch.cuda.empty_cache()
    return {"workload": "pair-count stress sweep (BASELINE configs[4]), per GPU: (n_h, n) x images per step; weak scaling over %d GPU(s)" % world,
            "rows": rows} if rank == 0 else None


def measure_kg(torch, dist, dev, rank, world):
    """BASELINE configs[3]: TransH embedding training on the HOI label graph (80 objects, 117 verbs) with negative sampling
    -- the device sampler (skg_kg_sample, bit-exact with OpenKE's C++ sampler) inside Trainer.run, OpenKE defaults
    (nbatches = 1, 25 negatives per positive, bern, dim 50, margin 4, SGD) -- and the sampler alone on a large batch.
    N > 1: the triple batch is sharded over the ranks (same global batch on every rank, each scores its slice, ONE
    62.8 KB all-reduce of the table gradients per step): "strong" scaling of a fixed global batch."""
    from skghoi_b200 import shard, synthetic
    from skghoi_b200.kg_data import KGIndex, TrainDataSampler
    from skghoi_b200.kg_train import NegativeSamplingStrategy, TrainDataLoader, Trainer
    from skghoi_b200.kge import MarginLoss, TransH
    otv = synthetic.hico_object_to_verb()
    h, t, r = [], [], []
    for o in range(80):
        for v in otv[o]:
            h.append(49); r.append(int(v)); t.append(o)
    idx = KGIndex(h, t, r, 80, 117, device=dev)
    seeds = [12345 + 977 * i for i in range(8)]
    loader = TrainDataLoader(idx, nbatches=1, neg_ent=25, bern_flag=True, threads=8, seeds=seeds)
    torch.manual_seed(7)                                                    # identical initial tables on every rank
    model = TransH(80, 117, dim=50, p_norm=1, norm_flag=True).to(dev)
    strat = NegativeSamplingStrategy(model, MarginLoss(margin=4.0), batch_size=loader.get_batch_size()).to(dev)
    steps = 50
    tr = Trainer(strat, loader, train_times=steps, alpha=0.5, distributed=world > 1)
    tr.run(); torch.cuda.synchronize(dev)                                   # warm-up epoch block
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); hist = tr.run(); e1.record(); torch.cuda.synchronize(dev)
    ms = shard.max_over_ranks([e0.elapsed_time(e1) / steps], device=dev)[0]
    trip = loader.get_batch_size() * 26
    # algorithmic bytes per triple of one step: sampler writes 3 int64 + 1 fp32 (28 B), the scorer reads the 3 indices (24 B)
    # and writes the score (4 B), its backward re-reads them (24 B) and scatters 4 table-row gradients of 50 fp32 (800 B)
    bytes_per_triple = 28 + 28 + 24 + 800
    out = {"workload": "TransH KG embedding training on the HOI label graph (80 objects, 117 verbs, %d triples), %d positives x 25 "
                       "negatives per step (BASELINE configs[3])" % (idx.train_total, loader.get_batch_size()),
           "value": trip / (ms * 1e-3), "unit": "triples/s", "ms_per_step": ms, "steps": steps, "n_gpus": world,
           "scaling": "strong (fixed global batch, triple batch sharded over ranks, tables replicated)" if world > 1 else "single GPU",
           "step": "device negative sampling (skg_kg_sample) + TransH forward + margin loss + backward + %sSGD"
                   % ("one all-reduce of the table gradients (62.8 KB) + " if world > 1 else ""),
           "loss_first_last": [float(hist[0]), float(hist[-1])],
           "roofline": {"bound": "hbm (nominal: the step is launch-latency bound, ~25 launches for < 20 MB of traffic)",
                        "bytes_per_triple": bytes_per_triple, "achieved": trip * bytes_per_triple / (ms * 1e-3) / 1e9, "unit": "GB/s"}}
    if world > 1:
        return out if rank == 0 else None
    # the same loop with whole epochs replayed as one CUDA graph (Trainer(graph=True)): capture cost included
    trg = Trainer(strat, loader, train_times=500, alpha=0.5, optimizer=tr.optimizer, graph=True, distributed=False)
    torch.cuda.synchronize(dev); t0 = time.perf_counter(); hist_g = trg.run(); torch.cuda.synchronize(dev)
    ms_g = (time.perf_counter() - t0) / 500 * 1e3
    bs = 1 << 16
    smp = TrainDataSampler(idx, bs, 25, True, 8, seeds)
    for _ in range(3):
        smp.sample()
    e0.record()
    for _ in range(20):
        smp.sample()
    e1.record(); torch.cuda.synchronize(dev)
    us = e0.elapsed_time(e1) / 20 * 1e3
    # ... and as ONE persistent cooperative kernel (Trainer(fused=True), csrc/kgtrain.cu): 500 steps per launch
    trf = Trainer(strat, loader, train_times=500, alpha=0.5, optimizer=tr.optimizer, fused=True, distributed=False)
    trf.run(); torch.cuda.synchronize(dev)
    e0.record(); hist_f = trf.run(); e1.record(); torch.cuda.synchronize(dev)
    ms_f = e0.elapsed_time(e1) / 500
    out["fused_kernel"] = {"ms_per_step": ms_f, "value": trip / (ms_f * 1e-3), "unit": "triples/s", "steps": 500,
                           "loss_last": float(hist_f[-1]),
                           "roofline": {"bound": "latency (4 grid-wide barriers per step; tables and gradients in shared memory)",
                                        "bytes_per_triple": bytes_per_triple, "achieved": trip * bytes_per_triple / (ms_f * 1e-3) / 1e9,
                                        "unit": "GB/s"},
                           "note": "skg_kg_train_steps: sampler + TransH forward + margin loss + backward + SGD for every step "
                                   "inside one launch; same batches as the eager loop, updates equal up to fp32 atomic order"}
    out["graph_replay"] = {"ms_per_step": ms_g, "value": trip / (ms_g * 1e-3), "steps": 500, "loss_last": float(hist_g[-1]),
                           "note": "Trainer(graph=True): two steps per CUDA graph, wall clock incl. warm-up steps and capture"}
    out["sampler_alone"] = {"batch": bs, "neg_rate": 25, "us_per_call": us, "triples_per_s": bs * 26 / (us * 1e-6)}
    return out


def measure_training(torch, dist, args, dev, rank, world, flush):
    """BASELINE configs[2]: interaction-head training step, batch 16 images / GPU, data-parallel.  One step = zero_grad,
    forward (train mode, targets), backward of the three losses, bucketed gradient all-reduce (NCCL when world > 1) and
    an SGD update.  Device-resident inputs, CUDA events per step, L2 flushed between steps, max over ranks."""
    from skghoi_b200 import shard, synthetic
    from skghoi_b200.head import build_interaction_head
    otv = synthetic.hico_object_to_verb()
    # (diagnostic knobs, tools/ only: SKG_BENCH_NO_NP_AR drops the rank average of the loss normalisers,
    #  SKG_BENCH_NO_GRAD_AR the gradient all-reduce -- to attribute the N > 1 step time)
    ih = build_interaction_head(otv, seed=42, distributed=world > 1 and not os.environ.get("SKG_BENCH_NO_NP_AR"))
    ih.box_pair_head.transh_head.set_tables(*synthetic.make_transh_tables(42))
    ih = ih.to(dev).train()
    ih.box_pair_head.transh_head.transh.to(dev)
    ih.box_pair_head.transh_head.negatives_on_device = True     # same distribution as head.py:939, device generator, batched
    n_img = args.train_images
    batch = synthetic.make_batch([(args.humans, args.boxes)] * n_img, seed=1042 + rank, with_targets=True)
    n_per = [int(b.shape[0]) for b in batch["box_coords"]]
    # detections as per-image views of three packed device tensors (what preprocess_detections_device hands out)
    d = dict(f={"3": batch["features"]["3"].to(dev)}, bf=batch["box_features"].to(dev),
             boxes=list(torch.cat(batch["box_coords"]).to(dev).split(n_per)), labels=list(torch.cat(batch["box_labels"]).to(dev).split(n_per)),
             scores=list(torch.cat(batch["box_scores"]).to(dev).split(n_per)),
             targets=[{k: v.to(dev) for k, v in t.items()} for t in batch["targets"]])
    from skghoi_b200 import engine
    meta = engine._host_meta(batch["box_labels"], batch["box_scores"], 49)
    params = [p for p in ih.parameters() if p.requires_grad]
    # gradients AND parameters live in flat buckets: stable .grad addresses for the native backward's unpack table, one
    # all-reduce + one scale per bucket when N > 1, one fused SGD kernel per bucket instead of ~400 per-tensor updates
    buckets = shard.GradBuckets(params).attach(flatten_params=True, module=ih)
    opt = torch.optim.SGD(buckets.flat_params, lr=1e-4, momentum=0.9, fused=True)

    if world > 1 and not os.environ.get("SKG_BENCH_NO_GRAD_AR"):
        # the packed gradient groups are all-reduced INSIDE loss.backward(), each behind its stage event of
        # skg_head_backward_staged, while the rest of the backward runs (train._GradAllReduce)
        from skghoi_b200 import train as _train
        _train.enable_gradient_allreduce(ih)

    def step():
        buckets.zero_grad()
        res = ih.forward_from_pooled(d["f"], batch["image_shapes"], d["bf"], d["boxes"], d["labels"], d["scores"],
                                     d["targets"], meta=meta)
        ls = res[-1]
        (ls["hoi_loss"] + ls["interactiveness_loss"] + ls["transH_loss"].sum()).backward()
        opt.step()
        return ls
    for _ in range(6):
        ls = step(); flush.zero_()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.train_steps)]
    for s_, e in ev:
        s_.record(); ls = step(); e.record(); flush.zero_()
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    t_ms = shard.max_over_ranks([sum(s_.elapsed_time(e) for s_, e in ev)], device=dev)[0]
    raw_steps = [s_.elapsed_time(e) for s_, e in ev]
    per_step = sorted(raw_steps)
    slow = [(i, round(t, 3)) for i, t in enumerate(raw_steps) if t > 1.15 * per_step[len(per_step) // 2]]
    pairs = n_img * args.humans * (args.boxes - 1)
    return {"workload": "interaction-head training step, batch %d images/GPU x (%d humans, %d boxes), data-parallel "
                        "(BASELINE configs[2])" % (n_img, args.humans, args.boxes),
            "value": pairs * world * args.train_steps / (t_ms * 1e-3), "unit": UNIT, "ms_per_step": t_ms / args.train_steps,
            "steps": args.train_steps, "n_gpus": world, "pairs_per_step_per_gpu": pairs,
            "ms_per_step_rank0_min_median_max": [per_step[0], per_step[len(per_step) // 2], per_step[-1]],
            "slow_steps_rank0": slow,
            "negatives": "TransH negatives drawn on the device (uniform without replacement per image, as head.py:939)",
            "step": "zero_grad + skg_head_forward(TRAIN) + post / losses + skg_head_backward (every dX / dW on the tcgen05 GEMM, trainops kernels) + %s + gradient unpack + fused SGD over flat parameter buffers"
                    % ("NCCL all-reduce of the 4 packed gradient groups, each started behind its stage event inside the backward "
                       "(the 81 MB node-side group overlaps the cell-side chain)" if world > 1 else "no all-reduce (1 GPU)"),
            "grad_bytes": sum(p.numel() * 4 for p in params),
            "final_losses": {k: float(v.detach().reshape(-1)[0]) for k, v in ls.items()}}


def _time_kernel(torch, fn, flush, reps=20):
    fn(); torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        tot += s.elapsed_time(e)
    return tot / reps * 1e-3


def _time_rotating(torch, fns, rounds=10):
    """Average launch duration over back-to-back launches that rotate through ``fns`` -- closures on DISJOINT operand /
    output sets whose combined footprint exceeds the 126 MB L2, so no launch finds its operands cached -- with ONE
    CUDA-event pair around the whole timed region (launch gaps included)."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(rounds):
        for f in fns:
            f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) * 1e-3 / (rounds * len(fns))


def measure_rooflines(torch, _lib, engine, ih, dev, flush, hbm_peak, tf_peak, peak_src):
    """Times the dominant kernels alone (CUDA events on the launch stream, L2 flushed before each launch)."""
    run = ih.__dict__["_skg_last_run"]
    g = engine.dominant_gemm_closure(ih, run)
    sec_flush = _time_kernel(torch, g["launch"], flush)
    # headline protocol: rotate private operand sets (> L2 in total), back-to-back launches, one event pair
    sets = [engine.dominant_gemm_closure(ih, run, private_copy=True) for _ in range(max(3, int(300e6 // g["footprint"]) + 1))]
    sec = _time_rotating(torch, [c["launch"] for c in sets])
    rot_note = "%d private operand/output sets of %.0f MB rotated (%.0f MB > 126 MB L2), back-to-back launches, one CUDA-event pair" % (
        len(sets), g["footprint"] / 1e6, len(sets) * g["footprint"] / 1e6)
    del sets
    roof = {"kernel": "gemm_bf16_kernel<256> (fused fc_2 of attention_head|obj_to_sub|sub_to_obj|attention_head_g, "
                      "M=%d N=%d K=%d, Hadamard+ReLU epilogue)" % (g["M"], g["N"], g["K"]),
            "bound": "tensor", "achieved": g["flops"] / sec / 1e12, "peak": tf_peak, "unit": "TFLOP/s",
            "frac": g["flops"] / sec / 1e12 / tf_peak, "traffic": None, "peak_source": peak_src + ", burst bf16",
            "us_per_launch": sec * 1e6, "flops_per_launch": g["flops"], "l2": rot_note,
            "us_per_launch_l2_flushed_single": sec_flush * 1e6,
            "frac_l2_flushed_single": g["flops"] / sec_flush / 1e12 / tf_peak}
    extra = {}
    # spatial encoding at 2^23 cells (SURVEY 8(d) asks >= 2^20): enc + x_keep/y_keep only; ONE launch writes 1.7 GB, 13 x the
    # 126 MB L2, so write absorption by the L2 cannot flatter the number
    from skghoi_b200 import synthetic
    B = 16384
    gen = torch.Generator().manual_seed(1)
    bx, _, _ = synthetic.make_detections([(16, 32)], [(800, 1216)], gen)
    boxes = bx[0].repeat(B, 1).to(dev)
    seg = engine.Segments([16] * B, [32] * B, [(800, 1216)] * B, dev)
    C, Pn = seg.total_cells, seg.total_pairs
    enc = torch.empty(C, 46, dtype=torch.float32, device=dev)
    xk = torch.empty(Pn, dtype=torch.int64, device=dev); yk = torch.empty(Pn, dtype=torch.int64, device=dev)
    nanf = torch.zeros(B, dtype=torch.int32, device=dev)
    p = _lib.ptr

    def sp_launch():
        _lib.call("skg_pairs_spatial_encode", p(boxes), p(seg.d_box_off), p(seg.d_n_h), p(seg.d_cell_off), p(seg.d_pair_off),
                  p(seg.d_hum_off), p(seg.d_hw), B, C, 23, 1e-10, 0, p(enc), None, None, p(xk), p(yk), None, None, None,
                  None, None, p(nanf), _lib.stream_ptr(dev))
    sec_flush = _time_kernel(torch, sp_launch, flush)
    enc_b = torch.empty_like(enc); xk_b = torch.empty_like(xk); yk_b = torch.empty_like(yk)

    def sp_launch_b():
        _lib.call("skg_pairs_spatial_encode", p(boxes), p(seg.d_box_off), p(seg.d_n_h), p(seg.d_cell_off), p(seg.d_pair_off),
                  p(seg.d_hum_off), p(seg.d_hw), B, C, 23, 1e-10, 0, p(enc_b), None, None, p(xk_b), p(yk_b), None, None, None,
                  None, None, p(nanf), _lib.stream_ptr(dev))
    sec = _time_rotating(torch, [sp_launch, sp_launch_b])
    del enc_b, xk_b, yk_b
    nbytes = 184 * C + 16 * Pn + 16 * seg.total_boxes
    extra["spatial_roofline"] = {"kernel": "pairs_spatial_kernel<23> (%d cells, enc + x_keep/y_keep)" % C, "bound": "hbm",
                                 "achieved": nbytes / sec / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                 "frac": nbytes / sec / 1e9 / hbm_peak, "traffic": None, "bytes_per_launch": nbytes,
                                 "us_per_launch": sec * 1e6, "peak_source": peak_src,
                                 "l2": "two output sets rotated (2 x %.0f MB written > 126 MB L2), back-to-back launches, one CUDA-event pair" % (nbytes / 1e6),
                                 "us_per_launch_l2_flushed_single": sec_flush * 1e6,
                                 "frac_l2_flushed_single": nbytes / sec_flush / 1e9 / hbm_peak}
    del enc, xk, yk
    # TransH: score-only grid mode (4 B / triple) and API-faithful mode (804 B written / triple)
    tr = ih.box_pair_head.transh_head.transh
    e, r, w = tr.ent_embeddings.weight.detach(), tr.rel_embeddings.weight.detach(), tr.norm_vector.weight.detach()
    cells = 1 << 21
    tails = (torch.arange(cells, device=dev) % 32).to(torch.int64)
    score = torch.empty(cells * 117, dtype=torch.float32, device=dev)

    def th_score():
        _lib.call("skg_transh_forward_grid", p(e), p(r), p(w), 80, 117, 50, 49, p(tails), cells, 2, 1, None, None, None, None,
                  p(score), _lib.stream_ptr(dev))
    sec_flush = _time_kernel(torch, th_score, flush, reps=10)
    sec = _time_rotating(torch, [th_score], rounds=10)     # one launch writes > 126 MB: nothing of it survives in L2
    nb = 4 * cells * 117 + 8 * cells
    extra["transh_score_roofline"] = {"kernel": "transh_grid_kernel<false> (grid mode, score only: [80,117] score table broadcast, %d triples)" % (cells * 117),
                                      "bound": "hbm", "achieved": nb / sec / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                      "frac": nb / sec / 1e9 / hbm_peak, "traffic": None, "bytes_per_launch": nb,
                                      "us_per_launch": sec * 1e6, "triples_per_s": cells * 117 / sec,
                                      "us_per_launch_l2_flushed_single": sec_flush * 1e6}
    del score
    cells2 = 16384
    T = cells2 * 117
    outs = [torch.empty(T, 50, dtype=torch.float32, device=dev) for _ in range(4)]
    score = torch.empty(T, dtype=torch.float32, device=dev)

    def th_full():
        _lib.call("skg_transh_forward_grid", p(e), p(r), p(w), 80, 117, 50, 49, p(tails), cells2, 2, 1, p(outs[0]), p(outs[1]),
                  p(outs[2]), p(outs[3]), p(score), _lib.stream_ptr(dev))
    sec_flush = _time_kernel(torch, th_full, flush, reps=10)
    sec = _time_rotating(torch, [th_full], rounds=10)     # one launch writes > 126 MB: nothing of it survives in L2
    nb = 804 * T + 8 * cells2
    extra["transh_rows_roofline"] = {"kernel": "transh_grid_kernel<true> (grid mode, API-faithful h_/r/w/t_ rows + score, %d triples)" % T,
                                     "bound": "hbm", "achieved": nb / sec / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                     "frac": nb / sec / 1e9 / hbm_peak, "traffic": None, "bytes_per_launch": nb,
                                     "us_per_launch": sec * 1e6, "triples_per_s": T / sec,
                                     "us_per_launch_l2_flushed_single": sec_flush * 1e6}
    del outs, score
    # general kernel, explicit int64 index arrays, API-faithful rows: 804 B written + 24 B read per triple
    T2 = 1 << 20
    g2 = torch.Generator(device=dev).manual_seed(3)
    hd = torch.randint(0, 80, (T2,), device=dev, generator=g2); rl = torch.randint(0, 117, (T2,), device=dev, generator=g2)
    tt_ = torch.randint(0, 80, (T2,), device=dev, generator=g2)
    outs = [torch.empty(T2, 50, dtype=torch.float32, device=dev) for _ in range(4)]
    score = torch.empty(T2, dtype=torch.float32, device=dev)

    def th_explicit():
        _lib.call("skg_transh_forward", p(e), p(r), p(w), 80, 117, 50, p(hd), p(rl), p(tt_), T2, 2, 1, p(outs[0]), p(outs[1]),
                  p(outs[2]), p(outs[3]), p(score), _lib.stream_ptr(dev))
    sec_flush = _time_kernel(torch, th_explicit, flush, reps=10)
    sec = _time_rotating(torch, [th_explicit], rounds=10)     # one launch writes > 126 MB: nothing of it survives in L2
    nb = 828 * T2
    extra["transh_explicit_rows_roofline"] = {"kernel": "transh_forward_kernel (explicit int64 indices, h_/r/w/t_ rows + score, %d triples)" % T2,
                                              "bound": "hbm", "achieved": nb / sec / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                              "frac": nb / sec / 1e9 / hbm_peak, "traffic": None, "bytes_per_launch": nb,
                                              "us_per_launch": sec * 1e6, "triples_per_s": T2 / sec,
                                              "us_per_launch_l2_flushed_single": sec_flush * 1e6}

    def th_explicit_score():
        _lib.call("skg_transh_forward", p(e), p(r), p(w), 80, 117, 50, p(hd), p(rl), p(tt_), T2, 2, 1, None, None, None, None,
                  p(score), _lib.stream_ptr(dev))
    sec = _time_kernel(torch, th_explicit_score, flush, reps=10)
    extra["transh_explicit_score"] = {"kernel": "transh_forward_kernel (explicit indices, score only: issue-bound, 28 B / triple)",
                                      "triples_per_s": T2 / sec, "us_per_launch": sec * 1e6, "achieved_gbs": 28 * T2 / sec / 1e9}
    return roof, extra


def _quiet_stdout():
    """Libraries (NCCL's version banner, torchrun notices) write to fd 1; the contract is ONE JSON line on stdout.
    Point fd 1 at stderr for the run and hand back a file object on the real stdout for the final line."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


if __name__ == "__main__":
    a = parse_args()
    _real_stdout = _quiet_stdout()
    sys.stdout = _real_stdout
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
    sys.stdout.flush()
