#!/usr/bin/env python
"""Benchmark of the PATHS hierarchical forward+backward step (BASELINE.json metric: slides/sec fwd+bwd, 5 magnifications,
K=20; % of roofline; vs host CPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16x3|bf16|fp32]
                    [--workload cfg3|cfg2|cfg4|cfg5] [--slides-per-gpu B] [--graph]

Workloads (BASELINE.json:configs).  The default, and the line the driver records, is cfg 3:
  cfg3  BRCA-shaped synthetic pyramids (level-0 grid W0~U{5..10} x H0~U{4..8}, 5 levels, x2 per level, 1024-d features, 30 %
        background), models/brca_paths_0 architecture with random weights, survival NLL, fwd+bwd, slide-level data
        parallelism with the bucketed NCCL all-reduce of the flat gradient buffer (weak scaling: fixed slides per GPU).
        `--slides-per-gpu 32` is the reference's own batch size (the small-batch regime).
  cfg2  the same model, inference (`model.eval()`, no_grad) on 64 slides: hazards + per-level importance.
  cfg4  one 142x141 = 20 022-cell slide mixed with 63 BRCA-shaped slides in one varlen batch, K=64, fwd+bwd.
  cfg5  6 magnifications (0.3125x-10x), K=32, W0~U{3..5} x H0~U{2..4}, fwd+bwd.
One "step" = one pass (forward+backward, or forward for cfg2) over one batch.

`value`  : whole-job slides/s with the feature pyramids already resident in HBM (the dataset is loaded once, as the
           reference pre-loads it into host RAM), timed with CUDA events, max over ranks.
`e2e`    : the same metric through the reference-facing API over HOST batches -- the reference's loop body (train.py:58-65):
           utils.inference_end2end(batch) + loss.backward() + loss.item(), the level-0 feature rows and labels copied
           pinned-host -> device every step, the loss read back every step.  It is "host batch to gradients": the optimizer
           step is a separate kernel (`e2e.with_optimizer` reports the rate with opt.step() included).
`--impl reference`: the reference's CPU path on the host cores, same metric: the unmodified reference where a checkout is
           reachable ($PATHS_REFERENCE or /root/reference; kind "reference"), else its restatement oracle/paths_oracle.py
           (kind "port" -- the GPU box has no checkout: the reference is pure Python and cannot travel).
"""
import argparse
import copy
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

BRCA_CONFIG = {
    "wsi_dir": "/x", "preprocess_dir": "/x", "csv_path": "/x", "seed": 0, "batch_size": 32, "save_epochs": 10,
    "eval_epochs": 1, "lr": 2e-5, "lr_decay_per_epoch": 0.99, "base_power": 0.625, "magnification_factor": 2,
    "num_levels": 5, "num_epochs": 40, "min_epochs": 0, "top_k_patches": 20, "early_stopping": False,
    "hipt_splits": True, "hipt_val_proportion": 0.0, "model_type": "PATHS",
    "model_config": {"hierarchical_ctx": True, "slide_ctx_mode": "residual", "patch_embed_dim": 1024, "dropout": 0.0,
                     "importance_mode": "mul", "trans_dim": 128, "trans_heads": 4, "trans_layers": 2,
                     "pos_encoding_mode": "2d", "importance_mlp_hidden_dim": 128,
                     "hierarchical_ctx_mlp_hidden_dim": 256, "lstm": True},
    "root_name": "brca_paths",
}
WORKLOADS = {
    "cfg3": dict(desc="cfg3 BRCA-shaped survival training fwd+bwd (models/brca_paths_0, L=5, K=20, D=1024)",
                 over={}, wrange=(5, 10), hrange=(4, 8), train=True, slides=672, metric="slides_per_sec_fwd_bwd"),
    "cfg2": dict(desc="cfg2 models/brca_paths_0 inference (eval, no_grad), batch of 64 BRCA-shaped slides, hazards + per-level importance",
                 over={}, wrange=(5, 10), hrange=(4, 8), train=False, slides=64, metric="slides_per_sec_fwd"),
    "cfg4": dict(desc="cfg4 large-bag stress: one 142x141 = 20 022-cell slide + 63 BRCA-shaped slides in one varlen batch, L=5, K=64, fwd+bwd",
                 over={"top_k_patches": 64}, wrange=(5, 10), hrange=(4, 8), train=True, slides=64, big=(142, 141),
                 metric="slides_per_sec_fwd_bwd"),
    "cfg5": dict(desc="cfg5 deep hierarchy: 6 magnifications (0.3125x-10x), K=32, W0~U{3..5} x H0~U{2..4}, fwd+bwd",
                 over={"num_levels": 6, "base_power": 0.3125, "top_k_patches": 32}, wrange=(3, 5), hrange=(2, 4), train=True,
                 slides=448, metric="slides_per_sec_fwd_bwd"),
}
DTYPES = {"fp32": "f32", "bf16x3": "bf16x3 (split bf16, fp32 accumulate; fp32-parity 1e-4)", "bf16": "bf16"}


def workload_config(name):
    data = copy.deepcopy(BRCA_CONFIG)
    data.update(WORKLOADS[name]["over"])
    return data


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  NVML from a thread of this
    process (first sample immediately, then every 10 ms: a 20-step region is ~0.3 s, and a freshly spawned nvidia-smi can take longer
    than that to print its first row); `nvidia-smi -lms` as the fallback when NVML cannot be loaded."""
    REASONS = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]

    def __init__(self, index):
        self.rows, self.p, self.f, self.th = [], None, None, None
        self.stop_flag = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = torch.cuda.get_device_properties(index).uuid
            try:
                h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
            except Exception:
                vis = os.environ.get("CUDA_VISIBLE_DEVICES")
                h = pynvml.nvmlDeviceGetHandleByIndex(int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))

            def loop():
                while True:
                    try:
                        self.rows.append((float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)),
                                          int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                                          if hasattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons")
                                          else int(pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h))))
                    except Exception:
                        pass
                    if self.stop_flag.wait(0.01):
                        return
            self.th = threading.Thread(target=loop, daemon=True)
            self.th.start()
            return
        except Exception:
            self.th = None
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "50",
                                       "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if self.th is not None:
            self.stop_flag.set()
            self.th.join(timeout=2)
            sm = sorted(r[0] for r in self.rows)
            bits = 0
            for r in self.rows:
                bits |= r[1]
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.mx,
                    "reasons": sorted(n for n, m in self.REASONS if bits & m), "samples": len(sm), "source": "nvml"}
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[4:8]):
                    if v.strip().lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# ---------------------------------------------------------------------------------------------------------------------
CPU_SAMPLE_SLIDES, CPU_SAMPLE_STEPS = 32, 30      # the reference's own batch size (larger batches run slower per slide)
CPU_ALT_SLIDES, CPU_ALT_STEPS = 128, 2            # ... and a larger batch, reported beside it


def cpu_rate(workload, n_slides, steps, warmup, threads=None):
    """slides/s of the reference's CPU path on a bounded sample of the workload: the unmodified reference when a checkout is
    reachable, else its restatement (oracle/paths_oracle.py).  Returns (rate, s/step, threads, kind)."""
    from oracle import ref_runner
    from paths_b200 import synth
    w = WORKLOADS[workload]
    if "big" in w:
        return None
    if threads is None:                      # torchrun exports OMP_NUM_THREADS=1: use every core this process may run on
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(max(1, threads))
    data = workload_config(workload)
    slides = synth.make_brca_batch(range(n_slides), data["num_levels"], data["model_config"]["patch_embed_dim"],
                                   wrange=w["wrange"], hrange=w["hrange"])
    ref = ref_runner.reference_dir()
    if ref is not None:
        runner = ref_runner.ReferenceRunner(ref, data, slides)
        step = lambda: runner.step(backward=w["train"])  # noqa: E731
        kind = "reference"
    else:
        from oracle import paths_oracle as po
        from paths_b200.config import Config
        cfg = Config.from_dict(data)
        torch.manual_seed(cfg.seed)
        model = cfg.get_model()
        sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
        ocfg = po.OracleConfig.from_config(cfg)
        ybin = torch.tensor([s.survival_bin for s in slides]); cens = torch.tensor([s.censored for s in slides])
        grids = [s.grids for s in slides]
        if w["train"]:
            step = lambda: po.grads(sd, ocfg, grids, ybin, cens)  # noqa: E731
        else:
            def step():
                with torch.no_grad():
                    po.end2end(sd, ocfg, grids, ybin, cens)
        kind = "port"
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = time.perf_counter() - t0
    return n_slides * steps / dt, dt / steps, torch.get_num_threads(), kind


def _kind_text(kind):
    return "unmodified reference, imported from a checkout" if kind == "reference" else "oracle/paths_oracle.py, the torch-CPU restatement"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = WORKLOADS[args.workload]
    n = 32                                   # the reference's own batch_size (models/brca_paths_0/config.json:6)
    steps = max(1, min(args.steps, 200)); warm = max(0, min(args.warmup, 50))     # ~0.3 s per step: the driver's K and W as given
    got = cpu_rate(args.workload, n, steps, warm)
    if got is None:
        print(json.dumps({"impl": "reference", "unavailable": f"{args.workload}: the 28 GB pyramid of the 20 022-cell slide is not materialised on the host"}),
              file=_OUT, flush=True)
        return
    rate, spp, cores, kind = got
    what = "fwd+bwd" if w["train"] else "forward (no_grad)"
    line = {
        "metric": w["metric"], "value": rate, "unit": "slides/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": spp * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": w["desc"], "slides_per_step": n, "parallelism": "host CPU threads",
                   "note": "32 slides per step = the reference's own batch_size; the GPU arm runs the same workload at its own per-GPU batch"},
        "cpu_baseline": {"value": rate, "unit": "slides/s", "cores": cores, "kind": kind,
                         "sample": f"{n} slides per step, {what}, {steps} steps ({_kind_text(kind)})"},
        "e2e": {"value": rate, "unit": "slides/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_OUT, flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# The JSON line is the only thing that may reach stdout: libraries that print there (NCCL's version banner under torchrun)
# are sent to stderr for the duration of the run.
_OUT = sys.stdout


def _claim_stdout():
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


# Reference-equivalent algorithmic work: no credit for skipping the depth-0 h = 0 half of the gate contraction
# or for pruning the last decoder layer to the special token.  MACs per packed row, forward:
MAC_LSTM = 2 * 1024 * (3 * 256 + 1024) + 256 * 1024 + (1024 * 128 + 128) + 1024 * 128      # LSTM cell + importance + proj_in
MAC_DEC = 2 * (4 * 128 * 128 + 2 * 128 * 512)                                               # two decoder layers
BYTES_ROW_FWD = 4096 + 5120 + 5120 + 4                                                       # feature, parent state, state, alpha


def step_roofline(bags, train, peak_tflops, hbm_gbs, live_grad_bytes=28.9e6):
    """bags[l] = per-slide bag sizes of level l.  Returns flops / bytes / roofline seconds of one step, whole and per region."""
    mul = 3.0 if train else 1.0
    rows = sum(sum(b) for b in bags)
    attn = sum(2.0 * 2 * 128 * n * (n + 1) for b in bags for n in b)          # 2 layers x (QK^T + PV) = 512 T MAC per token
    f_lstm = 2.0 * MAC_LSTM * rows * mul
    f_dec = (2.0 * MAC_DEC * rows + 2.0 * attn) * mul
    by = BYTES_ROW_FWD * rows * (2.0 if train else 1.0) + (2 * live_grad_bytes if train else 0.0)
    flops = f_lstm + f_dec
    t = max(flops / (peak_tflops * 1e12), by / (hbm_gbs * 1e9))
    return {"flops": flops, "bytes": by, "t_roof_s": t, "flops_lstm": f_lstm, "flops_decoder": f_dec, "rows": rows}


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("PATHS_B200_PRECISION", "bf16x3"), choices=["fp32", "bf16x3", "bf16"])
    ap.add_argument("--workload", default="cfg3", choices=sorted(WORKLOADS))
    ap.add_argument("--slides-per-gpu", type=int, default=None)
    ap.add_argument("--dropout", type=float, default=0.0,
                    help="aggregator dropout of the training step (BASELINE.json's workload is quoted at 0.0; the shipped configs train at 0.05)")
    ap.add_argument("--store-dtype", default="fp32", choices=["fp32", "bf16"],
                    help="element type of the HBM-resident pyramid store (bf16: features rounded once at load, half the footprint)")
    ap.add_argument("--graph", action="store_true", help="replay the resident step as one CUDA graph (ModelRuntime.graph_step)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-fast-mode", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import torch.distributed as dist
    from paths_b200 import synth
    from paths_b200.config import Config
    from paths_b200.data_utils.dataset import collate_fn
    from paths_b200.data_utils.prefetch import DevicePrefetcher
    from paths_b200.data_utils.slide import load_patch_preprocessed_slide
    from paths_b200.engine import PyramidStore
    from paths_b200.parallel import DataParallelStep
    from paths_b200.utils import inference_end2end

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W = WORKLOADS[args.workload]
    train = W["train"]
    B = args.slides_per_gpu or W["slides"]
    cfg = Config.from_dict(workload_config(args.workload))
    cfg.model_config.dropout = args.dropout
    torch.manual_seed(cfg.seed)
    model = cfg.get_model().to(dev)
    model.train() if train else model.eval()
    rt = model.runtime()
    rt.set_precision(args.precision)
    D, Lv = cfg.model_config.patch_embed_dim, cfg.num_levels

    # ---- the rank's shard of the dataset: pyramids resident in HBM
    store = PyramidStore(Lv, D, dev, dtype=torch.bfloat16 if args.store_dtype == "bf16" else torch.float32)
    items = []
    for i in range(B):
        sid = rank * B + i
        shape = W["big"] if ("big" in W and i == 0) else synth.brca_shape(sid, wrange=W["wrange"], hrange=W["hrange"])
        s = synth.make_device_slide_fast(sid, *shape, Lv, D, 0.3, dev)
        sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(),
                                           Lv, grids=s.grids, store=store)
        items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored, "slide": sl})
    batch = collate_fn(items, packed=True)
    batch["fts_packed"] = batch["fts_packed"].pin_memory()
    padded_bytes = batch["fts"].numel() * 4
    batch["fts"] = batch["fts"][:, :1].clone()                 # the padded copy never leaves the host; keep a stub of the schema key
    ids = [it.store_index for it in batch["slide"]]
    caps, maxseg = store.capacities(ids, cfg.top_k_patches)
    eng = rt.engine_for(B, caps, maxseg, cfg.top_k_patches)
    eng.bind_store(store)
    slide_ids = torch.tensor(ids, dtype=torch.int32, device=dev)
    ybin = batch["survival_bin"].to(dev, torch.int32); cens = batch["censored"].to(dev, torch.int32)
    scale = 1.0 / (B * world)
    dp = DataParallelStep(model)
    dp.time_exposed_comm(world > 1)
    exposed = []

    def step_resident():
        if args.dropout > 0:
            rt._arm_dropout(eng)                               # a fresh seed per step, as inference_end2end does
        eng.forward(slide_ids, ybin, cens, scale, True)
        over, need = eng.overflow_poll()                       # the product path's check (waits for the last select only)
        assert not over, f"row capacity overflow ({need})"
        if train:
            dp.backward(eng)                                   # whole-step backward + bucketed all-reduce behind it

    graph_info = None
    step_fn = step_resident
    if args.graph:
        assert world == 1 and args.dropout == 0, "--graph: single-GPU resident step without dropout"
        gs = rt.graph_step(eng, slide_ids, ybin, cens, scale, backward=train)
        step_fn = gs.replay
        graph_info = {"nodes": gs.nodes}

    from paths_b200.optim import FlatAdamW
    opt = FlatAdamW(model, lr=cfg.lr)

    prefetcher = DevicePrefetcher([], dev)

    def run_e2e(steps, with_opt=False):
        # the reference's train-loop body (train.py:58-66) over host batches; the loader wrapper stages batch i+1 into HBM
        # (pinned host -> device, on a copy stream) while step i runs.  Every step copies its own inputs and reads its loss.
        losses = []
        if train:
            dp.attach(global_batch=B * world)
        try:
            prefetcher.loader = (batch for _ in range(steps))      # one prefetcher for the run, as around a DataLoader: its device
            for b in prefetcher:                                   # ring is allocated once, not per epoch
                if train:
                    opt.zero_grad()                                # train.py:59
                    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
                    loss.backward()                                # all-reduce inside (attached DataParallelStep)
                    if with_opt:
                        opt.step()                                 # train.py:66
                else:
                    with torch.no_grad():
                        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
                losses.append(float(loss.item()))
                step_wall.append(time.perf_counter())
                if world > 1 and train:
                    exposed.append(dp.last_exposed_comm())
        finally:
            dp.detach()
        if not with_opt:
            e2e_losses.extend(losses)
        return losses

    host_ms = [0.0]
    e2e_losses = []
    step_wall = []
    # The synthetic dataset is hundreds of thousands of live Python objects; a full garbage-collector pass over them costs tens of
    # milliseconds, and the end-to-end loop (which waits for the loss every step) pays such a pause in full.  Freeze what exists
    # now (the objects of set-up never die): later collections only look at what the steps themselves allocate.
    import gc
    gc.collect()
    gc.freeze()

    def timed(fn, steps, warmup, profile=False, multi=False):
        if multi:
            fn(warmup)
        else:
            for _ in range(warmup):
                fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        sampler = ClockSampler(local) if (rank == 0 and profile) else None
        if profile:
            eng.profile(True)
        l0 = eng.launch_count()
        cuprof = profile and os.environ.get("PATHS_BENCH_CUPROF") == "1"    # ncu --profile-from-start off
        if cuprof:
            torch.cuda.cudart().cudaProfilerStart()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        h0 = time.perf_counter()
        if multi:
            fn(steps)
        else:
            for _ in range(steps):
                fn()
        host_ms[0] = (time.perf_counter() - h0) * 1e3 / steps      # host time per step inside the loop (enqueue + the overflow poll)
        b.record()
        torch.cuda.synchronize()
        if cuprof:
            torch.cuda.cudart().cudaProfilerStop()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        if profile:
            eng.profile(False)
        clocks = sampler.stop() if sampler else None
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        n_coll = (Lv + 1) * steps if (world > 1 and train) else 0      # NCCL kernels of the bucketed all-reduce
        return ms, eng.launch_count() - l0 + n_coll, clocks

    use_prof = not args.graph                                   # (event records inside a captured graph would be replayed, not re-timed)
    ms, launches, clocks = timed(step_fn, args.steps, args.warmup, profile=use_prof)
    if args.graph:
        launches = graph_info["nodes"] * args.steps
    host_enqueue_ms = host_ms[0]
    exposed_resident = dp.last_exposed_comm() if (world > 1 and train) else 0.0
    loss_resident = float(eng.buffer(0, "LOSS").item())       # read after the timed region
    over, need = eng.overflowed()
    assert not over, f"row capacity overflow ({need})"
    counts = eng.counts()
    bags = []
    for l in range(Lv):
        seg = eng.buffer(l, "SEG")[:B + 1].tolist()
        bags.append([seg[i + 1] - seg[i] for i in range(B)])
    gate_ms = dw_ms = 0.0
    regions = {}
    if use_prof:
        gate_ms, gate_n = eng.profile_read(0)
        dw_ms, dw_n = eng.profile_read(1)
        eng.profile_read(2)
        # in-situ region times (CUDA events on the stream each region runs on; decoder regions overlap the LSTM chain)
        for tag, name in [(3, "lstm_fwd"), (4, "decoder_fwd"), (5, "select_gather"), (6, "decoder_bwd"), (7, "lstm_bwd")]:
            r_ms, _ = eng.profile_read(tag)
            regions[name] = r_ms / args.steps
    value = B * world * args.steps / (ms / 1e3)

    # ---- roofline of the dominant kernel: the LSTM gate contraction G = [x || h_parent] . Wg^T  (tcgen05, TMEM accumulators)
    C = cfg.model_config.hierarchical_ctx_mlp_hidden_dim
    flops_per_step = sum(2.0 * m * (3 * C + D) * (D if l == 0 else 2 * D) for l, m in enumerate(counts))
    # what the kernels execute: in the tensor-core modes a level with fewer parent slots (slides x kept rows) than rows computes
    # W_h h_parent once per slot and adds it in the gate epilogue (engine.cu: Level::shared) -- the algorithmic count above stays the
    # reference's (no credit for the shortcut in `achieved`), the issue fraction below is of the executed products
    topk = list(cfg.top_k_patches)
    slots = [0] + [B * min(topk[l - 1], eng.maxseg[l - 1]) if topk[l - 1] >= 0 else B * eng.maxseg[l - 1] for l in range(1, Lv)]
    shared = [args.precision != "fp32" and l >= 1 and slots[l] <= eng.caps[l] and "PATHS_B200_NO_SHARED_PARENT" not in os.environ
              and bool(cfg.model_config.lstm) for l in range(Lv)]
    executed_flops = sum(2.0 * (3 * C + D) * D * ((m + slots[l]) if shared[l] else (m if l == 0 else 2 * m)) for l, m in enumerate(counts))
    pk, pk_src = peaks()
    peak = pk.get("bf16_tflops_sustained", pk.get("bf16_tflops"))
    hbm = pk.get("hbm_gbs", 6650.0)
    achieved = flops_per_step * args.steps / (gate_ms / 1e3) / 1e12 if gate_ms > 0 else None
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp) and args.workload == "cfg3" and B == 672:
        traffic = json.load(open(tp)).get(args.precision)
    passes = {"fp32": None, "bf16x3": 3, "bf16": 1}[args.precision]
    sr = step_roofline(bags, train, peak, hbm)
    step_frac = sr["t_roof_s"] / (ms / 1e3 / args.steps)
    region_frac = {}
    if regions:
        fl = {"lstm_fwd": sr["flops_lstm"] / (3 if train else 1), "lstm_bwd": sr["flops_lstm"] * 2 / 3,
              "decoder_fwd": sr["flops_decoder"] / (3 if train else 1), "decoder_bwd": sr["flops_decoder"] * 2 / 3}
        for k, f in fl.items():
            if regions.get(k, 0) > 0:
                region_frac[k] = f / (peak * 1e12) / (regions[k] / 1e3)
    roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak if (peak and achieved) else None,
                "traffic": traffic, "kernel": "gemm_tma_kernel (LSTM gate contraction, fwd)" if passes else "gemm_nt_kernel (fp32 SIMT)",
                "peak_source": pk_src + ", bf16 dense sustained", "kernel_ms_per_step": gate_ms / args.steps,
                "kernel_share_of_step": gate_ms / ms, "algorithmic_flops_per_step": flops_per_step,
                "executed_flops_per_step": executed_flops, "shared_parent_levels": [l for l in range(Lv) if shared[l]],
                "mma_passes": passes,
                "tensor_issue_frac": (achieved * executed_flops / flops_per_step * passes / peak) if (passes and peak and achieved) else None,
                "dW_kernel_tflops": (flops_per_step * args.steps / (dw_ms / 1e3) / 1e12) if dw_ms > 0 else None,
                "region_ms_per_step": regions,
                # whole step against the roofline, reference-equivalent algorithmic work (9.18 MFLOP + 1024 T per row
                # forward, x3 for fwd+bwd; 14.3 KB per row forward, x2): step_frac = max(flops / peak, bytes / HBM) / ms_per_step.
                # In bf16x3 mode every algorithmic MAC of a contraction is three tensor-core passes: the fraction is capped near 1/3.
                "step_frac": step_frac, "step_flops": sr["flops"], "step_bytes": sr["bytes"], "step_bound": "tensor" if sr["flops"] / (peak * 1e12) >= sr["bytes"] / (hbm * 1e9) else "hbm",
                "step_tensor_issue_frac": step_frac * passes if passes else None,
                "region_frac": region_frac}

    # ---- end-to-end through the reference-facing API with host buffers
    e2e = None
    if not args.no_e2e:
        exposed.clear()
        e_ms, _, _ = timed(run_e2e, args.steps, max(5, args.warmup), multi=True)
        w = step_wall[-args.steps:]
        per = sorted((b - a) * 1e3 for a, b in zip(w[:-1], w[1:]))
        h2d = batch["fts_packed"].numel() * 4 + B * 4 * 2
        e2e = {"value": B * world * args.steps / (e_ms / 1e3), "unit": "slides/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 4 + 8, "ms_per_step": e_ms / args.steps,
               "api": "for batch in DevicePrefetcher(host batches): paths_b200.utils.inference_end2end(batch) + loss.backward() + "
                      "loss.item() -- host batch to gradients (no optimizer step); level-0 rows cross PCIe unpadded (collate_fn(packed=True): "
                      f"{h2d} bytes against {padded_bytes} for the padded tensor); the pinned-host -> HBM copy of step i+1 overlaps step i; "
                      "d2h = the loss + the 8-byte overflow flag",
               "exposed_comm_ms_per_step": (sum(exposed[-args.steps:]) / max(1, len(exposed[-args.steps:]))) if exposed else 0.0,
               "host_wall_ms_per_step": {"min": per[0], "median": per[len(per) // 2], "max": per[-1]} if per else None}
        if train:
            o_ms, _, _ = timed(lambda n: run_e2e(n, with_opt=True), args.steps, max(5, args.warmup), multi=True)
            e2e["with_optimizer"] = {"value": B * world * args.steps / (o_ms / 1e3), "unit": "slides/s", "ms_per_step": o_ms / args.steps,
                                     "what": "the same loop with opt.step() (fused flat AdamW, train.py:66) after loss.backward()"}

    # ---- the stated-tolerance fast mode (single-pass bf16 operands, 1e-2), same workload, reported beside the headline
    fast = None
    if args.precision == "bf16x3" and not args.no_fast_mode and not args.graph:
        rt.set_precision("bf16")
        f_ms, _, _ = timed(step_resident, args.steps, 3, profile=True)
        fg_ms, _ = eng.profile_read(0)
        for tag in range(1, 8):
            eng.profile_read(tag)
        f_ach = flops_per_step * args.steps / (fg_ms / 1e3) / 1e12 if fg_ms > 0 else 0.0
        fast = {"dtype": "bf16 (single pass, stated tolerance 1e-2)", "value": B * world * args.steps / (f_ms / 1e3), "unit": "slides/s",
                "ms_per_step": f_ms / args.steps, "gate_kernel_tflops": f_ach, "gate_kernel_frac_of_peak": f_ach / peak if peak else None,
                "step_frac": sr["t_roof_s"] / (f_ms / 1e3 / args.steps)}
        rt.set_precision(args.precision)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        got = cpu_rate(args.workload, CPU_SAMPLE_SLIDES, CPU_SAMPLE_STEPS, 1)          # ~10 s of host work
        if got is not None:
            rate, spp, cores, kind = got
            what = "fwd+bwd" if train else "forward (no_grad)"
            cpu = {"value": rate, "unit": "slides/s", "cores": cores, "kind": kind,
                   "sample": f"{CPU_SAMPLE_SLIDES} slides per step (the reference's batch_size), {what}, {CPU_SAMPLE_STEPS} timed steps after 1 warm-up ({_kind_text(kind)})"}
            alt = cpu_rate(args.workload, CPU_ALT_SLIDES, CPU_ALT_STEPS, 1)
            cpu["larger_batch"] = {"value": alt[0], "slides_per_step": CPU_ALT_SLIDES, "steps": CPU_ALT_STEPS}

    # full-size consistency (no oracle at this size): the resident path (features gathered from the HBM store) and the
    # end-to-end path (the same features arriving as host rows) must compute the same loss, step after step
    consistency = None
    if e2e_losses:
        consistency = {"loss_resident": loss_resident, "loss_e2e": e2e_losses[-1],
                       "rel_diff": abs(e2e_losses[-1] - loss_resident) / max(abs(loss_resident), 1e-30),
                       "e2e_step_spread": max(e2e_losses) - min(e2e_losses)}
    if rank == 0:
        rows = sum(counts)
        line = {
            "metric": W["metric"], "value": value, "unit": "slides/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": DTYPES[args.precision], "data": "synthetic",
            "config": {"workload": W["desc"], "slides_per_gpu": B, "global_batch": B * world, "levels": Lv, "top_k": list(cfg.top_k_patches)[0],
                       "packed_rows_per_step_per_gpu": rows, "rows_per_level": counts, "largest_bag_per_level": [max(b) for b in bags],
                       "parallelism": f"dp{world}", "dropout": args.dropout, "cuda_graph": bool(args.graph), "store_dtype": args.store_dtype,
                       "l2": "per-step working set (activations of the packed rows, gathers from the HBM-resident pyramid store) exceeds the 126 MB L2: no flush needed"
                             if rows * 100e3 > 126e6 else "working set below L2 at this batch size: the per-step weight re-pack and activation writes evict little; stated, not flushed"},
            "host_ms_per_step": host_enqueue_ms, "exposed_comm_ms_per_step": exposed_resident, "consistency": consistency,
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "fast_mode": fast, "gpu_launches": launches, "clocks": clocks,
        }
        print(json.dumps(line), file=_OUT, flush=True)
    if os.environ.get("PATHS_B200_PROF_DUMP"):               # developer aid: per-call-site timings are written at engine destruction
        torch.cuda.synchronize()
        rt.engines.clear(); rt.last_engine = None
        del eng
        import gc; gc.collect()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
