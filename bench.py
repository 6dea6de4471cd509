#!/usr/bin/env python
"""Benchmark of the PATHS hierarchical forward+backward step (BASELINE.json metric: slides/sec fwd+bwd, 5 magnifications,
K=20; % of roofline; vs host CPU).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--precision bf16x3|bf16|fp32]

Workload (SURVEY.md §8(d) cfg 3): BRCA-shaped synthetic pyramids (level-0 grid W0~U{5..10} x H0~U{4..8}, 5 levels, x2 per
level, 1024-d features, 30 % background), models/brca_paths_0 architecture with random weights, survival NLL loss,
slide-level data parallelism with one NCCL all-reduce of the flat gradient buffer per step (weak scaling: fixed slides per
GPU).  One "step" = one forward+backward pass over one batch.

`value`  : whole-job slides/s with the feature pyramids already resident in HBM (the dataset is loaded once, as the
           reference pre-loads it into host RAM), timed with CUDA events, max over ranks.
`e2e`    : the same metric through the reference-facing API -- utils.inference_end2end(batch) on the collated HOST batch
           (pinned level-0 features + labels copied host->device every step) + loss.backward() + loss.item().
`--impl reference`: the CPU oracle (torch-CPU restatement of the reference's own code path; the reference is pure Python
           on PyTorch and cannot travel to the GPU box) on the host cores, same metric.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

WORKLOAD = "cfg3 BRCA-shaped survival training fwd+bwd (models/brca_paths_0, L=5, K=20, D=1024)"
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
DTYPES = {"fp32": "f32", "bf16x3": "bf16x3 (split bf16, fp32 accumulate; fp32-parity 1e-4)", "bf16": "bf16"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "50",
                                       "-i", str(index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
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
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
CPU_SAMPLE_SLIDES, CPU_SAMPLE_STEPS = 32, 30      # the reference's own batch size (larger oracle batches run slower per slide)


def cpu_oracle_rate(n_slides, steps, warmup, threads=None):
    """slides/s of the CPU oracle (fwd+bwd) on a bounded sample of the same workload."""
    from oracle import paths_oracle as po
    from paths_b200 import synth
    from paths_b200.config import Config
    if threads is None:                      # torchrun exports OMP_NUM_THREADS=1: use every core this process may run on
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(max(1, threads))
    cfg = Config.from_dict(BRCA_CONFIG)
    torch.manual_seed(cfg.seed)
    model = cfg.get_model()
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    ocfg = po.OracleConfig.from_config(cfg)
    slides = synth.make_brca_batch(range(n_slides), cfg.num_levels, cfg.model_config.patch_embed_dim)
    ybin = torch.tensor([s.survival_bin for s in slides]); cens = torch.tensor([s.censored for s in slides])
    grids = [s.grids for s in slides]
    for _ in range(warmup):
        po.grads(sd, ocfg, grids, ybin, cens)
    t0 = time.perf_counter()
    for _ in range(steps):
        po.grads(sd, ocfg, grids, ybin, cens)
    dt = time.perf_counter() - t0
    return n_slides * steps / dt, dt / steps, torch.get_num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = 32                                   # the reference's own batch_size (models/brca_paths_0/config.json:6)
    steps = max(1, min(args.steps, 200)); warm = max(0, min(args.warmup, 50))     # ~0.3 s per step: the driver's K and W as given
    rate, spp, cores = cpu_oracle_rate(n, steps, warm)
    line = {
        "metric": "slides_per_sec_fwd_bwd", "value": rate, "unit": "slides/s", "n_gpus": args.gpus, "steps": steps, "warmup": warm,
        "ms_per_step": spp * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "impl": "reference",
        "config": {"workload": WORKLOAD, "slides_per_step": n, "parallelism": "host CPU threads"},
        "cpu_baseline": {"value": rate, "unit": "slides/s", "cores": cores, "kind": "port",
                         "sample": f"{n} BRCA-shaped slides per step, fwd+bwd, {steps} steps (oracle/paths_oracle.py, torch CPU)"},
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


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("PATHS_B200_PRECISION", "bf16x3"), choices=["fp32", "bf16x3", "bf16"])
    ap.add_argument("--slides-per-gpu", type=int, default=672)
    ap.add_argument("--dropout", type=float, default=0.0,
                    help="aggregator dropout of the training step (BASELINE.json's workload is quoted at 0.0; the shipped configs train at 0.05)")
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
    from paths_b200.utils import inference_end2end

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.slides_per_gpu
    cfg = Config.from_dict(BRCA_CONFIG)
    cfg.model_config.dropout = args.dropout
    torch.manual_seed(cfg.seed)
    model = cfg.get_model().to(dev)
    model.train()
    rt = model.runtime()
    rt.set_precision(args.precision)
    rt.global_batch = B * world
    D, Lv = cfg.model_config.patch_embed_dim, cfg.num_levels

    # ---- the rank's shard of the dataset: pyramids resident in HBM
    store = PyramidStore(Lv, D, dev)
    items = []
    for i in range(B):
        sid = rank * B + i
        s = synth.make_device_slide_fast(sid, *synth.brca_shape(sid), Lv, D, 0.3, dev)
        sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(),
                                           Lv, grids=s.grids, store=store)
        items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored, "slide": sl})
    batch = collate_fn(items)
    batch["fts"] = batch["fts"].pin_memory()
    ids = [it.store_index for it in batch["slide"]]
    caps, maxseg = store.capacities(ids, cfg.top_k_patches)
    eng = rt.engine_for(B, caps, maxseg, cfg.top_k_patches)
    eng.bind_store(store)
    slide_ids = torch.tensor(ids, dtype=torch.int32, device=dev)
    ybin = batch["survival_bin"].to(dev, torch.int32); cens = batch["censored"].to(dev, torch.int32)
    scale = 1.0 / (B * world)
    gflat = rt.layout.gflat

    def step_resident():
        if args.dropout > 0:
            rt._arm_dropout(eng)                               # a fresh seed per step, as inference_end2end does
        eng.forward(slide_ids, ybin, cens, scale, True)
        eng.backward(True)
        if world > 1:
            dist.all_reduce(gflat)

    from paths_b200.optim import FlatAdamW
    opt = FlatAdamW(model, lr=cfg.lr) if hasattr(cfg, "lr") else FlatAdamW(model)

    def run_e2e(steps):
        # the reference's train-loop body (train.py:58-65) over host batches; the loader wrapper stages batch i+1 into HBM
        # (pinned host -> device, on a copy stream) while step i runs.  Every step copies its own inputs and reads its loss.
        losses = []
        for b in DevicePrefetcher((batch for _ in range(steps)), dev):
            opt.zero_grad()                                    # train.py:59
            hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
            loss.backward()
            if world > 1:
                dist.all_reduce(gflat)
            losses.append(float(loss.item()))
        e2e_losses.extend(losses)
        return losses

    host_ms = [0.0]
    e2e_losses = []

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
        host_ms[0] = (time.perf_counter() - h0) * 1e3 / steps      # host time to enqueue one step (no device sync inside)
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
        return ms, eng.launch_count() - l0 + (steps if world > 1 else 0), clocks

    ms, launches, clocks = timed(step_resident, args.steps, args.warmup, profile=True)
    host_enqueue_ms = host_ms[0]
    loss_resident = float(eng.buffer(0, "LOSS").item())       # read after the timed region
    over, need = eng.overflowed()
    assert not over, f"row capacity overflow ({need})"
    counts = eng.counts()
    gate_ms, gate_n = eng.profile_read(0)
    dw_ms, dw_n = eng.profile_read(1)
    eng.profile_read(2)
    # in-situ region times (CUDA events on the stream each region runs on; decoder regions overlap the LSTM chain)
    regions = {}
    for tag, name in [(3, "lstm_fwd"), (4, "decoder_fwd"), (5, "select_gather"), (6, "decoder_bwd"), (7, "lstm_bwd")]:
        r_ms, _ = eng.profile_read(tag)
        regions[name] = r_ms / args.steps
    value = B * world * args.steps / (ms / 1e3)

    # ---- roofline of the dominant kernel: the LSTM gate contraction G = [x || h_parent] . Wg^T  (tcgen05, TMEM accumulators)
    C = cfg.model_config.hierarchical_ctx_mlp_hidden_dim
    flops_per_step = sum(2.0 * m * (3 * C + D) * (D if l == 0 else 2 * D) for l, m in enumerate(counts))
    pk, pk_src = peaks()
    peak = pk.get("bf16_tflops_sustained", pk.get("bf16_tflops"))
    achieved = flops_per_step * args.steps / (gate_ms / 1e3) / 1e12 if gate_ms > 0 else 0.0
    traffic = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        traffic = json.load(open(tp)).get(args.precision)
    passes = {"fp32": None, "bf16x3": 3, "bf16": 1}[args.precision]
    roofline = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                "traffic": traffic, "kernel": "gemm_tma_kernel (LSTM gate contraction, fwd)" if passes else "gemm_nt_kernel (fp32 SIMT)",
                "peak_source": pk_src + ", bf16 dense sustained", "kernel_ms_per_step": gate_ms / args.steps,
                "kernel_share_of_step": gate_ms / ms, "algorithmic_flops_per_step": flops_per_step,
                "mma_passes": passes, "tensor_issue_frac": (achieved * passes / peak) if passes and peak else None,
                "dW_kernel_tflops": (flops_per_step * args.steps / (dw_ms / 1e3) / 1e12) if dw_ms > 0 else None,
                "region_ms_per_step": regions}

    # ---- end-to-end through the reference-facing API with host buffers
    e2e = None
    if not args.no_e2e:
        e_ms, _, _ = timed(run_e2e, args.steps, 3, multi=True)
        h2d = batch["fts"].numel() * 4 + B * 4 * 3
        e2e = {"value": B * world * args.steps / (e_ms / 1e3), "unit": "slides/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": 4, "ms_per_step": e_ms / args.steps,
               "api": "for batch in DevicePrefetcher(host batches): paths_b200.utils.inference_end2end(batch) + loss.backward() + "
                      "loss.item(); the pinned-host -> HBM copy of step i+1 overlaps step i"}

    # ---- the stated-tolerance fast mode (single-pass bf16 operands, 1e-2), same workload, reported beside the headline
    fast = None
    if args.precision == "bf16x3" and not args.no_fast_mode:
        rt.set_precision("bf16")
        f_ms, _, _ = timed(step_resident, args.steps, 3, profile=True)
        fg_ms, _ = eng.profile_read(0)
        for tag in range(1, 8):
            eng.profile_read(tag)
        f_ach = flops_per_step * args.steps / (fg_ms / 1e3) / 1e12 if fg_ms > 0 else 0.0
        fast = {"dtype": "bf16 (single pass, stated tolerance 1e-2)", "value": B * world * args.steps / (f_ms / 1e3), "unit": "slides/s",
                "ms_per_step": f_ms / args.steps, "gate_kernel_tflops": f_ach, "gate_kernel_frac_of_peak": f_ach / peak if peak else None}
        rt.set_precision(args.precision)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        rate, spp, cores = cpu_oracle_rate(CPU_SAMPLE_SLIDES, CPU_SAMPLE_STEPS, 1)          # ~10 s of host work
        cpu = {"value": rate, "unit": "slides/s", "cores": cores, "kind": "port",
               "sample": f"{CPU_SAMPLE_SLIDES} BRCA-shaped slides per step, fwd+bwd, {CPU_SAMPLE_STEPS} timed steps after 1 warm-up (oracle/paths_oracle.py, torch CPU)"}

    # full-size consistency (no oracle at this size): the resident path (features gathered from the HBM store) and the
    # end-to-end path (the same features arriving as the padded host batch) must compute the same loss, step after step
    consistency = None
    if e2e_losses:
        consistency = {"loss_resident": loss_resident, "loss_e2e": e2e_losses[-1],
                       "rel_diff": abs(e2e_losses[-1] - loss_resident) / max(abs(loss_resident), 1e-30),
                       "e2e_step_spread": max(e2e_losses) - min(e2e_losses)}
    if rank == 0:
        rows = sum(counts)
        line = {
            "metric": "slides_per_sec_fwd_bwd", "value": value, "unit": "slides/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": DTYPES[args.precision], "data": "synthetic",
            "config": {"workload": WORKLOAD, "slides_per_gpu": B, "global_batch": B * world, "levels": Lv, "top_k": 20,
                       "packed_rows_per_step_per_gpu": rows, "rows_per_level": counts, "parallelism": f"dp{world}",
                       "dropout": args.dropout,
                       "l2": "per-step working set (activations of ~181 k packed rows, gathers from a ~94 GB pyramid store) far exceeds the 126 MB L2: no flush needed"},
            "host_enqueue_ms_per_step": host_enqueue_ms, "consistency": consistency, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "fast_mode": fast, "gpu_launches": launches, "clocks": clocks,
        }
        print(json.dumps(line), file=_OUT, flush=True)
    if os.environ.get("PATHS_B200_PROF_DUMP"):               # developer aid: per-call-site timings are written at engine destruction
        torch.cuda.synchronize()
        rt.engines.clear(); rt.last_engine = None; rt._pending = None
        del eng
        import gc; gc.collect()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
