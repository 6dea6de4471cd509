"""Two-rank check of the data-parallel step on real GPUs (run under torchrun on a 2-GPU box; output kept in profiles/):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/dp_check.py

Every rank builds the same model and the same batch of BRCA-shaped slides.  Rank r runs DataParallelStep on its shard
(bucketed all-reduce behind the whole-step backward, NCCL); rank 0 also runs the whole batch alone.  The summed 2-rank
gradient must equal the 1-rank gradient of the concatenated batch, and the reduced loss the single-rank loss."""
import copy
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from paths_b200 import synth  # noqa: E402
from paths_b200.parallel import DataParallelStep, shard_slides  # noqa: E402
from tests.common import BASE_CONFIG, product_model, product_batch  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from paths_b200.config import Config
    data = copy.deepcopy(BASE_CONFIG); data["model_config"]["dropout"] = 0.0
    cfg = Config.from_dict(data, test_mode=True)
    n = 48
    slides = synth.make_brca_batch(range(n), cfg.num_levels, cfg.model_config.patch_embed_dim)
    shapes = {k: tuple(v.shape) for k, v in cfg.get_model().state_dict().items()}
    sd = synth.synth_state_dict(shapes, seed=3)
    out = {"world": world, "slides": n}
    for prec, pname in [(0, "fp32"), (1, "bf16x3")]:
        model = product_model(cfg, sd, dev)
        rt = model.runtime(); rt.set_precision(prec)
        mine = shard_slides(n, rank, world)
        batch, _ = product_batch(cfg, model, [slides[i] for i in mine], device=dev)
        dp = DataParallelStep(model); dp.time_exposed_comm(True)
        hz, loss = dp.step(batch, cfg.top_k_patches, global_batch=n)
        gl = dp.global_loss(loss)
        g_dp = rt.layout.gflat.clone()
        exposed = dp.last_exposed_comm()
        if rank == 0:
            model1 = product_model(cfg, sd, dev)
            rt1 = model1.runtime(); rt1.set_precision(prec)
            full, _ = product_batch(cfg, model1, slides, device=dev)
            from paths_b200.utils import inference_end2end
            hz1, loss1 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model1, cfg.base_power, full, cfg.task)
            loss1.backward(); torch.cuda.synchronize()
            g1 = rt1.layout.gflat
            scale = float(g1.abs().max())
            out[pname] = {"max_abs_diff_over_gmax": float((g_dp - g1).abs().max()) / scale,
                          "rel_l2": float((g_dp - g1).norm() / g1.norm()),
                          "loss_2rank": float(gl), "loss_1rank": float(loss1), "exposed_comm_ms": exposed,
                          "hazards_rank0_equal": bool(torch.equal(hz, hz1.detach()[:len(mine)]))}
        dist.barrier()
    if rank == 0:
        print(json.dumps(out, indent=1))
        ok = all(out[p]["max_abs_diff_over_gmax"] < (1e-6 if p == "fp32" else 1e-5) and
                 abs(out[p]["loss_2rank"] - out[p]["loss_1rank"]) < 1e-6 for p in ("fp32", "bf16x3"))
        print("DP_CHECK", "PASS" if ok else "FAIL")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
