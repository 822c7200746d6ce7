"""Worker of the 2-rank data-parallel tests (spawned by tests/test_gpu_ddp.py with torch.multiprocessing):
every rank trains on its half of a batch with DataParallelTrainer (SyncBatchNorm statistics + flat gradient all-reduce);
rank 0 also runs the whole batch alone and checks that the synchronised gradients are the single-GPU gradients."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)


def build(dev):
    import recipe
    import speech_anonymization_b200 as sab
    m = sab.ConvAutoencoder()
    w = recipe.make_weights(recipe.CONVAE_SHAPES, 8886)
    m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
    return m.to(dev).train()


def loss_of(m, x, noise, y):
    rec, logp = m(x, pool_noise=noise)
    B = x.shape[0]
    return 1.0 * torch.nn.MSELoss()(rec.view(B, -1), x.view(B, -1)) + 0.8 * torch.nn.NLLLoss()(logp, y)


def run(rank, world, port, out_path):
    import recipe
    import speech_anonymization_b200 as sab
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    Bl, T, accum = 4, 36, 2
    result = {}
    m = build(dev)
    dp = sab.DataParallelTrainer(m)
    dp.prepare()
    full = []
    for step in range(accum):                       # gradient accumulation: only the last micro-batch communicates
        xg = torch.from_numpy(recipe.make_feats(300 + step, world * Bl, T)).to(dev)
        ng = torch.from_numpy(recipe.make_noise(300 + step, world * Bl, 128)).to(dev)
        yg = torch.from_numpy((np.arange(world * Bl) + step) % 2).long().to(dev)
        full.append((xg, ng, yg))
        sl = slice(rank * Bl, (rank + 1) * Bl)
        ctx = dp.no_sync() if step < accum - 1 else torch.autograd.profiler.record_function("sync")
        with ctx:
            (loss_of(m, xg[sl].contiguous(), ng[sl].contiguous(), yg[sl]) / accum).backward()
            dp.finish_backward()
    synced = dp.grads.flat.clone()
    # every rank must now hold the same averaged gradient
    other = [torch.empty_like(synced) for _ in range(world)]
    dist.all_gather(other, synced)
    result["ranks_agree"] = float(max((o - synced).abs().max() for o in other))
    # global-batch MI inputs
    Xg, yg2 = sab.gather_for_mi(torch.full((3, 5), float(rank), device=dev), torch.full((3,), rank, device=dev))
    result["gather_ok"] = bool(Xg.shape == (3 * world, 5) and float(Xg[3 * (world - 1)][0]) == world - 1 and int(yg2[-1]) == world - 1)
    if rank == 0:
        ref = build(dev)                            # single GPU, whole batch, plain (local) BatchNorm statistics
        rg = sab.FlatGradients(ref)
        for xg, ng, yg in full:
            (loss_of(ref, xg, ng, yg) / accum).backward()
        scale = float(rg.flat.abs().max())
        result["max_abs_diff_over_max"] = float((synced - rg.flat).abs().max()) / scale
        worst = 0.0
        for n, v in dp.grads.views.items():
            r = rg.views[n]
            worst = max(worst, float((v - r).abs().max()) / max(float(r.abs().max()), 1e-4))
        result["worst_param_rel"] = worst
        bufs = dict(m.named_buffers())
        result["running_stats_rel"] = max(float((b - bufs[k]).abs().max() / b.abs().max().clamp_min(1e-6))
                                          for k, b in ref.named_buffers() if "running" in k)
        import json
        json.dump(result, open(out_path, "w"))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    run(int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
