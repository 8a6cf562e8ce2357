"""Multi-GPU check of the DP engine: W ranks x batch 1 must track 1 rank x batch W.
Run: python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/dp_check.py
Rank 0 also runs the single-process batch-W reference on its own GPU first and compares losses / parameters."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.config import ModelConfig  # noqa: E402
from fault_tolerant_llm_training_b200.parallel.env import DistEnv, init_dist  # noqa: E402
from fault_tolerant_llm_training_b200.trainer import make_trainer  # noqa: E402

CFG = ModelConfig(dim=512, n_layers=4, n_heads=8, n_kv_heads=2, multiple_of=256, seq_len=256, vocab_size=8192)
STEPS = 6


def batches(world):
    g = torch.Generator().manual_seed(5)
    out = []
    for _ in range(STEPS):
        tok = torch.randint(0, CFG.vocab_size, (world, CFG.seq_len + 1), generator=g)
        lab = tok[:, 1:].clone()
        lab[:, :7] = -100
        out.append((tok[:, :-1].contiguous(), lab))
    return out


def main():
    env = init_dist("cuda")
    W = env.world_size
    data = batches(W)
    kw = dict(dtype=torch.bfloat16, lr=1e-3, warmup_steps=2, grad_max_norm=1.0, seed=3)
    ref_losses, ref_params = None, None
    if env.rank == 0:
        solo = make_trainer(CFG, DistEnv(device=env.device), parallel="none", **kw)
        ref_losses = [solo.step(x, y).item() for x, y in data]
        ref_params = torch.cat([p.detach().float().reshape(-1) for p in solo.model.parameters()])
        del solo
    env.barrier()
    tr = make_trainer(CFG, env, parallel="dp", **kw)
    losses = []
    for x, y in data:
        h = tr.step(x[env.rank:env.rank + 1].contiguous(), y[env.rank:env.rank + 1].contiguous())
        losses.append(h.item())
    tr.drain()
    params = torch.cat([p.detach().float().reshape(-1) for p in tr.model.parameters()])
    # every rank must hold identical parameters after the fused all-gather
    digest = params.double().sum().item()
    digests = env.gather_objects(digest)
    # timing of the collective part on a bigger flat buffer is left to bench.py; here: correctness
    if env.rank == 0:
        rel = float((params - ref_params).norm() / ref_params.norm())
        out = {"world": W, "losses_dp": [round(l, 4) for l in losses], "losses_1gpu_batchW": [round(l, 4) for l in ref_losses],
               "max_loss_diff": max(abs(a - b) for a, b in zip(losses, ref_losses)), "param_rel_diff": rel,
               "ranks_identical": len(set(digests)) == 1, "grad_norm": tr.last_grad_norm}
        out["ok"] = bool(out["ranks_identical"] and out["max_loss_diff"] < 5e-2 and rel < 2e-2)
        print(json.dumps(out), flush=True)
    # checkpoint path: gather moments, save on rank 0, reload everywhere
    save = tr.save_fn()
    save("/tmp/ftl_dp_check", "dp", tr.model, tr.optimizer, tr.lr_scheduler, STEPS, rank=env.rank, world_size=W)
    path = tr.finalize_fn()()       # rank 0 consolidates the W shard files into the reference layout
    env.barrier()
    if env.rank == 0:
        ck = torch.load(path, weights_only=False)
        ref_sd = {k: v.float() for k, v in tr.model.state_dict().items()}
        assert all(torch.equal(ck["model"][k].float(), ref_sd[k].cpu()) for k in ref_sd)
        nz = sum(int(s["exp_avg"].abs().sum() > 0) for s in ck["optimizer"]["state"].values())
        print(json.dumps({"ckpt_tensors_with_moments": nz, "of": len(ck["optimizer"]["state"])}), flush=True)
    env.shutdown()


if __name__ == "__main__":
    main()
