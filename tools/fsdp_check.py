"""Multi-GPU check of the fully-sharded engine: W ranks x batch 1 (FSDP, gather fused into the QKV / gate-up GEMMs)
must track 1 rank x batch W, and a sharded checkpoint must round-trip.
Run: python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/fsdp_check.py"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.config import ModelConfig  # noqa: E402
from fault_tolerant_llm_training_b200.parallel.env import DistEnv, init_dist  # noqa: E402
from fault_tolerant_llm_training_b200.trainer import make_trainer  # noqa: E402

CFG = ModelConfig(dim=1024, n_layers=3, n_heads=8, n_kv_heads=4, multiple_of=1024, seq_len=256, vocab_size=8192)  # ffn 3072: row shards of 128 at W = 2, 4, 8
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
        solo.drain()
        ref_params = {n: p.detach().float().clone() for n, p in solo.model.named_parameters()}
        del solo
    env.barrier()
    tr = make_trainer(CFG, env, parallel="fsdp", keep_shards=True, **kw)
    losses = [tr.step(x[env.rank:env.rank + 1].contiguous(), y[env.rank:env.rank + 1].contiguous()).item() for x, y in data]
    tr.drain()
    full = tr.full_parameters()
    if env.rank == 0:
        num = den = 0.0
        for name, t in full.items():
            if name.endswith("attention.wqkv"):
                pre = name[:-len("wqkv")]
                ref = torch.cat([ref_params[pre + k + ".weight"] for k in ("wq", "wk", "wv")])
            else:
                ref = ref_params[name]
            num += float((t.float().reshape(-1) - ref.reshape(-1)).pow(2).sum())
            den += float(ref.pow(2).sum())
        rel = (num / den) ** 0.5
        out = {"world": W, "fused_gather_gemms": tr.engine.fused_gather is not None,
               "losses_fsdp": [round(l, 4) for l in losses], "losses_1gpu_batchW": [round(l, 4) for l in ref_losses],
               "max_loss_diff": max(abs(a - b) for a, b in zip(losses, ref_losses)), "param_rel_diff": rel,
               "own_numel": tr.own_numel, "full_numel": tr.flat.numel, "committed": tr.committed_updates()}
        out["ok"] = bool(out["max_loss_diff"] < 5e-2 and rel < 2e-2 and out["committed"] == STEPS)
        print(json.dumps(out), flush=True)
    # sharded checkpoint -> consolidated file -> reload into a fresh FSDP trainer -> identical next loss
    tr.save_fn()("/tmp/ftl_fsdp_check", "fsdp", tr.model, tr.optimizer, tr.lr_scheduler, STEPS, rank=env.rank, world_size=W)
    path = tr.finalize_fn()()
    env.barrier()
    from fault_tolerant_llm_training_b200.ckpt.layout import load_checkpoint
    from fault_tolerant_llm_training_b200.ckpt.sharded import load_sharded, shard_dir
    x, y = data[0]
    want = tr.step(x[env.rank:env.rank + 1].contiguous(), y[env.rank:env.rank + 1].contiguous()).item()
    got = []
    for ck in (load_sharded(shard_dir("/tmp/ftl_fsdp_check", "fsdp")), load_checkpoint("/tmp/ftl_fsdp_check", "fsdp")):
        t2 = make_trainer(CFG, env, parallel="fsdp", **dict(kw, seed=99))
        t2.load_state(ck)
        got.append(t2.step(x[env.rank:env.rank + 1].contiguous(), y[env.rank:env.rank + 1].contiguous()).item())
        t2.drain()
        del t2
    if env.rank == 0:
        print(json.dumps({"resume_from_shards_equal": got[0] == want, "resume_from_consolidated_equal": got[1] == want,
                          "want": want, "got": got}), flush=True)
    env.shutdown()


if __name__ == "__main__":
    main()
