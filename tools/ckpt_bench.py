"""SIGUSR1-to-durable-checkpoint time of the full 8.05 B model (BASELINE.json metric 2; the reference took 33.6 s
to torch.save 48 GB on its cluster, logs/output_444664.out:94-95).  Run alone for 1 GPU or under torchrun."""
import json
import os
import shutil
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.ckpt.layout import load_checkpoint  # noqa: E402
from fault_tolerant_llm_training_b200.config import ModelConfig  # noqa: E402
from fault_tolerant_llm_training_b200.optim.lr import build_lr_scheduler  # noqa: E402,F401
from fault_tolerant_llm_training_b200.parallel.env import init_dist  # noqa: E402
from fault_tolerant_llm_training_b200.trainer import make_trainer  # noqa: E402


def main():
    env = init_dist("cuda")
    out_dir = sys.argv[1] if len(sys.argv) > 1 else "/tmp/ftl_ckpt_bench"
    consolidate = "--consolidate" in sys.argv
    cfg = ModelConfig(seq_len=2048, vocab_size=131072)
    tr = make_trainer(cfg, env, dtype=torch.bfloat16, lr=5e-5, warmup_steps=100, keep_shards=True)
    g = torch.Generator().manual_seed(env.rank)
    tok = torch.randint(0, cfg.vocab_size, (1, 2049), generator=g)
    for _ in range(2):
        h = tr.step(tok[:, :-1].contiguous(), tok[:, 1:].contiguous())
    h.item()
    tr.drain()
    if env.is_main:
        shutil.rmtree(out_dir, ignore_errors=True)
    env.barrier()
    t0 = time.perf_counter()
    tr.save_fn()(out_dir, "bench", tr.model, tr.optimizer, tr.lr_scheduler, 2, data_state={"batches_consumed": 2},
                 rank=env.rank, world_size=env.world_size)
    durable = time.perf_counter() - t0
    res = {"world": env.world_size, "signal_to_durable_s": round(env.max_float(durable), 3),
           "per_rank": {k: round(v, 3) for k, v in tr.checkpointer.timings.items()}}
    if env.is_main:
        print(json.dumps(res), flush=True)
    if consolidate and env.is_main:
        # the box has room for one 48 GB copy on disk, so the consolidated file goes to tmpfs here
        from fault_tolerant_llm_training_b200.ckpt.sharded import consolidate as do_consolidate, shard_dir as sd
        t1 = time.perf_counter()
        out = do_consolidate(sd(out_dir, "bench"), "/dev/shm/ftl_checkpoint_bench.ckpt")
        res["consolidate_s"] = round(time.perf_counter() - t1, 3)
        res["consolidated_bytes"] = os.path.getsize(out)
        os.remove(out)
    # resume from the shards: disk -> device through the native reader
    env.barrier()
    t2 = time.perf_counter()
    from fault_tolerant_llm_training_b200.ckpt.sharded import load_sharded, shard_dir
    tr.load_state(load_sharded(shard_dir(out_dir, "bench")))
    res["load_from_shards_s"] = round(env.max_float(time.perf_counter() - t2), 3)
    if env.is_main:
        print(json.dumps(res), flush=True)
        shutil.rmtree(out_dir, ignore_errors=True)
    env.shutdown()


if __name__ == "__main__":
    main()
