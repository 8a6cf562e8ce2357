"""Per-kernel device time of whole training steps in their real schedule (torch profiler / CUPTI: warm caches, real
clocks, overlapped streams) — the complement of the serialised ncu launch list in profiles/step_launches_*.md.
Usage: python tools/step_kernels_time.py [--steps 3] [--seq 2048] > profiles/step_kernels_<tag>.md"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from fault_tolerant_llm_training_b200.config import ModelConfig
from fault_tolerant_llm_training_b200.parallel.env import init_dist
from fault_tolerant_llm_training_b200.trainer import make_trainer

p = argparse.ArgumentParser()
p.add_argument("--steps", type=int, default=3)
p.add_argument("--seq", type=int, default=2048)
p.add_argument("--layers", type=int, default=0)
p.add_argument("--parallel", default="auto")
a = p.parse_args()
env = init_dist("cuda")
cfg = ModelConfig(seq_len=a.seq, vocab_size=131072)
if a.layers:
    cfg.n_layers = a.layers
trainer = make_trainer(cfg, env, dtype=torch.bfloat16, lr=5e-5, warmup_steps=100, grad_max_norm=1.0, clip_mode="clip", seed=0,
                       parallel=a.parallel)
g = torch.Generator().manual_seed(0)
batches = [(torch.randint(0, cfg.vocab_size, (1, a.seq), generator=g).pin_memory(),
            torch.randint(0, cfg.vocab_size, (1, a.seq), generator=g).pin_memory()) for _ in range(2)]
for i in range(3):
    last = trainer.step(*batches[i % 2])
last.item()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    e0.record()
    for i in range(a.steps):
        last = trainer.step(*batches[i % 2])
    e1.record()
    torch.cuda.synchronize()
if not env.is_main:   # every rank profiles (so that no rank runs ahead); rank 0 reports
    env.barrier()
    sys.exit(0)
env.barrier()
rows = [(e.key, e.device_time_total / a.steps / 1e3, e.count // a.steps) for e in prof.key_averages() if e.device_time_total > 0]
total = sum(r[1] for r in rows)
print(f"# Kernel time per training step on rank 0 of {env.world_size} ({trainer.parallel}; profiled under CUPTI, {a.steps} steps, seq {a.seq}): step {e0.elapsed_time(e1) / a.steps:.2f} ms, "
      f"sum of kernel durations {total:.2f} ms\n")
print("| ms / step | launches | share | kernel |\n|---:|---:|---:|---|")
for k, t, n in sorted(rows, key=lambda r: -r[1])[:32]:
    print(f"| {t:.3f} | {n} | {100 * t / total:.1f}% | `{k[:70]}` |")
