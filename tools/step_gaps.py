"""Where is the compute stream idle inside a training step?  Profiles a few steps under CUPTI, takes the kernels of the
stream that carries the GEMMs and lists the largest gaps between consecutive kernels (with their neighbours), the total
idle time per step and the histogram of small gaps (launch-bound stretches)."""
import argparse, json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from fault_tolerant_llm_training_b200.config import ModelConfig
from fault_tolerant_llm_training_b200.parallel.env import init_dist
from fault_tolerant_llm_training_b200.trainer import make_trainer

p = argparse.ArgumentParser()
p.add_argument("--steps", type=int, default=3)
p.add_argument("--seq", type=int, default=2048)
a = p.parse_args()
env = init_dist("cuda")
cfg = ModelConfig(seq_len=a.seq, vocab_size=131072)
trainer = make_trainer(cfg, env, dtype=torch.bfloat16, lr=5e-5, warmup_steps=100, grad_max_norm=1.0, clip_mode="clip", seed=0,
                       parallel="auto")
g = torch.Generator().manual_seed(0)
batches = [(torch.randint(0, cfg.vocab_size, (1, a.seq), generator=g).pin_memory(),
            torch.randint(0, cfg.vocab_size, (1, a.seq), generator=g).pin_memory()) for _ in range(2)]
for i in range(3):
    last = trainer.step(*batches[i % 2])
last.item()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for i in range(a.steps):
        last = trainer.step(*batches[i % 2])
    torch.cuda.synchronize()
path = os.path.join(tempfile.mkdtemp(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") == "kernel"]
streams = {}
for e in ev:
    streams.setdefault(e["args"].get("stream"), []).append(e)
main = max(streams, key=lambda s: sum(e["dur"] for e in streams[s] if "gemm" in e["name"]))
ks = sorted(streams[main], key=lambda e: e["ts"])
span = ks[-1]["ts"] + ks[-1]["dur"] - ks[0]["ts"]
busy = sum(e["dur"] for e in ks)
gaps = [(ks[i + 1]["ts"] - (ks[i]["ts"] + ks[i]["dur"]), ks[i]["name"][:48], ks[i + 1]["name"][:48]) for i in range(len(ks) - 1)]
print(f"# Compute-stream idle time ({a.steps} steps, seq {a.seq})\n")
print(f"span {span / 1e3 / a.steps:.2f} ms/step, busy {busy / 1e3 / a.steps:.2f} ms/step, idle {(span - busy) / 1e3 / a.steps:.2f} ms/step, "
      f"{len(ks) // a.steps} kernels/step on this stream; other streams: "
      + ", ".join(f"{sum(e['dur'] for e in v) / 1e3 / a.steps:.2f} ms" for s, v in streams.items() if s != main) + "\n")
for lo, hi in ((0, 2), (2, 5), (5, 10), (10, 50), (50, 200), (200, 1e9)):
    sel = [x[0] for x in gaps if lo <= x[0] < hi]
    print(f"- gaps of {lo}-{hi if hi < 1e9 else 'inf'} us: {len(sel) // a.steps} per step, {sum(sel) / 1e3 / a.steps:.2f} ms per step")
print("\n| gap us | after | before |\n|---:|---|---|")
for gdur, x, y in sorted(gaps, key=lambda t: -t[0])[:24]:
    print(f"| {gdur:.0f} | `{x}` | `{y}` |")

# what the other streams were doing during the largest gap (the step boundary)
gi = max(range(len(ks) - 1), key=lambda i: ks[i + 1]["ts"] - (ks[i]["ts"] + ks[i]["dur"]))
g0, g1 = ks[gi]["ts"] + ks[gi]["dur"], ks[gi + 1]["ts"]
print(f"\nOther-stream kernels overlapping the largest gap ({(g1 - g0):.0f} us; times relative to its start):\n")
print("| start us | dur us | kernel |\n|---:|---:|---|")
for e in sorted(ev, key=lambda e: e["ts"]):
    if e["args"].get("stream") != main and e["ts"] + e["dur"] > g0 - 500 and e["ts"] < g1:
        print(f"| {e['ts'] - g0:.0f} | {e['dur']:.0f} | `{e['name'][:60]}` |")

for pat in ("dp_adamw_allgather", "dp_reduce_scatter"):
    d = [e["dur"] for e in sorted(ev, key=lambda e: e["ts"]) if pat in e["name"]]
    n = len(d) // a.steps
    print(f"\n`{pat}` launch durations of the last step (us): " + " ".join(f"{x:.0f}" for x in d[-n:]))

# timeline window in the forward phase: AdamW launches 10..12 of the last step and every kernel that overlaps them
ad = [e for e in sorted(ev, key=lambda e: e["ts"]) if "dp_adamw_allgather" in e["name"]]
n = len(ad) // a.steps
w0, w1 = ad[-n + 10]["ts"], ad[-n + 12]["ts"] + ad[-n + 12]["dur"]
print(f"\nTimeline of all streams while AdamW launches 10-12 of the last step run (us from the window start; C = compute stream):\n")
print("| stream | start | end | kernel |\n|---|---:|---:|---|")
for e in sorted(ev, key=lambda e: e["ts"]):
    if e["ts"] + e["dur"] > w0 and e["ts"] < w1:
        print(f"| {'C' if e['args'].get('stream') == main else 'comm'} | {e['ts'] - w0:.0f} | {e['ts'] + e['dur'] - w0:.0f} | `{e['name'][:50]}` |")
