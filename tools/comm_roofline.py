"""Roofline of the fused compute+collective kernels (BASELINE.json: "each fused path is reported as achieved fraction of its
roofline, the slower of compute-at-peak and bytes-over-NVLink").

    python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/comm_roofline.py [--mb 256] [--nvls]

Every kernel is timed with CUDA events on its launching stream (max over ranks, median of `--iters` after warm-up, a
device barrier in front so that all ranks start together).  Denominators: NVLink 770 GB/s per direction per GPU (the
measured peer-copy figure of B200_PROFILING.md; 900 nominal), HBM `hbm_gbs` of MEASURED_PEAKS.json, tensor cores
`bf16_tflops` (burst: the kernels are timed in isolation).  Rank 0 prints one JSON line per kernel.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from fault_tolerant_llm_training_b200.ops import native as N  # noqa: E402
from fault_tolerant_llm_training_b200.parallel.env import init_dist  # noqa: E402
from fault_tolerant_llm_training_b200.parallel.symm import make_symmetric  # noqa: E402

NVLINK_GBS = 770.0


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), float(p["bf16_tflops"]), "measured"
    except (OSError, KeyError, ValueError):
        return 6650.0, 1590.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=256, help="bytes of ONE rank's shard, in MiB")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--nvls", action="store_true")
    a = ap.parse_args()
    env = init_dist("cuda")
    W, r, dev = env.world_size, env.rank, env.device
    hbm, tflops, src = peaks()
    shard = a.mb * (1 << 20) // 2          # elements per rank
    total = shard * W
    flag_words = N.lib().ftl_comm_flag_words()
    use_mc = a.nvls and W > 1
    symm = make_symmetric(env, {"p": total * 2, "g": total * 2, "flags": flag_words * 4, "slots": 64},
                          multicast=("p", "g") if use_mc else ())
    p, g = symm.tensor("p", torch.bfloat16), symm.tensor("g", torch.bfloat16)
    g.copy_((torch.randn(total, device=dev) * 0.01).to(torch.bfloat16))
    p.copy_((torch.randn(total, device=dev, generator=torch.Generator(device=dev).manual_seed(1)) * 0.02).to(torch.bfloat16))
    m = torch.zeros(shard, dtype=torch.bfloat16, device=dev)
    v = torch.zeros(shard, dtype=torch.bfloat16, device=dev)
    mfull, vfull = torch.zeros(total, dtype=torch.bfloat16, device=dev), torch.zeros(total, dtype=torch.bfloat16, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    parts = torch.zeros(N.dp_partials(0), device=dev)
    hyper = torch.tensor([1e-4, 0.9, 0.999, 1e-8, 1e-2, 0.1, 0.0316, 1.0], device=dev)
    stats = torch.tensor([1.0, 1.0, 1.0, 1.0], device=dev)
    lo = r * shard
    epoch = [0]

    def bar():
        if W > 1:
            epoch[0] += 1
            N.barrier(symm.table("flags"), r, W, epoch[0], status, 60000)

    def timed(fn):
        ts = []
        for i in range(a.iters + 3):
            bar()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            if i >= 3:
                ts.append(env.max_float(e0.elapsed_time(e1)))
        ts.sort()
        return ts[len(ts) // 2]

    def report(name, ms, nvlink_bytes, hbm_bytes, flops=0.0, note=""):
        t_nv = nvlink_bytes / (NVLINK_GBS * 1e9) * 1e3
        t_hbm = hbm_bytes / (hbm * 1e9) * 1e3
        t_fl = flops / (tflops * 1e12) * 1e3
        roof = max(t_nv, t_hbm, t_fl)
        bound = "nvlink" if roof == t_nv else ("hbm" if roof == t_hbm else "tensor")
        if r == 0:
            print(json.dumps({"kernel": name, "world": W, "shard_mib": a.mb, "ms": round(ms, 3),
                              "nvlink_gbs_bottleneck_dir": round(nvlink_bytes / ms / 1e6, 1),
                              "hbm_gbs": round(hbm_bytes / ms / 1e6, 1), "roofline_ms": round(roof, 3), "bound": bound,
                              "fraction_of_roofline": round(roof / ms, 3), "denominators": src, "note": note}), flush=True)

    sb = shard * 2   # bytes of one shard
    # reduce-scatter ⊕ norm²: W-1 remote shards in, every local copy of every shard is read once by someone
    ms = timed(lambda: N.dp_reduce_scatter(symm.table("g"), r, W, lo, shard, parts, status))
    report(f"dp_reduce_scatter_kernel<{W}> (peer loads)", ms, (W - 1) * sb, W * sb + (sb if W > 1 else 0),
           note="inbound (W-1) shards per GPU; HBM: own buffer read once by the W readers + own shard written")
    # clip ⊕ AdamW ⊕ all-gather: W-1 remote copies out
    ms = timed(lambda: N.dp_adamw_allgather(symm.table("p"), g, mfull, vfull, r, W, lo, shard, hyper, stats, status))
    report(f"dp_adamw_allgather_kernel<{W}> (peer stores)", ms, (W - 1) * sb, 4 * sb + 3 * sb + (W - 1) * sb,
           note="outbound (W-1) shards per GPU; HBM: p,g,m,v read, p,m,v written, W-1 incoming shards written")
    if use_mc:
        ms = timed(lambda: N.nvls_reduce_scatter(symm.mc_ptr["g"], g, lo, shard, parts, status))
        report("nvls_reduce_scatter_kernel (multimem.ld_reduce)", ms, (W - 1) * sb, W * sb + sb,
               note="outbound (W-1) shards per GPU (the switch pulls them), inbound ONE reduced shard")
        ms = timed(lambda: N.nvls_adamw_allgather(symm.mc_ptr["p"], p, g, mfull, vfull, lo, shard, hyper, stats, status))
        report("nvls_adamw_allgather_kernel (multimem.st)", ms, (W - 1) * sb, 4 * sb + 2 * sb + W * sb,
               note="outbound ONE shard per GPU, inbound W-1 (the switch replicates)")
    # plain parameter all-gather (FSDP)
    dst = torch.empty(total, dtype=torch.bfloat16, device=dev)
    own = make_symmetric(env, {"own": sb})
    own.tensor("own", torch.bfloat16).copy_(p[lo:lo + shard])
    ms = timed(lambda: N.fsdp_gather(own.table("own"), 0, dst, shard, r, W, status))
    report(f"fsdp_gather_kernel<{W}>", ms, (W - 1) * sb, W * sb + W * sb, note="inbound (W-1) shards; HBM: read + write of every shard")
    # gather-fused gate|up GEMM at the model's shape (x [2048, 4096] x w1|w3 [14336, 4096]) vs the same GEMM on local weights
    T, d, F = 2048, 4096, 14336
    if F % (128 * W) == 0 and W > 1:
        wsh = make_symmetric(env, {"own": 2 * (F // W) * d * 2})
        wo = wsh.tensor("own", torch.bfloat16)
        wo.copy_((torch.randn(wo.numel(), device=dev) * 0.02).to(torch.bfloat16))
        x = (torch.randn(T, d, device=dev) * 0.5).to(torch.bfloat16)
        w1, w3 = torch.empty(F, d, dtype=torch.bfloat16, device=dev), torch.empty(F, d, dtype=torch.bfloat16, device=dev)
        act, gate, up = (torch.empty(T, F, dtype=torch.bfloat16, device=dev) for _ in range(3))
        scratch = torch.zeros(1 + 2 * (F // 128) + 8, dtype=torch.int32, device=dev)
        ms_f = timed(lambda: N.gemm_swiglu_gather(x, w1, w3, act, gate, up, wsh.table("own"), 0, (F // W) * d, F // W, r, W, scratch))
        # the same launch with every "owner" pointing at this rank's own shard: the mechanism without NVLink
        local_tab = torch.tensor([wsh.local_ptr["own"]] * W, dtype=torch.int64, device=dev)
        ms_loc = timed(lambda: N.gemm_swiglu_gather(x, w1, w3, act, gate, up, local_tab, 0, (F // W) * d, F // W, r, W, scratch))
        ms_l = timed(lambda: N.gemm_swiglu(x, w1, w3, act, gate, up))
        fl = 2.0 * T * d * 2 * F
        nv = (W - 1) * 2 * (F // W) * d * 2
        report(f"gemm_swiglu_gather (all-gather fused into the gate|up GEMM, W={W})", ms_f, nv, 2 * F * d * 2 * 2 + 3 * T * F * 2, fl,
               note=f"same GEMM on local weights: {ms_l:.3f} ms; fused gather with all sources local: {ms_loc:.3f} ms; "
                    f"FTL_GATHER_WARPS={os.environ.get('FTL_GATHER_WARPS', '4')}; a separate gather of these weights would add "
                    f"{nv / (NVLINK_GBS * 1e9) * 1e3:.3f} ms at 770 GB/s")
    env.shutdown()


if __name__ == "__main__":
    main()
