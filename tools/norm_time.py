#!/usr/bin/env python
"""Isolated timing of the RMSNorm kernels at the training shape (rows 2048, dim 4096): CUDA events, L2 flushed between
iterations.  `python tools/norm_time.py [rows] [dim]` prints one JSON line; run under ncu for the per-kernel detail."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.ops import native as N  # noqa: E402


def timed(fn, flush, iters=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
    dim = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
    dev = torch.device("cuda:0")
    bf = lambda *s: torch.randn(*s, device=dev, dtype=torch.bfloat16)
    x, delta, w = bf(rows, dim), bf(rows, dim), bf(dim)
    h, y, dx, dw = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x), torch.empty_like(w)
    rstd = torch.empty(rows, device=dev, dtype=torch.float32)
    scratch = torch.empty(N.rmsnorm_partials(rows) * dim, device=dev, dtype=torch.float32)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {"rows": rows, "dim": dim}
    out["fwd_add_us"] = timed(lambda: N.rmsnorm_fwd(x, w, 1e-5, y, rstd, delta=delta, h_out=h), flush)
    out["fwd_plain_us"] = timed(lambda: N.rmsnorm_fwd(x, w, 1e-5, y, rstd), flush)
    out["bwd_res_us"] = timed(lambda: N.rmsnorm_bwd(y, x, w, rstd, dx, dw, scratch, dres=h), flush)
    mb = rows * dim * 2 / 1e6
    out["fwd_add_gbs"] = round(4 * mb / out["fwd_add_us"] * 1e3, 1)
    out["bwd_res_gbs"] = round(4 * mb / out["bwd_res_us"] * 1e3, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
