"""Cross-GPU race stress of the in-kernel data plane (run by tests/test_t5_multigpu.py):

  1. `--epochs` barrier epochs: before each one every rank waits a random time on the device, publishes the epoch
     number in its own peer-mapped buffer, passes the device barrier and reads every peer's value back through
     NVLink — a stale value means the barrier let a rank through before a peer's earlier write was visible;
  2. `--gemms` gather-fused QKV GEMMs (csrc/gemm_tc.cu): the owners rewrite their weight shards every iteration, random
     delays in front, result compared bit for bit with the same GEMM on a locally assembled weight;
  3. the barrier's time budget: one rank stays away -> the others come back with status 1, and an update launched with
     that status leaves parameters and moments untouched;
  4. the abort word: one rank raises it while the others wait -> status 2;
  5. the row-sparse reduce-scatter of the embedding gradient against the dense fixed-order one: bit-equal shards over
     three steps with changing token lists (rows of earlier steps must end up zero again).

    python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/comm_stress.py
"""
import argparse
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fault_tolerant_llm_training_b200.ops import native as N  # noqa: E402
from fault_tolerant_llm_training_b200.parallel.env import init_dist  # noqa: E402
from fault_tolerant_llm_training_b200.parallel.symm import SymmetricBuffers, _RawCuda  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=10000)
    ap.add_argument("--gemms", type=int, default=200)
    a = ap.parse_args()
    env = init_dist("cuda")
    W, r, dev = env.world_size, env.rank, env.device
    K, rows_per_rank = 512, 128
    Nq = rows_per_rank * W
    flag_words = N.lib().ftl_comm_flag_words()
    symm = SymmetricBuffers(env, {"flags": flag_words * 4, "val": 64 * 4, "own": rows_per_rank * K * 2,
                                  "flags2": flag_words * 4, "flags3": flag_words * 4})
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    rng = torch.Generator().manual_seed(1000 + r)
    peers_val = [torch.as_tensor(_RawCuda(p, 64 * 4), device=dev).view(torch.int32) for p in symm.peer_ptr["val"]]
    mine = symm.tensor("val", torch.int32)

    # ---- 1. barrier epochs -----------------------------------------------------------------------------------
    bad = torch.zeros(1, dtype=torch.int64, device=dev)
    delays = torch.randint(0, 20000, (a.epochs,), generator=rng).tolist()
    epoch = 0
    t0 = time.perf_counter()
    for e in range(1, a.epochs + 1):
        if delays[e - 1] > 15000:
            torch.cuda._sleep(delays[e - 1] * 20)           # up to ~0.25 ms of skew on this rank
        mine[r:r + 1].fill_(e)
        epoch += 1
        N.barrier(symm.table("flags"), r, W, epoch, status, 20000)
        for q in range(W):
            bad += (peers_val[q][q:q + 1] != e).sum()
        epoch += 1
        N.barrier(symm.table("flags"), r, W, epoch, status, 20000)   # nobody overwrites before everyone has read
    torch.cuda.synchronize()
    barrier_s = time.perf_counter() - t0
    barrier_ok = int(bad.item()) == 0 and int(status.item()) == 0

    # ---- 2. gather-fused GEMM under rewritten shards ---------------------------------------------------------------
    B, S, H, KV, hd = 1, 256, max(W // 2, 1) * 2, 1, 64
    # N = (H + 2 KV) * hd must equal rows_per_rank * W: choose heads accordingly
    H = (Nq // hd) - 2 * KV
    rot = (H + KV) * hd
    own = symm.tensor("own", torch.bfloat16).view(rows_per_rank, K)
    x = (torch.randn(B * S, K, generator=torch.Generator().manual_seed(7)) * 0.5).to(dev, torch.bfloat16)
    half = hd // 2
    ang = torch.arange(S)[:, None].float() * (1.0 / (10000 ** (torch.arange(half).float() / half)))[None, :]
    cos, sin = ang.cos().to(dev).contiguous(), ang.sin().to(dev).contiguous()
    stage = torch.zeros(Nq, K, dtype=torch.bfloat16, device=dev)
    full = torch.empty(Nq, K, dtype=torch.bfloat16, device=dev)
    out, ref = (torch.empty(B * S, Nq, dtype=torch.bfloat16, device=dev) for _ in range(2))
    scratch = torch.zeros(1 + 2 * (Nq // 128) + 8, dtype=torch.int32, device=dev)
    gemm_bad = 0
    gd = torch.randint(0, 20000, (a.gemms,), generator=rng).tolist()
    for it in range(a.gemms):
        g = torch.Generator().manual_seed(50000 + it)             # every rank derives the same full weight
        wfull = (torch.randn(Nq, K, generator=g) * 0.1).to(torch.bfloat16)
        full.copy_(wfull)
        own.copy_(full[r * rows_per_rank:(r + 1) * rows_per_rank])
        epoch += 1
        N.barrier(symm.table("flags"), r, W, epoch, status, 20000)     # all shards rewritten
        if gd[it] > 12000:
            torch.cuda._sleep(gd[it] * 30)
        stage.zero_()
        N.gemm_qkv_rope_gather(x, stage, out, cos, sin, S, hd, rot, symm.table("own"), 0, rows_per_rank, r, W, scratch)
        N.gemm_qkv_rope(x, full, ref, cos, sin, S, hd, rot)
        gemm_bad += int((~torch.eq(out, ref)).sum().item()) + int((~torch.eq(stage, full)).sum().item())
        epoch += 1
        N.barrier(symm.table("flags"), r, W, epoch, status, 20000)     # peers have finished reading my shard
    torch.cuda.synchronize()
    gemm_ok = gemm_bad == 0 and int(status.item()) == 0

    # ---- 3. barrier time budget: rank 0 stays away ---------------------------------------------------------------
    st2 = torch.zeros(1, dtype=torch.int32, device=dev)
    env.barrier()
    if r != 0:
        N.barrier(symm.table("flags2"), r, W, 1, st2, 300)
    torch.cuda.synchronize()
    p = torch.randn(4096, device=dev).to(torch.bfloat16)
    gbuf, m, v = torch.randn_like(p.float()).to(torch.bfloat16), torch.zeros_like(p), torch.zeros_like(p)
    p0 = p.clone()
    hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 1e-2, 0.1, 0.0316, 1.0], device=dev)
    stats = torch.tensor([1.0, 1.0, 1.0, 1.0], device=dev)
    N.adamw_shard(p, gbuf, m, v, hyper, stats, st2 if r != 0 else torch.ones(1, dtype=torch.int32, device=dev))
    torch.cuda.synchronize()
    untouched = bool(torch.equal(p, p0) and int(m.abs().sum().item()) == 0)
    timeout_status = int(st2.item())
    env.barrier()

    # ---- 4. abort word --------------------------------------------------------------------------------------
    st3 = torch.zeros(1, dtype=torch.int32, device=dev)
    if r != 0:
        N.barrier(symm.table("flags3"), r, W, 1, st3, 20000)
    else:
        time.sleep(0.5)
        N.raise_abort(symm.table("flags3"), W)
    torch.cuda.synchronize()
    abort_status = int(st3.item())
    # ---- 5. sparse vs dense embedding reduce-scatter ----------------------------------------------------------------
    V, D, Tn, cap = 4096, 256, 512, 1024
    es = SymmetricBuffers(env, {"gA": V * D * 2, "gB": V * D * 2, "toks": 3 * W * cap * 8, "fl": flag_words * 4})
    gA, gB = es.tensor("gA", torch.bfloat16).view(V, D), es.tensor("gB", torch.bfloat16).view(V, D)
    toks = es.tensor("toks", torch.int64)
    st5 = torch.zeros(1, dtype=torch.int32, device=dev)
    rows = V // W
    lo, hi = r * rows * D, (r + 1) * rows * D
    partsA = torch.zeros(N.dp_partials(0), device=dev)
    partsB = torch.zeros(W * Tn, device=dev)
    sparse_ok, ep, prev_tok = True, 0, None
    for step in range(3):
        g5 = torch.Generator().manual_seed(900 + 10 * step + r)
        tok = torch.randint(0, V, (Tn,), generator=g5)
        tok[::5] = int(tok[0])                              # repeats inside a rank
        tok[1::7] = (step * 37) % V                         # an id every rank shares
        tok = tok.to(dev)
        dout = (torch.randn(Tn, D, generator=g5) * 0.3).to(dev, torch.bfloat16)
        gA.zero_()
        N.embedding_bwd_direct(tok, None, dout, gA)
        N.embedding_bwd_direct(tok, prev_tok, dout, gB)     # gB is never memset: only the previous rows are cleared
        slot = step % 3
        N.emb_tokens_push(tok, es.table("toks"), slot * W * cap, r, W, cap)
        ep += 1
        N.barrier(es.table("fl"), r, W, ep, st5, 20000)
        N.dp_reduce_scatter(es.table("gA"), r, W, lo, hi - lo, partsA, st5)
        cur = toks[slot * W * cap:]
        prv = toks[((step - 1) % 3) * W * cap:] if step else None
        N.emb_sparse_reduce(es.table("gB"), r, W, cur, prv, Tn, cap, 0, D, V, lo, hi, partsB, st5)
        torch.cuda.synchronize()
        dense_shard, sparse_shard = gA.view(-1)[lo:hi], gB.view(-1)[lo:hi]
        sparse_ok &= bool(torch.equal(dense_shard, sparse_shard))
        sparse_ok &= bool(torch.allclose(partsA.sum(), partsB.sum(), rtol=1e-4, atol=1e-4))
        ep += 1
        N.barrier(es.table("fl"), r, W, ep, st5, 20000)      # peers have finished reading my rows before the next step
        torch.cuda.synchronize()
        prev_tok = tok
    res = env.gather_objects({"barrier_ok": barrier_ok, "gemm_ok": gemm_ok, "timeout": timeout_status,
                              "untouched": untouched, "abort": abort_status, "sparse_ok": sparse_ok})
    if r == 0:
        others = res[1:]
        print(json.dumps({"world": W, "epochs": a.epochs, "barrier_s": round(barrier_s, 2),
                          "barrier_epochs_ok": all(x["barrier_ok"] for x in res),
                          "gather_gemm_ok": all(x["gemm_ok"] for x in res), "gemms": a.gemms,
                          "timeout_status": max(x["timeout"] for x in others),
                          "untouched_after_timeout": all(x["untouched"] for x in res),
                          "abort_status": min(x["abort"] for x in others),
                          "sparse_embedding_reduce_ok": all(x["sparse_ok"] for x in res)}), flush=True)
    env.shutdown()


if __name__ == "__main__":
    main()
