"""Per-tile timeline of the attention forward kernel (clock64 probes inside CTA 0, first work item).

    python tools/attn_probe.py build     # here: nvcc -DFTL_ATTN_DEBUG csrc/attn_tc.cu -> build/libattn_dbg.so
    python tools/attn_probe.py run       # on the GPU box: runs the kernel once and prints the deltas

Slots (cycles relative to the first probe of the item):
  math thread (warp 2 lane 0): 0 loop top, 1 S ready, 2 scores in registers, 3 row max exchanged,
                               4 P computed, 5 P published, 6 O(t) ready (fold), 7 fold done
  MMA thread:                  8 K/V landed, 9 QK^T issued, 10 P(t) seen, 11 O buffer free, 12 P V issued
"""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "build", "libattn_dbg.so")


def build():
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "--expt-relaxed-constexpr", "-DFTL_ATTN_DEBUG", *os.environ.get("FTL_PROBE_DEFS", "").split(), "-Xcompiler", "-fPIC", "-shared", "-cudart", "static",
           "-I", os.path.join(ROOT, "csrc"), "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "csrc", "attn_tc.cu"), "-o", LIB, "-lcuda"]
    subprocess.check_call(cmd)
    print("built", LIB)


def run():
    import torch
    B, S, H, KV, hd = 1, 2048, 32, 8, 128
    if len(sys.argv) > 3:
        B, S = int(sys.argv[2]), int(sys.argv[3])
    lib = ctypes.CDLL(LIB)
    torch.manual_seed(0)
    qkv = (torch.randn(B * S, (H + 2 * KV) * hd, device="cuda") * 0.7).to(torch.bfloat16)
    out = torch.empty(B * S, H * hd, device="cuda", dtype=torch.bfloat16)
    lse = torch.empty(B * H * S, device="cuda")
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    lib.ftl_attn_fwd.argtypes = [vp, i64, vp, i64, vp, i32, i32, i32, i32, i32, vp]
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        rc = lib.ftl_attn_fwd(qkv.data_ptr(), qkv.stride(0), out.data_ptr(), out.stride(0), lse.data_ptr(), B, S, H, KV, hd, st)
        assert rc == 0, rc
    torch.cuda.synchronize()
    host = (ctypes.c_longlong * (64 * 16))()
    lib.ftl_attn_dbg_read.argtypes = [vp]
    assert lib.ftl_attn_dbg_read(host) == 0
    n = S // 128
    t0 = host[0]
    names = ["top", "Srdy", "Sreg", "max", "Pcmp", "Ppub", "Ordy", "fold", "KVld", "Siss", "Pseen", "Ofree", "PViss"]
    if S % 256 == 0 and os.environ.get("FTL_ATTN_FWD") != "1":      # attn_fwd2_kernel: softmax A, softmax B, MMA warp (tile A)
        names = ["A.top", "A.Srdy", "A.max", "A.exp", "A.pub", "B.Srdy", "B.max", "B.exp", "B.pub", "-", "M.Pseen", "M.iss", "-"]
    print("tile " + " ".join(f"{x:>6}" for x in names))
    n = max(t for t in range(64) if host[t * 16]) + 1
    for t in range(n):
        print(f"{t:4d} " + " ".join(f"{host[t * 16 + k] - t0 if host[t * 16 + k] else -1:6d}" for k in range(13)))
    per = (host[(n - 1) * 16 + 0] - host[2 * 16 + 0]) / (n - 3)
    print("cycles per tile (steady state):", per)


def run_bwd():
    """Timeline of the first work unit of CTA 0 in attn_bwd_dkv_kernel (per 128-key x 64-query inner iteration)."""
    import torch
    B, S, H, KV, hd = 1, 2048, 32, 8, 128
    lib = ctypes.CDLL(LIB)
    torch.manual_seed(0)
    W = (H + 2 * KV) * hd
    qkv = (torch.randn(B * S, W, device="cuda") * 0.7).to(torch.bfloat16)
    dout = (torch.randn(B * S, H * hd, device="cuda") * 0.5).to(torch.bfloat16)
    out = torch.empty(B * S, H * hd, device="cuda", dtype=torch.bfloat16)
    lse = torch.empty(B * H * S, device="cuda"); delta = torch.empty_like(lse)
    dqkv = torch.zeros_like(qkv)
    vp, i64, i32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int
    st = torch.cuda.current_stream().cuda_stream
    lib.ftl_attn_fwd.argtypes = [vp, i64, vp, i64, vp, i32, i32, i32, i32, i32, vp]
    assert lib.ftl_attn_fwd(qkv.data_ptr(), qkv.stride(0), out.data_ptr(), out.stride(0), lse.data_ptr(), B, S, H, KV, hd, st) == 0
    lib.ftl_attn_bwd_scratch_bytes.restype = i64
    lib.ftl_attn_bwd_scratch_bytes.argtypes = [i32] * 5
    scratch = torch.empty(max(lib.ftl_attn_bwd_scratch_bytes(B, S, H, KV, hd) // 4, 1), device="cuda")
    lib.ftl_attn_bwd.argtypes = [vp, i64, vp, vp, i64, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp]
    for _ in range(2):
        rc = lib.ftl_attn_bwd(qkv.data_ptr(), qkv.stride(0), out.data_ptr(), dout.data_ptr(), out.stride(0), lse.data_ptr(),
                              delta.data_ptr(), dqkv.data_ptr(), None, None, scratch.data_ptr(), B, S, H, KV, hd, st)
        assert rc == 0, rc
    torch.cuda.synchronize()
    host = (ctypes.c_longlong * (64 * 16))()
    lib.ftl_attn_dbg_read.argtypes = [vp]
    assert lib.ftl_attn_dbg_read(host) == 0
    which = sys.argv[2] if len(sys.argv) > 2 else "dkv"
    names = {"dkv": ["top", "bar", "Srdy", "Tfree", "ld", "math", "pub", "-", "M.Qld", "M.top", "M.Tseen", "M.iss"],
             # dq: one row per work ITEM of CTA 0: item start, first S ready, key-tile loop done, next rows requested,
             # last MMA done, next rows published, epilogue done; last column = key tiles of the item
             "dq": ["start", "S0rdy", "loopend", "loaded", "accfull", "publ", "epi", "tiles", "e.ld0", "e.st0", "e.ld1", "e.st1"]}[which]
    t0 = min(host[k] for k in range(16) if host[k])
    n = max(t for t in range(64) if host[t * 16]) + 1
    print("iter " + " ".join(f"{x:>7}" for x in names))
    for t in range(n):
        print(f"{t:4d} " + " ".join(f"{(host[t * 16 + k] - t0 if names[k] != 'tiles' else host[t * 16 + k]) if host[t * 16 + k] else -1:7d}" for k in range(len(names))))
    if n > 6 and which == "dkv":
        print("cycles per iteration (steady state):", (host[(n - 2) * 16] - host[2 * 16]) / (n - 4))


if __name__ == "__main__":
    {"build": build, "run": run, "bwd": run_bwd}[sys.argv[1]]()
