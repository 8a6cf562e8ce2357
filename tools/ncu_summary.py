"""Markdown summary of an .ncu-rep (run here, no GPU needed): python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x.md"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "time"), ("sm__cycles_elapsed.avg", "SM cycles elapsed"), ("smsp__cycles_active.avg", "SMSP cycles active"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU (MUFU) pipe %"),
        ("sm__issue_active.avg.pct_of_peak_sustained_elapsed", "issue slots busy %"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed", "shared-memory wavefronts %"),
        ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM written"),
        ("launch__registers_per_thread", "registers / thread"), ("launch__block_size", "block size"), ("launch__grid_size", "grid"),
        ("launch__shared_mem_per_block_dynamic", "dynamic smem / block"), ("smsp__inst_executed.sum", "warp instructions")]
print(f"# ncu --set full summary of `{rep}`\n")
print("| kernel | " + " | ".join(n for k, n in want if k in idx) + " |")
print("|---|" + "---:|" * sum(1 for k, _ in want if k in idx))
for r in rows[2:]:
    cells = []
    for k, _ in want:
        if k in idx:
            v, u = r[idx[k]], units[idx[k]]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:,.0f}" if f >= 1000 else f"{f:.1f}"
            except ValueError:
                pass
            cells.append(f"{v} {u}".strip())
    print(f"| `{r[idx['Kernel Name']][:44]}` | " + " | ".join(cells) + " |")
