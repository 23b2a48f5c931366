"""per-kernel summary table of an ncu report (--set full) + optional launch list csv"""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.split("\n")))
hdr, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__registers_per_thread", "regs"), ("gpu__time_duration.sum", "time"),
        ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts"),
        ("l1tex__data_pipe_lsu_wavefronts.sum.pct_of_peak_sustained_elapsed", "lsu_pipe%"),
        ("smsp__inst_executed.sum", "warp_inst")]
idx = [(hdr.index(c) if c in hdr else None, n) for c, n in cols]
print(" | ".join(n + ("[" + units[i] + "]" if i is not None and units[i] else "") for i, n in idx))
for r in rows[2:]:
    if len(r) < len(hdr): continue
    out = []
    for i, n in idx:
        v = r[i] if i is not None else "-"
        if n == "kernel": v = v.split("(")[0].replace("void ", "")[:24]
        out.append(v)
    print(" | ".join(out))
