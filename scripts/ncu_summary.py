"""Per-kernel summary of an ncu report (raw page): duration, DRAM bytes, DRAM %, L1TEX %, issue %, occupancy."""
import csv, subprocess, sys
rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
keys = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"), ("dram__bytes_write.sum", "dram_wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%"),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps%"),
        ("launch__registers_per_thread", "regs"), ("launch__grid_size", "grid"),
        ("smsp__inst_executed.sum", "inst"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts")]
for r in rows[2:]:
    name = r[h.index("Kernel Name")].replace("void unnamed>::", "")
    print(name[:90])
    print("    " + "  ".join("%s=%s%s" % (lbl, r[h.index(k)][:10], u[h.index(k)].replace("register/thread", "").replace("byte", "B")[:6])
                            for k, lbl in keys if k in h))
