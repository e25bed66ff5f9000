"""Per-kernel summary of an `ncu --set full` report: duration, DRAM traffic, pipe utilisation, occupancy, top stalls."""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout.splitlines()
rows = list(csv.reader(raw))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_throughput_pct"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_pipe_pct"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor_pipe_pct"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps_active_pct"),
        ("launch__registers_per_thread", "regs"), ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct")]
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0]
    parts = []
    for key, label in want:
        if key in idx and r[idx[key]] not in ("", "n/a"):
            parts.append(f"{label}={r[idx[key]]} {units[idx[key]]}".strip())
    print(f"{name}: " + ", ".join(parts))
