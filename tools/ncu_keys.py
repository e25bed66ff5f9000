"""Key metrics per kernel from an .ncu-rep: python tools/ncu_keys.py gpurun_out/x.ncu-rep"""
import csv, subprocess, sys, io
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
keys = {"gpu__time_duration.sum": "us", "sm__cycles_elapsed.max": "cyc", "smsp__inst_executed.sum": "inst",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue%", "sm__warps_active.avg.pct_of_peak_sustained_active": "warps%",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum": "smem_wf", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "bank_conf",
        "sm__inst_executed_pipe_fma.sum": "fma_inst", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma%",
        "launch__registers_per_thread": "regs", "launch__grid_size": "grid", "launch__block_size": "block",
        "dram__bytes_read.sum": "dram_rd", "dram__bytes_write.sum": "dram_wr",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio": "st_long",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio": "st_short",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio": "st_mio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio": "st_bar",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio": "st_wait",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio": "st_math",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio": "st_notsel",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio": "st_lg",
        "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio": "st_sleep",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio": "st_membar",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio": "st_branch",
        "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio": "st_disp",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio": "st_noinst",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor%"}
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    name = r[ki].split("(")[0]
    parts = []
    for k, short in keys.items():
        if k in hdr:
            v = r[hdr.index(k)]
            try:
                f = float(v.replace(",", ""))
                v = f"{f:.3g}" if abs(f) < 1e4 else f"{f:.4g}"
            except ValueError:
                pass
            parts.append(f"{short}={v}")
    print(name, " ".join(parts))
