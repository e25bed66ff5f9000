"""Top stalled SASS instructions of one kernel from an ncu report: python tools/ncu_hot.py rep.ncu-rep kernel_regex [N]"""
import csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kern}", "--print-source", "sass"],
                     capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(out) if l.startswith('"Address"'))
rows = list(csv.DictReader(out[start:]))
tot = sum(int(r["# Samples"] or 0) for r in rows)
stall_cols = [c for c in rows[0] if c.startswith("stall_") and "Not Issued" not in c]
agg = {c: sum(int(r[c] or 0) for r in rows) for c in stall_cols}
print("total samples", tot, {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:7]})
rows.sort(key=lambda r: -int(r["# Samples"] or 0))
for r in rows[:n]:
    top = sorted(((int(r[c] or 0), c) for c in stall_cols), reverse=True)[:2]
    print(f'{int(r["# Samples"]):6d} {100 * int(r["# Samples"]) / max(tot, 1):5.1f}%  {r["Source"][:70]:70s} {top}')
