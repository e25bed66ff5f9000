"""Print the headline and per-kernel microseconds of bench JSON lines: python tools/show_bench.py gpurun_out/a.json ..."""
import json, sys
for n in sys.argv[1:]:
    try:
        d = json.loads([l for l in open(n).read().splitlines() if l.startswith("{")][-1])
        print(n, round(d["value"] / 1e6, 2), "M/s", round(d["ms_per_step"] * 1e3, 1), "us  e2e", round(d["e2e"]["value"] / 1e6, 2),
              {k: round(v, 1) for k, v in d.get("roofline", {}).get("kernel_us", {}).items()}, "loss", round(d.get("final_loss", 0), 4))
    except Exception as e:
        print(n, "ERR", e)
