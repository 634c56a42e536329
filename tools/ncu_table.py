"""Markdown table of the metrics DESIGN.md quotes out of an .ncu-rep (`ncu --set full`), one row per captured launch.

    python tools/ncu_table.py gpurun_out/X.ncu-rep > profiles/rNN_X_ncu.md
"""
import csv
import io
import subprocess
import sys

COLS = [
    ("time", "gpu__time_duration.sum", 1.0),
    ("SM clk", "sm__cycles_elapsed.avg.per_second", 1.0),
    ("DRAM rd", "dram__bytes_read.sum", 1.0),
    ("DRAM wr", "dram__bytes_write.sum", 1.0),
    ("DRAM %", "dram__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("tensor pipe (hmma) % active", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", 1.0),
    ("MUFU (xu) % active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", 1.0),
    ("issue slots % active", "sm__inst_issued.avg.pct_of_peak_sustained_active", 1.0),
    ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    ("regs", "launch__registers_per_thread", 1.0),
    ("grid", "launch__grid_size", 1.0),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, data = rows[0], rows[1], rows[2:]

    def find(metric):
        if metric in hdr:
            return hdr.index(metric)
        for i, h in enumerate(hdr):
            if h.endswith("." + metric):
                return i
        return None

    idx = [(name, find(metric)) for name, metric, _ in COLS]
    ki = hdr.index("Kernel Name")
    print(f"# {path}: `ncu --set full --clock-control none` (per-launch, serialised, cold cache; NOT bench numbers)\n")
    print("| # | kernel | " + " | ".join(n for n, i in idx) + " |")
    print("|---|---|" + "---|" * len(idx))
    for k, r in enumerate(data):
        name = r[ki].replace("<unnamed>::", "").replace("void ", "").split("(")[0]
        vals = []
        for n, i in idx:
            if i is None or r[i] == "":
                vals.append("-")
                continue
            try:
                v = float(r[i].replace(",", ""))
                t = f"{v:.0f}" if n in ("regs", "grid") or v >= 100 else (f"{v:.2f}" if v < 10 else f"{v:.1f}")
                if n in ("time", "SM clk", "DRAM rd", "DRAM wr"):
                    t += " " + units[i]
                vals.append(t)
            except ValueError:
                vals.append(r[i])
        print(f"| {k} | `{name}` | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1])
