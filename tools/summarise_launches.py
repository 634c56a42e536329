"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: one bench step
(the rows between two consecutive decode_filter launches) aggregated per kernel.

    python tools/summarise_launches.py gpurun_out/launches.csv [step_index] > profiles/rNN_launches.txt
"""
import collections
import csv
import sys


def main(path, which=2):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    dec = [i for i, r in enumerate(rows) if "decode_filter" in r["Kernel Name"]]
    if len(dec) < 2:
        step = rows
    else:
        which = min(which, len(dec) - 2)
        step = rows[dec[which]:dec[which + 1]]
    agg = collections.OrderedDict()
    for r in step:
        n = r["Kernel Name"].split("(")[0][:64]
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += float(r["Metric Value"]) / 1e3
    tot = sum(a[1] for a in agg.values())
    print(f"# one bench step out of {path}: {len(step)} launches, {tot:.1f} us serialised (cold-cache, per-launch)")
    print("# per-launch times under ncu are serialised: compare SHARES with bench.py's stages_ms_per_step")
    print(f"{'kernel':66s} {'launches':>8s} {'us':>10s} {'share':>6s}")
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:66s} {a[0]:8d} {a[1]:10.1f} {a[1] / tot:6.3f}")
    g = [float(r["Metric Value"]) / 1e3 for r in step if "gemm_bf16" in r["Kernel Name"]]
    if len(g) >= 6:
        print("# first GEMM launches of the step (patch-embed, then qkv/proj/fc1/fc2 per block), us:")
        print("# " + " ".join(f"{x:.1f}" for x in g[:9]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 2)
