"""Top stall sites of one kernel out of `ncu -i X.ncu-rep --page source --csv --kernel-name K` (SASS view)."""
import csv
import sys


def main(path, ntop=40, section=0):
    rows = list(csv.reader(open(path)))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
    hdr = rows[starts[section] + 1]
    data = [r for r in rows[starts[section] + 2:starts[section + 1]] if len(r) == len(hdr)]
    si, ie = hdr.index("# Samples"), hdr.index("Instructions Executed")
    stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[si] or 0) for r in data)
    print(f"{rows[starts[section]][1][:80]}: {tot} samples over {len(data)} SASS instructions")
    agg = {hdr[i]: sum(int(r[i] or 0) for r in data) for i in stall}
    print("  ".join(f"{k[6:]}={v / tot:.2f}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:9]))
    top = sorted(range(len(data)), key=lambda j: -int(data[j][si] or 0))[:ntop]
    for j in sorted(top):
        r = data[j]
        st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall), reverse=True)[:2]
        print(f"{j:5d} {r[1].strip()[:72]:72s} {int(r[si]):6d} {r[ie]:>8s} {st}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40, int(sys.argv[3]) if len(sys.argv) > 3 else 0)
