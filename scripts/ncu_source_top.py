"""Top stall sites of a kernel from an `ncu --page source --csv` export (optionally gzipped):
`python scripts/ncu_source_top.py file.csv[.gz] [kernel index] [rows]`."""
import collections
import csv
import gzip
import io
import sys


def load(fn):
    txt = (gzip.open(fn, "rt") if fn.endswith(".gz") else open(fn)).read()
    out = []
    for b in txt.split('"Kernel Name",')[1:]:
        lines = b.split("\n")
        rows = list(csv.reader(io.StringIO("\n".join(lines[1:]))))
        hdr = rows[0]
        out.append((lines[0], hdr, [r for r in rows[1:] if len(r) == len(hdr)]))
    return out


def main():
    ks = load(sys.argv[1])
    which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    print("kernels in file:", [k[0][:70] for k in ks])
    name, hdr, data = ks[which]
    idx = {h: i for i, h in enumerate(hdr)}
    tot = sum(int(r[idx["# Samples"]]) for r in data)
    print(name[:100], "| SASS instructions", len(data), "| samples", tot)
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    agg = collections.Counter()
    for r in data:
        for h in stall_cols:
            agg[h] += int(r[idx[h]] or 0)
    print("stall reasons %:", {k[6:]: round(100 * v / tot, 1) for k, v in agg.most_common(9)})
    for r in sorted(data, key=lambda r: -int(r[idx["# Samples"]]))[:n]:
        st = {h[6:]: int(r[idx[h]] or 0) for h in stall_cols}
        print(f"{int(r[idx['# Samples']]):6d} {100 * int(r[idx['# Samples']]) / tot:5.1f}% {max(st, key=st.get):12s} {r[idx['Source']].strip()[:100]}")
    mix = collections.Counter()
    for r in data:
        op = r[idx["Source"]].strip().split()
        if op:
            o = op[1] if op[0].startswith("@") else op[0]
            mix[o.split(".")[0]] += int(r[idx["Instructions Executed"]] or 0)
    te = sum(mix.values())
    print("executed warp instructions", te, {k: round(100 * v / te, 1) for k, v in mix.most_common(16)})


if __name__ == "__main__":
    main()
