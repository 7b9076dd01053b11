"""Generates stereo_matching_b200/csrc/median25.cuh: a compare-exchange network that puts the median of 25 values
into v[12].  Batcher's odd-even merge sort for 32 inputs; comparators that touch the seven padding inputs (+inf,
already in place) are no-ops and dropped, then a backward liveness pass keeps only those that can reach output 12.
Checked against sorted() on random inputs (with many ties) before the file is written."""
import os
import random


def batcher(n):
    pairs, p = [], 1
    while p < n:
        k = p
        while k >= 1:
            for j in range(k % p, n - k, 2 * k):
                for i in range(min(k, n - j - k)):
                    if (i + j) // (2 * p) == (i + j + k) // (2 * p):
                        pairs.append((i + j, i + j + k))
            k //= 2
        p *= 2
    return pairs


def main():
    pairs = [(a, b) for a, b in batcher(32) if a < 25 and b < 25]
    live, keep = {12}, []
    for a, b in reversed(pairs):
        if a in live or b in live:
            keep.append((a, b))
            live.update((a, b))
    keep.reverse()
    for _ in range(20000):
        v = [random.choice([0.0, 1.0, 2.0, random.random()]) for _ in range(25)]
        w = v[:]
        for a, b in keep:
            if w[a] > w[b]:
                w[a], w[b] = w[b], w[a]
        assert w[12] == sorted(v)[12]
    lines = ["// Median of 25 values in registers: Batcher's odd-even merge sort for 32 inputs with the comparators that touch the",
             "// seven padding inputs (+inf, already in place) or cannot reach output 12 removed - %d compare-exchanges." % len(keep),
             "// Generated and checked against sorted() on 20 000 random inputs by scripts/gen_median25.py.",
             "#pragma once", "", "namespace stereo {", "", "__device__ __forceinline__ float median25(float (&v)[25])", "{",
             "#define STEREO_CE(a, b) { const float lo = fminf(v[a], v[b]), hi = fmaxf(v[a], v[b]); v[a] = lo; v[b] = hi; }"]
    row = []
    for a, b in keep:
        row.append("STEREO_CE(%d, %d)" % (a, b))
        if len(row) == 6:
            lines.append("    " + " ".join(row))
            row = []
    if row:
        lines.append("    " + " ".join(row))
    lines += ["#undef STEREO_CE", "    return v[12];", "}", "", "}  // namespace stereo", ""]
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stereo_matching_b200", "csrc", "median25.cuh")
    with open(out, "w") as f:
        f.write("\n".join(lines))
    print(out, len(keep), "compare-exchanges")


if __name__ == "__main__":
    main()
