#!/usr/bin/env python
"""Group the SASS of one kernel in an .ncu-rep (captured with --set full --import-source on) into runs of instructions
that executed about equally often, and print each run's share of executed instructions and of stall samples with its
commonest opcodes: a quick map of which loop costs what.   usage: ncu_sass_regions.py report.ncu-rep kernel_substring"""
import collections
import csv
import subprocess
import sys


def main():
    rep, pat = sys.argv[1], sys.argv[2]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    sections, cur = [], None
    for r in csv.reader(out.splitlines()):
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "data": []}
            sections.append(cur)
        elif cur is not None and cur["hdr"] is None:
            cur["hdr"] = r
        elif cur is not None and len(r) == len(cur["hdr"]):
            cur["data"].append(r)
    for sec in sections:
        if pat not in sec["name"]:
            continue
        hdr, data = sec["hdr"], sec["data"]
        isrc, iex, ismp = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
        total = sum(int(r[iex]) for r in data)
        samples = sum(int(r[ismp]) for r in data)
        print(sec["name"], "| warp instructions", total, "| samples", samples, "| SASS lines", len(data))
        runs = []
        for i, r in enumerate(data):
            ex, sm = int(r[iex]), int(r[ismp])
            if runs and abs(runs[-1]["ex"] - ex) <= 0.15 * max(runs[-1]["ex"], ex, 1):
                g = runs[-1]
                g["n"] += 1
                g["tot"] += ex
                g["smp"] += sm
                g["end"] = i
            else:
                runs.append({"start": i, "end": i, "ex": ex, "n": 1, "tot": ex, "smp": sm})
        for g in runs:
            if g["tot"] > 0.01 * total or g["smp"] > 0.02 * max(1, samples):
                ops = []
                for j in range(g["start"], g["end"] + 1):
                    w = data[j][isrc].split()
                    ops.append((w[1] if w[0].startswith("@") and len(w) > 1 else w[0]).split(".")[0])
                top = collections.Counter(ops).most_common(6)
                print("%5d-%-5d n=%-4d exec=%-9d inst %.3f  samples %.3f  %s" % (g["start"], g["end"], g["n"], g["ex"], g["tot"] / max(1, total),
                                                                               g["smp"] / max(1, samples), top))
        return
    print("no kernel matching", pat)


if __name__ == "__main__":
    main()
