#!/usr/bin/env python
"""One table row per kernel launch of an .ncu-rep captured with --set full: duration, DRAM bytes, issue utilisation, occupancy,
registers and the top stall reasons (cycles per issued instruction).   usage: ncu_summary.py report.ncu-rep [more.ncu-rep ...]"""
import csv
import json
import subprocess
import sys

WANT = {"gpu__time_duration.sum": "ms", "dram__bytes_read.sum": "rd_GB", "dram__bytes_write.sum": "wr_GB",
        "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue%", "sm__warps_active.avg.pct_of_peak_sustained_active": "warps%",
        "launch__registers_per_thread": "regs", "smsp__inst_executed.sum": "warp_inst",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed": "L2%", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed": "L1%"}
STALL = "smsp__average_warps_issue_stalled_"


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")][:70]}
        stalls = {}
        for i, h in enumerate(hdr):
            if h in WANT:
                v = float(r[i].replace(",", "")) if r[i] else 0.0
                u = units[i].lower()
                if WANT[h] == "ms":
                    v = v / 1e6 if u.startswith("ns") else (v / 1e3 if u.startswith("us") else v)
                if WANT[h].endswith("GB"):
                    v = {"byte": v / 1e9, "kbyte": v / 1e6, "mbyte": v / 1e3, "gbyte": v}.get(u, v)
                d[WANT[h]] = round(v, 3)
            elif h.startswith(STALL) and h.endswith("_per_issue_active.ratio") and r[i]:
                stalls[h[len(STALL):-len("_per_issue_active.ratio")]] = float(r[i])
        d["stalls"] = ", ".join("%s %.1f" % kv for kv in sorted(stalls.items(), key=lambda kv: -kv[1])[:4])
        yield d


def main():
    traffic = {}
    print("| kernel | ms | dram rd GB | dram wr GB | issue % | warps % | regs | warp instr | L2 % | top stalls (cycles per issue) |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for rep in sys.argv[1:]:
        for d in rows_of(rep):
            print("| `%s` | %s | %s | %s | %s | %s | %s | %.3g | %s | %s |" % (d["kernel"], d.get("ms"), d.get("rd_GB"), d.get("wr_GB"), d.get("issue%"),
                                                                            d.get("warps%"), int(d.get("regs", 0)), d.get("warp_inst", 0), d.get("L2%"), d["stalls"]))
            traffic[d["kernel"]] = int((d.get("rd_GB", 0) + d.get("wr_GB", 0)) * 1e9)
    print()
    print(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()
