#!/usr/bin/env python
"""Merge the per-(world, size) files of benchmarks/micro_grid.sh into one JSON + a markdown table."""
import glob
import json
import os
import re
import sys


def main(src="gpurun_out", out_json="profiles/micro_grid_r2.json", out_md="profiles/micro_grid_r2.md"):
    rows = []
    for path in sorted(glob.glob(os.path.join(src, "micro_grid_w*_*.json"))):
        m = re.search(r"micro_grid_w(\d+)_(\d+)_(ours|reference)\.json", path)
        if not m:
            continue
        for r in json.load(open(path)):
            r = dict(r, world=int(m.group(1)), elements=int(m.group(2)), impl=m.group(3))
            rows.append(r)
    json.dump(rows, open(out_json, "w"), indent=1)
    key = lambda r: (r["elements"], r["world"], r["codec"])
    ours = {key(r): r for r in rows if r["impl"] == "ours"}
    ref = {key(r): r for r in rows if r["impl"] == "reference"}
    lines = ["| elements | W | codec | ours ms | reference ms | speed-up | NCCL all-reduce ms | clocks ours (MHz, reasons) |", "|---|---|---|---|---|---|---|---|"]
    for k in sorted(ours):
        o = ours[k]
        if o["codec"] in ("nccl-allreduce",):
            continue
        base = k[2].split("-")[0]
        r = ref.get((k[0], k[1], base))
        nccl = ours.get((k[0], k[1], "nccl-allreduce")) or ref.get((k[0], k[1], "nccl-allreduce-flat"))
        ck = o.get("clocks") or {}
        lines.append("| %d | %d | %s | %.3f | %s | %s | %s | %s %s |" % (
            k[0], k[1], k[2], o["ms"], "%.3f" % r["ms"] if r else "-", "%.1fx" % (r["ms"] / o["ms"]) if r else "-",
            "%.3f" % nccl["ms"] if nccl else "-", ck.get("sm_mhz"), ",".join(ck.get("reasons") or []) or "none"))
    open(out_md, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines[:12]))
    print("... %d rows -> %s, %s" % (len(rows), out_json, out_md))


if __name__ == "__main__":
    main(*sys.argv[1:])
