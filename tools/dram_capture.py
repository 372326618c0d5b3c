#!/usr/bin/env python
"""Turns an ncu launch list (CSV) of ONE step of bench.py into profiles/<name>.json: DRAM bytes and serialised time per
kernel and per phase. bench.py reads `phase_bytes` of the newest capture of its workload for `roofline.traffic`.

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
        --log-file gpurun_out/x.csv python bench.py --workload config3 --steps 1 --warmup 0 --no-e2e --no-cpu
    python tools/dram_capture.py gpurun_out/x.csv profiles/r02_dram_config3_n1 "config3, N=1"

(writes <out>.csv = the launch list without ncu's banner lines, and <out>.json)
"""
import csv
import json
import re
import sys

PHASE_OF = [("k1_", "transpose"), ("k2_", "stats"), ("k3_plan", "sort"), ("k3_order", "sort"), ("k3_sort", "sort"),
            ("k3_walk", "count"), ("k3_ties", "count"), ("k4_", "effects"), ("k5_summaries", "summaries"),
            ("k5_", "minrank")]
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
         "ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "nsecond": 1e-6, "s": 1e3, "second": 1e3}


def main():
    src, out, what = sys.argv[1], sys.argv[2], sys.argv[3]
    lines = [ln for ln in open(src) if ln.startswith('"')]
    rows = list(csv.DictReader(lines))
    per_launch = {}
    for r in rows:
        m = re.search(r"(k\d_\w+)", r["Kernel Name"])
        if not m:
            continue  # the synthetic-data kernels and anything not part of a step
        name = m.group(1)
        val = float(r["Metric Value"].replace(",", "")) * SCALE[r["Metric Unit"]]
        d = per_launch.setdefault(int(r["ID"]), {"name": name})
        d[r["Metric Name"]] = val
    agg, phase_b, phase_ms = {}, {}, {}
    for d in per_launch.values():
        a = agg.setdefault(d["name"], {"launches": 0, "ms": 0.0, "dram_read": 0.0, "dram_write": 0.0})
        a["launches"] += 1
        a["ms"] += d.get("gpu__time_duration.sum", 0.0)
        a["dram_read"] += d.get("dram__bytes_read.sum", 0.0)
        a["dram_write"] += d.get("dram__bytes_write.sum", 0.0)
        ph = next(p for pre, p in PHASE_OF if d["name"].startswith(pre))
        phase_b[ph] = phase_b.get(ph, 0.0) + d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        phase_ms[ph] = phase_ms.get(ph, 0.0) + d.get("gpu__time_duration.sum", 0.0)
    open(out + ".csv", "w").writelines(lines)
    doc = {
        "source": "ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none, "
                  "one step of bench.py (%s): %s.csv" % (what, out),
        "phase_bytes": {k: int(v) for k, v in phase_b.items()},
        "phase_ms_serialised": {k: round(v, 3) for k, v in phase_ms.items()},
        "kernels": {k: {"launches": a["launches"], "ms": round(a["ms"], 3), "dram_read": int(a["dram_read"]),
                        "dram_write": int(a["dram_write"])} for k, a in agg.items()},
        "total_ms_serialised": round(sum(phase_ms.values()), 2),
        "total_dram_bytes": int(sum(phase_b.values())),
    }
    json.dump(doc, open(out + ".json", "w"), indent=1)
    print(json.dumps({k: doc[k] for k in ("phase_bytes", "phase_ms_serialised", "total_ms_serialised", "total_dram_bytes")}))


if __name__ == "__main__":
    main()
