"""Summarise one `ncu --csv` metrics pass over a profiled step: per-kernel launches, time, DRAM bytes, tensor-pipe
activity -> JSON (the file bench.py reads `roofline.traffic` from) and a share-of-step table on stdout.

    ncu --profile-from-start off --clock-control none --csv --log-file launches.csv \
        --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
        python tools/profile_step.py
    python tools/ncu_launch_summary.py launches.csv profiles/r2_gemm_traffic_2layer_step.json
"""
import collections
import csv
import json
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1], errors="replace")))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            hdr, start = r, i + 1
            break
    ik, im, iv, iu, iid = (hdr.index(c) for c in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
    per = collections.OrderedDict()            # launch id -> {name, metric: value}
    for r in rows[start:]:
        if len(r) <= iv:
            continue
        d = per.setdefault(r[iid], {"name": r[ik]})
        v = float(r[iv].replace(",", "")) if r[iv] not in ("", "n/a") else 0.0
        u = r[iu].lower()
        if r[im].startswith("dram__bytes"):
            v *= {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        if r[im] == "gpu__time_duration.sum":
            v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "nsecond": 1e-3, "usecond": 1, "msecond": 1e3}.get(u, 1)
        d[r[im]] = v
    agg = collections.OrderedDict()
    for d in per.values():
        m = re.search(r"([A-Za-z_0-9]+)(<[^(]*>)?\(", d["name"].replace("void ", "").replace("<unnamed>::", ""))
        key = (m.group(1) + (m.group(2) or "")) if m else d["name"][:60]
        a = agg.setdefault(key, {"launches": 0, "us": 0.0, "dram_bytes": 0.0, "_tw": 0.0})
        us = d.get("gpu__time_duration.sum", 0.0)
        a["launches"] += 1
        a["us"] += us
        a["dram_bytes"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        a["_tw"] += us * d.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 0.0)
    total = sum(a["us"] for a in agg.values())
    out = collections.OrderedDict()
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
        out[k] = {"launches": a["launches"], "us": round(a["us"], 3), "dram_bytes": a["dram_bytes"],
                  "tensor_active_pct": a["_tw"] / a["us"] if a["us"] else 0.0, "share_of_step": a["us"] / total}
        print(f"{a['us']:9.1f} us {a['launches']:4d}  {100 * a['us'] / total:5.1f}%  {a['dram_bytes'] / 1e6:9.1f} MB  {k[:80]}")
    print(f"total {total:.1f} us over {len(per)} launches")
    if len(sys.argv) > 2:
        json.dump(out, open(sys.argv[2], "w"), indent=1)


if __name__ == "__main__":
    main()
