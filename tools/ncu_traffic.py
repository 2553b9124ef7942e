"""Summarise the per-launch DRAM traffic pass of tools/ncu_capture.sh (no GPU needed):
    python tools/ncu_traffic.py gpurun_out/traffic_<tag>.csv profiles/<name>.json
Per kernel name: launches, total ms, DRAM bytes (read + write) per launch and in total.  bench.py reads the JSON to fill
`roofline.traffic` (ncu numbers are cold-cache and serialised: they describe traffic, never time)."""
import collections
import csv
import json
import re
import sys


def main():
    src, dst = sys.argv[1], sys.argv[2]
    rows = list(csv.reader(open(src)))
    for i, r in enumerate(rows):
        if "Kernel Name" in r:
            hdr, start = r, i + 1
            break
    ki, mi, vi, ui, idi = (hdr.index(k) for k in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
    per = collections.defaultdict(dict)
    names = {}
    for r in rows[start:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        u = r[ui]
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1)
        per[r[idi]][r[mi]] = v * scale
        names[r[idi]] = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("wc::", "")
    agg = collections.defaultdict(lambda: {"launches": 0, "ms": 0.0, "dram_bytes": 0.0})
    for k, m in per.items():
        a = agg[names[k]]
        a["launches"] += 1
        a["ms"] += m.get("gpu__time_duration.sum", 0.0)
        a["dram_bytes"] += m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0)
    out = {}
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
        out[k] = {"launches": a["launches"], "ms_under_ncu": round(a["ms"], 3), "dram_bytes_total": int(a["dram_bytes"]),
                  "dram_bytes_per_launch": int(a["dram_bytes"] / a["launches"])}
    json.dump({"source": src, "note": "ncu --clock-control none, one msc+flip step (8 dataset images), cold-cache serialised "
               "launches", "kernels": out}, open(dst, "w"), indent=1)
    for k, a in list(out.items())[:14]:
        print(f"{k[:48]:48s} x{a['launches']:4d} {a['ms_under_ncu']:8.3f} ms  {a['dram_bytes_per_launch'] / 1e6:9.2f} MB/launch")


if __name__ == "__main__":
    main()
