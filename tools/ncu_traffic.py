#!/usr/bin/env python
"""Extract per-launch DRAM traffic and duration of named kernels from an ncu report:
   python tools/ncu_traffic.py report.ncu-rep kernel_substring [...] -> JSON on stdout
(merge the output into profiles/r01_ncu_traffic.json, which bench.py reads for roofline.traffic)."""
import csv
import json
import subprocess
import sys

rep, names = sys.argv[1], sys.argv[2:]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


out = {}
for r in rows[2:]:
    k = r[col["Kernel Name"]]
    for n in names:
        if n in k and n not in out:
            rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
            wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
            out[n] = {"dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr,
                      "duration_ms_under_ncu": float(r[col["gpu__time_duration.sum"]]) *
                      {"ms": 1, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(units[col["gpu__time_duration.sum"]], 1),
                      "grid": r[col["Grid Size"]] if "Grid Size" in col else None, "report": rep.split("/")[-1]}
print(json.dumps(out, indent=1))
