#!/usr/bin/env python
"""ncu_traffic.py <file.ncu-rep> -- DRAM bytes and duration per captured launch (ncu --page raw --csv) as JSON."""
import csv
import io
import json
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
col = {n: i for i, n in enumerate(hdr)}
res = []
for r in rows[2:]:
    def f(name):
        try:
            return float(r[col[name]].replace(",", ""))
        except Exception:
            return None
    unit_r = rows[1][col["dram__bytes_read.sum"]] if "dram__bytes_read.sum" in col else ""
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit_r, 1)
    unit_w = rows[1][col["dram__bytes_write.sum"]] if "dram__bytes_write.sum" in col else ""
    scale_w = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit_w, 1)
    res.append({"kernel": r[col["Kernel Name"]][:60], "grid": r[col["Grid Size"]] if "Grid Size" in col else None,
                "dram_read_bytes": (f("dram__bytes_read.sum") or 0) * scale, "dram_write_bytes": (f("dram__bytes_write.sum") or 0) * scale_w,
                "duration_under_ncu (unit of the report, us)": f("gpu__time_duration.sum")})
print(json.dumps({"source": sys.argv[1], "launches": res}, indent=1))
